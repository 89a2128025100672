"""bench.py -- the headline metric of BASELINE.json on B200: rays/sec at 64+128 samples per ray (render, train).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--mode MODE] [--precision fp16|fp32]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N --steps K --warmup W

Modes (every mode prints ONE JSON line on rank 0; `--mode train` is the default the driver runs):
  train   BASELINE configs[1]: lego full-res 800x800 training step, N_rand = 1024 rays per GPU per step (weak scaling;
          `--scaling strong` splits ONE 1024-ray batch over the ranks), 8x256 coarse + fine MLP, 64 stratified + 128
          importance samples, forward + loss + backward + [gradient all-reduce] + Adam.  A step = one iteration.
          The line also carries a "render" object: configs[0] measured in the same run (same N, row-block sharded).
  render  BASELINE configs[0]: lego 400x400 render_only, chunk 32768, the 40 pose_spherical turntable poses, 8x256
          coarse + fine, through render_path (frames land in pinned host memory).  A step = one pose (frame).
  render800  the same at 800x800 (640 000 rays per pose).
  wheel   configs[2]: robot wheel scene 1280x720, K with fx != fy from the parser's transforms.json, NDC, 60 video poses.
  fern    configs[3]: LLFF fern 504x378, NDC + use_viewdirs, 120-pose spiral render, plus an N_rand = 4096 train step.
  sweep   configs[4]: 2^12 .. 2^22 rays, end-to-end render and per-kernel rates (MLP tensor roofline, compositing HBM).
At N > 1 rendering is partitioned by contiguous row blocks (ray blocks for `sweep`) with no collective: strong scaling.

`value`  : rays/s over all ranks with the step's inputs already resident in HBM.
`e2e`    : the same through the public API with HOST inputs: train = pinned rays/targets H2D + loss D2H every step
           (TrainStep.step); render = host poses in, frames out to pinned host memory (render_path).
`roofline`: the dominant kernel timed alone with CUDA events on its launch stream, against MEASURED_PEAKS.json.
`cpu_baseline` / `--impl reference`: the reference's OWN PyTorch code (oracle/_ref, staged by oracle/build_ref.py;
           the 8x256 graph is the nn.Module restatement of oracle/ref_arm.py) on this box's host cores.
"""
import argparse
import ctypes as C
import json
import math
import os
import subprocess
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

N_SAMPLES, N_IMPORTANCE = 64, 128
FLOPS_PER_POINT = 1186816                                   # 8x256 graph, 2*MAC, unpadded
PTS_PER_RAY = N_SAMPLES + N_SAMPLES + N_IMPORTANCE          # 64 coarse + 192 fine evaluations
FLOPS_PER_RAY_RENDER = FLOPS_PER_POINT * PTS_PER_RAY        # 303.82 MFLOP
FLOPS_PER_RAY_TRAIN = 3 * FLOPS_PER_RAY_RENDER              # fwd + dgrad + wgrad: 911.47 MFLOP
METRIC = "rays/sec at 64+128 samples/ray"

WORKLOADS = {
    "train": "lego 800x800 train step, N_rand=1024 rays/GPU, 8x256 coarse+fine, 64+128 samples (BASELINE configs[1])",
    "render": "lego 400x400 render_only, chunk 32768, 40 turntable poses, 8x256 coarse+fine, 64+128 samples (BASELINE configs[0])",
    "render800": "lego 800x800 render, chunk 32768, turntable poses, 8x256 coarse+fine, 64+128 samples",
    "wheel": "robot wheel 1280x720 pose path (parser_instant_ngp transforms.json, fx != fy, NDC), 60 video poses (BASELINE configs[2])",
    "fern": "fern LLFF 504x378, NDC + use_viewdirs, 8x256, 120-pose spiral render + N_rand=4096 train step (BASELINE configs[3])",
    "sweep": "ray-batch sweep 2^12..2^22 rays, 8x256 coarse+fine, 64+128 samples (BASELINE configs[4])",
}


def peaks():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        p["source"] = "measured"
        return p
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


def shared_config(mode, args, world):
    """The workload description BOTH arms print (the driver compares the two `config` objects)."""
    cfg = {"workload": WORKLOADS[mode], "n_samples": N_SAMPLES, "n_importance": N_IMPORTANCE, "mode": mode}
    if mode == "train":
        per = args.n_rand // world if args.scaling == "strong" else args.n_rand
        cfg.update(rays_per_step_per_gpu=per, perturb=1, raw_noise_std=0.0, optimizer="Adam lr 5e-4, inside the timed region",
                   l2="per-step working set (activation records of 262144 points) exceeds the 126 MB L2; a different ray batch every step")
    else:
        cfg.update(chunk=32768, l2="per-chunk working set (raw network output of 6.3 M points = 100 MB, depths, weights) "
                                   "exceeds the 126 MB L2 together with the next chunk's; a different pose every step")
    return cfg


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx, self.rows, self.proc = str(gpu_index), [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "20", "-i", self.idx], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.05)
        self.proc.terminate()
        sm = [float(r[1]) for r in self.rows if len(r) >= 8 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 8 and r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 8 for i in range(4) if r[4 + i].lower() == "active"})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def gpu_index(local):
    if "CUDA_VISIBLE_DEVICES" in os.environ:
        return os.environ["CUDA_VISIBLE_DEVICES"].split(",")[local]
    return torch.cuda.current_device()


# ====================================================================================== reference / CPU arm
def cpu_reference_train(n_rays, steps, warmup, budget_s):
    """rays/s of the reference's own run_one_iter_of_nerf + loss + backward + Adam on the host cores (oracle/ref_arm.py).
    The step is the full n_rays batch unless the measured warm-up step says K + W of them would overrun budget_s; then
    a bounded sample of the batch (fewer rays per step, stated in `sample`)."""
    from oracle import ref_arm
    st = ref_arm.ReferenceTrainStep()
    n = n_rays
    t_probe, _ = st.step(min(n, 128))                      # also pays the one-off lazy initialisation
    t_probe, _ = st.step(min(n, 128))
    per_ray = t_probe / min(n, 128)
    total = steps + max(0, warmup - 1)
    if per_ray * n * total > budget_s:
        n = max(64, int(budget_s / (per_ray * total)) // 64 * 64)
    for _ in range(max(0, warmup - 1)):
        st.step(n)
    times = [st.step(n)[0] for _ in range(steps)]
    return {"rate": n / float(np.mean(times)), "times": times, "n": n, "threads": st.threads,
            "sample": ("the full %d-ray step" % n if n == n_rays else "%d rays per step of the %d-ray batch" % (n, n_rays))
                      + " (fwd + loss + bwd + Adam) through the reference's run_one_iter_of_nerf, %d threads" % st.threads}


def cpu_reference_render(H, W, focal, steps, warmup, budget_s):
    """rays/s of the reference rendering row blocks of config 1 poses on the host cores."""
    from oracle import ref_arm
    from oracle.nerf_oracle import pose_spherical
    rr = ref_arm.ReferenceRender(H, W, focal)
    c2w = pose_spherical(30.0, -30.0, 4.0)
    pr = min(8, H)
    t_probe, _ = rr.render_rows(c2w, 0, pr)
    t_probe, _ = rr.render_rows(c2w, 0, pr)
    per_ray = 1.3 * t_probe / (pr * W)   # small batches run a little faster per ray than full chunks
    total = steps + max(0, warmup - 1)
    rows = H
    if per_ray * H * W * total > budget_s:
        rows = max(1, int(budget_s / (per_ray * W * total)))
    for _ in range(max(0, warmup - 1)):
        rr.render_rows(c2w, 0, rows)
    times = [rr.render_rows(pose_spherical(-180.0 + 9.0 * i, -30.0, 4.0), 0, rows)[0] for i in range(steps)]
    return {"rate": rows * W / float(np.mean(times)), "times": times, "n": rows * W, "threads": rr.threads,
            "sample": ("the full frame" if rows == H else "rows [0,%d) of each %dx%d pose" % (rows, H, W))
                      + " through the reference's get_ray_bundle + run_one_iter_of_nerf, %d threads" % rr.threads}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    world = int(os.environ.get("WORLD_SIZE", "1"))
    mode = args.mode
    if mode == "train":
        r = cpu_reference_train(args.n_rand, max(1, args.steps), args.warmup, budget_s=150.0)
    elif mode in ("render", "render800"):
        hw = 400 if mode == "render" else 800
        r = cpu_reference_render(hw, hw, 555.555 * hw / 400.0, max(1, args.steps), args.warmup, budget_s=150.0)
    else:
        print(json.dumps({"impl": "reference", "unavailable": "the reference arm covers --mode train / render / render800"}))
        return
    out = {
        "impl": "reference", "metric": METRIC + (" (train step)" if mode == "train" else " (render)"), "value": r["rate"],
        "unit": "rays/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * float(np.mean(r["times"])), "higher_is_better": True,
        "scaling": args.scaling if mode == "train" else "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": shared_config(mode, args, world),
        "cpu_baseline": {"value": r["rate"], "unit": "rays/s", "cores": r["threads"], "kind": "reference", "sample": r["sample"]},
        "e2e": {"value": r["rate"], "unit": "rays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(out), flush=True)


# ====================================================================================== kernel rooflines
def _profiled_traffic(fname):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch from a committed `ncu --set full` summary; None if absent."""
    try:
        tot = 0.0
        for line in open(os.path.join(ROOT, "profiles", fname)):
            parts = line.strip().split(",")
            if len(parts) == 3 and parts[0] in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                tot += float(parts[2]) * {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1.0}[parts[1]]
        return tot or None
    except Exception:
        return None


def _profiled_traffic_backward():
    """DRAM bytes (read + write) of the fine pass's dgrad + wgrad launches from the committed ncu list of one train step
    (profiles/r2_train_kernels_ncu.csv, the section captured after the 1-bit masks); None if absent."""
    try:
        unit = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}
        lines = open(os.path.join(ROOT, "profiles", "r2_train_kernels_ncu.csv")).read().split("# AFTER")[1].splitlines()[1:]
        launches, cur = [], None
        for ln in lines:
            k, m, u, v = ln.rsplit(",", 3)
            if cur is None or m in cur["seen"] or k != cur["k"]:
                cur = {"k": k, "t": 0.0, "b": 0.0, "seen": set()}
                launches.append(cur)
            cur["seen"].add(m)
            if m == "gpu__time_duration.sum":
                cur["t"] = float(v)
            elif m.startswith("dram__bytes"):
                cur["b"] += float(v) * unit[u]
        tot = 0.0
        for name in ("mlp_tc_dgrad", "mlp_tc_wgrad"):
            tot += max((l for l in launches if l["k"].startswith(name)), key=lambda l: l["t"])["b"]
        return tot or None
    except Exception:
        return None


def _events_time(fn, reps, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e-3 / reps


def forward_kernel_roofline(model, prec_name, n_rays=32768, reps=5):
    """The fused MLP forward kernel alone over one fine-pass chunk (n_rays x 192 points); TFLOP/s, seconds, points."""
    from nerf_kinematics_b200 import ops
    from nerf_kinematics_b200._lib import check, lib
    dev = model._flat.device
    S = N_SAMPLES + N_IMPORTANCE
    n = n_rays
    o = torch.randn(n, 3, device=dev) * 0.5 + torch.tensor([0, 0, 4.0], device=dev)
    d = torch.nn.functional.normalize(torch.randn(n, 3, device=dev), dim=-1)
    rays = ops.make_ray_batch(o, d, 2.0, 6.0, use_viewdirs=True)
    z = ops.sample_coarse(rays, S)
    raw = torch.empty(n, S, 4, device=dev)
    h, prec = model.prepare(prec_name)
    nbytes = lib().nb_mlp_workspace_bytes(h, prec, n * S, 0)
    ws = torch.empty(max(nbytes, 1), dtype=torch.uint8, device=dev)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def run():
        check(lib().nb_mlp_forward(h, prec, C.c_void_p(rays.data_ptr()), 11, C.c_void_p(z.data_ptr()), n, S,
                                   C.c_void_p(raw.data_ptr()), C.c_void_p(ws.data_ptr()), nbytes, 0, st))
    t = _events_time(run, reps)
    return n * S * FLOPS_PER_POINT / t / 1e12, t, n * S


def train_kernel_rooflines(model, prec_name, n_rays, reps=10):
    """The train step's own big launches, timed alone on the fine pass of one step (n_rays x 192 points):
    forward with activation records, and the backward call (dgrad chain + wgrad + reduce)."""
    from nerf_kinematics_b200 import ops
    from nerf_kinematics_b200._lib import check, lib
    dev = model._flat.device
    S = N_SAMPLES + N_IMPORTANCE
    n = n_rays
    o = torch.randn(n, 3, device=dev) * 0.5 + torch.tensor([0, 0, 4.0], device=dev)
    d = torch.nn.functional.normalize(torch.randn(n, 3, device=dev), dim=-1)
    rays = ops.make_ray_batch(o, d, 2.0, 6.0, use_viewdirs=True)
    z = ops.sample_coarse(rays, S)
    raw = torch.empty(n, S, 4, device=dev)
    d_master = torch.randn(n * S, 4, device=dev) * 1e-3
    d_raw = d_master.clone()
    grad = torch.zeros(model.n_params, device=dev)
    h, prec = model.prepare(prec_name)
    nbytes = lib().nb_mlp_workspace_bytes(h, prec, n * S, 1)
    ws = torch.empty(max(nbytes, 1), dtype=torch.uint8, device=dev)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def fwd():
        check(lib().nb_mlp_forward(h, prec, C.c_void_p(rays.data_ptr()), 11, C.c_void_p(z.data_ptr()), n, S,
                                   C.c_void_p(raw.data_ptr()), C.c_void_p(ws.data_ptr()), nbytes, 1, st))

    def bwd():
        check(lib().nb_mlp_backward(h, prec, n * S, C.c_void_p(d_raw.data_ptr()), C.c_void_p(ws.data_ptr()), nbytes,
                                    C.c_void_p(grad.data_ptr()), st))
    t_f = _events_time(fwd, reps)
    t_b = _events_time(bwd, reps)
    fl = n * S * FLOPS_PER_POINT
    return {"points": n * S, "forward_ms": 1e3 * t_f, "forward_tflops": fl / t_f / 1e12,
            "backward_ms": 1e3 * t_b, "backward_tflops": 2 * fl / t_b / 1e12}


# ====================================================================================== render measurement
def lego_models(dev, lx=10, seed=0):
    import nerf_kinematics_b200 as nk
    torch.manual_seed(seed)  # identical replicas on every rank
    mc = nk.NeRF(input_ch=3 + 6 * lx).to(dev)
    mf = nk.NeRF(input_ch=3 + 6 * lx).to(dev)
    return mc, mf


def render_kwargs_for(mc, mf, prec_name, *, ndc, near, far, use_viewdirs=True, white_bkgd=True):
    return dict(network_query_fn=None, perturb=0.0, N_importance=N_IMPORTANCE, network_fine=mf, N_samples=N_SAMPLES,
                network_fn=mc, use_viewdirs=use_viewdirs, white_bkgd=white_bkgd, raw_noise_std=0.0, ndc=ndc, lindisp=False,
                near=near, far=far, precision=prec_name)


def measure_render(H, W, K, poses, kw, chunk, steps, warmup, rank, world, barrier, dist_max, clocks=True, local=0):
    """Render `steps` frames (cycling through `poses`), this rank's row block only.
    resident: poses already on the device, frames stay in HBM (FrameRenderer);  e2e: render_path with host poses,
    frames copied to pinned host memory inside the timed region.  Returns a dict of seconds / counts."""
    import nerf_kinematics_b200 as nk
    from nerf_kinematics_b200 import ops, parallel
    from nerf_kinematics_b200.run_nerf import _frame_renderer
    dev = next(kw["network_fn"].parameters()).device
    r0, r1 = parallel.row_block(H, rank, world)
    rk = {k: v for k, v in kw.items() if k not in ("ndc", "near", "far", "use_viewdirs")}
    with torch.no_grad():
        fr = _frame_renderer(H, W, K, chunk, kw["ndc"], kw["near"], kw["far"], kw["use_viewdirs"], rk, rows=(r0, r1))
    assert fr is not None
    M = len(poses)
    poses_dev = [torch.as_tensor(np.asarray(p, dtype=np.float32)[:3, :4]).to(dev) for p in poses]
    sel = [(i % M) for i in range(warmup + steps)]
    with torch.no_grad():
        for i in sel[:warmup]:
            fr.render(poses_dev[i])
        barrier()
        ops.launch_count_reset()
        rep0 = fr.replays
        sampler = ClockSampler(gpu_index(local)) if clocks else None
        if sampler:
            sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in sel[warmup:]:
            fr.render(poses_dev[i])
        e1.record()
        barrier()
        clk = sampler.stop() if sampler else None
        t_res = dist_max(e0.elapsed_time(e1) * 1e-3)
        launches = ops.launch_count() + (fr.launches_per_frame or 0) * (fr.replays - rep0)
        # ---- e2e: the public call, host poses -> pinned host frames
        host_poses = [np.asarray(poses[i], dtype=np.float32) for i in sel[warmup:]]
        nk.render_path(host_poses[:max(1, min(2, len(host_poses)))], (H, W, K), chunk, kw, row_block=(rank, world))
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        rgbs, disps = nk.render_path(host_poses, (H, W, K), chunk, kw, row_block=(rank, world))
        e1.record()
        barrier()
        t_e2e = dist_max(e0.elapsed_time(e1) * 1e-3)
    rays_rank = (r1 - r0) * W
    return {"t_res": t_res, "t_e2e": t_e2e, "launches": int(launches), "clocks": clk, "rays_per_frame": H * W,
            "rays_rank": rays_rank, "h2d": 48, "d2h": rays_rank * 16, "frame_mean": float(rgbs.mean()),
            "finite": float(np.isfinite(rgbs).mean())}   # fraction of finite rgb values (NDC-degenerate poses give NaN rays, like the reference)


def render_summary(m, steps, world, peak):
    sustained = peaks().get("bf16_tflops_sustained")
    rate = m["rays_per_frame"] * steps / m["t_res"]
    e2e = m["rays_per_frame"] * steps / m["t_e2e"]
    return {"value": rate, "unit": "rays/s", "ms_per_frame": 1e3 * m["t_res"] / steps,
            "e2e": {"value": e2e, "unit": "rays/s", "h2d_bytes_per_step": m["h2d"] * world,
                    "d2h_bytes_per_step": m["rays_per_frame"] * 16, "ms_per_step": 1e3 * m["t_e2e"] / steps},
            "roofline": {"bound": "tensor", "kernel": "whole frame (all kernels of render())",
                         "achieved": rate / world * FLOPS_PER_RAY_RENDER / 1e12, "peak": peak, "unit": "TFLOP/s",
                         "frac": rate / world * FLOPS_PER_RAY_RENDER / 1e12 / peak,
                         "peak_sustained": sustained,   # a frame is a long step: the power-capped clock applies
                         "frac_of_sustained": (rate / world * FLOPS_PER_RAY_RENDER / 1e12 / sustained) if sustained else None},
            "frames": steps, "gpu_launches": m["launches"], "finite": m["finite"]}


# ====================================================================================== our arm
def setup_dist(args):
    import torch.distributed as dist
    from nerf_kinematics_b200 import parallel
    rank, local, world = parallel.init_distributed("nccl")
    assert world == args.gpus or world == 1, "launch with torchrun --nproc-per-node %d" % args.gpus
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def dist_max(x):
        t = torch.tensor([x], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)

    def finish():
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
    return rank, local, world, dev, barrier, dist_max, finish


def run_train(args):
    import torch.distributed as dist
    import nerf_kinematics_b200 as nk
    from nerf_kinematics_b200 import ops
    from nerf_kinematics_b200.poses import blender_render_poses, pose_spherical
    from nerf_kinematics_b200.trainer import TrainStep

    threads = os.cpu_count() or 1
    world_env = int(os.environ.get("WORLD_SIZE", "1"))
    cpu = None
    if world_env == 1 and not args.no_cpu_baseline:
        # the CPU leg runs BEFORE any GPU / process-group work so nothing else competes for the host cores
        cpu = cpu_reference_train(args.n_rand, 2, 1, budget_s=25.0)
    rank, local, world, dev, barrier, dist_max, finish = setup_dist(args)
    H = W = 800
    FOCAL = 1111.111
    K, Wm = args.steps, max(3, args.warmup)
    n_rand = args.n_rand // world if args.scaling == "strong" else args.n_rand
    mc, mf = lego_models(dev)
    prec_name = args.precision
    step = TrainStep(mc, mf, N_samples=N_SAMPLES, N_importance=N_IMPORTANCE, perturb=1.0, raw_noise_std=0.0,
                     white_bkgd=True, precision=prec_name, lrate=5e-4, lrate_decay=500, world_size=world,
                     cuda_graph=not args.no_graph, seed=1234 + rank)  # jitter + inverse-CDF uniforms drawn in-kernel (f4)
    # ---- synthetic inputs of the named shape: one 800x800 pose, N_rand rays per step without replacement
    c2w = torch.from_numpy(pose_spherical(30.0 + 7.0 * rank, -30.0, 4.0))
    ro, rd = nk.get_rays(H, W, FOCAL, c2w)
    ro_flat, rd_flat = ro.reshape(-1, 3), rd.reshape(-1, 3)
    POOL = 64                                   # distinct ray batches, cycled (each 1024 of 640 000 rays)
    R = args.repeats
    rs = np.random.RandomState(1234 + rank)
    idx = torch.from_numpy(np.stack([rs.choice(H * W, n_rand, replace=False) for _ in range(POOL)])).to(dev)
    tgt_pool = torch.rand(POOL, n_rand, 3, generator=torch.Generator().manual_seed(1 + rank)).to(dev)
    # resident arm: packed rays of every pooled batch already in HBM (no gather inside the timed region)
    rays_pool = torch.stack([ops.make_ray_batch(ro_flat[idx[i]], rd_flat[idx[i]], 2.0, 6.0, use_viewdirs=True) for i in range(POOL)])
    # e2e arm: one pinned host block of 9 * N_rand floats per step ([o | d | target]) -> ONE host->device copy; the
    # step's loss goes back through an asynchronous device->host copy into its own pinned slot and is READ one step
    # later, after the next step has been enqueued, so the GPU never waits for the host (every loss is read inside
    # the timed region).
    h_in = [torch.cat([ro_flat[idx[i]].reshape(-1), rd_flat[idx[i]].reshape(-1), tgt_pool[i].reshape(-1)]).cpu().pin_memory()
            for i in range(POOL)]
    h_loss = torch.zeros((POOL, 2), dtype=torch.float32).pin_memory()
    ev = [torch.cuda.Event() for _ in range(POOL)]
    e2e_seen = []

    def resident_step(i):
        return step.step(rays_pool[i % POOL], tgt_pool[i % POOL])

    def e2e_step(i):
        j = i % POOL
        blk = h_in[j].to(dev, non_blocking=True)
        m = 3 * n_rand
        rays = ops.make_ray_batch(blk[0:m].view(n_rand, 3), blk[m:2 * m].view(n_rand, 3), 2.0, 6.0, use_viewdirs=True)
        loss = step.step(rays, blk[2 * m:3 * m].view(n_rand, 3))
        h_loss[j].copy_(loss, non_blocking=True)
        ev[j].record()
        if e2e_step.prev is not None:
            ev[e2e_step.prev].synchronize()
            e2e_seen.append(float(h_loss[e2e_step.prev, 0]))   # D2H result of the previous step, read on the host
        e2e_step.prev = j
        return loss
    e2e_step.prev = None

    def timed(fn):
        """R back-to-back repeats of EXACTLY K steps each (device-timed, max over ranks); the median repeat is reported,
        the clock sampler watches all of them."""
        c = 0
        for _ in range(Wm):
            fn(c)
            c += 1
        barrier()
        ops.launch_count_reset()
        replays0 = step.graph_replays
        sampler = ClockSampler(gpu_index(local))
        sampler.start()
        ts, last = [], None
        for _ in range(R):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(K):
                last = fn(c)
                c += 1
            if fn is e2e_step:   # the last step's loss is read on the host before the clock stops
                ev[e2e_step.prev].synchronize()
                e2e_seen.append(float(h_loss[e2e_step.prev, 0]))
                e2e_step.prev = None
            e1.record()
            barrier()
            ts.append(dist_max(e0.elapsed_time(e1) * 1e-3))
        clocks = sampler.stop()
        # kernels inside a replayed CUDA graph are counted once at capture: add them per replay
        launches = (ops.launch_count() + (step.launches_per_step or 0) * (step.graph_replays - replays0)) // R
        return float(np.median(ts)), ts, launches, clocks, last

    t_res, ts_res, launches, clocks, last_loss = timed(resident_step)
    t_e2e, ts_e2e, _, _, _ = timed(e2e_step)
    value = world * n_rand * K / t_res
    e2e = world * n_rand * K / t_e2e
    # ---- the other half of the metric in the same run: config 1 render at this N (row blocks, no collective)
    kw = render_kwargs_for(mc, mf, prec_name, ndc=False, near=2.0, far=6.0)
    render_obj = None
    if not args.no_render:
        try:
            frames = 8
            m = measure_render(400, 400, 555.555, blender_render_poses(40), kw, 32768, frames, 3, rank, world, barrier,
                               dist_max, clocks=False, local=local)
            render_obj = render_summary(m, frames, world, peaks()["bf16_tflops"])
            render_obj["workload"] = WORKLOADS["render"]
        except Exception as ex:  # reported, never fatal for the headline
            render_obj = {"failed": "%s: %s" % (type(ex).__name__, ex)}
    if rank == 0:
        pk = peaks()
        peak = pk["bf16_tflops"]
        tf, t_kernel, pts = forward_kernel_roofline(mf, prec_name)
        tk = train_kernel_rooflines(mf, prec_name, n_rand) if prec_name != "fp32" else None
        out = {
            "metric": METRIC + " (train step)", "value": value, "unit": "rays/s", "n_gpus": world,
            "steps": K, "warmup": Wm, "ms_per_step": 1e3 * t_res / K, "higher_is_better": True, "scaling": args.scaling,
            "vs_baseline": None, "dtype": "f16" if prec_name != "fp32" else "f32", "data": "synthetic",
            "config": shared_config("train", args, world),
            "impl_detail": {"mlp_precision": "fp16 operands (tcgen05 kind::f16), fp32 accumulation" if prec_name != "fp32" else "fp32",
                            "cuda_graph": bool(step.use_graph),
                            "graph_shape": ("forward, then the two networks' backward passes as parallel branches, Adam"
                                            if getattr(step, "_bwd_streams", False) else "one chain")
                            + ("; one all-reduce of the whole gradient bucket before Adam" if world > 1 else ""),
                            "timed_region": "median of %d back-to-back repeats of exactly "
                            "K steps" % R, "repeat_ms_per_step": [1e3 * t / K for t in ts_res]},
            "e2e": {"value": e2e, "unit": "rays/s", "h2d_bytes_per_step": n_rand * 9 * 4 * world,
                    "d2h_bytes_per_step": 8 * world, "ms_per_step": 1e3 * t_e2e / K},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": None,
            "cpu_baseline": None,
            "render": render_obj,
            "extra": {"loss_last": [float(x) for x in last_loss.tolist()]},
        }
        step_frac = value / world * FLOPS_PER_RAY_TRAIN / 1e12 / peak
        if tk is not None:
            # the step's own dominant launch: the backward (dgrad chain + wgrad) of the fine pass
            out["roofline"] = {"bound": "tensor", "kernel": "train step backward of the fine pass (dgrad chain + wgrad + reduce), %d points" % tk["points"],
                               "achieved": tk["backward_tflops"], "peak": peak, "unit": "TFLOP/s",
                               "frac": tk["backward_tflops"] / peak, "traffic": _profiled_traffic_backward(),
                               "peak_source": pk["source"], "ms_per_launch": tk["backward_ms"],
                               "step_frac_of_tensor_peak": step_frac}
            out["roofline_kernels"] = {
                "train_forward_with_records": {"ms": tk["forward_ms"], "tflops": tk["forward_tflops"], "frac": tk["forward_tflops"] / peak},
                "train_backward": {"ms": tk["backward_ms"], "tflops": tk["backward_tflops"], "frac": tk["backward_tflops"] / peak},
                "render_forward_32768x192": {"ms": 1e3 * t_kernel, "tflops": tf, "frac": tf / peak,
                                             "traffic": _profiled_traffic("r2_mlp_tc_kernel_ncu_full.csv")}}
        else:
            out["roofline"] = {"bound": "tensor", "kernel": "fused MLP forward (%s), %d points" % (prec_name, pts),
                               "achieved": tf, "peak": peak, "unit": "TFLOP/s", "frac": tf / peak, "traffic": None,
                               "peak_source": pk["source"], "ms_per_launch": 1e3 * t_kernel,
                               "step_frac_of_tensor_peak": step_frac}
        if cpu is not None:
            out["cpu_baseline"] = {"value": cpu["rate"], "unit": "rays/s", "cores": cpu["threads"], "kind": "reference",
                                   "sample": cpu["sample"] + ", 2 steps, timed before the GPU work starts"}
        else:
            out["cpu_baseline"] = {"value": None, "unit": "rays/s", "cores": threads, "kind": "reference",
                                   "sample": "measured at N=1 only (rank 0 alone on the host)"}
        print(json.dumps(out), flush=True)
    step.close()
    finish()


def run_render(args, mode):
    import nerf_kinematics_b200 as nk  # noqa: F401
    from nerf_kinematics_b200.poses import blender_render_poses
    world_env = int(os.environ.get("WORLD_SIZE", "1"))
    hw = 400 if mode == "render" else 800
    focal = 555.555 * hw / 400.0
    cpu = None
    if world_env == 1 and not args.no_cpu_baseline:
        cpu = cpu_reference_render(hw, hw, focal, 1, 1, budget_s=20.0)
    rank, local, world, dev, barrier, dist_max, finish = setup_dist(args)
    mc, mf = lego_models(dev)
    kw = render_kwargs_for(mc, mf, args.precision, ndc=False, near=2.0, far=6.0)
    K, Wm = args.steps, max(3, args.warmup)
    m = measure_render(hw, hw, focal, blender_render_poses(40), kw, 32768, K, Wm, rank, world, barrier, dist_max, local=local)
    if rank == 0:
        pk = peaks()
        peak = pk["bf16_tflops"]
        s = render_summary(m, K, world, peak)
        tf, t_kernel, pts = forward_kernel_roofline(mf, args.precision)
        out = {"metric": METRIC + " (render)", "value": s["value"], "unit": "rays/s", "n_gpus": world, "steps": K,
               "warmup": Wm, "ms_per_step": s["ms_per_frame"], "higher_is_better": True, "scaling": "strong",
               "vs_baseline": None, "dtype": "f16" if args.precision != "fp32" else "f32", "data": "synthetic",
               "config": shared_config(mode, args, world),
               "impl_detail": {"partition": "contiguous row blocks per rank, no collective", "cuda_graph": "one per frame",
                               "frame_roofline": s["roofline"]},
               "e2e": s["e2e"], "gpu_launches": s["gpu_launches"], "clocks": m["clocks"],
               "roofline": {"bound": "tensor", "kernel": "fused MLP forward (%s), %d points" % (args.precision, pts),
                            "achieved": tf, "peak": peak, "unit": "TFLOP/s", "frac": tf / peak,
                            "traffic": _profiled_traffic("r2_mlp_tc_kernel_ncu_full.csv"), "peak_source": pk["source"],
                            "ms_per_launch": 1e3 * t_kernel, "frame_frac_of_tensor_peak": s["roofline"]["frac"]},
               "cpu_baseline": ({"value": cpu["rate"], "unit": "rays/s", "cores": cpu["threads"], "kind": "reference",
                                 "sample": cpu["sample"]} if cpu else
                                {"value": None, "unit": "rays/s", "cores": os.cpu_count(), "kind": "reference",
                                 "sample": "measured at N=1 only"}),
               "extra": {"finite": m["finite"], "frame_mean": m["frame_mean"]}}
        print(json.dumps(out), flush=True)
    finish()


def run_wheel(args):
    """configs[2]: poses.txt -> (unmodified) parser_instant_ngp.py -> transforms_test_video.json (committed fixture of the
    parser's output) -> load_transforms -> render_path with the 3x3 K (fx != fy), NDC, at 1280x720."""
    import nerf_kinematics_b200 as nk
    from nerf_kinematics_b200.poses import load_transforms
    rank, local, world, dev, barrier, dist_max, finish = setup_dist(args)
    tr = load_transforms(os.path.join(ROOT, "tests", "golden", "wheel_transforms_test_video.json"))
    H, W, Kmat, poses = tr["H"], tr["W"], tr["K"], tr["poses"]
    ck = np.load(os.path.join(ROOT, "tests", "golden", "ckpt_wheel_robot6.npz"))
    res = {}
    K, Wm = min(args.steps, len(poses)), 2
    for name in ("A_8x256_random", "B_4x128_wheel_robot6"):
        if name.startswith("A"):
            mc, mf = lego_models(dev, lx=6)
        else:
            mc = nk.model_from_state_dict({k[2:]: torch.from_numpy(ck[k]) for k in ck.files if k.startswith("c.")}).to(dev)
            mf = nk.model_from_state_dict({k[2:]: torch.from_numpy(ck[k]) for k in ck.files if k.startswith("f.")}).to(dev)
        kw = render_kwargs_for(mc, mf, args.precision, ndc=True, near=0.0, far=1.0, white_bkgd=False)
        m = measure_render(H, W, Kmat, list(poses), kw, 32768, K, Wm, rank, world, barrier, dist_max, clocks=False, local=local)
        fl = mc.flops_per_point * PTS_PER_RAY
        rate = m["rays_per_frame"] * K / m["t_res"]
        res[name] = {"value": rate, "e2e": m["rays_per_frame"] * K / m["t_e2e"], "ms_per_frame": 1e3 * m["t_res"] / K,
                     "frac_of_tensor_peak": rate / world * fl / 1e12 / peaks()["bf16_tflops"], "finite": m["finite"],
                     "flops_per_ray": fl}
    if rank == 0:
        a = res["A_8x256_random"]
        print(json.dumps({"metric": METRIC + " (render)", "value": a["value"], "unit": "rays/s", "n_gpus": world, "steps": K,
                          "warmup": Wm, "ms_per_step": a["ms_per_frame"], "higher_is_better": True, "scaling": "strong",
                          "vs_baseline": None, "dtype": "f16", "data": "synthetic weights / real robot poses",
                          "config": dict(shared_config("wheel", args, world), H=H, W=W, fx=float(Kmat[0, 0]), fy=float(Kmat[1, 1]),
                                         poses=len(poses)),
                          "e2e": {"value": a["e2e"], "unit": "rays/s", "h2d_bytes_per_step": 48 * world, "d2h_bytes_per_step": H * W * 16},
                          "models": res}), flush=True)
    finish()


def run_fern(args):
    """configs[3]: fern 504x378, NDC + viewdirs, Lx = 10, 8x256; 120-pose spiral render and an N_rand = 4096 train step with
    raw_noise_std = 1 (paper_configs/llff_config.txt:18-20)."""
    import nerf_kinematics_b200 as nk
    from nerf_kinematics_b200 import ops
    from nerf_kinematics_b200.poses import render_path_spiral
    from nerf_kinematics_b200.trainer import TrainStep
    rank, local, world, dev, barrier, dist_max, finish = setup_dist(args)
    H, W, focal = 378, 504, 407.57
    c2w = np.eye(4, dtype=np.float32)[:3]
    poses = render_path_spiral(c2w, np.array([0.0, 1.0, 0.0]), [0.3, 0.3, 0.1], 3.0, 0.2, 0.5, 2, 120)
    mc, mf = lego_models(dev)
    kw = render_kwargs_for(mc, mf, args.precision, ndc=True, near=0.0, far=1.0, white_bkgd=False)
    K = min(args.steps, len(poses))
    m = measure_render(H, W, focal, poses, kw, 32768, K, 3, rank, world, barrier, dist_max, clocks=False, local=local)
    rate = m["rays_per_frame"] * K / m["t_res"]
    # train step: N_rand = 4096 rays of one pose, NDC, noise 1.0, perturb on
    step = TrainStep(mc, mf, N_samples=N_SAMPLES, N_importance=N_IMPORTANCE, perturb=1.0, raw_noise_std=1.0, white_bkgd=False,
                     precision=args.precision, lrate=5e-4, lrate_decay=250, world_size=world, cuda_graph=True, seed=7 + rank)
    ro, rd = nk.get_rays(H, W, focal, torch.from_numpy(np.asarray(poses[0])))
    n_rand = 4096
    rs = np.random.RandomState(rank)
    batches = []
    for _ in range(8):
        sel = torch.from_numpy(rs.choice(H * W, n_rand, replace=False)).to(dev)
        batches.append(ops.make_ray_batch(ro.reshape(-1, 3)[sel], rd.reshape(-1, 3)[sel], 0.0, 1.0, use_viewdirs=True, ndc=True,
                                          H=H, W=W, focal=focal))
    tgt = torch.rand(n_rand, 3, device=dev)
    for i in range(3):
        step.step(batches[i % 8], tgt)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    Kt = 20
    for i in range(Kt):
        loss = step.step(batches[i % 8], tgt)
    e1.record()
    barrier()
    t_tr = dist_max(e0.elapsed_time(e1) * 1e-3)
    if rank == 0:
        peak = peaks()["bf16_tflops"]
        tr_rate = world * n_rand * Kt / t_tr
        print(json.dumps({"metric": METRIC + " (render)", "value": rate, "unit": "rays/s", "n_gpus": world, "steps": K, "warmup": 3,
                          "ms_per_step": 1e3 * m["t_res"] / K, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                          "dtype": "f16", "data": "synthetic", "config": dict(shared_config("fern", args, world), H=H, W=W, focal=focal),
                          "e2e": {"value": m["rays_per_frame"] * K / m["t_e2e"], "unit": "rays/s", "h2d_bytes_per_step": 48 * world,
                                  "d2h_bytes_per_step": H * W * 16},
                          "render_frac_of_tensor_peak": rate / world * FLOPS_PER_RAY_RENDER / 1e12 / peak, "finite": m["finite"],
                          "train": {"value": tr_rate, "unit": "rays/s", "n_rand": n_rand, "raw_noise_std": 1.0, "ms_per_step": 1e3 * t_tr / Kt,
                                    "frac_of_tensor_peak": tr_rate / world * FLOPS_PER_RAY_TRAIN / 1e12 / peak,
                                    "loss": [float(x) for x in loss.tolist()]}}), flush=True)
    finish()


def run_sweep(args):
    """configs[4]: N in 2^12 .. 2^22 rays (o ~ N(0,1)*0.5 + [0,0,4], d unit * U[0.8,1.2], near 2, far 6, seed 0), model (A):
    end-to-end render through the captured chunk chain, plus the stage kernels alone at each N.  At world > 1 every rank
    takes a contiguous block of N / world rays (no collective)."""
    from nerf_kinematics_b200 import ops
    from nerf_kinematics_b200.frame import FrameRenderer
    rank, local, world, dev, barrier, dist_max, finish = setup_dist(args)
    mc, mf = lego_models(dev)
    pk = peaks()
    rows = []
    g = torch.Generator(device="cpu").manual_seed(0)
    for e in range(12, 23, 2 if not args.full_sweep else 1):
        N = 1 << e
        n = N // world
        o = (torch.randn(N, 3, generator=g) * 0.5 + torch.tensor([0, 0, 4.0]))[rank * n:(rank + 1) * n].to(dev)
        d = torch.nn.functional.normalize(torch.randn(N, 3, generator=g), dim=-1) * (0.8 + 0.4 * torch.rand(N, 1, generator=g))
        d = d[rank * n:(rank + 1) * n].to(dev)
        packed = ops.make_ray_batch(o, d, 2.0, 6.0, use_viewdirs=True)
        fr = FrameRenderer(1, n, 1.0, mc, mf, chunk=32768, near=2.0, far=6.0, ndc=False, use_viewdirs=True, N_samples=N_SAMPLES,
                           N_importance=N_IMPORTANCE, white_bkgd=True, precision=args.precision, external_rays=True)
        reps = max(2, min(50, (1 << 21) // N))
        with torch.no_grad():
            for _ in range(2):
                fr.render_packed(packed)
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                fr.render_packed(packed)
            e1.record()
            barrier()
        t = dist_max(e0.elapsed_time(e1) * 1e-3) / reps
        row = {"rays": N, "rays_per_gpu": n, "ms": 1e3 * t, "rays_per_s": N / t,
               "frac_of_tensor_peak": N / t / world * FLOPS_PER_RAY_RENDER / 1e12 / pk["bf16_tflops"]}
        if rank == 0:
            # stage kernels alone on this rank's block (HBM rooflines, SURVEY 8d bytes per ray)
            m = min(n, 1 << 20)
            rays = packed[:m]
            zc = ops.sample_coarse(rays, N_SAMPLES)
            rawc = torch.randn(m, N_SAMPLES, 4, device=dev)
            rawf = torch.randn(m, N_SAMPLES + N_IMPORTANCE, 4, device=dev)
            _, _, _, wc, _ = ops.composite_forward(rawc, zc, rays, white_bkgd=True)
            zf, _, _ = ops.sample_fine(zc, wc, N_IMPORTANCE)
            kr = max(3, min(50, (1 << 22) // m))
            tc_ = _events_time(lambda: ops.composite_forward(rawc, zc, rays, white_bkgd=True), kr)
            tf_ = _events_time(lambda: ops.composite_forward(rawf, zf, rays, white_bkgd=True, want_weights=False), kr)
            ts_ = _events_time(lambda: ops.sample_fine(zc, wc, N_IMPORTANCE), kr)
            h, prec = mf.prepare(args.precision)
            tfl, tk, pts = forward_kernel_roofline(mf, args.precision, n_rays=m, reps=3) if m * 192 * 16 < 8e9 else (None, None, None)
            row["kernels"] = {"rays": m,
                              "composite_coarse_frac_hbm": m * 1572 / tc_ / 1e9 / pk["hbm_gbs"],
                              "composite_fine_frac_hbm": m * 3876 / tf_ / 1e9 / pk["hbm_gbs"],
                              "sample_fine_frac_hbm": m * 1272 / ts_ / 1e9 / pk["hbm_gbs"],
                              "mlp_fine_frac_tensor": (tfl / pk["bf16_tflops"]) if tfl else None}
            del rawc, rawf
        rows.append(row)
        del fr, packed
        torch.cuda.empty_cache()
    if rank == 0:
        best = max(r["rays_per_s"] for r in rows)
        print(json.dumps({"metric": METRIC + " (render)", "value": best, "unit": "rays/s", "n_gpus": world, "steps": len(rows),
                          "warmup": 2, "ms_per_step": None, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                          "dtype": "f16", "data": "synthetic", "config": shared_config("sweep", args, world), "sweep": rows}), flush=True)
    finish()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--mode", default="train", choices=sorted(WORKLOADS))
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"], help="train: rays per GPU fixed (weak) or one batch split (strong)")
    ap.add_argument("--n-rand", type=int, default=1024)
    ap.add_argument("--repeats", type=int, default=5, help="back-to-back repeats of the K-step timed region (median reported)")
    ap.add_argument("--precision", default="fp16", choices=["fp16", "bf16", "fp32"])
    ap.add_argument("--no-graph", action="store_true", help="launch the train step kernel by kernel (no CUDA graph)")
    ap.add_argument("--no-render", action="store_true", help="train mode: skip the config-1 render object")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--full-sweep", action="store_true", help="sweep: every power of two instead of every other")
    args = ap.parse_args()
    if args.precision == "bf16":
        args.precision = "fp16"   # historical alias: the tensor-core path computes in IEEE fp16
    if args.impl == "reference":
        run_reference(args)
    elif args.mode == "train":
        run_train(args)
    elif args.mode in ("render", "render800"):
        run_render(args, args.mode)
    elif args.mode == "wheel":
        run_wheel(args)
    elif args.mode == "fern":
        run_fern(args)
    else:
        run_sweep(args)


if __name__ == "__main__":
    main()
