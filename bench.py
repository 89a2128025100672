"""bench.py -- the headline metric of BASELINE.json on B200: rays/sec at 64+128 samples per ray.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--precision auto|bf16|fp32]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N --steps K --warmup W

Workload (N=1 and every N, weak scaling): BASELINE.json configs[1] -- lego full-res 800x800 training step,
N_rand = 1024 rays per GPU per step drawn without replacement from one 800x800 pose, 8x256 coarse + fine MLP
(the init_nerf_model graph), 64 stratified + 128 importance samples, forward + loss + backward +
[gradient all-reduce] + Adam step.  Synthetic rays/targets, random-init weights.

`value`  : rays/s over all ranks with the step's inputs already resident in HBM.
`e2e`    : the same step fed from pinned HOST buffers (H2D of ray origins/directions/targets and D2H of the
           loss inside the timed region) through the public TrainStep API.
`roofline`: the dominant kernel (the fused MLP kernel of the precision in use), timed alone with CUDA events on
           its launch stream, against the measured bf16 tensor peak of MEASURED_PEAKS.json.
`cpu_baseline`: the CPU oracle port (oracle/nerf_oracle.py, a restatement pinned to the reference) timed on this
           box's host cores on a bounded sample of the same step.
`--impl reference`: the reference's CPU algorithm (the oracle port; the reference itself is PyTorch/TF Python
           that cannot travel to the GPU box) on all host threads, same metric/config, bounded sample per step.
"""
import argparse
import ctypes as C
import json
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

H = W = 800
FOCAL = 1111.111
N_RAND = 1024
N_SAMPLES, N_IMPORTANCE = 64, 128
FLOPS_PER_POINT = 1186816            # 8x256 graph, 2*MAC, unpadded
FLOPS_PER_RAY_TRAIN = 3 * FLOPS_PER_POINT * (N_SAMPLES + N_SAMPLES + N_IMPORTANCE)  # fwd + dgrad + wgrad
WORKLOAD = "lego 800x800 train step, N_rand=1024 rays/GPU, 8x256 coarse+fine, 64+128 samples (BASELINE configs[1])"


def peaks():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        p["source"] = "measured"
        return p
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


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
                                          "-lms", "100", "-i", self.idx], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = [float(r[1]) for r in self.rows if len(r) >= 8 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 8 and r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 8 for i in range(4) if r[4 + i].lower() == "active"})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ------------------------------------------------------------------------------------------ reference / CPU arm
def cpu_train_step_rate(n_rays, reps, threads):
    """rays/s of the CPU oracle port on one bounded train step (8x256 coarse+fine, 64+128 samples, fwd+bwd)."""
    from oracle import nerf_oracle as O
    torch.set_num_threads(threads)
    sdc, sdf = O.init_tf8x256_state(seed=0), O.init_tf8x256_state(seed=1)
    c2w = O.pose_spherical(30.0, -30.0, 4.0)
    ro, rd = O.pixel_rays(H, W, FOCAL, c2w)
    sel = np.random.RandomState(0).choice(H * W, n_rays, replace=False)
    rays = O.pack_rays(ro.reshape(-1, 3)[sel], rd.reshape(-1, 3)[sel], 2.0, 6.0, ndc=False, use_viewdirs=True)
    tgt = torch.rand(n_rays, 3, generator=torch.Generator().manual_seed(1))
    g = torch.Generator().manual_seed(2)
    times = []
    for i in range(reps + 1):
        t_rand = torch.rand(n_rays, N_SAMPLES, generator=g)
        u_fine = torch.rand(n_rays, N_IMPORTANCE, generator=g)
        t0 = time.perf_counter()
        O.train_step_grads("tf", sdc, sdf, rays, tgt, n_coarse=N_SAMPLES, n_fine=N_IMPORTANCE, n_freq_xyz=10,
                           n_freq_dir=4, white_bkgd=True, t_rand=t_rand, u_fine=u_fine)
        if i > 0:
            times.append(time.perf_counter() - t0)
    return n_rays / float(np.median(times)), times


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    n = 64
    from oracle import nerf_oracle as O  # noqa: F401
    torch.set_num_threads(threads)
    for _ in range(max(0, args.warmup - 1)):
        cpu_train_step_rate(n, 0, threads)
    t0 = time.perf_counter()
    rate, times = cpu_train_step_rate(n, max(1, args.steps), threads)
    ms = 1e3 * float(np.mean(times))
    out = {
        "impl": "reference", "metric": "rays/sec at 64+128 samples/ray (train step)", "value": rate, "unit": "rays/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "sample": "%d rays per step of the 1024-ray batch" % n},
        "cpu_baseline": {"value": rate, "unit": "rays/s", "cores": threads, "kind": "port",
                         "sample": "%d-ray train step (fwd+bwd, no optimizer), oracle/nerf_oracle.py on %d threads" % (n, threads)},
        "e2e": {"value": rate, "unit": "rays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(out), flush=True)


# ------------------------------------------------------------------------------------------ our arm
def profiled_traffic_bytes():
    """dram__bytes_read.sum + dram__bytes_write.sum of the roofline kernel's launch (same 32 768-ray x 192-sample
    workload) from the committed `ncu --set full` capture under profiles/; None if the summary is missing."""
    try:
        tot = 0.0
        for line in open(os.path.join(ROOT, "profiles", "r1c_mlp_tc_pair_kernel_ncu_full.csv")):
            parts = line.strip().split(",")
            if len(parts) == 3 and parts[0] in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                tot += float(parts[2]) * {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1.0}[parts[1]]
        return tot or None
    except Exception:
        return None


def kernel_roofline(model, prec_name, n_rays, reps):
    """The dominant kernel alone: fused MLP forward over the fine pass of one step (n_rays x 192 points)."""
    import nerf_kinematics_b200 as nk  # noqa: F401
    from nerf_kinematics_b200 import ops
    from nerf_kinematics_b200._lib import check, lib
    dev = model._flat.device
    S = N_SAMPLES + N_IMPORTANCE
    n = max(n_rays, 32768)  # enough tiles to fill 148 SMs for several waves
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
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        run()
    e1.record()
    torch.cuda.synchronize()
    t = e0.elapsed_time(e1) * 1e-3 / reps
    return n * S * FLOPS_PER_POINT / t / 1e12, t, n * S


def render_rate(models, prec_name, reps=3):
    """Extra: config 1 (lego 400x400 render, chunk 32768) through render(); rays/s, device-timed."""
    import nerf_kinematics_b200 as nk
    from nerf_kinematics_b200.poses import pose_spherical
    mc, mf = models
    kw = dict(network_query_fn=None, perturb=0.0, N_importance=N_IMPORTANCE, network_fine=mf, N_samples=N_SAMPLES,
              network_fn=mc, use_viewdirs=True, white_bkgd=True, raw_noise_std=0.0, ndc=False, lindisp=False,
              near=2.0, far=6.0, precision=prec_name)
    c2w = torch.from_numpy(pose_spherical(30.0, -30.0, 4.0))[:3, :4]
    with torch.no_grad():
        nk.render(400, 400, 555.555, chunk=32768, c2w=c2w, **kw)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            nk.render(400, 400, 555.555, chunk=32768, c2w=c2w, **kw)
        e1.record()
        torch.cuda.synchronize()
    return 160000 * reps / (e0.elapsed_time(e1) * 1e-3)


def run_ours(args):
    import torch.distributed as dist
    import nerf_kinematics_b200 as nk
    from nerf_kinematics_b200 import ops, parallel
    from nerf_kinematics_b200.poses import pose_spherical
    from nerf_kinematics_b200.trainer import TrainStep

    rank, local, world = parallel.init_distributed("nccl")
    assert world == args.gpus or world == 1, "launch with torchrun --nproc-per-node %d" % args.gpus
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    K, Wm = args.steps, max(3, args.warmup)
    torch.manual_seed(0)  # identical replicas on every rank
    mc, mf = nk.NeRF().to(dev), nk.NeRF().to(dev)
    prec_name = args.precision
    if prec_name == "auto":
        prec_name = "bf16" if getattr(nk, "BF16_BACKWARD", False) else "fp32"
    step = TrainStep(mc, mf, N_samples=N_SAMPLES, N_importance=N_IMPORTANCE, perturb=1.0, raw_noise_std=0.0,
                     white_bkgd=True, precision=prec_name, lrate=5e-4, lrate_decay=500, world_size=world,
                     cuda_graph=not args.no_graph, seed=1234 + rank)  # jitter + inverse-CDF uniforms drawn in-kernel (f4)
    # ---- synthetic inputs of the named shape: one 800x800 pose, N_rand rays per step without replacement
    c2w = torch.from_numpy(pose_spherical(30.0 + 7.0 * rank, -30.0, 4.0))
    ro, rd = nk.get_rays(H, W, FOCAL, c2w)
    rays_all = ops.make_ray_batch(ro, rd, 2.0, 6.0, use_viewdirs=True)          # [640000, 11] resident in HBM
    total = Wm + K
    rs = np.random.RandomState(1234 + rank)
    idx = torch.from_numpy(np.stack([rs.choice(H * W, N_RAND, replace=False) for _ in range(total)])).to(dev)
    tgt_all = torch.rand(total, N_RAND, 3, generator=torch.Generator().manual_seed(1 + rank)).to(dev)
    ro_flat, rd_flat = ro.reshape(-1, 3), rd.reshape(-1, 3)
    # pinned host copies for the e2e arm
    h_ro = [ro_flat[idx[i]].cpu().pin_memory() for i in range(total)]
    h_rd = [rd_flat[idx[i]].cpu().pin_memory() for i in range(total)]
    h_tg = [tgt_all[i].cpu().pin_memory() for i in range(total)]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def resident_step(i):
        return step.step(rays_all[idx[i]], tgt_all[i])

    # e2e: one pinned host block of 9 * N_rand floats per step ([o | d | target], each contiguous) -> ONE host->device copy; the step's loss goes back
    # through an asynchronous device->host copy into its own pinned slot and is READ one step later, after the next
    # step has been enqueued, so the GPU never waits for the host (every loss is still read inside the timed region).
    h_in = [torch.cat([h_ro[i].reshape(-1), h_rd[i].reshape(-1), h_tg[i].reshape(-1)]).pin_memory() for i in range(total)]
    h_loss = torch.zeros((total, 2), dtype=torch.float32).pin_memory()
    ev = [torch.cuda.Event() for _ in range(total)]
    e2e_seen = []

    def e2e_step(i):
        blk = h_in[i].to(dev, non_blocking=True)
        m = 3 * N_RAND
        rays = ops.make_ray_batch(blk[0:m].view(N_RAND, 3), blk[m:2 * m].view(N_RAND, 3), 2.0, 6.0, use_viewdirs=True)
        loss = step.step(rays, blk[2 * m:3 * m].view(N_RAND, 3))
        h_loss[i].copy_(loss, non_blocking=True)
        ev[i].record()
        if i > 0:
            ev[i - 1].synchronize()
            e2e_seen.append(float(h_loss[i - 1, 0]))   # D2H result of the previous step, read on the host
        return loss

    def timed(fn):
        for i in range(Wm):
            fn(i)
        barrier()
        ops.launch_count_reset()
        replays0 = step.graph_replays
        sampler = ClockSampler(torch.cuda.current_device() if "CUDA_VISIBLE_DEVICES" not in os.environ else
                               os.environ["CUDA_VISIBLE_DEVICES"].split(",")[local])
        sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(Wm, Wm + K):
            last = fn(i)
        if fn is e2e_step:   # the last step's loss is read on the host before the clock stops
            ev[Wm + K - 1].synchronize()
            e2e_seen.append(float(h_loss[Wm + K - 1, 0]))
        e1.record()
        barrier()
        clocks = sampler.stop()
        t = torch.tensor([e0.elapsed_time(e1) * 1e-3], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        # kernels inside a replayed CUDA graph are counted once at capture: add them per replay
        launches = ops.launch_count() + (step.launches_per_step or 0) * (step.graph_replays - replays0)
        return float(t), launches, clocks, last

    t_res, launches, clocks, last_loss = timed(resident_step)
    t_e2e, _, _, _ = timed(e2e_step)
    value = world * N_RAND * K / t_res
    e2e = world * N_RAND * K / t_e2e
    if rank == 0:
        pk = peaks()
        tf, t_kernel, pts = kernel_roofline(mf, prec_name, N_RAND, 5)
        peak = pk["bf16_tflops"]
        try:
            render = render_rate((mc, mf), "bf16")
        except Exception as ex:  # reported, never fatal for the headline
            render = "failed: %s" % ex
        threads = os.cpu_count() or 1
        cpu_rate, cpu_times = cpu_train_step_rate(64, 2, threads)
        out = {
            "metric": "rays/sec at 64+128 samples/ray (train step)", "value": value, "unit": "rays/s", "n_gpus": world,
            "steps": K, "warmup": Wm, "ms_per_step": 1e3 * t_res / K, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16" if prec_name == "bf16" else "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "rays_per_step_per_gpu": N_RAND, "n_samples": N_SAMPLES,
                       "n_importance": N_IMPORTANCE, "mlp_precision": prec_name, "optimizer": "fused Adam (in the timed region)",
                       "cuda_graph": bool(step.use_graph),
                       "l2": "per-step working set (saved activations of 262144 points) exceeds the 126 MB L2; "
                             "a different ray batch every step"},
            "e2e": {"value": e2e, "unit": "rays/s", "h2d_bytes_per_step": N_RAND * 9 * 4 * world,
                    "d2h_bytes_per_step": 8 * world, "ms_per_step": 1e3 * t_e2e / K},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"bound": "tensor", "kernel": "fused MLP forward (%s), %d points" % (prec_name, pts),
                         "achieved": tf, "peak": peak, "unit": "TFLOP/s", "frac": tf / peak,
                         "traffic": profiled_traffic_bytes() if prec_name == "bf16" else None,
                         "peak_source": pk["source"], "ms_per_launch": 1e3 * t_kernel,
                         "step_frac_of_tensor_peak": value / world * FLOPS_PER_RAY_TRAIN / 1e12 / peak},
            "cpu_baseline": {"value": cpu_rate, "unit": "rays/s", "cores": threads, "kind": "port",
                             "sample": "64-ray train step (fwd+bwd) x2, oracle/nerf_oracle.py"},
            "extra": {"render_rays_per_s_400x400_bf16": render, "loss_last": [float(x) for x in last_loss.tolist()]},
        }
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--precision", default="auto", choices=["auto", "bf16", "fp32"])
    ap.add_argument("--no-graph", action="store_true", help="launch the train step kernel by kernel (no CUDA graph)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
