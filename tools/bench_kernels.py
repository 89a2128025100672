"""Stand-alone kernel timings (CUDA events on the launching stream) for the roofline table in DESIGN.md.

    python tools/bench_kernels.py [--rays 32768] [--reps 10]

Prints one line per kernel: achieved TFLOP/s (MLP) or GB/s (HBM-bound stages) and the fraction of the
measured peak in MEASURED_PEAKS.json."""
import argparse
import ctypes as C
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import nerf_kinematics_b200 as nk  # noqa: E402
from nerf_kinematics_b200 import ops  # noqa: E402
from nerf_kinematics_b200._lib import check, lib  # noqa: E402


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "fallback": True}


def timeit(fn, reps, flush=None, inner=1):
    """Median / best seconds per call.  inner > 1: that many back-to-back calls between the two events, so the CPU-side
    launch path (output allocation, ctypes) of a short kernel does not show up as GPU idle time inside the timed
    region; only for working sets larger than the 126 MB L2 (no flush between the inner calls)."""
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        if flush is not None:
            flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(inner):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e-3 / inner)
    ts.sort()
    return ts[len(ts) // 2], ts[0]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rays", type=int, default=32768)
    ap.add_argument("--reps", type=int, default=10)
    ap.add_argument("--model", default="tf", choices=["tf", "flexible"])
    ap.add_argument("--only-mlp", action="store_true", help="just the 16-bit fused MLP kernel (short run for ncu)")
    ap.add_argument("--only-stages", action="store_true", help="skip the MLP kernels")
    a = ap.parse_args()
    pk = peaks()
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    n, S, I = a.rays, 64, 128
    model = (nk.NeRF() if a.model == "tf" else nk.FlexibleNeRFModel(num_encoding_fn_xyz=10)).to(dev)
    o = torch.randn(n, 3, device=dev) * 0.5 + torch.tensor([0, 0, 4.0], device=dev)
    d = torch.nn.functional.normalize(torch.randn(n, 3, device=dev), dim=-1) * (0.8 + 0.4 * torch.rand(n, 1, device=dev))
    rays = ops.make_ray_batch(o, d, 2.0, 6.0, use_viewdirs=True)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    res = {}
    for S_ in (() if a.only_stages else (S + I,) if a.only_mlp else (S, S + I)):
        z = ops.sample_coarse(rays, S_)
        raw = torch.empty(n, S_, 4, device=dev)
        for prec_name in (("bf16",) if a.only_mlp else ("bf16", "fp32")):
            h, prec = model.prepare(prec_name)
            P = n * S_
            if prec_name == "fp32" and P > (1 << 21):
                continue
            nbytes = lib().nb_mlp_workspace_bytes(h, prec, P, 0)
            ws = torch.empty(max(nbytes, 1), dtype=torch.uint8, device=dev)
            st = C.c_void_p(torch.cuda.current_stream().cuda_stream)

            def run():
                check(lib().nb_mlp_forward(h, prec, C.c_void_p(rays.data_ptr()), 11, C.c_void_p(z.data_ptr()), n, S_,
                                           C.c_void_p(raw.data_ptr()), C.c_void_p(ws.data_ptr()), nbytes, 0, st))
            med, best = timeit(run, a.reps, flush)
            tf = P * model.flops_per_point / med / 1e12
            print("mlp_forward %-4s %s S=%3d  P=%9d  %8.3f ms  %8.1f TFLOP/s  frac(of measured bf16 %.0f)=%.3f  [best %.3f ms]"
                  % (prec_name, a.model, S_, P, med * 1e3, tf, pk["bf16_tflops"], tf / pk["bf16_tflops"], best * 1e3))
            res["mlp_%s_%d" % (prec_name, S_)] = tf
    if a.only_mlp:
        return
    # HBM-bound stages (use >= 1M rays so the kernels run long enough to measure bandwidth)
    n = max(n, 1 << 20)
    o = torch.randn(n, 3, device=dev) * 0.5 + torch.tensor([0, 0, 4.0], device=dev)
    d = torch.nn.functional.normalize(torch.randn(n, 3, device=dev), dim=-1)
    rays = ops.make_ray_batch(o, d, 2.0, 6.0, use_viewdirs=True)
    for S_, tag in ((S, "coarse"), (S + I, "fine")):
        z = ops.sample_coarse(rays, S_)
        raw = torch.randn(n, S_, 4, device=dev)
        med, _ = timeit(lambda: ops.composite_forward(raw, z, rays, None, True, want_weights=(tag == "coarse")), a.reps, flush, inner=4)
        bytes_ray = (S_ * 16 + S_ * 4 + 12) + (S_ * 4 if tag == "coarse" else 0) + 24
        gbs = n * bytes_ray / med / 1e9
        print("composite_fwd %-6s %8.3f ms  %8.1f GB/s  frac(of measured %.0f)=%.3f" % (tag, med * 1e3, gbs, pk["hbm_gbs"], gbs / pk["hbm_gbs"]))
        if True:
            g = torch.randn(n, 3, device=dev)
            med, _ = timeit(lambda: ops.composite_backward(raw, z, rays, None, True, g), a.reps, flush, inner=4)
            gbs = n * (S_ * 20 + 24 + S_ * 16) / med / 1e9
            print("composite_bwd %-6s %8.3f ms  %8.1f GB/s  frac=%.3f" % (tag, med * 1e3, gbs, gbs / pk["hbm_gbs"]))
    z = ops.sample_coarse(rays, S)
    w = torch.rand(n, S, device=dev) ** 4
    med, _ = timeit(lambda: ops.sample_fine(z, w, I), a.reps, flush, inner=4)
    gbs = n * 1272 / med / 1e9
    print("sample_fine (pdf+merge) %8.3f ms  %8.1f GB/s  frac=%.3f" % (med * 1e3, gbs, gbs / pk["hbm_gbs"]))
    c2w = torch.eye(4)
    med, _ = timeit(lambda: ops.get_rays(2160, 3840, 1000.0, c2w), a.reps, flush, inner=8)
    gbs = 2160 * 3840 * 24 / med / 1e9
    print("get_rays 3840x2160 %8.3f ms  %8.1f GB/s  frac=%.3f" % (med * 1e3, gbs, gbs / pk["hbm_gbs"]))


if __name__ == "__main__":
    main()
