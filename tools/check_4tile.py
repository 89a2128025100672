"""Four-tile forward kernel (mlp_tc4_kernel, graphs with every stage N <= 128) against the two-tile kernel on the same
inputs: must agree bit for bit (same MMAs in the same order per row), for tile counts that exercise full rounds,
partly filled last rounds and ragged tiles.  Then times both.  Run under `timeout`: a barrier protocol bug is a hang."""
import ctypes as C
import os
import sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nerf_kinematics_b200 as nk  # noqa: E402
from nerf_kinematics_b200 import ops  # noqa: E402
from nerf_kinematics_b200._lib import check, lib  # noqa: E402

dev = torch.device("cuda")
torch.manual_seed(0)
models = {"4x128_L6": nk.FlexibleNeRFModel(num_encoding_fn_xyz=6).to(dev),
          "4x128_L10": nk.FlexibleNeRFModel(num_encoding_fn_xyz=10).to(dev),
          "4x128_noview": nk.FlexibleNeRFModel(num_encoding_fn_xyz=6, use_viewdirs=False).to(dev)}


def fwd(m, rays, z, four, sync=True):
    lib().nb_set_option(b"fwd_4tile", int(four))
    n, S = z.shape
    raw = torch.full((n, S, 4), float("nan"), device=dev)
    h, prec = m.prepare("fp16")
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    check(lib().nb_mlp_forward(h, prec, C.c_void_p(rays.data_ptr()), rays.shape[1], C.c_void_p(z.data_ptr()), n, S,
                               C.c_void_p(raw.data_ptr()), C.c_void_p(0), 0, 0, st))
    if sync:
        torch.cuda.synchronize()
    return raw


def fwd32(m, rays, z):
    pts = rays[:, None, 0:3] + rays[:, None, 3:6] * z[..., None]
    with torch.no_grad():
        return m.query(pts, rays[:, 8:11] if m.use_viewdirs else None, precision="fp32")


def batch(m, n, S):
    o = torch.randn(n, 3, device=dev) * 0.5 + torch.tensor([0, 0, 4.0], device=dev)
    d = torch.nn.functional.normalize(torch.randn(n, 3, device=dev), dim=-1)
    rays = ops.make_ray_batch(o, d, 2.0, 6.0, use_viewdirs=bool(m.use_viewdirs))
    return rays, ops.sample_coarse(rays, S)


ok = True
for name, m in models.items():
    for n, S in ((1, 64), (2, 64), (3, 64), (5, 64), (7, 192), (9, 64), (64, 64), (148 * 2 + 1, 64), (148 * 4 + 3, 64),
                 (1024, 192), (4099, 64), (20011, 192)):
        rays, z = batch(m, n, S)
        a, b, b2 = fwd(m, rays, z, 0), fwd(m, rays, z, 1), fwd(m, rays, z, 1)
        same = torch.equal(a, b) and bool(torch.isfinite(b).all())
        ref = fwd32(m, rays, z)
        ea, eb = float((a - ref).abs().mean()), float((b - ref).abs().mean())
        close = bool(torch.isfinite(b).all()) and float((a - b).abs().max()) <= 4e-4 * max(1.0, float(ref.abs().max())) \
            and abs(ea - eb) <= 0.05 * max(ea, 1e-7)
        ok &= same
        print("%-13s rays %5d x %3d (%5d tiles): %s  max|diff| %.3e  frac differing %.4f | err vs fp32: 2-tile %.3e 4-tile %.3e"
              " | run-to-run %s" % (name, n, S, (n * S + 127) // 128, "equal" if same else ("close" if close else "DIFFERENT"),
                                    float((a - b).abs().max()), float((a != b).any(-1).float().mean()), ea, eb,
                                    "deterministic" if torch.equal(b, b2) else "VARIES"), flush=True)
        if not same and n * S <= 4096:
            dif = (a != b).reshape(-1, 4)
            print("      channels differing:", dif.sum(0).tolist(), flush=True)
        if not same and n * S <= 4096:
            rows = torch.nonzero((a != b).reshape(-1, 4).any(-1)).flatten()
            print("      differing points (index, index %% 128):", [(int(r), int(r) % 128) for r in rows[:24]], flush=True)
m = models["4x128_L6"]
flops = 2 * sum(p.numel() for n_, p in m.named_parameters() if n_.endswith("weight"))
for n, S in ((32768, 192), (65536, 64)):
    rays, z = batch(m, n, S)
    for four in (0, 1, 0, 1):
        fwd(m, rays, z, four)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            fwd(m, rays, z, four, sync=False)
        e1.record()
        torch.cuda.synchronize()
        t = e0.elapsed_time(e1) / 5
        print("points %d  4tile=%d  %.3f ms  %.1f M points/s  %.1f TFLOP/s (model flops)" % (
            n * S, four, t, n * S / t / 1e3, n * S * flops / t / 1e9), flush=True)
lib().nb_set_option(b"fwd_4tile", 1)
print("ALL EQUAL" if ok else "MISMATCH")
sys.exit(0 if ok else 1)
