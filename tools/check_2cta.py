"""CTA-pair forward kernel (mlp_tc2_kernel) against the single-CTA kernel on the same inputs: must agree bit for bit
(same MMAs in the same order per row), for tile counts that exercise full rounds, a lone stream and dummy tiles.
Then times both.  Run under `timeout`: a protocol bug in a barrier chain shows up as a hang."""
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
models = {"8x256": nk.NeRF().to(dev), "4x128": nk.FlexibleNeRFModel(num_encoding_fn_xyz=6).to(dev),
          "8x256_noview": nk.NeRF(use_viewdirs=False, input_ch_views=3).to(dev)}


def fwd(m, rays, z, two):
    lib().nb_set_option(b"fwd_2cta", int(two))
    n, S = z.shape
    raw = torch.full((n, S, 4), float("nan"), device=dev)
    h, prec = m.prepare("fp16")
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    check(lib().nb_mlp_forward(h, prec, C.c_void_p(rays.data_ptr()), rays.shape[1], C.c_void_p(z.data_ptr()), n, S,
                               C.c_void_p(raw.data_ptr()), C.c_void_p(0), 0, 0, st))
    torch.cuda.synchronize()
    return raw


ok = True
for name, m in models.items():
    for n, S in ((1, 64), (2, 64), (3, 64), (5, 64), (7, 192), (64, 64), (148 * 3 + 1, 64), (1024, 192), (4099, 64)):
        o = torch.randn(n, 3, device=dev) * 0.5 + torch.tensor([0, 0, 4.0], device=dev)
        d = torch.nn.functional.normalize(torch.randn(n, 3, device=dev), dim=-1)
        rays = ops.make_ray_batch(o, d, 2.0, 6.0, use_viewdirs=bool(m.use_viewdirs))
        z = ops.sample_coarse(rays, S)
        a, b, b2 = fwd(m, rays, z, 0), fwd(m, rays, z, 1), fwd(m, rays, z, 1)
        same = torch.equal(a, b)
        ok &= same
        print("%-13s rays %5d x %3d (%5d tiles): %s  max|diff| %.3e  | 2cta run-to-run %s  frac differing %.4f" % (
            name, n, S, (n * S + 127) // 128, "equal" if same else "DIFFERENT", float((a - b).abs().max()),
            "deterministic" if torch.equal(b, b2) else "VARIES", float((a != b).float().mean())), flush=True)
        if not same and n * S <= 2048:
            rows = torch.nonzero((a != b).reshape(-1, 4).any(-1)).flatten()
            print("      differing points (index, index %% 128):", [(int(r), int(r) % 128) for r in rows[:24]], flush=True)
m = models["8x256"]
n, S = 32768, 192
o = torch.randn(n, 3, device=dev) * 0.5 + torch.tensor([0, 0, 4.0], device=dev)
d = torch.nn.functional.normalize(torch.randn(n, 3, device=dev), dim=-1)
rays = ops.make_ray_batch(o, d, 2.0, 6.0, use_viewdirs=True)
z = ops.sample_coarse(rays, S)
for two in (0, 1, 0, 1):
    fwd(m, rays, z, two)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        fwd(m, rays, z, two)
    e1.record()
    torch.cuda.synchronize()
    t = e0.elapsed_time(e1) / 5
    print("2cta=%d  %.3f ms  %.1f TFLOP/s" % (two, t, n * S * 1186816 / t / 1e9), flush=True)
print("ALL EQUAL" if ok else "MISMATCH")
sys.exit(0 if ok else 1)
