"""Fixed per-launch cost of the persistent tensor-core kernels: time the inference forward at 1, 2, 4, ... tiles per CTA
(148 CTAs x 128-point tiles) and fit T = c + r * pair_rounds.  python tools/fixed_cost_sweep.py"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import nerf_kinematics_b200 as nk
from nerf_kinematics_b200 import ops
from nerf_kinematics_b200._lib import check, lib
dev = torch.device("cuda:0")
torch.manual_seed(0)
model = nk.NeRF().to(dev)
h, prec = model.prepare("bf16")
S = 64
res = []
for tiles_per_cta in (2, 4, 8, 16, 32):
    n = 148 * tiles_per_cta * 128 // S
    o = torch.randn(n, 3, device=dev) * 0.5 + torch.tensor([0, 0, 4.0], device=dev)
    d = torch.nn.functional.normalize(torch.randn(n, 3, device=dev), dim=-1)
    rays = ops.make_ray_batch(o, d, 2.0, 6.0, use_viewdirs=True)
    z = ops.sample_coarse(rays, S)
    raw = torch.empty(n, S, 4, device=dev)
    nbytes = lib().nb_mlp_workspace_bytes(h, prec, n * S, 0)
    ws = torch.empty(max(nbytes, 1), dtype=torch.uint8, device=dev)
    def run():
        st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        check(lib().nb_mlp_forward(h, prec, C.c_void_p(rays.data_ptr()), 11, C.c_void_p(z.data_ptr()), n, S,
                                   C.c_void_p(raw.data_ptr()), C.c_void_p(ws.data_ptr()), nbytes, 0, st))
    g = torch.cuda.CUDAGraph()
    run(); torch.cuda.synchronize()
    with torch.cuda.graph(g):
        for _ in range(20):
            run()
    g.replay(); torch.cuda.synchronize()
    ts = []
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) / 20 * 1e3)
    t = sorted(ts)[2]
    res.append((tiles_per_cta, t))
    print("tiles/CTA %3d  %8.1f us per launch (20 launches per graph replay)" % (tiles_per_cta, t))
(x0, y0), (x1, y1) = res[1], res[-1]
r = (y1 - y0) / (x1 - x0)
print("per tile %.2f us, intercept (fixed cost incl. the in-graph launch gap) %.1f us" % (r, y0 - r * x0))
