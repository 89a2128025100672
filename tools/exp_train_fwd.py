"""Timing experiment (needs a -DNB_TIMING_EXPERIMENTS build): the train-step forward of the fine pass (1024 rays x 192)
with NB_FWD_DEBUG bits: 4 = no X-record stores, 8 = stores but no wait before the accumulator hand-off, 16 = no mask bits."""
import ctypes as C, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nerf_kinematics_b200 as nk
from nerf_kinematics_b200 import ops
from nerf_kinematics_b200._lib import check, lib
dev = torch.device("cuda"); torch.manual_seed(0)
m = nk.NeRF().to(dev)
n, S = 1024, 192
o = torch.randn(n, 3, device=dev) * 0.5 + torch.tensor([0, 0, 4.0], device=dev)
d = torch.nn.functional.normalize(torch.randn(n, 3, device=dev), dim=-1)
rays = ops.make_ray_batch(o, d, 2.0, 6.0, use_viewdirs=True); z = ops.sample_coarse(rays, S)
raw = torch.empty(n, S, 4, device=dev)
h, prec = m.prepare("fp16")
for save in (0, 1):
    nbytes = lib().nb_mlp_workspace_bytes(h, prec, n * S, save)
    ws = torch.empty(max(nbytes, 1), dtype=torch.uint8, device=dev)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    run = lambda: check(lib().nb_mlp_forward(h, prec, C.c_void_p(rays.data_ptr()), 11, C.c_void_p(z.data_ptr()), n, S, C.c_void_p(raw.data_ptr()), C.c_void_p(ws.data_ptr()), nbytes, save, st))
    for _ in range(5): run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50): run()
    e1.record(); torch.cuda.synchronize()
    print("NB_FWD_DEBUG=%s save=%d  %.1f us" % (os.environ.get("NB_FWD_DEBUG", "0"), save, e0.elapsed_time(e1) * 20))
