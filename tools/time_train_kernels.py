"""Per-kernel CUDA-event-free timing helper: runs K train steps (for use under `ncu --metrics gpu__time_duration.sum`)
or prints torch-profiler-free step time.  python tools/time_train_kernels.py [steps]"""
import os, sys, time
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R)
import torch
import nerf_kinematics_b200 as nk
from nerf_kinematics_b200 import ops
from nerf_kinematics_b200.trainer import TrainStep
from nerf_kinematics_b200.poses import pose_spherical
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 10
dev = torch.device("cuda")
torch.manual_seed(0)
mc, mf = nk.NeRF().to(dev), nk.NeRF().to(dev)
st = TrainStep(mc, mf, N_samples=64, N_importance=128, perturb=1.0, white_bkgd=True, precision="bf16")
ro, rd = nk.get_rays(800, 800, 1111.11, torch.from_numpy(pose_spherical(30.0, -30.0, 4.0)))
rays_all = ops.make_ray_batch(ro, rd, 2.0, 6.0, use_viewdirs=True)
g = torch.Generator(device="cuda").manual_seed(0)
def batch():
    idx = torch.randint(0, 640000, (1024,), device=dev, generator=g)
    return rays_all[idx], torch.rand(1024, 3, device=dev, generator=g)
for _ in range(3):
    st.step(*batch())
torch.cuda.synchronize()
bs = [batch() for _ in range(steps)]
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for b in bs:
    st.step(*b)
e1.record()
torch.cuda.synchronize()
print("NB_WG_DEBUG=%s  %.3f ms/step" % (os.environ.get("NB_WG_DEBUG", "0"), e0.elapsed_time(e1) / steps))
