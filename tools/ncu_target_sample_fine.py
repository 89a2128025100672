"""Short target for `ncu -k regex:sample_fine`: the fused pdf + merge kernel on 262 144 rays (64 depths, 128 samples),
deterministic u.  NB_SAMPLE_FINE2=0 selects the one-ray-per-warp kernel."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_kinematics_b200 import ops
dev = torch.device("cuda:0")
torch.manual_seed(0)
n = 1 << 18
o = torch.randn(n, 3, device=dev) * 0.5 + torch.tensor([0, 0, 4.0], device=dev)
d = torch.nn.functional.normalize(torch.randn(n, 3, device=dev), dim=-1)
rays = ops.make_ray_batch(o, d, 2.0, 6.0, use_viewdirs=True)
z = ops.sample_coarse(rays, 64)
w = torch.rand(n, 64, device=dev) ** 4
for _ in range(4):
    ops.sample_fine(z, w, 128)
torch.cuda.synchronize()
print("ok")
