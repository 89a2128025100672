"""ncu target: two launches of the four-tile forward kernel (FlexibleNeRFModel 4x128, 32 768 rays x 192 samples)."""
import os
import sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nerf_kinematics_b200 as nk  # noqa: E402
from nerf_kinematics_b200 import ops  # noqa: E402

dev = torch.device("cuda")
torch.manual_seed(0)
m = nk.FlexibleNeRFModel(num_encoding_fn_xyz=6).to(dev)
n, S = 32768, 192
o = torch.randn(n, 3, device=dev) * 0.5 + torch.tensor([0, 0, 4.0], device=dev)
d = torch.nn.functional.normalize(torch.randn(n, 3, device=dev), dim=-1)
rays = ops.make_ray_batch(o, d, 2.0, 6.0, use_viewdirs=True)
z = ops.sample_coarse(rays, S)
pts = rays[:, None, 0:3] + rays[:, None, 3:6] * z[..., None]
with torch.no_grad():
    for _ in range(2):
        m.query(pts, rays[:, 8:11], precision="fp16")
torch.cuda.synchronize()
print("done")
