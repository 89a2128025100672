"""Small, complete pass over every kernel family for compute-sanitizer (memcheck / racecheck / synccheck):
    compute-sanitizer --tool memcheck python tools/sanitize_target.py
render (fp16 tensor-core and fp32 paths, ragged tile counts), one captured-free train step (forward with records, 1-bit
masks, dgrad chain, wgrad, reduce, fused Adam), the stage kernels at odd sizes."""
import os
import sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nerf_kinematics_b200 as nk  # noqa: E402
from nerf_kinematics_b200 import ops  # noqa: E402
from nerf_kinematics_b200.render_core import render_chunk  # noqa: E402
from nerf_kinematics_b200.trainer import TrainStep  # noqa: E402

dev = torch.device("cuda")
torch.manual_seed(0)
for mk in (lambda: nk.NeRF(), lambda: nk.FlexibleNeRFModel(num_encoding_fn_xyz=6)):
    mc, mf = mk().to(dev), mk().to(dev)
    n = 37
    o = torch.randn(n, 3, device=dev) * 0.5 + torch.tensor([0, 0, 4.0], device=dev)
    d = torch.nn.functional.normalize(torch.randn(n, 3, device=dev), dim=-1)
    rays = ops.make_ray_batch(o, d, 2.0, 6.0, use_viewdirs=True)
    with torch.no_grad():
        for prec in ("fp16", "fp32"):
            out = render_chunk(rays, mc, mf, N_samples=27, N_importance=45, white_bkgd=True, precision=prec)
            assert torch.isfinite(out["rgb"]).all()
    st = TrainStep(mc, mf, N_samples=27, N_importance=45, perturb=1.0, raw_noise_std=0.5, white_bkgd=True, precision="fp16",
                   cuda_graph=False)
    tgt = torch.rand(n, 3, device=dev)
    l0 = st.step(rays, tgt).clone()
    l1 = st.step(rays, tgt).clone()
    torch.cuda.synchronize()
    print(type(mc).__name__, "loss", l0.tolist(), "->", l1.tolist())
ro, rd = nk.get_rays(7, 9, 11.0, torch.eye(4))
ops.ndc_rays(7, 9, 11.0, 1.0, ro.reshape(-1, 3) + torch.tensor([0, 0, 2.0], device=dev), rd.reshape(-1, 3))
ops.searchsorted(torch.sort(torch.rand(5, 13, device=dev), -1)[0], torch.rand(5, 7, device=dev), side="right")
ops.posenc(torch.rand(33, 3, device=dev), 10)
torch.cuda.synchronize()
print("sanitize target ok")
