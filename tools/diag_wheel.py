"""Diagnostic: where do the fp16 and fp32 renders of the wheel_robot6 checkpoint differ at 1280x720 / NDC?"""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import nerf_kinematics_b200 as nk
from nerf_kinematics_b200 import ops
from nerf_kinematics_b200.poses import load_transforms
from nerf_kinematics_b200.render_core import render_chunk
tr = load_transforms(os.path.join(ROOT, "tests", "golden", "wheel_transforms_test_video.json"))
H, W, K, poses = tr["H"], tr["W"], tr["K"], tr["poses"]
ck = np.load(os.path.join(ROOT, "tests", "golden", "ckpt_wheel_robot6.npz"))
mc = nk.model_from_state_dict({k[2:]: torch.from_numpy(ck[k]) for k in ck.files if k.startswith("c.")}).cuda()
mf = nk.model_from_state_dict({k[2:]: torch.from_numpy(ck[k]) for k in ck.files if k.startswith("f.")}).cuda()
c2w = torch.from_numpy(poses[17])
ro, rd = nk.get_rays(H, W, K, c2w)
rays = ops.make_ray_batch(ro, rd, 0.0, 1.0, use_viewdirs=True, ndc=True, H=H, W=W, focal=float(K[0, 0]))
outs = {}
with torch.no_grad():
    for prec in ("fp32", "fp16"):
        parts = [render_chunk(rays[i:i + 32768], mc, mf, N_samples=64, N_importance=128, precision=prec, want_raw=True, want_z=True)
                 for i in range(0, rays.shape[0], 32768)]
        outs[prec] = {k: torch.cat([p[k] for p in parts]) for k in ("rgb", "acc", "rgb0", "raw", "z_vals")}
d = (outs["fp16"]["rgb"] - outs["fp32"]["rgb"]).abs().max(-1)[0]
d0 = (outs["fp16"]["rgb0"] - outs["fp32"]["rgb0"]).abs().max(-1)[0]
print("fine : max %.3e mean %.3e  frac>1e-2 %.3e  frac>1e-3 %.3e" % (d.max(), d.mean(), (d > 1e-2).float().mean(), (d > 1e-3).float().mean()))
print("coarse: max %.3e mean %.3e  frac>1e-2 %.3e" % (d0.max(), d0.mean(), (d0 > 1e-2).float().mean()))
print("raw32 absmax", outs["fp32"]["raw"].abs().amax(dim=(0, 1)).tolist(), "raw16 absmax", outs["fp16"]["raw"].abs().amax(dim=(0, 1)).tolist())
i = int(d.argmax())
print("worst ray", i, divmod(i, W), "rgb32", outs["fp32"]["rgb"][i].tolist(), "rgb16", outs["fp16"]["rgb"][i].tolist())
zz = (outs["fp16"]["z_vals"][i] - outs["fp32"]["z_vals"][i]).abs()
print("z diff max on that ray", float(zz.max()), "n z differing >1e-4:", int((zz > 1e-4).sum()))
r32, r16 = outs["fp32"]["raw"][i], outs["fp16"]["raw"][i]
j = int((r32 - r16).abs().max(-1)[0].argmax())
print("largest raw diff at sample", j, r32[j].tolist(), r16[j].tolist())
print("sigma32 top", torch.topk(r32[:, 3], 5)[0].tolist(), "sigma16 top", torch.topk(r16[:, 3], 5)[0].tolist())
# hidden activation magnitudes of the fine net on that ray (fp32 torch)
sd = {k[2:]: torch.from_numpy(ck[k]) for k in ck.files if k.startswith("f.")}
sys.path.insert(0, ROOT)
from oracle import nerf_oracle as O
rr = rays[i].cpu()
pts = rr[None, 0:3] + rr[None, 3:6] * outs["fp32"]["z_vals"][i].cpu()[:, None]
x = O.posenc(pts, mf.n_freq_xyz)
h = torch.nn.functional.linear(x, sd["layer1.weight"], sd["layer1.bias"])
print("layer1 |h| max", float(h.abs().max()))
for l in range(3):
    h = torch.relu(torch.nn.functional.linear(h, sd["layers_xyz.%d.weight" % l], sd["layers_xyz.%d.bias" % l]))
    print("layers_xyz.%d max" % l, float(h.max()))
print("weights absmax", {k: float(v.abs().max()) for k, v in sd.items() if k.endswith("weight")})
