import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from test_gpu_tc_train import _mlp_grads, _inputs, CASES
import nerf_kinematics_b200 as nk
case = sys.argv[1] if len(sys.argv) > 1 else "tf8x256"
n, S = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (37, 27)
sd = CASES[case]()
uv = any(k.startswith(("views_linears", "layers_dir")) for k in sd)
m = nk.model_from_state_dict(sd).to("cuda")
rays, z, d_raw = _inputs(n, S, seed=5, use_viewdirs=uv)
raw32, g32 = _mlp_grads(m, "fp32", rays, z, d_raw)
raw16, g16 = _mlp_grads(m, "bf16", rays, z, d_raw)
print("raw err", float((raw16 - raw32).abs().max()), "scale", float(raw32.abs().max()))
for key, shape, off in m._layout:
    cnt = 1
    for s_ in shape: cnt *= s_
    a, b = g16[off:off + cnt], g32[off:off + cnt]
    cos = float(torch.dot(a, b) / (a.norm() * b.norm()).clamp_min(1e-30))
    print("%-28s %-12s err %.3e ref %.3e rel %.4f cos %.5f norm ratio %.4f" % (key, tuple(shape), float((a - b).abs().max()), float(b.abs().max()),
          float((a - b).abs().max() / b.abs().max().clamp_min(1e-30)), cos, float(a.norm() / b.norm().clamp_min(1e-30))))
