"""Debug: check dgrad / wgrad of the tensor-core path exactly against the saved X record (TF graph with view dirs)."""
import sys, os, ctypes as C
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import torch
from test_gpu_tc_train import _inputs, CASES
import nerf_kinematics_b200 as nk
from nerf_kinematics_b200._lib import check, lib
from nerf_kinematics_b200.ops import _ptr, _stream

def unswizzle(rec, n_tiles, nblk):
    t = rec.view(torch.float16).view(n_tiles, nblk, 128, 8, 8)
    r = torch.arange(128, device=t.device) & 7
    c = torch.arange(8, device=t.device)
    idx = (c[None, :] ^ r[:, None])  # [128, 8] physical chunk of logical chunk c
    idx = idx[None, None, :, :, None].expand(n_tiles, nblk, 128, 8, 8)
    return torch.gather(t, 3, idx).reshape(n_tiles, nblk, 128, 64)

case = sys.argv[1] if len(sys.argv) > 1 else "tf8x256"
n, S = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (37, 27)
sd = CASES[case][0]()
m = nk.model_from_state_dict(sd).to("cuda")
rays, z, d_raw = _inputs(n, S, seed=5)
h, pid = m.prepare("bf16")
P = n * S
raw = torch.empty((P, 4), device="cuda")
grad = torch.zeros((m.n_params,), device="cuda")
nbytes = lib().nb_mlp_workspace_bytes(h, pid, P, 1)
ws = torch.zeros((nbytes,), dtype=torch.uint8, device="cuda")
check(lib().nb_mlp_forward(h, pid, _ptr(rays), rays.shape[1], _ptr(z), n, S, _ptr(raw), _ptr(ws), nbytes, 1, _stream()))
check(lib().nb_mlp_backward(h, pid, P, _ptr(d_raw.clone()), _ptr(ws), nbytes, _ptr(grad), _stream()))
torch.cuda.synchronize()
xo, xb, do, db = C.c_int64(), C.c_int(), C.c_int64(), C.c_int()
check(lib().nb_mlp_record_layout(h, P, _ptr(ws), C.byref(xo), C.byref(xb), C.byref(do), C.byref(db)))
nt = (P + 127) // 128
X = unswizzle(ws[xo.value: xo.value + nt * xb.value * 16384], nt, xb.value).double()
Dz = unswizzle(ws[do.value: do.value + nt * db.value * 16384], nt, db.value).double()
scale_bits = ws[:2048].view(torch.int32)
base = (-ws.data_ptr()) % 1024
mx = ws[base: base + 4].view(torch.float32).item()
import math
scale = 2.0 ** math.floor(math.log2(1024.0 / mx))
print("tiles", nt, "xrec blocks", xb.value, "drec blocks", db.value, "absmax", mx, "scale", scale)
def blocks(T, b0, nb):  # -> [P_pad, 64*nb]
    return T[:, b0:b0 + nb].permute(0, 2, 1, 3).reshape(nt * 128, nb * 64)
D_, W_ = 8, 256
q = lambda k: sd[k].half().double().cuda()
encx, encd = blocks(X, 0, 1), blocks(X, 1, 1)
hs = [blocks(X, 2 + 4 * i, 4) for i in range(D_)]
feat, dirh = blocks(X, 2 + 4 * D_, 4), blocks(X, 2 + 4 * D_ + 4, 2)
dr = torch.zeros(nt * 128, 4, dtype=torch.float64, device="cuda"); dr[:P] = d_raw.double()
def cmp(name, got, ref):
    err = (got - ref).abs().max().item(); rm = ref.abs().max().item()
    cos = (got * ref).sum().item() / max(got.norm().item() * ref.norm().item(), 1e-300)
    print("%-10s err %.3e ref %.3e rel %.5f cos %.7f" % (name, err, rm, err / max(rm, 1e-300), cos))
# reference chain
dz_dir = (dr[:, :3] @ q("rgb_linear.weight")) * (dirh > 0)
cmp("dz_dir", blocks(Dz, 0, 2) / scale, dz_dir)
dz_feat = dz_dir @ q("views_linears.0.weight")[:, :256]
cmp("dz_feat", blocks(Dz, 2, 4) / scale, dz_feat)
dz = (dz_feat @ q("bottleneck_linear.weight") + dr[:, 3:4] * q("alpha_linear.weight")) * (hs[7] > 0)
cmp("dz_7", blocks(Dz, 6, 4) / scale, dz)
dzs = {7: dz}
for i in range(7, 0, -1):
    Wi = q("pts_linears.%d.weight" % i)
    off = 63 if i - 1 == 4 else 0
    dz = (dzs[i] @ Wi[:, off:off + 256]) * (hs[i - 1] > 0)
    dzs[i - 1] = dz
    cmp("dz_%d" % (i - 1), blocks(Dz, 6 + 4 * (8 - i), 4) / scale, dz)
# wgrads from the KERNEL's dz (isolates the wgrad kernel) and from the reference dz
lay = {k: (sh, off) for k, sh, off in m._layout}
def G(k):
    sh, off = lay[k]; cnt = 1
    for s_ in sh: cnt *= s_
    return grad[off:off + cnt].view(sh).double()
for i in range(8):
    kdz = blocks(Dz, 6 + 4 * (7 - i), 4) / scale
    xin = encx[:, :63] if i == 0 else (torch.cat([encx[:, :63], hs[4]], 1) if i == 5 else hs[i - 1])
    cmp("W%d|kdz" % i, G("pts_linears.%d.weight" % i), kdz.t() @ xin)
    cmp("W%d|ref" % i, G("pts_linears.%d.weight" % i), dzs[i].t() @ xin)
    cmp("b%d|kdz" % i, G("pts_linears.%d.bias" % i), kdz.sum(0))
kd = blocks(Dz, 2, 4) / scale
cmp("Wfeat|kdz", G("bottleneck_linear.weight"), kd.t() @ hs[7])
kd = blocks(Dz, 0, 2) / scale
cmp("Wdir|kdz", G("views_linears.0.weight"), kd.t() @ torch.cat([feat, encd[:, :27]], 1))
cmp("Wrgb", G("rgb_linear.weight"), dr[:, :3].t() @ dirh)
cmp("Walpha", G("alpha_linear.weight"), dr[:, 3:4].t() @ hs[7])
