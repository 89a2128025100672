"""GPU: the tensor-core (fp16 operand, fp32 accumulate) TRAIN path -- forward with the saved X record, fused dgrad
chain, split-K wgrad, head gradients -- against the fp32 kernels and the oracle's autograd.

Reference being matched: tf.GradientTape / loss.backward() through run_network + the network
(nerf/run_nerf.py:807-826, nerf-pytorch/train_nerf.py:286)."""
import ctypes as C
import math

import pytest
import torch
import torch.nn.functional as F

from conftest import load_golden, maxdiff
from oracle import nerf_oracle as O

pytestmark = pytest.mark.gpu

import nerf_kinematics_b200 as nk  # noqa: E402
from nerf_kinematics_b200._lib import check, lib  # noqa: E402
from nerf_kinematics_b200.ops import _ptr, _stream  # noqa: E402

DEV = "cuda"


def _mlp_grads(model, prec, rays, z, d_raw):
    """nb_mlp_forward(save) + nb_mlp_backward through the C-ABI; returns (raw, flat grad)."""
    h, pid = model.prepare(prec)
    n, S = z.shape
    P = n * S
    raw = torch.empty((P, 4), dtype=torch.float32, device=DEV)
    grad = torch.zeros((model.n_params,), dtype=torch.float32, device=DEV)
    nbytes = lib().nb_mlp_workspace_bytes(h, pid, P, 1)
    ws = torch.empty((max(int(nbytes), 1),), dtype=torch.uint8, device=DEV)
    check(lib().nb_mlp_forward(h, pid, _ptr(rays), rays.shape[1], _ptr(z), n, S, _ptr(raw), _ptr(ws), nbytes, 1,
                               _stream()), "nb_mlp_forward")
    d = d_raw.clone()
    check(lib().nb_mlp_backward(h, pid, P, _ptr(d), _ptr(ws), nbytes, _ptr(grad), _stream()), "nb_mlp_backward")
    torch.cuda.synchronize()
    return raw, grad


def _inputs(n, S, seed, use_viewdirs=True):
    g = torch.Generator().manual_seed(seed)
    o = torch.randn(n, 3, generator=g) * 0.3
    d = torch.nn.functional.normalize(torch.randn(n, 3, generator=g), dim=-1)
    rays = torch.cat([o, d, torch.full((n, 1), 2.0), torch.full((n, 1), 6.0)] + ([d] if use_viewdirs else []), -1)
    z = torch.sort(torch.rand(n, S, generator=g) * 4 + 2, dim=-1).values
    d_raw = torch.randn(n * S, 4, generator=g) * 1e-4 * torch.rand(n * S, 1, generator=g) ** 4
    return rays.contiguous().to(DEV), z.contiguous().to(DEV), d_raw.contiguous().to(DEV)


CASES = {
    "tf8x256": (lambda: O.init_tf8x256_state(seed=0), (4,)),
    "tf8x256_noviews": (lambda: O.init_tf8x256_state(use_viewdirs=False, seed=1), (4,)),
    "flexible4x128": (lambda: O.init_flexible_state(seed=2), ()),
    "tf4x128_skip1": (lambda: O.init_tf8x256_state(D=4, W=128, skips=(1,), seed=3), (1,)),
}


def _q16(t):
    """Round to fp16 with a straight-through gradient (what an MMA operand tile holds)."""
    return t + (t.detach().half().to(t.dtype) - t.detach())


def _emulated_16bit_grads(sd, skips, rays, z, d_raw, uv):
    """Autograd (fp64 accumulate, CPU) of the network the tensor-core kernels evaluate: weights, encodings and every
    layer's activations rounded to fp16, biases and accumulation in full precision.  Same ReLU masks as the kernel,
    which the fp32 network does not have: a pre-activation within ~5e-4 of zero flips between the two."""
    rays, z, d_raw = rays.cpu(), z.cpu(), d_raw.cpu()
    pts = (rays[:, None, 0:3] + rays[:, None, 3:6] * z[..., None]).reshape(-1, 3)
    ex = _q16(O.posenc(pts, 10)).double()
    ed = None
    if uv:
        S = z.shape[1]
        ed = _q16(O.posenc(rays[:, 8:11], 4))[:, None, :].expand(-1, S, -1).reshape(-1, 27).double()
    P = {k: v.clone().double().requires_grad_(True) for k, v in sd.items()}
    lin = lambda x, name: F.linear(x, _q16(P[name + ".weight"]), P[name + ".bias"])
    tf = "pts_linears.0.weight" in sd
    if tf:
        h, i = ex, 0
        while f"pts_linears.{i}.weight" in sd:
            h = _q16(F.relu(lin(h, f"pts_linears.{i}")))
            if i in skips:
                h = torch.cat((ex, h), -1)
            i += 1
        if uv:
            alpha = lin(h, "alpha_linear")
            y = torch.cat((_q16(lin(h, "bottleneck_linear")), ed), -1)
            y = _q16(F.relu(lin(y, "views_linears.0")))
            raw = torch.cat((lin(y, "rgb_linear"), alpha), -1)
        else:
            raw = lin(h, "output_linear")
    else:
        h = _q16(lin(ex, "layer1"))
        i = 0
        while f"layers_xyz.{i}.weight" in sd:
            h = _q16(F.relu(lin(h, f"layers_xyz.{i}")))
            i += 1
        alpha = lin(h, "fc_alpha")
        y = torch.cat((_q16(F.relu(lin(h, "fc_feat"))), ed), -1)
        y = _q16(F.relu(lin(y, "layers_dir.0")))
        raw = torch.cat((lin(y, "fc_rgb"), alpha), -1)
    (raw * d_raw.double()).sum().backward()
    return raw.detach(), {k: v.grad for k, v in P.items()}


@pytest.mark.parametrize("case", list(CASES))
@pytest.mark.parametrize("n,S", [(37, 27), (128, 64)])  # 999 points = ragged last tile; 8192 points = 64 tiles
def test_tc_backward_matches_16bit_network_autograd(case, n, S):
    sd = CASES[case][0]()
    uv = any(k.startswith(("views_linears", "layers_dir")) for k in sd)
    m = nk.model_from_state_dict(sd).to(DEV)
    rays, z, d_raw = _inputs(n, S, seed=5, use_viewdirs=uv)
    raw16, g16 = _mlp_grads(m, "bf16", rays, z, d_raw)
    assert torch.isfinite(g16).all()
    raw_e, g_e = _emulated_16bit_grads(sd, CASES[case][1], rays, z, d_raw, uv)
    assert maxdiff(raw16, raw_e) <= 2e-3 * max(1.0, float(raw_e.abs().max()))
    # The emulation rounds exact sin/cos to fp16, the kernel rounds its double-angle recurrence: a few encodings differ by
    # one fp16 ulp, which moves ~1e-4 of the ReLU units across zero -> ~1-3 % relative L2 on the trunk gradients (the
    # relative L2 error is sqrt(fraction of flipped units)).  Exactness of the backward itself is pinned by
    # test_tc_backward_exact_given_saved_activations below.
    for key, shape, off in m._layout:
        cnt = 1
        for s_ in shape:
            cnt *= s_
        a, b = g16[off:off + cnt].cpu().double(), g_e[key].reshape(-1)
        cos = float(torch.dot(a, b) / (a.norm() * b.norm()).clamp_min(1e-300))
        assert cos >= 0.995 and 0.97 <= float(a.norm() / b.norm()) <= 1.03, (key, cos)


def _unswizzle(rec, n_tiles, nblk):
    """[n_tiles * nblk * 16 KB] SWIZZLE_128B images -> [n_tiles, nblk, 128, 64] fp16 (16-byte chunk c of row r sits at
    chunk position c ^ (r & 7))."""
    t = rec.view(torch.float16).view(n_tiles, nblk, 128, 8, 8)
    r = torch.arange(128, device=t.device) & 7
    c = torch.arange(8, device=t.device)
    idx = (c[None, :] ^ r[:, None])[None, None, :, :, None].expand(n_tiles, nblk, 128, 8, 8)
    return torch.gather(t, 3, idx).reshape(n_tiles, nblk, 128, 64)


@pytest.mark.parametrize("case", list(CASES))
@pytest.mark.parametrize("n,S", [(37, 27), (128, 64)])  # 999 points = ragged last tile; 8192 points = 64 tiles
def test_tc_backward_exact_given_saved_activations(case, n, S):
    """dgrad chain and wgrad against an fp64 backward that starts from the SAME saved activations (X record): removes
    the ReLU-flip ambiguity, leaving only the fp16 rounding of the gradient tiles (2^-11 per layer)."""
    sd = CASES[case][0]()
    skips = CASES[case][1]
    uv = any(k.startswith(("views_linears", "layers_dir")) for k in sd)
    tf = "pts_linears.0.weight" in sd
    m = nk.model_from_state_dict(sd).to(DEV)
    rays, z, d_raw = _inputs(n, S, seed=6, use_viewdirs=uv)
    h, pid = m.prepare("bf16")
    P = n * S
    raw = torch.empty((P, 4), device=DEV)
    grad = torch.zeros((m.n_params,), device=DEV)
    nbytes = lib().nb_mlp_workspace_bytes(h, pid, P, 1)
    ws = torch.zeros((nbytes,), dtype=torch.uint8, device=DEV)
    check(lib().nb_mlp_forward(h, pid, _ptr(rays), rays.shape[1], _ptr(z), n, S, _ptr(raw), _ptr(ws), nbytes, 1, _stream()))
    check(lib().nb_mlp_backward(h, pid, P, _ptr(d_raw.clone()), _ptr(ws), nbytes, _ptr(grad), _stream()))
    torch.cuda.synchronize()
    xo, xb, do, db = C.c_int64(), C.c_int(), C.c_int64(), C.c_int()
    check(lib().nb_mlp_record_layout(h, P, _ptr(ws), C.byref(xo), C.byref(xb), C.byref(do), C.byref(db)))
    nt = (P + 127) // 128
    X = _unswizzle(ws[xo.value: xo.value + nt * xb.value * 16384], nt, xb.value).double()
    # blocks 0 / 1 are the encodings; activation blocks hold kActScale * h (a power of two: exact to undo)
    X[:, 2:] /= float(lib().nb_mlp_activation_scale())
    # the ReLU masks the dgrad chain used: 1 bit per pre-activation, taken from the fp32 accumulator's sign (a positive
    # value too small for the 16-bit record still passes its gradient, like in the fp32 reference)
    mo, mb = C.c_int64(), C.c_int()
    check(lib().nb_mlp_mask_record_layout(h, P, _ptr(ws), C.byref(mo), C.byref(mb)))
    assert mb.value == xb.value - 2
    words = ws[mo.value: mo.value + nt * mb.value * 128 * 8].view(torch.int32).view(nt, mb.value, 128, 2).to(torch.int64) & 0xFFFFFFFF
    sh = (31 - torch.arange(32, device=DEV)).view(1, 1, 1, 1, 32)
    M = ((words.unsqueeze(-1) >> sh) & 1).reshape(nt, mb.value, 128, 64).double()     # [tile, block, row, element]
    mask = lambda b0, nb: blocks(M, b0 - 2, nb) > 0                                   # by X-record block index
    blocks = lambda T, b0, nb: T[:, b0:b0 + nb].permute(0, 2, 1, 3).reshape(nt * 128, nb * 64)
    q = lambda k: sd[k].half().double().to(DEV)
    names = ([f"pts_linears.{i}" for i in range(64) if f"pts_linears.{i}.weight" in sd] if tf else
             ["layer1"] + [f"layers_xyz.{i}" for i in range(64) if f"layers_xyz.{i}.weight" in sd])
    D, W = len(names), sd[names[0] + ".weight"].shape[0]
    wb = W // 64
    encx, encd = blocks(X, 0, 1)[:, :63], blocks(X, 1, 1)[:, :27]
    hs = [blocks(X, 2 + wb * i, wb) for i in range(D)]
    relu_after = lambda i: tf or i > 0
    dr = torch.zeros(nt * 128, 4, dtype=torch.float64, device=DEV)
    dr[:P] = d_raw.double()
    ref = {}
    if uv:
        n_alpha, n_feat, n_dir, n_rgb = (("alpha_linear", "bottleneck_linear", "views_linears.0", "rgb_linear") if tf else
                                         ("fc_alpha", "fc_feat", "layers_dir.0", "fc_rgb"))
        fw = sd[n_feat + ".weight"].shape[0]
        fb, dbk = fw // 64, (W // 2) // 64
        feat, dirh = blocks(X, 2 + wb * D, fb), blocks(X, 2 + wb * D + fb, dbk)
        assert xb.value == 2 + wb * D + fb + dbk
        dz_dir = (dr[:, :3] @ q(n_rgb + ".weight")) * mask(2 + wb * D + fb, dbk)
        dz_feat = dz_dir @ q(n_dir + ".weight")[:, :fw]
        if not tf:
            dz_feat = dz_feat * mask(2 + wb * D, fb)
        dz = dz_feat @ q(n_feat + ".weight") + dr[:, 3:4] * q(n_alpha + ".weight")
        ref[n_rgb + ".weight"], ref[n_rgb + ".bias"] = dr[:, :3].t() @ dirh, dr[:, :3].sum(0)
        ref[n_alpha + ".weight"], ref[n_alpha + ".bias"] = dr[:, 3:4].t() @ hs[-1], dr[:, 3:4].sum(0)
        ref[n_dir + ".weight"], ref[n_dir + ".bias"] = dz_dir.t() @ torch.cat([feat, encd], 1), dz_dir.sum(0)
        ref[n_feat + ".weight"], ref[n_feat + ".bias"] = dz_feat.t() @ hs[-1], dz_feat.sum(0)
    else:
        dz = dr @ q("output_linear.weight")
        ref["output_linear.weight"], ref["output_linear.bias"] = dr.t() @ hs[-1], dr.sum(0)
    if relu_after(D - 1):
        dz = dz * mask(2 + wb * (D - 1), wb)
    for i in range(D - 1, -1, -1):
        cat = i >= 1 and (i - 1) in skips
        xin = encx if i == 0 else (torch.cat([encx, hs[i - 1]], 1) if cat else hs[i - 1])
        ref[names[i] + ".weight"], ref[names[i] + ".bias"] = dz.t() @ xin, dz.sum(0)
        if i > 0:
            off = 63 if cat else 0
            dz = dz @ q(names[i] + ".weight")[:, off:off + W]
            if relu_after(i - 1):
                dz = dz * mask(2 + wb * (i - 1), wb)
    bad = []
    for key, shape, off in m._layout:
        cnt = 1
        for s_ in shape:
            cnt *= s_
        a, b = grad[off:off + cnt].double(), ref[key].reshape(-1)
        err, rm = float((a - b).abs().max()), float(b.abs().max())
        # bias gradients are sums of signed fp16-rounded terms (cancellation): 1 %; weights 0.4 %
        if err > (1e-2 if key.endswith(".bias") else 4e-3) * rm + 1e-14:
            bad.append((key, err, rm))
    assert not bad, bad



@pytest.mark.parametrize("case", ["tf8x256", "flexible4x128"])
def test_tc_backward_close_to_fp32_kernels(case):
    """Against the fp32 network the difference is dominated by ReLU units whose pre-activation the 16-bit forward
    moves across zero (~0.05 % of units -> ~2-4 % relative L2): direction and norm must agree."""
    sd = CASES[case][0]()
    m = nk.model_from_state_dict(sd).to(DEV)
    rays, z, d_raw = _inputs(512, 64, seed=5)
    raw32, g32 = _mlp_grads(m, "fp32", rays, z, d_raw)
    raw16, g16 = _mlp_grads(m, "bf16", rays, z, d_raw)
    assert maxdiff(raw16, raw32) <= 2e-2 * max(1.0, float(raw32.abs().max()))
    for key, shape, off in m._layout:
        cnt = 1
        for s_ in shape:
            cnt *= s_
        a, b = g16[off:off + cnt], g32[off:off + cnt]
        cos = float(torch.dot(a, b) / (a.norm() * b.norm()).clamp_min(1e-30))
        assert cos >= 0.995 and 0.97 <= float(a.norm() / b.norm()) <= 1.03, (key, cos)


def test_tc_backward_accumulates_and_is_deterministic():
    sd = CASES["tf8x256"][0]()
    m = nk.model_from_state_dict(sd).to(DEV)
    rays, z, d_raw = _inputs(64, 64, seed=9)
    _, g1 = _mlp_grads(m, "bf16", rays, z, d_raw)
    _, g2 = _mlp_grads(m, "bf16", rays, z, d_raw)
    # the split-K partials are reduced in a fixed order; only the 640 head entries use atomics
    big = [(k, sh, off) for k, sh, off in m._layout if "pts_linears" in k or "bottleneck" in k or "views" in k]
    for k, sh, off in big:
        cnt = 1
        for s_ in sh:
            cnt *= s_
        assert torch.equal(g1[off:off + cnt], g2[off:off + cnt]), k
    # scaling d_raw by a power of two scales the gradient exactly (the per-call loss scale follows max|d_raw|)
    _, g4 = _mlp_grads(m, "bf16", rays, z, d_raw * 4.0)
    k, sh, off = big[3]
    assert torch.equal(g4[off:off + 100], g1[off:off + 100] * 4.0)


def test_tc_train_step_against_oracle_autograd():
    """Whole train step (render + loss + backward) with the tensor-core MLPs vs the oracle's autograd."""
    sdc, sdf = O.init_tf8x256_state(seed=0), O.init_tf8x256_state(seed=1)
    mc, mf = nk.model_from_state_dict(sdc).to(DEV), nk.model_from_state_dict(sdf).to(DEV)
    g = load_golden("train_flexible")
    rays, tgt = g["rays"][:96], g["target"][:96]
    loss, gc, gf, _ = O.train_step_grads("tf", sdc, sdf, rays, tgt, n_coarse=64, n_fine=128, n_freq_xyz=10,
                                         n_freq_dir=4, white_bkgd=True)
    from nerf_kinematics_b200.render_core import render_chunk
    out = render_chunk(rays.to(DEV), mc, mf, N_samples=64, N_importance=128, white_bkgd=True, precision="bf16")
    l2 = nk.img2mse(out["rgb0"], tgt.to(DEV)) + nk.img2mse(out["rgb"], tgt.to(DEV))
    l2.backward()
    assert abs(float(l2.detach()) - float(loss)) <= 2e-3
    for name, m, ref in (("coarse", mc, gc), ("fine", mf, gf)):
        num = 0.0
        for k, p in m.named_parameters():
            assert p.grad is not None and torch.isfinite(p.grad).all(), k
            num += float((p.grad.cpu().double() * ref[k].double()).sum())
        gn = sum(float(p.grad.double().norm() ** 2) for p in m.parameters()) ** 0.5
        rn = sum(float(v.double().norm() ** 2) for v in ref.values()) ** 0.5
        cos = num / max(gn * rn, 1e-30)
        # the fine network's samples depend on the coarse weights through a discontinuous sampler: direction, not bits
        assert cos >= (0.995 if name == "coarse" else 0.97), (name, cos)
        assert 0.9 <= gn / rn <= 1.1, (name, gn, rn)


# ------------------------------------------------------------------------------------ fused Adam + CUDA-graph step
def _train_setup(seed=0, **kw):
    from nerf_kinematics_b200.trainer import TrainStep
    sdc, sdf = O.init_flexible_state(seed=seed), O.init_flexible_state(seed=seed + 1)
    mc, mf = nk.model_from_state_dict(sdc).to(DEV), nk.model_from_state_dict(sdf).to(DEV)
    g = load_golden("train_flexible")
    rays, tgt = g["rays"][:64].to(DEV), g["target"][:64].to(DEV)
    kw.setdefault("perturb", 0.0)
    st = TrainStep(mc, mf, N_samples=64, N_importance=128, white_bkgd=True, lrate=5e-3, lrate_decay=1, **kw)
    return st, mc, mf, rays, tgt


@pytest.mark.parametrize("n", [4096, 1000003])
def test_adam_kernel_matches_torch_adam(n):
    """nb_adam_step against torch.optim.Adam (PT train_nerf.py:288-289) and the TF formula (run_nerf.py:827) on
    identical gradients, 3 steps with changing learning rate."""
    g = torch.Generator().manual_seed(n)
    p0 = torch.randn(n, generator=g).to(DEV)
    grads = [(torch.randn(n, generator=g) * 10 ** float(torch.randint(-6, 0, (1,), generator=g))).to(DEV) for _ in range(3)]
    ref = torch.nn.Parameter(p0.clone())
    opt = torch.optim.Adam([ref], lr=1e-3, betas=(0.9, 0.999), eps=1e-8, foreach=False)
    for tf_style in (0, 1):
        p, m, v = p0.clone(), torch.zeros(n, device=DEV), torch.zeros(n, device=DEV)
        pt, mt, vt = p0.double().clone(), torch.zeros(n, device=DEV, dtype=torch.float64), torch.zeros(n, device=DEV, dtype=torch.float64)
        for t, gr in enumerate(grads, 1):
            lr = 1e-3 * 0.5 ** (t - 1)
            hyper = torch.tensor([lr, 0.9, 0.999, 1e-8, 1 - 0.9 ** t, 1 - 0.999 ** t, 1 - 0.9, 1 - 0.999], dtype=torch.float32, device=DEV)
            gbuf = gr.clone()
            check(lib().nb_adam_step(_ptr(p), _ptr(gbuf), _ptr(m), _ptr(v), n, _ptr(hyper), tf_style, 1, _stream()))
            assert float(gbuf.abs().max()) == 0.0  # zero_grad fused
            if tf_style == 0:
                opt.param_groups[0]["lr"] = lr
                ref.grad = gr.clone()
                opt.step()
            else:
                mt = 0.9 * mt + 0.1 * gr.double()
                vt = 0.999 * vt + 0.001 * gr.double() ** 2
                pt = pt - lr * math.sqrt(1 - 0.999 ** t) / (1 - 0.9 ** t) * mt / (vt.sqrt() + 1e-8)
        if tf_style == 0:
            st = opt.state[ref]
            assert maxdiff(p, ref.detach()) <= 2e-6
            assert maxdiff(m, st["exp_avg"]) <= 1e-6 * float(st["exp_avg"].abs().max())
            assert maxdiff(v, st["exp_avg_sq"]) <= 1e-6 * float(st["exp_avg_sq"].abs().max())
        else:
            assert maxdiff(p, pt) <= 2e-6


@pytest.mark.parametrize("prec", ["fp32", "bf16"])
def test_fused_adam_train_step_and_checkpoint_layout(prec):
    """TrainStep with the fused optimizer against TrainStep driving torch.optim.Adam: same loss trajectory (early Adam
    updates are sign-like, so parameters whose gradient is ~0 may legitimately differ), same learning-rate schedule,
    and an optimizer state dict in the reference checkpoint's layout (train_nerf.py:401-414) that round-trips."""
    a, mca, mfa, rays, tgt = _train_setup(precision=prec, fused_adam=True)
    b, mcb, mfb, _, _ = _train_setup(precision=prec, fused_adam=False)
    for _ in range(4):
        la = a.step(rays, tgt).clone()
        lb = b.step(rays, tgt).clone()
        assert maxdiff(la, lb) <= 2e-3
    assert abs(a.opt.param_groups[0]["lr"] - b.opt.param_groups[0]["lr"]) < 1e-12
    sa, sb = a.optimizer_state_dict(), b.opt.state_dict()
    assert sa["state"].keys() == sb["state"].keys()
    assert sa["param_groups"][0]["params"] == sb["param_groups"][0]["params"]
    for k in sa["state"]:
        assert float(sa["state"][k]["step"]) == float(sb["state"][k]["step"]) == 4.0
        assert sa["state"][k]["exp_avg"].shape == sb["state"][k]["exp_avg"].shape
    c, mcc, mfc, _, _ = _train_setup(precision=prec, fused_adam=True)
    c.load_optimizer_state_dict(sb)
    assert c.global_step == 4
    k0 = next(iter(sb["state"]))
    assert maxdiff(c.exp_avg[: sb["state"][k0]["exp_avg"].numel()].view_as(sb["state"][k0]["exp_avg"]), sb["state"][k0]["exp_avg"]) == 0.0


def test_cuda_graph_step_equals_eager_step():
    """The captured iteration (repack, render, loss, backward, Adam in ONE graph launch) replays to the same
    parameters as the eager sequence; perturb=0 so both see identical samples."""
    a, mca, mfa, rays, tgt = _train_setup(precision="bf16", cuda_graph=True)
    b, mcb, mfb, _, _ = _train_setup(precision="bf16", cuda_graph=False)
    losses = []
    for i in range(5):
        r = torch.roll(rays, i, 0)
        t = torch.roll(tgt, i, 0)
        la, lb = a.step(r, t).clone(), b.step(r, t).clone()
        losses.append(float(la[0]))
        assert maxdiff(la, lb) <= 1e-6, i
    assert a.graph_replays == 5 and a.launches_per_step and a.launches_per_step >= 10
    for (k, pa), (_, pb) in zip(list(mca.named_parameters()) + list(mfa.named_parameters()),
                                list(mcb.named_parameters()) + list(mfb.named_parameters())):
        assert maxdiff(pa, pb) <= 1e-6 + 1e-5 * float(pb.detach().abs().max()), k
    assert losses[-1] < losses[0]
    # rendering with the updated weights works after graph steps (the 16-bit images are repacked on demand)
    from nerf_kinematics_b200.render_core import render_chunk
    with torch.no_grad():
        oa = render_chunk(rays, mca, mfa, N_samples=64, N_importance=128, white_bkgd=True, precision="bf16")
        ob = render_chunk(rays, mcb, mfb, N_samples=64, N_importance=128, white_bkgd=True, precision="bf16")
    assert maxdiff(oa["rgb"], ob["rgb"]) <= 1e-4


def test_two_chain_graph_step_is_bit_identical_to_the_single_chain_step():
    """The captured single-GPU iteration runs the coarse network's backward on a side stream, forked right after the coarse
    loss (trainer._device_iteration_two_streams, nb_render_rays_part); the chains share no buffer, so losses and every
    updated parameter must equal the single-chain graph's bit for bit -- with in-kernel jitter and noise on, too."""
    res = []
    for streams in (True, False):
        st, mc, mf, rays, tgt = _train_setup(precision="bf16", cuda_graph=True, perturb=1.0, raw_noise_std=0.5, seed=3)
        st._bwd_streams = streams
        losses = [st.step(torch.roll(rays, i, 0), torch.roll(tgt, i, 0)).clone() for i in range(4)]
        torch.cuda.synchronize()
        res.append((losses, [p.detach().clone() for p in list(mc.parameters()) + list(mf.parameters())]))
        st.close()
    for la, lb in zip(res[0][0], res[1][0]):
        assert torch.equal(la, lb)
    for pa, pb in zip(res[0][1], res[1][1]):
        assert torch.equal(pa, pb)


def test_step_pixels_equals_step_on_gathered_rays():
    """selecting pixels on the device and generating only their rays gives the same iteration as
    building the whole image's bundle and indexing it (train_nerf.py:240-253)."""
    import numpy as np
    from nerf_kinematics_b200 import ops
    H, W, f = 40, 56, 70.0
    c2w = O.pose_spherical(25.0, -30.0, 4.0)
    img = torch.rand(H, W, 3, generator=torch.Generator().manual_seed(3)).to(DEV)
    sel = torch.from_numpy(np.random.RandomState(1).choice(H * W, 64, replace=False)).to(DEV)
    a, mca, mfa, _, _ = _train_setup(precision="bf16", cuda_graph=True)
    b, mcb, mfb, _, _ = _train_setup(precision="bf16", cuda_graph=True)
    ro, rd = ops.get_rays(H, W, f, c2w)
    rays = ops.make_ray_batch(ro.reshape(-1, 3)[sel], rd.reshape(-1, 3)[sel], 2.0, 6.0, use_viewdirs=True)
    for _ in range(2):
        la = a.step_pixels(H, W, f, c2w, img, sel, 2.0, 6.0).clone()
        lb = b.step(rays, img.reshape(-1, 3)[sel]).clone()
        assert torch.equal(la, lb)
    for (k, pa), (_, pb) in zip(mfa.named_parameters(), mfb.named_parameters()):
        assert torch.equal(pa, pb), k
