"""CPU oracle for the classic-NeRF hot path.  TEST INFRASTRUCTURE ONLY.

This file is a plain PyTorch-CPU (fp32) restatement of the reference's algorithm for the
path in SURVEY.md section 8.  It is never imported by the product package
(``nerf_kinematics_b200``); only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may use it, and only as the
checker / CPU baseline.

Parity status: PINNED.  Every function below is checked bit-for-bit (or to the stated
tolerance) against the reference's own PyTorch implementation imported from
``/root/reference/nerf-pytorch`` (``oracle/reference_shim.py``) by
``oracle/make_golden.py``; the resulting vectors live in ``tests/golden/*.npz`` and are
re-checked on every ``pytest -m "not gpu"`` run (tests/test_oracle_golden.py).  The
TensorFlow original (``nerf/run_nerf.py``) cannot run here (no TF); its 8x256 graph and its
``render``/``render_rays`` plumbing are restated from the source and their stage numerics
are delegated to the same stage functions, which are expression-for-expression identical
between ``nerf/run_nerf_helpers.py`` and ``nerf-pytorch/nerf/nerf_helpers.py``.

Citations are relative to /root/reference.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.nn.functional as F

Tensor = torch.Tensor


# --------------------------------------------------------------------------------------
# a1  ray generation
# --------------------------------------------------------------------------------------
def pixel_rays(H: int, W: int, focal, c2w: Tensor, K=None) -> Tuple[Tensor, Tensor]:
    """Rays through every pixel, shape [H, W, 3] each.

    Follows nerf-pytorch/nerf/nerf_helpers.py:67-110 (get_ray_bundle) which is the twin of
    nerf/run_nerf_helpers.py:123-130 (get_rays): camera direction
    ``[(i-W/2)/f, -(j-H/2)/f, -1]`` (no half-pixel offset), rotated by multiply-broadcast
    and a sum over the last axis (not a matmul), origin = c2w[:3, 3] broadcast.
    ``K`` (3x3) generalises focal/centre as yenchenlin's K-API does: fx, fy, cx, cy.
    """
    c2w = c2w.to(torch.float32)
    if K is not None:
        fx, fy, cx, cy = float(K[0][0]), float(K[1][1]), float(K[0][2]), float(K[1][2])
    else:
        fx = fy = float(focal)
        cx, cy = W * 0.5, H * 0.5
    col = torch.arange(W, dtype=torch.float32)
    row = torch.arange(H, dtype=torch.float32)
    jj, ii = torch.meshgrid(row, col, indexing="ij")  # ii = column index, jj = row index
    cam = torch.stack([(ii - cx) / fx, -(jj - cy) / fy, -torch.ones_like(ii)], dim=-1)
    rays_d = torch.sum(cam[..., None, :] * c2w[:3, :3], dim=-1)
    rays_o = c2w[:3, -1].expand(rays_d.shape)
    return rays_o, rays_d


# --------------------------------------------------------------------------------------
# a2  unit view directions (computed from world-space directions, before NDC)
# --------------------------------------------------------------------------------------
def unit_viewdirs(rays_d: Tensor) -> Tensor:
    """nerf-pytorch/nerf/train_utils.py:143-148 / nerf/run_nerf.py:297-307."""
    v = rays_d / rays_d.norm(p=2, dim=-1).unsqueeze(-1)
    return v.reshape(-1, 3)


# --------------------------------------------------------------------------------------
# a3  NDC rays
# --------------------------------------------------------------------------------------
def to_ndc(H: int, W: int, focal: float, near: float, rays_o: Tensor, rays_d: Tensor):
    """nerf-pytorch/nerf/nerf_helpers.py:170-197 == nerf/run_nerf_helpers.py:143-178."""
    t = -(near + rays_o[..., 2]) / rays_d[..., 2]
    rays_o = rays_o + t[..., None] * rays_d
    sx = -1.0 / (W / (2.0 * focal))
    sy = -1.0 / (H / (2.0 * focal))
    o0 = sx * rays_o[..., 0] / rays_o[..., 2]
    o1 = sy * rays_o[..., 1] / rays_o[..., 2]
    o2 = 1.0 + 2.0 * near / rays_o[..., 2]
    d0 = sx * (rays_d[..., 0] / rays_d[..., 2] - rays_o[..., 0] / rays_o[..., 2])
    d1 = sy * (rays_d[..., 1] / rays_d[..., 2] - rays_o[..., 1] / rays_o[..., 2])
    d2 = -2.0 * near / rays_o[..., 2]
    return torch.stack([o0, o1, o2], -1), torch.stack([d0, d1, d2], -1)


# --------------------------------------------------------------------------------------
# a5  stratified depths
# --------------------------------------------------------------------------------------
def coarse_depths(near: Tensor, far: Tensor, n_samples: int, lindisp: bool = False,
                  t_rand: Optional[Tensor] = None) -> Tensor:
    """nerf-pytorch/nerf/train_utils.py:45-65 == nerf/run_nerf.py:181-199.

    ``near``/``far`` are [N, 1].  ``t_rand`` ([N, n_samples] in [0,1)) turns on the
    stratified jitter (perturb); parity is only defined for a supplied ``t_rand``.
    """
    t = torch.linspace(0.0, 1.0, n_samples, dtype=near.dtype)
    if not lindisp:
        z = near * (1.0 - t) + far * t
    else:
        z = 1.0 / (1.0 / near * (1.0 - t) + 1.0 / far * t)
    z = z.expand(near.shape[0], n_samples)
    if t_rand is not None:
        mids = 0.5 * (z[..., 1:] + z[..., :-1])
        upper = torch.cat((mids, z[..., -1:]), dim=-1)
        lower = torch.cat((z[..., :1], mids), dim=-1)
        z = lower + (upper - lower) * t_rand
    return z


def points_on_rays(rays_o: Tensor, rays_d: Tensor, z: Tensor) -> Tensor:
    """pts = o + d*z  (train_utils.py:67 / run_nerf.py:202)."""
    return rays_o[..., None, :] + rays_d[..., None, :] * z[..., :, None]


# --------------------------------------------------------------------------------------
# a6  positional encoding
# --------------------------------------------------------------------------------------
def posenc(x: Tensor, n_freqs: int, include_input: bool = True) -> Tensor:
    """nerf-pytorch/nerf/nerf_helpers.py:113-157 == Embedder nerf/run_nerf_helpers.py:22-56.

    Channel order [x, sin(2^0 x), cos(2^0 x), ..., sin(2^(L-1) x), cos(2^(L-1) x)];
    bands are exact powers of two (2**linspace(0, L-1, L)).
    """
    parts = [x] if include_input else []
    bands = 2.0 ** torch.linspace(0.0, n_freqs - 1, n_freqs, dtype=x.dtype) if n_freqs > 0 else []
    for f in bands:
        parts.append(torch.sin(x * f))
        parts.append(torch.cos(x * f))
    return parts[0] if len(parts) == 1 else torch.cat(parts, dim=-1)


# --------------------------------------------------------------------------------------
# a7 / a8  the two MLP graphs, as pure functions of a state_dict
# --------------------------------------------------------------------------------------
def mlp_flexible(sd: Dict[str, Tensor], x: Tensor, dim_xyz: int, use_viewdirs: bool = True) -> Tensor:
    """FlexibleNeRFModel.forward at its constructor defaults (4 layers, skip never fires):
    nerf-pytorch/nerf/models.py:233-256.  ``layer1`` has NO ReLU (:238); feat = ReLU(fc_feat)
    (:248); alpha = fc_alpha on the trunk output (:249); dir layer on [feat, view] (:250-252).
    """
    xyz, view = x[..., :dim_xyz], x[..., dim_xyz:]
    h = F.linear(xyz, sd["layer1.weight"], sd["layer1.bias"])
    i = 0
    while f"layers_xyz.{i}.weight" in sd:
        h = F.relu(F.linear(h, sd[f"layers_xyz.{i}.weight"], sd[f"layers_xyz.{i}.bias"]))
        i += 1
    if not use_viewdirs:
        return F.linear(h, sd["fc_out.weight"], sd["fc_out.bias"])
    feat = F.relu(F.linear(h, sd["fc_feat.weight"], sd["fc_feat.bias"]))
    alpha = F.linear(h, sd["fc_alpha.weight"], sd["fc_alpha.bias"])
    y = torch.cat((feat, view), dim=-1)
    j = 0
    while f"layers_dir.{j}.weight" in sd:
        y = F.relu(F.linear(y, sd[f"layers_dir.{j}.weight"], sd[f"layers_dir.{j}.bias"]))
        j += 1
    rgb = F.linear(y, sd["fc_rgb.weight"], sd["fc_rgb.bias"])
    return torch.cat((rgb, alpha), dim=-1)


def mlp_tf8x256(sd: Dict[str, Tensor], x: Tensor, input_ch: int, use_viewdirs: bool = True,
                skips: Sequence[int] = (4,)) -> Tensor:
    """The Keras functional graph of init_nerf_model, nerf/run_nerf_helpers.py:80-118.

    Trunk: D x (Dense(W)+ReLU); after layer index i in ``skips`` the activations become
    concat([inputs_pts, h]) (:97-100).  Heads with view dirs (:102-113): alpha = Dense(1)(h),
    bottleneck = Dense(256, linear)(h), Dense(W//2)+ReLU on [bottleneck, views], rgb = Dense(3);
    output [rgb, alpha].  Without: Dense(4) (:115).  Weights here use nn.Linear's [out, in]
    convention under the keys pts_linears.{i}, alpha_linear, bottleneck_linear,
    views_linears.0, rgb_linear, output_linear.
    """
    pts, views = x[..., :input_ch], x[..., input_ch:]
    h = pts
    i = 0
    while f"pts_linears.{i}.weight" in sd:
        h = F.relu(F.linear(h, sd[f"pts_linears.{i}.weight"], sd[f"pts_linears.{i}.bias"]))
        if i in skips:
            h = torch.cat((pts, h), dim=-1)
        i += 1
    if not use_viewdirs:
        return F.linear(h, sd["output_linear.weight"], sd["output_linear.bias"])
    alpha = F.linear(h, sd["alpha_linear.weight"], sd["alpha_linear.bias"])
    bott = F.linear(h, sd["bottleneck_linear.weight"], sd["bottleneck_linear.bias"])
    y = torch.cat((bott, views), dim=-1)
    y = F.relu(F.linear(y, sd["views_linears.0.weight"], sd["views_linears.0.bias"]))
    rgb = F.linear(y, sd["rgb_linear.weight"], sd["rgb_linear.bias"])
    return torch.cat((rgb, alpha), dim=-1)


def init_tf8x256_state(D=8, W=256, input_ch=63, input_ch_views=27, use_viewdirs=True,
                       skips=(4,), seed=0) -> Dict[str, Tensor]:
    """Random-init weights with nn.Linear's default init, seeded."""
    g = torch.Generator().manual_seed(seed)
    sd: Dict[str, Tensor] = {}

    def lin(name, fin, fout):
        bound = 1.0 / math.sqrt(fin)
        sd[name + ".weight"] = (torch.rand(fout, fin, generator=g) * 2 - 1) * bound
        sd[name + ".bias"] = (torch.rand(fout, generator=g) * 2 - 1) * bound

    fin = input_ch
    for i in range(D):
        lin(f"pts_linears.{i}", fin, W)
        fin = W + input_ch if i in skips else W
    if use_viewdirs:
        lin("alpha_linear", fin, 1)
        lin("bottleneck_linear", fin, 256)
        lin("views_linears.0", 256 + input_ch_views, W // 2)
        lin("rgb_linear", W // 2, 3)
    else:
        lin("output_linear", fin, 4)
    return sd


def init_flexible_state(hidden=128, n_layers=4, dim_xyz=63, dim_dir=27, seed=0) -> Dict[str, Tensor]:
    """Random-init FlexibleNeRFModel-shaped weights (models.py:186-231 default graph)."""
    g = torch.Generator().manual_seed(seed)
    sd: Dict[str, Tensor] = {}

    def lin(name, fin, fout):
        bound = 1.0 / math.sqrt(fin)
        sd[name + ".weight"] = (torch.rand(fout, fin, generator=g) * 2 - 1) * bound
        sd[name + ".bias"] = (torch.rand(fout, generator=g) * 2 - 1) * bound

    lin("layer1", dim_xyz, hidden)
    for i in range(n_layers - 1):
        lin(f"layers_xyz.{i}", hidden, hidden)
    lin("layers_dir.0", dim_dir + hidden, hidden // 2)
    lin("fc_alpha", hidden, 1)
    lin("fc_rgb", hidden // 2, 3)
    lin("fc_feat", hidden, hidden)
    return sd


def run_mlp(kind: str, sd, pts: Tensor, viewdirs: Optional[Tensor], n_freq_xyz: int, n_freq_dir: int,
            netchunk: int = 65536) -> Tensor:
    """run_network: train_utils.py:8-25 == run_nerf.py:30-45.  pts [N,S,3], viewdirs [N,3]."""
    flat = pts.reshape(-1, 3)
    emb = posenc(flat, n_freq_xyz)
    dim_xyz = emb.shape[-1]
    if viewdirs is not None:
        dirs = viewdirs[:, None, :].expand(pts.shape).reshape(-1, 3)
        emb = torch.cat((emb, posenc(dirs, n_freq_dir)), dim=-1)
    outs = []
    for s in range(0, emb.shape[0], netchunk):
        blk = emb[s:s + netchunk]
        if kind == "flexible":
            outs.append(mlp_flexible(sd, blk, dim_xyz, viewdirs is not None))
        else:
            outs.append(mlp_tf8x256(sd, blk, dim_xyz, viewdirs is not None))
    raw = torch.cat(outs, dim=0)
    return raw.reshape(list(pts.shape[:-1]) + [raw.shape[-1]])


# --------------------------------------------------------------------------------------
# a9  compositing
# --------------------------------------------------------------------------------------
def composite(raw: Tensor, z: Tensor, rays_d: Tensor, noise: Optional[Tensor] = None,
              white_bkgd: bool = False):
    """volume_render_radiance_field, nerf-pytorch/nerf/volume_rendering_utils.py:6-53
    == raw2outputs nerf/run_nerf.py:93-163.  ``noise`` is the already-scaled additive
    sigma noise (None == raw_noise_std 0).  Returns rgb, disp, acc, weights, depth.
    """
    dists = torch.cat((z[..., 1:] - z[..., :-1],
                       torch.full_like(z[..., :1], 1e10)), dim=-1)
    dists = dists * rays_d[..., None, :].norm(p=2, dim=-1)
    rgb = torch.sigmoid(raw[..., :3])
    sig = raw[..., 3] if noise is None else raw[..., 3] + noise
    sigma = F.relu(sig)
    alpha = 1.0 - torch.exp(-sigma * dists)
    trans = torch.cumprod(1.0 - alpha + 1e-10, dim=-1)
    trans = torch.roll(trans, 1, -1)
    trans[..., 0] = 1.0
    weights = alpha * trans
    rgb_map = (weights[..., None] * rgb).sum(dim=-2)
    depth_map = (weights * z).sum(dim=-1)
    acc_map = weights.sum(dim=-1)
    disp_map = 1.0 / torch.max(1e-10 * torch.ones_like(depth_map), depth_map / acc_map)
    if white_bkgd:
        rgb_map = rgb_map + (1.0 - acc_map[..., None])
    return rgb_map, disp_map, acc_map, weights, depth_map


# --------------------------------------------------------------------------------------
# a10  hierarchical sampling
# --------------------------------------------------------------------------------------
def pdf_cdf(weights: Tensor) -> Tensor:
    """cdf with leading zero: nerf_helpers.py:265-271 == run_nerf_helpers.py:186-189."""
    w = weights + 1e-5
    pdf = w / torch.sum(w, dim=-1, keepdim=True)
    cdf = torch.cumsum(pdf, dim=-1)
    return torch.cat([torch.zeros_like(cdf[..., :1]), cdf], dim=-1)


def searchsorted_right(cdf: Tensor, u: Tensor) -> Tensor:
    """torchsearchsorted.searchsorted(cdf, u, side='right') (nerf_helpers.py:288).  The
    extension (github.com/aliutkus/torchsearchsorted, unpinned HEAD, requirements.txt:9) is
    not vendored; its documented semantics are numpy.searchsorted(side='right') per row,
    i.e. the count of cdf entries <= u."""
    return torch.searchsorted(cdf.contiguous(), u.contiguous(), right=True)


def invert_cdf(bins: Tensor, cdf: Tensor, u: Tensor) -> Tuple[Tensor, Tensor]:
    """nerf_helpers.py:286-300 == run_nerf_helpers.py:199-209.  Returns (samples, inds)."""
    inds = searchsorted_right(cdf, u)
    below = torch.clamp(inds - 1, min=0)
    above = torch.clamp(inds, max=cdf.shape[-1] - 1)
    cdf_b = torch.gather(cdf, -1, below)
    cdf_a = torch.gather(cdf, -1, above)
    bin_b = torch.gather(bins, -1, below)
    bin_a = torch.gather(bins, -1, above)
    denom = cdf_a - cdf_b
    denom = torch.where(denom < 1e-5, torch.ones_like(denom), denom)
    t = (u - cdf_b) / denom
    return bin_b + t * (bin_a - bin_b), inds


def sample_pdf(bins: Tensor, weights: Tensor, n_samples: int, det: bool = True,
               u: Optional[Tensor] = None) -> Tensor:
    """sample_pdf_2, nerf_helpers.py:260-302 (== TF sample_pdf run_nerf_helpers.py:183-211)."""
    cdf = pdf_cdf(weights)
    if u is None:
        if not det:
            raise ValueError("oracle needs explicit u for the stochastic branch")
        u = torch.linspace(0.0, 1.0, steps=n_samples, dtype=weights.dtype)
        u = u.expand(list(cdf.shape[:-1]) + [n_samples])
    samples, _ = invert_cdf(bins, cdf, u.contiguous())
    return samples


# --------------------------------------------------------------------------------------
# a5..a11 per-chunk renderer, a12 chunk driver
# --------------------------------------------------------------------------------------
def render_chunk(kind: str, sd_coarse, sd_fine, rays: Tensor, *, n_coarse: int, n_fine: int,
                 n_freq_xyz: int, n_freq_dir: int, lindisp=False, white_bkgd=False,
                 netchunk=65536, t_rand=None, u_fine=None, noise_c=None, noise_f=None) -> Dict[str, Tensor]:
    """predict_and_render_radiance (train_utils.py:28-127) == render_rays (run_nerf.py:48-243).

    ``rays`` is [N, 8] or [N, 11] = [o, d, near, far, (viewdir)].
    """
    ro, rd = rays[..., :3], rays[..., 3:6]
    near, far = rays[..., 6:7], rays[..., 7:8]
    vd = rays[..., 8:11] if rays.shape[-1] > 8 else None
    z = coarse_depths(near, far, n_coarse, lindisp, t_rand)
    raw = run_mlp(kind, sd_coarse, points_on_rays(ro, rd, z), vd, n_freq_xyz, n_freq_dir, netchunk)
    rgb_c, disp_c, acc_c, w, depth_c = composite(raw, z, rd, noise_c, white_bkgd)
    out = dict(rgb_coarse=rgb_c, disp_coarse=disp_c, acc_coarse=acc_c, depth_coarse=depth_c,
               weights_coarse=w, z_coarse=z, raw_coarse=raw)
    if n_fine > 0:
        z_mid = 0.5 * (z[..., 1:] + z[..., :-1])
        z_samp = sample_pdf(z_mid, w[..., 1:-1], n_fine, det=(u_fine is None), u=u_fine).detach()
        z_all, _ = torch.sort(torch.cat((z, z_samp), dim=-1), dim=-1)
        sd = sd_fine if sd_fine is not None else sd_coarse
        raw_f = run_mlp(kind, sd, points_on_rays(ro, rd, z_all), vd, n_freq_xyz, n_freq_dir, netchunk)
        rgb_f, disp_f, acc_f, w_f, depth_f = composite(raw_f, z_all, rd, noise_f, white_bkgd)
        out.update(rgb_fine=rgb_f, disp_fine=disp_f, acc_fine=acc_f, depth_fine=depth_f,
                   z_samples=z_samp, z_fine=z_all, raw_fine=raw_f, weights_fine=w_f,
                   z_std=z_samp.std(dim=-1, unbiased=False))
    return out


def pack_rays(rays_o: Tensor, rays_d: Tensor, near: float, far: float, *, H=None, W=None, focal=None,
              ndc: bool = False, use_viewdirs: bool = True) -> Tensor:
    """Ray packing: run_one_iter_of_nerf train_utils.py:141-168 == render run_nerf.py:297-325."""
    vd = unit_viewdirs(rays_d) if use_viewdirs else None
    if ndc:
        ro, rd = to_ndc(H, W, focal, 1.0, rays_o, rays_d)
    else:
        ro, rd = rays_o, rays_d
    ro = ro.reshape(-1, 3).to(torch.float32)
    rd = rd.reshape(-1, 3).to(torch.float32)
    nr = near * torch.ones_like(rd[..., :1])
    fr = far * torch.ones_like(rd[..., :1])
    rays = torch.cat((ro, rd, nr, fr), dim=-1)
    if use_viewdirs:
        rays = torch.cat((rays, vd), dim=-1)
    return rays


def render_rays_chunked(kind, sd_coarse, sd_fine, rays: Tensor, chunk: int, **kw) -> Dict[str, Tensor]:
    """batchify_rays run_nerf.py:246-257 == train_utils.py:170-185."""
    acc: Dict[str, List[Tensor]] = {}
    for s in range(0, rays.shape[0], chunk):
        r = render_chunk(kind, sd_coarse, sd_fine, rays[s:s + chunk], **kw)
        for k, v in r.items():
            acc.setdefault(k, []).append(v)
    return {k: torch.cat(v, dim=0) for k, v in acc.items()}


def render_image(kind, sd_coarse, sd_fine, H, W, focal, c2w, *, near, far, ndc=False,
                 use_viewdirs=True, chunk=32768, K=None, **kw) -> Dict[str, Tensor]:
    """render(H, W, focal, c2w=...) run_nerf.py:260-336 / eval_nerf.py:157-166."""
    ro, rd = pixel_rays(H, W, focal, c2w, K)
    rays = pack_rays(ro, rd, near, far, H=H, W=W, focal=focal, ndc=ndc, use_viewdirs=use_viewdirs)
    out = render_rays_chunked(kind, sd_coarse, sd_fine, rays, chunk, **kw)
    return {k: v.reshape([H, W] + list(v.shape[1:])) for k, v in out.items()}


# --------------------------------------------------------------------------------------
# a13  loss + backward (autograd on CPU)
# --------------------------------------------------------------------------------------
def train_step_grads(kind, sd_coarse, sd_fine, rays: Tensor, target: Tensor, **kw):
    """loss = mse(rgb_coarse, tgt) + mse(rgb_fine, tgt)  (train_nerf.py:271-286 ==
    run_nerf.py:815-826); returns (loss, grads_coarse, grads_fine) with grads keyed like
    the state dicts.  No gradient flows through sample_pdf (train_utils.py:103)."""
    pc = {k: v.detach().clone().requires_grad_(True) for k, v in sd_coarse.items()}
    pf = {k: v.detach().clone().requires_grad_(True) for k, v in sd_fine.items()} if sd_fine is not None else None
    out = render_chunk(kind, pc, pf, rays, **kw)
    loss = F.mse_loss(out["rgb_coarse"], target)
    if "rgb_fine" in out:
        loss = loss + F.mse_loss(out["rgb_fine"], target)
    loss.backward()
    gc = {k: (v.grad if v.grad is not None else torch.zeros_like(v)) for k, v in pc.items()}
    gf = {k: (v.grad if v.grad is not None else torch.zeros_like(v)) for k, v in pf.items()} if pf is not None else None
    return loss.detach(), gc, gf, out


# --------------------------------------------------------------------------------------
# camera paths used by the bench configs (host side, data-free)
# --------------------------------------------------------------------------------------
def pose_spherical(theta_deg: float, phi_deg: float, radius: float) -> Tensor:
    """nerf-pytorch/nerf/load_blender.py:9-37 (trans_t, rot_phi, rot_theta, pose_spherical)."""
    t = torch.eye(4)
    t[2, 3] = radius
    phi = phi_deg / 180.0 * math.pi
    rp = torch.tensor([[1, 0, 0, 0],
                       [0, math.cos(phi), -math.sin(phi), 0],
                       [0, math.sin(phi), math.cos(phi), 0],
                       [0, 0, 0, 1]], dtype=torch.float32)
    th = theta_deg / 180.0 * math.pi
    rt = torch.tensor([[math.cos(th), 0, -math.sin(th), 0],
                       [0, 1, 0, 0],
                       [math.sin(th), 0, math.cos(th), 0],
                       [0, 0, 0, 1]], dtype=torch.float32)
    c2w = rt @ (rp @ t)
    flip = torch.tensor([[-1, 0, 0, 0], [0, 0, 1, 0], [0, 1, 0, 0], [0, 0, 0, 1]], dtype=torch.float32)
    return flip @ c2w
