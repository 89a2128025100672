"""torch.Tensor-level wrappers of the C-ABI stage kernels (include/nerf_b200.h).

Every function here takes CUDA float32 tensors, allocates its outputs with torch.empty, and enqueues
ONE library call on torch's current stream.  There is no CPU implementation behind any of them: a CPU
tensor raises.  Reference citations are on the C declarations in include/nerf_b200.h.
"""
import ctypes as C

import torch

from . import _lib
from ._lib import check, lib

FP32, BF16 = 0, 1  # nb_precision


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _f32(t: torch.Tensor, name: str = "tensor") -> torch.Tensor:
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise _lib.NerfB200Error("%s must be a CUDA tensor: nerf_kinematics_b200 has no CPU fallback" % name)
    if t.dtype != torch.float32:
        t = t.to(torch.float32)
    return t.contiguous()


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None and t.numel() > 0 else C.c_void_p(0)


def _optptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


# ------------------------------------------------------------------------------------------ a1
def get_rays(H: int, W: int, K, c2w, row0: int = 0, row1: int = None, device=None):
    """Rows [row0,row1) of the pixel ray bundle.  K: 3x3 intrinsics (or a scalar focal length)."""
    _lib.require_cuda()
    if row1 is None:
        row1 = H
    fx, fy, cx, cy = intrinsics(H, W, K)
    on_dev = isinstance(c2w, torch.Tensor) and c2w.is_cuda
    if device is None:
        device = c2w.device if on_dev else torch.device("cuda")
    with torch.cuda.device(device):
        ro = torch.empty((row1 - row0, W, 3), dtype=torch.float32, device=device)
        rd = torch.empty_like(ro)
        if on_dev:   # pose already on the device: the kernel reads it there, no device-to-host synchronisation
            pose = c2w.detach()[:3, :4].to(device=device, dtype=torch.float32).contiguous()
            check(lib().nb_get_rays_dev(H, W, fx, fy, cx, cy, _ptr(pose), row0, row1, _ptr(ro), _ptr(rd), _stream()),
                  "nb_get_rays_dev")
        else:
            arr = (C.c_float * 12)(*_host_pose(c2w))
            check(lib().nb_get_rays(H, W, fx, fy, cx, cy, arr, row0, row1, _ptr(ro), _ptr(rd), _stream()), "nb_get_rays")
    return ro, rd


def _host_pose(c2w):
    """12 floats (3x4 row-major) from a host pose (numpy array, nested list or CPU tensor)."""
    import numpy as np
    a = c2w.detach().numpy() if isinstance(c2w, torch.Tensor) else np.asarray(c2w)
    return [float(v) for v in np.asarray(a, dtype=np.float32)[:3, :4].reshape(-1)]


def get_rays_at(H: int, W: int, K, c2w, pixel_index, device=None):
    """Rays of the selected pixels only: pixel_index int64 [n] = row * W + col.  Bit-identical to
    `get_rays(H, W, K, c2w)` flattened with reshape(-1, 3) and indexed (train_nerf.py:240-253)."""
    _lib.require_cuda()
    fx, fy, cx, cy = intrinsics(H, W, K)
    on_dev = isinstance(c2w, torch.Tensor) and c2w.is_cuda
    if device is None:
        device = pixel_index.device if isinstance(pixel_index, torch.Tensor) and pixel_index.is_cuda else torch.device("cuda")
    idx = torch.as_tensor(pixel_index, dtype=torch.int64).to(device).contiguous()
    with torch.cuda.device(device):
        ro = torch.empty((idx.numel(), 3), dtype=torch.float32, device=device)
        rd = torch.empty_like(ro)
        if on_dev:
            pose = c2w.detach()[:3, :4].to(device=device, dtype=torch.float32).contiguous()
            check(lib().nb_get_rays_at_dev(H, W, fx, fy, cx, cy, _ptr(pose), _ptr(idx), idx.numel(), _ptr(ro), _ptr(rd),
                                           _stream()), "nb_get_rays_at_dev")
        else:
            arr = (C.c_float * 12)(*_host_pose(c2w))
            check(lib().nb_get_rays_at(H, W, fx, fy, cx, cy, arr, _ptr(idx), idx.numel(), _ptr(ro), _ptr(rd), _stream()),
                  "nb_get_rays_at")
    return ro, rd


def to8b_psnr(x, gt=None, want_u8=True):
    """render_path's output stage on the device: uint8 image (255 * clip(x, 0, 1), truncated like
    numpy astype) and, against `gt`, the PSNR the reference prints (run_nerf.py:364-366).  Returns (u8 or None, psnr or None)."""
    _lib.require_cuda()
    x = _f32(x, "x")
    out8 = torch.empty(x.shape, dtype=torch.uint8, device=x.device) if want_u8 else None
    acc = None
    if gt is not None:
        gt = _f32(gt.to(x.device), "gt")
        if gt.shape != x.shape:
            raise ValueError("to8b_psnr: gt shape %s != image shape %s" % (tuple(gt.shape), tuple(x.shape)))
        acc = torch.zeros((1,), dtype=torch.float64, device=x.device)
    with torch.cuda.device(x.device):
        check(lib().nb_to8b(_ptr(x), _optptr(gt), x.numel(), _optptr(out8), _optptr(acc), _stream()), "nb_to8b")
    psnr = None
    if acc is not None:
        psnr = -10.0 * torch.log10(acc / x.numel()).to(torch.float32)
    return out8, psnr


def intrinsics(H, W, K):
    """(fx, fy, cx, cy) from a 3x3 K or a scalar focal (then cx = W/2, cy = H/2 like the reference)."""
    if isinstance(K, torch.Tensor) and K.numel() == 1:
        K = float(K)
    if isinstance(K, (int, float)):
        return float(K), float(K), W * 0.5, H * 0.5
    try:
        import numpy as np
        if isinstance(K, np.ndarray) and K.size == 1:
            f = float(K)
            return f, f, W * 0.5, H * 0.5
    except ImportError:  # pragma: no cover
        pass
    return float(K[0][0]), float(K[1][1]), float(K[0][2]), float(K[1][2])


# ------------------------------------------------------------------------------------------ a3
def ndc_rays(H, W, focal, near, rays_o, rays_d):
    ro, rd = _f32(rays_o, "rays_o"), _f32(rays_d, "rays_d")
    oo, od = torch.empty_like(ro), torch.empty_like(rd)
    n = ro.numel() // 3
    with torch.cuda.device(ro.device):
        check(lib().nb_ndc_rays(int(H), int(W), float(focal), float(near), _ptr(ro), _ptr(rd), n, _ptr(oo), _ptr(od),
                                _stream()), "nb_ndc_rays")
    return oo, od


def make_ray_batch(rays_o, rays_d, near, far, *, use_viewdirs, ndc=False, H=0, W=0, focal=1.0, viewdir_src=None,
                   near_arr=None, far_arr=None):
    """[n, 8|11] packed rays: [o, d, near, far, (unit viewdir from the PRE-NDC direction)]."""
    ro, rd = _f32(rays_o, "rays_o").reshape(-1, 3), _f32(rays_d, "rays_d").reshape(-1, 3)
    n = ro.shape[0]
    vs = _f32(viewdir_src).reshape(-1, 3) if viewdir_src is not None else None
    na = _f32(near_arr).reshape(-1) if near_arr is not None else None
    fa = _f32(far_arr).reshape(-1) if far_arr is not None else None
    out = torch.empty((n, 11 if use_viewdirs else 8), dtype=torch.float32, device=ro.device)
    with torch.cuda.device(ro.device):
        check(lib().nb_make_ray_batch(_ptr(ro), _ptr(rd), _optptr(vs), n, int(bool(use_viewdirs)), int(bool(ndc)),
                                      int(H), int(W), float(focal), float(near), float(far), _optptr(na), _optptr(fa),
                                      _ptr(out), _stream()), "nb_make_ray_batch")
    return out


# ------------------------------------------------------------------------------------------ a5
def sample_coarse(rays, n_samples, lindisp=False, t_rand=None):
    rays = _f32(rays, "rays")
    n = rays.shape[0]
    tr = _f32(t_rand) if t_rand is not None else None
    z = torch.empty((n, n_samples), dtype=torch.float32, device=rays.device)
    with torch.cuda.device(rays.device):
        check(lib().nb_sample_coarse(_ptr(rays), rays.shape[1], n, n_samples, int(bool(lindisp)), _optptr(tr), _ptr(z),
                                     _stream()), "nb_sample_coarse")
    return z


# ------------------------------------------------------------------------------------------ a6
def posenc(x, n_freqs, include_input=True):
    x = _f32(x, "x")
    shp = x.shape[:-1]
    x2 = x.reshape(-1, 3)
    Cc = (3 if include_input else 0) + 6 * n_freqs
    out = torch.empty((x2.shape[0], Cc), dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        check(lib().nb_posenc(_ptr(x2), x2.shape[0], n_freqs, int(bool(include_input)), _ptr(out), Cc, _stream()),
              "nb_posenc")
    return out.reshape(*shp, Cc)


# ------------------------------------------------------------------------------------------ a9
def composite_forward(raw, z_vals, rays, noise=None, white_bkgd=False, want_weights=True):
    raw, z, rays = _f32(raw, "raw"), _f32(z_vals, "z_vals"), _f32(rays, "rays")
    n, S = z.shape
    dev = raw.device
    rgb = torch.empty((n, 3), dtype=torch.float32, device=dev)
    disp = torch.empty((n,), dtype=torch.float32, device=dev)
    acc = torch.empty_like(disp)
    depth = torch.empty_like(disp)
    w = torch.empty((n, S), dtype=torch.float32, device=dev) if want_weights else None
    nz = _f32(noise) if noise is not None else None
    with torch.cuda.device(dev):
        check(lib().nb_composite(_ptr(raw), _ptr(z), _ptr(rays), rays.shape[1], n, S, _optptr(nz), int(bool(white_bkgd)),
                                 _ptr(rgb), _ptr(disp), _ptr(acc), _ptr(depth), _optptr(w), _stream()), "nb_composite")
    return rgb, disp, acc, w, depth


def composite_backward(raw, z_vals, rays, noise, white_bkgd, d_rgb=None, d_disp=None, d_acc=None, d_depth=None):
    raw, z, rays = _f32(raw), _f32(z_vals), _f32(rays)
    n, S = z.shape
    d_raw = torch.empty((n, S, 4), dtype=torch.float32, device=raw.device)
    gs = [(_f32(g) if g is not None else None) for g in (d_rgb, d_disp, d_acc, d_depth)]
    nz = _f32(noise) if noise is not None else None
    with torch.cuda.device(raw.device):
        check(lib().nb_composite_backward(_ptr(raw), _ptr(z), _ptr(rays), rays.shape[1], n, S, _optptr(nz),
                                          int(bool(white_bkgd)), _optptr(gs[0]), _optptr(gs[1]), _optptr(gs[2]),
                                          _optptr(gs[3]), _ptr(d_raw), _stream()), "nb_composite_backward")
    return d_raw


class _Composite(torch.autograd.Function):
    @staticmethod
    def forward(ctx, raw, z_vals, rays, noise, white_bkgd):
        rgb, disp, acc, w, depth = composite_forward(raw, z_vals, rays, noise, white_bkgd)
        ctx.save_for_backward(raw, z_vals, rays, noise if noise is not None else torch.empty(0, device=raw.device))
        ctx.white = bool(white_bkgd)
        ctx.has_noise = noise is not None
        ctx.mark_non_differentiable(w)
        return rgb, disp, acc, w, depth

    @staticmethod
    def backward(ctx, g_rgb, g_disp, g_acc, g_w, g_depth):
        raw, z, rays, noise = ctx.saved_tensors
        d_raw = composite_backward(raw, z, rays, noise if ctx.has_noise else None, ctx.white, g_rgb, g_disp, g_acc,
                                   g_depth)
        return d_raw.reshape(raw.shape), None, None, None, None


def composite(raw, z_vals, rays, noise=None, white_bkgd=False):
    """raw2outputs with autograd w.r.t. raw.  Returns (rgb, disp, acc, weights, depth); weights carry no grad."""
    return _Composite.apply(raw, z_vals, rays, noise, white_bkgd)


# ------------------------------------------------------------------------------------------ a10 / a11
def searchsorted(a, v, side="left"):
    a, v = _f32(a, "a"), _f32(v, "v")
    a2 = a.reshape(-1, a.shape[-1])
    v2 = v.reshape(-1, v.shape[-1])
    out = torch.empty(v2.shape, dtype=torch.int64, device=v.device)
    with torch.cuda.device(v.device):
        check(lib().nb_searchsorted(_ptr(a2), a2.shape[0], a2.shape[1], _ptr(v2), v2.shape[0], v2.shape[1],
                                    int(side == "right"), _ptr(out), _stream()), "nb_searchsorted")
    return out.reshape(v.shape)


def sample_pdf(bins, weights, n_samples, det=False, u=None, return_cdf=False):
    bins, weights = _f32(bins, "bins"), _f32(weights, "weights")
    n, nb = bins.shape
    if weights.shape[1] != nb - 1:
        raise ValueError("sample_pdf: weights must have one column fewer than bins")
    if u is None and not det:
        u = torch.rand((n, n_samples), dtype=torch.float32, device=bins.device)
    uu = _f32(u) if u is not None else None
    out = torch.empty((n, n_samples), dtype=torch.float32, device=bins.device)
    cdf = torch.empty((n, nb), dtype=torch.float32, device=bins.device) if return_cdf else None
    with torch.cuda.device(bins.device):
        check(lib().nb_sample_pdf(_ptr(bins), nb, _ptr(weights), nb - 1, n, nb, _optptr(uu), n_samples, _ptr(out),
                                  _optptr(cdf), _stream()), "nb_sample_pdf")
    return (out, cdf) if return_cdf else out


def sample_fine(z_vals, weights, n_importance, u=None, want_samples=False, want_std=False):
    z, w = _f32(z_vals, "z_vals"), _f32(weights, "weights")
    n, S = z.shape
    uu = _f32(u) if u is not None else None
    dev = z.device
    zs = torch.empty((n, n_importance), dtype=torch.float32, device=dev) if want_samples else None
    zf = torch.empty((n, S + n_importance), dtype=torch.float32, device=dev)
    sd = torch.empty((n,), dtype=torch.float32, device=dev) if want_std else None
    with torch.cuda.device(dev):
        check(lib().nb_sample_fine(_ptr(z), _ptr(w), n, S, _optptr(uu), n_importance, _optptr(zs), _ptr(zf),
                                   _optptr(sd), _stream()), "nb_sample_fine")
    return zf, zs, sd


def merge_sorted(a, b):
    a, b = _f32(a), _f32(b)
    n = a.shape[0]
    out = torch.empty((n, a.shape[1] + b.shape[1]), dtype=torch.float32, device=a.device)
    with torch.cuda.device(a.device):
        check(lib().nb_merge_sorted(_ptr(a), a.shape[1], _ptr(b), b.shape[1], n, _ptr(out), _stream()),
              "nb_merge_sorted")
    return out


# ------------------------------------------------------------------------------------------ a13
def mse_loss(pred, target, grad_scale=1.0, want_grad=True):
    pred, target = _f32(pred), _f32(target)
    loss = torch.zeros((1,), dtype=torch.float32, device=pred.device)
    d = torch.empty_like(pred) if want_grad else None
    with torch.cuda.device(pred.device):
        check(lib().nb_mse_loss(_ptr(pred), _ptr(target), pred.numel(), float(grad_scale), _ptr(loss), _optptr(d),
                                _stream()), "nb_mse_loss")
    return loss[0], d


def launch_count() -> int:
    return int(lib().nb_launch_count())


def launch_count_reset():
    lib().nb_launch_count_reset()
