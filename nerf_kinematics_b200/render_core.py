"""One chunk of rays through the whole hot path, as ONE native call with autograd.

Python-side twin of render_rays (nerf/run_nerf.py:48-243) / predict_and_render_radiance
(nerf-pytorch/nerf/train_utils.py:28-127): it only allocates outputs + workspace and calls
nb_render_rays; during training the call is wrapped in a torch.autograd.Function whose backward is
nb_render_rays_backward, so the reference's `loss.backward(); optimizer.step()` loop works unchanged.
"""
import ctypes as C

import torch

from . import _lib
from ._lib import RenderCfg, RenderGrads, RenderOut, RenderRand, check, lib
from .models import KernelMLP, precision_id
from .ops import _f32, _optptr, _stream

OUT_KEYS = ("rgb", "disp", "acc", "depth", "rgb0", "disp0", "acc0", "depth0", "z_std", "raw", "z_vals", "weights")


def _alloc_outputs(n, S, I, dev, want_raw, want_z, want_w):
    f = lambda *shape: torch.empty(shape, dtype=torch.float32, device=dev)
    T = S + I
    out = {"rgb": f(n, 3), "disp": f(n), "acc": f(n), "depth": f(n)}
    if I > 0:
        out.update(rgb0=f(n, 3), disp0=f(n), acc0=f(n), depth0=f(n), z_std=f(n))
    if want_raw:
        out["raw"] = f(n, T, 4)
    if want_z:
        out["z_vals"] = f(n, T)
    if want_w:
        out["weights"] = f(n, T)
    return out


def _random_inputs(n, S, I, dev, perturb, raw_noise_std, rnd):
    """The reference draws these from the framework RNG (run_nerf.py:197, helpers:195, run_nerf.py:131);
    tests may inject them through `rnd` to get a defined comparison."""
    rnd = dict(rnd or {})
    if perturb and "t_rand" not in rnd:
        rnd["t_rand"] = torch.rand((n, S), dtype=torch.float32, device=dev)
    if perturb and I > 0 and "u_fine" not in rnd:
        rnd["u_fine"] = torch.rand((n, I), dtype=torch.float32, device=dev)
    if raw_noise_std and raw_noise_std > 0.0:
        if "noise_c" not in rnd:
            rnd["noise_c"] = torch.randn((n, S), dtype=torch.float32, device=dev) * float(raw_noise_std)
        if I > 0 and "noise_f" not in rnd:
            rnd["noise_f"] = torch.randn((n, S + I), dtype=torch.float32, device=dev) * float(raw_noise_std)
    return {k: (_f32(v) if v is not None else None) for k, v in rnd.items()}


def _cfg(S, I, lindisp, white_bkgd, prec, save):
    return RenderCfg(int(S), int(I), int(bool(lindisp)), int(bool(white_bkgd)), int(prec), int(bool(save)))


def _rand_struct(rnd):
    return RenderRand(_optptr(rnd.get("t_rand")), _optptr(rnd.get("u_fine")), _optptr(rnd.get("noise_c")),
                      _optptr(rnd.get("noise_f")))


def _forward_native(rays, coarse, fine, S, I, lindisp, white_bkgd, prec, save, rnd, want_raw, want_z, want_w):
    n = rays.shape[0]
    dev = rays.device
    hc, _ = coarse.prepare(prec)
    hf = fine.prepare(prec)[0] if (fine is not None and I > 0) else None
    out = _alloc_outputs(n, S, I, dev, want_raw, want_z, want_w)
    cfg = _cfg(S, I, lindisp, white_bkgd, prec, save)
    with torch.cuda.device(dev):
        nbytes = lib().nb_render_workspace_bytes(hc, hf, rays.shape[1], n, C.byref(cfg))
        ws = torch.empty((max(int(nbytes), 1),), dtype=torch.uint8, device=dev)
        o = RenderOut(*[_optptr(out.get(k)) for k in OUT_KEYS])
        rs = _rand_struct(rnd)
        check(lib().nb_render_rays(hc, hf, C.c_void_p(rays.data_ptr()), rays.shape[1], n, C.byref(cfg), C.byref(rs),
                                   C.byref(o), C.c_void_p(ws.data_ptr()), nbytes, _stream()), "nb_render_rays")
    return out, ws, nbytes


_GRAD_KEYS = ("rgb", "disp", "acc", "depth", "rgb0", "disp0", "acc0", "depth0")


class _RenderChunkFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, rays, opts, coarse, fine, *params):
        S, I = opts["S"], opts["I"]
        out, ws, nbytes = _forward_native(rays, coarse, fine, S, I, opts["lindisp"], opts["white_bkgd"], opts["prec"],
                                          True, opts["rnd"], opts["want_raw"], opts["want_z"], opts["want_w"])
        ctx.rays, ctx.opts, ctx.coarse, ctx.fine, ctx.ws, ctx.nbytes = rays, opts, coarse, fine, ws, nbytes
        ctx.keys = [k for k in OUT_KEYS if k in out]
        nondiff = [out[k] for k in ctx.keys if k not in _GRAD_KEYS]
        ctx.mark_non_differentiable(*nondiff)
        return tuple(out[k] for k in ctx.keys)

    @staticmethod
    def backward(ctx, *gouts):
        opts, coarse, fine, rays = ctx.opts, ctx.coarse, ctx.fine, ctx.rays
        g = {k: (_f32(go) if go is not None else None) for k, go in zip(ctx.keys, gouts) if k in _GRAD_KEYS}
        S, I = opts["S"], opts["I"]
        if I == 0:  # the single pass wrote rgb/disp/acc/depth
            pass
        dev = rays.device
        hc = coarse._handle
        use_fine = fine is not None and I > 0
        hf = fine._handle if use_fine else None
        gc = torch.zeros((coarse.n_params,), dtype=torch.float32, device=dev)
        gf = torch.zeros((fine.n_params,), dtype=torch.float32, device=dev) if use_fine else None
        cfg = _cfg(S, I, opts["lindisp"], opts["white_bkgd"], opts["prec"], True)
        gs = RenderGrads(*[_optptr(g.get(k)) for k in _GRAD_KEYS])
        rs = _rand_struct(opts["rnd"])
        with torch.cuda.device(dev):
            check(lib().nb_render_rays_backward(hc, hf, C.c_void_p(rays.data_ptr()), rays.shape[1], rays.shape[0],
                                                C.byref(cfg), C.byref(rs), C.byref(gs), C.c_void_p(ctx.ws.data_ptr()),
                                                ctx.nbytes, C.c_void_p(gc.data_ptr()), _optptr(gf), _stream()),
                  "nb_render_rays_backward")
        ctx.ws = None
        grads = [v for _, v in coarse.grad_views(gc)]
        if use_fine:
            grads += [v for _, v in fine.grad_views(gf)]
        return (None, None, None, None) + tuple(grads)


def render_chunk(rays, coarse: KernelMLP, fine, *, N_samples, N_importance=0, lindisp=False, perturb=0.0,
                 white_bkgd=False, raw_noise_std=0.0, precision=None, rnd=None, want_raw=False, want_z=False,
                 want_weights=False):
    """Render one chunk of packed rays [n, 8|11].  Returns a dict with the keys of OUT_KEYS that apply."""
    _lib.require_cuda()
    rays = _f32(rays, "rays")
    if rays.dim() != 2 or rays.shape[1] not in (8, 11):
        raise ValueError("rays must be [n, 8] or [n, 11], got %s" % (tuple(rays.shape),))
    n = rays.shape[0]
    S, I = int(N_samples), int(N_importance)
    prec = precision_id(coarse.precision if precision is None else precision)
    rnd = _random_inputs(n, S, I, rays.device, perturb, raw_noise_std, rnd)
    if fine is coarse:
        fine = None
    params = coarse.ordered_parameters() + (fine.ordered_parameters() if (fine is not None and I > 0) else [])
    need_grad = torch.is_grad_enabled() and any(p.requires_grad for p in params)
    if n == 0:
        return _alloc_outputs(0, S, I, rays.device, want_raw, want_z, want_weights)
    if not need_grad:
        out, _, _ = _forward_native(rays, coarse, fine, S, I, lindisp, white_bkgd, prec, False, rnd, want_raw, want_z,
                                    want_weights)
        return out
    coarse.prepare(prec)
    if fine is not None:
        fine.prepare(prec)
    opts = dict(S=S, I=I, lindisp=lindisp, white_bkgd=white_bkgd, prec=prec, rnd=rnd, want_raw=want_raw, want_z=want_z,
                want_w=want_weights)
    keys = [k for k in OUT_KEYS if k in _alloc_keys(S, I, want_raw, want_z, want_weights)]
    vals = _RenderChunkFn.apply(rays, opts, coarse, fine, *params)
    return dict(zip(keys, vals))


def _alloc_keys(S, I, want_raw, want_z, want_w):
    keys = {"rgb", "disp", "acc", "depth"}
    if I > 0:
        keys |= {"rgb0", "disp0", "acc0", "depth0", "z_std"}
    if want_raw:
        keys.add("raw")
    if want_z:
        keys.add("z_vals")
    if want_w:
        keys.add("weights")
    return keys
