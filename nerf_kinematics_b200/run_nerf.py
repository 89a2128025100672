"""The north_star API surface: create_nerf / render / render_path / render_rays / batchify_rays / run_network.

Semantics follow nerf/run_nerf.py line for line (citations per function); the arithmetic is the CUDA
library behind include/nerf_b200.h.  Differences a caller can observe, all deliberate:
  * tensors are torch CUDA tensors instead of TF eager tensors (render_path still returns numpy);
  * the `focal` position also accepts a 3x3 intrinsics matrix K (yenchenlin-style) -- a scalar gives
    K = [[f,0,W/2],[0,f,H/2],[0,0,1]], bit-identical to get_rays (run_nerf_helpers.py:123-130);
  * `chunk` bounds live scratch only; `netchunk` is accepted and ignored (the network kernel tiles
    internally, "Does not affect final results", run_nerf.py:271-272);
  * `precision=` ('fp32' | 'bf16') may be passed through render kwargs to pick the kernel family;
  * NaN guard: TF's check_numerics (run_nerf.py:240-241) is opt-in via check_numerics=True.
"""
import os
import time

import numpy as np
import torch

from . import _lib, ops
from .frame import FrameRenderer
from .models import NeRF
from .nerf_helpers import get_embedder, to8b
from .render_core import render_chunk


def run_network(inputs, viewdirs, fn, embed_fn=None, embeddirs_fn=None, netchunk=1024 * 64):
    """run_nerf.py:30-45.  The embedding is fused into the network kernel, so embed_fn / embeddirs_fn /
    netchunk are accepted for signature compatibility only."""
    return fn.query(inputs, viewdirs)


def render_rays(ray_batch, network_fn, network_query_fn, N_samples, retraw=False, lindisp=False, perturb=0.0,
                N_importance=0, network_fine=None, white_bkgd=False, raw_noise_std=0.0, verbose=False,
                precision=None, check_numerics=False, rnd=None):
    """run_nerf.py:48-243 -> dict(rgb_map, disp_map, acc_map[, raw][, rgb0, disp0, acc0, z_std])."""
    out = render_chunk(ray_batch, network_fn, network_fine, N_samples=N_samples, N_importance=N_importance,
                       lindisp=lindisp, perturb=perturb, white_bkgd=white_bkgd, raw_noise_std=raw_noise_std,
                       precision=precision, rnd=rnd, want_raw=retraw)
    ret = {"rgb_map": out["rgb"], "disp_map": out["disp"], "acc_map": out["acc"]}
    if retraw:
        ret["raw"] = out["raw"]
    if N_importance > 0:
        ret["rgb0"], ret["disp0"], ret["acc0"], ret["z_std"] = out["rgb0"], out["disp0"], out["acc0"], out["z_std"]
    if check_numerics:  # TF check_numerics (run_nerf.py:240-241): one device flag for all outputs, read back once
        flag = torch.zeros(1, dtype=torch.int32, device=ray_batch.device)
        keys = list(ret)
        with torch.cuda.device(ray_batch.device):
            for bit, k in enumerate(keys):
                v = ret[k].contiguous()
                _lib.check(_lib.lib().nb_check_finite(ops._ptr(v), v.numel(), ops._ptr(flag), 1 << bit, ops._stream()),
                           "nb_check_finite")
        bits = int(flag.item())
        if bits:
            bad = [k for b, k in enumerate(keys) if bits & (1 << b)]
            raise FloatingPointError("output %s : Tensor had NaN or Inf values" % ", ".join(bad))
    return ret


def batchify_rays(rays_flat, chunk=1024 * 32, **kwargs):
    """run_nerf.py:246-257."""
    all_ret = {}
    for i in range(0, rays_flat.shape[0], chunk):
        ret = render_rays(rays_flat[i:i + chunk], **kwargs)
        for k, v in ret.items():
            all_ret.setdefault(k, []).append(v)
    return {k: (torch.cat(v, 0) if len(v) > 1 else v[0]) for k, v in all_ret.items()}


def _focal_scalar(K):
    fx, fy, _, _ = ops.intrinsics(1, 1, K)
    return fx


_FRAME_CACHE_MAX = 4


def _frame_renderer(H, W, focal, chunk, ndc, near, far, use_viewdirs, kwargs, rows=None):
    """A cached FrameRenderer when this render() call is the plain per-pose inference form (pose given, no gradient
    wanted, no random inputs, no retraw): the whole frame is then one CUDA-graph replay.  None otherwise."""
    fn = kwargs.get("network_fn")
    if fn is None or not hasattr(fn, "prepare"):
        return None
    if kwargs.get("retraw") or kwargs.get("rnd") or kwargs.get("perturb") or (kwargs.get("raw_noise_std") or 0.0) > 0.0:
        return None
    if isinstance(near, torch.Tensor) or isinstance(far, torch.Tensor):
        return None
    fine = kwargs.get("network_fine")
    nets = [fn] + ([fine] if fine is not None else [])
    if torch.is_grad_enabled() and any(p.requires_grad for m in nets for p in m.parameters()):
        return None
    fx, fy, cx, cy = ops.intrinsics(H, W, focal)
    r0, r1 = rows if rows is not None else (0, H)
    key = (H, W, fx, fy, cx, cy, int(chunk), bool(ndc), float(near), float(far), bool(use_viewdirs), id(fine),
           int(kwargs.get("N_samples")), int(kwargs.get("N_importance", 0)), bool(kwargs.get("lindisp", False)),
           bool(kwargs.get("white_bkgd", False)), kwargs.get("precision") or fn.precision,
           bool(kwargs.get("check_numerics", False)), r0, r1)
    cache = fn.__dict__.setdefault("_frame_cache", {})
    fr = cache.get(key)
    if fr is None:
        if len(cache) >= _FRAME_CACHE_MAX:
            cache.pop(next(iter(cache)))
        fr = FrameRenderer(H, W, focal, fn, fine, chunk=chunk, row0=r0, row1=r1, near=near, far=far, ndc=ndc,
                           use_viewdirs=use_viewdirs, N_samples=kwargs["N_samples"],
                           N_importance=kwargs.get("N_importance", 0), lindisp=kwargs.get("lindisp", False),
                           white_bkgd=kwargs.get("white_bkgd", False), precision=kwargs.get("precision"),
                           check_numerics=kwargs.get("check_numerics", False))
        cache[key] = fr
    return fr


def render(H, W, focal, chunk=1024 * 32, rays=None, c2w=None, ndc=True, near=0.0, far=1.0, use_viewdirs=False,
           c2w_staticcam=None, _static_outputs=False, _rows=None, **kwargs):
    """run_nerf.py:260-336 -> [rgb_map, disp_map, acc_map, extras].  `focal` may be a 3x3 K.
    With a pose and inference options the frame runs as one captured CUDA graph (frame.FrameRenderer); results are
    identical to the chunk-by-chunk path (same kernels, same order per ray)."""
    H, W = int(H), int(W)
    if c2w is not None and c2w_staticcam is None:
        fr = _frame_renderer(H, W, focal, chunk, ndc, near, far, use_viewdirs, kwargs, rows=_rows)
        if fr is not None:
            out = fr.render(c2w, clone=not _static_outputs)
            extras = {k: out[k] for k in ("rgb0", "disp0", "acc0", "z_std") if k in out}
            return [out["rgb"], out["disp"], out["acc"], extras]
    if c2w is not None:
        r0, r1 = _rows if _rows is not None else (0, H)   # a rank's row block (multi-GPU render: SURVEY 8e)
        rays_o, rays_d = ops.get_rays(H, W, focal, c2w, row0=r0, row1=r1)
    else:
        rays_o, rays_d = rays[0], rays[1]
    viewdir_src = None
    if use_viewdirs and c2w_staticcam is not None:
        viewdir_src = rays_d
        rays_o, rays_d = ops.get_rays(H, W, focal, c2w_staticcam)
    sh = rays_d.shape
    near_arr = near if isinstance(near, torch.Tensor) and near.numel() > 1 else None
    far_arr = far if isinstance(far, torch.Tensor) and far.numel() > 1 else None
    packed = ops.make_ray_batch(rays_o, rays_d, 0.0 if near_arr is not None else float(near),
                                0.0 if far_arr is not None else float(far), use_viewdirs=use_viewdirs, ndc=ndc, H=H, W=W,
                                focal=_focal_scalar(focal), viewdir_src=viewdir_src, near_arr=near_arr, far_arr=far_arr)
    all_ret = batchify_rays(packed, chunk, **kwargs)
    for k in all_ret:
        all_ret[k] = all_ret[k].reshape(list(sh[:-1]) + list(all_ret[k].shape[1:]))
    k_extract = ["rgb_map", "disp_map", "acc_map"]
    ret_list = [all_ret[k] for k in k_extract]
    ret_dict = {k: v for k, v in all_ret.items() if k not in k_extract}
    return ret_list + [ret_dict]


def render_path(render_poses, hwf, chunk, render_kwargs, gt_imgs=None, savedir=None, render_factor=0, verbose=False,
                row_block=None):
    """run_nerf.py:339-375 -> (rgbs [M,H,W,3], disps [M,H,W]) as numpy arrays.
    row_block=(rank, world) (extension): this process renders only its contiguous block of image rows of
    every pose and returns [M, rows, W, ...]; ranks share nothing, there is no collective."""
    H, W, focal = hwf
    if render_factor != 0:
        H, W = H // render_factor, W // render_factor
        focal = focal / render_factor if not hasattr(focal, "shape") or np.ndim(focal) == 0 else _scale_K(focal, render_factor)
    # the reference converts every frame with .numpy() as soon as it is rendered (run_nerf.py:358),
    # which serialises host and device.  Here each frame is copied into a pinned ring slot asynchronously (the next
    # poses render while it drains; _FrameDrain), to8b / PSNR run on the device, the host only ever waits for a frame
    # that is R = 4 frames old.
    rows = None
    if row_block is not None:
        rank, world = row_block
        base, rem = divmod(int(H), int(world))
        r0 = rank * base + min(rank, rem)
        rows = (r0, r0 + base + (1 if rank < rem else 0))
    poses_list = [(c if isinstance(c, torch.Tensor) else torch.as_tensor(np.asarray(c)))[:3, :4] for c in render_poses]
    M = len(poses_list)
    rgbs = disps = None          # the results: ordinary numpy arrays, like the reference's np.stack(rgbs, 0)
    ring = None                  # _FrameDrain: a few pinned frame slots the device copies into asynchronously
    psnrs = []
    t = time.time()
    with torch.no_grad():
        for i, c2w in enumerate(poses_list):
            if verbose:
                print(i, time.time() - t)
                t = time.time()
            # _static_outputs: the frame's buffers are copied to the host right below, before the next replay on the
            # same stream overwrites them -- no per-frame clone
            rgb, disp, acc, _ = render(H, W, focal, chunk=chunk, c2w=c2w, _static_outputs=True, _rows=rows,
                                       **render_kwargs)
            if rgbs is None:
                rgbs = np.empty((M,) + tuple(rgb.shape), np.float32)
                disps = np.empty((M,) + tuple(disp.shape), np.float32)
                ring = _FrameDrain(tuple(rgb.shape), tuple(disp.shape), savedir is not None, rgb.is_cuda)
            u8 = None
            want_psnr = gt_imgs is not None and render_factor == 0 and verbose
            if want_psnr or savedir is not None:
                gt = torch.as_tensor(np.asarray(gt_imgs[i]), dtype=torch.float32) if want_psnr else None
                if gt is not None and rows is not None:
                    gt = gt[rows[0]:rows[1]]
                u8, psnr = ops.to8b_psnr(rgb, gt, want_u8=savedir is not None)
                if psnr is not None:
                    psnrs.append(psnr)
            ring.push(i, rgb, disp, u8, rgbs, disps, savedir)
    if rgbs is None:
        return np.zeros((0, H, W, 3), np.float32), np.zeros((0, H, W), np.float32)
    ring.finish(rgbs, disps, savedir)
    for p in psnrs:
        print(float(p))
    return rgbs, disps


_DRAIN_CACHE = {}


class _FrameDrain:
    """Frames leave the device through a ring of pinned slots (allocated once per frame shape: pinning M whole frames per
    call costs more than rendering them at small sizes).  Frame i is copied device -> slot i % R asynchronously; when the
    slot comes round again -- R frames later, so the host never waits for the frame being rendered -- its event is
    awaited and the frame is moved into the caller's numpy arrays (and written as a PNG when asked)."""
    R = 4

    def __init__(self, rgb_shape, disp_shape, want_u8, cuda):
        key = (rgb_shape, disp_shape, bool(want_u8), bool(cuda))
        ent = _DRAIN_CACHE.get(key)
        if ent is None:
            if len(_DRAIN_CACHE) >= 4:
                _DRAIN_CACHE.clear()
            ent = {"rgb": torch.empty((self.R,) + rgb_shape, dtype=torch.float32, pin_memory=cuda),
                   "disp": torch.empty((self.R,) + disp_shape, dtype=torch.float32, pin_memory=cuda),
                   "u8": torch.empty((self.R,) + rgb_shape, dtype=torch.uint8, pin_memory=cuda) if want_u8 else None}
            _DRAIN_CACHE[key] = ent
        self.buf = ent
        self.ev = [None] * self.R
        self.frame = [None] * self.R
        self.cuda = cuda

    def _retire(self, s, rgbs, disps, savedir):
        i = self.frame[s]
        if i is None:
            return
        if self.ev[s] is not None:
            self.ev[s].synchronize()
        rgbs[i] = self.buf["rgb"][s].numpy()
        disps[i] = self.buf["disp"][s].numpy()
        if savedir is not None and self.buf["u8"] is not None:
            _write_png(os.path.join(savedir, "{:03d}.png".format(i)), self.buf["u8"][s].numpy())
        self.frame[s] = None

    def push(self, i, rgb, disp, u8, rgbs, disps, savedir):
        s = i % self.R
        self._retire(s, rgbs, disps, savedir)
        self.buf["rgb"][s].copy_(rgb, non_blocking=True)
        self.buf["disp"][s].copy_(disp, non_blocking=True)
        if u8 is not None:
            self.buf["u8"][s].copy_(u8, non_blocking=True)
        if self.cuda:
            if self.ev[s] is None:
                self.ev[s] = torch.cuda.Event()
            self.ev[s].record()
        self.frame[s] = i

    def finish(self, rgbs, disps, savedir):
        order = sorted((f, s) for s, f in enumerate(self.frame) if f is not None)
        for _, s in order:
            self._retire(s, rgbs, disps, savedir)


def _scale_K(K, factor):
    K = np.array(K, dtype=np.float64, copy=True)
    K[:2, :] /= factor
    return K


def _write_png(path, rgb8):
    try:
        from PIL import Image
        Image.fromarray(rgb8).save(path)
    except ImportError:  # pragma: no cover
        np.save(path + ".npy", rgb8)


def create_nerf(args, device=None):
    """run_nerf.py:378-458 -> (render_kwargs_train, render_kwargs_test, start, grad_vars, models)."""
    device = torch.device(device if device is not None else "cuda")
    embed_fn, input_ch = get_embedder(args.multires, getattr(args, "i_embed", 0))
    if getattr(args, "i_embed", 0) == -1:
        raise NotImplementedError("i_embed=-1 (no positional encoding) is not used by any reference config")
    input_ch_views, embeddirs_fn = 0, None
    if args.use_viewdirs:
        embeddirs_fn, input_ch_views = get_embedder(args.multires_views, getattr(args, "i_embed", 0))
    skips = [4]
    model = NeRF(D=args.netdepth, W=args.netwidth, input_ch=input_ch, output_ch=4, skips=skips,
                 input_ch_views=input_ch_views if args.use_viewdirs else 3, use_viewdirs=args.use_viewdirs).to(device)
    grad_vars = list(model.parameters())
    models = {"model": model}
    model_fine = None
    if args.N_importance > 0:
        model_fine = NeRF(D=args.netdepth_fine, W=args.netwidth_fine, input_ch=input_ch, output_ch=4, skips=skips,
                          input_ch_views=input_ch_views if args.use_viewdirs else 3,
                          use_viewdirs=args.use_viewdirs).to(device)
        grad_vars += list(model_fine.parameters())
        models["model_fine"] = model_fine
    precision = getattr(args, "precision", "fp32")
    model.precision = precision
    if model_fine is not None:
        model_fine.precision = precision

    def network_query_fn(inputs, viewdirs, network_fn):
        return run_network(inputs, viewdirs, network_fn, embed_fn=embed_fn, embeddirs_fn=embeddirs_fn,
                           netchunk=getattr(args, "netchunk", 1024 * 64))

    render_kwargs_train = {
        "network_query_fn": network_query_fn,
        "perturb": args.perturb,
        "N_importance": args.N_importance,
        "network_fine": model_fine,
        "N_samples": args.N_samples,
        "network_fn": model,
        "use_viewdirs": args.use_viewdirs,
        "white_bkgd": args.white_bkgd,
        "raw_noise_std": args.raw_noise_std,
    }
    if args.dataset_type != "llff" or args.no_ndc:
        render_kwargs_train["ndc"] = False
        render_kwargs_train["lindisp"] = args.lindisp
    render_kwargs_test = dict(render_kwargs_train)
    render_kwargs_test["perturb"] = False
    render_kwargs_test["raw_noise_std"] = 0.0

    start = 0
    basedir, expname = getattr(args, "basedir", None), getattr(args, "expname", None)
    ft_path = getattr(args, "ft_path", None)
    if ft_path is not None and ft_path != "None":
        ckpts = [ft_path]
    elif basedir is not None and expname is not None and os.path.isdir(os.path.join(basedir, expname)):
        d = os.path.join(basedir, expname)
        ckpts = [os.path.join(d, f) for f in sorted(os.listdir(d))
                 if ("model_" in f and "fine" not in f and "optimizer" not in f)]
    else:
        ckpts = []
    if len(ckpts) > 0 and not getattr(args, "no_reload", False):
        ft_weights = ckpts[-1]
        model.set_weights(list(np.load(ft_weights, allow_pickle=True)))
        start = int(ft_weights[-10:-4]) + 1
        if model_fine is not None:
            ft_weights_fine = "{}_fine_{}".format(ft_weights[:-11], ft_weights[-10:])
            model_fine.set_weights(list(np.load(ft_weights_fine, allow_pickle=True)))
    return render_kwargs_train, render_kwargs_test, start, grad_vars, models


def save_weights(models, basedir, expname, i):
    """The np.save half of the TF checkpoint layout (run_nerf.py:835-843): model_{i:06d}.npy, model_fine_{i:06d}.npy."""
    os.makedirs(os.path.join(basedir, expname), exist_ok=True)
    for k, net in models.items():
        path = os.path.join(basedir, expname, "{}_{:06d}.npy".format(k, i))
        arr = np.empty(len(net.get_weights()), dtype=object)
        for j, w in enumerate(net.get_weights()):
            arr[j] = w
        np.save(path, arr, allow_pickle=True)
