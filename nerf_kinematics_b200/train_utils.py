"""run_one_iter_of_nerf / predict_and_render_radiance / run_network with the signatures of
nerf-pytorch/nerf/train_utils.py, on top of the CUDA library.

One reference quirk is reproduced on purpose and is switchable: run_one_iter_of_nerf does not forward
`mode` to predict_and_render_radiance (train_utils.py:171-180), so sampling / noise / background
options always come from options.nerf.train even when mode == "validation"; only the chunk size and
the final reshape follow `mode`.  Set REFERENCE_MODE_FORWARDING = False to use options.nerf.<mode>.
"""
import torch

from . import ops
from .render_core import render_chunk

REFERENCE_MODE_FORWARDING = True


def run_network(network_fn, pts, ray_batch, chunksize, embed_fn=None, embeddirs_fn=None):
    """train_utils.py:8-25: pts [N,S,3], ray_batch [N, 8|11] -> [N,S,4].  Encoders / chunksize are fused away."""
    viewdirs = ray_batch[..., -3:] if (embeddirs_fn is not None or ray_batch.shape[-1] > 8) else None
    if not getattr(network_fn, "use_viewdirs", True):
        viewdirs = None
    return network_fn.query(pts, viewdirs)


def predict_and_render_radiance(ray_batch, model_coarse, model_fine, options, mode="train", encode_position_fn=None,
                                encode_direction_fn=None, precision=None, rnd=None):
    """train_utils.py:28-127 -> (rgb_coarse, disp_coarse, acc_coarse, rgb_fine, disp_fine, acc_fine)."""
    o = getattr(options.nerf, mode)
    out = render_chunk(ray_batch, model_coarse, model_fine if model_fine else None, N_samples=o.num_coarse,
                       N_importance=o.num_fine, lindisp=o.lindisp, perturb=float(bool(o.perturb)),
                       white_bkgd=o.white_background, raw_noise_std=o.radiance_field_noise_std, precision=precision,
                       rnd=rnd)
    if o.num_fine > 0:
        return out["rgb0"], out["disp0"], out["acc0"], out["rgb"], out["disp"], out["acc"]
    return out["rgb"], out["disp"], out["acc"], None, None, None


def run_one_iter_of_nerf(height, width, focal_length, model_coarse, model_fine, ray_origins, ray_directions, options,
                         mode="train", encode_position_fn=None, encode_direction_fn=None, precision=None, rnd=None):
    """train_utils.py:130-202."""
    restore_shapes = [ray_directions.shape, ray_directions.shape[:-1], ray_directions.shape[:-1]]
    if model_fine:
        restore_shapes += restore_shapes
    use_viewdirs = bool(options.nerf.use_viewdirs)
    ndc = options.dataset.no_ndc is False
    rays = ops.make_ray_batch(ray_origins, ray_directions, options.dataset.near, options.dataset.far,
                              use_viewdirs=use_viewdirs, ndc=ndc, H=int(height), W=int(width),
                              focal=float(focal_length))
    chunksize = getattr(options.nerf, mode).chunksize
    inner_mode = "train" if REFERENCE_MODE_FORWARDING else mode
    pred = [predict_and_render_radiance(rays[i:i + chunksize], model_coarse, model_fine, options, mode=inner_mode,
                                        encode_position_fn=encode_position_fn,
                                        encode_direction_fn=encode_direction_fn, precision=precision,
                                        rnd=_slice_rnd(rnd, i, chunksize))
            for i in range(0, rays.shape[0], chunksize)]
    images = list(zip(*pred))
    images = [(torch.cat(im, dim=0) if len(im) > 1 else im[0]) if im[0] is not None else None for im in images]
    if mode == "validation":
        images = [im.view(shape) if im is not None else None for im, shape in zip(images, restore_shapes)]
        if model_fine:
            return tuple(images)
        return tuple(images + [None, None, None])
    return tuple(images)


def _slice_rnd(rnd, i, n):
    if not rnd:
        return rnd
    return {k: (v[i:i + n] if v is not None else None) for k, v in rnd.items()}
