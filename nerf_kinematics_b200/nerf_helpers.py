"""Reference-named math helpers, each a thin call into the CUDA library (no torch re-implementation).

Names and argument order follow nerf-pytorch/nerf/nerf_helpers.py and nerf/run_nerf_helpers.py so
reference call sites keep working: get_ray_bundle / get_rays, ndc_rays, positional_encoding,
get_embedding_function / get_embedder, sample_pdf / sample_pdf_2, img2mse, mse2psnr, to8b.
"""
import math

import torch

from . import ops


def img2mse(img_src, img_tgt):
    """nerf_helpers.py:9 / run_nerf_helpers.py:11 (plain torch: the callers differentiate through it)."""
    return torch.nn.functional.mse_loss(img_src, img_tgt)


def mse2psnr(mse):
    """nerf_helpers.py:13-18 / run_nerf_helpers.py:14."""
    mse = float(mse)
    if mse == 0:
        mse = 1e-5
    return -10.0 * math.log10(mse)


def to8b(x):
    """run_nerf_helpers.py:17: (255*clip(x,0,1)).astype(uint8) on a numpy array or tensor."""
    import numpy as np
    if isinstance(x, torch.Tensor):
        return (255.0 * x.clamp(0.0, 1.0)).to(torch.uint8)
    return (255 * np.clip(x, 0, 1)).astype(np.uint8)


def get_ray_bundle(height, width, focal_length, tform_cam2world):
    """nerf_helpers.py:67-110 -> (ray_origins, ray_directions), each [H, W, 3]."""
    return ops.get_rays(int(height), int(width), focal_length, tform_cam2world)


def get_rays(H, W, K, c2w):
    """run_nerf_helpers.py:123-130; K may be the scalar focal of the TF code or a 3x3 intrinsics matrix."""
    return ops.get_rays(int(H), int(W), K, c2w)


def ndc_rays(H, W, focal, near, rays_o, rays_d):
    """run_nerf_helpers.py:143-178 == nerf_helpers.py:170-197."""
    return ops.ndc_rays(H, W, focal, near, rays_o, rays_d)


def positional_encoding(tensor, num_encoding_functions=6, include_input=True, log_sampling=True):
    """nerf_helpers.py:113-157."""
    if not log_sampling:
        raise NotImplementedError("log_sampling=False is not used by any reference config")
    return ops.posenc(tensor, num_encoding_functions, include_input)


def get_embedding_function(num_encoding_functions=6, include_input=True, log_sampling=True):
    """nerf_helpers.py:160-167."""
    return lambda x: positional_encoding(x, num_encoding_functions, include_input, log_sampling)


def get_embedder(multires, i=0):
    """run_nerf_helpers.py:59-75 -> (embed_fn, out_dim); i == -1 is the identity embedder."""
    if i == -1:
        return (lambda x: x), 3
    return (lambda x: ops.posenc(x, multires, True)), 3 + 6 * multires


def sample_pdf(bins, weights, N_samples, det=False):
    """run_nerf_helpers.py:183-211 (also the dead nerf_helpers.py:222 variant's signature)."""
    return ops.sample_pdf(bins, weights, N_samples, det=det)


sample_pdf_2 = sample_pdf  # nerf_helpers.py:260-302, the one train_utils.py:4 imports


def cumprod_exclusive(tensor):
    """nerf_helpers.py:43-64.  Kept for API completeness; the compositing kernel does this scan itself."""
    c = torch.cumprod(tensor, dim=-1)
    c = torch.roll(c, 1, -1)
    c[..., 0] = 1.0
    return c


def get_minibatches(inputs, chunksize=1024 * 8):
    """nerf_helpers.py:20-25."""
    return [inputs[i:i + chunksize] for i in range(0, inputs.shape[0], chunksize)]
