"""volume_render_radiance_field with the reference signature (nerf-pytorch/nerf/volume_rendering_utils.py:6-53)."""
import torch

from . import ops


def volume_render_radiance_field(radiance_field, depth_values, ray_directions, radiance_field_noise_std=0.0,
                                 white_background=False):
    raw = radiance_field
    n, S = depth_values.shape[0], depth_values.shape[-1]
    dev = raw.device
    # the kernel reads the direction from packed rays [n, 8]
    rays = torch.zeros((n, 8), dtype=torch.float32, device=dev)
    rays[:, 3:6] = ray_directions.reshape(-1, 3)
    noise = None
    if radiance_field_noise_std > 0.0:
        noise = torch.randn((n, S), dtype=torch.float32, device=dev) * radiance_field_noise_std
    rgb, disp, acc, w, depth = ops.composite(raw.reshape(n, S, 4), depth_values.reshape(n, S), rays, noise,
                                             white_background)
    return rgb, disp, acc, w, depth
