"""nerf_kinematics_b200 -- the classic-NeRF hot path of tomeupm/nerf-kinematics, B200-native.

Host side: Python/PyTorch mirroring the reference's API (run_nerf.create_nerf / render / render_path,
train_utils.run_one_iter_of_nerf, nerf_helpers.*, models.FlexibleNeRFModel / NeRF, checkpoint layouts).
Device side: libnerf_b200.so, a plain C-ABI (include/nerf_b200.h) over hand-written sm_100a kernels.
There is no CPU fallback: without the library or a CUDA device every compute call raises.
"""
from . import _lib  # noqa: F401  (loads / builds libnerf_b200.so; raises if it cannot)
from .models import FlexibleNeRFModel, NeRF, model_from_state_dict  # noqa: F401
from .nerf_helpers import (get_embedder, get_embedding_function, get_ray_bundle, get_rays, img2mse, mse2psnr,  # noqa: F401
                           ndc_rays, positional_encoding, sample_pdf, sample_pdf_2, to8b)
from .run_nerf import batchify_rays, create_nerf, render, render_path, render_rays, run_network  # noqa: F401
from .train_utils import predict_and_render_radiance, run_one_iter_of_nerf  # noqa: F401
from .volume_rendering_utils import volume_render_radiance_field  # noqa: F401

__version__ = "0.1.0"
BF16_BACKWARD = True  # the tensor-core path has a backward (csrc/mlp_tc_bwd.cu): bench.py trains with it
