"""ctypes binding of the C-ABI in include/nerf_b200.h (libnerf_b200.so).

There is deliberately no CPU fallback: if the library is missing it is built with nvcc
(nerf_kinematics_b200/_build.py); if that fails, or a compute entry point is called without a CUDA
device, the call raises.  Every symbol declared in include/nerf_b200.h is bound here with its
argument types, so a missing export fails at import time.
"""
import ctypes as C
import os
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libnerf_b200.so")

p = C.c_void_p
i32 = C.c_int
i64 = C.c_int64
f32 = C.c_float
f64 = C.c_double
sz = C.c_size_t



class RenderCfg(C.Structure):  # nb_render_cfg
    _fields_ = [("n_samples", i32), ("n_importance", i32), ("lindisp", i32), ("white_bkgd", i32),
                ("precision", i32), ("save_for_backward", i32)]


class RenderRand(C.Structure):  # nb_render_rand
    _fields_ = [("t_rand", p), ("u_fine", p), ("noise_c", p), ("noise_f", p), ("philox_state", p), ("perturb", C.c_int),
                ("raw_noise_std", C.c_float)]


class RenderOut(C.Structure):  # nb_render_out
    _fields_ = [(k, p) for k in ("rgb", "disp", "acc", "depth", "rgb0", "disp0", "acc0", "depth0", "z_std", "raw",
                                 "z_vals", "weights")]


class RenderGrads(C.Structure):  # nb_render_grads
    _fields_ = [(k, p) for k in ("d_rgb", "d_disp", "d_acc", "d_depth", "d_rgb0", "d_disp0", "d_acc0", "d_depth0")]


# name -> (restype, argtypes); mirrors include/nerf_b200.h one to one
SIGNATURES = {
    "nb_abi_version": (i32, []),
    "nb_last_error": (C.c_char_p, []),
    "nb_device_count": (i32, []),
    "nb_set_option": (i32, [C.c_char_p, i32]),
    "nb_launch_count": (i64, []),
    "nb_launch_count_reset": (None, []),
    "nb_model_create": (i32, [i32, i32, i32, i32, i32, i32, i32, C.POINTER(p)]),
    "nb_model_destroy": (i32, [p]),
    "nb_model_param_count": (i64, [p]),
    "nb_model_num_tensors": (i32, [p]),
    "nb_model_tensor_info": (i32, [p, i32, C.c_char_p, i32, C.POINTER(i64), C.POINTER(i64), C.POINTER(i64)]),
    "nb_model_set_params": (i32, [p, p, i64]),
    "nb_model_pack": (i32, [p, i32, p]),
    "nb_model_flops_per_point": (i64, [p]),
    "nb_get_rays": (i32, [i32, i32, f32, f32, f32, f32, C.POINTER(f32), i32, i32, p, p, p]),
    "nb_get_rays_at": (i32, [i32, i32, f32, f32, f32, f32, C.POINTER(f32), p, i64, p, p, p]),
    "nb_get_rays_dev": (i32, [i32, i32, f32, f32, f32, f32, p, i32, i32, p, p, p]),
    "nb_get_rays_at_dev": (i32, [i32, i32, f32, f32, f32, f32, p, p, i64, p, p, p]),
    "nb_check_finite": (i32, [p, i64, p, i32, p]),
    "nb_to8b": (i32, [p, p, i64, p, p, p]),
    "nb_rng_advance": (i32, [p, p]),
    "nb_rng_fill": (i32, [p, i32, i32, f32, i64, i32, p, p]),
    "nb_ndc_rays": (i32, [i32, i32, f64, f64, p, p, i64, p, p, p]),
    "nb_make_ray_batch": (i32, [p, p, p, i64, i32, i32, i32, i32, f64, f64, f64, p, p, p, p]),
    "nb_sample_coarse": (i32, [p, i32, i64, i32, i32, p, p, p]),
    "nb_posenc": (i32, [p, i64, i32, i32, p, i32, p]),
    "nb_mlp_workspace_bytes": (sz, [p, i32, i64, i32]),
    "nb_mlp_forward": (i32, [p, i32, p, i32, p, i64, i32, p, p, sz, i32, p]),
    "nb_mlp_forward_points": (i32, [p, i32, p, p, i64, i32, p, p, sz, p]),
    "nb_mlp_backward": (i32, [p, i32, i64, p, p, sz, p, p]),
    "nb_mlp_activation_scale": (f32, []),
    "nb_mlp_mask_record_layout": (i32, [p, i64, p, C.POINTER(i64), C.POINTER(i32)]),
    "nb_mlp_record_layout": (i32, [p, i64, p, C.POINTER(i64), C.POINTER(i32), C.POINTER(i64), C.POINTER(i32)]),
    "nb_composite": (i32, [p, p, p, i32, i64, i32, p, i32, p, p, p, p, p, p]),
    "nb_composite_backward": (i32, [p, p, p, i32, i64, i32, p, i32, p, p, p, p, p, p]),
    "nb_searchsorted": (i32, [p, i64, i32, p, i64, i32, i32, p, p]),
    "nb_sample_pdf": (i32, [p, i32, p, i32, i64, i32, p, i32, p, p, p]),
    "nb_sample_fine": (i32, [p, p, i64, i32, p, i32, p, p, p, p]),
    "nb_merge_sorted": (i32, [p, i32, p, i32, i64, p, p]),
    "nb_mse_loss": (i32, [p, p, i64, f32, p, p, p]),
    "nb_adam_step": (i32, [p, p, p, p, i64, p, i32, i32, p]),
    "nb_render_workspace_bytes": (sz, [p, p, i32, i64, C.POINTER(RenderCfg)]),
    "nb_render_rays": (i32, [p, p, p, i32, i64, C.POINTER(RenderCfg), C.POINTER(RenderRand), C.POINTER(RenderOut), p,
                             sz, p]),
    "nb_render_rays_backward": (i32, [p, p, p, i32, i64, C.POINTER(RenderCfg), C.POINTER(RenderRand),
                                      C.POINTER(RenderGrads), p, sz, p, p, p]),
    "nb_render_static_scratch_bytes": (sz, []),
    "nb_render_rays_static": (i32, [p, p, p, i32, i64, C.POINTER(RenderCfg), C.POINTER(RenderRand), C.POINTER(RenderOut),
                                    p]),
}

_lock = threading.Lock()
_lib = None


class NerfB200Error(RuntimeError):
    pass


def _load():
    global _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            from . import _build
            _build.build()
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)  # AttributeError here == the library does not export the header
            fn.restype = res
            fn.argtypes = args
        if lib.nb_abi_version() != 3:
            raise NerfB200Error("libnerf_b200.so ABI version %d, binding expects 3" % lib.nb_abi_version())
        _lib = lib
        return lib


def lib():
    return _lib if _lib is not None else _load()


def last_error() -> str:
    return lib().nb_last_error().decode("utf-8", "replace")


def check(rc: int, what: str = ""):
    if rc != 0:
        raise NerfB200Error("%s failed (%d): %s" % (what or "libnerf_b200 call", rc, last_error()))


def require_cuda():
    n = lib().nb_device_count()
    if n <= 0:
        raise NerfB200Error("no CUDA device: nerf_kinematics_b200 has no CPU fallback (%s)" % last_error())
    return n
