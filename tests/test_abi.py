"""CPU: the C-ABI library loads, exports every symbol include/nerf_b200.h declares, the ctypes binding
covers them all, the host-side model description matches the reference's layouts, and compute entry
points fail loudly (no CPU fallback) when there is no CUDA device."""
import ctypes as C
import os
import re

import pytest
import torch

from conftest import ROOT, load_golden

import nerf_kinematics_b200 as nk
from nerf_kinematics_b200 import _lib, ops


def _declared():
    src = open(os.path.join(ROOT, "include", "nerf_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(nb_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    names = _declared()
    assert len(names) >= 30
    lib = C.CDLL(_lib.LIB_PATH)
    for n in names:
        assert hasattr(lib, n), "libnerf_b200.so does not export %s" % n
    assert sorted(_lib.SIGNATURES) == names, set(names) ^ set(_lib.SIGNATURES)
    assert _lib.lib().nb_abi_version() == 3


def test_no_torch_types_in_abi():
    src = open(os.path.join(ROOT, "include", "nerf_b200.h")).read()
    assert 'extern "C"' in src
    code = re.sub(r"/\*.*?\*/", "", src, flags=re.S)  # comments cite the reference's torch calls; code must not
    assert "torch" not in code.lower() and "at::" not in code and "#include <cuda" not in code


def test_model_layouts_match_reference_state_dicts():
    ck = load_golden("ckpt_lego_lowres3")
    sd = {k[2:]: v for k, v in ck.items() if k.startswith("c.")}
    m = nk.model_from_state_dict(sd)
    assert isinstance(m, nk.FlexibleNeRFModel)
    assert m.n_params == 84548 and m.flops_per_point == 167680
    assert list(m.state_dict().keys()) == list(sd.keys())  # same keys, same registration order
    for k, v in m.state_dict().items():
        assert torch.equal(v, sd[k])
    ckw = load_golden("ckpt_wheel_robot6")
    mw = nk.model_from_state_dict({k[2:]: v for k, v in ckw.items() if k.startswith("f.")})
    assert mw.n_params == 81476 and mw.dim_xyz == 39
    tf = nk.NeRF()
    assert tf.n_params == 595844 and tf.flops_per_point == 1186816
    assert tf.pts_linears[5].weight.shape == (256, 319)
    w = tf.get_weights()
    assert len(w) == 24 and w[0].shape == (63, 256) and w[10].shape == (319, 256)  # Keras [in, out]
    tf2 = nk.NeRF()
    tf2.set_weights(w)
    for a, b in zip(tf.state_dict().values(), tf2.state_dict().values()):
        assert torch.equal(a, b)


def test_keras_weight_list_order_matches_real_npy_files():
    """A Keras functional Model lists layers by graph depth: trunk, bottleneck, views, rgb, and alpha LAST (the
    official model_200000.npy / model_fine_*.npy layout read by the load_weights_from_keras converters): index
    2D = [W,256], 2D+2 = [256+27, W/2], 2D+4 = [W/2,3], 2D+6 = [W,1]."""
    D, W = 8, 256
    tf = nk.NeRF(D=D, W=W)
    w = tf.get_weights()
    shapes = [tuple(a.shape) for a in w]
    assert shapes[2 * D] == (W, 256) and shapes[2 * D + 1] == (256,)
    assert shapes[2 * D + 2] == (256 + 27, W // 2) and shapes[2 * D + 4] == (W // 2, 3)
    assert shapes[2 * D + 6] == (W, 1) and shapes[2 * D + 7] == (1,)
    # a list in the real file's shapes loads; alpha lands on alpha_linear
    import numpy as np
    rs = np.random.RandomState(0)
    real = [rs.randn(*s).astype(np.float32) for s in shapes]
    tf.set_weights(real)
    assert torch.equal(tf.alpha_linear.weight.detach().cpu(), torch.from_numpy(real[2 * D + 6]).t())
    assert torch.equal(tf.bottleneck_linear.bias.detach().cpu(), torch.from_numpy(real[2 * D + 1]))
    # the old (wrong) order [alpha, bottleneck, views, rgb] is rejected instead of silently mis-assigned
    wrong = real[:2 * D] + real[2 * D + 6:] + real[2 * D:2 * D + 6]
    with pytest.raises(ValueError):
        tf.set_weights(wrong)
    # without view directions: trunk + output_linear
    nv = nk.NeRF(use_viewdirs=False, input_ch_views=3)
    assert [tuple(a.shape) for a in nv.get_weights()][-2:] == [(W, 4), (4,)]


def test_flexible_with_active_skip_is_rejected_like_the_reference_crash():
    with pytest.raises(NotImplementedError):
        nk.FlexibleNeRFModel(num_layers=8, hidden_size=256)


def test_invalid_arguments_return_errors_not_crashes():
    lib = _lib.lib()
    h = C.c_void_p()
    assert lib.nb_model_create(7, 8, 256, 10, 4, 1, 4, C.byref(h)) == -1
    assert b"unknown graph" in lib.nb_last_error()
    assert lib.nb_searchsorted(None, 3, 4, None, 5, 4, 1, None, None) == -1  # row mismatch
    assert lib.nb_sample_pdf(None, 4, None, 3, 0, 1, None, 8, None, None, None) == -1  # nb < 2


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU behaviour")
def test_compute_fails_loudly_without_gpu():
    assert _lib.lib().nb_device_count() < 0 or _lib.lib().nb_device_count() == 0
    with pytest.raises(_lib.NerfB200Error):
        ops.get_rays(4, 4, 10.0, torch.eye(4))
    with pytest.raises(_lib.NerfB200Error):
        ops.posenc(torch.zeros(3, 3), 4)
    m = nk.FlexibleNeRFModel()
    with pytest.raises(_lib.NerfB200Error):
        m.query(torch.zeros(2, 4, 3), torch.zeros(2, 3))
    with pytest.raises(NotImplementedError):
        m(torch.zeros(2, 66))


def test_options_round_trip_without_a_gpu():
    """nb_set_option returns the previous value and rejects unknown names (include/nerf_b200.h); no device needed."""
    L = _lib.lib()
    for name, new in ((b"fwd_4tile", 0), (b"fwd_2cta", 1), (b"bwd_sm_reserve", 5), (b"composite_lc", 0), (b"sample_fine2", 0)):
        old = L.nb_set_option(name, new)
        assert old >= 0
        assert L.nb_set_option(name, old) == new
        assert L.nb_set_option(name, old) == old
    assert L.nb_set_option(b"no_such_option", 1) == -1
    assert b"no_such_option" in L.nb_last_error()
    assert abs(L.nb_mlp_activation_scale() - 2.0 ** -5) == 0.0
