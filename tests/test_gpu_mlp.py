"""GPU: the MLP kernels (fp32 FFMA path and bf16 tcgen05 path) against the oracle graphs."""
import pytest
import torch

from conftest import load_golden, maxdiff
from oracle import nerf_oracle as O

pytestmark = pytest.mark.gpu

import nerf_kinematics_b200 as nk  # noqa: E402

DEV = "cuda"


def _pts(n, S, seed):
    g = torch.Generator().manual_seed(seed)
    pts = (torch.rand(n, S, 3, generator=g) * 2 - 1) * 3.0
    dirs = torch.nn.functional.normalize(torch.randn(n, 3, generator=g), dim=-1)
    return pts, dirs


def test_flexible_fp32_against_reference_checkpoint_output():
    g = load_golden("mlp_flexible")
    ck = load_golden("ckpt_lego_lowres3")
    m = nk.model_from_state_dict({k[2:]: v for k, v in ck.items() if k.startswith("c.")}).to(DEV)
    raw = m.query(g["pts"].to(DEV)[:, None, :], g["dirs"].to(DEV))
    assert raw.shape == (200, 1, 4)
    # the seeded points lie outside the trained volume, raw reaches ~5e3: compare relative to magnitude
    rel = (raw[:, 0].cpu() - g["raw"]).abs() / g["raw"].abs().clamp_min(1.0)
    assert float(rel.max()) <= 1e-4


@pytest.mark.parametrize("kind", ["flexible", "tf"])
@pytest.mark.parametrize("uv", [True, False])
def test_fp32_graphs_against_oracle(kind, uv):
    pts, dirs = _pts(37, 13, 4)
    if kind == "flexible":
        sd = O.init_flexible_state(seed=2)
        if not uv:
            sd = {k: v for k, v in sd.items() if k.startswith(("layer1", "layers_xyz"))}
            sd["fc_out.weight"], sd["fc_out.bias"] = torch.randn(4, 128) * 0.1, torch.randn(4) * 0.1
        ref = O.run_mlp("flexible", sd, pts, dirs if uv else None, 10, 4)
    else:
        sd = O.init_tf8x256_state(use_viewdirs=uv, seed=3)
        ref = O.run_mlp("tf", sd, pts, dirs if uv else None, 10, 4)
    m = nk.model_from_state_dict(sd).to(DEV)
    got = m.query(pts.to(DEV), dirs.to(DEV) if uv else None)
    assert maxdiff(got, ref) <= 1e-4


def test_tf_graph_small_widths_and_depths():
    pts, dirs = _pts(9, 5, 7)
    sd = O.init_tf8x256_state(D=6, W=64, seed=5)
    # the bottleneck is Dense(256) whatever W is (run_nerf_helpers.py:104)
    assert sd["bottleneck_linear.weight"].shape == (256, 64)
    ref = O.run_mlp("tf", sd, pts, dirs, 10, 4)
    got = nk.model_from_state_dict(sd).to(DEV).query(pts.to(DEV), dirs.to(DEV))
    assert maxdiff(got, ref) <= 1e-4


def test_bf16_tensor_core_path_against_fp32_oracle():
    pts, dirs = _pts(300, 64, 11)
    for kind, sd in (("tf", O.init_tf8x256_state(seed=0)), ("flexible", O.init_flexible_state(seed=1))):
        ref = O.run_mlp(kind, sd, pts, dirs, 10, 4)
        m = nk.model_from_state_dict(sd).to(DEV)
        m.precision = "bf16"
        got = m.query(pts.to(DEV), dirs.to(DEV))
        err = maxdiff(got, ref)
        scale = float(ref.abs().max())
        assert err <= 2e-2 * max(1.0, scale), (kind, err, scale)


@pytest.mark.parametrize("n_points", [77, 128 * 3 + 5, 128 * 333 + 1])
def test_bf16_pair_kernel_ragged_tile_counts(n_points):
    """The forward kernel takes 128-point tiles two at a time from a per-CTA contiguous range: cover a lone partial
    tile, an even split and ranges with a trailing single tile (odd tiles per CTA), for both graphs."""
    g = torch.Generator().manual_seed(n_points)
    pts = (torch.rand(n_points, 1, 3, generator=g) * 2 - 1) * 3.0
    dirs = torch.nn.functional.normalize(torch.randn(n_points, 3, generator=g), dim=-1)
    for kind, sd in (("tf", O.init_tf8x256_state(seed=2)), ("flexible", O.init_flexible_state(seed=3))):
        ref = O.run_mlp(kind, sd, pts, dirs, 10, 4)
        m = nk.model_from_state_dict(sd).to(DEV)
        m.precision = "bf16"
        got = m.query(pts.to(DEV), dirs.to(DEV))
        assert torch.isfinite(got).all()
        err = maxdiff(got, ref)
        assert err <= 2e-2 * max(1.0, float(ref.abs().max())), (kind, n_points, err)


@pytest.mark.parametrize("cfg", [dict(use_viewdirs=False), dict(D=6, W=128), dict(D=8, W=256)])
def test_bf16_biases_ride_in_the_mma_other_graphs_and_large_biases(cfg):
    """The 16-bit kernel adds no bias in its epilogue: hidden-layer biases are element 63 of the encoding weights
    (times the encoding block's constant-one channel) or one extra K=16 step.  Cover the graph without view
    directions (output_linear head), a narrower and shallower trunk, and biases large enough
    (x 20) that a dropped or misplaced one shows far above the 16-bit tolerance."""
    pts, dirs = _pts(257, 3, 23)
    sd = O.init_tf8x256_state(seed=9, **cfg)
    sd = {k: (v * 20.0 if k.endswith(".bias") else v) for k, v in sd.items()}
    uv = cfg.get("use_viewdirs", True)
    ref = O.run_mlp("tf", sd, pts, dirs if uv else None, 10, 4)
    m = nk.model_from_state_dict(sd).to(DEV)
    m.precision = "bf16"
    got = m.query(pts.to(DEV), dirs.to(DEV) if uv else None)
    scale = float(ref.abs().max())
    assert scale > 0.5, scale                   # the biases matter in this network
    zero_b = {k: (torch.zeros_like(v) if k.endswith(".bias") else v) for k, v in sd.items()}
    assert maxdiff(O.run_mlp("tf", zero_b, pts, dirs if uv else None, 10, 4), ref) > 0.1 * scale
    assert maxdiff(got, ref) <= 1e-2 * max(1.0, scale), (maxdiff(got, ref), scale)
