"""GPU: the three tensor-core forward kernels are one function.  mlp_tc_kernel (two tiles per CTA), mlp_tc4_kernel
(four tiles per CTA, graphs with every stage N <= 128: the FlexibleNeRFModel 4x128 of every shipped checkpoint) and
mlp_tc2_kernel (CTA pair, optional) issue the same MMAs in the same order for a row and share the encoding and the
epilogue arithmetic, so their outputs must be equal bit for bit -- for single tiles, partly filled last rounds, ragged
last tiles and sizes that give every CTA several rounds.  Selected through nb_set_option (include/nerf_b200.h)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

import nerf_kinematics_b200 as nk  # noqa: E402
from nerf_kinematics_b200 import ops  # noqa: E402
from nerf_kinematics_b200._lib import lib  # noqa: E402

DEV = "cuda"
SIZES = ((1, 64), (3, 64), (7, 192), (9, 64), (148 * 4 + 3, 64), (4099, 64), (6007, 192))


def _batch(m, n, S, seed):
    g = torch.Generator().manual_seed(seed)
    o = (torch.randn(n, 3, generator=g) * 0.5 + torch.tensor([0, 0, 4.0])).to(DEV)
    d = torch.nn.functional.normalize(torch.randn(n, 3, generator=g), dim=-1).to(DEV)
    rays = ops.make_ray_batch(o, d, 2.0, 6.0, use_viewdirs=bool(m.use_viewdirs))
    z = ops.sample_coarse(rays, S)
    pts = rays[:, None, 0:3] + rays[:, None, 3:6] * z[..., None]
    return pts, (rays[:, 8:11] if m.use_viewdirs else None)


def _query(m, pts, vd, **opts):
    old = {k: lib().nb_set_option(k.encode(), v) for k, v in opts.items()}
    try:
        with torch.no_grad():
            out = m.query(pts, vd, precision="fp16").clone()
        torch.cuda.synchronize()
    finally:
        for k, v in old.items():
            lib().nb_set_option(k.encode(), v)
    return out


@pytest.mark.parametrize("kind", ["L6_viewdirs", "L10_viewdirs", "L6_noview"])
def test_four_tile_kernel_equals_two_tile_kernel(kind):
    torch.manual_seed(11)
    m = {"L6_viewdirs": lambda: nk.FlexibleNeRFModel(num_encoding_fn_xyz=6),
         "L10_viewdirs": lambda: nk.FlexibleNeRFModel(num_encoding_fn_xyz=10),
         "L6_noview": lambda: nk.FlexibleNeRFModel(num_encoding_fn_xyz=6, use_viewdirs=False)}[kind]().to(DEV)
    for i, (n, S) in enumerate(SIZES):
        pts, vd = _batch(m, n, S, i)
        a = _query(m, pts, vd, fwd_4tile=0, fwd_2cta=0)
        b = _query(m, pts, vd, fwd_4tile=1, fwd_2cta=0)
        assert bool(torch.isfinite(b).all())
        assert torch.equal(a, b), "%s %dx%d: four-tile kernel differs, max %.3e" % (kind, n, S, float((a - b).abs().max()))
        ref = m.query(pts, vd, precision="fp32")
        assert float((b - ref).abs().max()) <= 5e-3 * max(1.0, float(ref.abs().max()))


def test_wide_graph_is_not_taken_by_the_four_tile_kernel_and_cta_pair_kernel_is_equal():
    """8x256 is outside the four-tile kernel's shape (N = 256, 4 activation blocks): the option must not change it;
    the CTA-pair kernel (off by default) computes the same bits."""
    torch.manual_seed(12)
    m = nk.NeRF().to(DEV)
    for i, (n, S) in enumerate(SIZES[:5]):
        pts, vd = _batch(m, n, S, 100 + i)
        a = _query(m, pts, vd, fwd_4tile=0, fwd_2cta=0)
        assert torch.equal(a, _query(m, pts, vd, fwd_4tile=1, fwd_2cta=0))
        assert torch.equal(a, _query(m, pts, vd, fwd_4tile=0, fwd_2cta=1))
