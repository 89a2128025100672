"""GPU: the corners of the hot path through the reference-facing API -- empty and single-ray batches, ragged sample
counts (not multiples of the warp), no fine network, chunk sizes that leave a one-ray tail, per-ray near / far.  Every
case is compared with the fp32 CPU oracle (oracle/nerf_oracle.py, which restates run_nerf.py:48-243) on the same inputs."""
import pytest
import torch

from conftest import maxdiff
from oracle import nerf_oracle as O

pytestmark = pytest.mark.gpu

import nerf_kinematics_b200 as nk  # noqa: E402

DEV = "cuda"


def _scene(seed):
    sdc, sdf = O.init_flexible_state(dim_xyz=39, seed=seed), O.init_flexible_state(dim_xyz=39, seed=seed + 1)
    for sd in (sdc, sdf):    # a volume with structure (default init renders near-empty space)
        sd["fc_alpha.weight"] = sd["fc_alpha.weight"] * 40.0
    return sdc, sdf


def _rays(n, seed, near=2.0, far=6.0):
    g = torch.Generator().manual_seed(seed)
    o = torch.randn(n, 3, generator=g) * 0.3 + torch.tensor([0.0, 0.0, 4.0])
    d = torch.randn(n, 3, generator=g) * 0.3 + torch.tensor([0.0, 0.0, -1.0])
    return o, d


def _kw(mc, mf, S, I, **over):
    kw = dict(network_query_fn=None, perturb=0.0, N_importance=I, network_fine=mf, N_samples=S, network_fn=mc,
              use_viewdirs=True, white_bkgd=False, raw_noise_std=0.0, ndc=False, lindisp=False, near=2.0, far=6.0)
    kw.update(over)
    return kw


def _oracle(sdc, sdf, o, d, S, I, chunk=1 << 20, **kw):
    rays = O.pack_rays(o, d, 2.0, 6.0, use_viewdirs=True)
    return O.render_rays_chunked("flexible", sdc, sdf, rays, chunk, n_coarse=S, n_fine=I, n_freq_xyz=6, n_freq_dir=4, **kw)


@pytest.mark.parametrize("S,I", [(64, 128), (33, 17), (4, 3), (7, 0), (64, 0), (96, 160)])
def test_ragged_sample_counts_match_oracle(S, I):
    sdc, sdf = _scene(3)
    mc, mf = nk.model_from_state_dict(sdc).to(DEV), nk.model_from_state_dict(sdf).to(DEV)
    o, d = _rays(37, 1)
    ref = _oracle(sdc, sdf if I > 0 else None, o, d, S, I)
    with torch.no_grad():
        rgb, disp, acc, ex = nk.render(1, 37, 1.0, chunk=4096, rays=(o.to(DEV), d.to(DEV)),
                                       **_kw(mc, mf if I > 0 else None, S, I), precision="fp32")
    assert rgb.shape == (37, 3) and disp.shape == (37,) and acc.shape == (37,)
    key = "fine" if I > 0 else "coarse"
    assert maxdiff(rgb, ref["rgb_" + key]) <= 1e-4 and maxdiff(acc, ref["acc_" + key]) <= 1e-4
    if I > 0:
        assert set(ex) == {"rgb0", "disp0", "acc0", "z_std"}
        assert maxdiff(ex["rgb0"], ref["rgb_coarse"]) <= 1e-4
        assert maxdiff(ex["z_std"], ref["z_std"]) <= 1e-4
    else:
        assert ex == {}


def test_empty_and_single_ray_batches():
    sdc, sdf = _scene(5)
    mc, mf = nk.model_from_state_dict(sdc).to(DEV), nk.model_from_state_dict(sdf).to(DEV)
    o, d = _rays(9, 2)
    kw = _kw(mc, mf, 64, 128)
    for prec in ("fp32", "fp16"):
        with torch.no_grad():
            full = nk.render(1, 9, 1.0, chunk=4096, rays=(o.to(DEV), d.to(DEV)), precision=prec, **kw)
            one = nk.render(1, 1, 1.0, chunk=4096, rays=(o[4:5].to(DEV), d[4:5].to(DEV)), precision=prec, **kw)
            # no rays: batchify_rays collects nothing and render() indexes all_ret['rgb_map'] -- the reference raises
            # KeyError there (run_nerf.py:246-257, :329-331), and so does the mirror; the C-ABI call itself accepts n = 0
            with pytest.raises(KeyError):
                nk.render(1, 0, 1.0, chunk=4096, rays=(o[:0].to(DEV), d[:0].to(DEV)), precision=prec, **kw)
        # a ray's result does not depend on what else is in the batch
        assert torch.equal(one[0][0], full[0][4]) and torch.equal(one[2][0], full[2][4])
        assert torch.equal(one[3]["rgb0"][0], full[3]["rgb0"][4])
    # n = 0 straight through the C-ABI: NB_OK, nothing launched
    import ctypes as C
    from nerf_kinematics_b200 import ops
    from nerf_kinematics_b200._lib import RenderCfg, RenderOut, RenderRand, check, lib
    hc, prec_id = mc.prepare("fp32")
    hf, _ = mf.prepare("fp32")
    cfg, rs, ro = RenderCfg(64, 128, 0, 0, prec_id, 0), RenderRand(None, None, None, None, None, 0, 0.0), RenderOut()
    before = ops.launch_count()
    check(lib().nb_render_rays(hc, hf, C.c_void_p(0), 11, 0, C.byref(cfg), C.byref(rs), C.byref(ro), C.c_void_p(0), 0,
                               C.c_void_p(0)), "nb_render_rays")
    assert ops.launch_count() == before


def test_chunk_with_one_ray_tail_and_per_ray_bounds():
    """batchify_rays with a chunk that leaves a single-ray last piece; near / far given per ray (render(rays=...) with
    `near * ones_like` replaced by the packed-ray form the reference builds at run_nerf.py:319-321)."""
    sdc, sdf = _scene(7)
    mc, mf = nk.model_from_state_dict(sdc).to(DEV), nk.model_from_state_dict(sdf).to(DEV)
    o, d = _rays(41, 3)
    kw = _kw(mc, mf, 64, 128)
    with torch.no_grad():
        a = nk.render(1, 41, 1.0, chunk=1 << 15, rays=(o.to(DEV), d.to(DEV)), precision="fp32", **kw)
        b = nk.render(1, 41, 1.0, chunk=8, rays=(o.to(DEV), d.to(DEV)), precision="fp32", **kw)     # 5 x 8 + 1
        c = nk.render(1, 41, 1.0, chunk=1, rays=(o.to(DEV), d.to(DEV)), precision="fp32", **kw)
    for x, y, z in zip(a[:3], b[:3], c[:3]):
        assert torch.equal(x, y) and torch.equal(x, z)
    ref = _oracle(sdc, sdf, o, d, 64, 128, chunk=8)
    assert maxdiff(a[0], ref["rgb_fine"]) <= 1e-4


def test_degenerate_hierarchical_config_fails_loudly():
    """N_samples < 4 with importance sampling leaves sample_pdf no interior weights (weights[..., 1:-1],
    run_nerf.py:213-215): the library refuses instead of rendering something."""
    sdc, sdf = _scene(9)
    mc, mf = nk.model_from_state_dict(sdc).to(DEV), nk.model_from_state_dict(sdf).to(DEV)
    o, d = _rays(5, 4)
    with pytest.raises(nk.NerfB200Error if hasattr(nk, "NerfB200Error") else Exception):
        nk.render(1, 5, 1.0, rays=(o.to(DEV), d.to(DEV)), precision="fp32", **_kw(mc, mf, 3, 2))
