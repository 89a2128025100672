"""GPU: the 16-bit tensor-core path on the 8x256 graph (the one the headline times) against the fp32 CPU oracle,
end to end at >= 10 000 rays: the north_star bound (<= 1e-2 max-abs on rgb / acc, <= 0.05 dB PSNR) for the plain,
the NDC + viewdirs (fern shape scaled down) and the lindisp configurations, plus what happens beyond the fp16 range.

The scene is a seeded random-init 8x256 pair whose density head is sharpened (alpha weights x300, bias -3) so that
acc spreads over [0.3, 1] instead of the near-empty volume a default init renders: the importance sampling and the
compositing then see realistic, peaky weights and every error of the network is amplified 300x into sigma."""
import numpy as np
import pytest
import torch

from conftest import maxdiff
from oracle import nerf_oracle as O

pytestmark = pytest.mark.gpu

import nerf_kinematics_b200 as nk  # noqa: E402
from nerf_kinematics_b200 import ops  # noqa: E402

DEV = "cuda"


def _scene(seed, lx=10):
    sd = {k: v.clone() for k, v in O.init_tf8x256_state(seed=seed, input_ch=3 + 6 * lx).items()}
    sd["alpha_linear.weight"] *= 300.0
    sd["alpha_linear.bias"] = torch.full_like(sd["alpha_linear.bias"], -3.0)
    sd["rgb_linear.weight"] *= 8.0
    return sd


def _kw(mc, mf, **over):
    kw = dict(network_query_fn=None, perturb=0.0, N_importance=128, network_fine=mf, N_samples=64, network_fn=mc,
              use_viewdirs=True, white_bkgd=True, raw_noise_std=0.0, ndc=False, lindisp=False, near=2.0, far=6.0)
    kw.update(over)
    return kw


def _psnr_delta(img16, img32, sigma=0.03):
    """PSNR cost of the 16-bit path at a ~30.5 dB operating point: the target is the fp32 image plus N(0, sigma)."""
    tgt = img32 + sigma * torch.randn(img32.shape, generator=torch.Generator().manual_seed(5))
    psnr = lambda im: float(-10.0 * torch.log10(torch.mean((im - tgt) ** 2)))
    return abs(psnr(img16) - psnr(img32)), psnr(img32)


TAU = 0.5


def _compare(H, W, focal, c2w, *, ndc, near, far, lindisp, white):
    """Three comparisons against the fp32 CPU oracle, each on what is well defined:
    (1) the COARSE pass end to end (fixed depths): the whole 16-bit network + compositing on 64 samples;
    (2) the FINE network + compositing at the oracle's own 192 sample depths (the depths are a bit-exact function of
        the coarse weights -- test_sample_pdf_golden -- but a random-init network weighs sin(2^9 x) like sin(x), so a
        6e-6 shift of a sample moves its sigma by O(0.5): end to end the fp32 GPU path itself is 3e-3 .. 1e-2 away
        from the oracle on ~2 % of the rays; a trained network is smooth at that scale, see the shipped-checkpoint
        tests in test_gpu_e2e.py);
    (3) the fine pass end to end: mean error and the fraction of rays inside the bound.
    Rays whose far-plane sigma is within TAU of zero are left out of (1) and (2): raw2outputs gives the last sample the
    distance 1e10 (run_nerf.py:118-125), so its alpha is a step function of that sigma and the reference is discontinuous
    in its own rounding there (a trained scene has an empty or a solid far plane, a random one straddles zero)."""
    sdc, sdf = _scene(0), _scene(1)
    mc, mf = nk.model_from_state_dict(sdc).to(DEV), nk.model_from_state_dict(sdf).to(DEV)
    ref = O.render_image("tf", sdc, sdf, H, W, focal, c2w, near=near, far=far, ndc=ndc, n_coarse=64, n_fine=128,
                         n_freq_xyz=10, n_freq_dir=4, white_bkgd=white, lindisp=lindisp)
    kw = _kw(mc, mf, ndc=ndc, near=near, far=far, lindisp=lindisp, white_bkgd=white)
    with torch.no_grad():
        rgb32, _, acc32, ex32 = nk.render(H, W, focal, c2w=c2w[:3, :4], precision="fp32", **kw)
        rgb16, _, acc16, ex16 = nk.render(H, W, focal, c2w=c2w[:3, :4], precision="fp16", **kw)
    assert float(ref["acc_fine"].std()) > 0.05, "degenerate scene"
    assert bool(torch.isfinite(rgb16).all()) and bool(torch.isfinite(rgb32).all())
    well_c = ref["raw_coarse"][..., -1, 3].abs() > TAU
    well_f = ref["raw_fine"][..., -1, 3].abs() > TAU
    assert float(well_c.float().mean()) > 0.6 and float(well_f.float().mean()) > 0.6
    # (1) coarse pass end to end
    c32 = maxdiff(ex32["rgb0"].cpu()[well_c], ref["rgb_coarse"][well_c])
    c16 = maxdiff(ex16["rgb0"].cpu()[well_c], ref["rgb_coarse"][well_c])
    ca16 = maxdiff(ex16["acc0"].cpu()[well_c], ref["acc_coarse"][well_c])
    dps_c, p32 = _psnr_delta(ex16["rgb0"].cpu()[well_c], ref["rgb_coarse"][well_c])
    # (2) fine network + compositing at the oracle's depths
    ro, rd = O.pixel_rays(H, W, focal, c2w)
    rays = O.pack_rays(ro, rd, near, far, H=H, W=W, focal=focal, ndc=ndc, use_viewdirs=True)
    zf = ref["z_fine"].reshape(-1, 192)
    pts = O.points_on_rays(rays[:, 0:3], rays[:, 3:6], zf)
    with torch.no_grad():
        raw16 = mf.query(pts.to(DEV), rays[:, 8:11].to(DEV), precision="fp16")
        f_rgb, _, f_acc, _, _ = ops.composite_forward(raw16, zf.to(DEV), rays.to(DEV), white_bkgd=white)
    wf = well_f.reshape(-1)
    f16 = maxdiff(f_rgb.cpu()[wf], ref["rgb_fine"].reshape(-1, 3)[wf])
    fa16 = maxdiff(f_acc.cpu()[wf], ref["acc_fine"].reshape(-1)[wf])
    dps_f, _ = _psnr_delta(f_rgb.cpu()[wf], ref["rgb_fine"].reshape(-1, 3)[wf])
    # (3) fine pass end to end
    d16 = (rgb16.cpu() - ref["rgb_fine"]).abs().max(-1)[0][well_f & well_c]
    d32 = (rgb32.cpu() - ref["rgb_fine"]).abs().max(-1)[0][well_f & well_c]
    msg = ("coarse e2e: fp32 %.2e fp16 rgb %.2e acc %.2e dPSNR %.4f dB | fine at oracle depths: fp16 rgb %.2e acc %.2e dPSNR "
           "%.4f dB | fine e2e: fp32 max %.2e, fp16 mean %.2e, inside 1e-2: %.2f %% (fp32: %.2f %%)" % (
               c32, c16, ca16, dps_c, f16, fa16, dps_f, float(d32.max()), float(d16.mean()), 100 * float((d16 <= 1e-2).float().mean()),
               100 * float((d32 <= 1e-2).float().mean())))
    print(msg)
    assert c32 <= 1e-3 and c16 <= 1e-2 and ca16 <= 1e-2 and dps_c <= 0.05, msg
    assert f16 <= 1e-2 and fa16 <= 1e-2 and dps_f <= 0.05, msg
    assert float(d16.mean()) <= 2e-3 and float((d16 <= 1e-2).float().mean()) >= 0.95, msg
    return msg


def test_tc_8x256_render_100x100_within_bound():
    _compare(100, 100, 138.9, O.pose_spherical(30.0, -30.0, 4.0), ndc=False, near=2.0, far=6.0, lindisp=False, white=True)


def test_tc_8x256_render_lindisp_within_bound():
    _compare(72, 72, 100.0, O.pose_spherical(-50.0, -20.0, 4.0), ndc=False, near=2.0, far=6.0, lindisp=True, white=False)


def test_tc_8x256_ndc_viewdirs_fern_shape_within_bound():
    """fern LLFF 504 x 378 / 4: NDC rays (near plane 1, bounds 0..1), use_viewdirs, Lx = 10 (llff_config.txt:18-20)."""
    c2w = torch.eye(4)
    c2w[:3, 3] = torch.tensor([0.05, -0.03, 0.02])
    _compare(94, 126, 101.89, c2w, ndc=True, near=0.0, far=1.0, lindisp=False, white=False)


def test_tc_activations_beyond_fp16_range_clamp_instead_of_nan():
    """Deviation from the fp32 reference, stated in DESIGN.md: the tensor-core path stores activations as IEEE fp16 with
    a saturating conversion.  A network whose hidden activations exceed 65 504 renders finite values (clamped), where
    bf16 would have kept the range and a plain fp16 conversion would have produced inf -> NaN."""
    sd = {k: v.clone() for k, v in O.init_tf8x256_state(seed=3).items()}
    sd["pts_linears.0.weight"] *= 3.0e4      # first-layer activations ~1e4..1e5
    sd["pts_linears.1.weight"] *= 30.0       # second layer far beyond 65 504 before the clamp
    m = nk.model_from_state_dict(sd).to(DEV)
    pts = (torch.rand(4096, 1, 3, generator=torch.Generator().manual_seed(0)) * 4 - 2).to(DEV)
    vd = torch.nn.functional.normalize(torch.randn(4096, 3, generator=torch.Generator().manual_seed(1)), dim=-1).to(DEV)
    with torch.no_grad():
        raw32 = m.query(pts, vd, precision="fp32")
        raw16 = m.query(pts, vd, precision="fp16")
    x = O.posenc(pts.cpu().reshape(-1, 3), 10)
    h1 = torch.relu(torch.nn.functional.linear(x, sd["pts_linears.0.weight"], sd["pts_linears.0.bias"]))
    h2 = torch.relu(torch.nn.functional.linear(h1, sd["pts_linears.1.weight"], sd["pts_linears.1.bias"]))
    assert float(h2.max()) > 65504.0 * 4                      # the premise: the fp32 network leaves the fp16 range
    assert bool(torch.isfinite(raw32).all())
    assert bool(torch.isfinite(raw16).all()), "saturating conversion must clamp, not overflow to inf/NaN"
    assert float(raw16.abs().max()) < float(raw32.abs().max()) * 1.5   # clamped activations can only shrink the output


def test_tc_points_query_per_point_viewdirs():
    """KernelMLP.query: 2-D pts [N, 3] with per-point viewdirs [N, 3] (extract_mesh-style callers) equals the
    [N, 1, 3] form; a mismatched direction count raises."""
    sd = _scene(2)
    m = nk.model_from_state_dict(sd).to(DEV)
    g = torch.Generator().manual_seed(0)
    pts = (torch.rand(1000, 3, generator=g) * 2 - 1).to(DEV)
    vd = torch.nn.functional.normalize(torch.randn(1000, 3, generator=g), dim=-1).to(DEV)
    with torch.no_grad():
        a = m.query(pts, vd, precision="fp32")
        b = m.query(pts[:, None, :], vd, precision="fp32")
    assert a.shape == (1000, 4) and torch.equal(a, b[:, 0])
    ref = O.run_mlp("tf", sd, pts.cpu()[:, None, :], vd.cpu(), 10, 4)[:, 0]
    assert maxdiff(a, ref) <= 2e-3 * max(1.0, float(ref.abs().max()))
    with pytest.raises(ValueError):
        m.query(pts, vd[:1], precision="fp32")
