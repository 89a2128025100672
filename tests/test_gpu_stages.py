"""GPU: every stage kernel, called through the C-ABI, against the golden vectors generated from the
reference and against the CPU oracle on seeded inputs.  Tolerances are the north_star's:
rays 1e-6, searchsorted indices bit-exact, compositing 1e-3 (observed ~1e-6)."""
import numpy as np
import pytest
import torch

from conftest import load_golden, maxdiff
from oracle import nerf_oracle as O

pytestmark = pytest.mark.gpu

import nerf_kinematics_b200 as nk  # noqa: E402
from nerf_kinematics_b200 import ops, torchsearchsorted  # noqa: E402

DEV = "cuda"


def test_get_rays_kat_golden_and_K():
    g = load_golden("rays")
    o, d = nk.get_ray_bundle(3, 3, 10.0, g["kat_c2w"])
    assert maxdiff(d, g["kat_d"]) <= 1e-6 and maxdiff(o, g["kat_o"]) == 0
    H, W, f = g["lego_H"], g["lego_W"], g["lego_focal"]
    o, d = nk.get_ray_bundle(H, W, f, g["lego_c2w"].to(DEV))
    assert o.shape == (H, W, 3)  # (H, W, 3), not the docstring's (W, H, 3)
    assert maxdiff(d, g["lego_d"]) <= 1e-6 and maxdiff(o, g["lego_o"]) == 0
    # row-block form used by the multi-GPU partitioning: identical values
    o2, d2 = ops.get_rays(H, W, f, g["lego_c2w"], row0=5, row1=11)
    assert torch.equal(d2, d[5:11]) and torch.equal(o2, o[5:11])
    # K-API with fx != fy and an off-centre principal point (wheel: fl_x 674.42, fl_y 649.46)
    K = np.array([[674.419, 0, 640.0], [0, 649.457, 360.0], [0, 0, 1]])
    c2w = O.pose_spherical(20.0, -30.0, 4.0)
    ro, rd = O.pixel_rays(72, 128, None, c2w, K=K / 10)
    o3, d3 = nk.get_rays(72, 128, K / 10, c2w)
    assert maxdiff(d3, rd) <= 1e-6 and maxdiff(o3, ro) == 0
    # empty row block
    o4, d4 = ops.get_rays(H, W, f, g["lego_c2w"], row0=3, row1=3)
    assert o4.shape == (0, W, 3)


def test_ndc_rays_golden():
    g = load_golden("ndc")
    o, d = nk.ndc_rays(g["H"], g["W"], g["focal"], 1.0, g["o"].to(DEV), g["d"].to(DEV))
    assert maxdiff(o, g["ndc_o"]) <= 1e-6 and maxdiff(d, g["ndc_d"]) <= 1e-6


def test_ray_packing_matches_oracle():
    g = load_golden("ndc")
    for ndc in (False, True):
        for uv in (False, True):
            ref = O.pack_rays(g["o"], g["d"], 0.0 if ndc else 2.0, 1.0 if ndc else 6.0, H=g["H"], W=g["W"],
                              focal=g["focal"], ndc=ndc, use_viewdirs=uv)
            got = ops.make_ray_batch(g["o"].to(DEV), g["d"].to(DEV), 0.0 if ndc else 2.0, 1.0 if ndc else 6.0,
                                     use_viewdirs=uv, ndc=ndc, H=g["H"], W=g["W"], focal=g["focal"])
            assert got.shape == ref.shape
            assert maxdiff(got, ref) <= 1e-6


def test_stratified_depths():
    g = torch.Generator().manual_seed(5)
    n = 77
    near = torch.rand(n, 1, generator=g) + 1.5
    far = near + torch.rand(n, 1, generator=g) * 4 + 0.5
    rays = torch.zeros(n, 8)
    rays[:, 6:7], rays[:, 7:8] = near, far
    tr = torch.rand(n, 64, generator=g)
    for lindisp in (False, True):
        for t in (None, tr):
            ref = O.coarse_depths(near, far, 64, lindisp, t)
            got = ops.sample_coarse(rays.to(DEV), 64, lindisp, t.to(DEV) if t is not None else None)
            assert maxdiff(got, ref) <= (0 if (t is None and not lindisp) else 1e-6), (lindisp, t is None)
    for S in (2, 33, 128):  # torch.linspace bit pattern for several lengths
        assert maxdiff(ops.sample_coarse(rays.to(DEV), S), O.coarse_depths(near, far, S)) == 0


def test_posenc_golden():
    g = load_golden("posenc")
    for L in (10, 6, 4):
        got = nk.positional_encoding(g["x"].to(DEV), L)
        assert got.shape == g["e%d" % L].shape
        # sin/cos arguments reach 2^9*4: CUDA sinf/cosf vs the CPU libm differ in the last ulp
        assert maxdiff(got, g["e%d" % L]) <= 5e-7


def test_composite_golden_and_nan_rays():
    for tag in ("c", "f"):
        g = load_golden("composite_" + tag)
        n = g["raw"].shape[0]
        rays = torch.zeros(n, 8)
        rays[:, 3:6] = g["d"]
        for white, key in ((False, "rgb"), (True, "rgb_white")):
            rgb, disp, acc, w, depth = ops.composite_forward(g["raw"].to(DEV), g["z"].to(DEV), rays.to(DEV), None, white)
            assert maxdiff(rgb, g[key]) <= 1e-5
            assert maxdiff(acc, g["acc"]) <= 1e-5 and maxdiff(depth, g["depth"]) <= 2e-5
            assert maxdiff(w, g["weights"]) <= 1e-6
            assert torch.isnan(disp[:7]).all()  # sigma <= 0 everywhere -> 0/0 like the reference
            d_ok = ~torch.isnan(g["disp"])
            rel = ((disp.cpu()[d_ok] - g["disp"][d_ok]).abs() / g["disp"][d_ok].abs().clamp_min(1e-6)).max()
            assert float(rel) <= 1e-4
    # reference signature wrapper + empty batch
    g = load_golden("composite_c")
    out = nk.volume_render_radiance_field(g["raw"].to(DEV), g["z"].to(DEV), g["d"].to(DEV), 0.0, True)
    assert maxdiff(out[0], g["rgb_white"]) <= 1e-5
    e = ops.composite_forward(torch.zeros(0, 64, 4, device=DEV), torch.zeros(0, 64, device=DEV),
                              torch.zeros(0, 8, device=DEV))
    assert e[0].shape == (0, 3)


def test_composite_backward_matches_autograd_of_the_oracle():
    g = torch.Generator().manual_seed(21)
    for S, white in ((64, True), (192, False), (40, True)):
        n = 50
        raw = (torch.randn(n, S, 4, generator=g) * 1.5).requires_grad_(True)
        z, _ = torch.sort(torch.rand(n, S, generator=g) * 4 + 2, dim=-1)
        d = torch.randn(n, 3, generator=g)
        noise = torch.randn(n, S, generator=g) * 0.3
        rgb, disp, acc, w, depth = O.composite(raw, z, d, noise, white)
        gr, ga, gd = torch.randn(n, 3, generator=g), torch.randn(n, generator=g), torch.randn(n, generator=g)
        (rgb * gr).sum().add((acc * ga).sum()).add((depth * gd).sum()).backward()
        rays = torch.zeros(n, 8)
        rays[:, 3:6] = d
        got = ops.composite_backward(raw.detach().to(DEV), z.to(DEV), rays.to(DEV), noise.to(DEV), white,
                                     gr.to(DEV), None, ga.to(DEV), gd.to(DEV))
        scale = float(raw.grad.abs().max())
        assert maxdiff(got, raw.grad) <= 2e-5 * max(1.0, scale)
    # disparity gradient
    raw = (torch.randn(20, 64, 4, generator=g) + 1.0).requires_grad_(True)
    z, _ = torch.sort(torch.rand(20, 64, generator=g) * 4 + 2, dim=-1)
    d = torch.randn(20, 3, generator=g)
    disp = O.composite(raw, z, d)[1]
    gdisp = torch.randn(20, generator=g)
    (disp * gdisp).sum().backward()
    rays = torch.zeros(20, 8)
    rays[:, 3:6] = d
    got = ops.composite_backward(raw.detach().to(DEV), z.to(DEV), rays.to(DEV), None, False, None, gdisp.to(DEV))
    assert maxdiff(got, raw.grad) <= 2e-5 * max(1.0, float(raw.grad.abs().max()))


def test_searchsorted_bit_exact_and_shim_signature():
    g = load_golden("sample_pdf")
    cdf = g["cdf"].to(DEV)
    for key_u, key_i in (("u_det", "inds_det"), ("u_rnd", "inds_rnd")):
        inds = torchsearchsorted.searchsorted(cdf, g[key_u].to(DEV), side="right")
        assert inds.dtype == torch.int64
        assert torch.equal(inds.cpu().to(torch.int32), g[key_i])
    a = torch.tensor([[0.0, 0.25, 0.25, 1.0]], device=DEV)
    v = torch.tensor([[0.0, 0.25, 0.5, 1.0]], device=DEV)
    assert torchsearchsorted.searchsorted(a, v, side="right").tolist() == [[1, 3, 3, 4]]
    assert torchsearchsorted.searchsorted(a, v, side="left").tolist() == [[0, 1, 3, 3]]
    out = torch.empty(1, 4, dtype=torch.int64, device=DEV)
    torchsearchsorted.searchsorted(a, v, out=out, side="right")
    assert out.tolist() == [[1, 3, 3, 4]]
    # ragged: one row of a broadcast over many rows of v
    vv = torch.rand(9, 5, device=DEV)
    ref = torch.searchsorted(a.cpu().expand(9, 4).contiguous(), vv.cpu(), right=True)
    assert torch.equal(torchsearchsorted.searchsorted(a, vv, side="right").cpu(), ref)


def _close_fraction(a, b, tol=1e-4):
    return float(((a.cpu() - b.cpu()).abs() <= tol).float().mean())


def test_sample_pdf_golden():
    """sample_pdf_2 (nerf_helpers.py:260-302) bit for bit: the kernel normalises the pdf with torch.sum's own fp32
    summation order (csrc/volume.cu warp_sum_torch_order) and accumulates the CDF in double like torch.cumsum, so EVERY
    row's fp32 CDF equals the reference's and every sample follows; the residual is asserted, not tolerated."""
    g = load_golden("sample_pdf")
    bins, w = g["bins"].to(DEV), g["weights"].to(DEV)
    s, cdf = ops.sample_pdf(bins, w, 128, det=True, return_cdf=True)
    same = (cdf.cpu() == g["cdf"]).all(dim=-1)
    assert int((~same).sum()) == 0, "%d of %d rows have a CDF that differs from torch's" % (int((~same).sum()), same.numel())
    assert int((s.cpu() != g["samples_det"]).sum()) == 0
    s2 = ops.sample_pdf(bins, w, 128, u=g["u_rnd"].to(DEV))
    assert int((s2.cpu() != g["samples_rnd"]).sum()) == 0
    assert torch.equal(nk.sample_pdf_2(bins, w, 128, det=True).cpu(), g["samples_det"])


def test_pdf_normalisation_reproduces_torch_sum_order():
    """torch.sum(x, -1) in fp32 is not an exactly rounded sum: the kernel must follow ATen's 8-lane x 4-accumulator
    order for every row length (vector groups, left-over vectors, scalar tail, rows shorter than one vector)."""
    gen = torch.Generator().manual_seed(11)
    for nb in (3, 5, 8, 9, 10, 17, 31, 33, 40, 63, 64, 96, 127, 191, 255):
        n = 257
        w = torch.rand(n, nb - 1, generator=gen) ** 4
        w[::7] *= 1e-4                                   # rows dominated by the +1e-5 floor
        bins = torch.sort(torch.rand(n, nb, generator=gen), -1)[0]
        ww = w + 1e-5
        ref_cdf = torch.cat((torch.zeros(n, 1), torch.cumsum(ww / torch.sum(ww, -1, keepdim=True), -1)), -1)
        _, cdf = ops.sample_pdf(bins.to(DEV), w.to(DEV), 16, det=True, return_cdf=True)
        bad = int((cdf.cpu() != ref_cdf).any(dim=-1).sum())
        assert bad == 0, "nb=%d: %d of %d rows differ" % (nb, bad, n)


def test_sample_fine_merge_and_std():
    g = load_golden("sample_pdf")
    z = g["z"].to(DEV)
    n = z.shape[0]
    wfull = torch.zeros(n, 64)
    wfull[:, 1:-1] = g["weights"]
    zf, zs, sd = ops.sample_fine(z, wfull.to(DEV), 128, want_samples=True, want_std=True)
    # the fused kernel is the standalone sample_pdf + merge, bit for bit
    s_alone = ops.sample_pdf(g["bins"].to(DEV), g["weights"].to(DEV), 128, det=True)
    assert torch.equal(zs, s_alone)
    assert torch.equal(zs.cpu(), g["samples_det"])
    ref, _ = torch.sort(torch.cat((g["z"], zs.cpu()), -1), -1)
    assert torch.equal(zf.cpu(), ref)
    assert bool((zf[:, 1:] >= zf[:, :-1]).all())
    assert maxdiff(sd, zs.cpu().std(dim=-1, unbiased=False)) <= 1e-5
    # stochastic u: samples are unsorted, the merge must still be the full sort
    zf2, zs2, _ = ops.sample_fine(z, wfull.to(DEV), 128, u=g["u_rnd"].to(DEV), want_samples=True)
    ref, _ = torch.sort(torch.cat((g["z"], zs2.cpu()), -1), -1)
    assert torch.equal(zf2.cpu(), ref)
    m = ops.merge_sorted(z, torch.sort(zs2, -1)[0])
    assert torch.equal(m.cpu(), ref)


def test_sample_fine_two_ray_kernel_equals_one_ray_kernel():
    """The two-rays-per-warp kernel (64 depths, 128 / 64 deterministic samples; csrc/volume.cu sample_fine2_kernel) against
    the one-ray-per-warp kernel it replaces, bit for bit: peaked, flat and floor-dominated weights, repeated depths (the
    rank fix-up walk), an odd ray count (a warp's second ray missing) and every optional output."""
    from nerf_kinematics_b200._lib import lib
    gen = torch.Generator().manual_seed(5)
    n = 4097
    z = torch.sort(torch.rand(n, 64, generator=gen) * 4 + 2, -1)[0]
    z[::5] = torch.linspace(2.0, 6.0, 64)                       # the render path's depths
    z[3::11, 20:24] = z[3::11, 20:21]                           # repeated depths
    z[7::13, 40:] = z[7::13, 40:41]                             # a constant tail (zero-width bins)
    w = torch.rand(n, 64, generator=gen) ** 6
    w[::7] = 0.0                                                # only the 1e-5 floor: a uniform pdf
    w[1::7] *= 1e-6
    w[2::7, :30] = 0.0                                          # mass behind an empty front
    w[4::7, 33] = 50.0                                          # one dominant bin
    z, w = z.to(DEV), w.to(DEV)
    for ns in (128, 64):
        res = {}
        for v in (0, 1):
            old = lib().nb_set_option(b"sample_fine2", v)
            try:
                res[v] = (ops.sample_fine(z, w, ns, want_samples=True, want_std=True), ops.sample_fine(z, w, ns))
            finally:
                lib().nb_set_option(b"sample_fine2", old)
        (zf0, zs0, sd0), (zf0b, _, _) = res[0]
        (zf1, zs1, sd1), (zf1b, _, _) = res[1]
        assert torch.equal(zf1, zf0) and torch.equal(zs1, zs0) and torch.equal(zf1b, zf0)
        assert maxdiff(sd1, sd0) <= 2e-6                        # a different summation order
        ref, _ = torch.sort(torch.cat((z, zs1), -1), -1)
        assert torch.equal(zf1, ref)
        # random u (the train step): the two-ray kernel sorts the uniforms first, the one-ray kernel sorts the samples
        u = torch.rand(n, ns, generator=gen)
        u[5] = 0.0
        u[6, ::2] = u[6, 1::2]                                  # repeated uniforms
        u = u.to(DEV)
        rnd = {}
        for v in (0, 1):
            old = lib().nb_set_option(b"sample_fine2", v)
            try:
                rnd[v] = ops.sample_fine(z, w, ns, u=u, want_std=True)
            finally:
                lib().nb_set_option(b"sample_fine2", old)
        assert torch.equal(rnd[1][0], rnd[0][0])
        assert maxdiff(rnd[1][2], rnd[0][2]) <= 2e-6
        _, zs_u, _ = ops.sample_fine(z, w, ns, u=u, want_samples=True)   # (one-ray kernel: samples in the caller's order)
        ref, _ = torch.sort(torch.cat((z, zs_u), -1), -1)
        assert torch.equal(rnd[1][0], ref)


def test_mse_loss():
    g = torch.Generator().manual_seed(3)
    a, b = torch.rand(1000, 3, generator=g), torch.rand(1000, 3, generator=g)
    loss, d = ops.mse_loss(a.to(DEV), b.to(DEV))
    assert abs(float(loss) - float(torch.nn.functional.mse_loss(a, b))) <= 1e-6
    assert maxdiff(d, 2 * (a - b) / a.numel()) <= 1e-9


def test_get_rays_wide_rows_and_selected_pixels():
    """Images wider than one 256-pixel block with a ragged last block (1280x720 wheel frame, 504x378 fern), and the
    device-side ray selection of SURVEY 8(f) f1: bit-identical to gathering the full bundle."""
    c2w = O.pose_spherical(17.0, -30.0, 4.0)
    for (H, W, K) in ((72, 1280, torch.tensor([[674.419, 0, 640.0], [0, 649.457, 36.0], [0, 0, 1]])), (38, 504, 407.57)):
        f = K if not isinstance(K, torch.Tensor) else None
        o, d = ops.get_rays(H, W, K, c2w)
        if f is not None:
            ro, rd = O.pixel_rays(H, W, f, c2w)
            assert maxdiff(d, rd) <= 1e-6 and maxdiff(o, ro) == 0
        else:  # K-API: columns / rows independent -> check against the closed form in fp64
            jj, ii = torch.meshgrid(torch.arange(H, dtype=torch.float64), torch.arange(W, dtype=torch.float64), indexing="ij")
            dirs = torch.stack([(ii - 640.0) / 674.419, -(jj - 36.0) / 649.457, -torch.ones_like(ii)], -1)
            ref = (dirs[..., None, :] * torch.as_tensor(c2w, dtype=torch.float64)[:3, :3]).sum(-1)
            assert maxdiff(d.double(), ref) <= 1e-6
        o2, d2 = ops.get_rays(H, W, K, c2w, row0=3, row1=H - 2)
        assert torch.equal(d2, d[3:H - 2]) and torch.equal(o2, o[3:H - 2])
        sel = torch.from_numpy(np.random.RandomState(0).choice(H * W, 1024, replace=False)).to(DEV)
        o3, d3 = ops.get_rays_at(H, W, K, c2w, sel)
        assert torch.equal(d3, d.reshape(-1, 3)[sel]) and torch.equal(o3, o.reshape(-1, 3)[sel])


def test_to8b_and_psnr_output_stage():
    """render_path's output stage (run_nerf.py:358-373): to8b truncates like numpy astype(uint8); PSNR vs gt."""
    g = torch.Generator().manual_seed(5)
    img = (torch.rand(37, 53, 3, generator=g) * 1.4 - 0.2)
    gt = torch.rand(37, 53, 3, generator=g)
    u8, psnr = ops.to8b_psnr(img.to(DEV), gt.to(DEV))
    ref8 = (255 * np.clip(img.numpy(), 0, 1)).astype(np.uint8)
    assert np.array_equal(u8.cpu().numpy(), ref8)
    ref_psnr = -10.0 * np.log10(np.mean(np.square(img.numpy() - gt.numpy())))
    assert abs(float(psnr) - ref_psnr) <= 1e-4
    u8b, none = ops.to8b_psnr(img.to(DEV))
    assert none is None and torch.equal(u8b, u8)
