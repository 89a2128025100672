"""CPU: the oracle restatement (oracle/nerf_oracle.py) against the golden vectors generated from the
reference's own PyTorch code (oracle/make_golden.py).  Bit-exact unless a tolerance is written."""
import torch

from conftest import load_golden, maxdiff
from oracle import nerf_oracle as O

torch.set_num_threads(max(1, min(8, torch.get_num_threads())))


def test_rays_kat_and_lego():
    g = load_golden("rays")
    o, d = O.pixel_rays(3, 3, 10.0, g["kat_c2w"])
    assert maxdiff(d, g["kat_d"]) == 0 and maxdiff(o, g["kat_o"]) == 0
    o, d = O.pixel_rays(g["lego_H"], g["lego_W"], g["lego_focal"], g["lego_c2w"])
    assert maxdiff(d, g["lego_d"]) == 0 and maxdiff(o, g["lego_o"]) == 0
    # K-API: scalar focal == K with centred principal point
    K = [[g["lego_focal"], 0, g["lego_W"] / 2], [0, g["lego_focal"], g["lego_H"] / 2], [0, 0, 1]]
    o2, d2 = O.pixel_rays(g["lego_H"], g["lego_W"], None, g["lego_c2w"], K=K)
    assert maxdiff(d2, g["lego_d"]) == 0


def test_ndc():
    g = load_golden("ndc")
    o, d = O.to_ndc(g["H"], g["W"], g["focal"], 1.0, g["o"], g["d"])
    assert maxdiff(o, g["ndc_o"]) == 0 and maxdiff(d, g["ndc_d"]) == 0


def test_posenc():
    g = load_golden("posenc")
    for L in (10, 6, 4):
        assert maxdiff(O.posenc(g["x"], L), g["e%d" % L]) == 0


def test_flexible_mlp_with_checkpoint():
    g = load_golden("mlp_flexible")
    ck = load_golden("ckpt_lego_lowres3")
    sd = {k[2:]: v for k, v in ck.items() if k.startswith("c.")}
    x = torch.cat((O.posenc(g["pts"], 10), O.posenc(g["dirs"], 4)), -1)
    assert maxdiff(O.mlp_flexible(sd, x, 63, True), g["raw"]) <= 1e-6


def test_composite():
    for tag in ("c", "f"):
        g = load_golden("composite_" + tag)
        rgb, disp, acc, w, depth = O.composite(g["raw"], g["z"], g["d"], None, False)
        for nm, v in (("rgb", rgb), ("disp", disp), ("acc", acc), ("weights", w), ("depth", depth)):
            assert maxdiff(v, g[nm]) == 0, nm
        assert maxdiff(O.composite(g["raw"], g["z"], g["d"], None, True)[0], g["rgb_white"]) == 0
        # rays with sigma <= 0 everywhere: acc == 0 and disp is NaN, like the reference
        assert torch.isnan(disp[:7]).all() and float(acc[:7].abs().max()) == 0.0


def test_sample_pdf_and_searchsorted():
    g = load_golden("sample_pdf")
    assert maxdiff(O.pdf_cdf(g["weights"]), g["cdf"]) == 0
    s, i = O.invert_cdf(g["bins"], g["cdf"], g["u_det"])
    assert maxdiff(s, g["samples_det"]) == 0 and bool((i.to(torch.int32) == g["inds_det"]).all())
    s, i = O.invert_cdf(g["bins"], g["cdf"], g["u_rnd"])
    assert maxdiff(s, g["samples_rnd"]) == 0 and bool((i.to(torch.int32) == g["inds_rnd"]).all())
    assert maxdiff(O.sample_pdf(g["bins"], g["weights"], 128, det=True), g["samples_det"]) == 0
    merged, _ = torch.sort(torch.cat((g["z"], g["samples_det"]), -1), -1)
    assert maxdiff(merged, g["merged_det"]) == 0
    kat = O.searchsorted_right(torch.tensor([[0.0, 0.25, 0.25, 1.0]]), torch.tensor([[0.0, 0.25, 0.5, 1.0]]))
    assert kat.tolist() == [[1, 3, 3, 4]]


def _ckpt(name):
    ck = load_golden(name)
    return ({k[2:]: v for k, v in ck.items() if k.startswith("c.")},
            {k[2:]: v for k, v in ck.items() if k.startswith("f.")})


def test_end_to_end_lego_and_wheel():
    g = load_golden("e2e_lego")
    sc, sf = _ckpt("ckpt_lego_lowres3")
    with torch.no_grad():
        out = O.render_image("flexible", sc, sf, g["H"], g["W"], g["focal"], g["c2w"], near=2.0, far=6.0, ndc=False,
                             use_viewdirs=True, chunk=200, n_coarse=64, n_fine=128, n_freq_xyz=10, n_freq_dir=4)
    for k in ("rgb_coarse", "disp_coarse", "acc_coarse", "rgb_fine", "disp_fine", "acc_fine"):
        assert maxdiff(out[k], g[k]) <= 2e-6, k
    g = load_golden("e2e_wheel_ndc")
    sc, sf = _ckpt("ckpt_wheel_robot6")
    with torch.no_grad():
        out = O.render_image("flexible", sc, sf, g["H"], g["W"], g["focal"], g["c2w"], near=0.0, far=1.0, ndc=True,
                             use_viewdirs=True, chunk=4096, n_coarse=64, n_fine=128, n_freq_xyz=6, n_freq_dir=4)
    for k in ("rgb_coarse", "disp_coarse", "acc_coarse", "rgb_fine", "disp_fine", "acc_fine"):
        assert maxdiff(out[k], g[k]) <= 2e-6, k


def test_train_step_gradients():
    g = load_golden("train_flexible")
    sc, sf = O.init_flexible_state(seed=11), O.init_flexible_state(seed=12)
    loss, gc, gf, _ = O.train_step_grads("flexible", sc, sf, g["rays"], g["target"], n_coarse=64, n_fine=128,
                                         n_freq_xyz=10, n_freq_dir=4, white_bkgd=True)
    assert abs(float(loss) - float(g["loss"])) <= 1e-6
    for k in sc:
        assert maxdiff(gc[k], g["gc." + k]) <= 1e-6, k
        assert maxdiff(gf[k], g["gf." + k]) <= 1e-6, k


def test_tf_graph_shapes_and_flops():
    sd = O.init_tf8x256_state()
    n = sum(v.numel() for v in sd.values())
    assert n == 595844
    x = torch.randn(5, 90)
    assert O.mlp_tf8x256(sd, x, 63).shape == (5, 4)
    assert sd["pts_linears.5.weight"].shape == (256, 319)
