"""Generate tests/golden/*.npz from the reference's own PyTorch code and pin the oracle.

Run in the build container only (needs /root/reference):

    python oracle/make_golden.py

For each stage it (1) runs the UNMODIFIED reference function imported through
oracle/reference_shim.py on seeded inputs, (2) runs the oracle restatement on the same
inputs and asserts equality (bit-exact unless a tolerance is printed), (3) stores inputs and
reference outputs as a small .npz fixture.  tests/test_oracle_golden.py re-checks the oracle
against these files on every CPU test run; tests/test_gpu_*.py check the CUDA path against
them on the GPU box, where /root/reference does not exist.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)

from oracle import nerf_oracle as O  # noqa: E402
from oracle import reference_shim  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
os.makedirs(GOLD, exist_ok=True)
R = reference_shim.load()
torch.set_num_threads(1)


def save(name, **arrs):
    out = {}
    for k, v in arrs.items():
        if isinstance(v, torch.Tensor):
            v = v.detach().cpu().numpy()
        out[k] = np.asarray(v)
    np.savez_compressed(os.path.join(GOLD, name + ".npz"), **out)
    sz = os.path.getsize(os.path.join(GOLD, name + ".npz"))
    print("  wrote %-28s %7.1f KB" % (name + ".npz", sz / 1024))


def check(tag, a, b, tol=0.0):
    a, b = torch.as_tensor(a), torch.as_tensor(b)
    assert a.shape == b.shape, (tag, a.shape, b.shape)
    both_nan = torch.isnan(a) & torch.isnan(b)
    d = torch.where(both_nan, torch.zeros_like(a, dtype=torch.float64), (a.double() - b.double()).abs())
    m = float(d.max()) if d.numel() else 0.0
    print("  %-44s max|diff| = %.3e (tol %.1e)" % (tag, m, tol))
    assert m <= tol, tag


def lego_pose(theta=30.0):
    return O.pose_spherical(theta, -30.0, 4.0)


# ---------------------------------------------------------------- a1 rays
def g_rays():
    print("[a1] rays")
    # KAT resurrected from the commented block nerf_helpers.py:324-343: H=W=3, f=10, c2w=2*I
    c2w = torch.eye(4)
    c2w[:3, :3] *= 2
    ro, rd = R.get_ray_bundle(3, 3, 10.0, c2w)
    oo, od = O.pixel_rays(3, 3, 10.0, c2w)
    check("kat 3x3 rays_d", od, rd)
    check("kat 3x3 rays_o", oo, ro)
    # numpy formula of the commented KAT (get_rays_np)
    i, j = np.meshgrid(np.arange(3, dtype=np.float32), np.arange(3, dtype=np.float32), indexing="xy")
    dirs = np.stack([(i - 1.5) / 10.0, -(j - 1.5) / 10.0, -np.ones_like(i)], -1)
    rd_np = np.sum(dirs[..., None, :] * c2w[:3, :3].numpy(), -1)
    check("kat 3x3 numpy vs reference", torch.from_numpy(rd_np.astype(np.float32)), rd, 1e-7)
    pose = lego_pose()
    ro2, rd2 = R.get_ray_bundle(20, 24, 555.5555, pose)
    oo2, od2 = O.pixel_rays(20, 24, 555.5555, pose)
    check("lego 20x24 rays_d", od2, rd2)
    check("lego 20x24 rays_o", oo2, ro2)
    save("rays", kat_c2w=c2w, kat_o=ro, kat_d=rd, lego_c2w=pose, lego_H=20, lego_W=24,
         lego_focal=555.5555, lego_o=ro2, lego_d=rd2)


# ---------------------------------------------------------------- a3 ndc
def g_ndc():
    print("[a3] ndc")
    g = torch.Generator().manual_seed(3)
    n = 257
    o = torch.randn(n, 3, generator=g) * 0.3
    d = torch.randn(n, 3, generator=g)
    d[:, 2] = -d[:, 2].abs() - 0.2
    H, W, f = 378, 504, 407.57
    ro, rd = R.ndc_rays(H, W, f, 1.0, o, d)
    oo, od = O.to_ndc(H, W, f, 1.0, o, d)
    check("ndc o", oo, ro)
    check("ndc d", od, rd)
    save("ndc", o=o, d=d, H=H, W=W, focal=f, ndc_o=ro, ndc_d=rd)


# ---------------------------------------------------------------- a6 encoding
def g_posenc():
    print("[a6] positional encoding")
    g = torch.Generator().manual_seed(6)
    x = (torch.rand(300, 3, generator=g) * 2 - 1) * 4.0
    for L in (10, 6, 4):
        ref = R.positional_encoding(x, num_encoding_functions=L)
        check("posenc L=%d" % L, O.posenc(x, L), ref)
    save("posenc", x=x, e10=R.positional_encoding(x, 10), e6=R.positional_encoding(x, 6),
         e4=R.positional_encoding(x, 4))


# ---------------------------------------------------------------- a8 flexible MLP + ckpt
def load_ckpt(rel):
    return torch.load(os.path.join(reference_shim.REF_ROOT, rel), map_location="cpu", weights_only=False)


def flexible_module(sd):
    dim_xyz = sd["layer1.weight"].shape[1]
    m = R.models.FlexibleNeRFModel(num_encoding_fn_xyz=(dim_xyz - 3) // 6)
    m.load_state_dict(sd)
    return m.eval()


def g_mlp_flexible():
    print("[a8] FlexibleNeRFModel (lego-lowres3 checkpoint)")
    ck = load_ckpt("trained_models/nerf-pytorch/lego-lowres3/checkpoint110000.ckpt")
    sd = ck["model_coarse_state_dict"]
    g = torch.Generator().manual_seed(8)
    pts = (torch.rand(200, 3, generator=g) * 2 - 1) * 2.5
    dirs = torch.nn.functional.normalize(torch.randn(200, 3, generator=g), dim=-1)
    x = torch.cat((R.positional_encoding(pts, 10), R.positional_encoding(dirs, 4)), -1)
    with torch.no_grad():
        ref = flexible_module(sd)(x)
        got = O.mlp_flexible(sd, x, 63, True)
    check("flexible forward", got, ref)
    fixture = {"c." + k: v for k, v in ck["model_coarse_state_dict"].items()}
    fixture.update({"f." + k: v for k, v in ck["model_fine_state_dict"].items()})
    save("ckpt_lego_lowres3", **fixture)
    save("mlp_flexible", pts=pts, dirs=dirs, raw=ref)
    ckw = load_ckpt("trained_models/nerf-pytorch/wheel_robot6/checkpoint110000.ckpt")
    fixture = {"c." + k: v for k, v in ckw["model_coarse_state_dict"].items()}
    fixture.update({"f." + k: v for k, v in ckw["model_fine_state_dict"].items()})
    save("ckpt_wheel_robot6", **fixture)


# ---------------------------------------------------------------- a9 compositing
def g_composite():
    print("[a9] compositing")
    g = torch.Generator().manual_seed(9)
    for S, tag in ((64, "c"), (192, "f")):
        n = 130
        raw = torch.randn(n, S, 4, generator=g) * 2.0
        raw[:7, :, 3] = -1.0  # rays with sigma<=0 everywhere -> acc=0, disp NaN (SURVEY hard parts)
        z, _ = torch.sort(torch.rand(n, S, generator=g) * 4 + 2, dim=-1)
        d = torch.randn(n, 3, generator=g)
        for wb in (False, True):
            ref = R.volume_render_radiance_field(raw, z, d, 0.0, wb)
            got = O.composite(raw, z, d, None, wb)
            for nm, a, b in zip(("rgb", "disp", "acc", "weights", "depth"), got, ref):
                check("composite S=%d wb=%d %s" % (S, wb, nm), a, b)
        ref = R.volume_render_radiance_field(raw, z, d, 0.0, False)
        refw = R.volume_render_radiance_field(raw, z, d, 0.0, True)
        save("composite_" + tag, raw=raw, z=z, d=d, rgb=ref[0], disp=ref[1], acc=ref[2], weights=ref[3],
             depth=ref[4], rgb_white=refw[0])


# ---------------------------------------------------------------- a10 sample_pdf
def g_sample_pdf():
    print("[a10] sample_pdf_2 / searchsorted")
    g = torch.Generator().manual_seed(10)
    n = 300
    z, _ = torch.sort(torch.rand(n, 64, generator=g) * 4 + 2, dim=-1)
    bins = 0.5 * (z[:, 1:] + z[:, :-1])
    w = torch.rand(n, 62, generator=g) ** 4
    w[:10] = 0.0          # flat pdf
    w[10:20, 5:] = 0.0    # mass in few bins -> many denom<1e-5 bins
    ref = R.sample_pdf_2(bins, w, 128, det=True)
    got = O.sample_pdf(bins, w, 128, det=True)
    check("sample_pdf det", got, ref)
    cdf = O.pdf_cdf(w)
    u_det = torch.linspace(0.0, 1.0, 128).expand(n, 128).contiguous()
    u_rnd = torch.rand(n, 128, generator=g)
    # searchsorted semantics check of SURVEY 8a diff #5: [0,.25,.25,1] vs [0,.25,.5,1] -> [1,3,3,4]
    kat = O.searchsorted_right(torch.tensor([[0.0, 0.25, 0.25, 1.0]]), torch.tensor([[0.0, 0.25, 0.5, 1.0]]))
    assert kat.tolist() == [[1, 3, 3, 4]], kat
    s_det, i_det = O.invert_cdf(bins, cdf, u_det)
    s_rnd, i_rnd = O.invert_cdf(bins, cdf, u_rnd)
    check("invert_cdf det vs reference", s_det, ref)
    # stochastic branch of the reference with the same u: monkeypatch torch.rand
    orig = torch.rand
    torch.rand = lambda *a, **k: u_rnd.clone()
    try:
        ref_rnd = R.sample_pdf_2(bins, w, 128, det=False)
    finally:
        torch.rand = orig
    check("sample_pdf rand(u given)", s_rnd, ref_rnd)
    merged, _ = torch.sort(torch.cat((z, ref), -1), -1)
    save("sample_pdf", z=z, bins=bins, weights=w, cdf=cdf, u_det=u_det, u_rnd=u_rnd,
         inds_det=i_det.to(torch.int32), inds_rnd=i_rnd.to(torch.int32), samples_det=ref,
         samples_rnd=ref_rnd, merged_det=merged)


# ---------------------------------------------------------------- end to end (B) + NDC
def g_end_to_end():
    print("[e2e] run_one_iter_of_nerf, lego-lowres3 checkpoint, 24x24, 64+128")
    ck = load_ckpt("trained_models/nerf-pytorch/lego-lowres3/checkpoint110000.ckpt")
    mc, mf = flexible_module(ck["model_coarse_state_dict"]), flexible_module(ck["model_fine_state_dict"])
    H = W = 24
    focal = 555.5555 * 24 / 400
    pose = lego_pose(40.0)
    ro, rd = R.get_ray_bundle(H, W, focal, pose)
    opts = reference_shim.make_options(R, no_ndc=True, near=2.0, far=6.0, num_coarse=64, num_fine=128,
                                       chunksize=200)
    enc_x, enc_d = R.get_embedding_function(10), R.get_embedding_function(4)
    with torch.no_grad():
        ref = R.run_one_iter_of_nerf(H, W, focal, mc, mf, ro, rd, opts, mode="validation",
                                     encode_position_fn=enc_x, encode_direction_fn=enc_d)
        got = O.render_image("flexible", ck["model_coarse_state_dict"], ck["model_fine_state_dict"], H, W,
                             focal, pose, near=2.0, far=6.0, ndc=False, use_viewdirs=True, chunk=200,
                             n_coarse=64, n_fine=128, n_freq_xyz=10, n_freq_dir=4)
    names = ("rgb_coarse", "disp_coarse", "acc_coarse", "rgb_fine", "disp_fine", "acc_fine")
    for nm, r in zip(names, ref):
        check("lego e2e " + nm, got[nm], r, 0.0)
    save("e2e_lego", H=H, W=W, focal=focal, c2w=pose, near=2.0, far=6.0,
         **{nm: r for nm, r in zip(names, ref)}, depth_fine=got["depth_fine"], z_fine=got["z_fine"])

    print("[e2e] NDC + Lx=6, wheel_robot6 checkpoint, 18x32")
    ckw = load_ckpt("trained_models/nerf-pytorch/wheel_robot6/checkpoint110000.ckpt")
    mc, mf = flexible_module(ckw["model_coarse_state_dict"]), flexible_module(ckw["model_fine_state_dict"])
    H, W, focal = 18, 32, 33.0
    pose = torch.eye(4)
    pose[:3, 3] = torch.tensor([0.05, -0.02, 0.1])
    ro, rd = R.get_ray_bundle(H, W, focal, pose)
    opts = reference_shim.make_options(R, no_ndc=False, near=0.0, far=1.0, num_coarse=64, num_fine=128,
                                       chunksize=4096, white_background=False)
    enc_x, enc_d = R.get_embedding_function(6), R.get_embedding_function(4)
    with torch.no_grad():
        ref = R.run_one_iter_of_nerf(H, W, focal, mc, mf, ro, rd, opts, mode="validation",
                                     encode_position_fn=enc_x, encode_direction_fn=enc_d)
        got = O.render_image("flexible", ckw["model_coarse_state_dict"], ckw["model_fine_state_dict"], H, W,
                             focal, pose, near=0.0, far=1.0, ndc=True, use_viewdirs=True, chunk=4096,
                             n_coarse=64, n_fine=128, n_freq_xyz=6, n_freq_dir=4)
    for nm, r in zip(names, ref):
        check("wheel ndc e2e " + nm, got[nm], r, 0.0)
    save("e2e_wheel_ndc", H=H, W=W, focal=focal, c2w=pose, near=0.0, far=1.0,
         **{nm: r for nm, r in zip(names, ref)}, depth_fine=got["depth_fine"])


# ---------------------------------------------------------------- a13 train-step gradients
def g_train():
    print("[a13] train step loss + gradients (reference autograd), 64 rays, 64+128")
    sd_c = O.init_flexible_state(seed=11)
    sd_f = O.init_flexible_state(seed=12)
    mc, mf = flexible_module(sd_c).train(), flexible_module(sd_f).train()
    H = W = 40
    focal = 55.0
    pose = lego_pose(10.0)
    ro, rd = R.get_ray_bundle(H, W, focal, pose)
    sel = np.random.RandomState(0).choice(H * W, 64, replace=False)
    ro_s, rd_s = ro.reshape(-1, 3)[sel], rd.reshape(-1, 3)[sel]
    tgt = torch.rand(64, 3, generator=torch.Generator().manual_seed(1))
    opts = reference_shim.make_options(R, no_ndc=True, near=2.0, far=6.0, num_coarse=64, num_fine=128,
                                       chunksize=4096, white_background=True)
    enc_x, enc_d = R.get_embedding_function(10), R.get_embedding_function(4)
    out = R.run_one_iter_of_nerf(H, W, focal, mc, mf, ro_s, rd_s, opts, mode="train",
                                 encode_position_fn=enc_x, encode_direction_fn=enc_d)
    loss = torch.nn.functional.mse_loss(out[0], tgt) + torch.nn.functional.mse_loss(out[3], tgt)
    loss.backward()
    rays = O.pack_rays(ro_s, rd_s, 2.0, 6.0, ndc=False, use_viewdirs=True)
    oloss, gc, gf, _ = O.train_step_grads("flexible", sd_c, sd_f, rays, tgt, n_coarse=64, n_fine=128,
                                          n_freq_xyz=10, n_freq_dir=4, white_bkgd=True)
    check("train loss", oloss, loss.detach(), 0.0)
    fx = dict(rays=rays, target=tgt, loss=loss.detach())
    for k, p in mc.named_parameters():
        check("grad coarse " + k, gc[k], p.grad, 1e-7)
        fx["gc." + k] = p.grad
    for k, p in mf.named_parameters():
        check("grad fine " + k, gf[k], p.grad, 1e-7)
        fx["gf." + k] = p.grad
    save("train_flexible", **fx)


if __name__ == "__main__":
    g_rays()
    g_ndc()
    g_posenc()
    g_mlp_flexible()
    g_composite()
    g_sample_pdf()
    g_end_to_end()
    g_train()
    print("oracle pinned against the reference on all stages.")
