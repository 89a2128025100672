"""Diagnostic: fp16 / fp32 GPU renders of the synthetic 8x256 scenes vs the CPU oracle, by conditioning mask."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import nerf_kinematics_b200 as nk
from oracle import nerf_oracle as O

def scene(seed, aw, ab):
    sd = {k: v.clone() for k, v in O.init_tf8x256_state(seed=seed).items()}
    sd["alpha_linear.weight"] *= aw
    sd["alpha_linear.bias"] = torch.full_like(sd["alpha_linear.bias"], ab)
    sd["rgb_linear.weight"] *= 8.0
    return sd

H = W = 64
for aw, ab in ((100, -1.0), (300, -3.0)):
    for ndc in (False, True):
        sdc, sdf = scene(0, aw, ab), scene(1, aw, ab)
        mc, mf = nk.model_from_state_dict(sdc).cuda(), nk.model_from_state_dict(sdf).cuda()
        if ndc:
            c2w = torch.eye(4); c2w[:3, 3] = torch.tensor([0.05, -0.03, 0.02]); near, far, focal = 0.0, 1.0, 52.0
        else:
            c2w = O.pose_spherical(30.0, -30.0, 4.0); near, far, focal = 2.0, 6.0, 88.0
        ref = O.render_image("tf", sdc, sdf, H, W, focal, c2w, near=near, far=far, ndc=ndc, n_coarse=64, n_fine=128, n_freq_xyz=10,
                             n_freq_dir=4, white_bkgd=True)
        kw = dict(network_query_fn=None, perturb=0.0, N_importance=128, network_fine=mf, N_samples=64, network_fn=mc, use_viewdirs=True,
                  white_bkgd=True, raw_noise_std=0.0, ndc=ndc, lindisp=False, near=near, far=far)
        with torch.no_grad():
            r32 = nk.render(H, W, focal, c2w=c2w[:3, :4], precision="fp32", **kw)
            r16 = nk.render(H, W, focal, c2w=c2w[:3, :4], precision="fp16", **kw)
        sc, sf = ref["raw_coarse"][..., -1, 3], ref["raw_fine"][..., -1, 3]
        d32 = (r32[0].cpu() - ref["rgb_fine"]).abs().max(-1)[0]
        d16 = (r16[0].cpu() - ref["rgb_fine"]).abs().max(-1)[0]
        a16 = (r16[2].cpu() - ref["acc_fine"]).abs()
        c16 = (r16[3]["rgb0"].cpu() - ref["rgb_coarse"]).abs().max(-1)[0]
        print("scene x%d b%.1f ndc=%d: sigma std %.2f; all rays: fp32 max %.2e fp16 max %.2e mean %.2e" % (aw, ab, ndc, float(ref["raw_fine"][..., 3].std()), d32.max(), d16.max(), d16.mean()))
        for tau in (0.0, 0.3, 1.0, 2.0):
            w = (sc.abs() > tau) & (sf.abs() > tau)
            if w.sum() == 0: continue
            print("   tau %.1f well %.2f: fp32 max %.2e | fp16 rgb max %.2e (>1e-2: %d of %d) acc max %.2e rgb0 max %.2e" % (
                tau, float(w.float().mean()), d32[w].max(), d16[w].max(), int((d16[w] > 1e-2).sum()), int(w.sum()), a16[w].max(), c16[w].max()))
