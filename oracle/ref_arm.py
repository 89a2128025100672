"""The reference arm of bench.py: the reference's OWN PyTorch code for the hot path, timed on the host cores.

TEST / BASELINE INFRASTRUCTURE ONLY (imported by bench.py's `--impl reference` and `cpu_baseline` legs and by tests).

Everything numerical below runs inside the unmodified reference package imported by oracle/reference_shim.py
(/root/reference, or its verbatim copy oracle/_ref staged by oracle/build_ref.py): `get_ray_bundle`
(nerf_helpers.py:67-110), `get_embedding_function` (:134-157), `run_one_iter_of_nerf` (train_utils.py:130-202) and
through it `predict_and_render_radiance`, `sample_pdf_2`, `volume_render_radiance_field`; loss / backward /
Adam are torch, as in train_nerf.py:271-297.  The one piece that cannot be imported is the 8x256 Keras graph of
BASELINE configs[0..1] (nerf/run_nerf_helpers.py:80-118 needs TensorFlow): it is the plain nn.Module restatement
below, pushed through the imported pipeline as `model_coarse` / `model_fine`.
"""
import os
import time

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

from . import reference_shim


class TF8x256(nn.Module):
    """init_nerf_model(D=8, W=256, input_ch=63, input_ch_views=27, skips=[4], use_viewdirs=True) as an nn.Module
    taking the concatenated embedding [P, 63 + 27] like the reference's models do (nerf/run_nerf_helpers.py:80-118:
    trunk Dense+ReLU, re-concat [pts, h] after layer 4, alpha / bottleneck(linear) / views(ReLU) / rgb heads)."""

    def __init__(self, D=8, W=256, input_ch=63, input_ch_views=27, skips=(4,), use_viewdirs=True):
        super().__init__()
        self.input_ch, self.skips, self.use_viewdirs = input_ch, tuple(skips), use_viewdirs
        ins = [input_ch] + [(W + input_ch) if i in skips else W for i in range(D - 1)]
        self.pts_linears = nn.ModuleList([nn.Linear(ins[i], W) for i in range(D)])
        if use_viewdirs:
            self.alpha_linear = nn.Linear(W, 1)
            self.bottleneck_linear = nn.Linear(W, 256)
            self.views_linears = nn.ModuleList([nn.Linear(256 + input_ch_views, W // 2)])
            self.rgb_linear = nn.Linear(W // 2, 3)
        else:
            self.output_linear = nn.Linear(W, 4)

    def forward(self, x):
        pts, views = x[..., :self.input_ch], x[..., self.input_ch:]
        h = pts
        for i, lin in enumerate(self.pts_linears):
            h = F.relu(lin(h))
            if i in self.skips:
                h = torch.cat((pts, h), dim=-1)
        if not self.use_viewdirs:
            return self.output_linear(h)
        alpha = self.alpha_linear(h)
        y = torch.cat((self.bottleneck_linear(h), views), dim=-1)
        y = F.relu(self.views_linears[0](y))
        return torch.cat((self.rgb_linear(y), alpha), dim=-1)


def available():
    return reference_shim.available()


class ReferenceTrainStep:
    """BASELINE configs[1] through the reference: one 800x800 pose, N_rand rays per step, 8x256 coarse + fine,
    64 + 128 samples, perturb on, loss + backward + Adam (lr 5e-4)."""

    def __init__(self, H=800, W=800, focal=1111.111, n_samples=64, n_importance=128, threads=None, seed=0):
        self.R = reference_shim.load()
        R = self.R
        self.threads = threads or (os.cpu_count() or 1)
        torch.set_num_threads(self.threads)
        torch.manual_seed(seed)
        self.H, self.W, self.focal = H, W, focal
        self.mc, self.mf = TF8x256(), TF8x256()
        self.opt = torch.optim.Adam(list(self.mc.parameters()) + list(self.mf.parameters()), lr=5e-4)
        self.enc_x = R.get_embedding_function(num_encoding_functions=10, include_input=True, log_sampling=True)
        self.enc_d = R.get_embedding_function(num_encoding_functions=4, include_input=True, log_sampling=True)
        self.options = reference_shim.make_options(R, use_viewdirs=True, no_ndc=True, near=2.0, far=6.0,
                                                   num_coarse=n_samples, num_fine=n_importance, chunksize=32768,
                                                   perturb=True, lindisp=False, noise=0.0, white_background=True)
        from .nerf_oracle import pose_spherical
        c2w = pose_spherical(30.0, -30.0, 4.0)
        ro, rd = R.get_ray_bundle(H, W, focal, c2w)           # the reference rebuilds this every step (train_nerf.py:240)
        self.ro, self.rd = ro.reshape(-1, 3), rd.reshape(-1, 3)
        self.rs = np.random.RandomState(1234)
        self.g = torch.Generator().manual_seed(1)

    def step(self, n_rays):
        R = self.R
        sel = torch.from_numpy(self.rs.choice(self.H * self.W, n_rays, replace=False))
        tgt = torch.rand(n_rays, 3, generator=self.g)
        t0 = time.perf_counter()
        out = R.run_one_iter_of_nerf(self.H, self.W, self.focal, self.mc, self.mf, self.ro[sel], self.rd[sel],
                                     self.options, mode="train", encode_position_fn=self.enc_x,
                                     encode_direction_fn=self.enc_d)
        rgb_c, rgb_f = out[0], out[3]
        loss = F.mse_loss(rgb_c[..., :3], tgt) + F.mse_loss(rgb_f[..., :3], tgt)   # train_nerf.py:271-285
        self.opt.zero_grad()
        loss.backward()
        self.opt.step()
        return time.perf_counter() - t0, float(loss)


class ReferenceRender:
    """BASELINE configs[0] through the reference: lego 400x400 render_only, chunk 32768, 64 + 128 samples."""

    def __init__(self, H=400, W=400, focal=555.555, n_samples=64, n_importance=128, threads=None, seed=0):
        self.R = reference_shim.load()
        R = self.R
        self.threads = threads or (os.cpu_count() or 1)
        torch.set_num_threads(self.threads)
        torch.manual_seed(seed)
        self.H, self.W, self.focal = H, W, focal
        self.mc, self.mf = TF8x256(), TF8x256()
        self.enc_x = R.get_embedding_function(num_encoding_functions=10, include_input=True, log_sampling=True)
        self.enc_d = R.get_embedding_function(num_encoding_functions=4, include_input=True, log_sampling=True)
        self.options = reference_shim.make_options(R, use_viewdirs=True, no_ndc=True, near=2.0, far=6.0,
                                                   num_coarse=n_samples, num_fine=n_importance, chunksize=32768,
                                                   perturb=False, lindisp=False, noise=0.0, white_background=True)

    def render_rows(self, c2w, row0, row1):
        """Rows [row0, row1) of one pose (a bounded sample of the frame); returns seconds, rgb_fine."""
        R = self.R
        t0 = time.perf_counter()
        with torch.no_grad():
            ro, rd = R.get_ray_bundle(self.H, self.W, self.focal, torch.as_tensor(c2w))
            ro, rd = ro[row0:row1].reshape(-1, 3), rd[row0:row1].reshape(-1, 3)
            out = R.run_one_iter_of_nerf(self.H, self.W, self.focal, self.mc, self.mf, ro, rd, self.options,
                                         mode="train", encode_position_fn=self.enc_x, encode_direction_fn=self.enc_d)
        return time.perf_counter() - t0, out[3]
