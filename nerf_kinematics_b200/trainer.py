"""One training iteration of the reference loop as a single object.

`TrainStep.step(rays, target)` does what the body of the reference loops does
(nerf/run_nerf.py:807-827; nerf-pytorch/train_nerf.py:256-297): render the ray batch with the coarse and
fine networks, loss = mse(rgb_fine, tgt) + mse(rgb_coarse, tgt), backward, [all-reduce], Adam step,
exponential learning-rate decay.  Unlike `loss.backward()` through render() (which also works), it
drives nb_render_rays / nb_mse_loss / nb_render_rays_backward directly into ONE flat fp32 gradient
bucket holding both networks, which is what the data-parallel all-reduce (NCCL over NVLink) sends.
"""
import ctypes as C

import torch

from . import ops
from ._lib import RenderCfg, RenderGrads, RenderOut, RenderRand, check, lib
from .models import precision_id
from .ops import _f32, _optptr, _stream
from .render_core import OUT_KEYS, _random_inputs


class TrainStep:
    def __init__(self, model_coarse, model_fine, *, N_samples=64, N_importance=128, perturb=1.0, raw_noise_std=0.0,
                 white_bkgd=False, lindisp=False, precision=None, lrate=5e-4, lrate_decay=250, decay_rate=0.1,
                 eps=1e-8, process_group=None, world_size=1, optimizer=None):
        self.mc, self.mf = model_coarse, (model_fine if N_importance > 0 else None)
        self.S, self.I = int(N_samples), int(N_importance)
        self.perturb, self.noise, self.white, self.lindisp = perturb, raw_noise_std, white_bkgd, lindisp
        self.prec = precision_id(model_coarse.precision if precision is None else precision)
        self.lrate, self.decay_steps, self.decay_rate = lrate, lrate_decay * 1000, decay_rate
        self.pg, self.world = process_group, int(world_size)
        self.mc.flatten_parameters()
        nc = self.mc.n_params
        nf = self.mf.n_params if self.mf is not None else 0
        dev = self.mc._flat.device
        if self.mf is not None:
            self.mf.flatten_parameters()
        # one bucket for both networks: [coarse | fine]
        self.bucket = torch.zeros((nc + nf,), dtype=torch.float32, device=dev)
        self.gc, self.gf = self.bucket[:nc], (self.bucket[nc:] if nf else None)
        params = []
        for m, gflat in ((self.mc, self.gc), (self.mf, self.gf)):
            if m is None:
                continue
            for p, gv in m.grad_views(gflat):
                p.grad = gv
                params.append(p)
        self.params = params
        self.opt = optimizer if optimizer is not None else torch.optim.Adam(params, lr=lrate, betas=(0.9, 0.999), eps=eps)
        self.global_step = 0
        self.loss_buf = torch.zeros((2,), dtype=torch.float32, device=dev)
        self._ws = None
        self._ws_key = None

    def _workspace(self, n, stride, cfg, hc, hf, dev):
        key = (n, stride)
        if self._ws_key != key:
            nbytes = lib().nb_render_workspace_bytes(hc, hf, stride, n, C.byref(cfg))
            self._ws = torch.empty((max(int(nbytes), 1),), dtype=torch.uint8, device=dev)
            self._ws_key = key
            f = lambda *s: torch.empty(s, dtype=torch.float32, device=dev)
            self._out = {"rgb": f(n, 3), "rgb0": f(n, 3)}
            self._d = {"rgb": f(n, 3), "rgb0": f(n, 3)}
        return self._ws

    def forward_backward(self, rays, target, rnd=None):
        """Render + loss + backward into the flat bucket.  Returns the (device) loss tensor [2] = (fine, coarse)."""
        rays, target = _f32(rays, "rays"), _f32(target, "target")
        n, stride, dev = rays.shape[0], rays.shape[1], rays.device
        hc, _ = self.mc.prepare(self.prec)
        hf = self.mf.prepare(self.prec)[0] if self.mf is not None else None
        cfg = RenderCfg(self.S, self.I, int(self.lindisp), int(self.white), self.prec, 1)
        ws = self._workspace(n, stride, cfg, hc, hf, dev)
        rnd = _random_inputs(n, self.S, self.I, dev, self.perturb, self.noise, rnd)
        rs = RenderRand(_optptr(rnd.get("t_rand")), _optptr(rnd.get("u_fine")), _optptr(rnd.get("noise_c")),
                        _optptr(rnd.get("noise_f")))
        hier = self.I > 0
        out = {"rgb": self._out["rgb"]}
        if hier:
            out["rgb0"] = self._out["rgb0"]
        o = RenderOut(*[_optptr(out.get(k)) for k in OUT_KEYS])
        scale = 1.0 / self.world  # the mean over the global batch: each rank contributes its share
        with torch.cuda.device(dev):
            st = _stream()
            check(lib().nb_render_rays(hc, hf, C.c_void_p(rays.data_ptr()), stride, n, C.byref(cfg), C.byref(rs),
                                       C.byref(o), C.c_void_p(ws.data_ptr()), ws.numel(), st), "nb_render_rays")
            self.loss_buf.zero_()
            self.bucket.zero_()
            check(lib().nb_mse_loss(C.c_void_p(out["rgb"].data_ptr()), C.c_void_p(target.data_ptr()), n * 3, scale,
                                    C.c_void_p(self.loss_buf.data_ptr()), C.c_void_p(self._d["rgb"].data_ptr()), st))
            gk = {"d_rgb": self._d["rgb"]}
            if hier:
                check(lib().nb_mse_loss(C.c_void_p(out["rgb0"].data_ptr()), C.c_void_p(target.data_ptr()), n * 3, scale,
                                        C.c_void_p(self.loss_buf.data_ptr() + 4), C.c_void_p(self._d["rgb0"].data_ptr()),
                                        st))
                gk["d_rgb0"] = self._d["rgb0"]
            gs = RenderGrads(*[_optptr(gk.get(k)) for k in ("d_rgb", "d_disp", "d_acc", "d_depth", "d_rgb0", "d_disp0",
                                                            "d_acc0", "d_depth0")])
            check(lib().nb_render_rays_backward(hc, hf, C.c_void_p(rays.data_ptr()), stride, n, C.byref(cfg),
                                                C.byref(rs), C.byref(gs), C.c_void_p(ws.data_ptr()), ws.numel(),
                                                C.c_void_p(self.gc.data_ptr()), _optptr(self.gf), st),
                  "nb_render_rays_backward")
        return self.loss_buf

    def allreduce(self):
        if self.world > 1:
            import torch.distributed as dist
            dist.all_reduce(self.bucket, op=dist.ReduceOp.SUM, group=self.pg)

    def step(self, rays, target, rnd=None):
        """One full iteration; returns the device loss tensor [2] (fine, coarse) of this rank's batch."""
        loss = self.forward_backward(rays, target, rnd)
        self.allreduce()
        # the optimizer reads p.grad, which are views of the bucket
        for m, gflat in ((self.mc, self.gc), (self.mf, self.gf)):
            if m is None:
                continue
            for p, gv in m.grad_views(gflat):
                if p.grad is None or p.grad.data_ptr() != gv.data_ptr():
                    p.grad = gv
        self.opt.step()
        self.global_step += 1
        new_lr = self.lrate * (self.decay_rate ** (self.global_step / self.decay_steps))
        for gparam in self.opt.param_groups:
            gparam["lr"] = new_lr
        return loss
