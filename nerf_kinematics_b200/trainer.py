"""One training iteration of the reference loop as a single object.

`TrainStep.step(rays, target)` does what the body of the reference loops does
(nerf/run_nerf.py:807-827; nerf-pytorch/train_nerf.py:256-297): render the ray batch with the coarse and
fine networks, loss = mse(rgb_fine, tgt) + mse(rgb_coarse, tgt), backward, [all-reduce], Adam step,
exponential learning-rate decay.  Unlike `loss.backward()` through render() (which also works), it
drives nb_render_rays / nb_mse_loss / nb_render_rays_backward directly into ONE flat fp32 gradient
bucket holding both networks, which is what the data-parallel all-reduce (NCCL over NVLink) sends.

Two B200-side additions over the reference loop, both optional:
  * fused Adam (nb_adam_step): one launch per network instead of torch's ~7 multi-tensor launches; the moments live
    in flat buffers that a regular torch.optim.Adam object aliases, so `optimizer.state_dict()` keeps the reference's
    checkpoint layout (train_nerf.py:401-414);
  * CUDA graph: the whole device side of an iteration (weight repack, render, loss, backward, Adam) is captured once
    per batch shape and replayed with one launch; learning rate and bias corrections are read from a small device
    array refreshed before each replay.  Inside the graph the two networks' backward passes are parallel branches
    (_device_iteration_two_streams); with more than one rank the graph ends before Adam, the whole bucket is
    all-reduced once, and a second graph holds Adam.
"""
import ctypes as C
import os
import math

import torch

from . import ops
from ._lib import RenderCfg, RenderGrads, RenderOut, RenderRand, check, lib
from .models import precision_id
from .ops import _f32, _optptr, _stream
from .render_core import OUT_KEYS, _random_inputs


def _round_up(a, b):
    return (a + b - 1) // b * b


class TrainStep:
    def __init__(self, model_coarse, model_fine, *, N_samples=64, N_importance=128, perturb=1.0, raw_noise_std=0.0,
                 white_bkgd=False, lindisp=False, precision=None, lrate=5e-4, lrate_decay=250, decay_rate=0.1,
                 eps=1e-8, betas=(0.9, 0.999), process_group=None, world_size=1, optimizer=None, fused_adam=None,
                 tf_adam=False, cuda_graph=False, device_rng=True, seed=0):
        self.mc, self.mf = model_coarse, (model_fine if N_importance > 0 else None)
        self.S, self.I = int(N_samples), int(N_importance)
        self.perturb, self.noise, self.white, self.lindisp = perturb, raw_noise_std, white_bkgd, lindisp
        self.prec = precision_id(model_coarse.precision if precision is None else precision)
        self.lrate, self.decay_steps, self.decay_rate = lrate, lrate_decay * 1000, decay_rate
        self.betas, self.eps, self.tf_adam = tuple(betas), float(eps), bool(tf_adam)
        self.pg, self.world = process_group, int(world_size)
        self.fused = (optimizer is None) if fused_adam is None else bool(fused_adam)
        self.use_graph = bool(cuda_graph) and self.fused
        self.mc.flatten_parameters()
        nc = self.mc.n_params
        nf = self.mf.n_params if self.mf is not None else 0
        dev = self.mc._flat.device
        self.dev = dev
        if self.mf is not None:
            self.mf.flatten_parameters()
        # one bucket for both networks: [coarse | pad | fine]; the fine part starts 256-byte aligned
        off_f = _round_up(nc, 64)
        self.bucket = torch.zeros((off_f + nf,), dtype=torch.float32, device=dev)
        self.gc, self.gf = self.bucket[:nc], (self.bucket[off_f:off_f + nf] if nf else None)
        # The optimizer sees the parameters in REGISTRATION order, coarse then fine, exactly like the reference's
        # `list(model_coarse.parameters()) + list(model_fine.parameters())` (train_nerf.py:165-167): optimizer state
        # dicts index parameters by position, so this is what makes a shipped `optimizer_state_dict` line up.  The flat
        # buffers keep the native order; aliasing goes by view, not by position.
        params = []
        for m, gflat in ((self.mc, self.gc), (self.mf, self.gf)):
            if m is None:
                continue
            for p, gv in m.grad_views(gflat):
                p.grad = gv
            params += list(m.parameters())
        self.params = params
        self.opt = optimizer if optimizer is not None else torch.optim.Adam(params, lr=lrate, betas=self.betas, eps=eps)
        self.global_step = 0
        self.lr_offset = 0   # schedule index = global_step + lr_offset (train_nerf.py's loop lags one iteration, resumes shift it)
        self.loss_buf = torch.zeros((2,), dtype=torch.float32, device=dev)
        # jitter / inverse-CDF uniforms / sigma noise are drawn inside the kernels (Philox, counter =
        # ray, sample, input, step) unless the caller injects arrays; {seed, step} lives on the device so that a
        # captured step draws fresh numbers on every replay.  Ranks get different seeds (different rays anyway).
        self.device_rng = bool(device_rng)
        self.rng_state = torch.tensor([int(seed), 0], dtype=torch.int64, device=dev)
        self._ws = None
        self._ws_cache = {}
        self._graphs = {}
        # world > 1: the all-reduce runs eagerly between two captured halves (forward+backward | Adam).  Capturing the
        # NCCL collective itself (NB_SPLIT_GRAPH=0) measured the same step time on 2 B200 but left the process group
        # unable to shut down cleanly.
        self._split_graph = bool(int(os.environ.get("NB_SPLIT_GRAPH", "1")))
        self._bwd_streams = bool(int(os.environ.get("NB_BWD_STREAMS", "1")))
        self._side = None
        self._overlap = bool(int(os.environ.get("NB_OVERLAP_ALLREDUCE", "1")))
        # world > 1: 1 = forward + both backward chains as ONE graph, one all-reduce of the whole bucket, Adam;
        # 0 = the three-graph form (fine all-reduce under the coarse backward)
        self._dp_chains = bool(int(os.environ.get("NB_DP_CHAINS", "1"))) and self._bwd_streams
        # the backward kernels are persistent, one 215 KB CTA per SM: leave a few SMs to NCCL so that the fine network's
        # all-reduce really runs under the coarse network's backward (NB_BWD_SM_RESERVE, default 8 when world > 1)
        if self.world > 1 and self.mf is not None and self._overlap and not self._dp_chains:
            lib().nb_set_option(b"bwd_sm_reserve", int(os.environ.get("NB_BWD_SM_RESERVE", "8")))
        self.launches_per_step = None  # kernels inside one captured iteration (graph mode)
        self.graph_replays = 0
        if self.fused:
            # Adam moments in the bucket's layout; the torch optimizer's per-parameter state aliases them
            self.exp_avg = torch.zeros_like(self.bucket)
            self.exp_avg_sq = torch.zeros_like(self.bucket)
            self._segments = [(self.mc, 0, nc)] + ([(self.mf, off_f, nf)] if nf else [])
            self._step_t = torch.zeros((), dtype=torch.float32)  # torch keeps `step` as a CPU fp32 scalar tensor
            self._alias_optimizer_state()
            # pinned staging ring for the per-step hyper-parameters (the host runs ahead of the GPU: a single slot could
            # be overwritten before its asynchronous copy has executed)
            self.hyper_host = torch.zeros((16, 8), dtype=torch.float32).pin_memory() if dev.type == "cuda" else None
            self._hyper_ev = [None] * 16
            self._hyper_i = 0
            self.hyper = torch.zeros((8,), dtype=torch.float32, device=dev)

    # ------------------------------------------------------------------ optimizer state (checkpoint layout)
    def _alias_optimizer_state(self):
        for m, off, _ in self._segments:
            for (p, _gv), (_p2, mv), (_p3, vv) in zip(m.grad_views(self.bucket[off:off + m.n_params]),
                                                      m.grad_views(self.exp_avg[off:off + m.n_params]),
                                                      m.grad_views(self.exp_avg_sq[off:off + m.n_params])):
                self.opt.state[p] = {"step": self._step_t, "exp_avg": mv, "exp_avg_sq": vv}

    def optimizer_state_dict(self):
        """torch.optim.Adam-format state (the reference checkpoint's `optimizer_state_dict`, train_nerf.py:408)."""
        if self.fused:
            self._step_t.fill_(float(self.global_step))
        return self.opt.state_dict()

    def load_optimizer_state_dict(self, sd):
        """Resume from a reference checkpoint's optimizer state (train_nerf.py:184-190)."""
        self.opt.load_state_dict(sd)
        if not self.fused:
            return
        step = 0
        for m, off, _ in self._segments:
            for (p, _g), (_p2, mv), (_p3, vv) in zip(m.grad_views(self.bucket[off:off + m.n_params]),
                                                     m.grad_views(self.exp_avg[off:off + m.n_params]),
                                                     m.grad_views(self.exp_avg_sq[off:off + m.n_params])):
                st = self.opt.state.get(p)
                if st:
                    mv.copy_(st["exp_avg"])
                    vv.copy_(st["exp_avg_sq"])
                    step = max(step, int(float(st["step"])))
        self.global_step = step
        self._alias_optimizer_state()

    def learning_rate(self, step_index):
        """lr used by iteration `step_index` (0-based): lr0 * decay_rate ** (i / decay_steps) (run_nerf.py:705-709,
        train_nerf.py:292-297 set it after iteration i-1 for iteration i)."""
        return self.lrate * (self.decay_rate ** (max(step_index + self.lr_offset, 0) / self.decay_steps))

    # ------------------------------------------------------------------ device side of one iteration
    def _workspace(self, n, stride, cfg, hc, hf, dev):
        """Workspace and output buffers per batch shape (kept alive: captured graphs hold their addresses)."""
        key = (n, stride)
        ent = self._ws_cache.get(key)
        if ent is None:
            nbytes = lib().nb_render_workspace_bytes(hc, hf, stride, n, C.byref(cfg))
            f = lambda *s: torch.empty(s, dtype=torch.float32, device=dev)
            ent = (torch.empty((max(int(nbytes), 1),), dtype=torch.uint8, device=dev),
                   {"rgb": f(n, 3), "rgb0": f(n, 3)}, {"rgb": f(n, 3), "rgb0": f(n, 3)})
            if len(self._ws_cache) >= 4 and not self._graphs:
                self._ws_cache.clear()
            self._ws_cache[key] = ent
        self._ws, self._out, self._d = ent
        return self._ws

    def forward_backward(self, rays, target, rnd=None, force_pack=False, zero_bucket=True, phase=None):
        """Render + loss + backward into the flat bucket.  Returns the (device) loss tensor [2] = (fine, coarse).
        phase splits the call: 1 = forward, losses and the fine network's backward, 2 = the coarse network's backward
        (three-graph data-parallel step: the fine gradients' all-reduce runs under it); "fwd" = forward and losses,
        "fine" / "coarse" = one network's backward (the two-chain step); None = everything."""
        rays, target = _f32(rays, "rays"), _f32(target, "target")
        n, stride, dev = rays.shape[0], rays.shape[1], rays.device
        hc, _ = self.mc.prepare(self.prec, force=force_pack)
        hf = self.mf.prepare(self.prec, force=force_pack)[0] if self.mf is not None else None
        cfg = RenderCfg(self.S, self.I, int(self.lindisp), int(self.white), self.prec, 1)
        ws = self._workspace(n, stride, cfg, hc, hf, dev)
        in_kernel = self.device_rng and rnd is None and (bool(self.perturb) or (self.noise or 0.0) > 0.0)
        if in_kernel:
            rs = RenderRand(None, None, None, None, C.c_void_p(self.rng_state.data_ptr()), int(bool(self.perturb)),
                            float(self.noise or 0.0))
        else:
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
            if phase not in (2, "coarse", "fine"):
                check(lib().nb_render_rays(hc, hf, C.c_void_p(rays.data_ptr()), stride, n, C.byref(cfg), C.byref(rs),
                                           C.byref(o), C.c_void_p(ws.data_ptr()), ws.numel(), st), "nb_render_rays")
                self.loss_buf.zero_()
                if zero_bucket:
                    self.bucket.zero_()
                check(lib().nb_mse_loss(C.c_void_p(out["rgb"].data_ptr()), C.c_void_p(target.data_ptr()), n * 3, scale,
                                        C.c_void_p(self.loss_buf.data_ptr()), C.c_void_p(self._d["rgb"].data_ptr()), st))
                if hier:
                    check(lib().nb_mse_loss(C.c_void_p(out["rgb0"].data_ptr()), C.c_void_p(target.data_ptr()), n * 3,
                                            scale, C.c_void_p(self.loss_buf.data_ptr() + 4),
                                            C.c_void_p(self._d["rgb0"].data_ptr()), st))
            gk = {}
            if phase in (None, 1, "fine"):
                gk["d_rgb"] = self._d["rgb"]
            if hier and phase in (None, 2, "coarse"):
                gk["d_rgb0"] = self._d["rgb0"]
            if gk:
                gs = RenderGrads(*[_optptr(gk.get(k)) for k in ("d_rgb", "d_disp", "d_acc", "d_depth", "d_rgb0",
                                                                "d_disp0", "d_acc0", "d_depth0")])
                check(lib().nb_render_rays_backward(hc, hf, C.c_void_p(rays.data_ptr()), stride, n, C.byref(cfg),
                                                    C.byref(rs), C.byref(gs), C.c_void_p(ws.data_ptr()), ws.numel(),
                                                    C.c_void_p(self.gc.data_ptr()), _optptr(self.gf), st),
                      "nb_render_rays_backward")
            if in_kernel and phase in (None, 2):
                check(lib().nb_rng_advance(C.c_void_p(self.rng_state.data_ptr()), st), "nb_rng_advance")
        return self.loss_buf

    def allreduce(self):
        if self.world > 1:
            import torch.distributed as dist
            dist.all_reduce(self.bucket, op=dist.ReduceOp.SUM, group=self.pg)

    def _fused_adam(self, which=None):
        """nb_adam_step per network (which = 0 coarse / 1 fine / None both); gradients are cleared by the same pass
        (the next backward accumulates)."""
        with torch.cuda.device(self.dev):
            st = _stream()
            for i, (m, off, cnt) in enumerate(self._segments):
                if which is not None and i != which:
                    continue
                check(lib().nb_adam_step(C.c_void_p(m._flat.data_ptr()), C.c_void_p(self.bucket.data_ptr() + 4 * off),
                                         C.c_void_p(self.exp_avg.data_ptr() + 4 * off),
                                         C.c_void_p(self.exp_avg_sq.data_ptr() + 4 * off), cnt,
                                         C.c_void_p(self.hyper.data_ptr()), int(self.tf_adam), 1, st), "nb_adam_step")
                m.mark_params_changed()

    def _device_iteration(self, rays, target, rnd=None):
        """Everything of one iteration that runs on the GPU, in stream order (capturable in a CUDA graph)."""
        loss = self.forward_backward(rays, target, rnd, force_pack=True, zero_bucket=False)
        self.allreduce()
        self._fused_adam()
        return loss

    def _device_iteration_two_streams(self, rays, target, with_adam=True):
        """Iteration with the two networks' backward passes as two parallel chains (after the losses they share nothing:
        own d_raw, records, partial sums and gradient buffers).  Every heavy kernel is a persistent one-CTA-per-SM launch
        over a static split of its tiles, so a kernel's SMs fall idle one by one at its end (a CTA has 10 or 11 tiles of
        the fine pass, 3 or 4 of the coarse one); with a second, independent chain in the graph the block scheduler fills
        those tails, and the gaps of the small kernels, with the other chain's CTAs.  Measured on one B200, same box:
        1.52 -> 1.46 ms per step; forking earlier (coarse backward under the fine FORWARD, the chunk renderer split in two calls) was
        slower, 1.495 vs 1.470, and a stream priority for the fine chain changed nothing.  NB_BWD_STREAMS=0 keeps the
        single chain.  with_adam=False: the data-parallel step, which all-reduces the bucket between this and Adam."""
        main = torch.cuda.current_stream(self.dev)
        if self._side is None:
            self._side = torch.cuda.Stream(self.dev)
        loss = self.forward_backward(rays, target, force_pack=True, zero_bucket=False, phase="fwd")
        fork = torch.cuda.Event()
        fork.record(main)
        self._side.wait_event(fork)
        with torch.cuda.stream(self._side):
            self.forward_backward(rays, target, force_pack=False, zero_bucket=False, phase="coarse")
            join = torch.cuda.Event()
            join.record(self._side)
        self.forward_backward(rays, target, force_pack=False, zero_bucket=False, phase="fine")
        main.wait_event(join)
        if self.device_rng and (bool(self.perturb) or (self.noise or 0.0) > 0.0):
            with torch.cuda.device(self.dev):
                check(lib().nb_rng_advance(C.c_void_p(self.rng_state.data_ptr()), _stream()), "nb_rng_advance")
        if with_adam:
            self._fused_adam()
        return loss

    def _device_iteration_overlapped(self, rays, target):
        """Data-parallel iteration as ONE capturable stream program: forward + fine backward, the fine bucket's
        all-reduce on NCCL's stream under the coarse backward, the coarse bucket's all-reduce under the fine
        network's Adam, then the coarse network's Adam.  (NB_SPLIT_GRAPH=0; the default keeps the collectives eager
        between three captured parts.)"""
        import torch.distributed as dist
        loss = self.forward_backward(rays, target, force_pack=True, zero_bucket=False, phase=1)
        w_fine = dist.all_reduce(self.gf, op=dist.ReduceOp.SUM, group=self.pg, async_op=True)
        self.forward_backward(rays, target, force_pack=False, zero_bucket=False, phase=2)
        w_coarse = dist.all_reduce(self.gc, op=dist.ReduceOp.SUM, group=self.pg, async_op=True)
        w_fine.wait()
        self._fused_adam(which=1)
        w_coarse.wait()
        self._fused_adam(which=0)
        return loss

    def close(self):
        """Drop the captured graphs (they pin NCCL resources: do this before destroy_process_group)."""
        self._graphs.clear()
        torch.cuda.synchronize(self.dev)

    def _set_hyper(self):
        """lr and bias corrections of the iteration about to run -> device (read by nb_adam_step)."""
        t = self.global_step + 1
        b1, b2 = self.betas
        vals = [self.learning_rate(self.global_step), b1, b2, self.eps, 1.0 - b1 ** t, 1.0 - b2 ** t, 1.0 - b1, 1.0 - b2]
        if self.hyper_host is not None:
            k = self._hyper_i
            self._hyper_i = (k + 1) % self.hyper_host.shape[0]
            if self._hyper_ev[k] is not None:
                self._hyper_ev[k].synchronize()
            slot = self.hyper_host[k]
            for i, v in enumerate(vals):
                slot[i] = v
            self.hyper.copy_(slot, non_blocking=True)
            if self._hyper_ev[k] is None:
                self._hyper_ev[k] = torch.cuda.Event()
            self._hyper_ev[k].record()
        else:
            self.hyper.copy_(torch.tensor(vals, dtype=torch.float32))

    def _graph_for(self, rays, target):
        key = (tuple(rays.shape), tuple(target.shape))
        g = self._graphs.get(key)
        if g is not None:
            return g
        s_rays, s_tgt = torch.empty_like(rays), torch.empty_like(target)
        s_rays.copy_(rays)
        s_tgt.copy_(target)
        # one eager iteration's worth of lazy setup (library buffers, function attributes) must precede the capture:
        # run the forward/backward once without touching parameters or moments
        self.forward_backward(s_rays, s_tgt, force_pack=True, zero_bucket=False)
        self.bucket.zero_()
        # NCCL creates its communicator (allocations, host syncs) on the first collective: that cannot happen inside a
        # capture, so run one all-reduce of the zeroed bucket eagerly first
        self.allreduce()
        torch.cuda.synchronize(self.dev)
        before = ops.launch_count()
        g = None
        if self.world == 1 or not self._split_graph:
            try:
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph):
                    # (random arrays drawn by torch cannot be shared between the forward and two backward calls: that
                    # configuration keeps the single call)
                    host_rng = not self.device_rng and (bool(self.perturb) or (self.noise or 0.0) > 0.0)
                    if self.world == 1 and self.mf is not None and self.I > 0 and self._bwd_streams and not host_rng:
                        self._device_iteration_two_streams(s_rays, s_tgt)
                    elif self.world == 1 or self.mf is None or not self._overlap:
                        self._device_iteration(s_rays, s_tgt)
                    else:
                        self._device_iteration_overlapped(s_rays, s_tgt)
                g = (graph, None, s_rays, s_tgt, None)
            except Exception:  # collective not capturable on this stack: capture the local parts instead
                if self.world == 1:
                    raise
                self._split_graph = True
                torch.cuda.synchronize(self.dev)
                self.bucket.zero_()
                before = ops.launch_count()
        if g is None:
            # forward + fine backward | coarse backward | Adam: the fine gradients' all-reduce runs under the coarse
            # backward on NCCL's stream, only the coarse gradients' all-reduce is exposed
            ga, gb = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
            # (host-drawn random arrays cannot be shared between two calls: that configuration keeps one backward)
            host_rng = not self.device_rng and (bool(self.perturb) or (self.noise or 0.0) > 0.0)
            chains = self._dp_chains and self.mf is not None and self.I > 0 and not host_rng
            gm = torch.cuda.CUDAGraph() if (self.mf is not None and self._overlap and not host_rng and not chains) else None
            with torch.cuda.graph(ga):
                if chains:   # both backward chains in one graph, ONE all-reduce of the whole bucket after it
                    self._device_iteration_two_streams(s_rays, s_tgt, with_adam=False)
                else:
                    self.forward_backward(s_rays, s_tgt, force_pack=True, zero_bucket=False, phase=1 if gm is not None else None)
            if gm is not None:
                with torch.cuda.graph(gm):
                    self.forward_backward(s_rays, s_tgt, force_pack=False, zero_bucket=False, phase=2)
            with torch.cuda.graph(gb):
                self._fused_adam()
            g = (ga, gb, s_rays, s_tgt, gm)
        self.launches_per_step = ops.launch_count() - before
        self._graphs[key] = g
        return g

    # ------------------------------------------------------------------ public
    def step(self, rays, target, rnd=None):
        """One full iteration; returns the device loss tensor [2] (fine, coarse) of this rank's batch."""
        if not self.fused:
            return self._step_torch_optimizer(rays, target, rnd)
        self._set_hyper()
        if self.use_graph and rnd is None:
            rays, target = _f32(rays, "rays"), _f32(target, "target")
            graph, graph_opt, s_rays, s_tgt, graph_mid = self._graph_for(rays, target)
            s_rays.copy_(rays, non_blocking=True)
            s_tgt.copy_(target, non_blocking=True)
            graph.replay()
            if graph_opt is not None:  # split mode: the all-reduces run eagerly between the captured parts
                if graph_mid is not None:
                    import torch.distributed as dist
                    w_fine = dist.all_reduce(self.gf, op=dist.ReduceOp.SUM, group=self.pg, async_op=True)
                    graph_mid.replay()
                    w_coarse = dist.all_reduce(self.gc, op=dist.ReduceOp.SUM, group=self.pg, async_op=True)
                    w_fine.wait()
                    w_coarse.wait()
                else:
                    self.allreduce()
                graph_opt.replay()
            self.graph_replays += 1
            loss = self.loss_buf
        else:
            loss = self._device_iteration(rays, target, rnd)
        self.global_step += 1
        lr = self.learning_rate(self.global_step)
        for gparam in self.opt.param_groups:
            gparam["lr"] = lr
        return loss

    def step_pixels(self, H, W, K, c2w, target_image, pixel_index, near, far, *, ndc=False):
        """One iteration on selected pixels of one posed image (the loop body of
        nerf-pytorch/train_nerf.py:240-297 / nerf/run_nerf.py:762-826): only the N_rand selected rays are generated
        (nb_get_rays_at) instead of the whole image's bundle, targets are gathered on the device.
        target_image: [H, W, 3] or [H*W, 3] device tensor; pixel_index: int64 [N_rand] = row * W + col."""
        idx = torch.as_tensor(pixel_index, dtype=torch.int64).to(self.dev)
        ro, rd = ops.get_rays_at(H, W, K, c2w, idx, device=self.dev)
        focal = ops.intrinsics(H, W, K)[0]
        rays = ops.make_ray_batch(ro, rd, near, far, use_viewdirs=bool(self.mc.use_viewdirs), ndc=ndc, H=H, W=W,
                                  focal=focal)
        tgt = _f32(target_image, "target_image").reshape(-1, 3)[idx]
        return self.step(rays, tgt)

    def _step_torch_optimizer(self, rays, target, rnd=None):
        """The reference's own optimizer object (any torch.optim) on the parameters' .grad views of the bucket."""
        loss = self.forward_backward(rays, target, rnd)
        self.allreduce()
        for m, gflat in ((self.mc, self.gc), (self.mf, self.gf)):
            if m is None:
                continue
            for p, gv in m.grad_views(gflat):
                if p.grad is None or p.grad.data_ptr() != gv.data_ptr():
                    p.grad = gv
        self.opt.step()
        self.global_step += 1
        new_lr = self.learning_rate(self.global_step)
        for gparam in self.opt.param_groups:
            gparam["lr"] = new_lr
        return loss
