"""One frame (or one row block of a frame) of render() as a single replayable CUDA graph.

What the reference does per pose in render_path (nerf/run_nerf.py:339-375 -> render :260-336 -> batchify_rays
:246-257 -> render_rays :48-243): build the pixel rays, pack them, walk them in `chunk`-sized pieces, concatenate.
On the reference every piece allocates its intermediates and launches hundreds of framework kernels; here the whole
chain -- nb_get_rays_dev, nb_make_ray_batch, nb_render_rays per chunk writing straight into the frame's output
arrays -- is captured ONCE per (H, W, K, rows, chunk, options) with static buffers and replayed per pose: one
graph launch per frame, no allocation, no host synchronisation.  The pose is read from device memory by the ray
kernel, so a replay renders whatever 3x4 matrix was copied there (asynchronously, from pinned memory) before it.

Multi-GPU: a rank constructs it with its own row block [row0, row1); there is no collective.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib, ops
from ._lib import RenderCfg, RenderOut, RenderRand, check, lib
from .models import precision_id
from .render_core import OUT_KEYS

_WIDTH = {"rgb": 3, "disp": 1, "acc": 1, "depth": 1, "rgb0": 3, "disp0": 1, "acc0": 1, "depth0": 1, "z_std": 1}


class FrameRenderer:
    def __init__(self, H, W, K, coarse, fine, *, chunk=1024 * 32, row0=0, row1=None, near=0.0, far=1.0, ndc=True,
                 use_viewdirs=False, N_samples=64, N_importance=0, lindisp=False, white_bkgd=False, precision=None,
                 check_numerics=False, cuda_graph=True, device=None, external_rays=False):
        _lib.require_cuda()
        self.H, self.W = int(H), int(W)
        self.fx, self.fy, self.cx, self.cy = ops.intrinsics(self.H, self.W, K)
        self.row0, self.row1 = int(row0), int(self.H if row1 is None else row1)
        self.rows = self.row1 - self.row0
        self.n = self.rows * self.W
        self.chunk = max(1, int(chunk))
        self.mc = coarse
        self.mf = fine if (fine is not None and fine is not coarse and N_importance > 0) else None
        self.S, self.I = int(N_samples), int(N_importance)
        self.near, self.far, self.ndc, self.use_viewdirs = float(near), float(far), bool(ndc), bool(use_viewdirs)
        self.lindisp, self.white = bool(lindisp), bool(white_bkgd)
        self.prec = precision_id(coarse.precision if precision is None else precision)
        self.check_numerics = bool(check_numerics)
        self.use_graph = bool(cuda_graph)
        self.external_rays = bool(external_rays)   # rays are handed in already packed (render(rays=...) form): H = 1, W = n
        coarse.flatten_parameters()
        dev = torch.device(device) if device is not None else coarse._flat.device
        self.dev = dev
        self.stride = 11 if self.use_viewdirs else 8
        f = lambda *s: torch.empty(s, dtype=torch.float32, device=dev)
        n = max(self.n, 1)
        self.pose_dev = torch.zeros(12, dtype=torch.float32, device=dev)
        # host -> device pose staging: a ring of pinned slots, each guarded by an event, so that the host running ahead of
        # the GPU (frames are GPU-bound) never overwrites a pose whose asynchronous copy has not executed yet
        self._pose_ring = torch.zeros(8, 12, dtype=torch.float32).pin_memory()
        self._pose_ev = [None] * 8
        self._pose_i = 0
        self.ro, self.rd, self.packed = f(n, 3), f(n, 3), f(n, self.stride)
        keys = ["rgb", "disp", "acc"] + (["rgb0", "disp0", "acc0", "z_std"] if self.I > 0 else [])
        self.out = {k: (f(n, 3) if _WIDTH[k] == 3 else f(n)) for k in keys}
        self.flag = torch.zeros(1, dtype=torch.int32, device=dev)
        self.flag_pin = torch.zeros(1, dtype=torch.int32).pin_memory()
        self._cfg = RenderCfg(self.S, self.I, int(self.lindisp), int(self.white), self.prec, 0)
        hc, hf = self._handles()
        with torch.cuda.device(dev):
            nbytes = lib().nb_render_workspace_bytes(hc, hf, self.stride, min(self.chunk, n), C.byref(self._cfg))
        self.ws = torch.empty((max(int(nbytes), 1),), dtype=torch.uint8, device=dev)
        self._graph = None
        self.launches_per_frame = None
        self.replays = 0

    def _handles(self):
        hc, _ = self.mc.prepare(self.prec)
        hf = self.mf.prepare(self.prec)[0] if self.mf is not None else None
        return hc, hf

    def _enqueue(self, hc, hf):
        """The frame's kernels in stream order (capturable: static pointers only)."""
        n, st = self.n, ops._stream()
        if not self.external_rays:
            check(lib().nb_get_rays_dev(self.H, self.W, self.fx, self.fy, self.cx, self.cy,
                                        C.c_void_p(self.pose_dev.data_ptr()), self.row0, self.row1,
                                        C.c_void_p(self.ro.data_ptr()), C.c_void_p(self.rd.data_ptr()), st), "nb_get_rays_dev")
            check(lib().nb_make_ray_batch(C.c_void_p(self.ro.data_ptr()), C.c_void_p(self.rd.data_ptr()), C.c_void_p(0), n,
                                          int(self.use_viewdirs), int(self.ndc), self.H, self.W, float(self.fx), self.near,
                                          self.far, C.c_void_p(0), C.c_void_p(0), C.c_void_p(self.packed.data_ptr()), st),
                  "nb_make_ray_batch")
        rs = RenderRand(None, None, None, None, None, 0, 0.0)
        for i in range(0, n, self.chunk):
            m = min(self.chunk, n - i)
            ptrs = []
            for k in OUT_KEYS:
                t = self.out.get(k)
                ptrs.append(C.c_void_p(t.data_ptr() + 4 * i * _WIDTH[k]) if t is not None else C.c_void_p(0))
            o = RenderOut(*ptrs)
            check(lib().nb_render_rays(hc, hf, C.c_void_p(self.packed.data_ptr() + 4 * i * self.stride), self.stride, m,
                                       C.byref(self._cfg), C.byref(rs), C.byref(o), C.c_void_p(self.ws.data_ptr()),
                                       self.ws.numel(), st), "nb_render_rays")
        if self.check_numerics:
            for bit, k in enumerate(self.out):
                t = self.out[k]
                check(lib().nb_check_finite(C.c_void_p(t.data_ptr()), t.numel(), C.c_void_p(self.flag.data_ptr()), 1 << bit, st),
                      "nb_check_finite")

    def render_packed(self, packed, clone=False):
        """external_rays form: `packed` [n, 8|11] rays (ops.make_ray_batch) -> dict of [n, ...] outputs."""
        assert self.external_rays and tuple(packed.shape) == (self.n, self.stride)
        with torch.cuda.device(self.dev):
            self.packed.copy_(packed, non_blocking=True)
        out = self.render(None, clone=clone)
        return {k: v.reshape([self.n] + list(v.shape[2:])) for k, v in out.items()}

    def render(self, c2w, clone=False):
        """Render the row block for the 3x4 (or 4x4) camera-to-world `c2w` (host array / CPU tensor preferred: a CUDA
        tensor is copied device-to-device without a sync).  Returns dict(rgb [rows, W, 3], disp, acc [rows, W][,
        rgb0, disp0, acc0, z_std]) -- views of the static buffers (overwritten by the next call) unless clone=True."""
        if self.n == 0:
            return {k: v[:0].reshape([0, self.W] + list(v.shape[1:])) for k, v in self.out.items()}
        with torch.cuda.device(self.dev):
            if c2w is None:
                assert self.external_rays
            elif isinstance(c2w, torch.Tensor) and c2w.is_cuda:
                self.pose_dev.copy_(c2w[:3, :4].to(torch.float32).reshape(-1))
            else:
                i = self._pose_i
                self._pose_i = (i + 1) % self._pose_ring.shape[0]
                if self._pose_ev[i] is not None:
                    self._pose_ev[i].synchronize()
                self._pose_ring[i].copy_(torch.as_tensor(np.asarray(c2w, dtype=np.float32))[:3, :4].reshape(-1))
                self.pose_dev.copy_(self._pose_ring[i], non_blocking=True)
                if self._pose_ev[i] is None:
                    self._pose_ev[i] = torch.cuda.Event()
                self._pose_ev[i].record()
            hc, hf = self._handles()   # repacks the 16-bit images only when the parameters changed (eager, outside the graph)
            if self.check_numerics:
                self.flag.zero_()
            if not self.use_graph:
                self._enqueue(hc, hf)
            else:
                if self._graph is None:
                    self._enqueue(hc, hf)   # lazy library setup must precede the capture
                    torch.cuda.synchronize(self.dev)
                    before = ops.launch_count()
                    g = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(g):
                        self._enqueue(hc, hf)
                    self.launches_per_frame = ops.launch_count() - before
                    self._graph = g
                    if self.check_numerics:
                        self.flag.zero_()
                self._graph.replay()
                self.replays += 1
            if self.check_numerics:
                self.flag_pin.copy_(self.flag, non_blocking=True)
                torch.cuda.current_stream().synchronize()   # the reference's check raises per render_rays call
                bits = int(self.flag_pin[0])
                if bits:
                    bad = [k for b, k in enumerate(self.out) if bits & (1 << b)]
                    raise FloatingPointError("output %s : Tensor had NaN or Inf values" % ", ".join(bad))
        res = {k: v[:self.n].reshape([self.rows, self.W] + list(v.shape[1:])) for k, v in self.out.items()}
        return {k: v.clone() for k, v in res.items()} if clone else res
