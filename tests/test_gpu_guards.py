"""GPU: out-of-bounds writes.  compute-sanitizer is closed on this pool (profiles/r2_sanitizer_unavailable.txt), so every
caller-provided buffer of the C-ABI entry points is placed between two guard bands filled with a canary pattern, sized
EXACTLY as the header / the *_workspace_bytes query says, for ragged sizes (tile counts not multiples of the CTA pair, S not a
multiple of 32, a last chunk shorter than the others); after the call the bands must be untouched and the results finite."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

import nerf_kinematics_b200 as nk  # noqa: E402
from nerf_kinematics_b200 import ops  # noqa: E402
from nerf_kinematics_b200._lib import RenderCfg, RenderGrads, RenderOut, RenderRand, check, lib  # noqa: E402
from nerf_kinematics_b200.render_core import OUT_KEYS  # noqa: E402

DEV = "cuda"
GUARD = 4096          # bytes on each side
CANARY = 0x5A


class Guarded:
    """`nbytes` usable bytes (256-byte aligned start) between two canary bands."""

    def __init__(self, nbytes):
        self.n = int(nbytes)
        pad = (256 - self.n % 256) % 256
        self.buf = torch.full((GUARD + self.n + pad + GUARD,), CANARY, dtype=torch.uint8, device=DEV)
        base = self.buf.data_ptr()
        self.off = GUARD + (256 - (base + GUARD) % 256) % 256
        assert self.off + self.n + GUARD <= self.buf.numel() + 256
        self.ptr = base + self.off

    def view(self, dtype, *shape):
        return self.buf[self.off:self.off + self.n].view(dtype).view(*shape)

    def intact(self):
        lo = self.buf[:self.off]
        hi = self.buf[self.off + self.n:]
        return bool((lo == CANARY).all()) and bool((hi == CANARY).all())


def _rays(n, use_viewdirs=True):
    g = torch.Generator().manual_seed(n)
    o = torch.randn(n, 3, generator=g) * 0.5 + torch.tensor([0, 0, 4.0])
    d = torch.nn.functional.normalize(torch.randn(n, 3, generator=g), dim=-1)
    return ops.make_ray_batch(o.to(DEV), d.to(DEV), 2.0, 6.0, use_viewdirs=use_viewdirs)


@pytest.mark.parametrize("prec", ["fp16", "fp32"])
@pytest.mark.parametrize("n,S,I", [(37, 27, 45), (301, 64, 128), (5, 9, 7)])
def test_render_rays_and_backward_stay_inside_their_buffers(prec, n, S, I):
    torch.manual_seed(0)
    mc, mf = nk.NeRF().to(DEV), nk.NeRF().to(DEV)
    hc, pid = mc.prepare(prec)
    hf, _ = mf.prepare(prec)
    rays = _rays(n)
    cfg = RenderCfg(S, I, 0, 1, pid, 1)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    ws = Guarded(lib().nb_render_workspace_bytes(hc, hf, 11, n, C.byref(cfg)))
    outs = {"rgb": Guarded(n * 12), "disp": Guarded(n * 4), "acc": Guarded(n * 4), "depth": Guarded(n * 4), "rgb0": Guarded(n * 12),
            "disp0": Guarded(n * 4), "acc0": Guarded(n * 4), "depth0": Guarded(n * 4), "z_std": Guarded(n * 4),
            "raw": Guarded(n * (S + I) * 16), "z_vals": Guarded(n * (S + I) * 4), "weights": Guarded(n * (S + I) * 4)}
    o = RenderOut(*[C.c_void_p(outs[k].ptr) for k in OUT_KEYS])
    rs = RenderRand(None, None, None, None, None, 0, 0.0)
    check(lib().nb_render_rays(hc, hf, C.c_void_p(rays.data_ptr()), 11, n, C.byref(cfg), C.byref(rs), C.byref(o),
                               C.c_void_p(ws.ptr), ws.n, st), "nb_render_rays")
    gc, gf = Guarded(mc.n_params * 4), Guarded(mf.n_params * 4)
    gc.view(torch.float32, -1).zero_()
    gf.view(torch.float32, -1).zero_()
    d_rgb = torch.randn(n, 3, device=DEV) * 1e-3
    gs = RenderGrads(C.c_void_p(d_rgb.data_ptr()), None, None, None, C.c_void_p(d_rgb.data_ptr()), None, None, None)
    check(lib().nb_render_rays_backward(hc, hf, C.c_void_p(rays.data_ptr()), 11, n, C.byref(cfg), C.byref(rs), C.byref(gs),
                                        C.c_void_p(ws.ptr), ws.n, C.c_void_p(gc.ptr), C.c_void_p(gf.ptr), st),
          "nb_render_rays_backward")
    torch.cuda.synchronize()
    for k, g in list(outs.items()) + [("workspace", ws), ("grad_coarse", gc), ("grad_fine", gf)]:
        assert g.intact(), "%s: guard band overwritten" % k
    assert bool(torch.isfinite(outs["rgb"].view(torch.float32, n, 3)).all())
    assert bool(torch.isfinite(gf.view(torch.float32, -1)).all()) and float(gf.view(torch.float32, -1).abs().max()) > 0
    # a workspace one byte short is refused, not overrun
    rc = lib().nb_render_rays(hc, hf, C.c_void_p(rays.data_ptr()), 11, n, C.byref(cfg), C.byref(rs), C.byref(o),
                              C.c_void_p(ws.ptr), ws.n - 1, st)
    assert rc == -4


@pytest.mark.parametrize("n,S,ns", [(33, 64, 128), (7, 9, 5), (129, 33, 256)])
def test_stage_kernels_stay_inside_their_buffers(n, S, ns):
    rays = _rays(n)
    z = ops.sample_coarse(rays, S)
    raw = torch.randn(n, S, 4, device=DEV)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    w, rgb, acc = Guarded(n * S * 4), Guarded(n * 12), Guarded(n * 4)
    check(lib().nb_composite(C.c_void_p(raw.data_ptr()), C.c_void_p(z.data_ptr()), C.c_void_p(rays.data_ptr()), 11, n, S, None, 1,
                             C.c_void_p(rgb.ptr), None, C.c_void_p(acc.ptr), None, C.c_void_p(w.ptr), st))
    zf, zs, sd = Guarded(n * (S + ns) * 4), Guarded(n * ns * 4), Guarded(n * 4)
    check(lib().nb_sample_fine(C.c_void_p(z.data_ptr()), C.c_void_p(w.ptr), n, S, None, ns, C.c_void_p(zs.ptr), C.c_void_p(zf.ptr),
                               C.c_void_p(sd.ptr), st))
    u = torch.rand(n, ns, device=DEV)
    zf2 = Guarded(n * (S + ns) * 4)
    check(lib().nb_sample_fine(C.c_void_p(z.data_ptr()), C.c_void_p(w.ptr), n, S, C.c_void_p(u.data_ptr()), ns, None,
                               C.c_void_p(zf2.ptr), None, st))
    draw = Guarded(n * S * 16)
    g = torch.randn(n, 3, device=DEV)
    check(lib().nb_composite_backward(C.c_void_p(raw.data_ptr()), C.c_void_p(z.data_ptr()), C.c_void_p(rays.data_ptr()), 11, n, S,
                                      None, 1, C.c_void_p(g.data_ptr()), None, None, None, C.c_void_p(draw.ptr), st))
    H, W = 7, 13
    ro, rd = Guarded(H * W * 12), Guarded(H * W * 12)
    pose = (C.c_float * 12)(*[1, 0, 0, 0.1, 0, 1, 0, 0.2, 0, 0, 1, 0.3])
    check(lib().nb_get_rays(H, W, 9.0, 8.0, 6.5, 3.5, pose, 0, H, C.c_void_p(ro.ptr), C.c_void_p(rd.ptr), st))
    torch.cuda.synchronize()
    for k, gd in (("weights", w), ("rgb", rgb), ("acc", acc), ("z_fine", zf), ("z_samples", zs), ("z_std", sd), ("z_fine(u)", zf2),
                  ("d_raw", draw), ("rays_o", ro), ("rays_d", rd)):
        assert gd.intact(), "%s: guard band overwritten" % k
    merged = zf.view(torch.float32, n, S + ns)
    assert bool((merged[:, 1:] >= merged[:, :-1]).all())
    m2 = zf2.view(torch.float32, n, S + ns)
    assert bool((m2[:, 1:] >= m2[:, :-1]).all())
