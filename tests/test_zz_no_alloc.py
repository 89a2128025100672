"""GPU: nb_render_rays_static (include/nerf_b200.h) -- the allocation-free variant of nb_render_rays whose intermediates
live in a static 16 MiB device array inside the library -- against nb_render_rays with a caller workspace, through the
C-ABI, on two calls in a row (the second call finds the first one's contents in the shared array), for both precisions;
and its size limit.  Collected last on purpose."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

import nerf_kinematics_b200 as nk  # noqa: E402
from nerf_kinematics_b200 import ops  # noqa: E402
from nerf_kinematics_b200._lib import NerfB200Error, RenderCfg, RenderOut, RenderRand, check, last_error, lib  # noqa: E402
from nerf_kinematics_b200.render_core import OUT_KEYS  # noqa: E402

DEV = "cuda"
KEYS = ("rgb", "disp", "acc", "rgb0", "disp0", "acc0", "z_std")


def _models():
    torch.manual_seed(17)
    mc, mf = nk.NeRF().to(DEV), nk.NeRF().to(DEV)
    with torch.no_grad():   # a volume with structure (see tests/test_gpu_tc_parity.py)
        for m in (mc, mf):
            m.alpha_linear.weight.mul_(20.0)
            m.alpha_linear.bias.fill_(-1.0)
            m.rgb_linear.weight.mul_(8.0)
    return mc, mf


def _rays(n, seed):
    g = torch.Generator().manual_seed(seed)
    o = (torch.randn(n, 3, generator=g) * 0.3 + torch.tensor([0.0, 0.0, 4.0])).to(DEV)
    d = (torch.randn(n, 3, generator=g) * 0.3 + torch.tensor([0.0, 0.0, -1.0])).to(DEV)
    return ops.make_ray_batch(o, d, 2.0, 6.0, use_viewdirs=True)


def _outputs(n):
    f = lambda *s: torch.full(s, float("nan"), dtype=torch.float32, device=DEV)
    t = {k: (f(n, 3) if k.startswith("rgb") else f(n)) for k in KEYS}
    ptrs = [C.c_void_p(t[k].data_ptr()) if k in t else C.c_void_p(0) for k in OUT_KEYS]
    return t, RenderOut(*ptrs)


def _call(mc, mf, rays, precision, static):
    n = rays.shape[0]
    hc, prec = mc.prepare(precision)
    hf, _ = mf.prepare(precision)
    cfg = RenderCfg(64, 128, 0, 1, prec, 0)
    rs = RenderRand(None, None, None, None, None, 0, 0.0)
    t, o = _outputs(n)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    if static:
        check(lib().nb_render_rays_static(hc, hf, C.c_void_p(rays.data_ptr()), rays.shape[1], n, C.byref(cfg), C.byref(rs),
                                          C.byref(o), st), "nb_render_rays_static")
    else:
        nbytes = lib().nb_render_workspace_bytes(hc, hf, rays.shape[1], n, C.byref(cfg))
        ws = torch.empty((max(int(nbytes), 1),), dtype=torch.uint8, device=DEV)
        check(lib().nb_render_rays(hc, hf, C.c_void_p(rays.data_ptr()), rays.shape[1], n, C.byref(cfg), C.byref(rs),
                                   C.byref(o), C.c_void_p(ws.data_ptr()), ws.numel(), st), "nb_render_rays")
    torch.cuda.synchronize()
    return t


@pytest.mark.parametrize("precision", ["fp16", "fp32"])
def test_static_scratch_render_equals_workspace_render_two_calls_in_a_row(precision):
    assert lib().nb_render_static_scratch_bytes() == 16 << 20
    mc, mf = _models()
    # two different batches of different sizes back to back: the second call runs over the first one's leftovers
    # (fp32 keeps its layer activations in the scratch as well, hence the smaller batches)
    sizes = (1500, 777) if precision == "fp16" else (16, 9)
    for i, n in enumerate(sizes):
        rays = _rays(n, 40 + i)
        ref = _call(mc, mf, rays, precision, static=False)
        got = _call(mc, mf, rays, precision, static=True)
        for k in KEYS:
            assert bool(torch.isfinite(ref[k]).all()), k
            # same kernels in the same order on the same inputs: the repository's render-vs-render tests use torch.equal
            assert torch.equal(got[k], ref[k]), (precision, n, k, float((got[k] - ref[k]).abs().max()))
        assert float(ref["rgb"].std()) > 1e-3 and float(ref["z_std"].min()) > 0.0   # an image, not a constant


def test_static_scratch_rejects_batches_beyond_its_size():
    mc, mf = _models()
    rays = _rays(8192, 3)   # 8 192 rays x (64 + 192) samples x 20 B >> 16 MiB
    with pytest.raises(NerfB200Error) as e:
        _call(mc, mf, rays, "fp16", static=True)
    msg = str(e.value)
    assert "nb_render_rays_static" in msg and "16777216" in msg, msg
    assert "static array" in last_error()
    # save_for_backward is refused too
    hc, prec = mc.prepare("fp16")
    hf, _ = mf.prepare("fp16")
    cfg = RenderCfg(64, 128, 0, 1, prec, 1)
    rs = RenderRand(None, None, None, None, None, 0, 0.0)
    _, o = _outputs(4)
    small = _rays(4, 5)
    rc = lib().nb_render_rays_static(hc, hf, C.c_void_p(small.data_ptr()), 11, 4, C.byref(cfg), C.byref(rs), C.byref(o),
                                     C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc != 0
