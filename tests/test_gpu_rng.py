"""the in-kernel Philox generator behind perturb / raw_noise_std.

The reference draws these numbers from the framework generator (run_nerf.py:192-199, run_nerf_helpers.py:192-197,
run_nerf.py:130-132), so the match is statistical; what IS exact is that the numbers drawn inside the kernels are
the ones nb_rng_fill reports: feeding those arrays through the explicit-array path must give the same bits."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _fill(state, inp, normal, scale, n, row):
    from nerf_kinematics_b200._lib import check, lib
    out = torch.empty((n, row), dtype=torch.float32, device=state.device)
    with torch.cuda.device(state.device):
        check(lib().nb_rng_fill(C.c_void_p(state.data_ptr()), inp, normal, scale, n, row, C.c_void_p(out.data_ptr()),
                                C.c_void_p(torch.cuda.current_stream().cuda_stream)), "nb_rng_fill")
    return out


def test_philox_streams_statistics_and_determinism():
    dev = torch.device("cuda:0")
    st = torch.tensor([7, 0], dtype=torch.int64, device=dev)
    u = _fill(st, 0, 0, 0.0, 4096, 192)
    assert float(u.min()) >= 0.0 and float(u.max()) < 1.0
    assert abs(float(u.mean()) - 0.5) < 2e-3 and abs(float(u.var()) - 1.0 / 12.0) < 2e-3
    # same (seed, step, input) -> same numbers; another input / step / seed -> different ones
    assert torch.equal(u, _fill(st, 0, 0, 0.0, 4096, 192))
    u1 = _fill(st, 1, 0, 0.0, 4096, 192)
    assert float((u == u1).float().mean()) < 1e-3
    from nerf_kinematics_b200._lib import check, lib
    check(lib().nb_rng_advance(C.c_void_p(st.data_ptr()), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    torch.cuda.synchronize()
    assert st.tolist() == [7, 1]
    assert float((u == _fill(st, 0, 0, 0.0, 4096, 192)).float().mean()) < 1e-3
    st2 = torch.tensor([8, 0], dtype=torch.int64, device=dev)
    assert float((u == _fill(st2, 0, 0, 0.0, 4096, 192)).float().mean()) < 1e-3
    # no correlation between neighbouring samples / rays (lag-1 autocorrelation ~ N(0, 1/sqrt(N)))
    c = u - 0.5
    assert abs(float((c[:, 1:] * c[:, :-1]).mean()) * 12.0) < 5e-3
    assert abs(float((c[1:] * c[:-1]).mean()) * 12.0) < 5e-3
    # uniform histogram (chi-square, 64 bins, ~786 K draws): mean 63, sd ~11
    h = torch.histc(u, bins=64, min=0.0, max=1.0).double()
    e = u.numel() / 64.0
    assert float(((h - e) ** 2 / e).sum()) < 140.0
    # normal deviates: scale, moments, tails
    g = _fill(st, 2, 1, 2.0, 4096, 192)
    assert abs(float(g.mean())) < 1e-2 and abs(float(g.std()) - 2.0) < 1e-2
    z = g / 2.0
    assert abs(float((z ** 3).mean())) < 2e-2 and abs(float((z ** 4).mean()) - 3.0) < 5e-2
    assert torch.isfinite(g).all()


def _models(dev):
    import nerf_kinematics_b200 as nk
    torch.manual_seed(3)
    mc = nk.NeRF(D=8, W=256, input_ch=63, input_ch_views=27, skips=[4], use_viewdirs=True).to(dev)
    mf = nk.NeRF(D=8, W=256, input_ch=63, input_ch_views=27, skips=[4], use_viewdirs=True).to(dev)
    return mc, mf


@pytest.mark.parametrize("noise", [0.0, 1.0])
def test_in_kernel_numbers_equal_the_explicit_array_path(noise):
    """TrainStep with the device generator == TrainStep fed the arrays nb_rng_fill reports (loss and gradients)."""
    from nerf_kinematics_b200.trainer import TrainStep
    dev = torch.device("cuda:0")
    n, S, I = 200, 64, 128
    g = torch.Generator().manual_seed(5)
    rays = torch.cat([torch.randn(n, 3, generator=g) * 0.2 + torch.tensor([0.0, 0.0, 4.0]),
                      torch.nn.functional.normalize(torch.randn(n, 3, generator=g) * 0.2 + torch.tensor([0.0, 0.0, -1.0]), dim=-1),
                      torch.full((n, 1), 2.0), torch.full((n, 1), 6.0)], -1)
    rays = torch.cat([rays, rays[:, 3:6]], -1).to(dev).contiguous()
    tgt = torch.rand(n, 3, generator=g).to(dev)
    res = []
    for mode in ("kernel", "arrays"):
        mc, mf = _models(dev)
        st = TrainStep(mc, mf, N_samples=S, N_importance=I, perturb=1.0, raw_noise_std=noise, white_bkgd=True,
                       precision="fp32", seed=11)
        rnd = None
        if mode == "arrays":
            rnd = {"t_rand": _fill(st.rng_state, 0, 0, 0.0, n, S), "u_fine": _fill(st.rng_state, 1, 0, 0.0, n, I)}
            if noise > 0:
                rnd["noise_c"] = _fill(st.rng_state, 2, 1, noise, n, S)
                rnd["noise_f"] = _fill(st.rng_state, 3, 1, noise, n, S + I)
        loss = st.forward_backward(rays, tgt, rnd=rnd).clone()
        torch.cuda.synchronize()
        res.append((loss.cpu(), st.bucket.clone().cpu(), st.rng_state.tolist(), st._out["rgb"].clone().cpu(),
                    st._out["rgb0"].clone().cpu()))
    assert res[0][2] == [11, 1] and res[1][2] == [11, 0]   # only the in-kernel path consumes a step
    # the rendered colours, the loss (one-block sum) and the gradients (split-K partials summed in a fixed order,
    # mlp_fp32.cu ordered_sum_kernel) are the same bits: nothing on the path accumulates with atomics
    assert torch.equal(res[0][3], res[1][3]) and torch.equal(res[0][4], res[1][4])
    assert torch.equal(res[0][0], res[1][0])
    assert torch.equal(res[0][1], res[1][1]), float((res[0][1] - res[1][1]).abs().max())
    assert float(res[0][1].abs().max()) > 0.0


def test_graph_replays_draw_fresh_numbers_and_perturbed_render_matches_oracle_statistics():
    """A captured step advances the device-side step counter on every replay; and the jittered coarse depths keep
    the reference's stratification: z[s] stays inside its bin [mid(s-1,s), mid(s,s+1)] (run_nerf.py:192-199)."""
    from nerf_kinematics_b200.trainer import TrainStep
    dev = torch.device("cuda:0")
    mc, mf = _models(dev)
    n = 256
    g = torch.Generator().manual_seed(6)
    rays = torch.cat([torch.randn(n, 3, generator=g) * 0.2 + torch.tensor([0.0, 0.0, 4.0]),
                      torch.nn.functional.normalize(torch.randn(n, 3, generator=g) * 0.2 + torch.tensor([0.0, 0.0, -1.0]), dim=-1),
                      torch.full((n, 1), 2.0), torch.full((n, 1), 6.0)], -1)
    rays = torch.cat([rays, rays[:, 3:6]], -1).to(dev).contiguous()
    tgt = torch.rand(n, 3, generator=g).to(dev)
    st = TrainStep(mc, mf, N_samples=64, N_importance=128, perturb=1.0, white_bkgd=True, precision="bf16",
                   cuda_graph=True, seed=3)
    losses = [float(st.step(rays, tgt)[0]) for _ in range(4)]
    torch.cuda.synchronize()
    assert st.rng_state.tolist()[0] == 3 and st.rng_state.tolist()[1] >= 4
    assert all(np.isfinite(losses))
    # stratification of the jitter the kernels draw
    u = _fill(st.rng_state, 0, 0, 0.0, n, 64)
    t = torch.linspace(0.0, 1.0, 64, device=dev)
    z = 2.0 * (1 - t) + 6.0 * t
    mids = 0.5 * (z[1:] + z[:-1])
    lower, upper = torch.cat([z[:1], mids]), torch.cat([mids, z[-1:]])
    zz = lower + (upper - lower) * u
    assert bool((zz >= lower - 1e-6).all()) and bool((zz <= upper + 1e-6).all())
    assert bool((zz[:, 1:] >= zz[:, :-1]).all())


@pytest.mark.parametrize("precision", ["fp32", "fp16"])
def test_train_step_is_reproducible_bit_for_bit(precision):
    """Same seed, same batch, two fresh TrainSteps: loss, gradients and the updated parameters are the same bits (no
    reduction on the path -- loss, bias / weight gradients, split-K partials -- depends on a race)."""
    from nerf_kinematics_b200.trainer import TrainStep
    dev = torch.device("cuda:0")
    n = 1024
    g = torch.Generator().manual_seed(8)
    rays = torch.cat([torch.randn(n, 3, generator=g) * 0.2 + torch.tensor([0.0, 0.0, 4.0]),
                      torch.nn.functional.normalize(torch.randn(n, 3, generator=g) * 0.2 + torch.tensor([0.0, 0.0, -1.0]), dim=-1),
                      torch.full((n, 1), 2.0), torch.full((n, 1), 6.0)], -1)
    rays = torch.cat([rays, rays[:, 3:6]], -1).to(dev).contiguous()
    tgt = torch.rand(n, 3, generator=g).to(dev)
    res = []
    for _ in range(2):
        mc, mf = _models(dev)
        st = TrainStep(mc, mf, N_samples=64, N_importance=128, perturb=1.0, raw_noise_std=1.0, white_bkgd=True,
                       precision=precision, seed=21, cuda_graph=False)
        losses = [st.step(rays, tgt).clone() for _ in range(2)]
        torch.cuda.synchronize()
        res.append((torch.stack(losses).cpu(), st.bucket.clone().cpu(), torch.cat([mc._flat, mf._flat]).cpu()))
    assert torch.equal(res[0][0], res[1][0]), (res[0][0], res[1][0])
    assert torch.equal(res[0][1], res[1][1]), float((res[0][1] - res[1][1]).abs().max())
    assert torch.equal(res[0][2], res[1][2])
