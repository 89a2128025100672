"""Worker of tests/test_gpu_multi.py, launched with torchrun (one process per GPU, NCCL).

SURVEY section 4 contract: "N-GPU render must equal 1-GPU render bit-for-bit; DP train must equal single-GPU with the
concatenated batch up to all-reduce summation order".  Each rank checks its own part and exits non-zero on failure.
    render : this rank's row block (parallel.render_rows / render_path(row_block=...)) == the same rows of the whole
             frame rendered on the same GPU, torch.equal; rank 0 additionally gathers all blocks and compares.
    train  : the three-graph overlapped data-parallel step (trainer.py: forward + fine backward | coarse backward |
             Adam, all-reduces between them) on this rank's half of a 2*n-ray batch, against ONE single-GPU step on the
             concatenated batch: loss, gradients (read from Adam's first-step exp_avg) and updated parameters.
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import nerf_kinematics_b200 as nk  # noqa: E402
from nerf_kinematics_b200 import ops, parallel  # noqa: E402
from nerf_kinematics_b200.poses import pose_spherical  # noqa: E402
from nerf_kinematics_b200.trainer import TrainStep  # noqa: E402


def models(dev, seed=0):
    torch.manual_seed(seed)
    mc, mf = nk.NeRF().to(dev), nk.NeRF().to(dev)
    with torch.no_grad():       # a scene with structure (see tests/test_gpu_tc_parity.py)
        for m in (mc, mf):
            m.alpha_linear.weight.mul_(300.0)
            m.alpha_linear.bias.fill_(-3.0)
            m.rgb_linear.weight.mul_(8.0)
    return mc, mf


def check_render(rank, world, dev):
    H, W, focal = 96, 120, 140.0
    mc, mf = models(dev)
    kw = dict(network_query_fn=None, perturb=0.0, N_importance=128, network_fine=mf, N_samples=64, network_fn=mc,
              white_bkgd=True, raw_noise_std=0.0, lindisp=False, precision="fp16")
    c2w = pose_spherical(35.0, -30.0, 4.0)
    with torch.no_grad():
        full = nk.render(H, W, focal, chunk=4096, c2w=c2w[:3, :4], use_viewdirs=True, ndc=False, near=2.0, far=6.0, **kw)
        part, (r0, r1) = parallel.render_rows(H, W, focal, c2w, rank, world, chunk=4096, near=2.0, far=6.0, ndc=False,
                                              use_viewdirs=True, **kw)
        assert (r0, r1) == parallel.row_block(H, rank, world)
        assert torch.equal(part["rgb_map"], full[0][r0:r1]), "row block differs from the single-GPU frame"
        assert torch.equal(part["acc_map"], full[2][r0:r1]) and torch.equal(part["rgb0"], full[3]["rgb0"][r0:r1])
        # the public pose-path call, sharded: [M, rows, W, 3]
        rgbs, disps = nk.render_path([c2w, pose_spherical(80.0, -30.0, 4.0)], (H, W, focal), 4096,
                                     dict(kw, use_viewdirs=True, ndc=False, near=2.0, far=6.0), row_block=(rank, world))
        assert rgbs.shape == (2, r1 - r0, W, 3) and np.array_equal(rgbs[0], full[0][r0:r1].cpu().numpy())
        # optional assembly on every rank: equals the single-GPU frame bit for bit
        whole = parallel.gather_rows(part["rgb_map"], H, rank, world)
        assert torch.equal(whole, full[0]), "gathered frame differs"
    return "render: rows [%d,%d) of %d equal the 1-GPU frame; gathered frame equal" % (r0, r1, H)


def check_train(rank, world, dev):
    n = 1024
    H = W = 800
    c2w = torch.from_numpy(pose_spherical(30.0, -30.0, 4.0))
    rs = np.random.RandomState(0)
    idx = torch.from_numpy(rs.choice(H * W, n * world, replace=False)).to(dev)
    ro, rd = ops.get_rays_at(H, W, 1111.111, c2w, idx, device=dev)
    rays_all = ops.make_ray_batch(ro, rd, 2.0, 6.0, use_viewdirs=True)
    tgt_all = torch.rand(n * world, 3, generator=torch.Generator().manual_seed(3)).to(dev)
    # ---- data parallel: this rank's slice, graph-captured overlapped step
    mc, mf = models(dev)
    dp = TrainStep(mc, mf, N_samples=64, N_importance=128, perturb=0.0, raw_noise_std=0.0, white_bkgd=True,
                   precision="fp16", lrate=5e-4, world_size=world, cuda_graph=True)
    sl = slice(rank * n, (rank + 1) * n)
    loss_dp = dp.step(rays_all[sl], tgt_all[sl]).clone()
    torch.cuda.synchronize()
    assert dp.use_graph and (world == 1 or dp._graphs), "the captured step did not run"
    if world > 1:
        dist.all_reduce(loss_dp, op=dist.ReduceOp.SUM)
        loss_dp /= world
    # ---- single GPU, concatenated batch (every rank does it on its own GPU)
    mc1, mf1 = models(dev)
    one = TrainStep(mc1, mf1, N_samples=64, N_importance=128, perturb=0.0, raw_noise_std=0.0, white_bkgd=True,
                    precision="fp16", lrate=5e-4, world_size=1, cuda_graph=True)
    loss_1 = one.step(rays_all, tgt_all).clone()
    torch.cuda.synchronize()
    assert torch.allclose(loss_dp, loss_1, rtol=2e-4, atol=1e-6), (loss_dp.tolist(), loss_1.tolist())
    # gradients: after Adam's first step exp_avg = (1 - beta1) * grad
    g_dp, g_1 = dp.exp_avg / 0.1, one.exp_avg / 0.1
    scale = float(g_1.abs().max())
    assert scale > 0
    err = float((g_dp - g_1).abs().max()) / scale
    cos = float(torch.nn.functional.cosine_similarity(g_dp, g_1, dim=0))
    # fp16 gradient tiles with a per-call power-of-two scale: the two batch splits round differently
    assert err <= 2e-2 and cos >= 0.9995, (err, cos)
    # all ranks hold identical parameters after the step (the all-reduce really summed everyone)
    flat = torch.cat([mc._flat, mf._flat])
    ref = flat.clone()
    if world > 1:
        dist.broadcast(ref, src=0)
    assert torch.equal(flat, ref), "replicas diverged"
    # and they moved like the single-GPU ones where the gradient is not at the sign-flip noise level
    flat1 = torch.cat([mc1._flat, mf1._flat])
    big = g_1.abs() > 1e-3 * scale
    same = (torch.sign(g_dp[big]) == torch.sign(g_1[big])).float().mean()
    assert float(same) >= 0.999
    return "train: loss %.6f/%.6f (dp/single), grad max-rel err %.2e, cos %.6f, replicas identical" % (
        float(loss_dp.sum()), float(loss_1.sum()), err, cos)


def main():
    rank, local, world = parallel.init_distributed("nccl")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    what = sys.argv[1] if len(sys.argv) > 1 else "all"
    msgs = []
    if what in ("all", "render"):
        msgs.append(check_render(rank, world, dev))
    if what in ("all", "train"):
        msgs.append(check_train(rank, world, dev))
    if world > 1:
        dist.barrier()
    print("[rank %d/%d] OK  %s" % (rank, world, " | ".join(msgs)), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
