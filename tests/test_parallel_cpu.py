"""CPU, world_size 2 over gloo: the host-side partitioning logic of the multi-GPU path (no data-path collective for
rendering; one flat-bucket all-reduce for training)."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from nerf_kinematics_b200 import parallel


def test_row_blocks_cover_the_image_exactly_once():
    for H in (1, 7, 400, 720, 801):
        for world in (1, 2, 3, 4, 8):
            blocks = [parallel.row_block(H, r, world) for r in range(world)]
            assert blocks[0][0] == 0 and blocks[-1][1] == H
            for a, b in zip(blocks, blocks[1:]):
                assert a[1] == b[0]
            sizes = [b[1] - b[0] for b in blocks]
            assert max(sizes) - min(sizes) <= 1
    assert sorted(sum((parallel.pose_block(11, r, 4) for r in range(4)), [])) == list(range(11))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, H):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    r, _, w = parallel.init_distributed("gloo")
    assert (r, w) == (rank, world)
    # render partition: every rank produces its own rows of a deterministic "image"; gather == the full image
    W = 5
    full = torch.arange(H * W * 3, dtype=torch.float32).reshape(H, W, 3)
    r0, r1 = parallel.row_block(H, rank, world)
    got = parallel.gather_rows(full[r0:r1].clone(), H, rank, world)
    assert torch.equal(got, full)
    # data-parallel gradient bucket: each rank seeds its loss gradient with 1/world, all-reduce(SUM) == global mean
    g = torch.Generator().manual_seed(100 + rank)
    local_grad = torch.randn(1000, generator=g)
    bucket = local_grad / world
    dist.all_reduce(bucket, op=dist.ReduceOp.SUM)
    ref = sum(torch.randn(1000, generator=torch.Generator().manual_seed(100 + k)) for k in range(world)) / world
    assert torch.allclose(bucket, ref, atol=1e-6)
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gloo_partition_and_bucket_allreduce():
    port = _free_port()
    mp.spawn(_worker, args=(2, port, 7), nprocs=2, join=True)
