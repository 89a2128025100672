"""Multi-GPU partitioning of the path: one process per GPU (torchrun), no data-path
collective for rendering (rays are independent: each rank renders a contiguous row block), one
all-reduce of the flat gradient bucket for training (trainer.TrainStep)."""
import os

import torch
import torch.distributed as dist

from . import ops
from .run_nerf import batchify_rays


def init_distributed(backend=None):
    """Reads RANK / LOCAL_RANK / WORLD_SIZE / MASTER_* from the environment (torchrun)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local)
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        dist.init_process_group(backend=backend, rank=rank, world_size=world)
    return rank, local, world


def row_block(H, rank, world):
    """Rows [r0, r1) of an H-row image owned by `rank`: contiguous, sizes differ by at most one."""
    base, rem = divmod(int(H), int(world))
    r0 = rank * base + min(rank, rem)
    return r0, r0 + base + (1 if rank < rem else 0)


def pose_block(n_poses, rank, world):
    """Round-robin pose assignment (better than row blocks for small images)."""
    return list(range(rank, n_poses, world))


def render_rows(H, W, K, c2w, rank, world, chunk=1024 * 32, near=0.0, far=1.0, ndc=True, use_viewdirs=False,
                **render_kwargs):
    """This rank's row block of render(H, W, K, c2w=...) -> dict of [rows, W, ...] tensors.  No collective."""
    r0, r1 = row_block(H, rank, world)
    from .run_nerf import _frame_renderer
    fr = _frame_renderer(int(H), int(W), K, chunk, ndc, near, far, use_viewdirs, render_kwargs, rows=(r0, r1))
    if fr is not None:   # inference form: the row block is one captured CUDA graph (frame.FrameRenderer)
        out = fr.render(c2w, clone=not render_kwargs.get("_static_outputs", False))
        ret = {"rgb_map": out["rgb"], "disp_map": out["disp"], "acc_map": out["acc"]}
        ret.update({k: out[k] for k in ("rgb0", "disp0", "acc0", "z_std") if k in out})
        return ret, (r0, r1)
    render_kwargs = {k: v for k, v in render_kwargs.items() if k != "_static_outputs"}
    ro, rd = ops.get_rays(H, W, K, c2w, row0=r0, row1=r1)
    fx, _, _, _ = ops.intrinsics(H, W, K)
    packed = ops.make_ray_batch(ro, rd, near, far, use_viewdirs=use_viewdirs, ndc=ndc, H=H, W=W, focal=fx)
    ret = batchify_rays(packed, chunk, **render_kwargs)
    return {k: v.reshape([r1 - r0, W] + list(v.shape[1:])) for k, v in ret.items()}, (r0, r1)


def gather_rows(local, H, rank, world, group=None, dst=0):
    """Optional assembly of the row blocks on `dst` (20 B/ray; not needed when each rank writes its own rows)."""
    if world == 1:
        return local
    blocks = [row_block(H, r, world) for r in range(world)]
    maxrows = max(b[1] - b[0] for b in blocks)
    pad = torch.zeros((maxrows,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[: local.shape[0]] = local
    out = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(out, pad, group=group)
    return torch.cat([o[: b[1] - b[0]] for o, b in zip(out, blocks)], 0)
