import sys, os, time, cProfile, pstats
sys.path.insert(0, "/root/repo")
import numpy as np, torch
import nerf_kinematics_b200 as nk
import bench
from nerf_kinematics_b200.poses import blender_render_poses
dev = torch.device("cuda:0")
mc, mf = bench.lego_models(dev)
kw = bench.render_kwargs_for(mc, mf, "bf16", ndc=False, near=2.0, far=6.0)
poses = [np.asarray(p, dtype=np.float32) for p in blender_render_poses(40)]
for world in (8, 1):
    nk.render_path(poses[:2], (400, 400, 555.555), 32768, kw, row_block=(0, world))
    torch.cuda.synchronize()
    for rep in range(2):
        t0 = time.perf_counter()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        rgbs, disps = nk.render_path(poses, (400, 400, 555.555), 32768, kw, row_block=(0, world))
        e1.record(); torch.cuda.synchronize()
        print("world", world, "e2e ms/frame", e0.elapsed_time(e1) / 40, "wall", (time.perf_counter() - t0) * 1e3 / 40)
    pr = cProfile.Profile(); pr.enable()
    nk.render_path(poses, (400, 400, 555.555), 32768, kw, row_block=(0, world))
    pr.disable()
    if world == 8:
        pstats.Stats(pr).sort_stats("cumulative").print_stats(22)
