"""Golden vectors for the load_robot twin, produced by the reference's own loader.

TEST INFRASTRUCTURE ONLY.  Run in the build container (needs /root/reference or oracle/_ref):

    python oracle/make_robot_golden.py

(1) A small synthetic capture (seeded images + a poses.txt in the robot's grammar) is written to a temp directory,
    the UNMODIFIED nerf-pytorch/nerf/load_robot.py:load_robot_data reads it, and every output is stored in
    tests/golden/robot_synth.npz.  tests/test_load_robot_cpu.py rebuilds the same directory with
    `make_synthetic_capture` and compares the twin's outputs with the stored ones.
(2) When the real wheel capture is present (datasets/nerf-raw/wheel_real_robot, 12 photos at 1280x720), the reference
    loader's poses / bounds / render path / test index at factor 8 go to tests/golden/robot_wheel.npz together with
    the strided 160x90 images as uint8.
imageio is absent from this image: the reference's `imageio.imread` is served by PIL (same decoded bytes).
"""
import os
import sys
import tempfile
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
GOLD = os.path.join(ROOT, "tests", "golden")


def make_synthetic_capture(dst, n=7, H=48, W=64, seed=0):
    """n PNG photos `TestNERF <i>.png` (with a gap in the numbering, like the real capture) + poses.txt."""
    from PIL import Image
    rs = np.random.RandomState(seed)
    os.makedirs(os.path.join(dst, "images_robot"), exist_ok=True)
    numbers = [0, 1, 2, 3, 5, 10, 11][:n]
    lines = []
    for i, num in enumerate(numbers):
        img = rs.randint(0, 256, size=(H, W, 3)).astype(np.uint8)
        Image.fromarray(img).save(os.path.join(dst, "images_robot", "TestNERF %d.png" % num))
        a = 2.0 * np.pi * i / n
        z = np.array([np.cos(a), 0.3, np.sin(a)])
        z /= np.linalg.norm(z)
        x = np.cross([0.0, 1.0, 0.0], z)
        x /= np.linalg.norm(x)
        y = np.cross(z, x)
        t = -1.7 * z + rs.randn(3) * 0.05
        m = np.eye(4)
        m[:3, 0], m[:3, 1], m[:3, 2], m[:3, 3] = x, y, z, t
        rows = [", ".join("%.6f" % v for v in m[r]) for r in range(4)]
        lines.append("T%d = [ %s ];" % (i, " ;\n  ".join(rows)))
    with open(os.path.join(dst, "poses.txt"), "w") as f:
        f.write("\n".join(lines) + "\n")
    return dst


def _reference_loader():
    from oracle import reference_shim
    from PIL import Image
    im = types.ModuleType("imageio")
    im.imread = lambda p: np.asarray(Image.open(p))
    sys.modules["imageio"] = im
    reference_shim.load()
    import nerf.load_robot as ref_lr
    return ref_lr


def main():
    ref = _reference_loader()
    with tempfile.TemporaryDirectory() as tmp:
        make_synthetic_capture(tmp)
        out = {}
        for tag, kw in (("f2", dict(factor=2)), ("f1_norecenter", dict(factor=1, recenter=False)),
                        ("f4_zflat", dict(factor=4, path_zflat=True, bd_factor=None))):
            imgs, poses, bds, rp, it = ref.load_robot_data(tmp, **kw)
            out.update({tag + "_imgs": (imgs * 255.0 + 0.5).astype(np.uint8), tag + "_poses": poses, tag + "_bds": bds,
                        tag + "_render_poses": rp, tag + "_i_test": np.int64(it)})
        np.savez_compressed(os.path.join(GOLD, "robot_synth.npz"), **out)
        print("wrote robot_synth.npz", os.path.getsize(os.path.join(GOLD, "robot_synth.npz")))
    real = os.path.join(os.environ.get("NERF_REFERENCE_ROOT", "/root/reference"), "datasets", "nerf-raw", "wheel_real_robot")
    if os.path.isdir(real):
        imgs, poses, bds, rp, it = ref.load_robot_data(real, factor=8)
        np.savez_compressed(os.path.join(GOLD, "robot_wheel.npz"), imgs=(imgs * 255.0 + 0.5).astype(np.uint8), poses=poses,
                            bds=bds, render_poses=rp, i_test=np.int64(it))
        print("wrote robot_wheel.npz", os.path.getsize(os.path.join(GOLD, "robot_wheel.npz")), imgs.shape, poses[0, :, 4])


if __name__ == "__main__":
    main()
