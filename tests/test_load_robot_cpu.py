"""CPU: the load_robot twin (nerf_kinematics_b200/load_robot.py) against the reference loader's own
outputs (tests/golden/robot_*.npz, produced by oracle/make_robot_golden.py from the unmodified
nerf-pytorch/nerf/load_robot.py): strided downsample, doubly divided focal, invented bounds, recentring, spiral path."""
import os
import tempfile

import numpy as np
import pytest

from conftest import GOLD
from oracle.make_robot_golden import make_synthetic_capture

from nerf_kinematics_b200.load_robot import load_robot_data, parse_poses_file


def _check(got, g, tag):
    imgs, poses, bds, rp, it = got
    assert imgs.dtype == np.float32 and poses.dtype == np.float32 and bds.dtype == np.float32 and rp.dtype == np.float32
    assert np.array_equal((imgs * 255.0 + 0.5).astype(np.uint8), g[tag + "imgs"])
    for name, a in (("poses", poses), ("bds", bds), ("render_poses", rp)):
        assert a.shape == g[tag + name].shape, name
        assert np.array_equal(a, g[tag + name]), (name, float(np.abs(a - g[tag + name]).max()))
    assert int(it) == int(g[tag + "i_test"])


def test_synthetic_capture_matches_the_reference_loader():
    g = np.load(os.path.join(GOLD, "robot_synth.npz"))
    with tempfile.TemporaryDirectory() as tmp:
        make_synthetic_capture(tmp)
        _check(load_robot_data(tmp, factor=2), g, "f2_")
        _check(load_robot_data(tmp, factor=1, recenter=False), g, "f1_norecenter_")
        _check(load_robot_data(tmp, factor=4, path_zflat=True, bd_factor=None), g, "f4_zflat_")
        imgs, poses, bds, rp, _ = load_robot_data(tmp, factor=2)
        # the quirks, spelled out: strided subsample, focal from the downsampled width divided by the factor again
        assert imgs.shape == (7, 24, 32, 3) and rp.shape == (120, 3, 5)
        assert poses[0, 2, 4] == np.float32(32 / (2.0 * np.tan(87.0 * np.pi / 360.0)) / 2)
        assert np.allclose(bds[:, 1] / bds[:, 0], 30.0) and np.isclose(bds.min(), 1.0 / 0.75)


@pytest.mark.skipif(not os.path.isdir("/root/reference/datasets/nerf-raw/wheel_real_robot"), reason="real capture not on this box")
def test_real_wheel_capture_matches_the_reference_loader():
    g = np.load(os.path.join(GOLD, "robot_wheel.npz"))
    got = load_robot_data("/root/reference/datasets/nerf-raw/wheel_real_robot", factor=8)
    _check(got, {("" + k): g[k] for k in g.files}, "")
    assert got[0].shape == (12, 90, 160, 3)


def test_error_behaviour_is_the_reference_s():
    with pytest.raises(SystemExit) as e:
        load_robot_data("/nonexistent/dir")
    assert e.value.code == 1
    with tempfile.TemporaryDirectory() as tmp:
        make_synthetic_capture(tmp, n=3)
        with open(os.path.join(tmp, "poses.txt"), "a") as f:      # one pose too many
            f.write("[1 0 0 0; 0 1 0 0; 0 0 1 0; 0 0 0 1]\n")
        with pytest.raises(SystemExit):
            load_robot_data(tmp)
        assert len(parse_poses_file(os.path.join(tmp, "poses.txt"))) == 4
    assert parse_poses_file("/nonexistent/poses.txt") == []
