"""CPU: the host-side readers of the robot pipeline against the output of the unmodified reference
parser on the repo's real capture (fixtures made by oracle/make_transforms_golden.py)."""
import os

import numpy as np

from nerf_kinematics_b200 import poses

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_load_transforms_wheel_capture():
    t = poses.load_transforms(os.path.join(GOLD, "wheel_transforms.json"))
    assert (t["H"], t["W"]) == (720, 1280) and t["poses"].shape == (11, 4, 4) and t["poses"].dtype == np.float32
    # 87 x 58 degree field of view: fl_x != fl_y (parser_instant_ngp.py:266-272), principal point at the centre
    assert abs(t["K"][0, 0] - 674.4193) < 1e-3 and abs(t["K"][1, 1] - 649.4572) < 1e-3
    assert t["K"][0, 2] == 640.0 and t["K"][1, 2] == 360.0
    # image index 7 is missing from the capture: frames are paired sequentially with the next existing image
    names = [os.path.basename(p) for p in t["file_paths"]]
    assert names[:7] == ["TestNERF %d.jpg" % i for i in (1, 2, 3, 4, 5, 6, 8)]
    # --recenter: camera centres normalised to the unit ball, Z column negated -> det(R) = -1
    c = t["poses"][:, :3, 3]
    assert abs(np.linalg.norm(c, axis=1).max() - 1.0) < 1e-5 and np.abs(c.mean(0)).max() < 1e-6
    assert np.allclose(np.linalg.det(t["poses"][:, :3, :3].astype(np.float64)), -1.0, atol=1e-4)
    assert t["aabb_scale"] == 2.0


def test_load_transforms_val_test_and_video_files():
    assert poses.load_transforms(os.path.join(GOLD, "wheel_transforms_val.json"))["poses"].shape == (1, 4, 4)
    tt = poses.load_transforms(os.path.join(GOLD, "wheel_transforms_test.json"))  # `transform_matrix_start` frames
    assert tt["poses"].shape == (8, 4, 4)
    tv = poses.load_transforms(os.path.join(GOLD, "wheel_transforms_test_video.json"))
    assert tv["poses"].shape == (60, 4, 4) and tv["file_paths"][0] is None
    # the generated orbits look at the scene centre: -z axis of the camera points to it
    fwd = -tv["poses"][:, :3, 2]
    to_centre = -tv["poses"][:, :3, 3] + np.array([0, 0, tv["poses"][:, 2, 3].mean()])
    cos = (fwd * to_centre).sum(1) / np.linalg.norm(to_centre, axis=1)
    assert cos.min() > 0.95


def test_parse_poses_txt_grammar(tmp_path):
    txt = """[ 0.730263, -0.682985, -0.015725, 1072.532608 ;
      -0.683050, -0.729519, -0.035290, 132.989927 ;
       0.012630,  0.036512, -0.999253, -53.612386 ;
       0.000000,  0.000000,  0.000000,   1.000000 ];


    [ 1, 0, 0, 1.5e1 ; 0, 1, 0, -2 ; 0, 0, 1, 3 ; 0, 0, 0, 1 ];
    [ 1, 2, 3 ; 4, 5, 6 ];
    """
    p = tmp_path / "poses.txt"
    p.write_text(txt)
    m = poses.parse_poses_txt(str(p))
    assert m.shape == (2, 4, 4)  # the malformed block is skipped like in the reference
    assert m[0, 0, 3] == 1072.532608 and m[0, 2, 2] == -0.999253 and m[1, 0, 3] == 15.0 and m[1, 1, 3] == -2.0
