"""Data-free camera paths the reference builds on the host (numpy): the Blender turntable and the LLFF spiral."""
import numpy as np


def pose_spherical(theta, phi, radius):
    """nerf/load_blender.py:15-37 == nerf-pytorch/nerf/load_blender.py:9-37 -> 4x4 float32 camera-to-world."""
    t = np.eye(4, dtype=np.float32)
    t[2, 3] = radius
    p = phi / 180.0 * np.pi
    rp = np.array([[1, 0, 0, 0], [0, np.cos(p), -np.sin(p), 0], [0, np.sin(p), np.cos(p), 0], [0, 0, 0, 1]], np.float32)
    th = theta / 180.0 * np.pi
    rt = np.array([[np.cos(th), 0, -np.sin(th), 0], [0, 1, 0, 0], [np.sin(th), 0, np.cos(th), 0], [0, 0, 0, 1]],
                  np.float32)
    c2w = rt @ (rp @ t)
    flip = np.array([[-1, 0, 0, 0], [0, 0, 1, 0], [0, 1, 0, 0], [0, 0, 0, 1]], np.float32)
    return (flip @ c2w).astype(np.float32)


def blender_render_poses(n=40):
    """The 40-view turntable of load_blender.py:79-85: theta = linspace(-180, 180, n+1)[:-1], phi = -30, r = 4."""
    return np.stack([pose_spherical(a, -30.0, 4.0) for a in np.linspace(-180, 180, n + 1)[:-1]], 0)


def _normalize(x):
    return x / np.linalg.norm(x)


def _viewmatrix(z, up, pos):
    vec2 = _normalize(z)
    vec0 = _normalize(np.cross(up, vec2))
    vec1 = _normalize(np.cross(vec2, vec0))
    return np.stack([vec0, vec1, vec2, pos], 1)


def render_path_spiral(c2w, up, rads, focal, zdelta, zrate, rots, N):
    """nerf-pytorch/nerf/load_llff.py:169-182 -> list of 3x4 poses (the hwf column is dropped)."""
    out = []
    rads = np.array(list(rads) + [1.0])
    for theta in np.linspace(0.0, 2.0 * np.pi * rots, N + 1)[:-1]:
        c = np.dot(c2w[:3, :4], np.array([np.cos(theta), -np.sin(theta), -np.sin(theta * zrate), 1.0]) * rads)
        z = _normalize(c - np.dot(c2w[:3, :4], np.array([0, 0, -focal, 1.0])))
        out.append(_viewmatrix(z, up, c).astype(np.float32))
    return out


# ------------------------------------------------------------------------------------------ robot poses (host side)
# SURVEY.md 8(f) f2.  Poses still come from the robot's forward kinematics: poses.txt -> parser_instant_ngp.py (unchanged,
# on the host) -> transforms.json.  These readers turn that output into what render() / render_path() take.
def parse_poses_txt(path):
    """The `[ a, b, c, d ; ... ; ... ; ... ];` blocks of the robot's poses.txt -> float64 [N, 4, 4]
    (same grammar as parser_instant_ngp.py:parse_poses_file / nerf-pytorch/nerf/load_robot.py:8-50)."""
    import re
    text = open(path, "r").read()
    mats = []
    for block in re.findall(r"\[\s*(.*?)\s*\]", text, re.DOTALL):
        rows = []
        for row in block.split(";"):
            nums = re.findall(r"-?\d+\.?\d*(?:[eE][-+]?\d+)?", row.replace(",", " "))
            if len(nums) == 4:
                rows.append([float(x) for x in nums])
        if len(rows) == 4:
            mats.append(rows)
    return np.asarray(mats, dtype=np.float64).reshape(-1, 4, 4)


def load_transforms(path_or_dict):
    """Read a transforms.json written by parser_instant_ngp.py (create_base_json :128-145, frames :297-355).

    Returns dict(H, W, K [3,3] float32 from fl_x / fl_y / cx / cy -- the wheel frames have fl_x != fl_y, so the scalar
    `focal` of the reference API is not enough --, poses [N,4,4] float32 camera-to-world from `transform_matrix` (or
    `transform_matrix_start` in the _test file), file_paths, sharpness, aabb_scale, camera_angle_x/y).
    Note: with --recenter the parser negates the rotation's Z column, so det(R) = -1 (instant-ngp warns about it);
    the matrices are passed through unchanged."""
    import json
    d = path_or_dict if isinstance(path_or_dict, dict) else json.load(open(path_or_dict, "r"))
    W, H = int(d["w"]), int(d["h"])
    K = np.array([[d["fl_x"], 0.0, d["cx"]], [0.0, d["fl_y"], d["cy"]], [0.0, 0.0, 1.0]], dtype=np.float32)
    poses, paths, sharp = [], [], []
    for fr in d.get("frames", []):
        m = fr.get("transform_matrix", fr.get("transform_matrix_start"))
        if m is None:
            raise ValueError("transforms.json frame without transform_matrix: %r" % (sorted(fr.keys()),))
        m = np.asarray(m, dtype=np.float64)
        if m.shape != (4, 4):
            raise ValueError("transform_matrix of shape %s" % (m.shape,))
        poses.append(m.astype(np.float32))
        paths.append(fr.get("file_path"))
        sharp.append(fr.get("sharpness"))
    poses = np.stack(poses, 0) if poses else np.zeros((0, 4, 4), np.float32)
    return {"H": H, "W": W, "K": K, "poses": poses, "file_paths": paths, "sharpness": sharp,
            "aabb_scale": d.get("aabb_scale"), "camera_angle_x": d.get("camera_angle_x"),
            "camera_angle_y": d.get("camera_angle_y")}
