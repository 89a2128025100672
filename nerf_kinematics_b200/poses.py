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
