"""Host-side loader for the robot scenes: twin of nerf-pytorch/nerf/load_robot.py:162-362.

The robot's forward kinematics writes one 4x4 camera-to-world matrix per photo into `poses.txt`; this loader pairs them
with `images_robot/*` and returns what the reference's train / eval scripts consume, in the LLFF layout:
`(images [N,H,W,3] float32 in [0,1], poses [N,3,5] = [R | t | (H,W,focal)], bds [N,2], render_poses [120,3,5], i_test)`.
Same signature, same return order, same error behaviour (message + SystemExit(1)).  It stays on the host (numpy): the
hot path starts at get_rays.  Quirks of the reference that are reproduced because checkpoints were trained with them:
  * images are downsampled by striding `img[::factor, ::factor]`, not by area averaging (:264);
  * the focal length is computed from the ALREADY downsampled width (87 degree horizontal field of view) and then
    divided by `factor` once more (:272-276), i.e. it is `factor` times too small;
  * the scene bounds are invented from the camera positions: near = 0.1 R, far = 3 R with R the largest distance of a
    camera from the mean camera position (:286-299), then rescaled by 1 / (near * bd_factor) like LLFF;
  * `spherify` is accepted and ignored.
"""
import os
import re
import sys

import numpy as np

from .poses import _normalize, _viewmatrix

_FOV_H_DEG = 87.0   # RealSense colour camera, horizontal (load_robot.py:272)


def _die(msg):
    print(msg)
    sys.exit(1)


def _imread(path):
    try:
        import imageio
        if hasattr(imageio, "imread"):
            return np.asarray(imageio.imread(path))
    except ImportError:
        pass
    from PIL import Image
    return np.asarray(Image.open(path))


def parse_poses_file(path):
    """`[ a, b, c, d ; ... ]` blocks -> list of float64 4x4 arrays (load_robot.py:8-50).  The reference's number pattern
    has no exponent part; neither has this one.  Returns [] (after printing) when the file cannot be read."""
    try:
        text = open(path, "r").read()
    except Exception as ex:
        print("Error al parsear el archivo de poses %s: %s" % (path, ex))
        return []
    out = []
    for block in re.findall(r"\[\s*(.*?)\s*\]", text, re.DOTALL):
        rows = []
        for row in block.split(";"):
            row = row.strip().replace(",", "")
            if not row:
                continue
            nums = re.findall(r"-?\d+\.?\d*", row)
            if len(nums) == 4:
                rows.append([float(v) for v in nums])
        if len(rows) == 4:
            out.append(np.array(rows))
    return out


def _poses_avg(poses):
    """Average camera frame of [N,3,5] (or [N,4,4]) poses, with the hwf column of the first one appended."""
    centre = poses[:, :3, 3].mean(0)
    z = _normalize(poses[:, :3, 2].sum(0))
    up = poses[:, :3, 1].sum(0)
    return np.concatenate([_viewmatrix(z, up, centre), poses[0, :3, -1:]], 1)


def _recenter(poses44):
    """Express every pose in the average camera's frame (load_robot.py:135-146)."""
    last = np.array([[0.0, 0.0, 0.0, 1.0]])
    avg = np.concatenate([_poses_avg(poses44)[:3, :4], last], -2)
    full = np.concatenate([poses44[:, :3, :4], np.tile(last[None], [poses44.shape[0], 1, 1])], -2)
    out = poses44 + 0
    out[:, :3, :4] = (np.linalg.inv(avg) @ full)[:, :3, :4]
    return out


def _spiral(c2w, up, rads, focal, zrate, rots, n):
    """load_robot.py:148-160: 3x5 poses on a spiral around c2w, looking at the point `focal` in front of it."""
    rads = np.array(list(rads) + [1.0])
    hwf = c2w[:, 4:5]
    path = []
    for theta in np.linspace(0.0, 2.0 * np.pi * rots, n + 1)[:-1]:
        pos = np.dot(c2w[:3, :4], np.array([np.cos(theta), -np.sin(theta), -np.sin(theta * zrate), 1.0]) * rads)
        look = _normalize(pos - np.dot(c2w[:3, :4], np.array([0, 0, -focal, 1.0])))
        path.append(np.concatenate([_viewmatrix(look, up, pos), hwf], 1))
    return path


def _trailing_number(name):
    m = re.search(r"(\d+)\.", name)
    return int(m.group(1)) if m else float("inf")


def load_robot_data(basedir, factor=8, recenter=True, bd_factor=0.75, spherify=False, path_zflat=False):
    if not os.path.isdir(basedir):
        _die("Error: El directorio base '%s' no existe." % basedir)
    img_dir = os.path.join(basedir, "images_robot")
    poses_path = os.path.join(basedir, "poses.txt")
    if not os.path.isfile(poses_path):
        _die("Error: Debe proporcionar un archivo poses.txt válido. Archivo proporcionado: %s" % poses_path)
    names = [f for f in os.listdir(img_dir)
             if os.path.isfile(os.path.join(img_dir, f)) and f.lower().endswith((".jpg", ".jpeg", ".png"))]
    if not names:
        _die("Error: No se encontraron imágenes en '%s'." % img_dir)
    names.sort(key=_trailing_number)            # stable: by the number in front of the extension
    mats = parse_poses_file(poses_path)
    if not mats:
        _die("Error: No se pudieron parsear matrices de poses del archivo.")
    if len(mats) != len(names):
        _die("Error: El número de poses (%d) no coincide con el número de imágenes (%d)" % (len(mats), len(names)))

    frames = []
    for f in names:
        im = _imread(os.path.join(img_dir, f))
        if im.ndim == 3 and im.shape[2] == 4:
            im = im[:, :, :3]
        elif im.ndim == 2:
            im = np.stack([im] * 3, axis=-1)
        frames.append(im)
    imgs = np.array(frames).astype(np.float32) / 255.0
    c2w_all = np.array(mats).astype(np.float32)
    H, W = imgs[0].shape[:2]
    if factor > 1:
        imgs = np.array([im[::factor, ::factor] for im in imgs])
        H, W = H // factor, W // factor
    focal = W / (2.0 * np.tan((_FOV_H_DEG * np.pi / 180.0) / 2.0))
    if factor > 1:
        focal = focal / factor                  # (sic) second division, see the module docstring

    poses = np.zeros((len(c2w_all), 3, 5), dtype=np.float32)
    poses[:, :, :4] = c2w_all[:, :3, :].astype(np.float32)
    poses[:, :, 4] = np.array([H, W, focal], dtype=np.float32)

    cam = c2w_all[:, :3, 3].astype(np.float32)
    radius = np.max(np.linalg.norm(cam - np.mean(cam, axis=0).astype(np.float32), axis=1).astype(np.float32)).astype(np.float32)
    near, far = (radius * 0.1).astype(np.float32), (radius * 3.0).astype(np.float32)
    bds = np.array([[near, far] for _ in range(len(c2w_all))], dtype=np.float32)
    if bd_factor is not None:
        sc = (1.0 / (bds.min() * bd_factor)).astype(np.float32)
        poses[:, :3, 3] *= sc
        bds *= sc
    if recenter:
        tmp = np.zeros((len(poses), 4, 4))
        tmp[:, :3, :] = poses[:, :, :4]
        tmp[:, 3, 3] = 1.0
        poses[:, :, :4] = _recenter(tmp)[:, :3, :]

    c2w = _poses_avg(poses)
    up = _normalize(poses[:, :3, 1].sum(0))
    close_depth, inf_depth = bds.min() * 0.9, bds.max() * 5.0
    dt = 0.75
    focal_spiral = 1.0 / ((1.0 - dt) / close_depth + dt / inf_depth)
    rads = np.percentile(np.abs(poses[:, :3, 3]), 90, 0)
    n_views, n_rots = 120, 2
    if path_zflat:
        c2w[:3, 3] = c2w[:3, 3] + (-close_depth * 0.1) * c2w[:3, 2]
        rads[2] = 0.0
        n_rots, n_views = 1, n_views // 2
    render_poses = np.array(_spiral(c2w, up, rads, focal_spiral, 0.5, n_rots, n_views), dtype=np.float32)

    centre = _poses_avg(poses)
    i_test = np.argmin(np.sum(np.square(centre[:3, 3] - poses[:, :3, 3]), -1))
    return imgs, poses, bds, render_poses, i_test
