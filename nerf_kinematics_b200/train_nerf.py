"""Thin train / eval drivers: the loop of nerf-pytorch/train_nerf.py:194-415 and the render loop of
eval_nerf.py:127-189 on top of trainer.TrainStep, with the reference's checkpoint layout so that a shipped `.ckpt`
(models AND optimizer state) resumes here and a checkpoint written here loads in the reference.

Dataset loading, TensorBoard and video writing stay with the caller (out of scope): `train` takes
images [N,H,W,3] / poses [N,3|4,4] tensors and the split, exactly the variables the reference's loop body reads.
Device-side differences from the reference loop, none of which change the arithmetic contract:
  * only the N_rand selected rays are generated (TrainStep.step_pixels) instead of the whole image's bundle;
  * the optimizer is the fused Adam over flat buffers; `optimizer.state_dict()` keeps torch.optim.Adam's format.
Reproduced quirks: models are built from the encoding fields of cfg.models.* only -- `num_layers` / `hidden_size` of
the YAML are ignored by the reference too (train_nerf.py:144-162), so every shipped checkpoint is the 4x128 default;
pixel k of np.random.choice(H * W) is (row k % H, col k // H) because `coords` comes from meshgrid_xy (:243-253);
iteration i runs with the learning rate set after iteration i - 1 (:292-297).
"""
import os

import numpy as np
import torch

from . import models as _models
from .nerf_helpers import mse2psnr
from .train_utils import run_one_iter_of_nerf
from .trainer import TrainStep


class CfgNode(dict):
    """Attribute access over a nested dict (stands in for nerf-pytorch/nerf/cfgnode.py: `CfgNode(yaml.safe_load(f))`)."""

    def __init__(self, d=None):
        super().__init__()
        for k, v in (d or {}).items():
            self[k] = CfgNode(v) if isinstance(v, dict) else v

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError:
            raise AttributeError(k)


def build_models(cfg, device="cuda"):
    """train_nerf.py:144-162 / eval_nerf.py:97-125."""
    def one(c):
        return getattr(_models, c.type)(num_encoding_fn_xyz=c.num_encoding_fn_xyz, num_encoding_fn_dir=c.num_encoding_fn_dir,
                                        include_input_xyz=c.include_input_xyz, include_input_dir=c.include_input_dir,
                                        use_viewdirs=c.use_viewdirs).to(device)
    coarse = one(cfg.models.coarse)
    fine = one(cfg.models.fine) if hasattr(cfg.models, "fine") else None
    return coarse, fine


def make_train_step(cfg, model_coarse, model_fine, precision="fp32", cuda_graph=True, **kw):
    t = cfg.nerf.train
    return TrainStep(model_coarse, model_fine, N_samples=t.num_coarse, N_importance=t.num_fine if model_fine is not None else 0,
                     perturb=float(bool(t.perturb)), raw_noise_std=t.radiance_field_noise_std, white_bkgd=t.white_background,
                     lindisp=t.lindisp, precision=precision, lrate=cfg.optimizer.lr, lrate_decay=cfg.scheduler.lr_decay,
                     decay_rate=cfg.scheduler.lr_decay_factor, cuda_graph=cuda_graph, **kw)


def load_checkpoint(path_or_dict, model_coarse, model_fine, train_step=None):
    """train_nerf.py:184-190 (and eval_nerf.py:127-142 without a train step).  Returns start_iter."""
    ck = path_or_dict if isinstance(path_or_dict, dict) else torch.load(path_or_dict, map_location="cpu", weights_only=False)
    model_coarse.load_state_dict(ck["model_coarse_state_dict"])
    if ck.get("model_fine_state_dict") and model_fine is not None:
        model_fine.load_state_dict(ck["model_fine_state_dict"])
    if train_step is not None and ck.get("optimizer_state_dict"):
        train_step.load_optimizer_state_dict(ck["optimizer_state_dict"])
    return int(ck.get("iter", 0))


def checkpoint_dict(i, model_coarse, model_fine, train_step, loss, psnr):
    """The dict train_nerf.py:401-409 saves (tensors on the CPU so the file loads anywhere)."""
    cpu = lambda sd: {k: v.detach().cpu().clone() for k, v in sd.items()}
    osd = train_step.optimizer_state_dict()
    osd = {"state": {k: {kk: (vv.detach().cpu().clone() if isinstance(vv, torch.Tensor) else vv) for kk, vv in st.items()}
                     for k, st in osd["state"].items()}, "param_groups": osd["param_groups"]}
    return {"iter": i, "model_coarse_state_dict": cpu(model_coarse.state_dict()),
            "model_fine_state_dict": None if model_fine is None else cpu(model_fine.state_dict()),
            "optimizer_state_dict": osd, "loss": loss, "psnr": psnr}


def train(cfg, images, poses, hwf, i_train, i_val=None, *, load_checkpoint_path="", iters=None, logdir=None,
          precision="fp32", device="cuda", log=print):
    """The loop of train_nerf.py:194-415.  Returns (model_coarse, model_fine, train_step, last checkpoint dict)."""
    H, W, focal = int(hwf[0]), int(hwf[1]), float(hwf[2])
    np.random.seed(cfg.experiment.randomseed)
    torch.manual_seed(cfg.experiment.randomseed)
    coarse, fine = build_models(cfg, device)
    ts = make_train_step(cfg, coarse, fine, precision=precision, seed=cfg.experiment.randomseed)
    start = 0
    if load_checkpoint_path and os.path.exists(load_checkpoint_path):
        start = load_checkpoint(load_checkpoint_path, coarse, fine, ts)
    n_iters = cfg.experiment.train_iters if iters is None else start + iters
    images = torch.as_tensor(images, dtype=torch.float32).to(device)
    ndc = cfg.dataset.no_ndc is False
    last = None
    for i in range(start, n_iters):
        ts.lr_offset = i - 1 - ts.global_step            # iteration i uses the rate set after iteration i - 1
        img_idx = int(np.random.choice(i_train))
        k = np.random.choice(H * W, size=(cfg.nerf.train.num_random_rays), replace=False)
        pix = torch.from_numpy((k % H) * W + (k // H))   # coords = meshgrid_xy(arange(H), arange(W)).reshape(-1, 2)
        loss2 = ts.step_pixels(H, W, focal, np.asarray(poses[img_idx], dtype=np.float32)[:3, :4], images[img_idx][..., :3],
                               pix, cfg.dataset.near, cfg.dataset.far, ndc=ndc)
        want_log = i % cfg.experiment.print_every == 0 or i == n_iters - 1
        want_save = i % cfg.experiment.save_every == 0 or i == n_iters - 1
        if want_log or want_save:
            loss = float(loss2.sum())
            psnr = mse2psnr(loss)
            if want_log:
                log("[TRAIN] Iter: %d Loss: %s PSNR: %s" % (i, loss, psnr))
            if want_save:
                last = checkpoint_dict(i, coarse, fine, ts, loss, psnr)
                if logdir is not None:
                    os.makedirs(logdir, exist_ok=True)
                    torch.save(last, os.path.join(logdir, "checkpoint" + str(i).zfill(5) + ".ckpt"))
        if i_val is not None and len(i_val) and (i % cfg.experiment.validate_every == 0 or i == n_iters - 1):
            j = int(np.random.choice(i_val))
            rgb_c, rgb_f = validate(cfg, coarse, fine, H, W, focal, poses[j], precision=precision)
            tgt = images[j][..., :3]
            vloss = float(torch.mean((rgb_c[..., :3] - tgt) ** 2)) + (float(torch.mean((rgb_f[..., :3] - tgt) ** 2)) if rgb_f is not None else 0.0)
            log("Validation loss: %s Validation PSNR: %s" % (vloss, mse2psnr(vloss)))
    return coarse, fine, ts, last


def validate(cfg, model_coarse, model_fine, H, W, focal, pose, precision="fp32"):
    """One full frame in validation mode (train_nerf.py:345-362 / eval_nerf.py:157-166) -> (rgb_coarse, rgb_fine)."""
    from . import ops
    with torch.no_grad():
        ro, rd = ops.get_rays(H, W, focal, np.asarray(pose, dtype=np.float32)[:3, :4])
        out = run_one_iter_of_nerf(H, W, focal, model_coarse, model_fine, ro, rd, cfg, mode="validation", precision=precision)
    return out[0], out[3]


def evaluate(cfg, checkpoint, render_poses, hwf, savedir=None, precision="fp32", device="cuda"):
    """eval_nerf.py:127-189: load a checkpoint, render every pose, optionally write <savedir>/<i:04d>.png.
    Returns the list of [H, W, 3] images (fine when there is a fine network)."""
    from .ops import to8b_psnr
    from .run_nerf import _write_png
    coarse, fine = build_models(cfg, device)
    load_checkpoint(checkpoint, coarse, fine)
    H, W, focal = int(hwf[0]), int(hwf[1]), float(hwf[2])
    frames = []
    for i, pose in enumerate(render_poses):
        rgb_c, rgb_f = validate(cfg, coarse, fine, H, W, focal, pose, precision=precision)
        rgb = rgb_f if rgb_f is not None else rgb_c
        frames.append(rgb[..., :3])
        if savedir is not None:
            os.makedirs(savedir, exist_ok=True)
            u8, _ = to8b_psnr(rgb[..., :3].contiguous())
            _write_png(os.path.join(savedir, "%04d.png" % i), u8.cpu().numpy())
    return frames
