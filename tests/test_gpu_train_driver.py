"""GPU: SURVEY 8(f) f1 -- resume the shipped lego-lowres3 checkpoint INCLUDING its optimizer_state_dict, run the thin
train loop (nerf_kinematics_b200/train_nerf.py, the body of nerf-pytorch/train_nerf.py:194-415), save, and load the
result in the reference's own model class / torch.optim.Adam (oracle/_ref)."""
import os
import tempfile

import numpy as np
import pytest
import torch

from conftest import GOLD, load_golden
from oracle import reference_shim

pytestmark = pytest.mark.gpu

import nerf_kinematics_b200 as nk  # noqa: E402
from nerf_kinematics_b200 import train_nerf as T  # noqa: E402

CFG = {
    "dataset": {"far": 6, "near": 2, "no_ndc": True, "type": "blender"},
    "experiment": {"id": "lego-lowres3", "logdir": "logs", "print_every": 5, "randomseed": 42, "save_every": 5000,
                   "train_iters": 200000, "validate_every": 1000000},
    "models": {k: {"hidden_size": 128, "include_input_dir": True, "include_input_xyz": True, "log_sampling_dir": True,
                   "log_sampling_xyz": True, "num_encoding_fn_dir": 4, "num_encoding_fn_xyz": 10, "num_layers": 8,
                   "skip_connect_every": 3, "type": "FlexibleNeRFModel", "use_viewdirs": True} for k in ("coarse", "fine")},
    "nerf": {"use_viewdirs": True,
             "train": {"chunksize": 131072, "lindisp": False, "num_coarse": 64, "num_fine": 64, "num_random_rays": 1024,
                       "perturb": True, "radiance_field_noise_std": 0.2, "white_background": False},
             "validation": {"chunksize": 131072, "lindisp": False, "num_coarse": 64, "num_fine": 64, "perturb": False,
                            "radiance_field_noise_std": 0.0, "white_background": False}},
    "optimizer": {"lr": 0.005, "type": "Adam"},
    "scheduler": {"lr_decay": 250, "lr_decay_factor": 0.1},
}


def _shipped_checkpoint():
    """The reference-format dict of trained_models/nerf-pytorch/lego-lowres3/checkpoint110000.ckpt from the two fixtures."""
    w = load_golden("ckpt_lego_lowres3")
    o = np.load(os.path.join(GOLD, "ckpt_lego_lowres3_optim.npz"))
    n = int(o["n"])
    state = {i: {"step": torch.tensor(float(o["step_%d" % i])), "exp_avg": torch.from_numpy(o["exp_avg_%d" % i]),
                 "exp_avg_sq": torch.from_numpy(o["exp_avg_sq_%d" % i])} for i in range(n)}
    group = {"lr": float(o["lr"]), "betas": (0.9, 0.999), "eps": 1e-08, "weight_decay": 0, "amsgrad": False, "maximize": False,
             "foreach": None, "capturable": False, "differentiable": False, "fused": None, "params": list(range(n))}
    return {"iter": int(o["iter"]), "model_coarse_state_dict": {k[2:]: v for k, v in w.items() if k.startswith("c.")},
            "model_fine_state_dict": {k[2:]: v for k, v in w.items() if k.startswith("f.")},
            "optimizer_state_dict": {"state": state, "param_groups": [group]}, "loss": float(o["loss"]), "psnr": float(o["psnr"])}


def test_shipped_optimizer_state_lands_on_the_right_tensors():
    cfg = T.CfgNode(CFG)
    ck = _shipped_checkpoint()
    coarse, fine = T.build_models(cfg)
    assert coarse.n_params == 84548                      # num_layers: 8 in the YAML is ignored, like the reference does
    ts = T.make_train_step(cfg, coarse, fine, precision="fp32", cuda_graph=False)
    assert T.load_checkpoint(ck, coarse, fine, ts) == 110000 and ts.global_step == 110002
    params = list(coarse.parameters()) + list(fine.parameters())    # train_nerf.py:165-167
    names = [n for n, _ in coarse.named_parameters()] * 2
    assert names[8] == "layers_dir.0.weight" and names[14] == "fc_feat.weight"      # registration, not native, order
    for i, p in enumerate(params):
        st = ts.opt.state[p]
        ref = ck["optimizer_state_dict"]["state"][i]
        assert tuple(st["exp_avg"].shape) == tuple(p.shape) == tuple(ref["exp_avg"].shape), (i, names[i])
        assert torch.equal(st["exp_avg"].cpu(), ref["exp_avg"]) and torch.equal(st["exp_avg_sq"].cpu(), ref["exp_avg_sq"])
        # ... and what the fused kernel updates is that very memory (views of the flat moment buffers)
        lo, hi = ts.exp_avg.data_ptr(), ts.exp_avg.data_ptr() + 4 * ts.exp_avg.numel()
        assert lo <= st["exp_avg"].data_ptr() < hi
    out = ts.optimizer_state_dict()
    assert out["param_groups"][0]["params"] == list(range(32)) and float(out["state"][0]["step"]) == 110002.0


def test_resume_train_save_and_reload_in_the_reference():
    cfg = T.CfgNode(CFG)
    ck = _shipped_checkpoint()
    H = W = 100
    g = torch.Generator().manual_seed(0)
    images = torch.rand(3, H, W, 3, generator=g)
    from nerf_kinematics_b200.poses import pose_spherical
    poses = np.stack([pose_spherical(a, -30.0, 4.0) for a in (0.0, 40.0, 80.0)])
    with tempfile.TemporaryDirectory() as tmp:
        src = os.path.join(tmp, "checkpoint110000.ckpt")
        torch.save(ck, src)
        lines = []
        coarse, fine, ts, last = T.train(cfg, images, poses, (H, W, 138.9), [0, 1], i_val=[2], load_checkpoint_path=src,
                                         iters=10, logdir=tmp, precision="fp32", log=lines.append)
        saved = os.path.join(tmp, "checkpoint110009.ckpt")
        assert os.path.exists(saved) and last["iter"] == 110009 and any(l.startswith("[TRAIN] Iter: 110005") for l in lines)
        new = torch.load(saved, map_location="cpu", weights_only=False)
    assert set(new) == set(ck)
    for part in ("model_coarse_state_dict", "model_fine_state_dict"):
        assert list(new[part]) == list(ck[part])
        moved = 0
        for k in ck[part]:
            assert new[part][k].shape == ck[part][k].shape and new[part][k].dtype == ck[part][k].dtype
            moved += int(not torch.equal(new[part][k], ck[part][k]))
        assert moved == len(ck[part]), "every tensor must have been trained"
    so, sn = ck["optimizer_state_dict"], new["optimizer_state_dict"]
    assert sn["param_groups"][0]["params"] == so["param_groups"][0]["params"] and len(sn["state"]) == 32
    for i in range(32):
        assert sn["state"][i]["exp_avg"].shape == so["state"][i]["exp_avg"].shape
        assert sn["state"][i]["exp_avg"].dtype == torch.float32 and float(sn["state"][i]["step"]) == 110012.0
        assert not torch.equal(sn["state"][i]["exp_avg"], so["state"][i]["exp_avg"])
    lr = sn["param_groups"][0]["lr"]
    assert abs(lr - 0.005 * 0.1 ** (110009 / 250000.0)) <= 1e-9     # the rate the reference's loop leaves behind (:292-297)
    assert np.isfinite(new["loss"]) and new["loss"] < 1.0
    if reference_shim.available():     # the reference's own classes accept the file (strict keys, optimizer by position)
        R = reference_shim.load()
        mc, mf = R.models.FlexibleNeRFModel(num_encoding_fn_xyz=10), R.models.FlexibleNeRFModel(num_encoding_fn_xyz=10)
        mc.load_state_dict(new["model_coarse_state_dict"])
        mf.load_state_dict(new["model_fine_state_dict"])
        opt = torch.optim.Adam(list(mc.parameters()) + list(mf.parameters()), lr=0.005)
        opt.load_state_dict(new["optimizer_state_dict"])
        for p, i in zip(opt.param_groups[0]["params"], range(32)):
            assert opt.state[p]["exp_avg"].shape == p.shape
        x = torch.rand(5, 63 + 27)
        assert torch.isfinite(mc(x)).all()
    else:
        pytest.fail("oracle/_ref is missing: run python oracle/build_ref.py before shipping the tree to the GPU box")
