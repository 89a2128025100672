"""tests/golden/ckpt_lego_lowres3_optim.npz: the `optimizer_state_dict` of the shipped checkpoint
trained_models/nerf-pytorch/lego-lowres3/checkpoint110000.ckpt (torch.optim.Adam over
list(model_coarse.parameters()) + list(model_fine.parameters()), train_nerf.py:165-169), so that the checkpoint
round trip of SURVEY 8(f) f1 can be tested where /root/reference does not exist.  TEST INFRASTRUCTURE ONLY.

    python oracle/make_optim_golden.py
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CKPT = os.path.join(os.environ.get("NERF_REFERENCE_ROOT", "/root/reference"), "trained_models", "nerf-pytorch", "lego-lowres3",
                    "checkpoint110000.ckpt")


def main():
    ck = torch.load(CKPT, weights_only=False, map_location="cpu")
    osd = ck["optimizer_state_dict"]
    out = {"iter": np.int64(ck["iter"]), "loss": np.float64(float(ck["loss"])), "psnr": np.float64(float(ck["psnr"])),
           "lr": np.float64(osd["param_groups"][0]["lr"]), "n": np.int64(len(osd["state"]))}
    for i, st in osd["state"].items():
        out["step_%d" % i] = np.float64(float(st["step"]))
        out["exp_avg_%d" % i] = st["exp_avg"].numpy()
        out["exp_avg_sq_%d" % i] = st["exp_avg_sq"].numpy()
    dst = os.path.join(ROOT, "tests", "golden", "ckpt_lego_lowres3_optim.npz")
    np.savez_compressed(dst, **out)
    print("wrote", dst, os.path.getsize(dst), "iter", ck["iter"], "lr", out["lr"], "step", out["step_0"])
    # the weights fixture (ckpt_lego_lowres3.npz, oracle/make_golden.py) must be the same checkpoint
    w = np.load(os.path.join(ROOT, "tests", "golden", "ckpt_lego_lowres3.npz"))
    for k, v in ck["model_coarse_state_dict"].items():
        assert np.array_equal(w["c." + k], v.numpy()), k


if __name__ == "__main__":
    main()
