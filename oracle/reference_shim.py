"""Import the reference's own PyTorch implementation, unmodified.

TEST / BASELINE INFRASTRUCTURE ONLY.  The package is imported from /root/reference when that exists (the build
container) and otherwise from oracle/_ref/, the verbatim copy staged by oracle/build_ref.py (git-ignored; it
travels to the GPU box with the snapshot).  Used by oracle/make_golden.py to generate tests/golden/*.npz, by the
tests that pin the oracle, and by bench.py's reference arm / cpu_baseline leg (oracle/ref_arm.py).

Two third-party modules the reference imports are absent from this image and are shimmed:
  * torchsearchsorted (nerf-pytorch/requirements.txt:9, unpinned git HEAD) ->
    torch.searchsorted(right=(side == 'right'))
  * imageio (only used by the dataset loaders) -> empty module
"""
import os
import sys
import types

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))


def _find_root():
    for c in (os.environ.get("NERF_REFERENCE_ROOT"), "/root/reference", os.path.join(_HERE, "_ref")):
        if c and os.path.isdir(os.path.join(c, "nerf-pytorch", "nerf")):
            return c
    return "/root/reference"


REF_ROOT = _find_root()
REF_PT = os.path.join(REF_ROOT, "nerf-pytorch")


def available() -> bool:
    return os.path.isdir(os.path.join(REF_PT, "nerf"))


def load():
    """Returns the imported reference package ``nerf`` (nerf-pytorch/nerf/__init__.py)."""
    if not available():
        raise RuntimeError("reference tree not present at %s" % REF_PT)
    if "torchsearchsorted" not in sys.modules:
        m = types.ModuleType("torchsearchsorted")

        def searchsorted(a, v, out=None, side="left"):
            return torch.searchsorted(a, v, right=(side == "right"))

        m.searchsorted = searchsorted
        sys.modules["torchsearchsorted"] = m
    if "imageio" not in sys.modules:
        sys.modules["imageio"] = types.ModuleType("imageio")
    if REF_PT not in sys.path:
        sys.path.insert(0, REF_PT)
    import nerf as ref_nerf  # noqa: E402  (the reference package)

    return ref_nerf


def make_options(ref_nerf, *, use_viewdirs=True, no_ndc=True, near=2.0, far=6.0, num_coarse=64,
                 num_fine=128, chunksize=32768, perturb=False, lindisp=False, noise=0.0,
                 white_background=False):
    """A CfgNode with exactly the fields run_one_iter_of_nerf reads (train_utils.py:130-202)."""
    mode = dict(chunksize=chunksize, num_coarse=num_coarse, num_fine=num_fine, perturb=perturb,
                lindisp=lindisp, radiance_field_noise_std=noise, white_background=white_background)
    cfg = dict(nerf=dict(use_viewdirs=use_viewdirs, train=dict(mode), validation=dict(mode)),
               dataset=dict(no_ndc=no_ndc, near=near, far=far))
    return ref_nerf.CfgNode(cfg)
