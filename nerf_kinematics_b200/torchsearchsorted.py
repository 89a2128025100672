"""Same-signature stand-in for the reference's only native dependency, torchsearchsorted
(nerf-pytorch/requirements.txt:9; call sites nerf_helpers.py:241-243, 286-288), backed by nb_searchsorted."""
from . import ops


def searchsorted(a, v, out=None, side="left"):
    res = ops.searchsorted(a, v, side=side)
    if out is not None:
        out.copy_(res)
        return out
    return res
