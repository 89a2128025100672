import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLD = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


def load_golden(name):
    d = np.load(os.path.join(GOLD, name + ".npz"))
    return {k: torch.from_numpy(np.asarray(d[k])) if d[k].ndim > 0 else d[k].item() for k in d.files}


@pytest.fixture
def golden():
    return load_golden


def maxdiff(a, b, nan_ok=True):
    a, b = torch.as_tensor(a).detach().cpu().double(), torch.as_tensor(b).detach().cpu().double()
    assert a.shape == b.shape, (a.shape, b.shape)
    if nan_ok:
        both = torch.isnan(a) & torch.isnan(b)
        assert bool((torch.isnan(a) == torch.isnan(b)).all()), "NaN pattern differs"
        a = torch.where(both, torch.zeros_like(a), a)
        b = torch.where(both, torch.zeros_like(b), b)
    return float((a - b).abs().max()) if a.numel() else 0.0
