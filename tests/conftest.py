import os
import sys

import numpy as np
import pytest
import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if REPO not in sys.path:
    sys.path.insert(0, REPO)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(REPO, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def load_golden(name):
    return dict(np.load(os.path.join(GOLDEN, name + ".npz")))


def sub(d, prefix, as_torch=True):
    n = len(prefix)
    out = {k[n:]: v for k, v in d.items() if k.startswith(prefix)}
    if as_torch:
        out = {k: torch.from_numpy(np.asarray(v)) for k, v in out.items()}
    return out


def rel_err(actual, ref):
    """max|actual-ref| / max|ref|  (the parity metric of BASELINE.md section 3)."""
    a = torch.as_tensor(actual).double().reshape(-1)
    r = torch.as_tensor(ref).double().reshape(-1)
    assert a.shape == r.shape, (a.shape, r.shape)
    if r.numel() == 0:
        return 0.0
    assert torch.isfinite(a).all(), "non-finite values"
    den = r.abs().max().item()
    return (a - r).abs().max().item() / (den if den > 0 else 1.0)


def assert_close(actual, ref, tol, what=""):
    e = rel_err(actual, ref)
    assert e <= tol, f"{what}: max|err|/max|ref| = {e:.3e} > {tol:.1e}"
    return e


@pytest.fixture(scope="session")
def golden():
    cache = {}

    def get(name):
        if name not in cache:
            cache[name] = load_golden(name)
        return cache[name]
    return get
