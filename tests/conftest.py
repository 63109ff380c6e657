import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
GOLD = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    import pyoracle
    return pyoracle.Oracle()


@pytest.fixture(scope="session")
def ref_strict():
    import pyoracle
    if not pyoracle.Ref.available(strict=True):
        pytest.skip("oracle/_ref not built (reference tree absent)")
    return pyoracle.Ref(strict=True)


@pytest.fixture(scope="session")
def gold():
    cache = {}

    def load(name):
        if name not in cache:
            cache[name] = np.load(os.path.join(GOLD, name + ".npz"))
        return cache[name]
    return load


@pytest.fixture(scope="session")
def dsp():
    """The product binding; building is the job of __graft_entry__.build()."""
    import crlot_dsp_b200
    return crlot_dsp_b200


@pytest.fixture(scope="session")
def cuda(dsp):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    dsp.lib()           # raises loudly if the extension is missing: no fallback
    return torch


def bits(a):
    a = np.ascontiguousarray(a)
    return a.view(np.uint64) if a.dtype == np.complex64 else a.view(np.uint32)


def rel_l2(a, b):
    a = np.asarray(a).astype(np.complex128 if np.iscomplexobj(a) else np.float64).ravel()
    b = np.asarray(b).astype(a.dtype).ravel()
    den = np.linalg.norm(b)
    return float(np.linalg.norm(a - b) / den) if den > 0 else float(np.linalg.norm(a - b))
