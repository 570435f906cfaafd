import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")
    config.addinivalue_line("markers", "tf32: runs in the default TF32 execution mode instead of exact mode")


@pytest.fixture(scope="session")
def golden_ops():
    import numpy as np
    d = os.path.join(GOLDEN, "ops")

    def load(name):
        return np.load(os.path.join(d, name + ".npy"))
    return load


FP32_TOL = 1e-5
TF32_TOL = 2e-3


@pytest.fixture(scope="session")
def capi():
    import mobilefinetuner_b200.capi as capi
    assert capi.lib.mft_device_check() == 0, capi.lib.mft_last_error()
    return capi


def dev(a, dtype=None):
    import torch
    dtype = dtype or torch.float32
    return torch.as_tensor(np.ascontiguousarray(a), dtype=dtype).cuda()


def rel(a, b):
    import torch
    a = a.detach().cpu().numpy() if torch.is_tensor(a) else np.asarray(a)
    a, b = a.astype(np.float64), np.asarray(b, np.float64)
    assert a.shape == b.shape, (a.shape, b.shape)
    if b.size == 0:
        return 0.0
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


def ws(nbytes):
    import torch
    return torch.empty(max(int(nbytes), 16), dtype=torch.uint8, device="cuda")


