import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def load_golden(name):
    """-> (OracleConfig kwargs, dict of numpy arrays)"""
    z = np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False)
    data = {k: z[k] for k in z.files}
    cfg = {}
    if "cfg_keys" in data:
        for k, v in zip(data["cfg_keys"], data["cfg_vals"]):
            cfg[str(k)] = float(v) if str(k) == "alpha" else int(v)
    return cfg, data


def golden_params(data, dtype=torch.float32):
    return {k[len("param."):]: torch.from_numpy(v).to(dtype) for k, v in data.items() if k.startswith("param.")}


def golden_grads(data):
    return {k[len("grad."):]: torch.from_numpy(v) for k, v in data.items() if k.startswith("grad.")}


def rel_err(a, b):
    """||a-b||_inf / ||b||_inf  (the parity norm of SURVEY.md §8c)."""
    a = torch.as_tensor(a, dtype=torch.float64)
    b = torch.as_tensor(b, dtype=torch.float64)
    denom = b.abs().max().item()
    return (a - b).abs().max().item() / (denom if denom > 0 else 1.0)


@pytest.fixture(scope="session")
def cuda_device():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda:0")
