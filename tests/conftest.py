"""Shared pytest fixtures.  ``-m gpu`` tests need a B200; everything else runs on CPU."""

import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_golden(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))


def golden_to_float(arr, dtype_name):
    """Golden array (uint16 bits for bf16) -> float32 numpy values."""
    from oracle.paged_attention_oracle import bf16_bits_to_f32

    if dtype_name == "bf16":
        return bf16_bits_to_f32(arr)
    return np.asarray(arr, dtype=np.float32)


def golden_to_torch(arr, dtype_name, device="cpu"):
    """Golden array -> torch tensor of the recorded dtype (bit-exact)."""
    import torch

    if dtype_name == "bf16":
        t = torch.from_numpy(np.ascontiguousarray(arr).view(np.int16)).view(torch.bfloat16)
    elif dtype_name == "f16":
        t = torch.from_numpy(np.ascontiguousarray(arr).view(np.int16)).view(torch.float16)
    else:
        t = torch.from_numpy(np.ascontiguousarray(arr))
    return t.to(device)


@pytest.fixture(scope="session")
def cuda_device():
    import torch

    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda:0")
