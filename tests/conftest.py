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
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def load_golden(name: str):
    return dict(np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False))


def from_bits(bits: np.ndarray, dtype: torch.dtype) -> torch.Tensor:
    """uint16 bit patterns -> bf16/fp16 tensor (same shape)."""
    return torch.from_numpy(bits.view(np.int16).copy()).view(dtype)


def to_bits(t: torch.Tensor) -> np.ndarray:
    return t.detach().cpu().contiguous().view(torch.int16).numpy().view(np.uint16)


def golden_logits(g):
    if "logits_f32" in g:
        return torch.from_numpy(g["logits_f32"].copy()), torch.from_numpy(g["dlogits_f32"].copy())
    dt = torch.bfloat16 if str(g["dtype"]) == "bf16" else torch.float16
    return from_bits(g["logits_bits"], dt), from_bits(g["dlogits_bits"], dt)


@pytest.fixture(scope="session")
def cuda():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda:0")
