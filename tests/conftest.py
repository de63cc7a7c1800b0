import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def ops():
    """The product op surface. Fails loudly (no fallback) if the extension is not built."""
    import torch
    assert torch.cuda.is_available(), "gpu-marked test collected without a GPU"
    import neural_magic_vllm_b200 as pkg
    return pkg.ops
