import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def golden(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False)


@pytest.fixture(scope="session")
def engine():
    """One libtcb200 context for the GPU tests. No skip-on-missing-library: on a GPU box a missing
    extension must fail loudly."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from triton_client_b200.gpu import Engine
    return Engine(0, workspace_bytes=512 << 20)
