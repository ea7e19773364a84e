import json
import os
import sys

import numpy as np
import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if REPO not in sys.path:
    sys.path.insert(0, REPO)
GOLDEN = os.path.join(REPO, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


def pytest_collection_modifyitems(config, items):
    """Every GPU test gets a 5-minute limit enforced from a watchdog thread: a kernel that never finishes blocks the
    main thread inside the CUDA runtime, where a signal-based timeout cannot fire, and would hold the GPU box until
    the outer limit."""
    for item in items:
        if item.get_closest_marker("gpu") is not None:
            item.add_marker(pytest.mark.timeout(300, method="thread"))


@pytest.fixture(scope="session")
def golden_small():
    with open(os.path.join(GOLDEN, "golden_small.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def golden_mnist():
    return dict(np.load(os.path.join(GOLDEN, "golden_mnist.npz")))


@pytest.fixture(scope="session")
def cuda_device():
    import torch
    if not torch.cuda.is_available():
        pytest.fail("GPU test selected but no CUDA device is visible (there is no CPU fallback)")
    return torch.device("cuda", 0)
