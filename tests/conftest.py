import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden", "golden_small.pt")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def golden():
    return torch.load(GOLDEN, map_location="cpu", weights_only=False)


@pytest.fixture(scope="session")
def golden_grads():
    """Parameter gradients of the unmodified reference for the golden_small.pt training steps
    (tests/golden/make_golden_grads.py)."""
    return torch.load(os.path.join(ROOT, "tests", "golden", "golden_grads.pt"), map_location="cpu", weights_only=False)


@pytest.fixture(scope="session")
def golden_grpo():
    """GRPO rollouts / objective of the unmodified reference (tests/golden/make_golden_grpo.py)."""
    return torch.load(os.path.join(ROOT, "tests", "golden", "golden_grpo.pt"), map_location="cpu", weights_only=False)
