import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    # GPU tests are skipped (not failed) where there is no device, so a bare `pytest tests/` works here.
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def fixture_forest():
    import numpy as np
    return dict(np.load(os.path.join(GOLDEN, "fixture_forest.npz"), allow_pickle=False))


@pytest.fixture(scope="session")
def oracle_small():
    import numpy as np
    return dict(np.load(os.path.join(GOLDEN, "oracle_small.npz"), allow_pickle=False))
