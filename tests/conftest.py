import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


# The pruned entry points hand small problems to the dense kernel (how small depends on n_obs: see ob200_score_pruned_multi).
# The parity tests want the pruned path itself at sizes an oracle can follow, so the fallback is pinned to its minimum (one
# wave) for the whole session; tests/test_gpu_bound.py::test_dense_fallback_policy removes the pin to check the default.
os.environ.setdefault("OB200_PRUNE_MIN_BATCHES", "1")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def engine():
    from metaopt_b200.engine import get_engine
    return get_engine(0)
