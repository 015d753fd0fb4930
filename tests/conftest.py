import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
PKG = os.path.join(ROOT, "alphazero-gomoku_b200")
for p in (ROOT, PKG, GOLDEN, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


def golden(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))


@pytest.fixture(scope="session")
def weights():
    from weights import make_state_dict
    return dict(pv=make_state_dict("pv", 1234), graphnet=make_state_dict("graphnet", 1234),
                dqn=make_state_dict("dqn", 1234))


@pytest.fixture(scope="session")
def oracle(weights):
    import oracle as orc
    return orc.Oracle(**weights)
