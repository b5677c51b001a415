import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests", "golden")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def svr2():
    """The product package.  The library is built in-tree; if it is missing this fails loudly (no fallback)."""
    import seedvr2_b200
    if not os.path.exists(seedvr2_b200._lib.LIB_PATH):
        seedvr2_b200.build_extension()
    seedvr2_b200._lib.lib()
    return seedvr2_b200


@pytest.fixture(scope="session")
def cuda(svr2):
    import torch
    assert torch.cuda.is_available(), "gpu-marked test needs a CUDA device"
    return torch.device("cuda:0")
