import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden_dir():
    return os.path.join(ROOT, "tests", "golden")


@pytest.fixture(scope="session")
def sc():
    """One MiningContext (one fa_ctx on cuda:0) shared by the GPU tests."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from fastapriori_b200 import MiningContext
    ctx = MiningContext(device=0)
    yield ctx
    ctx.close()
