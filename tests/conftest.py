import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "gaussian-splatting-npu_b200"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box: pytest -m gpu)")


def pytest_collection_modifyitems(config, items):
    # GPU tests must never silently pass without a GPU: outside `-m gpu` runs they are deselected by
    # the marker expression; if selected on a box without CUDA they fail loudly in their fixture.
    pass


@pytest.fixture(scope="session")
def cuda():
    import torch
    assert torch.cuda.is_available(), "this test needs a GPU (marked gpu)"
    return torch.device("cuda:0")
