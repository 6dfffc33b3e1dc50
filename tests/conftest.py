import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200; run with -m gpu on the GPU box")


@pytest.fixture(scope="session")
def oracle():
    import oracle_lib
    oracle_lib.lib()
    return oracle_lib


@pytest.fixture(scope="session")
def ctx():
    """One grit_ctx on cuda:0 for the GPU tests; fails loudly when the library or device is missing."""
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    from grit_b200.buffers import Context
    c = Context(0)
    yield c
    c.close()
