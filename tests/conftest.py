import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.fixture(scope="session")
def ctx():
    """One ccicp context on cuda:0 through the C ABI; fails loudly (no fallback) when unusable."""
    from slam_b200 import capi
    c = capi.Context(0)
    yield c
    c.close()


@pytest.fixture(scope="session")
def oracle():
    from oracle import refbind
    if not (refbind.available("reference") or refbind.available("port")):
        pytest.fail("no oracle library built: run python -c 'import __graft_entry__ as g; g.build()'")
    return refbind
