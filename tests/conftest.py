import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def _cuda_ok() -> bool:
    try:
        from modularevogym_b200 import capi
        return capi.load().mevo_device_count() > 0
    except Exception:
        return False


@pytest.fixture(scope="session")
def cuda_available():
    return _cuda_ok()


def pytest_collection_modifyitems(config, items):
    # GPU tests must FAIL (not skip) on a GPU box if the extension is missing; on a CPU-only box
    # they are deselected by `-m "not gpu"`.  If someone runs them here anyway, skip with a reason.
    if _cuda_ok():
        return
    skip = pytest.mark.skip(reason="no CUDA device visible")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)
