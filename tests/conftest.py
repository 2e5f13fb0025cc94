import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(autouse=True)
def _release_device_tensors():
    yield
    if "tests._gpu_util" not in sys.modules:
        return
    _gpu_util = sys.modules["tests._gpu_util"]
    bad = _gpu_util.check_guards()  # guard bands around every device buffer a kernel test handed to the C ABI
    _gpu_util._KEEP.clear()
    assert not bad, "a kernel wrote outside a buffer it was given: " + "; ".join(bad)
