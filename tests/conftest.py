import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: test needs a CUDA device (B200); run with -m gpu")


@pytest.fixture(scope="session")
def lib():
    from robomanipbaselines_b200 import _lib

    return _lib.load()


@pytest.fixture(scope="session")
def cuda():
    import torch

    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from robomanipbaselines_b200 import _lib

    _lib.check(_lib.load().rmb_check_device(0), "rmb_check_device")
    return torch.device("cuda:0")


@pytest.fixture(autouse=True)
def _no_device_fault(request):
    """After every GPU test: the kernels' fault flag (bounded mbarrier waits, see ptx.cuh) must still be clear."""
    yield
    if request.node.get_closest_marker("gpu") is None:
        return
    import torch

    if not torch.cuda.is_available():
        return
    from robomanipbaselines_b200 import _lib

    assert _lib.load().rmb_device_fault(1) == 0, "a TMA/tcgen05 pipeline timed out inside a kernel"
