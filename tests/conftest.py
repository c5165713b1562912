import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle():
    """The CPU checker (oracle/): test infrastructure, never the product."""
    import oracle as orc
    orc.build()
    return orc


@pytest.fixture(scope="session")
def rdl():
    """The product: ctypes view of libradler_b200.so. Fails loudly when it is not built."""
    import ska_sdp_func_radler_b200 as m
    m.lib()
    return m


@pytest.fixture(scope="session")
def torch_cuda():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return torch


@pytest.fixture(params=["persistent", "stepwise"])
def hogbom_loop_flavour(request, monkeypatch):
    """GenericClean path B has two drivers with the same arithmetic: one cooperative kernel for
    cubes that stay in L2 (the default up to 40 MB) and one fused kernel launch per iteration.
    Tests at small sizes run under both."""
    if request.param == "stepwise":
        monkeypatch.setenv("RDL_HOGBOM_PERSISTENT_MB", "0")
    else:
        monkeypatch.delenv("RDL_HOGBOM_PERSISTENT_MB", raising=False)
    return request.param
