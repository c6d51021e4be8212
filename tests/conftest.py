import os
import sys

import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if REPO not in sys.path:
    sys.path.insert(0, REPO)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def built():
    """Everything the suite loads: the CUDA engine, the host mirror, the oracle, the CPU probe."""
    import __graft_entry__ as ge
    ge.build()
    return True


@pytest.fixture(scope="session")
def orc(built):
    import oracle_lib
    return oracle_lib.Oracle()


@pytest.fixture(scope="session")
def ref(built):
    import oracle_lib
    if not os.path.exists(oracle_lib.LIB_REF):
        pytest.skip("oracle/_ref/libsspp_ref.so not built (needs /root/reference)")
    return oracle_lib.Oracle(oracle_lib.LIB_REF)


@pytest.fixture()
def eng(built):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from sspp_b200.engine import Engine
    e = Engine(0)
    yield e
    e.close()
