import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    # test infrastructure: RAIN_TEST_LIB points this test process at another build of the library (the cross-check build,
    # rain_b200/librainrast_xc.so); the product never reads it
    alt = os.environ.get("RAIN_TEST_LIB")
    if alt:
        from rain_b200 import _lib
        _lib.LIB_PATH = alt


@pytest.fixture(scope="session")
def oracle():
    from oracle.oracle_py import Oracle
    return Oracle()


@pytest.fixture(scope="session")
def refcuda():
    from oracle.oracle_py import RefCuda, REF_SO
    if not os.path.exists(REF_SO):
        pytest.skip("oracle/_ref/libref_rast.so not built (needs /root/reference at build time)")
    return RefCuda()


@pytest.fixture(scope="session")
def device():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda:0")
