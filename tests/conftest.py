import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def orc():
    import oracle
    return oracle.Oracle()


@pytest.fixture(scope="session")
def ref():
    import oracle
    oracle.build()
    if not oracle.Ref.available():
        pytest.skip("oracle/_ref/libsl_ref.so not built (no /root/reference here)")
    return oracle.Ref()


@pytest.fixture(scope="session")
def file0(orc):
    return orc.file0(want_rejections=True)
