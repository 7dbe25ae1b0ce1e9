import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    import oracle_ref
    oracle_ref.build_oracle()
    if not oracle_ref.available():
        pytest.skip("oracle/_ref not built and /root/reference absent")
    return oracle_ref


@pytest.fixture(scope="session")
def nusi():
    import __graft_entry__ as ge
    ge.build()
    import nusiprop_b200
    return nusiprop_b200
