import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def pkg():
    import b200pkg
    return b200pkg.load()


@pytest.fixture(scope="session")
def oracle_mod():
    from oracle import oracle as O
    if not os.path.exists(O.ORACLE_SO):
        O.build_oracle()
    return O


def have_ref():
    from oracle import oracle as O
    return os.path.exists(O.REF_SO)
