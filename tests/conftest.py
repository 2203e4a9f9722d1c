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
    """The CPU oracle (test infrastructure).  Built on demand with gcc."""
    from oracle import ff_oracle

    ff_oracle.build()
    return ff_oracle


@pytest.fixture(scope="session")
def nb():
    """The product library bound to cuda:0.  Fails loudly if the .so or the GPU is missing."""
    from netsdb_b200 import capi

    capi.init(int(os.environ.get("LOCAL_RANK", "0")))
    yield capi
    capi.shutdown()
