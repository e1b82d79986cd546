import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle():
    import orc
    return orc.Oracle()


@pytest.fixture(scope="session")
def refhost():
    import orc
    if not orc.RefHost.available():
        pytest.skip("oracle/_ref/libref_host.so not built (needs /root/reference)")
    return orc.RefHost()


@pytest.fixture(scope="session")
def rtb():
    """The product binding; building is __graft_entry__.build()'s job -- a missing library is an error, not a skip."""
    from cuda_raytracing_b200 import api
    api.lib()
    return api
