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
    from oracle.oracle import Oracle
    return Oracle("port")


@pytest.fixture(scope="session")
def ref():
    """The reference's own code (oracle/_ref).  Built here from /root/reference; on the GPU box
    only the prebuilt .so travels.  Skips when neither is available."""
    from oracle.oracle import Oracle, have_reference
    if not have_reference():
        pytest.skip("oracle/_ref not built and /root/reference absent")
    return Oracle("reference")


@pytest.fixture(scope="session", params=["port", "reference"])
def any_oracle(request):
    from oracle.oracle import Oracle, have_reference
    if request.param == "reference" and not have_reference():
        pytest.skip("oracle/_ref not available")
    return Oracle(request.param)
