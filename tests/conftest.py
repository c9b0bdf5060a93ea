import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def port():
    """The plain-C restatement oracle (always buildable)."""
    from oracle import bindings as B
    B.build(ref=False, port=True)
    return B.PortLib()


@pytest.fixture(scope="session")
def ref():
    """The compiled UNMODIFIED reference (oracle/_ref). Built when /root/reference exists, else the
    prebuilt .so that travelled with the repository snapshot; skipped when neither is there."""
    from oracle import bindings as B
    if os.path.isdir("/root/reference/src/vacancy"):
        B.build(ref=True, port=False)
    if not os.path.exists(B.REF_SO):
        pytest.skip("oracle/_ref/libvacancy_ref.so not available")
    return B.RefLib()


@pytest.fixture(scope="session")
def bunny():
    from tests import helpers
    return helpers.load_bunny()


@pytest.fixture(scope="session")
def cuda_lib():
    """The product library; GPU tests must fail loudly (not skip) when it is missing."""
    from vacancy_b200 import capi
    return capi.load_library()
