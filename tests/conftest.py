import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle():
    from oracle import Oracle
    return Oracle()


@pytest.fixture(scope="session")
def reference():
    """The unmodified reference compiled behind a C ABI (oracle/_ref); prebuilt files travel to
    the GPU box, /root/reference itself does not."""
    from oracle import Reference, have_reference, build
    if not have_reference() and Path("/root/reference/src").is_dir():
        build()
    if not have_reference():
        pytest.skip("oracle/_ref not built and /root/reference absent")
    return Reference()


@pytest.fixture(scope="session")
def femtc():
    import femtc_b200
    return femtc_b200
