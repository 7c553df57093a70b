import os
import sys
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200)")


@pytest.fixture(scope="session")
def ref():
    from oracle import refbind
    if not refbind.available():
        pytest.skip("oracle/_ref/libmobilert_ref.so not built (needs /root/reference at build time)")
    return refbind
