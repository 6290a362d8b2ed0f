import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def ref():
    """the real reference GDK (oracle/_ref), built by `make -C oracle ref`"""
    from oracle import refgdk
    if not refgdk.available():
        pytest.skip("oracle/_ref not built (needs /root/reference at build time)")
    refgdk.init(1)
    return refgdk


@pytest.fixture(scope="session")
def gdk():
    import monetdb_b200.gdk as g
    if g.lib.b200_device_count() < 1:
        pytest.skip("no CUDA device")
    g.init(0)
    return g
