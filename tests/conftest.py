import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def player():
    from piccolo_b200 import assets as A
    return A.load_player()


@pytest.fixture(scope="session")
def oracle():
    from oracle import pyoracle as O
    O.build(ref=True)  # no-op when up to date; oracle/_ref only when the reference tree is present
    return O.Oracle()


@pytest.fixture(scope="session")
def reflib_available():
    from oracle import pyoracle as O
    return O.RefLib.available()
