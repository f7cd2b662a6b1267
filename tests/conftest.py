import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import __graft_entry__ as ge  # noqa: E402


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def ug():
    return ge.load_package()


@pytest.fixture(scope="session")
def synth(ug):
    import importlib
    return importlib.import_module(ge.PKG_NAME + ".synth")


@pytest.fixture(scope="session")
def orc():
    from oracle import oracle
    oracle.lib()
    return oracle


@pytest.fixture(scope="session")
def cv2ref():
    from oracle import cv2_ref
    return cv2_ref


CASCADES = ["alt", "alt2", "default", "alt_tree"]


def rectset(a):
    return sorted(map(tuple, [[int(v) for v in r] for r in a]))
