import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import __graft_entry__ as entry  # noqa: E402


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def vpt():
    """The vpt_b200 package (ctypes binding + archive tools)."""
    return entry.load_package()


@pytest.fixture(scope="session")
def root():
    return ROOT


def scene_path(name):
    """Built scene archive (scenes/_built, produced by build()) or committed fixture."""
    for base in (os.path.join(ROOT, "scenes", "_built"), os.path.join(ROOT, "tests", "golden")):
        p = os.path.join(base, name + ".vptscene")
        if os.path.exists(p):
            return p
    return None


REF_RENDER = os.path.join(ROOT, "oracle", "_ref", "ref_render")


def have_ref():
    return os.path.exists(REF_RENDER)
