import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "docker-nvidia-glx-desktop_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu through gpurun)")
    # the built libraries are not in git: a fresh checkout builds them before the first test
    if not os.path.exists(os.path.join(PKG, "libb200vc.so")):
        import __graft_entry__
        __graft_entry__.build()


@pytest.fixture(scope="session")
def orc():
    """The CPU oracle (test infrastructure; never imported by the product)."""
    import oracle
    return oracle.get()


@pytest.fixture(scope="session")
def lib():
    import b200vc
    return b200vc.load()
