import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _build_once():
    import __graft_entry__ as g
    lib = os.path.join(ROOT, "kreeq_b200", "libkreeq_b200.so")
    tools = [os.path.join(ROOT, "kreeq_b200", t) for t in ("kreeq", "kreeq-feedcheck", "kreeq-decompressor")]
    if not os.path.exists(lib) or not all(os.path.exists(t) for t in tools):
        g.build()


@pytest.fixture(scope="session", autouse=True)
def built():
    _build_once()
    yield
