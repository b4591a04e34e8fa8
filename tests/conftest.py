import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on a B200)")


@pytest.fixture(scope="session")
def built():
    import __graft_entry__ as g
    g.build()
    return g


@pytest.fixture(scope="session")
def ctx(built):
    import p2p_llm_tunnel_b200 as p2pt
    return p2pt.Context(0)
