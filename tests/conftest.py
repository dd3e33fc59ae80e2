import os
import sys
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Build libb2pt.so, the oracle port and the test-only host build of the device headers once per session."""
    import __graft_entry__ as g
    if not os.path.exists(g.LIB) or os.environ.get("B2_REBUILD"):
        g.build()
    from oracle import oracle_py as O
    if not os.path.exists(O.PORT_SO):
        O.build()
    yield
