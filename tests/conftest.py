import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def llmc():
    """The ctypes binding with libllmc.so loaded (built in-tree if missing)."""
    import llm_consensus_b200.llmc as m
    if not os.path.exists(m.LIB_PATH):
        import llm_consensus_b200.build as b
        b.build()
    m.lib()
    return m


@pytest.fixture(scope="session")
def gpu(llmc):
    if llmc.device_count() < 1:
        pytest.fail("test is marked gpu but no CUDA device is visible -- libllmc has no CPU fallback")
    return llmc
