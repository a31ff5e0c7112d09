import os
import sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(ROOT, "open-gpu-kernel-modules_b200"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Builds the CPU checkers / emulator once per session if they are missing."""
    need = [
        os.path.join(ROOT, "oracle", "liboracle.so"),
        os.path.join(ROOT, "tests", "simt", "libsimt_kernels.so"),
        os.path.join(ROOT, "open-gpu-kernel-modules_b200", "libuvmtier.so"),
    ]
    if not all(os.path.exists(p) for p in need):
        sys.path.insert(0, ROOT)
        import __graft_entry__

        __graft_entry__.build()


@pytest.fixture(scope="session")
def ctx():
    """One uvmtier context on cuda:0 for the GPU tests."""
    import uvmtier

    c = uvmtier.Context(0)
    yield c
    c.close()
