import os
import sys

import pytest

# Before CUDA starts in this process: the exchange tests emulate several ranks with several
# contexts on ONE GPU, and a rank's stage 2 spins on the GPU for its peers. Streams that share
# a hardware queue (8 by default) could put a peer's work behind such a spin.
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Make sure the checker libraries exist (gcc, seconds). The CUDA library is built by
    __graft_entry__.build(); tests that need it load it through chronos_b200._ffi and fail
    loudly if it is missing."""
    from oracle import build_oracle

    build_oracle.build_oracle()
    build_oracle.build_ref()
    yield
