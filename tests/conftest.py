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
    """Build librdyb200.so and the CPU oracles once per session (no-op when up to date / prebuilt)."""
    from rdycore_b200 import build

    if not os.path.exists(build.LIB) or os.path.isdir(os.path.join(ROOT, "rdycore_b200", "csrc")) and build.stale():
        build.build_library()
    build.build_oracles()
    yield


def has_gpu() -> bool:
    try:
        import torch

        return torch.cuda.is_available()
    except Exception:
        return False
