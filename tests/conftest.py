"""pytest configuration: `gpu` marker (tests that need a B200) and the repo root on sys.path."""
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def built_lib():
    """Build (if stale) and dlopen libafx_b200.so; works without a GPU (nvcc cross-compiles)."""
    import __graft_entry__ as g
    g.build()
    from audio_feature_extraction_b200 import _native
    return _native.lib()


@pytest.fixture(scope="module")
def torch_cuda(built_lib):
    import torch
    if not torch.cuda.is_available():
        pytest.fail("-m gpu tests need a CUDA device (there is no CPU fallback)")
    return torch
