import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200, sm_100a); run with -m gpu")


def pytest_sessionstart(session):
    """Make sure libb200sft.so exists and is current before any test touches it (nvcc is present both in the
    build container and on the GPU box; a missing compiler is reported by the tests that need the library)."""
    try:
        from llm_fine_tune_distributed_b200 import build

        build.build()
    except Exception as e:  # pragma: no cover - surfaced by _cabi.load() in the tests
        print(f"[conftest] could not build libb200sft.so: {e}", file=sys.stderr)


@pytest.fixture(scope="session")
def cuda_device():
    import torch

    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda:0")
