import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def _cuda_ok():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _cuda_ok():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN


@pytest.fixture(scope="session")
def synth_weights():
    from mlx_audio_b200 import synth
    return synth.synth_decoder_weights(seed=2000)


@pytest.fixture(autouse=True)
def _dirty_allocator(request):
    """GPU tests start on DIRTY memory: blocks of assorted sizes are filled with NaN and handed back to torch's caching
    allocator, so that the ``torch.empty`` buffers a test (or the library's Python layer) allocates next do not happen to
    be zero.  compute-sanitizer's initcheck is closed on this pool (profiles/r02_sanitizer.txt); this is the stand-in that
    makes a kernel that reads what nobody wrote - or a store that was skipped - fail here instead of in production."""
    if "gpu" in request.keywords and _cuda_ok():
        import torch
        sizes = [1 << 12, 1 << 15, 1 << 17, 3 << 17, 1 << 19, 1 << 20, 5 << 20, 16 << 20]      # floats: 16 KB .. 64 MB
        junk = [torch.full((n,), float("nan"), device="cuda") for n in sizes for _ in range(2)]
        del junk
    yield
