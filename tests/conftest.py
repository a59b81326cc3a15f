import os
import sys

import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if REPO not in sys.path:
    sys.path.insert(0, REPO)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:  # pragma: no cover
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session", autouse=True)
def _debug_library_knobs():
    """PC_TEST_PREFILL_PP=1 with PC_LIB_PATH pointing at the PC_DEBUG library: run the suite with the
    two-query-tile ping-pong prefill kernel forced on (A/B of the two tcgen05 kernels)."""
    if os.environ.get("PC_TEST_PREFILL_PP") in ("0", "1"):
        import ctypes
        from vllm_paged_compression_b200 import _lib
        fn = getattr(_lib.load(), "pc_debug_set_prefill_pp", None)
        if fn is None:
            raise RuntimeError("PC_TEST_PREFILL_PP needs the PC_DEBUG library (PC_LIB_PATH=...libpagedevict_b200_debug.so)")
        fn.argtypes = [ctypes.c_int32]
        fn.restype = None
        fn(int(os.environ["PC_TEST_PREFILL_PP"]))
    if os.environ.get("PC_TEST_PREFILL_DQ") in ("0", "1"):
        import ctypes
        from vllm_paged_compression_b200 import _lib
        fn = getattr(_lib.load(), "pc_debug_set_prefill_dq", None)
        if fn is None:
            raise RuntimeError("PC_TEST_PREFILL_DQ needs the PC_DEBUG library (PC_LIB_PATH=...libpagedevict_b200_debug.so)")
        fn.argtypes = [ctypes.c_int32]
        fn.restype = None
        fn(int(os.environ["PC_TEST_PREFILL_DQ"]))
    yield
