"""ctypes binding of libgabril_b200.so (the C-ABI boundary, include/gabril_b200.h).

There is no CPU or PyTorch fallback: if the library is missing, or a call returns a non-zero
status, a RuntimeError is raised (the reference raises Python exceptions on this path too).
"""
import ctypes
import os

import torch

# GABRIL_B200_LIB: another build of the same library (A/B experiments of kernel variants)
_LIB_PATH = os.environ.get("GABRIL_B200_LIB") or os.path.join(os.path.dirname(os.path.abspath(__file__)),
                                                              "libgabril_b200.so")
_lib = None

c_void_p = ctypes.c_void_p
c_int = ctypes.c_int
c_float = ctypes.c_float
c_ll = ctypes.c_longlong


def lib_path() -> str:
    return _LIB_PATH


def load() -> ctypes.CDLL:
    """Load the shared library (no GPU needed for loading / symbol lookup)."""
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            raise RuntimeError(
                f"libgabril_b200.so not found at {_LIB_PATH}: build it with "
                "`python -c 'import __graft_entry__ as g; g.build()'` (make -C gabril-carla_b200/csrc). "
                "There is no CPU fallback.")
        _lib = ctypes.CDLL(_LIB_PATH)
        _lib.gabril_last_error.restype = ctypes.c_char_p
    return _lib


_inited = False


def init():
    """Load + check the device is sm_100 + resolve the TMA encoders."""
    global _inited
    lib = load()
    if not _inited:
        if not torch.cuda.is_available():
            raise RuntimeError("gabril-carla_b200 needs a CUDA device (sm_100a); no CPU fallback")
        check(lib.gabril_init(), "gabril_init")
        _inited = True
    return lib


def check(status: int, what: str = ""):
    if status != 0:
        msg = load().gabril_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"libgabril_b200 {what} failed with status {status}: {msg}")


def ptr(t):
    """Device pointer of a tensor (None -> NULL)."""
    if t is None:
        return c_void_p(0)
    return c_void_p(t.data_ptr())


def stream():
    return c_void_p(torch.cuda.current_stream().cuda_stream)


def poll_error():
    """Raise if a kernel reported a protocol timeout (device-side flag)."""
    v = load().gabril_poll_error()
    if v != 0:
        raise RuntimeError(f"libgabril_b200: device-side error flag {v} (mbarrier wait timeout)")
