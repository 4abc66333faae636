"""Loads libssb200.so (the CUDA library behind include/ssb200.h).  There is no fallback: if the
library is missing or no sm_100 device is present, the caller gets an exception."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SSB200_LIB") or os.path.join(os.path.dirname(_HERE), "csrc", "libssb200.so")   # env: A/B builds
_lib = None

STATUS_OK = 0


class SolverError(RuntimeError):
    """Raised for any non-zero ssb200_status; the message is the kernelpp wording the reference's
    binding turns into RuntimeError (bindings/python/sparsesolvers/binding.cpp:48-62)."""


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise SolverError(f"{LIB_PATH} is not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(there is no CPU fallback)")
        l = C.CDLL(LIB_PATH)
        l.ssb200_status_string.restype = C.c_char_p
        l.ssb200_status_string.argtypes = [C.c_int]
        l.ssb200_last_error.restype = C.c_char_p
        l.ssb200_device_count.restype = C.c_int
        _lib = l
    return _lib


def check(status):
    if status != STATUS_OK:
        l = lib()
        msg = l.ssb200_status_string(status).decode()
        detail = l.ssb200_last_error().decode()
        raise SolverError(f"{msg}: {detail}" if detail else msg)
