"""ctypes binding of libramp_b200.so (include/ramp_b200.h).  No CPU fallback: importing the
product on a machine where the library is missing or fails to load raises."""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
_PATH = os.environ.get("RAMP_B200_LIB") or os.path.join(_HERE, "libramp_b200.so")   # override: A/B kernel variants
_lib = None

c_void_p, c_int, c_float, c_size_t = ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_size_t


class RampError(RuntimeError):
    pass


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_PATH):
            raise RampError("libramp_b200.so is not built (run `python -m ramp_b200.build`); there is no fallback path")
        _lib = ctypes.CDLL(_PATH)
        _lib.ramp_last_error.restype = ctypes.c_char_p
    return _lib


def check(status, what=""):
    if status != 0:
        raise RampError("%s failed (%d): %s" % (what, status, lib().ramp_last_error().decode()))


def ptr(t):
    """Device (or host) pointer of a torch tensor, None -> NULL."""
    return c_void_p(0 if t is None else t.data_ptr())


def stream_ptr():
    import torch
    return c_void_p(torch.cuda.current_stream().cuda_stream)
