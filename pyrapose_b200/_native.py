"""ctypes binding of libpyrapose_b200.so (the C-ABI in include/pyrapose_b200.h).

The product path has no CPU fallback: if the shared library is missing this module raises at
first use instead of silently computing something else.
"""
import ctypes
import os
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_lib", "libpyrapose_b200.so")
_lock = threading.Lock()
_lib = None

PPB_ERR_INVALID = -1

c_void_p = ctypes.c_void_p
c_int = ctypes.c_int
c_float = ctypes.c_float
c_double = ctypes.c_double
c_longlong = ctypes.c_longlong


class NativeLibraryMissing(RuntimeError):
    pass


def lib_path():
    return _LIB_PATH


def load():
    """Returns the loaded CDLL; builds nothing, falls back to nothing."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is None:
            if not os.path.exists(_LIB_PATH):
                raise NativeLibraryMissing(
                    "%s not found: run `python -m pyrapose_b200.build` (or __graft_entry__.build()); "
                    "pyrapose_b200 has no CPU fallback" % _LIB_PATH)
            lib = ctypes.CDLL(_LIB_PATH)
            lib.ppb_last_error.restype = ctypes.c_char_p
            lib.ppb_launch_count.restype = c_longlong
            _lib = lib
    return _lib


def check(rc, what=""):
    """Maps a ppb_status to the exception the reference would raise for the same mistake."""
    if rc == 0:
        return
    msg = load().ppb_last_error().decode("utf-8", "replace")
    if rc == PPB_ERR_INVALID:
        raise ValueError(msg or what)
    raise RuntimeError("%s (ppb_status %d)" % (msg or what, rc))


def launch_count():
    return int(load().ppb_launch_count())


def ptr(t):
    """Device (or host) address of a torch tensor / numpy array as a c_void_p; None -> NULL."""
    if t is None:
        return c_void_p(0)
    if hasattr(t, "data_ptr"):
        return c_void_p(t.data_ptr())
    return c_void_p(t.ctypes.data)


def current_stream_ptr():
    import torch
    return c_void_p(torch.cuda.current_stream().cuda_stream)
