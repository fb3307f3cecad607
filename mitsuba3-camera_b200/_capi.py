"""ctypes binding of liblensrt.so (include/lensrt.h). The CUDA library is the product: importing this
module fails loudly if it has not been built, and every compute call fails with LRT_ENODEVICE when no
GPU is present — there is no CPU fallback."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "liblensrt.so")

c_f = C.c_float
FP = C.POINTER(C.c_float)
IP = C.POINTER(C.c_int32)


class LensrtError(RuntimeError):
    pass


class FilmGeom(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("crop_w", C.c_int32), ("crop_h", C.c_int32),
                ("crop_x", C.c_int32), ("crop_y", C.c_int32), ("diagonal_mm", C.c_float)]


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). lensrt has no CPU fallback.")
    return C.CDLL(LIB_PATH)


lib = _load()
lib.lrt_last_error.restype = C.c_char_p
lib.lrt_launch_count.restype = C.c_uint64
lib.lrt_radical_inverse.restype = C.c_float


def check(rc: int):
    if rc != 0:
        raise LensrtError(f"liblensrt error {rc}: {lib.lrt_last_error().decode()}")


def fp(a: np.ndarray):
    assert a.dtype == np.float32 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(FP)


def ip(a: np.ndarray):
    assert a.dtype == np.int32 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(IP)


def dev_ptr(t):
    """torch CUDA tensor -> void* (device pointer)."""
    assert t.is_cuda and t.is_contiguous()
    return C.c_void_p(t.data_ptr())


def stream_ptr():
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)
