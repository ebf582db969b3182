"""ctypes binding of libfiode_b200.so (the C ABI declared in include/fiode_b200.h).

There is NO fallback: if the shared library is missing or a call fails, this raises.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

_HERE = Path(__file__).resolve().parent
LIB_PATH = _HERE / 'libfiode_b200.so'

_lib = None


class FiodeError(RuntimeError):
    pass


class DynWeights(C.Structure):
    _fields_ = [('Q1', C.c_void_p), ('b1', C.c_void_p), ('Qx', C.c_void_p), ('bx', C.c_void_p),
                ('Q2', C.c_void_p), ('b2', C.c_void_p), ('Q3', C.c_void_p), ('b3', C.c_void_p),
                ('alpha_1', C.c_float), ('alpha_2', C.c_float), ('sigma_1', C.c_float),
                ('scale_nominal', C.c_int), ('max_iter', C.c_int), ('tol', C.c_float)]


class BackboneWeights(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ('mu', 'std', 'Qf0', 'b0', 'Q2', 'b2', 'Qf4', 'b4', 'Q6', 'b6',
                                          'Q9', 'b9', 'Q11', 'b11', 'Qfc', 'bfc')]


class SolveStats(C.Structure):
    _fields_ = [('n_accept', C.c_int), ('n_reject', C.c_int), ('nfe', C.c_int), ('status', C.c_int),
                ('t_final', C.c_double), ('last_dt', C.c_double), ('n_logged', C.c_int), ('pad_', C.c_int),
                ('dts', C.c_double * 64), ('ratios', C.c_float * 64)]


def lib() -> C.CDLL:
    """Load the library (once).  Raises FiodeError if it has not been built."""
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise FiodeError(f'{LIB_PATH} is missing: run `python -m fiode_b200.build` '
                             '(there is no CPU / PyTorch fallback for the fiode_b200 hot path)')
        _lib = C.CDLL(str(LIB_PATH))
        _lib.fiode_last_error.restype = C.c_char_p
        _lib.fiode_launch_count.restype = C.c_longlong
        for name in dir(_lib):
            pass
        for fn in SIZE_T_FUNCS:
            getattr(_lib, fn).restype = C.c_size_t
    return _lib


SIZE_T_FUNCS = ()


def check(rc: int, what: str = ''):
    if rc != 0:
        msg = lib().fiode_last_error().decode(errors='replace')
        raise FiodeError(f'{what} failed (rc={rc}): {msg}')


def ptr(t):
    """Raw device pointer of a contiguous CUDA tensor (or None)."""
    if t is None:
        return None
    assert t.is_cuda and t.is_contiguous(), 'fiode_b200 needs contiguous CUDA tensors'
    return C.c_void_p(t.data_ptr())


def stream_ptr():
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def launch_count() -> int:
    return int(lib().fiode_launch_count())


def reset_launch_count():
    lib().fiode_reset_launch_count()
