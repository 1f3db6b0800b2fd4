"""ctypes front for the C oracle (TEST INFRASTRUCTURE ONLY -- see b2dp_oracle.c).

May be imported only from tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  The product package never imports it.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liboracle.so")
_lib = None

_f32p = ctypes.POINTER(ctypes.c_float)


def build(force=False):
    src = os.path.join(_HERE, "b2dp_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", _HERE, "liboracle.so"])
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        L = ctypes.CDLL(_SO)
        L.orc_adam_alpha.restype = ctypes.c_float
        L.orc_adam_alpha.argtypes = [ctypes.c_float, ctypes.c_float, ctypes.c_float, ctypes.c_long]
        L.orc_sumsq_f64.restype = ctypes.c_double
        L.orc_sumsq_f64.argtypes = [_f32p, ctypes.c_size_t]
        L.orc_f32_to_bf16.restype = ctypes.c_uint16
        L.orc_f32_to_bf16.argtypes = [ctypes.c_float]
        L.orc_max_threads.restype = ctypes.c_int
        L.orc_addn_f32.argtypes = [ctypes.POINTER(_f32p), ctypes.c_int, ctypes.c_size_t, ctypes.c_float, _f32p]
        L.orc_addn_bf16wire.argtypes = L.orc_addn_f32.argtypes
        L.orc_cast_bf16.argtypes = [_f32p, ctypes.POINTER(ctypes.c_uint16), ctypes.c_size_t]
        L.orc_div_n_f32.argtypes = [_f32p, ctypes.c_size_t, ctypes.c_int]
        L.orc_scale_f32.argtypes = [_f32p, ctypes.c_size_t, ctypes.c_float]
        L.orc_adam_f32.argtypes = [_f32p, _f32p, _f32p, _f32p, ctypes.c_size_t, ctypes.c_float,
                                   ctypes.c_float, ctypes.c_float, ctypes.c_float, ctypes.c_long]
        L.orc_sgdm_f32.argtypes = [_f32p, _f32p, _f32p, ctypes.c_size_t, ctypes.c_float,
                                   ctypes.c_float, ctypes.c_int, ctypes.c_float]
        _lib = L
    return _lib


def _p(a):
    assert a.dtype == np.float32 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(_f32p)


def set_threads(t):
    lib().orc_set_threads(int(t))


def max_threads():
    return int(lib().orc_max_threads())


def cast_bf16(x):
    x = np.ascontiguousarray(x, dtype=np.float32)
    out = np.empty(x.shape, dtype=np.uint16)
    lib().orc_cast_bf16(_p(x), out.ctypes.data_as(ctypes.POINTER(ctypes.c_uint16)), x.size)
    return out


def addn(grads, prescale=1.0, bf16_wire=False, out=None):
    gs = [np.ascontiguousarray(g, dtype=np.float32) for g in grads]
    n, ln = len(gs), gs[0].size
    if out is None:
        out = np.empty(ln, dtype=np.float32)
    arr = (_f32p * n)(*[_p(g) for g in gs])
    fn = lib().orc_addn_bf16wire if bf16_wire else lib().orc_addn_f32
    fn(arr, n, ln, ctypes.c_float(prescale), _p(out))
    return out


def div_n(x, n):
    lib().orc_div_n_f32(_p(x), x.size, int(n))
    return x


def scale(x, s):
    lib().orc_scale_f32(_p(x), x.size, ctypes.c_float(s))
    return x


def adam_alpha(lr, beta1, beta2, step):
    return np.float32(lib().orc_adam_alpha(lr, beta1, beta2, int(step)))


def adam_(w, m, v, g, lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-7, step=1):
    """In place on w, m, v (float32 contiguous)."""
    lib().orc_adam_f32(_p(w), _p(m), _p(v), _p(g), w.size, lr, beta1, beta2, eps, int(step))


def sgdm_(w, a, g, lr=0.01, mu=0.9, nesterov=False, wd=0.0):
    lib().orc_sgdm_f32(_p(w), _p(a), _p(g), w.size, lr, mu, int(bool(nesterov)), wd)


def sumsq(g):
    g = np.ascontiguousarray(g, dtype=np.float32)
    return float(lib().orc_sumsq_f64(_p(g), g.size))
