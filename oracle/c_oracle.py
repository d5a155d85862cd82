"""ctypes loader for the C oracle (oracle/lk_oracle.c).  TEST INFRASTRUCTURE ONLY:
imported by tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs, never
by the product package."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liblk_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "lk_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B"])
    return _SO


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        L = C.CDLL(_SO)
        dp = C.POINTER(C.c_double)
        L.lko_get_max_layer.argtypes = [C.c_int, C.c_int]
        L.lko_lk_single.argtypes = [dp, dp, C.c_int, C.c_int, C.c_int, C.c_double, dp, dp, C.c_void_p]
        L.lko_ctf.argtypes = [dp, dp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double, dp, dp, C.c_void_p]
        L.lko_pyr_down.argtypes = [dp, C.c_int, C.c_int, dp]
        L.lko_pyr_down.restype = None
        L.lko_pyr_up.argtypes = [dp, C.c_int, C.c_int, dp]
        L.lko_pyr_up.restype = None
        L.lko_warp.argtypes = [dp, C.c_int, C.c_int, dp, dp, C.c_float, dp]
        L.lko_warp.restype = None
        L.lko_gesture_stats.argtypes = [dp, dp, C.c_int, C.c_int, C.c_double, dp, dp, C.POINTER(C.c_longlong)]
        L.lko_gesture_stats.restype = None
        _lib = L
    return _lib


def set_threads(n: int) -> int:
    """OpenMP team size of the port (the reference hard-codes 8, ParallelVideoPyr.cpp:164; the CPU
    baseline uses every host thread).  torchrun exports OMP_NUM_THREADS=1, so the bench sets it
    explicitly.  Returns the value in effect."""
    lib()
    g = C.CDLL("libgomp.so.1")
    g.omp_set_num_threads(int(n))
    g.omp_get_max_threads.restype = C.c_int
    return int(g.omp_get_max_threads())


def _p(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def get_max_layer(w, h):
    return lib().lko_get_max_layer(w, h)


def _check(rc):
    if rc == 2:
        raise ValueError("size not divisible by 2^(n_levels-1)")
    if rc:
        raise ValueError("C oracle error %d" % rc)


def lk_single(a, b, r=1, det_eps=0.0, want_mask=False):
    a, b = _f64(a), _f64(b)
    h, w = a.shape
    u, v = np.empty_like(a), np.empty_like(a)
    m = np.empty((h, w), np.uint8)
    _check(lib().lko_lk_single(_p(a), _p(b), w, h, r, det_eps, _p(u), _p(v), m.ctypes.data))
    return (u, v, m.astype(bool)) if want_mask else (u, v)


def coarse_to_fine(a, b, n_levels, r=1, det_eps=0.0, want_mask=False):
    a, b = _f64(a), _f64(b)
    if a.shape != b.shape:
        raise ValueError("frame sizes differ")
    h, w = a.shape
    u, v = np.empty_like(a), np.empty_like(a)
    m = np.empty((h, w), np.uint8)
    _check(lib().lko_ctf(_p(a), _p(b), w, h, n_levels, r, det_eps, _p(u), _p(v), m.ctypes.data))
    return (u, v, m.astype(bool)) if want_mask else (u, v)


def pyr_down(a):
    a = _f64(a)
    h, w = a.shape
    o = np.empty(((h + 1) // 2, (w + 1) // 2))
    lib().lko_pyr_down(_p(a), w, h, _p(o))
    return o


def pyr_up(a):
    a = _f64(a)
    h, w = a.shape
    o = np.empty((2 * h, 2 * w))
    lib().lko_pyr_up(_p(a), w, h, _p(o))
    return o


def warp(src, upu, upv, sign):
    src, upu, upv = _f64(src), _f64(upu), _f64(upv)
    h, w = src.shape
    o = np.empty_like(src)
    lib().lko_warp(_p(src), w, h, _p(upu), _p(upv), float(sign), _p(o))
    return o


def gesture_stats(u, v, mag_thresh=20.0, count_thresh=500):
    u, v = _f64(u), _f64(v)
    h, w = u.shape
    sx, sy, c = C.c_double(), C.c_double(), C.c_longlong()
    lib().lko_gesture_stats(_p(u), _p(v), w, h, mag_thresh, C.byref(sx), C.byref(sy), C.byref(c))
    return {"sum_x": sx.value, "sum_y": sy.value, "count": c.value, "detected": c.value > count_thresh}
