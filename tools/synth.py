"""Synthetic directed graphs of the shapes BASELINE.json names, through
tools/libgraphgen.so.  Input generation for tests and bench only."""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_lib = None

CONFIGS = {
    # name: (kind, n, seed, params)
    "c2_chunglu_279": (1, 279, 3, [3.1908433806834573, 0.55]),
    "c3_er_2000": (0, 2000, 1, [0.05]),
    "c4_geo_31346": (2, 31346, 7, [0.4328, 0.08, 0.4, 0.4, 2.0]),
}


def _load():
    global _lib
    if _lib is None:
        L = C.CDLL(os.path.join(HERE, "libgraphgen.so"))
        L.gg_generate.restype = C.c_uint64
        L.gg_generate.argtypes = [C.c_int, C.c_uint32, C.c_uint64, C.c_void_p, C.c_int, C.POINTER(C.c_void_p)]
        L.gg_copy.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p]
        L.gg_free.argtypes = [C.c_void_p]
        _lib = L
    return _lib


def generate(kind, n, seed, params, nthreads=0):
    """-> uint32 (E,2) edge list, row-major order (the order a .flag file would list)."""
    L = _load()
    p = np.asarray(params, dtype=np.float64)
    out = C.c_void_p()
    ne = L.gg_generate(int(kind), int(n), int(seed), p.ctypes.data_as(C.c_void_p), int(nthreads or (os.cpu_count() or 1)), C.byref(out))
    arr = np.zeros((ne, 2), dtype=np.uint32)
    if ne:
        L.gg_copy(out, ne, arr.ctypes.data_as(C.c_void_p))
    L.gg_free(out)
    return arr


def config(name):
    kind, n, seed, params = CONFIGS[name]
    return n, generate(kind, n, seed, params)


def er(n, p, seed):
    return generate(0, n, seed, [p])
