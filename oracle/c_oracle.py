"""ctypes access to oracle/c/geo_oracle.c (TEST INFRASTRUCTURE ONLY)."""
import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
_lib = None


def load():
    global _lib
    if _lib is None:
        so = HERE / "_build" / "libgeo_oracle.so"
        if not so.exists():
            subprocess.run(["make", "-C", str(HERE / "c")], check=True)
        _lib = C.CDLL(str(so))
        _lib.oracle_pairwise_sum_f32.restype = C.c_float
        _lib.oracle_kpp_draw.restype = C.c_int64
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def knn_bruteforce(z, kq):
    z = np.ascontiguousarray(z, np.float32)
    M, d = z.shape
    idx = np.empty((M, kq), np.int32)
    d2 = np.empty((M, kq), np.float64)
    load().oracle_knn_bruteforce(_p(z), C.c_int64(M), C.c_int(d), C.c_int(kq), _p(idx), _p(d2))
    return idx, d2


def dijkstra(W, source):
    W = W.tocsr()
    ip, ix, dt = W.indptr.astype(np.int64), W.indices.astype(np.int32), W.data.astype(np.float32)
    out = np.empty(W.shape[0], np.float64)
    rc = load().oracle_dijkstra(_p(ip), _p(ix), _p(dt), C.c_int64(W.shape[0]), C.c_int32(int(source)), _p(out))
    assert rc == 0
    return out


def pairwise_sum_f32(x):
    x = np.ascontiguousarray(x, np.float32)
    return np.float32(load().oracle_pairwise_sum_f32(_p(x), C.c_int64(x.size)))


def kpp_draw(dmin, is_center, u):
    dmin = np.ascontiguousarray(dmin, np.float32)
    c = np.ascontiguousarray(is_center, np.uint8)
    tot = C.c_float(0)
    i = load().oracle_kpp_draw(_p(dmin), _p(c), C.c_int64(dmin.size), C.c_double(u), C.byref(tot))
    return int(i), np.float32(tot.value)
