"""ctypes front-end of oracle/lapjv.c with the ``lap.lapjv`` call contract.

TEST INFRASTRUCTURE ONLY.  Mirrors the observable contract of lapx 0.5.11.post1
``lap.lapjv(cost, extend_cost=False, cost_limit=inf, return_cost=True) -> (opt, x, y)``
as used at boxmot/utils/matching.py:54-55, boxmot/utils/association.py:50-51,109 and
boxmot/trackers/boosttrack/assoc.py:113 (lapx source is absent from /root/reference:
parity unpinned for ties, see lapjv.c header).
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE, "liboracle_lap.so"])


def _lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "liboracle_lap.so")
        if not os.path.exists(path):
            build()
        lib = ctypes.CDLL(path)
        lib.oracle_lapjv.restype = ctypes.c_int
        lib.oracle_lapjv.argtypes = [
            ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_double,
            ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        lib.oracle_ddot_skx.restype = ctypes.c_double
        lib.oracle_ddot_skx.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_long]
        _LIB = lib
    return _LIB


def ddot(x, y):
    """x . y of float64 vectors in OpenBLAS' SkylakeX summation order (oracle/numerics.c)."""
    x = np.ascontiguousarray(x, dtype=np.float64)
    y = np.ascontiguousarray(y, dtype=np.float64)
    assert x.ndim == 1 and x.shape == y.shape
    return float(_lib().oracle_ddot_skx(x.ctypes.data, y.ctypes.data, x.size))


def norm(x):
    """np.linalg.norm of a float64 vector = sqrt(x.dot(x)) (numpy/linalg/_linalg.py)."""
    return float(np.sqrt(ddot(x, x)))


last_scans = 0


def lapjv(cost, extend_cost=False, cost_limit=np.inf, return_cost=True):
    global last_scans
    if cost_limit is None:
        # lapx types cost_limit as a C double: None raises TypeError there.
        raise TypeError("cost_limit must be a float")
    c = np.ascontiguousarray(cost, dtype=np.float64)
    if c.ndim != 2:
        raise ValueError("2-dimensional array expected")
    nr, nc = c.shape
    if nr != nc and not extend_cost and not cost_limit < np.inf:
        raise ValueError("Square cost array expected. If cost is intentionally "
                         "non-square, pass extend_cost=True.")
    x = np.empty(nr, dtype=np.int32)
    y = np.empty(nc, dtype=np.int32)
    scans = ctypes.c_long(0)
    lim = float(cost_limit) if cost_limit < np.inf else 1.7976931348623157e308
    rc = _lib().oracle_lapjv(c.ctypes.data, nr, nc, int(bool(extend_cost)), lim,
                             x.ctypes.data, y.ctypes.data, ctypes.byref(scans))
    if rc != 0:
        raise ValueError("lapjv failed")
    last_scans = scans.value
    opt = np.nan
    if return_cost:
        rows = np.nonzero(x >= 0)[0]
        opt = float(c[rows, x[rows]].sum()) if rows.size else 0.0
    x = x.astype(np.int64)
    y = y.astype(np.int64)
    return opt, x, y
