"""ctypes front end of ``oracle/longdouble_gpr.c``: the exact-GPR path in x87 extended precision.

Test infrastructure only.  Used by ``oracle/make_golden.py`` / ``tools/precision_envelope.py`` to produce the
reference of record for the conditioning sweep (``tests/golden/precision_envelope.npz``): at condition numbers
of 1e8 .. 1e11 the float64 NumPy oracle itself is off by more than the 1e-8 parity bar, so the sweep needs a
reference that is not float64.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
KIND = {"Matern12": 0, "Matern32": 1, "Matern52": 2, "RBF": 3}


def load():
    global _LIB
    if _LIB is None:
        so = os.path.join(_HERE, "_build", "liboracle_ld.so")
        src = os.path.join(_HERE, "longdouble_gpr.c")
        if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
            subprocess.run(["make", "-C", _HERE], check=True, capture_output=True)
        _LIB = C.CDLL(so)
        _LIB.ld_gpr.restype = C.c_int
    return _LIB


def gpr(kind, X, y, theta, Xs=None, with_grad=True):
    """(lml, grad or None, f* or None, f*_var or None) in extended precision, rounded to float64 at the end."""
    lib = load()
    X = np.ascontiguousarray(X, dtype=np.float64)
    n, D = X.shape
    y = np.ascontiguousarray(y, dtype=np.float64)
    theta = np.ascontiguousarray(theta, dtype=np.float64)
    m = 0 if Xs is None else len(Xs)
    Xs_ = np.ascontiguousarray(Xs, dtype=np.float64) if m else np.zeros((1, D))
    lml = C.c_double(0.0)
    grad = np.zeros(D + 2) if with_grad else None
    fm, fv = (np.zeros(m), np.zeros(m)) if m else (None, None)
    p = lambda a: None if a is None else a.ctypes.data_as(C.c_void_p)
    rc = lib.ld_gpr(KIND[kind], n, D, m, p(X), p(y), p(Xs_), p(theta), C.byref(lml), p(grad), p(fm), p(fv))
    if rc == 1:
        return -np.inf, None, None, None
    if rc != 0:
        raise MemoryError("ld_gpr: allocation failed")
    return float(lml.value), grad, fm, fv


def sgpr(kind, X, y, Z, theta, Xs=None, jitter=1e-6):
    """(elbo, f* or None, f*_var or None) of the sparse model in extended precision (``ld_sgpr``)."""
    lib = load()
    lib.ld_sgpr.restype = C.c_int
    lib.ld_sgpr.argtypes = [C.c_int] * 5 + [C.c_void_p] * 5 + [C.c_double] + [C.c_void_p] * 3
    X = np.ascontiguousarray(X, dtype=np.float64)
    n, D = X.shape
    y = np.ascontiguousarray(y, dtype=np.float64)
    Z = np.ascontiguousarray(Z, dtype=np.float64)
    theta = np.ascontiguousarray(theta, dtype=np.float64)
    m = 0 if Xs is None else len(Xs)
    Xs_ = np.ascontiguousarray(Xs, dtype=np.float64) if m else np.zeros((1, D))
    elbo = C.c_double(0.0)
    fm, fv = (np.zeros(m), np.zeros(m)) if m else (None, None)
    p = lambda a: None if a is None else a.ctypes.data
    rc = lib.ld_sgpr(KIND[kind], n, D, len(Z), m, p(X), p(y), p(Z), p(Xs_), p(theta), float(jitter),
                     C.addressof(elbo), p(fm), p(fv))
    if rc == 1:
        return -np.inf, None, None
    if rc != 0:
        raise MemoryError("ld_sgpr: allocation failed")
    return float(elbo.value), fm, fv


def sgpr_grad_fd(kind, X, y, Z, theta, rel_step=1e-6, jitter=1e-6):
    """d elbo / d theta by central differences of the extended-precision ELBO.  With 64-bit mantissas the
    round-off term (eps / h) is about 1e-13 and the truncation term (h^2 times the third derivative) about 1e-12
    relative, so the result is a reference at the 1e-10 level - two orders below what the float64 paths are
    compared at."""
    theta = np.asarray(theta, dtype=np.float64)
    g = np.zeros(len(theta))
    for p in range(len(theta)):
        h = rel_step * theta[p]
        tp, tm = theta.copy(), theta.copy()
        tp[p] += h
        tm[p] -= h
        h2 = tp[p] - tm[p]                      # the step actually taken in float64
        g[p] = (sgpr(kind, X, y, Z, tp, jitter=jitter)[0] - sgpr(kind, X, y, Z, tm, jitter=jitter)[0]) / h2
    return g
