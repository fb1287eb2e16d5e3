"""Extended-precision (numpy.longdouble, x87 64-bit mantissa) restatement of ``oracle/vff.py`` - the checker of the
VFF conditioning study (``oracle/make_vff_envelope_golden.py`` -> ``tests/golden/vff_envelope.npz``).

Test infrastructure only.  Same formulas as ``GPR_kron.elbo`` / ``predict_f`` (``GPSat/vff.py:612-675``) and
``make_Kuu`` / ``make_Kuf`` (``:381-515``); every constant (pi, sqrt 3, sqrt 5) and every intermediate is long double,
the Cholesky factorisation and the triangular solves are written out because LAPACK has no long double.
"""
from __future__ import annotations

from functools import reduce

import numpy as np

LD = np.longdouble
PI = LD(4) * np.arctan(LD(1))


def _ld(x):
    return np.asarray(x, dtype=LD)


def chol(A):
    """Lower Cholesky factor of a symmetric positive definite long-double matrix (column by column)."""
    A = _ld(A).copy()
    n = len(A)
    L = np.zeros_like(A)
    for j in range(n):
        s = A[j, j] - np.dot(L[j, :j], L[j, :j])
        if not s > 0:
            raise np.linalg.LinAlgError("not positive definite")
        L[j, j] = np.sqrt(s)
        if j + 1 < n:
            L[j + 1:, j] = (A[j + 1:, j] - L[j + 1:, :j] @ L[j, :j]) / L[j, j]
    return L


def solve_lower(L, B):
    """L^-1 B for lower triangular L (forward substitution, all right-hand sides at once)."""
    B = _ld(B)
    X = np.zeros_like(B)
    for i in range(len(L)):
        X[i] = (B[i] - L[i, :i] @ X[:i]) / L[i, i]
    return X


def inv_spd(K):
    L = chol(K)
    Li = solve_lower(L, np.eye(len(K), dtype=LD))
    return Li.T @ Li, LD(2) * np.sum(np.log(np.diag(L)))


def kuu_dim(kind, ell, var, a, b, ms):
    ell, var, a, b = LD(ell), LD(var), LD(a), LD(b)
    ms = _ld(ms)
    om = LD(2) * PI * ms / (b - a)
    oms = om[om != 0]
    nc, ns = len(om), len(oms)
    if kind == "Matern12":
        lamb = LD(1) / ell
        f = np.where(om == 0, LD(2), LD(4))
        d_cos = (b - a) * (lamb ** 2 + om ** 2) / lamb / var / f
        d_sin = (b - a) * (lamb ** 2 + oms ** 2) / lamb / var / LD(4)
        W = np.zeros((nc + ns, 1), dtype=LD)
        W[:nc, 0] = LD(1) / np.sqrt(var)
    elif kind == "Matern32":
        lamb = np.sqrt(LD(3)) / ell
        f = np.where(om == 0, LD(4), LD(8))
        d_cos = (b - a) * (lamb ** 2 + om ** 2) ** 2 / lamb ** 3 / var / f
        d_sin = (b - a) * (lamb ** 2 + oms ** 2) ** 2 / lamb ** 3 / var / LD(8)
        W = np.zeros((nc + ns, 2), dtype=LD)
        W[:nc, 0] = LD(1) / np.sqrt(var)
        W[nc:, 1] = oms / lamb / np.sqrt(var)
    elif kind == "Matern52":
        lamb = np.sqrt(LD(5)) / ell
        f = np.where(om == 0, LD(16), LD(32))
        d_cos = LD(3) * (b - a) / f / lamb ** 5 / var * (lamb ** 2 + om ** 2) ** 3
        d_sin = LD(3) * (b - a) / LD(32) / lamb ** 5 / var * (lamb ** 2 + oms ** 2) ** 3
        W = np.zeros((nc + ns, 3), dtype=LD)
        W[:nc, 0] = (LD(3) * (om / lamb) ** 2 - LD(1)) / np.sqrt(LD(8) * var)
        W[:nc, 1] = LD(1) / np.sqrt(var)
        W[nc:, 2] = np.sqrt(LD(3)) * oms / lamb / np.sqrt(var)
    else:
        raise NotImplementedError(kind)
    return np.diag(np.concatenate([d_cos, d_sin])) + W @ W.T


def kuf(kind, ell, x, a, b, ms, edges):
    """[2 len(ms) - 1, n] features of one dimension; ``edges``: the kernel's edge values outside [a, b]
    (prediction points, make_Kuf) or none (training points, make_Kuf_np)."""
    ell, a, b = LD(ell), LD(a), LD(b)
    x = _ld(x)
    om = LD(2) * PI * _ld(ms) / (b - a)
    oms = om[om != 0]
    cos = np.cos(om[:, None] * (x - a)[None, :])
    sin = np.sin(oms[:, None] * (x - a)[None, :])
    if edges:
        lt, gt = x < a, x > b
        if kind == "Matern12":
            sin[:, lt | gt] = 0
            cos[:, lt] = np.exp(-np.abs(x[lt] - a) / ell)[None, :]
            cos[:, gt] = np.exp(-np.abs(x[gt] - b) / ell)[None, :]
        elif kind == "Matern32":
            for mask, edge in ((lt, a), (gt, b)):
                arg = np.sqrt(LD(3)) * np.abs(x[mask] - edge) / ell
                cos[:, mask] = ((LD(1) + arg) * np.exp(-arg))[None, :]
                sin[:, mask] = ((x[mask] - edge) * np.exp(-arg))[None, :] * oms[:, None]
        else:
            assert not (lt.any() or gt.any()), "Edges not implemented for Matern52"
    return np.vstack([cos, sin])


def kvs(mats):
    return reduce(lambda A, B: np.repeat(A, B.shape[0], axis=0) * np.tile(B, [A.shape[0], 1]), mats)


class Precomputed:
    def __init__(self, kind, X, y, a, b, ms):
        self.kind = kind
        self.X, self.y = _ld(X), _ld(y)
        self.a, self.b, self.ms = list(a), list(b), [np.asarray(m) for m in ms]
        D = self.X.shape[1]
        self.Kuf = kvs([kuf(kind, 1.0, self.X[:, i], self.a[i], self.b[i], self.ms[i], False) for i in range(D)])
        self.KufY = self.Kuf @ self.y
        self.KufKfu = self.Kuf @ self.Kuf.T
        self.trYY = self.y @ self.y
        self.Ms = [2 * len(m) - 1 for m in self.ms]


def _factor(pre, theta):
    D = pre.X.shape[1]
    theta = _ld(theta)
    ell, var, s2 = theta[:D], theta[D:2 * D], theta[2 * D]
    Kuu = [kuu_dim(pre.kind, ell[i], var[i], pre.a[i], pre.b[i], pre.ms[i]) for i in range(D)]
    inv_ld = [inv_spd(K) for K in Kuu]
    P = pre.KufKfu / s2 + reduce(np.kron, Kuu)
    L = chol(P)
    c = solve_lower(L, pre.KufY[:, None])[:, 0] / s2
    return ell, var, s2, Kuu, inv_ld, L, c


def elbo(pre, theta) -> float:
    ell, var, s2, Kuu, inv_ld, L, c = _factor(pre, theta)
    n = LD(len(pre.y))
    M = int(np.prod(pre.Ms))
    kuu_logdet = sum(LD(M // pre.Ms[i]) * inv_ld[i][1] for i in range(len(Kuu)))
    kinv = reduce(np.kron, [k[0] for k in inv_ld])
    val = -LD(0.5) * n * np.log(LD(2) * PI * s2)
    val -= np.sum(np.log(np.diag(L)))
    val += LD(0.5) * kuu_logdet
    val -= LD(0.5) * pre.trYY / s2
    val += LD(0.5) * (c @ c)
    val -= LD(0.5) * n * np.prod(var) / s2
    val += LD(0.5) * np.sum(kinv * pre.KufKfu) / s2
    return float(val)


def predict(pre, theta, Xs):
    ell, var, s2, Kuu, inv_ld, L, c = _factor(pre, theta)
    Xs = _ld(Xs)
    D = pre.X.shape[1]
    Kus = kvs([kuf(pre.kind, ell[i], Xs[:, i], pre.a[i], pre.b[i], pre.ms[i], True) for i in range(D)])
    tmp = solve_lower(L, Kus)
    mean = tmp.T @ c
    kinv = reduce(np.kron, [k[0] for k in inv_ld])
    fvar = np.prod(var) + np.sum(tmp * tmp, axis=0) - np.sum((kinv @ Kus) * Kus, axis=0)
    return mean.astype(np.float64), fvar.astype(np.float64)


def grad_fd(pre, theta, rel_step=1e-6):
    """d elbo / d theta by central differences of the extended-precision ELBO (good to ~1e-10, see
    ``oracle/longdouble.py::sgpr_grad_fd``)."""
    theta = np.asarray(theta, dtype=np.float64)
    g = np.zeros(len(theta))
    for p in range(len(theta)):
        h = rel_step * theta[p]
        tp, tm = theta.copy(), theta.copy()
        tp[p] += h
        tm[p] -= h
        g[p] = (elbo(pre, tp) - elbo(pre, tm)) / (tp[p] - tm[p])
    return g
