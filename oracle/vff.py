"""NumPy float64 restatement of the variational-Fourier-features model (oracle; test infrastructure only).

Follows the reference's in-repo implementation ``/root/reference/GPSat/vff.py`` (adapted there
from st--/VFF; TensorFlow ops restated 1:1 in NumPy) as used by ``GPflowVFFModel``
(``/root/reference/GPSat/models/vff_model.py:14-264``):

* ``make_Kuu``     ``vff.py:381-454``  per-dimension Kuu = blockdiag(diag + low rank) for Matern-1/2, 3/2, 5/2
* ``make_Kuf_np``  ``vff.py:518-523``  cos / sin features of the data (theta independent)
* ``make_Kuf``     ``vff.py:457-515``  features of prediction points with the edge corrections outside [a, b]
* ``make_kvs_np``  ``vff.py:553-559``  Kronecker-vector stack (Khatri-Rao product) over the dimensions
* ``GPR_kron.elbo`` ``vff.py:612-643``, ``GPR_kron.predict_f`` ``vff.py:645-675``

theta layout: ``[lengthscale_1..D, variance_1..D, likelihood_variance]`` (one 1-D kernel per input
dimension, each with its own variance: ``vff_model.py:163``; ``get_kernel_variance`` is their product,
``:217-218``).  Pinning: the reference has no VFF test (``tests/test_localexperts.py:253-274`` is
commented out) and TensorFlow is not installable here, so the ELBO / prediction are **parity
unpinned** beyond (i) the NumPy-only pieces ``make_Kuf_np`` / ``make_kvs_np`` executed from the
reference source (``tests/golden/vff_features.npz``) and (ii) convergence of the 1-D ELBO to the exact
GPR log marginal likelihood as the number of features grows.
"""
from __future__ import annotations

from functools import reduce

import numpy as np
import scipy.linalg as sla
import scipy.optimize as sopt

from . import gpr

KINDS = ("Matern12", "Matern32", "Matern52")


def n_features(ms) -> int:
    """cos features for every frequency, sin features for the non-zero ones (vff.py:594-598)."""
    return 2 * len(ms) - 1


def _kuu_parts(kind, ell, var, a, b, ms):
    """d (diagonal, cos block then sin block) and W (low-rank factor, block structured) with
    Kuu = diag(d) + W W^T; vff.py:381-454."""
    ms = np.asarray(ms, dtype=np.float64)
    om = 2.0 * np.pi * ms / (b - a)
    oms = om[om != 0]
    nc, ns = len(om), len(oms)
    if kind == "Matern12":
        lamb = 1.0 / ell
        two_or_four = np.where(om == 0, 2.0, 4.0)
        d_cos = (b - a) * (lamb ** 2 + om ** 2) / lamb / var / two_or_four
        d_sin = (b - a) * (lamb ** 2 + oms ** 2) / lamb / var / 4.0
        W = np.zeros((nc + ns, 1))
        W[:nc, 0] = 1.0 / np.sqrt(var)
    elif kind == "Matern32":
        lamb = np.sqrt(3.0) / ell
        four_or_eight = np.where(om == 0, 4.0, 8.0)
        d_cos = (b - a) * (lamb ** 2 + om ** 2) ** 2 / lamb ** 3 / var / four_or_eight
        d_sin = (b - a) * (lamb ** 2 + oms ** 2) ** 2 / lamb ** 3 / var / 8.0
        W = np.zeros((nc + ns, 2))
        W[:nc, 0] = 1.0 / np.sqrt(var)
        W[nc:, 1] = oms / lamb / np.sqrt(var)
    elif kind == "Matern52":
        lamb = np.sqrt(5.0) / ell
        s16 = np.where(om == 0, 16.0, 32.0)
        d_cos = 3 * (b - a) / s16 / lamb ** 5 / var * (lamb ** 2 + om ** 2) ** 3
        d_sin = 3 * (b - a) / 32.0 / lamb ** 5 / var * (lamb ** 2 + oms ** 2) ** 3
        W = np.zeros((nc + ns, 3))
        W[:nc, 0] = (3 * (om / lamb) ** 2 - 1) / np.sqrt(8 * var)
        W[:nc, 1] = 1.0 / np.sqrt(var)
        W[nc:, 2] = np.sqrt(3.0) * oms / lamb / np.sqrt(var)
    else:
        raise NotImplementedError(kind)
    return np.concatenate([d_cos, d_sin]), W


def kuu_dim(kind, ell, var, a, b, ms):
    """Dense per-dimension Kuu (``BlockDiagMat.get()``)."""
    d, W = _kuu_parts(kind, ell, var, a, b, ms)
    return np.diag(d) + W @ W.T


def kuf_np(X1, a, b, ms):
    """make_Kuf_np (vff.py:518-523): X1 [n, 1] -> [2 len(ms) - 1, n]."""
    ms = np.asarray(ms, dtype=np.float64)
    om = 2.0 * np.pi * ms / (b - a)
    cos = np.cos(om * (X1 - a)).T
    oms = om[om != 0]
    sin = np.sin(oms * (X1 - a)).T
    return np.vstack([cos, sin])


def kuf_pred(kind, ell, X1, a, b, ms):
    """make_Kuf (vff.py:457-515): as kuf_np inside [a, b], with the kernel's edge values outside."""
    ms = np.asarray(ms, dtype=np.float64)
    om = 2.0 * np.pi * ms / (b - a)
    oms = om[om != 0]
    x = X1[:, 0]
    cos = np.cos(om[:, None] * (x - a)[None, :])
    sin = np.sin(oms[:, None] * (x - a)[None, :])
    lt, gt = x < a, x > b
    if kind == "Matern12":
        sin[:, lt | gt] = 0.0
        cos[:, lt] = np.exp(-np.abs(x[lt] - a) / ell)[None, :]
        cos[:, gt] = np.exp(-np.abs(x[gt] - b) / ell)[None, :]
    elif kind == "Matern32":
        for mask, edge in ((lt, a), (gt, b)):
            arg = np.sqrt(3.0) * np.abs(x[mask] - edge) / ell
            cos[:, mask] = ((1 + arg) * np.exp(-arg))[None, :]
            sin[:, mask] = ((x[mask] - edge) * np.exp(-arg))[None, :] * oms[:, None]
    elif kind == "Matern52":
        assert not (lt.any() or gt.any()), "Edges not implemented for Matern52"
    else:
        raise NotImplementedError(kind)
    return np.vstack([cos, sin])


def kvs(mats):
    """make_kvs_np (vff.py:553-559): row index = i_1 * M_2 * ... + i_2 * ... (first dimension slowest)."""
    return reduce(lambda A, B: np.repeat(A, B.shape[0], axis=0) * np.tile(B, [A.shape[0], 1]), mats)


def split_theta(theta, D):
    theta = np.asarray(theta, dtype=np.float64)
    assert theta.shape == (2 * D + 1,)
    return theta[:D], theta[D:2 * D], float(theta[2 * D])


class Precomputed:
    """Static quantities of GPR_kron.__init__ (vff.py:600-610)."""

    def __init__(self, X, y, a, b, ms):
        X = np.asarray(X, dtype=np.float64)
        self.X, self.y = X, np.asarray(y, dtype=np.float64)
        self.a, self.b, self.ms = list(a), list(b), [np.asarray(m) for m in ms]
        D = X.shape[1]
        assert all(np.all(X[:, i] > self.a[i]) and np.all(X[:, i] < self.b[i]) for i in range(D))
        self.Kuf = kvs([kuf_np(X[:, i:i + 1], self.a[i], self.b[i], self.ms[i]) for i in range(D)])
        self.KufY = self.Kuf @ self.y
        self.KufKfu = self.Kuf @ self.Kuf.T
        self.trYY = float(self.y @ self.y)
        self.Ms = [n_features(m) for m in self.ms]


def _kron(mats):
    return reduce(np.kron, mats)


def elbo(kind, pre: Precomputed, theta):
    """GPR_kron.elbo (vff.py:612-643)."""
    D = pre.X.shape[1]
    ell, var, s2 = split_theta(theta, D)
    n = len(pre.y)
    Kuu = [kuu_dim(kind, ell[i], var[i], pre.a[i], pre.b[i], pre.ms[i]) for i in range(D)]
    Kuu_solid = _kron(Kuu)
    Kuu_inv_solid = _kron([np.linalg.inv(K) for K in Kuu])
    P = pre.KufKfu / s2 + Kuu_solid
    try:
        L = np.linalg.cholesky(P)
    except np.linalg.LinAlgError:
        return -np.inf
    log_det_P = 2.0 * np.sum(np.log(np.diag(L)))
    c = sla.solve_triangular(L, pre.KufY, lower=True) / s2
    M = int(np.prod(pre.Ms))
    Kuu_logdet = sum((M / pre.Ms[i]) * np.linalg.slogdet(Kuu[i])[1] for i in range(D))
    val = -0.5 * n * np.log(2 * np.pi * s2)
    val -= 0.5 * log_det_P
    val += 0.5 * Kuu_logdet
    val -= 0.5 * pre.trYY / s2
    val += 0.5 * float(c @ c)
    val -= 0.5 * n * float(np.prod(var)) / s2
    val += 0.5 * float(np.sum(Kuu_inv_solid * pre.KufKfu)) / s2
    return float(val)


def predict(kind, pre: Precomputed, theta, Xs):
    """GPR_kron.predict_f (vff.py:645-675), full_cov=False, plus y_var = f*_var + likelihood variance."""
    D = pre.X.shape[1]
    ell, var, s2 = split_theta(theta, D)
    Xs = np.asarray(Xs, dtype=np.float64)
    Kuu = [kuu_dim(kind, ell[i], var[i], pre.a[i], pre.b[i], pre.ms[i]) for i in range(D)]
    Kuu_inv_solid = _kron([np.linalg.inv(K) for K in Kuu])
    P = pre.KufKfu / s2 + _kron(Kuu)
    L = np.linalg.cholesky(P)
    c = sla.solve_triangular(L, pre.KufY, lower=True) / s2
    Kus = kvs([kuf_pred(kind, ell[i], Xs[:, i:i + 1], pre.a[i], pre.b[i], pre.ms[i]) for i in range(D)])
    tmp = sla.solve_triangular(L, Kus, lower=True)
    mean = tmp.T @ c
    fvar = np.prod(var) + np.sum(tmp * tmp, axis=0) - np.sum((Kuu_inv_solid @ Kus) * Kus, axis=0)
    return {"f*": mean, "f*_var": fvar, "y_var": fvar + s2}


def _kuu_parts_grad(kind, ell, var, a, b, ms):
    """(d, W) and their derivatives w.r.t. the lengthscale and the variance of one dimension."""
    ms = np.asarray(ms, dtype=np.float64)
    om = 2.0 * np.pi * ms / (b - a)
    oms = om[om != 0]
    nc = len(om)
    allom = np.concatenate([om, oms])
    d, W = _kuu_parts(kind, ell, var, a, b, ms)
    p, q, c0 = {"Matern12": (1, 1, 1.0), "Matern32": (2, 3, np.sqrt(3.0)), "Matern52": (3, 5, np.sqrt(5.0))}[kind]
    lamb = c0 / ell
    dlamb = -lamb / ell                                   # d lambda / d ell
    dd_l = d * (2 * p * lamb / (lamb ** 2 + allom ** 2) - q / lamb) * dlamb
    dd_v = -d / var
    dW_l = np.zeros_like(W)
    if kind == "Matern32":
        dW_l[nc:, 1] = -W[nc:, 1] / lamb * dlamb
    elif kind == "Matern52":
        dW_l[:nc, 0] = -6 * om ** 2 / lamb ** 3 / np.sqrt(8 * var) * dlamb
        dW_l[nc:, 2] = -W[nc:, 2] / lamb * dlamb
    dW_v = -0.5 * W / var
    return d, W, dd_l, dd_v, dW_l, dW_v


def _dim_matrices(kind, ell, var, a, b, ms):
    """K, dK/dell, dK/dvar, K^-1, dK^-1/dell, dK^-1/dvar, logdet K for one dimension."""
    d, W, dd_l, dd_v, dW_l, dW_v = _kuu_parts_grad(kind, ell, var, a, b, ms)
    K = np.diag(d) + W @ W.T
    dK_l = np.diag(dd_l) + dW_l @ W.T + W @ dW_l.T
    dK_v = np.diag(dd_v) + dW_v @ W.T + W @ dW_v.T
    Ki = np.linalg.inv(K)
    return K, dK_l, dK_v, Ki, -Ki @ dK_l @ Ki, -Ki @ dK_v @ Ki, np.linalg.slogdet(K)[1]


def elbo_and_grad(kind, pre: Precomputed, theta):
    """(elbo, d elbo / d theta).  With T = P^-1 + beta beta^T, beta = P^-1 Kuf y / s2, Phi = Kuf Kuf^T:

        d/d theta_{d,p} = -1/2 sum T * kron(.., dK_d, ..) + 1/2 (M / M_d) tr(K_d^-1 dK_d)
                          + 1/(2 s2) sum Phi * kron(.., dK_d^-1, ..) - [p = var] n prod(var) / (2 s2 var_d)
        d/d s2 = -n/(2 s2) + sum(T * Phi)/(2 s2^2) - c'c / s2 + [y'y + n prod(var) - tr(Kuu^-1 Phi)] / (2 s2^2)
    """
    D = pre.X.shape[1]
    ell, var, s2 = split_theta(theta, D)
    n = len(pre.y)
    val = elbo(kind, pre, theta)
    if not np.isfinite(val):
        return val, np.full(2 * D + 1, np.nan)
    dims = [_dim_matrices(kind, ell[i], var[i], pre.a[i], pre.b[i], pre.ms[i]) for i in range(D)]
    Ks = [m[0] for m in dims]
    Kis = [m[3] for m in dims]
    P = pre.KufKfu / s2 + _kron(Ks)
    L = np.linalg.cholesky(P)
    Pinv = sla.cho_solve((L, True), np.eye(len(P)))
    beta = Pinv @ pre.KufY / s2
    T = Pinv + np.outer(beta, beta)
    M = len(P)
    Phi = pre.KufKfu
    grad = np.empty(2 * D + 1)
    pv = float(np.prod(var))
    for i in range(D):
        for p, (dK, dKi) in enumerate(((dims[i][1], dims[i][4]), (dims[i][2], dims[i][5]))):
            dKuu = _kron([dK if e == i else Ks[e] for e in range(D)])
            dKuu_inv = _kron([dKi if e == i else Kis[e] for e in range(D)])
            g = -0.5 * np.sum(T * dKuu) + 0.5 * (M / pre.Ms[i]) * np.sum(Kis[i] * dK)
            g += 0.5 * np.sum(Phi * dKuu_inv) / s2
            if p == 1:
                g -= 0.5 * n * pv / var[i] / s2
            grad[i + p * D] = g
    c = sla.solve_triangular(L, pre.KufY, lower=True) / s2
    trKiPhi = float(np.sum(_kron(Kis) * Phi))
    grad[2 * D] = (-0.5 * n / s2 + 0.5 * np.sum(T * Phi) / s2 ** 2 - float(c @ c) / s2
                   + 0.5 * (pre.trYY + n * pv - trKiPhi) / s2 ** 2)
    return val, grad


class Transform(gpr.Transform):
    """gpflow defaults for the 2 D + 1 parameters: softplus, likelihood variance 1e-6 + softplus."""

    @classmethod
    def vff_default(cls, D):
        t = gpr.Transform.gpflow_default(2 * D - 1)          # 2 D - 1 + 2 = 2 D + 1 entries, last shifted
        return t


def fit(kind, pre: Precomputed, theta0, transform=None, trainable=None, max_iter=10_000):
    """gpflow.optimizers.Scipy().minimize(-elbo) (inherited optimise_parameters, gpflow_models.py:290-329).
    Returns (theta, elbo, success, nfev)."""
    D = pre.X.shape[1]
    P = 2 * D + 1
    theta0 = np.asarray(theta0, dtype=np.float64).copy()
    transform = Transform.vff_default(D) if transform is None else transform
    trainable = np.ones(P, bool) if trainable is None else np.asarray(trainable, bool)
    xi0 = transform.inverse(theta0)

    def fun(xi_free):
        xi = xi0.copy()
        xi[trainable] = xi_free
        theta = np.where(trainable, transform.forward(xi), theta0)
        val, g = elbo_and_grad(kind, pre, theta)
        if not np.isfinite(val):
            return np.inf, np.zeros(int(trainable.sum()))
        return -val, (-g * transform.dtheta_dxi(xi))[trainable]

    if not trainable.any():
        return theta0, elbo(kind, pre, theta0), True, 0
    res = sopt.minimize(fun, xi0[trainable], jac=True, method="L-BFGS-B", options=dict(maxiter=max_iter))
    xi = xi0.copy()
    xi[trainable] = res.x
    theta = np.where(trainable, transform.forward(xi), theta0)
    return theta, elbo(kind, pre, theta), bool(res.success), int(res.nfev)
