"""NumPy/SciPy float64 restatement of the sparse GPR (Titsias 2009) path (oracle; test infrastructure only).

The reference builds ``gpflow.models.SGPR(data, kernel, inducing_variable=Z, ...)``
(``/root/reference/GPSat/models/gpflow_models.py:825-830``), optimises ``training_loss = -elbo``
(``:317-321`` inherited), reports ``+elbo`` as the objective (``:860-862``) and predicts with
``predict_f`` / ``predict_y`` (``:229-230`` inherited).  GPflow (``gpflow>=2.9``,
``gpflow/models/sgpr.py``) is not installable here; its published algorithm is:

    Kuu = k(Z, Z) + jitter I (jitter = 1e-6);  L = chol(Kuu);  A = L^-1 Kuf / sigma
    B = A A^T + I;  LB = chol(B);  c = LB^-1 A y / sigma
    elbo = -n/2 log 2pi - sum log diag LB - n/2 log sigma^2 - 1/2 y'y / sigma^2 + 1/2 c'c
           - 1/2 sum_i k_ii / sigma^2 + 1/2 tr(A A^T)
    predict_f: tmp1 = L^-1 Kus; tmp2 = LB^-1 tmp1; mean = tmp2^T c; var = k_ss + sum tmp2^2 - sum tmp1^2

(the same formulas appear in the reference's own VFF twin, ``GPSat/vff.py:612-675``).
Pinning: the reference's known-answer test ``tests/test_localexperts.py:229-251``
(``test_gpflow_sgpr``, M = N = 50, tolerance 1e-4 against the sklearn values) through
``tests/golden/kat_localexperts.npz``; with Z = X the bound is tight, so the exact-GPR oracle pins
the ELBO and the predictions to round-off.  M < n is parity-unpinned beyond these formulas.
theta layout: ``[lengthscale_1..D, kernel_variance, likelihood_variance]``.
"""
from __future__ import annotations

import numpy as np
import scipy.linalg as sla
import scipy.optimize as sopt

from . import gpr

JITTER = 1e-6   # gpflow.config.default_jitter()


def _factor(kind, X, y, Z, theta):
    n, D = X.shape
    ls, var, noise = gpr.split_theta(theta, D)
    sigma = np.sqrt(noise)
    Kuu = gpr.kernel_matrix(kind, Z, None, ls, var)
    Kuu[np.diag_indices(len(Z))] += JITTER
    Kuf = gpr.kernel_matrix(kind, Z, X, ls, var)
    L = np.linalg.cholesky(Kuu)
    A = sla.solve_triangular(L, Kuf, lower=True) / sigma
    B = A @ A.T + np.eye(len(Z))
    LB = np.linalg.cholesky(B)
    c = sla.solve_triangular(LB, A @ y, lower=True) / sigma
    return dict(L=L, A=A, LB=LB, c=c, Kuu=Kuu, Kuf=Kuf, ls=ls, var=var, noise=noise)


def elbo(kind, X, y, Z, theta):
    """gpflow.models.SGPR.elbo (zero mean function)."""
    kind = gpr.canonical_kernel(kind)
    n = len(X)
    try:
        f = _factor(kind, X, y, Z, theta)
    except np.linalg.LinAlgError:
        return -np.inf
    A, LB, c, var, noise = f["A"], f["LB"], f["c"], f["var"], f["noise"]
    bound = -0.5 * n * gpr.LOG2PI
    bound -= np.sum(np.log(np.diag(LB)))
    bound -= 0.5 * n * np.log(noise)
    bound -= 0.5 * (y @ y) / noise
    bound += 0.5 * (c @ c)
    bound -= 0.5 * n * var / noise
    bound += 0.5 * np.sum(A * A)
    return float(bound)


def elbo_and_grad(kind, X, y, Z, theta, with_z=False):
    """(elbo, d elbo / d theta), theta in constrained space, Z fixed (the reference freezes the
    inducing points unless ``train_inducing_points``, gpflow_models.py:896-897); ``with_z`` adds
    d elbo / d Z [M, D] (what ``set_trainable(inducing_variable.Z, True)`` makes TensorFlow differentiate):
        d elbo / d z_kd = sum_i G_C[k,i] dk(z_k, x_i)/dz_kd + 2 sum_j G_K[k,j] dk(z_k, z_j)/dz_kd
    (the jitter on the diagonal of Kuu does not depend on Z; dk/dz vanishes at z_j = z_k).

    With K = Kuu + jitter I, C = Kuf, Sigma = K + C C^T / s2, beta = Sigma^-1 C y / s2:
        G_K = 1/2 [K^-1 - Sigma^-1 - beta beta^T - K^-1 C C^T K^-1 / s2]
        G_C = [(K^-1 - Sigma^-1) C + beta (y - C^T beta)^T] / s2
        d elbo / d s2 = -n / (2 s2) + [ |y - C^T beta|^2 + n var - tr((K^-1 - Sigma^-1) C C^T) ] / (2 s2^2)
    and d elbo / d theta_p = sum G_K * dK/dtheta_p + sum G_C * dC/dtheta_p - n / (2 s2) * d var / d theta_p.
    """
    kind = gpr.canonical_kernel(kind)
    n, D = X.shape
    M = len(Z)
    try:
        f = _factor(kind, X, y, Z, theta)
    except np.linalg.LinAlgError:
        return -np.inf, np.full(D + 2, np.nan)
    L, A, LB, ls, var, s2 = f["L"], f["A"], f["LB"], f["ls"], f["var"], f["noise"]
    C = f["Kuf"]
    val = elbo(kind, X, y, Z, theta)
    Kinv = sla.cho_solve((L, True), np.eye(M))
    # Sigma = L B L^T
    LiT = sla.solve_triangular(L, np.eye(M), lower=True, trans="T")          # L^-T
    Binv = sla.cho_solve((LB, True), np.eye(M))
    Sinv = LiT @ Binv @ LiT.T
    beta = Sinv @ (C @ y) / s2
    Dm = Kinv - Sinv
    KC = Kinv @ C
    G_K = 0.5 * (Dm - np.outer(beta, beta) - KC @ KC.T / s2)
    resid = y - C.T @ beta
    G_C = (Dm @ C + np.outer(beta, resid)) / s2
    grad = np.empty(D + 2)
    r2_uu = gpr.scaled_sqdist(Z, None, ls)
    r2_uf = gpr.scaled_sqdist(Z, X, ls)
    g_uu = gpr.g_of_r2(kind, r2_uu, var)
    g_uf = gpr.g_of_r2(kind, r2_uf, var)
    GKg, GCg = G_K * g_uu, G_C * g_uf
    for d in range(D):
        dzz = Z[:, d][:, None] - Z[:, d][None, :]
        dzx = Z[:, d][:, None] - X[:, d][None, :]
        grad[d] = (np.sum(GKg * dzz * dzz) + np.sum(GCg * dzx * dzx)) / ls[d] ** 3
    Kuu_nojit = f["Kuu"] - JITTER * np.eye(M)
    grad[D] = (np.sum(G_K * Kuu_nojit) + np.sum(G_C * C)) / var - 0.5 * n / s2
    grad[D + 1] = -0.5 * n / s2 + 0.5 * (resid @ resid + n * var - np.sum(Dm * (C @ C.T))) / s2 ** 2
    if with_z:
        gz = np.empty((M, D))
        for d in range(D):
            dxz = X[:, d][None, :] - Z[:, d][:, None]          # x_i - z_k
            dzz = Z[:, d][None, :] - Z[:, d][:, None]          # z_j - z_k
            gz[:, d] = (np.sum(GCg * dxz, axis=1) + 2.0 * np.sum(GKg * dzz, axis=1)) / ls[d] ** 2
        return val, grad, gz
    return val, grad


def predict(kind, X, y, Z, theta, Xs, full_cov=False):
    """SGPR.predict_f and predict_y: ``f*``, ``f*_var``, ``y_var``; with ``full_cov`` also ``f*_cov`` =
    Kss + tmp2^T tmp2 - tmp1^T tmp1 and ``y_cov`` (gpflow/models/sgpr.py ``predict_f``)."""
    kind = gpr.canonical_kernel(kind)
    f = _factor(kind, X, y, Z, theta)
    Kus = gpr.kernel_matrix(kind, Z, Xs, f["ls"], f["var"])
    tmp1 = sla.solve_triangular(f["L"], Kus, lower=True)
    tmp2 = sla.solve_triangular(f["LB"], tmp1, lower=True)
    mean = tmp2.T @ f["c"]
    fvar = f["var"] + np.sum(tmp2 * tmp2, axis=0) - np.sum(tmp1 * tmp1, axis=0)
    out = {"f*": mean, "f*_var": fvar, "y_var": fvar + f["noise"]}
    if full_cov:
        cov = gpr.kernel_matrix(kind, Xs, None, f["ls"], f["var"]) + tmp2.T @ tmp2 - tmp1.T @ tmp1
        out["f*_cov"] = cov
        out["y_cov"] = cov + f["noise"] * np.eye(len(Xs))
    return out


def fit_z(kind, X, y, Z0, theta0, transform=None, trainable=None, max_iter=10_000):
    """gpflow.optimizers.Scipy().minimize(-elbo) over the free hyper-parameters AND the inducing points
    (``optimise_parameters(train_inducing_points=True)``, gpflow_models.py:864-899; Z is unconstrained).
    Returns (theta, Z, elbo, success, nfev)."""
    n, D = X.shape
    P = D + 2
    theta0 = np.asarray(theta0, dtype=np.float64).copy()
    Z0 = np.asarray(Z0, dtype=np.float64)
    transform = gpr.Transform.gpflow_default(D) if transform is None else transform
    trainable = np.ones(P, bool) if trainable is None else np.asarray(trainable, bool)
    xi0 = transform.inverse(theta0)
    nfree = int(trainable.sum())

    def unpack(v):
        xi = xi0.copy()
        xi[trainable] = v[:nfree]
        return xi, np.where(trainable, transform.forward(xi), theta0), v[nfree:].reshape(Z0.shape)

    def fun(v):
        xi, theta, Z = unpack(v)
        val, g, gz = elbo_and_grad(kind, X, y, Z, theta, with_z=True)
        if not np.isfinite(val):
            return np.inf, np.zeros(len(v))
        return -val, np.concatenate([(-g * transform.dtheta_dxi(xi))[trainable], -gz.ravel()])

    res = sopt.minimize(fun, np.concatenate([xi0[trainable], Z0.ravel()]), jac=True, method="L-BFGS-B",
                        options=dict(maxiter=max_iter))
    _, theta, Z = unpack(res.x)
    return theta, Z, elbo(kind, X, y, Z, theta), bool(res.success), int(res.nfev)


def fit(kind, X, y, Z, theta0, transform=None, trainable=None, max_iter=10_000):
    """gpflow.optimizers.Scipy().minimize(-elbo) with Z fixed.  Returns (theta, elbo, success, nfev)."""
    n, D = X.shape
    P = D + 2
    theta0 = np.asarray(theta0, dtype=np.float64).copy()
    transform = gpr.Transform.gpflow_default(D) if transform is None else transform
    trainable = np.ones(P, bool) if trainable is None else np.asarray(trainable, bool)
    xi0 = transform.inverse(theta0)

    def fun(xi_free):
        xi = xi0.copy()
        xi[trainable] = xi_free
        theta = np.where(trainable, transform.forward(xi), theta0)
        val, g = elbo_and_grad(kind, X, y, Z, theta)
        if not np.isfinite(val):
            return np.inf, np.zeros(int(trainable.sum()))
        return -val, (-g * transform.dtheta_dxi(xi))[trainable]

    if not trainable.any():
        return theta0, elbo(kind, X, y, Z, theta0), True, 0
    res = sopt.minimize(fun, xi0[trainable], jac=True, method="L-BFGS-B", options=dict(maxiter=max_iter))
    xi = xi0.copy()
    xi[trainable] = res.x
    theta = np.where(trainable, transform.forward(xi), theta0)
    return theta, elbo(kind, X, y, Z, theta), bool(res.success), int(res.nfev)
