"""NumPy/SciPy float64 restatement of the exact-GPR hot path (oracle; test infrastructure only).

Restates, function by function, what the reference calls into GPflow for
(``/root/reference/GPSat/models/gpflow_models.py``):

* ``kernel_matrix``      - ``gpflow.kernels.{Matern12,Matern32,Matern52,SquaredExponential}``
                           with ARD lengthscales (constructed at ``gpflow_models.py:116-135``)
* ``lml``                - ``GPR.log_marginal_likelihood`` (called at ``gpflow_models.py:337``)
* ``lml_and_grad``       - what TF autodiff returns for ``training_loss`` (``:317-321``), here
                           analytically: d lml/d theta_p = 1/2 tr((aa^T - Ks^-1) dKs/d theta_p)
* ``predict``            - ``GPR.predict_f`` / ``predict_y`` (``:229-230``) via
                           ``gpflow.conditionals.base_conditional``
* ``Transform``          - ``gpflow.utilities.positive`` (softplus, shifted softplus) and
                           ``tfp.bijectors.Sigmoid(low, high)`` (``:490-494``); NumPy twins of the
                           same functions live in the reference at ``GPSat/utils.py:2320-2400``
* ``fit``                - ``gpflow.optimizers.Scipy().minimize`` (``:317-321``), i.e. SciPy
                           L-BFGS-B with jac=True on the unconstrained vector, SciPy defaults.

theta layout everywhere: ``[lengthscale_1 .. lengthscale_D, kernel_variance, likelihood_variance]``.
"""
from __future__ import annotations

import numpy as np
import scipy.linalg as sla
import scipy.optimize as sopt

KERNELS = ("Matern12", "Matern32", "Matern52", "RBF")
KERNEL_ALIASES = {"SquaredExponential": "RBF", "Exponential": "Matern12"}
LOG2PI = float(np.log(2.0 * np.pi))
SQRT3 = float(np.sqrt(3.0))
SQRT5 = float(np.sqrt(5.0))
# gpflow.likelihoods.Gaussian default variance_lower_bound (gpflow/likelihoods/scalar_continuous.py)
VARIANCE_LOWER_BOUND = 1e-6


def canonical_kernel(name: str) -> str:
    name = KERNEL_ALIASES.get(name, name)
    if name not in KERNELS:
        raise NotImplementedError(f"kernel '{name}' not in {KERNELS}")
    return name


# ----------------------------------------------------------------------------------------------
# kernels
# ----------------------------------------------------------------------------------------------

def scaled_sqdist(X, X2, ls, mode="direct"):
    """r^2_ij = sum_d ((x_id - x'_jd)/l_d)^2.

    mode="gpflow": the exact expression GPflow evaluates (``gpflow/utilities/ops.py``
    ``square_distance``): ``|a|^2 + |b|^2 - 2 a.b`` on the scaled inputs (not clipped here; the
    Matern kernels clip at 1e-36 before the sqrt).  mode="direct": sum of squared differences,
    which has no cancellation error; the two agree to ~1e-16*|x/l|^2 absolute.
    """
    A = X / ls
    B = A if X2 is None else X2 / ls
    if mode == "gpflow":
        As = np.sum(A * A, axis=-1)
        Bs = np.sum(B * B, axis=-1)
        return -2.0 * (A @ B.T) + As[:, None] + Bs[None, :]
    diff = A[:, None, :] - B[None, :, :]
    return np.sum(diff * diff, axis=-1)


def k_of_r2(kind, r2, var):
    """Stationary kernel as a function of r^2 (gpflow/kernels/stationaries.py)."""
    if kind == "RBF":
        return var * np.exp(-0.5 * r2)
    r = np.sqrt(np.maximum(r2, 1e-36))
    if kind == "Matern12":
        return var * np.exp(-r)
    if kind == "Matern32":
        return var * (1.0 + SQRT3 * r) * np.exp(-SQRT3 * r)
    if kind == "Matern52":
        return var * (1.0 + SQRT5 * r + 5.0 / 3.0 * r * r) * np.exp(-SQRT5 * r)
    raise NotImplementedError(kind)


def g_of_r2(kind, r2, var):
    """g(r) such that dk/dl_d = g(r) * (x_d-x'_d)^2 / l_d^3  (zero where r == 0)."""
    if kind == "RBF":
        return var * np.exp(-0.5 * r2)
    r = np.sqrt(np.maximum(r2, 1e-36))
    if kind == "Matern12":
        # autodiff through sqrt(max(r2, 1e-36)): derivative is 0 on the clamped branch
        return np.where(r2 > 1e-36, var * np.exp(-r) / r, 0.0)
    if kind == "Matern32":
        return 3.0 * var * np.exp(-SQRT3 * r)
    if kind == "Matern52":
        return 5.0 / 3.0 * var * (1.0 + SQRT5 * r) * np.exp(-SQRT5 * r)
    raise NotImplementedError(kind)


def kernel_matrix(kind, X, X2, ls, var, mode="direct"):
    kind = canonical_kernel(kind)
    return k_of_r2(kind, scaled_sqdist(X, X2, np.asarray(ls, float), mode), var)


# ----------------------------------------------------------------------------------------------
# log marginal likelihood, gradient, prediction
# ----------------------------------------------------------------------------------------------

def split_theta(theta, D):
    theta = np.asarray(theta, dtype=np.float64)
    assert theta.shape == (D + 2,), f"theta must have D+2={D + 2} entries"
    return theta[:D], float(theta[D]), float(theta[D + 1])


def lml(kind, X, y, theta, mode="direct"):
    """GPR.log_marginal_likelihood with a zero mean function.

    Ks = K + s2 I; L = chol(Ks); lml = -1/2 |L^-1 y|^2 - sum log L_ii - n/2 log 2pi
    (gpflow/models/gpr.py + gpflow/logdensities.py::multivariate_normal).
    Returns -inf if Ks is not positive definite.
    """
    n, D = X.shape
    ls, var, noise = split_theta(theta, D)
    Ks = kernel_matrix(kind, X, None, ls, var, mode)
    Ks[np.diag_indices(n)] += noise
    try:
        L = np.linalg.cholesky(Ks)
    except np.linalg.LinAlgError:
        return -np.inf
    a = sla.solve_triangular(L, y, lower=True)
    return float(-0.5 * a @ a - np.sum(np.log(np.diag(L))) - 0.5 * n * LOG2PI)


def lml_and_grad(kind, X, y, theta, mode="direct"):
    """(lml, d lml / d theta) with theta in constrained space, analytic gradient."""
    kind = canonical_kernel(kind)
    n, D = X.shape
    ls, var, noise = split_theta(theta, D)
    r2 = scaled_sqdist(X, None, ls, mode)
    K = k_of_r2(kind, r2, var)
    Ks = K.copy()
    Ks[np.diag_indices(n)] += noise
    try:
        L = np.linalg.cholesky(Ks)
    except np.linalg.LinAlgError:
        return -np.inf, np.full(D + 2, np.nan)
    a = sla.solve_triangular(L, y, lower=True)
    val = float(-0.5 * a @ a - np.sum(np.log(np.diag(L))) - 0.5 * n * LOG2PI)
    alpha = sla.solve_triangular(L, a, lower=True, trans="T")
    Kinv = sla.cho_solve((L, True), np.eye(n))
    Q = np.outer(alpha, alpha) - Kinv
    g = g_of_r2(kind, r2, var)
    grad = np.empty(D + 2)
    Qg = Q * g
    for d in range(D):
        dx = X[:, d][:, None] - X[:, d][None, :]
        grad[d] = 0.5 * np.sum(Qg * dx * dx) / ls[d] ** 3
    grad[D] = 0.5 * np.sum(Q * K) / var
    grad[D + 1] = 0.5 * np.trace(Q)
    return val, grad


def predict(kind, X, y, theta, Xs, mode="direct", full_cov=False):
    """GPR.predict_f + predict_y (gpflow/models/gpr.py, conditionals/util.py::base_conditional).

    A = L^-1 Kxs; f*_var = k** - colsum(A^2); f* = (L^-T A)^T y; y_var = f*_var + s2.
    """
    kind = canonical_kernel(kind)
    n, D = X.shape
    ls, var, noise = split_theta(theta, D)
    Ks = kernel_matrix(kind, X, None, ls, var, mode)
    Ks[np.diag_indices(n)] += noise
    L = np.linalg.cholesky(Ks)
    Kxs = kernel_matrix(kind, X, Xs, ls, var, mode)
    A = sla.solve_triangular(L, Kxs, lower=True)
    out = {}
    if full_cov:
        Kss = kernel_matrix(kind, Xs, None, ls, var, mode)
        cov = Kss - A.T @ A
        out["f*_cov"] = cov
        fvar = np.diag(cov).copy()
    else:
        fvar = var - np.sum(A * A, axis=0)
    A2 = sla.solve_triangular(L, A, lower=True, trans="T")
    out["f*"] = A2.T @ y
    out["f*_var"] = fvar
    out["y_var"] = fvar + noise
    if full_cov:
        ycov = out["f*_cov"].copy()
        ycov[np.diag_indices(len(Xs))] += noise
        out["y_cov"] = ycov
    return out


# ----------------------------------------------------------------------------------------------
# parameter transforms (unconstrained xi <-> constrained theta)
# ----------------------------------------------------------------------------------------------

TRANSFORM_SOFTPLUS = 0   # theta = shift + softplus(xi)      (gpflow.utilities.positive(lower=shift))
TRANSFORM_SIGMOID = 1    # theta = low + (high-low)*sigmoid(xi)   (tfp.bijectors.Sigmoid)


def softplus(x):
    x = np.asarray(x, float)
    return np.where(x > 0, x + np.log1p(np.exp(-np.abs(x))), np.log1p(np.exp(-np.abs(x))))


def inv_softplus(y):
    y = np.asarray(y, float)
    # log(exp(y) - 1) = y + log(1 - exp(-y))
    return y + np.log(-np.expm1(-y))


def sigmoid(x):
    x = np.asarray(x, float)
    e = np.exp(-np.abs(x))
    return np.where(x >= 0, 1.0 / (1.0 + e), e / (1.0 + e))


class Transform:
    """Per-component bijection xi -> theta.  ``kind[p]`` in {SOFTPLUS, SIGMOID}; ``a[p]`` is the
    shift (softplus) or ``low`` (sigmoid); ``b[p]`` is unused (softplus) or ``high`` (sigmoid)."""

    def __init__(self, kind, a, b):
        self.kind = np.asarray(kind, dtype=np.int32)
        self.a = np.asarray(a, dtype=np.float64)
        self.b = np.asarray(b, dtype=np.float64)

    @classmethod
    def gpflow_default(cls, D):
        """softplus for lengthscales / kernel variance, 1e-6 + softplus for the likelihood
        variance (gpflow.likelihoods.Gaussian variance_lower_bound; the reference refers to it
        at gpflow_models.py:404-409)."""
        P = D + 2
        a = np.zeros(P)
        a[P - 1] = VARIANCE_LOWER_BOUND
        return cls(np.zeros(P, np.int32), a, np.zeros(P))

    def copy(self):
        return Transform(self.kind.copy(), self.a.copy(), self.b.copy())

    def set_sigmoid(self, idx, low, high):
        self.kind[idx] = TRANSFORM_SIGMOID
        self.a[idx] = low
        self.b[idx] = high

    def forward(self, xi):
        xi = np.asarray(xi, float)
        sp = self.a + softplus(xi)
        s = sigmoid(xi)
        sg = self.b * s + self.a * (1.0 - s)
        return np.where(self.kind == TRANSFORM_SIGMOID, sg, sp)

    def inverse(self, theta):
        theta = np.asarray(theta, float)
        with np.errstate(all="ignore"):
            sp = inv_softplus(theta - self.a)
            sg = np.log(theta - self.a) - np.log(self.b - theta)
        return np.where(self.kind == TRANSFORM_SIGMOID, sg, sp)

    def dtheta_dxi(self, xi):
        xi = np.asarray(xi, float)
        s = sigmoid(xi)
        return np.where(self.kind == TRANSFORM_SIGMOID, (self.b - self.a) * s * (1.0 - s), s)


# ----------------------------------------------------------------------------------------------
# priors (gpflow.Parameter.log_prior_density; the reference only ever passes prior=None through,
# gpflow_models.py:617,624, so this restates GPflow's published convention and is parity-unpinned)
# ----------------------------------------------------------------------------------------------

PRIOR_NONE, PRIOR_NORMAL, PRIOR_LOGNORMAL, PRIOR_GAMMA = 0, 1, 2, 3


def log_prior_xi(pkind, pa, pb, transform, xi, trainable):
    """sum over trainable parameters with a prior of  log p(theta) + log|d theta/d xi|
    (prior on the constrained value, GPflow's default ``prior_on``), and its gradient in xi."""
    import scipy.special as sps
    xi = np.asarray(xi, float)
    theta = transform.forward(xi)
    total, grad = 0.0, np.zeros_like(xi)
    for p in range(len(xi)):
        if not trainable[p] or pkind[p] == PRIOR_NONE:
            continue
        th, a, b = theta[p], pa[p], pb[p]
        if pkind[p] == PRIOR_NORMAL:
            lp = -0.5 * ((th - a) / b) ** 2 - np.log(b) - 0.5 * LOG2PI
            dlp = -(th - a) / b ** 2
        elif pkind[p] == PRIOR_LOGNORMAL:
            lp = -np.log(th) - np.log(b) - 0.5 * LOG2PI - 0.5 * ((np.log(th) - a) / b) ** 2
            dlp = -1.0 / th - (np.log(th) - a) / (b ** 2 * th)
        else:
            lp = a * np.log(b) - sps.gammaln(a) + (a - 1.0) * np.log(th) - b * th
            dlp = (a - 1.0) / th - b
        s = float(sigmoid(xi[p]))
        if transform.kind[p] == TRANSFORM_SIGMOID:
            ldj = np.log(transform.b[p] - transform.a[p]) + np.log(s) + np.log1p(-s)
            dldj = 1.0 - 2.0 * s
        else:
            ldj = -float(softplus(-xi[p]))
            dldj = 1.0 - s
        total += lp + ldj
        grad[p] = dlp * transform.dtheta_dxi(xi)[p] + dldj
    return total, grad


# ----------------------------------------------------------------------------------------------
# fit
# ----------------------------------------------------------------------------------------------

def objective_and_grad_xi(kind, X, y, xi_free, theta0, trainable, transform, mode="direct", priors=None):
    """-(lml + log prior) and its gradient w.r.t. the free unconstrained variables
    (training_loss + autodiff)."""
    xi_full = transform.inverse(theta0)
    xi_full[trainable] = xi_free
    theta = np.where(trainable, transform.forward(xi_full), theta0)
    val, g_theta = lml_and_grad(kind, X, y, theta, mode)
    if not np.isfinite(val):
        return np.inf, np.zeros(int(np.sum(trainable)))
    g_xi = -g_theta * transform.dtheta_dxi(xi_full)
    if priors is not None:
        lp, glp = log_prior_xi(priors[0], priors[1], priors[2], transform, xi_full, trainable)
        val += lp
        g_xi = g_xi - glp
    return -val, g_xi[trainable]


def fit(kind, X, y, theta0, transform=None, trainable=None, max_iter=10_000, mode="direct",
        return_result=False, priors=None):
    """gpflow.optimizers.Scipy().minimize(training_loss, trainable_variables,
    options=dict(maxiter=max_iter))  (reference call: gpflow_models.py:317-321).

    SciPy defaults that call inherits: method L-BFGS-B, maxcor 10, ftol 2.22e-9, gtol 1e-5,
    maxfun 15000, maxls 20.  Returns (theta_opt, lml_opt, success, n_evals).
    """
    n, D = X.shape
    P = D + 2
    theta0 = np.asarray(theta0, dtype=np.float64).copy()
    transform = Transform.gpflow_default(D) if transform is None else transform
    trainable = np.ones(P, bool) if trainable is None else np.asarray(trainable, bool)
    xi0 = transform.inverse(theta0)

    def fun(xi_free):
        return objective_and_grad_xi(kind, X, y, xi_free, theta0, trainable, transform, mode, priors)

    if not trainable.any():
        return theta0, lml(kind, X, y, theta0, mode), True, 0
    res = sopt.minimize(fun, xi0[trainable], jac=True, method="L-BFGS-B",
                        options=dict(maxiter=max_iter))
    xi = xi0.copy()
    xi[trainable] = res.x
    theta = np.where(trainable, transform.forward(xi), theta0)
    out = (theta, lml(kind, X, y, theta, mode), bool(res.success), int(res.nfev))
    if return_result:
        return out + (res,)
    return out
