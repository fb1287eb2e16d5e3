"""The device L-BFGS state machine (gpsat_b200/csrc/lbfgs.cuh), compiled for the host, against
scipy.optimize.minimize(method="L-BFGS-B") - the optimiser the reference reaches through
gpflow.optimizers.Scipy (GPSat/models/gpflow_models.py:317-321).  CPU only."""
import ctypes as C
import os

import numpy as np
import pytest
import scipy.optimize as sopt

from oracle import gpr

HERE = os.path.dirname(os.path.abspath(__file__))
CB = C.CFUNCTYPE(None, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double), C.c_void_p)
RC_NAMES = {1: "pgtol", 2: "ftol", 3: "maxiter", 4: "maxfun", 5: "abnormal", 6: "not_finite"}


@pytest.fixture(scope="module")
def lib(built):
    return C.CDLL(os.path.join(HERE, "helpers", "liblbfgs_host.so"))


def minimize(lib, fun, x0, maxiter=10000, maxfun=15000, maxls=20):
    n = len(x0)

    def cb(xp, fp, gp, _):
        x = np.array([xp[i] for i in range(n)])
        f, g = fun(x)
        fp[0] = f
        for i in range(n):
            gp[i] = g[i]

    x = np.array(x0, float)
    f, nfev, nit = C.c_double(), C.c_int(), C.c_int()
    trace = np.zeros((4000, n))
    pd = lambda a: a.ctypes.data_as(C.POINTER(C.c_double))
    rc = lib.lbfgs_host_minimize(n, pd(x), CB(cb), None, C.c_double(1e-5), C.c_double(2.220446049250313e-09),
                                 maxiter, maxfun, maxls, C.byref(f), C.byref(nfev), C.byref(nit), pd(trace), 4000)
    return x, f.value, rc, nfev.value, nit.value, trace[:nfev.value]


def rosen(x):
    return sopt.rosen(x), sopt.rosen_der(x)


@pytest.mark.parametrize("x0", [[-1.2, 1.0], [2.0, -3.0, 1.5, 0.3, -1.0], [0.0] * 7])
def test_rosenbrock_same_path_as_scipy(lib, x0):
    xs = []

    def f2(x):
        xs.append(x.copy())
        return rosen(x)

    r = sopt.minimize(f2, x0, jac=True, method="L-BFGS-B")
    x, f, rc, nfev, nit, tr = minimize(lib, rosen, x0)
    assert (nfev, nit) == (r.nfev, r.nit)
    assert rc in (1, 2) and r.success
    np.testing.assert_allclose(tr, np.array(xs), atol=1e-7)
    np.testing.assert_allclose(x, r.x, atol=1e-9)


@pytest.mark.parametrize("kind", gpr.KERNELS)
def test_gp_objective_same_path_as_scipy(lib, kind):
    rng = np.random.default_rng(1)
    n = 70
    X = rng.uniform(-4, 4, (n, 3))
    K = gpr.kernel_matrix("Matern32", X, None, [2.0, 1.5, 3.0], 0.7)
    y = np.linalg.cholesky(K + 1e-10 * np.eye(n)) @ rng.standard_normal(n) + 0.2 * rng.standard_normal(n)
    tr_ = gpr.Transform.gpflow_default(3)
    tr_.set_sigmoid([0, 1, 2], [1e-8] * 3, [12, 12, 9.0])
    th0 = np.ones(5)
    trainable = np.ones(5, bool)
    fun = lambda xi: gpr.objective_and_grad_xi(kind, X, y, xi, th0, trainable, tr_)
    xi0 = tr_.inverse(th0)
    r = sopt.minimize(fun, xi0, jac=True, method="L-BFGS-B", options=dict(maxiter=10000))
    x, f, rc, nfev, nit, _ = minimize(lib, fun, xi0)
    assert (nfev, nit) == (r.nfev, r.nit)
    assert (rc in (1, 2)) == bool(r.success)
    np.testing.assert_allclose(tr_.forward(x), tr_.forward(r.x), rtol=1e-7)
    assert abs(f - r.fun) <= 1e-9 * abs(r.fun)


def test_maxiter_semantics(lib):
    r = sopt.minimize(rosen, [-1.2, 1.0], jac=True, method="L-BFGS-B", options=dict(maxiter=5))
    x, f, rc, nfev, nit, _ = minimize(lib, rosen, [-1.2, 1.0], maxiter=5)
    assert rc == 3 and not r.success
    assert (nfev, nit) == (r.nfev, r.nit)
    np.testing.assert_allclose(x, r.x, atol=1e-12)


def test_already_converged_start(lib):
    f0 = lambda x: (float(x @ x), 2 * x)
    x, f, rc, nfev, nit, _ = minimize(lib, f0, [0.0, 0.0])
    assert rc == 1 and nfev == 1 and nit == 0


def test_non_finite_start_and_trial(lib):
    bad = lambda x: (np.inf, np.zeros_like(x))
    assert minimize(lib, bad, [1.0, 2.0])[2] == 6

    def wall(x):   # objective undefined (non-PD stand-in) right of x = 1.5
        if x[0] > 1.5:
            return np.inf, np.zeros_like(x)
        return float((x[0] - 2.0) ** 2), np.array([2 * (x[0] - 2.0)])

    x, f, rc, nfev, nit, _ = minimize(lib, wall, [-3.0])
    assert np.isfinite(f) and x[0] <= 1.5 and f < 25.0
