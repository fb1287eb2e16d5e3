"""Parity of the CUDA path (through the C ABI) with the CPU oracle.  Needs a B200: -m gpu.

Tolerances (BASELINE.json north_star): relative 1e-8 on LML and predictive mean / variance at
fixed hyper-parameters; 1e-4 on optimised hyper-parameters and the resulting LML.
"""
import os

import numpy as np
import pytest

from oracle import gpr

pytestmark = pytest.mark.gpu

RTOL_FIXED = 1e-8
RTOL_OPT = 1e-4


def _expert(rng, n, D, m=0, noise=0.2):
    X = rng.uniform(-4, 4, (n, D))
    K = gpr.kernel_matrix("Matern32", X, None, [2.0, 1.5, 3.0][:D], 0.7)
    y = np.linalg.cholesky(K + 1e-8 * np.eye(n)) @ rng.standard_normal(n) + noise * rng.standard_normal(n)
    return X, y, rng.uniform(-4, 4, (m, D))


def _batch(kind, experts, with_pred=True):
    from gpsat_b200.engine import RaggedBatch
    return RaggedBatch.from_lists(kind, [e[0] for e in experts], [e[1] for e in experts],
                                  [e[2] for e in experts] if with_pred else None)


def _theta(rng, D):
    return np.concatenate([rng.uniform(0.6, 2.5, D), [rng.uniform(0.3, 1.5)], [rng.uniform(0.02, 0.3)]])


@pytest.mark.parametrize("n", [1, 2, 5, 127, 128, 129, 255, 256, 257, 600])
def test_factor_matches_numpy(engine, n):
    rng = np.random.default_rng(n)
    X, y, _ = _expert(rng, n, 3)
    th = _theta(rng, 3)
    Ks = gpr.kernel_matrix("Matern32", X, None, th[:3], th[3]) + th[4] * np.eye(n)
    Lr = np.linalg.cholesky(Ks)
    a = np.linalg.solve(Lr, y)
    # default configuration: the low-rho tier of the precision rule (one slice-pair group fewer) is active at
    # these hyper-parameters for n > 128; the observation row of the factor, whose a-priori scale |y| / sigma_n is the
    # loosest, shows it first
    L, U = engine.factor_debug("Matern32", X, y, th)
    np.testing.assert_allclose(L[:n, :n], Lr, rtol=0, atol=1e-11 * np.abs(Lr).max())
    np.testing.assert_allclose(L[n, :n], a, rtol=0, atol=1e-9 * max(1.0, np.abs(a).max()))
    try:                                   # all slice pairs: the factor to 1e-12 of its largest entry
        engine.set_precision_low_tier(False)
        L, U = engine.factor_debug("Matern32", X, y, th)
    finally:
        engine.set_precision_low_tier(True)
    np.testing.assert_allclose(L[:n, :n], Lr, rtol=0, atol=1e-12 * np.abs(Lr).max())
    np.testing.assert_allclose(L[n, :n], a, rtol=0, atol=1e-11 * max(1.0, np.abs(a).max()))
    assert L[n, n] == 1.0
    # property: L L^T reproduces K + s2 I, U = L_aug^-T
    np.testing.assert_allclose(L[:n, :n] @ L[:n, :n].T, Ks, rtol=0, atol=1e-12 * np.abs(Ks).max())
    np.testing.assert_allclose(U.T @ L, np.eye(n + 1), rtol=0, atol=1e-9)
    assert np.all(np.triu(L, 1) == 0) and np.all(np.tril(U, -1) == 0)


@pytest.mark.parametrize("kind", gpr.KERNELS)
@pytest.mark.parametrize("D", [1, 2, 3])
def test_lml_grad_predict_ragged(engine, kind, D):
    rng = np.random.default_rng(100 * D + len(kind))
    ns = [3, 40, 130, 257, 420, 64]
    experts = [_expert(rng, n, D, m=7 + n // 3) for n in ns]
    experts[1] = (experts[1][0], experts[1][1], np.zeros((0, D)))       # an expert without prediction rows
    b = _batch(kind, experts)
    theta = np.stack([_theta(rng, D) for _ in ns])
    lml, grad, status = engine.lml_grad(b, theta)
    pred = engine.predict(b, theta)
    assert np.all(status == 0) and np.all(pred["status"] == 0)
    for e, (X, y, Xs) in enumerate(experts):
        v, g = gpr.lml_and_grad(kind, X, y, theta[e])
        assert abs(lml[e] - v) <= RTOL_FIXED * abs(v)
        assert abs(pred["lml"][e] - v) <= RTOL_FIXED * abs(v)
        np.testing.assert_allclose(grad[e], g, rtol=1e-7, atol=1e-8 * np.abs(g).max())
        if len(Xs):
            p = gpr.predict(kind, X, y, theta[e], Xs)
            sl = slice(b.off_m[e], b.off_m[e + 1])
            np.testing.assert_allclose(pred["f*"][sl], p["f*"], rtol=RTOL_FIXED, atol=1e-10)
            np.testing.assert_allclose(pred["f*_var"][sl], p["f*_var"], rtol=RTOL_FIXED, atol=1e-12)
            np.testing.assert_allclose(pred["y_var"][sl], p["y_var"], rtol=RTOL_FIXED, atol=1e-12)


def test_lml_only_matches(engine):
    rng = np.random.default_rng(3)
    experts = [_expert(rng, n, 3) for n in (50, 300)]
    b = _batch("Matern52", experts, with_pred=False)
    theta = np.stack([_theta(rng, 3) for _ in experts])
    a, _, _ = engine.lml_grad(b, theta, with_grad=False)
    c, _, _ = engine.lml_grad(b, theta, with_grad=True)
    assert np.array_equal(a, c)


def test_golden_pure_python_through_cuda(engine, golden_dir):
    """Reference's in-repo NumPy GPR (pure_python_gpr.py:439-553) vs the CUDA path."""
    from gpsat_b200.engine import RaggedBatch
    g = np.load(os.path.join(golden_dir, "pure_python_gpr.npz"))
    X, y, Xs = g["X"], g["y"], g["Xs"]
    k = len(g["thetas"])
    b = RaggedBatch.from_lists("Matern32", [X] * k, [y] * k, [Xs] * k)
    pred = engine.predict(b, g["thetas"])
    for e in range(k):
        nlz = float(np.ravel(g["nlZ"][e])[0])
        assert abs(-pred["lml"][e] - nlz) <= RTOL_FIXED * abs(nlz)
        sl = slice(b.off_m[e], b.off_m[e + 1])
        np.testing.assert_allclose(pred["f*"][sl], g["fmean"][e], rtol=RTOL_FIXED, atol=1e-10)
        np.testing.assert_allclose(pred["f*_var"][sl], g["fvar"][e], rtol=1e-7, atol=1e-12)


@pytest.mark.parametrize("kind", gpr.KERNELS)
def test_golden_sklearn_through_cuda(engine, golden_dir, kind):
    from gpsat_b200.engine import RaggedBatch
    g = np.load(os.path.join(golden_dir, "sklearn_ard.npz"))
    b = RaggedBatch.from_lists(kind, [g["X"]], [g["y"]], [g["Xs"]])
    th = g["theta"][None, :]
    lml, grad, _ = engine.lml_grad(b, th)
    pred = engine.predict(b, th)
    assert abs(lml[0] - float(g[f"{kind}_lml"])) <= RTOL_FIXED * abs(lml[0])
    np.testing.assert_allclose(grad[0, :3], g[f"{kind}_grad_ls"], rtol=1e-7)
    np.testing.assert_allclose(grad[0, 3], g[f"{kind}_grad_var"], rtol=1e-7)
    np.testing.assert_allclose(pred["f*"], g[f"{kind}_fmean"], rtol=RTOL_FIXED, atol=1e-10)
    np.testing.assert_allclose(pred["f*_var"], g[f"{kind}_fvar"], rtol=1e-7, atol=1e-12)


def test_golden_kat_fit_through_cuda(engine, golden_dir):
    """tests/test_localexperts.py::test_gpflow_gpr of the reference (tolerance 1e-6 there):
    likelihood variance fixed at eps^2, kernel variance fixed, lengthscale in [1e-10, 5]."""
    from gpsat_b200.engine import RaggedBatch
    g = np.load(os.path.join(golden_dir, "kat_localexperts.npz"))
    eps = float(g["eps"])
    b = RaggedBatch.from_lists("Matern32", [g["x_train"]], [g["y_train"][:, 0]], [g["x_test"]])
    res = engine.fit(b, [[1.0, 1.0, eps ** 2]], [[1, 0, 0]], [[1e-10, 0, 1e-6]], [[5.0, 0, 0]], [[1, 0, 0]])
    assert res.status[0] == 0
    assert abs(res.theta[0, 0] - float(g["ls"])) < 1e-6
    assert res.theta[0, 1] == 1.0 and res.theta[0, 2] == eps ** 2
    assert abs(res.lml[0] - float(g["ml"])) < 1e-6
    pred = engine.predict(b, res.theta)
    assert abs(pred["f*"][0] - g["pred_mean"][0]) < 1e-6
    assert abs(pred["f*_var"][0] - g["pred_std"][0] ** 2) < 1e-6


@pytest.mark.parametrize("kind", ["Matern32", "RBF", "Matern12", "Matern52"])
def test_fit_matches_scipy_lbfgsb(engine, kind):
    """More experts than slots (max_slots=3) so slots are retired and refilled on the device."""
    rng = np.random.default_rng(len(kind))
    ns = [60, 150, 300, 90, 200, 35, 131, 260]
    experts = [_expert(rng, n, 3) for n in ns]
    b = _batch(kind, experts, with_pred=False)
    E = b.E
    tr = gpr.Transform.gpflow_default(3)
    tr.set_sigmoid([0, 1, 2], [1e-8] * 3, [12, 12, 9.0])
    th0 = np.ones((E, 5))
    res = engine.fit(b, th0, np.tile(tr.kind, (E, 1)), np.tile(tr.a, (E, 1)), np.tile(tr.b, (E, 1)),
                     np.ones((E, 5), np.uint8), max_slots=3)
    for e, (X, y, _) in enumerate(experts):
        th, l, ok, nev = gpr.fit(kind, X, y, th0[e], tr)
        assert (res.status[e] == 0) == ok
        assert abs(res.lml[e] - l) <= RTOL_OPT * abs(l)
        # 1e-4 per component, skipped for components pinned at a bound +- tol (here the
        # likelihood variance can run into its floor 1e-6 + softplus, where the objective is flat and the last
        # digits of the softplus part are decided by rounding)
        pinned = (th - tr.a < 1e-2 * np.maximum(tr.a, 1e-300)) & (tr.kind == 0) & (tr.a > 0)
        np.testing.assert_allclose(res.theta[e][~pinned], th[~pinned], rtol=RTOL_OPT)
        np.testing.assert_allclose(res.theta[e][pinned], th[pinned], rtol=1e-2)
        assert abs(int(res.n_evals[e]) - nev) <= 2      # same optimisation path as SciPy


def test_fit_fixed_parameters_and_maxiter(engine):
    rng = np.random.default_rng(11)
    experts = [_expert(rng, n, 2) for n in (80, 140)]
    b = _batch("Matern32", experts, with_pred=False)
    tr = gpr.Transform.gpflow_default(2)
    th0 = np.array([[1.0, 1.0, 0.7, 0.04]] * 2)
    trainable = np.array([[1, 1, 0, 0]] * 2, np.uint8)
    res = engine.fit(b, th0, np.tile(tr.kind, (2, 1)), np.tile(tr.a, (2, 1)), np.tile(tr.b, (2, 1)), trainable)
    for e, (X, y, _) in enumerate(experts):
        th, l, ok, nev = gpr.fit("Matern32", X, y, th0[e], tr, trainable=trainable[e].astype(bool))
        np.testing.assert_allclose(res.theta[e], th, rtol=RTOL_OPT)
        assert np.array_equal(res.theta[e, 2:], th0[e, 2:])
    # maxiter reached => status MAXITER (SciPy success False), iterate equals SciPy's
    res2 = engine.fit(b, th0, np.tile(tr.kind, (2, 1)), np.tile(tr.a, (2, 1)), np.tile(tr.b, (2, 1)),
                      np.ones((2, 4), np.uint8), max_iter=3)
    for e, (X, y, _) in enumerate(experts):
        th, l, ok, nev = gpr.fit("Matern32", X, y, th0[e], tr, max_iter=3)
        assert not ok and res2.status[e] == 3 and res2.n_iters[e] == 3
        np.testing.assert_allclose(res2.theta[e], th, rtol=1e-7)


def test_shard_equality_is_bitwise(engine):
    """Experts never interact: an expert evaluated / optimised alone gives bit-identical results to
    the same expert inside a batch (what makes 1-GPU and N-GPU runs identical)."""
    rng = np.random.default_rng(21)
    experts = [_expert(rng, n, 3, m=20) for n in (70, 190, 310, 129)]
    b = _batch("Matern32", experts)
    E = b.E
    theta = np.stack([_theta(rng, 3) for _ in range(E)])
    lml, grad, _ = engine.lml_grad(b, theta)
    pred = engine.predict(b, theta)
    tr = gpr.Transform.gpflow_default(3)
    args = lambda k: (np.tile(tr.kind, (k, 1)), np.tile(tr.a, (k, 1)), np.tile(tr.b, (k, 1)), np.ones((k, 5), np.uint8))
    fit_all = engine.fit(b, np.ones((E, 5)), *args(E), max_iter=6)
    for e in range(E):
        sub = b.select([e])
        l1, g1, _ = engine.lml_grad(sub, theta[e:e + 1])
        p1 = engine.predict(sub, theta[e:e + 1])
        assert l1[0] == lml[e] and np.array_equal(g1[0], grad[e])
        sl = slice(b.off_m[e], b.off_m[e + 1])
        assert np.array_equal(p1["f*"], pred["f*"][sl]) and np.array_equal(p1["f*_var"], pred["f*_var"][sl])
        f1 = engine.fit(sub, np.ones((1, 5)), *args(1), max_iter=6)
        assert np.array_equal(f1.theta[0], fit_all.theta[e]) and f1.lml[0] == fit_all.lml[e]


def test_device_pointers_equal_host_path(engine):
    import torch
    rng = np.random.default_rng(31)
    experts = [_expert(rng, n, 3, m=15) for n in (50, 200)]
    b = _batch("Matern32", experts)
    theta = np.stack([_theta(rng, 3) for _ in experts])
    bd = b.to_device("cuda:0")
    l_h, g_h, _ = engine.lml_grad(b, theta)
    l_d, g_d, _ = engine.lml_grad(bd, theta)
    assert isinstance(l_d, torch.Tensor) and l_d.is_cuda
    assert np.array_equal(l_d.cpu().numpy(), l_h) and np.array_equal(g_d.cpu().numpy(), g_h)
    p_h, p_d = engine.predict(b, theta), engine.predict(bd, theta)
    assert np.array_equal(p_d["f*"].cpu().numpy(), p_h["f*"])


def test_duplicates_and_not_pd_status(engine):
    rng = np.random.default_rng(41)
    X, y, Xs = _expert(rng, 60, 3, m=5)
    X[10:20] = X[0:10]                       # exact duplicates (binned data has them)
    b = _batch("Matern12", [(X, y, Xs)])
    th = np.array([[1.5, 1.0, 2.0, 0.8, 0.05]])
    lml, grad, st = engine.lml_grad(b, th)
    v, g = gpr.lml_and_grad("Matern12", X, y, th[0])
    assert st[0] == 0 and abs(lml[0] - v) <= RTOL_FIXED * abs(v)
    np.testing.assert_allclose(grad[0], g, rtol=1e-7)
    th_bad = np.array([[1.5, 1.0, 2.0, 0.8, 0.0]])    # singular: duplicates and no noise
    lml, grad, st = engine.lml_grad(b, th_bad)
    assert st[0] == 1 and lml[0] == -np.inf
    good = _expert(rng, 40, 3, m=5)
    b2 = _batch("Matern12", [(X, y, Xs), good])       # a bad expert must not poison its neighbour
    lml2, _, st2 = engine.lml_grad(b2, np.vstack([th_bad, th]))
    assert list(st2) == [1, 0]
    assert abs(lml2[1] - gpr.lml("Matern12", good[0], good[1], th[0])) <= RTOL_FIXED * abs(lml2[1])


def test_empty_batch_and_api_misuse(engine):
    from gpsat_b200.engine import RaggedBatch
    b = RaggedBatch("Matern32", 3, np.zeros(1, np.int64), np.zeros((0, 3)), np.zeros(0),
                    np.zeros(1, np.int64), np.zeros((0, 3)))
    lml, grad, st = engine.lml_grad(b, np.zeros((0, 5)))
    assert lml.shape == (0,)
    assert engine.predict(b, np.zeros((0, 5)))["f*"].shape == (0,)
    bad = RaggedBatch("Matern32", 9, np.array([0, 4], np.int64), np.zeros((4, 9)), np.zeros(4))
    with pytest.raises(RuntimeError, match="D out of range"):
        engine.lml_grad(bad, np.ones((1, 11)))
    # the sparse and the Fourier-feature entry points: empty batch, bad inducing sets
    assert engine.sgpr_elbo(b, [], np.zeros((0, 5)))[0].shape == (0,)
    assert engine.sgpr_predict(b, [], np.zeros((0, 5)))["f*"].shape == (0,)
    assert engine.vff_elbo(b, (np.zeros((0, 3)), np.zeros((0, 3)), np.zeros((0, 3), np.int32)), np.zeros((0, 7)))[0].shape == (0,)
    one = RaggedBatch.from_lists("Matern32", [np.random.default_rng(0).uniform(-1, 1, (20, 3))], [np.zeros(20)])
    with pytest.raises(RuntimeError, match="inducing points"):
        engine.sgpr_elbo(one, [np.zeros((0, 3))], np.ones((1, 5)))
    with pytest.raises(ValueError, match="one inducing-point array per expert"):
        engine.sgpr_elbo(one, [], np.ones((1, 5)))


@pytest.mark.parametrize("n", [2000, 4000])
def test_full_size_expert(engine, n):
    """BASELINE config sizes (c3 / north-star): one expert checked directly against LAPACK, plus
    size-independent properties (symmetric gradient identity, reproducibility)."""
    from gpsat_b200 import synthetic as syn
    spec = syn.WorkloadSpec("t", 2, n, n, 64, 64, 400e3, 8, (50e3, 50e3, 1.0), (600e3, 600e3, 16.0))
    b = syn.make_batch(spec, seed=5)
    theta = np.tile(np.array([3.0, 3.5, 4.0, 0.06, 0.006]), (2, 1))
    lml, grad, st = engine.lml_grad(b, theta)
    pred = engine.predict(b, theta)
    X, y, Xs = b.X[:n], b.y[:n], b.Xs[:64]
    v, g = gpr.lml_and_grad("Matern32", X, y, theta[0])
    assert st[0] == 0 and abs(lml[0] - v) <= RTOL_FIXED * abs(v)
    np.testing.assert_allclose(grad[0], g, rtol=1e-6, atol=1e-7 * np.abs(g).max())
    p = gpr.predict("Matern32", X, y, theta[0], Xs)
    np.testing.assert_allclose(pred["f*"][:64], p["f*"], rtol=RTOL_FIXED, atol=1e-10)
    np.testing.assert_allclose(pred["f*_var"][:64], p["f*_var"], rtol=RTOL_FIXED, atol=1e-12)
    lml_b, grad_b, _ = engine.lml_grad(b, theta)
    assert np.array_equal(lml, lml_b) and np.array_equal(grad, grad_b)


def test_bitwise_repeatable_and_order_independent(engine):
    """Race / ordering proxy (compute-sanitizer is closed on this pool): every reduction in the sweep
    has a fixed order, so repeated calls and a permuted batch give bit-identical results."""
    rng = np.random.default_rng(77)
    ns = [517, 130, 700, 64, 385, 257, 1030]
    experts = [_expert(rng, n, 3, m=20 + n // 7) for n in ns]
    theta = np.stack([_theta(rng, 3) for _ in ns])
    b = _batch("Matern52", experts)
    l1, g1, _ = engine.lml_grad(b, theta)
    p1 = engine.predict(b, theta)
    for _ in range(2):
        l2, g2, _ = engine.lml_grad(b, theta)
        p2 = engine.predict(b, theta)
        assert np.array_equal(l1, l2) and np.array_equal(g1, g2)
        assert np.array_equal(p1["f*"], p2["f*"]) and np.array_equal(p1["f*_var"], p2["f*_var"])
    perm = [3, 6, 0, 5, 1, 4, 2]
    bp = _batch("Matern52", [experts[k] for k in perm])
    lp, gp, _ = engine.lml_grad(bp, theta[perm])
    assert np.array_equal(lp, l1[perm]) and np.array_equal(gp, g1[perm])


@pytest.mark.parametrize("n", [2000, 3100])
def test_full_size_single_expert(engine, n):
    """One expert at the sizes of configs c3 / north star (BASELINE.json): LML, gradient and
    predictions against the oracle (LAPACK) at the fixed-theta tolerance."""
    rng = np.random.default_rng(n)
    X = rng.uniform(-8, 8, (n, 3))
    f = np.sin(X[:, 0]) * np.cos(0.7 * X[:, 1]) + 0.3 * np.sin(0.5 * X[:, 2])
    y = f + 0.1 * rng.standard_normal(n)
    Xs = rng.uniform(-8, 8, (300, 3))
    th = np.array([[1.7, 2.1, 3.0, 0.6, 0.02]])
    b = _batch("Matern32", [(X, y, Xs)])
    lml, grad, st = engine.lml_grad(b, th)
    pred = engine.predict(b, th)
    v, g = gpr.lml_and_grad("Matern32", X, y, th[0])
    assert st[0] == 0 and abs(lml[0] - v) <= RTOL_FIXED * abs(v)
    np.testing.assert_allclose(grad[0], g, rtol=1e-7, atol=1e-8 * np.abs(g).max())
    p = gpr.predict("Matern32", X, y, th[0], Xs)
    np.testing.assert_allclose(pred["f*"], p["f*"], rtol=RTOL_FIXED, atol=1e-9)
    np.testing.assert_allclose(pred["f*_var"], p["f*_var"], rtol=RTOL_FIXED, atol=1e-11)


def test_property_random_shapes(engine):
    """Property test (hypothesis): any n in [1, 600], D in {1, 2, 3}, any of the four kernels, any
    hyper-parameters in a plausible box - LML, gradient and predictions agree with the oracle."""
    from hypothesis import given, settings, strategies as st, HealthCheck

    @settings(max_examples=30, deadline=None, derandomize=True, suppress_health_check=list(HealthCheck))
    @given(n=st.integers(1, 600), D=st.integers(1, 3), kind=st.sampled_from(list(gpr.KERNELS)),
           m=st.integers(0, 40), seed=st.integers(0, 2 ** 16),
           ls=st.floats(0.3, 4.0), var=st.floats(0.05, 3.0), noise=st.floats(1e-3, 0.5))
    def check(n, D, kind, m, seed, ls, var, noise):
        rng = np.random.default_rng(seed)
        X, y, Xs = _expert(rng, n, D, m=m)
        th = np.concatenate([np.full(D, ls) * rng.uniform(0.7, 1.4, D), [var, noise]])[None, :]
        b = _batch(kind, [(X, y, Xs)])
        lml, grad, st_ = engine.lml_grad(b, th)
        v, g = gpr.lml_and_grad(kind, X, y, th[0])
        assert st_[0] == 0
        assert abs(lml[0] - v) <= RTOL_FIXED * max(abs(v), 1.0)
        np.testing.assert_allclose(grad[0], g, rtol=1e-7, atol=1e-7 * max(np.abs(g).max(), 1.0))
        if m:
            pred = engine.predict(b, th)
            p = gpr.predict(kind, X, y, th[0], Xs)
            np.testing.assert_allclose(pred["f*"], p["f*"], rtol=RTOL_FIXED, atol=1e-10 * max(1.0, np.abs(y).max()))
            np.testing.assert_allclose(pred["f*_var"], p["f*_var"], rtol=RTOL_FIXED, atol=1e-11 * var)

    check()


def test_gradient_matches_finite_differences_of_the_cuda_lml(engine):
    """Size-independent property at a c3-sized expert: the analytic gradient equals central finite
    differences of the engine's own LML (no oracle involved)."""
    rng = np.random.default_rng(8)
    X, y, _ = _expert(rng, 2500, 3)
    th = np.array([1.8, 1.4, 2.6, 0.7, 0.08])
    b = _batch("Matern32", [(X, y, np.zeros((0, 3)))])
    _, grad, _ = engine.lml_grad(b, th[None, :])
    for p in range(5):
        h = 1e-5 * th[p]
        tp, tm = th.copy(), th.copy()
        tp[p] += h
        tm[p] -= h
        lp, _, _ = engine.lml_grad(b, tp[None, :], with_grad=False)
        lm, _, _ = engine.lml_grad(b, tm[None, :], with_grad=False)
        fd = (lp[0] - lm[0]) / (2 * h)
        assert abs(fd - grad[0][p]) <= 2e-5 * max(1.0, abs(grad[0][p])), (p, fd, grad[0][p])


@pytest.mark.parametrize("kind,D", [("Matern32", 3), ("RBF", 2), ("Matern52", 1)])
def test_i8_gradient_path_matches_dmma_path(engine, kind, D):
    """The K^-1 tile products of the gradient on the int8 tensor path (error-free slicing, i8mma.cuh)
    against the FP64 mma.sync path on the same factor, and against the oracle: ragged sizes around
    the tile and slab edges (k ranges of 1..5 blocks, partial last slabs, diagonal and off-diagonal tiles)."""
    t0 = engine.i8_timeouts()
    rng = np.random.default_rng(77 + D)
    ns = [1, 15, 16, 31, 100, 127, 128, 129, 200, 256, 383, 600]
    experts = [_expert(rng, n, D) for n in ns]
    b = _batch(kind, experts, with_pred=False)
    theta = np.stack([_theta(rng, D) for _ in ns])
    try:
        engine.set_lauum_path("dmma")
        lml0, g0, st0 = engine.lml_grad(b, theta)
        engine.set_lauum_path("i8")
        lml1, g1, st1 = engine.lml_grad(b, theta)
    finally:
        engine.set_lauum_path("i8")
    assert engine.i8_timeouts() == t0
    assert np.all(st0 == 0) and np.all(st1 == 0)
    np.testing.assert_array_equal(lml0, lml1)                  # the factorisation is the same code
    for e, (X, y, _) in enumerate(experts):
        scale = np.abs(g0[e]).max()
        np.testing.assert_allclose(g1[e], g0[e], rtol=1e-10, atol=1e-11 * scale)
        v, g = gpr.lml_and_grad(kind, X, y, theta[e])
        np.testing.assert_allclose(g1[e], g, rtol=1e-7, atol=1e-8 * np.abs(g).max())


@pytest.mark.parametrize("kind,D,noise", [("Matern32", 3, None), ("RBF", 2, None), ("Matern12", 1, None),
                                          ("Matern52", 3, 1e-4)])
def test_i8_step_path_matches_dmma_path(engine, kind, D, noise):
    """The potrf / trtri main loops on the int8 tensor path (8 exact slices with a-priori row scales,
    step_kernel.cuh) against the FP64 mma.sync loops and against the oracle: experts of 2..11 tile rows, ragged
    sizes around the tile edges, prediction rows, and a badly conditioned case (noise 1e-4, (var + noise) / noise
    up to 1.5e4) where the a-priori scale of U = L^-T is far above its typical entries.  Three configurations:
    FP64 mma.sync, int8-sliced forced on every slot (precision rule off), and the default (rule on: the badly
    conditioned experts run on FP64 mma.sync) - the default is held to the north-star tolerances in every case."""
    t0 = engine.i8_timeouts()
    rng = np.random.default_rng(177 + D)
    ns = [129, 200, 256, 257, 383, 640, 1000, 1300]
    experts = [_expert(rng, n, D, m=37 + (n % 200)) for n in ns]
    b = _batch(kind, experts, with_pred=True)
    theta = np.stack([_theta(rng, D) for _ in ns])
    if noise is not None:
        theta[:, -1] = noise
    out = {}
    try:
        for name, path, rule in (("dmma", "dmma", False), ("i8", "i8", False), ("default", "i8", True)):
            engine.set_step_path(path)
            engine.set_precision_rule(rule)
            engine.set_precision_low_tier(rule)            # "i8": all 36 slice pairs on every slot
            lml, g, st = engine.lml_grad(b, theta)
            p = engine.predict(b, theta)
            L, U = engine.factor_debug(kind, experts[3][0], experts[3][1], theta[3])
            out[name] = (lml, g, st, p, L, U)
    finally:
        engine.set_step_path("i8")
        engine.set_precision_rule(True)
        engine.set_precision_low_tier(True)
    assert engine.i8_timeouts() == t0
    (lml0, g0, st0, p0, L0, U0), (lml1, g1, st1, p1, L1, U1) = out["dmma"], out["i8"]
    lml2, g2, st2, p2, _, _ = out["default"]
    assert np.all(st0 == 0) and np.all(st1 == 0) and np.all(st2 == 0)
    assert not np.array_equal(lml0, lml1)                  # the two paths really are different arithmetic
    rho = (theta[:, D] + theta[:, D + 1]) / theta[:, D + 1]
    from gpsat_b200 import _lib
    rho_max = _lib.load().gpsat_engine_precision_rule_default()
    rho_low = _lib.load().gpsat_engine_precision_low_tier_default()
    for e in range(len(ns)):                               # the rule picks the path per expert
        if rho[e] <= rho_low:                              # low tier: 28 of the 36 slice pairs
            assert abs(lml2[e] - lml1[e]) <= 1e-9 * abs(lml1[e])
        else:
            assert lml2[e] == (lml1[e] if rho[e] <= rho_max else lml0[e])
    # factor of the 257-observation expert: elementwise against the row scales the slicing uses
    np.testing.assert_allclose(L1, L0, rtol=0, atol=1e-12 * np.abs(L0).max())
    np.testing.assert_allclose(U1, U0, rtol=0, atol=1e-10 * np.abs(U0).max())
    loose = noise is not None                              # forced int8 slicing at rho ~ 1e4: the regime the rule excludes
    np.testing.assert_allclose(lml1, lml0, rtol=1e-9 if loose else 1e-11)
    off = b.off_m
    for e, (X, y, Xs) in enumerate(experts):
        scale = np.abs(g0[e]).max()
        np.testing.assert_allclose(g1[e], g0[e], rtol=1e-6 if loose else 1e-8, atol=(1e-7 if loose else 1e-9) * scale)
        v, g = gpr.lml_and_grad(kind, X, y, theta[e])
        pr = gpr.predict(kind, X, y, theta[e], Xs)
        sl = slice(off[e], off[e + 1])
        for (lm_, g_, p_, tl) in ((lml1, g1, p1, 100.0 if loose else 1.0), (lml2, g2, p2, 1.0)):
            assert abs(lm_[e] - v) <= 1e-8 * abs(v)
            np.testing.assert_allclose(g_[e], g, rtol=1e-7 * tl, atol=1e-8 * tl * np.abs(g).max())
            tol = 1e-8 * tl
            np.testing.assert_allclose(p_["f*"][sl], pr["f*"], rtol=tol, atol=tol * np.abs(pr["f*"]).max())
            np.testing.assert_allclose(p_["f*_var"][sl], pr["f*_var"], rtol=tol, atol=tol * theta[e][D])
        tol = 1e-6 if loose else 1e-8
        np.testing.assert_allclose(p1["f*"][sl], p0["f*"][sl], rtol=tol / 10, atol=tol / 10 * np.abs(pr["f*"]).max())


def test_persistent_lauum_is_bitwise_equal(engine):
    """lauum_grad_i8q_kernel (persistent, warp-specialised, Q tile through an L2-resident scratch; GPSAT_LAUUM_I8Q)
    against lauum_grad_i8_kernel (one CTA per tile, the default): the same arithmetic, so LML and gradient must agree bit for bit - ragged
    sizes from one to eleven tile rows, more tiles than SMs, a slot the precision rule sends to FP64 mma.sync, and a
    slot with 8 planes (var > 255 noise)."""
    rng = np.random.default_rng(4242)
    ns = [1, 100, 128, 129, 257, 383, 640, 1000, 1300] + [520] * 30
    experts = [_expert(rng, n, 3) for n in ns]
    b = _batch("Matern32", experts, with_pred=False)
    theta = np.stack([_theta(rng, 3) for _ in ns])
    theta[5, 3:] = [1.0, 1e-4]          # rho = 1e4: FP64 path by the precision rule
    theta[6, 3:] = [1.5, 2e-3]          # rho = 750: 8 planes in the K^-1 products
    t0 = engine.i8_timeouts()
    try:
        engine.set_lauum_path("i8")
        lml0, g0, st0 = engine.lml_grad(b, theta)
        engine.set_lauum_path("i8q")
        lml1, g1, st1 = engine.lml_grad(b, theta)
        lml2, g2, st2 = engine.lml_grad(b, theta)
    finally:
        engine.set_lauum_path("i8")
    assert engine.i8_timeouts() == t0
    assert np.all(st0 == 0) and np.all(st1 == 0)
    np.testing.assert_array_equal(lml0, lml1)
    np.testing.assert_array_equal(g0, g1)
    np.testing.assert_array_equal(g1, g2)              # and repeatable
    v, g = gpr.lml_and_grad("Matern32", experts[7][0], experts[7][1], theta[7])
    np.testing.assert_allclose(g1[7], g, rtol=1e-7, atol=1e-8 * np.abs(g).max())


def _envelope_run(engine, cases, items, kind):
    b = _batch(kind, [(it[0], it[1], it[2]) for it in items], with_pred=True)
    theta = np.stack([it[3] for it in items])
    lml, grad, st = engine.lml_grad(b, theta)
    p = engine.predict(b, theta)
    return lml, grad, st, p, b.off_m


@pytest.mark.parametrize("kind", ["Matern32", "RBF"])
def test_i8_precision_envelope(engine, golden_dir, kind):
    """The default arithmetic over the whole conditioning range the reference's API can reach
    (likelihood variance 1e-2 .. GPflow's floor 1e-6, gpflow_models.py:404-411; kernel variance 0.05 .. 10;
    n 1000 .. 4000; 10 % exact duplicate rows) against the extended-precision oracle
    (tests/golden/precision_envelope.npz <- oracle/make_envelope_golden.py <- oracle/longdouble_gpr.c).
    Tolerances: the north star's 1e-8 on LML and predictions everywhere.  Gradient (relative to its largest
    component): 1e-7 up to rho = (var + noise) / noise = 5e5 and 2e-13 * rho above - at rho = 1e7 (noise at
    GPflow's floor, variance 10) LAPACK float64 itself is off by up to 5.9e-7 and FP64 mma.sync by 9.3e-7
    (profiles/r02_precision_envelope.md), so 1e-7 is not attainable by any float64 implementation there."""
    import envelope as env
    gold = np.load(os.path.join(golden_dir, "precision_envelope.npz"))
    cs = [c for c in env.cases() if c["kind"] == kind]
    items = [env.make_case(c) for c in cs]
    lml, grad, st, p, off = _envelope_run(engine, cs, items, kind)
    bad = []
    for k, c in enumerate(cs):
        ref = {q: gold[f"{c['name']}/{q}"] for q in ("lml", "grad", "fmean", "fvar")}
        ref["lml"] = float(ref["lml"])
        sl = slice(off[k], off[k + 1])
        e = env.rel_errors(lml[k], grad[k], p["f*"][sl], p["f*_var"][sl], ref, items[k][3])
        rho = (c["var"] + c["noise"]) / c["noise"]
        tol = dict(lml=1e-8, grad=max(1e-7, 2e-13 * rho), fmean=1e-8, fvar=1e-8)
        if st[k] != 0 or any(not (e[q] <= tol[q]) for q in tol):
            bad.append((c["name"], int(st[k]), e, tol))
    assert not bad, bad


@pytest.mark.parametrize("kind", ["Matern32", "RBF", "Matern12", "Matern52"])
def test_i8_low_tier_envelope(engine, golden_dir, kind):
    """Low-rho tier of the precision rule (slot_set_theta in engine.cu: while rho = (var + noise) / noise <= 32 - 12 for
    the squared-exponential kernel - the int8-sliced k-sums of the sweep drop their last slice-pair group: 28 instead of
    36 MMAs per 32-wide slab) against extended precision at rho = 11 and 31.3 (just below the threshold) and 51
    (outside: tier on == tier off), n = 1000 and 2000, 10 % duplicate rows, all four kernels: held to a TENTH of the
    north star's tolerances (1e-9 on LML and predictions, 1e-8 on the gradient), so that the tier keeps a margin of 10
    at its own edge.  The same cases with the tier switched off are written beside them to
    gpurun_out/low_tier_envelope.json (profiles/r02_low_tier_envelope.md)."""
    import json
    import os
    import envelope as env
    gold = np.load(os.path.join(golden_dir, "low_tier_envelope.npz"))
    cs = [c for c in env.low_tier_cases() if c["kind"] == kind]
    items = [env.make_case(c) for c in cs]
    runs = {}
    try:
        for name, tier in (("tier_on", True), ("tier_off", False)):
            engine.set_precision_low_tier(tier)
            runs[name] = _envelope_run(engine, cs, items, kind)
    finally:
        engine.set_precision_low_tier(True)
    rows, bad = [], []
    for k, c in enumerate(cs):
        ref = {q: gold[f"{c['name']}/{q}"] for q in ("lml", "grad", "fmean", "fvar")}
        ref["lml"] = float(ref["lml"])
        row = dict(case=c["name"], rho=(c["var"] + c["noise"]) / c["noise"])
        for name, (lml, grad, st, p, off) in runs.items():
            sl = slice(off[k], off[k + 1])
            row[name] = env.rel_errors(lml[k], grad[k], p["f*"][sl], p["f*_var"][sl], ref, items[k][3])
            row[name]["status"] = int(st[k])
        rows.append(row)
        e = row["tier_on"]
        tol = dict(lml=1e-9, grad=1e-8, fmean=1e-9, fvar=1e-9)
        if e["status"] != 0 or any(not (e[q] <= tol[q]) for q in tol):
            bad.append(row)
    os.makedirs("gpurun_out", exist_ok=True)
    path = os.path.join("gpurun_out", "low_tier_envelope.json")
    prev = json.load(open(path)) if os.path.exists(path) else []
    json.dump([r for r in prev if not r["case"].startswith(f"low-{kind}-")] + rows, open(path, "w"), indent=1)
    assert not bad, bad


def test_reference_ill_conditioned_smoke(engine, golden_dir):
    """The reference's own numerical-stability smoke (test_numerical_stability.py:49-91): RBF on a 20 x 20 grid in
    unscaled metre coordinates with lengthscales 50 000.  The reference only checks that the optimisation does not
    raise; here LML / gradient / predictions at its starting point also match extended precision, and the fit from
    that point ends with status OK."""
    import envelope as env
    from gpsat_b200.engine import RaggedBatch
    gold = np.load(os.path.join(golden_dir, "precision_envelope.npz"))
    X, y, Xs, theta = env.smoke_case()
    b = RaggedBatch.from_lists("RBF", [X], [y], [Xs])
    lml, grad, st = engine.lml_grad(b, theta[None, :])
    p = engine.predict(b, theta[None, :])
    ref = dict(lml=float(gold["smoke/lml"]), grad=gold["smoke/grad"], fmean=gold["smoke/fmean"], fvar=gold["smoke/fvar"])
    e = env.rel_errors(lml[0], grad[0], p["f*"], p["f*_var"], ref, theta)
    assert st[0] == 0 and e["lml"] <= 1e-8 and e["grad"] <= 1e-7 and e["fmean"] <= 1e-8 and e["fvar"] <= 1e-8, e
    tk = np.zeros((1, 4), np.int32)
    ta = np.zeros((1, 4)); ta[0, 3] = 1e-6
    res = engine.fit(b, theta[None, :], tk, ta, np.zeros((1, 4)), np.ones((1, 4), np.uint8))
    assert res.status[0] == 0 and np.isfinite(res.lml[0]) and res.lml[0] >= lml[0]


def test_i8_timeout_fails_the_call(engine):
    """A barrier wait of the int8 pipeline that times out must fail the product call (non-zero return code ->
    RuntimeError through the binding), not hand back unfinished tiles with status OK."""
    rng = np.random.default_rng(5)
    experts = [_expert(rng, n, 3, m=9) for n in (300, 520)]
    b = _batch("Matern32", experts, with_pred=True)
    theta = np.stack([_theta(rng, 3) for _ in experts])
    before = engine.i8_timeouts()
    th0 = np.ones((2, 5))
    tk = np.zeros((2, 5), np.int32)
    ta = np.zeros((2, 5)); ta[:, 4] = 1e-6
    try:
        engine.debug_set_i8_spin_limit(0)          # every wait completes, then reports a time-out
        with pytest.raises(RuntimeError, match="timed out"):
            engine.lml_grad(b, theta)
        with pytest.raises(RuntimeError, match="timed out"):
            engine.predict(b, theta)
        with pytest.raises(RuntimeError, match="timed out"):
            engine.fit(b, th0, tk, ta, np.zeros((2, 5)), np.ones((2, 5), np.uint8))
    finally:
        engine.debug_set_i8_spin_limit(1 << 22)
    assert engine.i8_timeouts() == before + 3
    lml, grad, st = engine.lml_grad(b, theta)      # the flag was cleared: the next call is clean
    assert np.all(st == 0) and engine.i8_timeouts() == before + 3
    for e, (X, y, _) in enumerate(experts):
        v, _ = gpr.lml_and_grad("Matern32", X, y, theta[e])
        assert abs(lml[e] - v) <= RTOL_FIXED * abs(v)


def test_low_tier_guard_keeps_the_search_path(engine):
    """During a fit the low-rho tier is offered only while the line search expects a decrease above 1e-5 |f|
    (lbfgs_kernel in engine.cu): its noise of ~1e-10 |f| must not cost trial steps in the last iterations.  48 c2-shaped
    experts (rho = 3 .. 35 at the optimum: all inside the tier), optimised with the tier on and off: same status, the
    optimised LML equal to 1e-6 (the north star asks 1e-4 of an optimum), and the number of evaluations per expert the same up to the +-few steps any change of
    rounding causes on a flat likelihood (without the guard single experts needed 20+ evaluations more)."""
    import dataclasses
    from gpsat_b200 import synthetic as syn
    spec = dataclasses.replace(syn.WORKLOADS["c2"], E=48, n_lo=300, n_hi=700, m_lo=5, m_hi=9)
    batch = syn.make_batch(spec, seed=3)
    th0, tk, ta, tb, tr = syn.default_fit_setup(spec, batch.E)
    res = {}
    try:
        for tier in (True, False):
            engine.set_precision_low_tier(tier)
            res[tier] = engine.fit(batch, th0, tk, ta, tb, tr)
    finally:
        engine.set_precision_low_tier(True)
    a, b = res[True], res[False]
    np.testing.assert_array_equal(a.status, b.status)
    assert np.all(a.status == 0)
    np.testing.assert_allclose(a.lml, b.lml, rtol=1e-6)      # optimised values: SciPy's ftol stop decides the last digits
    dn = np.abs(a.n_evals.astype(int) - b.n_evals.astype(int))
    assert dn.max() <= 6 and dn.mean() <= 0.5, (dn.max(), dn.mean(), a.n_evals, b.n_evals)
    rho = (a.theta[:, 3] + a.theta[:, 4]) / a.theta[:, 4]
    assert np.all(rho <= 64)                                  # the experts really live where the tier applies
