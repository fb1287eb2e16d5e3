"""Parity of the CUDA sparse-GPR path (C ABI: gpsat_sgpr_elbo / gpsat_sgpr_predict) with oracle/sgpr.py.

Kuu carries GPflow's jitter of 1e-6, so its condition number reaches ~1e6 and forward errors of
1e-16 * cond are expected on either side: the tolerance at fixed hyper-parameters is 1e-7 relative
(with an absolute floor scaled by the kernel variance), looser than the exact-GPR 1e-8.
"""
import numpy as np
import pytest

from oracle import gpr, sgpr

pytestmark = pytest.mark.gpu

RTOL = 1e-7


def _expert(rng, n, D, M, m):
    X = rng.uniform(-4, 4, (n, D))
    K = gpr.kernel_matrix("Matern32", X, None, [2.0, 1.5, 3.0][:D], 0.7)
    y = np.linalg.cholesky(K + 1e-8 * np.eye(n)) @ rng.standard_normal(n) + 0.2 * rng.standard_normal(n)
    Z = X[rng.choice(n, size=M, replace=False)] if M <= n else rng.uniform(-4, 4, (M, D))
    return X, y, rng.uniform(-4, 4, (m, D)), Z


def _theta(rng, D):
    return np.concatenate([rng.uniform(0.8, 2.5, D), [rng.uniform(0.3, 1.5)], [rng.uniform(0.02, 0.3)]])


def _batch(kind, experts):
    from gpsat_b200.engine import RaggedBatch
    return RaggedBatch.from_lists(kind, [e[0] for e in experts], [e[1] for e in experts], [e[2] for e in experts])


@pytest.mark.parametrize("kind", gpr.KERNELS)
@pytest.mark.parametrize("D", [1, 3])
def test_elbo_predict_ragged(engine, kind, D):
    rng = np.random.default_rng(7 * D + len(kind))
    shapes = [(300, 40, 17), (130, 127, 0), (700, 128, 140), (257, 129, 5), (64, 5, 3), (1000, 300, 260)]
    experts = [_expert(rng, n, D, M, m) for n, M, m in shapes]
    theta = np.stack([_theta(rng, D) for _ in experts])
    b = _batch(kind, experts)
    Z = [e[3] for e in experts]
    elbo, _, status = engine.sgpr_elbo(b, Z, theta)
    out = engine.sgpr_predict(b, Z, theta)
    assert np.all(status == 0) and np.all(out["status"] == 0)
    np.testing.assert_array_equal(out["elbo"], elbo)          # same factorisation, same bits
    for e, (X, y, Xs, Ze) in enumerate(experts):
        ref = sgpr.elbo(kind, X, y, Ze, theta[e])
        assert abs(elbo[e] - ref) <= RTOL * abs(ref), (e, elbo[e], ref)
        if len(Xs):
            p = sgpr.predict(kind, X, y, Ze, theta[e], Xs)
            sl = slice(int(b.off_m[e]), int(b.off_m[e + 1]))
            atol = RTOL * theta[e][D]
            np.testing.assert_allclose(out["f*"][sl], p["f*"], rtol=RTOL, atol=atol * max(1.0, np.abs(y).max()))
            np.testing.assert_allclose(out["f*_var"][sl], p["f*_var"], rtol=RTOL, atol=atol)
            np.testing.assert_allclose(out["y_var"][sl], p["y_var"], rtol=RTOL, atol=atol)


def test_inducing_at_data_equals_exact_gpr(engine):
    """Z = X makes the bound tight: ELBO -> LML, predictions -> exact GPR, up to the 1e-6 jitter
    (the reference's own SGPR known-answer test runs in this regime, tests/test_localexperts.py:229-251)."""
    rng = np.random.default_rng(3)
    X, y, Xs, _ = _expert(rng, 150, 2, 10, 30)
    theta = np.array([[1.7, 1.2, 0.8, 0.1]])
    b = _batch("Matern52", [(X, y, Xs)])
    out = engine.sgpr_predict(b, [X], theta)
    lml = gpr.lml("Matern52", X, y, theta[0])
    assert abs(out["elbo"][0] - lml) < 1e-4 * abs(lml)
    ref = gpr.predict("Matern52", X, y, theta[0], Xs)
    np.testing.assert_allclose(out["f*"], ref["f*"], rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(out["f*_var"], ref["f*_var"], rtol=1e-4, atol=1e-5)


def test_kat_sgpr(engine, golden_dir):
    """The reference's SGPR known-answer test (tests/test_localexperts.py:229-251): M = N = 50,
    Matern32, at the golden length scale the prediction matches the sklearn values to 1e-4."""
    import os
    g = np.load(os.path.join(golden_dir, "kat_localexperts.npz"))
    X, y = g["x_train"], g["y_train"][:, 0]
    theta = np.array([[float(g["ls"]), 1.0, float(g["eps"]) ** 2]])
    b = _batch("Matern32", [(X, y, g["x_test"])])
    out = engine.sgpr_predict(b, [X.copy()], theta)
    assert abs(out["f*"][0] - g["pred_mean"][0]) < 1e-4
    assert abs(out["f*_var"][0] - g["pred_std"][0] ** 2) < 1e-4
    ref = sgpr.predict("Matern32", X, y, X.copy(), theta[0], g["x_test"])
    np.testing.assert_allclose(out["f*"], ref["f*"], rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(out["f*_var"], ref["f*_var"], rtol=1e-6, atol=1e-7)
    e_ref = sgpr.elbo("Matern32", X, y, X.copy(), theta[0])
    assert abs(out["elbo"][0] - e_ref) < 1e-6 * abs(e_ref)


def test_not_pd_is_per_expert(engine):
    rng = np.random.default_rng(11)
    ok = _expert(rng, 200, 2, 30, 4)
    bad = _expert(rng, 200, 2, 30, 4)
    Zbad = np.repeat(bad[3][:1], 30, axis=0)            # 30 identical inducing points, RBF: rank 1 + jitter
    theta = np.array([[1.5, 1.5, 1.0, 0.1], [1.5, 1.5, 1e8, 0.1]])
    b = _batch("RBF", [ok[:3], bad[:3]])
    elbo, _, status = engine.sgpr_elbo(b, [ok[3], Zbad], theta)
    assert status[0] == 0 and np.isfinite(elbo[0])
    ref = sgpr.elbo("RBF", ok[0], ok[1], ok[3], theta[0])
    assert abs(elbo[0] - ref) <= RTOL * abs(ref)
    assert status[1] in (0, 1)
    if status[1] == 1:
        assert elbo[1] == -np.inf


def test_device_memory_batch(engine):
    torch = pytest.importorskip("torch")
    rng = np.random.default_rng(5)
    experts = [_expert(rng, n, 3, M, m) for n, M, m in [(200, 50, 9), (333, 64, 20)]]
    theta = np.stack([_theta(rng, 3) for _ in experts])
    b = _batch("Matern32", experts)
    Z = [e[3] for e in experts]
    host = engine.sgpr_predict(b, Z, theta)
    dev = engine.sgpr_predict(b.to_device("cuda:0"), Z, theta)
    for k in ("f*", "f*_var", "y_var", "elbo"):
        np.testing.assert_array_equal(dev[k].cpu().numpy(), host[k])


@pytest.mark.parametrize("kind", gpr.KERNELS)
@pytest.mark.parametrize("D", [1, 3])
def test_elbo_gradient_ragged(engine, kind, D):
    """d elbo / d theta against the oracle's analytic gradient (itself checked by finite differences
    in tests/test_oracle_golden.py)."""
    rng = np.random.default_rng(11 * D + len(kind))
    # many inducing points on a line make Kuu + 1e-6 I ill-conditioned (cond ~ 1e8), and then the
    # oracle's own gradient moves by 1e-3 under a 1e-15 perturbation of its inputs: keep M small in 1-D
    shapes = ([(300, 12), (130, 20), (700, 16), (257, 24), (64, 5), (900, 18)] if D == 1 else
              [(300, 40), (130, 127), (700, 128), (257, 129), (64, 5), (900, 300)])
    experts = [_expert(rng, n, D, M, 0) for n, M in shapes]
    theta = np.stack([_theta(rng, D) for _ in experts])
    b = _batch(kind, experts)
    Z = [e[3] for e in experts]
    elbo, grad, status = engine.sgpr_elbo(b, Z, theta, with_grad=True)
    assert np.all(status == 0)
    for e, (X, y, _, Ze) in enumerate(experts):
        v, g = sgpr.elbo_and_grad(kind, X, y, Ze, theta[e])
        assert abs(elbo[e] - v) <= RTOL * abs(v)
        # round-off sensitivity of the oracle itself (inputs perturbed in the last bit)
        _, g2 = sgpr.elbo_and_grad(kind, X * (1 + 1e-15), y, Ze * (1 + 1e-15), theta[e])
        atol = max(1e-6 * np.abs(g).max(), 100 * np.abs(g - g2).max())
        assert atol < 1e-3 * np.abs(g).max(), "test case too ill-conditioned to be meaningful"
        np.testing.assert_allclose(grad[e], g, rtol=1e-6, atol=atol)


def test_fit_matches_oracle(engine):
    """Optimised hyper-parameters and ELBO within 1e-4 of SciPy L-BFGS-B on the oracle's -elbo
    (the north_star tolerance for optimised quantities), experts of different sizes in one call."""
    rng = np.random.default_rng(21)
    D = 2
    experts = [_expert(rng, n, D, M, 0) for n, M in [(400, 60), (250, 40), (600, 130)]]
    E, P = len(experts), D + 2
    tr = gpr.Transform.gpflow_default(D)
    theta0 = np.tile(np.array([1.0, 1.0, 1.0, 1.0]), (E, 1))
    b = _batch("Matern32", experts)
    Z = [e[3] for e in experts]
    res = engine.fit(b, theta0, np.tile(tr.kind, (E, 1)), np.tile(tr.a, (E, 1)), np.tile(tr.b, (E, 1)),
                     np.ones((E, P), np.uint8), inducing=Z)
    assert np.all(res.status == 0)
    for e, (X, y, _, Ze) in enumerate(experts):
        th, val, ok, nfev = sgpr.fit("Matern32", X, y, Ze, theta0[e], tr)
        assert ok
        assert abs(res.lml[e] - val) <= 1e-4 * abs(val), (res.lml[e], val)
        np.testing.assert_allclose(res.theta[e], th, rtol=2e-3)
        assert abs(res.objective[e] + res.lml[e]) < 1e-12 * abs(val)


def test_sgpr_model_driver_sequence(engine, golden_dir):
    """B200SGPRModel through the call order of LocalExpertOI.run (local_experts.py:1043-1164) on the
    reference's SGPR known-answer case (tests/test_localexperts.py:229-251): lengthscale, ELBO-side
    prediction within the reference's 1e-4 of the sklearn values."""
    import os
    from gpsat_b200.models import B200SGPRModel
    g = np.load(os.path.join(golden_dir, "kat_localexperts.npz"))
    eps = float(g["eps"])
    # The inducing points are the first rows of a shuffle drawn from numpy's global generator, as in the reference
    # (gpflow_models.py:816-819).  With M = N = 50 and the 1e-6 jitter this ELBO is flat to rounding at the optimum,
    # and whether L-BFGS ends with "converged" or "abnormal line search" there depends on the summation order, i.e.
    # on the shuffle (about 5 of 24 orders end abnormal, with either exp implementation; the lengthscale is right in
    # all of them).  Fix the order so that the reference's own `assert result` is reproducible.
    np.random.seed(0)
    m = B200SGPRModel(coords=g["x_train"].copy(), obs=g["y_train"].copy(), kernel="Matern32",
                      num_inducing_points=50, noise_variance=eps ** 2, verbose=False)
    assert m.param_names == ["lengthscales", "kernel_variance", "likelihood_variance", "inducing_points"]
    assert m.get_inducing_points().shape == (50, 1)
    m.set_parameter_constraints({"lengthscales": {"low": float(g["constraint_low"]),
                                                  "high": float(g["constraint_high"])}},
                                move_within_tol=True, tol=1e-2)
    ok = m.optimise_parameters(fixed_params=["kernel_variance", "likelihood_variance"])
    assert ok
    assert abs(m.get_lengthscales()[0] - float(g["ls"])) < 1e-4
    th = np.concatenate([m.get_lengthscales(), [m.get_kernel_variance(), m.get_likelihood_variance()]])
    ref = sgpr.elbo("Matern32", g["x_train"], g["y_train"][:, 0], m.get_inducing_points(), th)
    assert abs(m.get_objective_function_value() - ref) < 1e-6 * abs(ref)         # +ELBO, not negated
    p = m.predict(g["x_test"])
    assert set(p) == {"f*", "f*_var", "y_var", "f_bar"}
    assert abs(p["f*"][0] - g["pred_mean"][0]) < 1e-4
    assert abs(p["f*_var"][0] - g["pred_std"][0] ** 2) < 1e-4


def test_batched_helper_equals_loop(engine):
    """fit_predict_models on sparse models == the per-model loop (same kernels, same order of sums)."""
    import pandas as pd
    from gpsat_b200.models import B200SGPRModel, fit_predict_models
    rng = np.random.default_rng(4)

    def build(seed):
        r = np.random.default_rng(seed)
        n = 300 + 40 * seed
        xy = r.uniform(-2, 2, (n, 2))
        df = pd.DataFrame({"x": xy[:, 0], "y": xy[:, 1],
                           "obs": np.sin(2 * xy[:, 0]) + 0.3 * xy[:, 1] + 0.1 * r.standard_normal(n)})
        np.random.seed(seed)
        return B200SGPRModel(data=df, obs_col="obs", coords_col=["x", "y"], obs_mean="local",
                             num_inducing_points=50, verbose=False)

    pred = rng.uniform(-1.8, 1.8, (25, 2))
    loop = [build(s) for s in (1, 2, 3)]
    loop_ok = [m.optimise_parameters() for m in loop]
    loop_obj = [m.get_objective_function_value() for m in loop]
    loop_pred = [m.predict(pred) for m in loop]
    batch = [build(s) for s in (1, 2, 3)]
    ok_b, obj_b, pred_b = fit_predict_models(batch, [pred] * 3)
    assert list(ok_b) == loop_ok
    for e in range(3):
        np.testing.assert_array_equal(batch[e].get_inducing_points(), loop[e].get_inducing_points())
        np.testing.assert_allclose(batch[e]._theta, loop[e]._theta, rtol=1e-12)
        assert abs(obj_b[e] - loop_obj[e]) <= 1e-12 * abs(loop_obj[e])
        np.testing.assert_allclose(pred_b[e]["f*"], loop_pred[e]["f*"], rtol=1e-12, atol=1e-14)
        np.testing.assert_allclose(pred_b[e]["f*_var"], loop_pred[e]["f*_var"], rtol=1e-12, atol=1e-14)


def test_full_size_c5_expert(engine):
    """One expert at the c5 shape (n = 20 000, M = 500, 3-D): ELBO, gradient and predictions against
    the oracle (a few seconds of LAPACK on the host)."""
    from gpsat_b200 import synthetic as syn
    spec = syn.WorkloadSpec("c5", 1, 20000, 20000, 300, 300, 500e3, 15, (25e3, 25e3, 1.0), (600e3, 600e3, 30.0))
    X, y, Xs = syn.make_expert(spec, 0, 5)
    Z = X[np.random.default_rng(1).permutation(len(X))[:500]].copy()
    theta = np.array([[4.0, 4.0, 5.0, 0.05, 0.005]])
    b = _batch("Matern32", [(X, y, Xs)])
    elbo, grad, status = engine.sgpr_elbo(b, [Z], theta, with_grad=True)
    v, g = sgpr.elbo_and_grad("Matern32", X, y, Z, theta[0])
    assert status[0] == 0 and abs(elbo[0] - v) <= 1e-7 * abs(v)
    _, g2 = sgpr.elbo_and_grad("Matern32", X * (1 + 1e-15), y, Z * (1 + 1e-15), theta[0])
    atol = max(1e-6 * np.abs(g).max(), 100 * np.abs(g - g2).max())
    np.testing.assert_allclose(grad[0], g, rtol=1e-6, atol=atol)
    out = engine.sgpr_predict(b, [Z], theta)
    p = sgpr.predict("Matern32", X, y, Z, theta[0], Xs)
    np.testing.assert_allclose(out["f*"], p["f*"], rtol=1e-6, atol=1e-8)
    np.testing.assert_allclose(out["f*_var"], p["f*_var"], rtol=1e-6, atol=1e-9)


def test_sgpr_against_extended_precision(engine, golden_dir):
    """ELBO / gradient / predictions of the sparse model against extended precision (oracle/longdouble_gpr.c ld_sgpr;
    gradient = central differences of that ELBO; goldens: oracle/make_sgpr_envelope_golden.py), on the c5 shape
    (n = 20 000, M = 500) and on three smaller cases with cond(Kuu + 1e-6 I) between 1e3 and 3e7.  Measured on B200
    (profiles/r02_sgpr_envelope.md): the CUDA path is within 4e-13 (ELBO), 5e-10 (gradient) and 8e-13 (predictions) of
    extended precision, so it is held to the north star's 1e-8 on all four quantities here.  The looser 1e-7 / 1e-6
    of the other tests in this file is the float64 NumPy oracle's own error (its gradient is off by 5.7e-6 on the
    RBF case, recorded in the goldens as f64_err), not the CUDA path's."""
    import json
    import os
    import envelope as env
    from gpsat_b200.engine import RaggedBatch
    gold = np.load(os.path.join(golden_dir, "sgpr_envelope.npz"))
    rows, bad = [], []
    for c in env.SGPR_CASES:
        X, y, Z, Xs, theta = env.make_sgpr_case(c)
        b = RaggedBatch.from_lists(c["kind"], [X], [y], [Xs])
        v, g, st = engine.sgpr_elbo(b, [Z], theta[None, :], with_grad=True)
        p = engine.sgpr_predict(b, [Z], theta[None, :])
        ref = {k: gold[f"{c['name']}/{k}"] for k in ("elbo", "grad", "fmean", "fvar", "f64_err", "cond_kuu")}
        err = np.array([abs(v[0] - float(ref["elbo"])) / abs(float(ref["elbo"])),
                        np.max(np.abs(g[0] - ref["grad"])) / np.max(np.abs(ref["grad"])),
                        np.max(np.abs(p["f*"] - ref["fmean"])) / np.max(np.abs(ref["fmean"])),
                        np.max(np.abs(p["f*_var"] - ref["fvar"])) / theta[3]])
        tol = np.full(4, 1e-8)
        rows.append(dict(case=c["name"], cond_kuu=float(ref["cond_kuu"]), gpu=err.tolist(), f64_oracle=ref["f64_err"].tolist(),
                         tol=tol.tolist(), status=int(st[0])))
        if st[0] != 0 or np.any(err > tol):
            bad.append(rows[-1])
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(rows, open(os.path.join("gpurun_out", "sgpr_envelope.json"), "w"), indent=1)
    assert not bad, bad


@pytest.mark.parametrize("kind", ["Matern32", "RBF"])
def test_sgpr_full_cov(engine, kind):
    """gpsat_sgpr_predict_cov (SGPR.predict_f(full_cov=True), gpflow_models.py:230, 245-263): the m x m blocks against
    the oracle, symmetric, diagonal = f*_var, ragged m including 0 and sizes around the 64-wide output tile; and through
    B200SGPRModel.predict(full_cov=True) with the exact model's keys."""
    from gpsat_b200.models import B200SGPRModel
    rng = np.random.default_rng(31 + len(kind))
    shapes = [(300, 40, 17), (500, 128, 64), (257, 129, 65), (200, 30, 0), (700, 200, 130)]
    experts = [_expert(rng, n, 3, M, m) for n, M, m in shapes]
    theta = np.stack([_theta(rng, 3) for _ in experts])
    b = _batch(kind, experts)
    out = engine.sgpr_predict(b, [e[3] for e in experts], theta, full_cov=True)
    ref_plain = engine.sgpr_predict(b, [e[3] for e in experts], theta)
    np.testing.assert_array_equal(out["f*"], ref_plain["f*"])
    off = b.off_m
    for e, (X, y, Xs, Z) in enumerate(experts):
        cov = out["f*_cov"][e]
        assert cov.shape == (len(Xs), len(Xs))
        if len(Xs) == 0:
            continue
        p = sgpr.predict(kind, X, y, Z, theta[e], Xs, full_cov=True)
        np.testing.assert_allclose(cov, p["f*_cov"], rtol=RTOL, atol=RTOL * theta[e][3])
        np.testing.assert_array_equal(cov, cov.T)
        np.testing.assert_allclose(np.diag(cov), out["f*_var"][off[e]:off[e + 1]], rtol=1e-10, atol=1e-12 * theta[e][3])
    X, y, Xs, Z = experts[0]
    m = B200SGPRModel(coords=X, obs=y[:, None], kernel=kind, num_inducing_points=len(Z), verbose=False)
    m.set_inducing_points(Z)
    m.set_parameters(lengthscales=theta[0][:3], kernel_variance=theta[0][3], likelihood_variance=theta[0][4])
    res = m.predict(Xs, full_cov=True)
    assert set(res) >= {"f*", "f*_var", "y_var", "f*_cov", "y_cov", "f_bar"}
    np.testing.assert_allclose(res["f*_cov"], out["f*_cov"][0], rtol=1e-12, atol=1e-14)
    np.testing.assert_allclose(res["y_cov"], res["f*_cov"] + theta[0][4] * np.eye(len(Xs)), rtol=1e-12, atol=1e-14)


@pytest.mark.parametrize("kind", gpr.KERNELS)
def test_inducing_point_gradient(engine, kind):
    """gpsat_sgpr_elbo_gradz: d elbo / d Z, what TensorFlow differentiates once train_inducing_points=True has made
    inducing_variable.Z trainable (gpflow_models.py:896-897).  Against the oracle's analytic gradient (itself checked
    by central differences in tests/test_oracle_golden.py) on ragged experts - M around the 128-wide tiles, Z off the
    data, duplicated inducing points - and against central differences of the CUDA ELBO itself; the ELBO and the theta
    gradient of the same call are bit-identical to gpsat_sgpr_elbo's."""
    rng = np.random.default_rng(5 + len(kind))
    shapes = [(300, 40), (130, 127), (700, 128), (257, 129), (64, 5), (1000, 300)]
    experts = [_expert(rng, n, 3, M, 0) for n, M in shapes]
    experts = [(X, y, Xs, Z + 0.05 * rng.standard_normal(Z.shape)) for X, y, Xs, Z in experts]
    experts[0][3][7] = experts[0][3][3]                       # two inducing points at the same place
    theta = np.stack([_theta(rng, 3) for _ in experts])
    b = _batch(kind, experts)
    Z = [e[3] for e in experts]
    elbo, grad, gz, status = engine.sgpr_elbo(b, Z, theta, with_z=True)
    elbo0, grad0, _ = engine.sgpr_elbo(b, Z, theta, with_grad=True)
    assert np.all(status == 0)
    np.testing.assert_array_equal(elbo, elbo0)
    np.testing.assert_array_equal(grad, grad0)
    off = np.concatenate([[0], np.cumsum([len(z) for z in Z])])
    assert gz.shape == (off[-1], 3)
    for e, (X, y, _, Ze) in enumerate(experts):
        v, g, ref = sgpr.elbo_and_grad(kind, X, y, Ze, theta[e], with_z=True)
        got = gz[off[e]:off[e + 1]]
        np.testing.assert_allclose(got, ref, rtol=1e-6, atol=1e-6 * np.abs(ref).max())
    # central differences of the CUDA ELBO on a few coordinates of expert 2
    X, y, _, Ze = experts[2]
    b1 = _batch(kind, [experts[2]])
    for k, d in ((0, 0), (17, 1), (127, 2)):
        h = 1e-5
        Zp, Zm = Ze.copy(), Ze.copy()
        Zp[k, d] += h
        Zm[k, d] -= h
        fp, _, _ = engine.sgpr_elbo(b1, [Zp], theta[2:3])
        fm, _, _ = engine.sgpr_elbo(b1, [Zm], theta[2:3])
        fd = (fp[0] - fm[0]) / (2 * h)
        ref = gz[off[2] + k, d]
        assert abs(fd - ref) <= 2e-5 * max(1.0, np.abs(gz[off[2]:off[3]]).max()), (k, d, fd, ref)


def test_train_inducing_points(engine):
    """B200SGPRModel.optimise_parameters(train_inducing_points=True) (gpflow_models.py:864-899): the search over
    theta and Z reaches a higher ELBO than the fixed-Z search from the same start, moves Z, and agrees with the
    oracle driven the same way (SciPy L-BFGS-B over [xi, Z] with the NumPy ELBO and gradients) on the objective."""
    from gpsat_b200.models import B200SGPRModel
    rng = np.random.default_rng(11)
    n, M = 400, 12
    X = rng.uniform(-3, 3, (n, 2))
    y = np.sin(1.5 * X[:, 0]) * np.cos(X[:, 1]) + 0.1 * rng.standard_normal(n)

    def build():
        np.random.seed(4)
        return B200SGPRModel(coords=X, obs=y[:, None], kernel="Matern52", num_inducing_points=M, verbose=False)

    fixed = build()
    assert fixed.optimise_parameters()
    elbo_fixed = fixed.get_objective_function_value()
    m = build()
    Z0, th0 = m.get_inducing_points(), m._theta.copy()
    ok = m.optimise_parameters(train_inducing_points=True)
    elbo_z = m.get_objective_function_value()
    assert m.opt_info["n_evals"] > 5
    assert elbo_z > elbo_fixed + 1.0, (elbo_z, elbo_fixed)
    assert np.abs(m.get_inducing_points() - Z0).max() > 1e-2
    assert abs(-m.opt_info["objective"] - elbo_z) <= 1e-9 * abs(elbo_z)
    tho, Zo, vo, oko, nfev = sgpr.fit_z("Matern52", X, y, Z0, th0)
    assert ok == oko or abs(elbo_z - vo) <= 1e-3 * abs(vo)
    assert abs(elbo_z - vo) <= 2e-3 * abs(vo), (elbo_z, vo, m.opt_info, nfev)
    p = m.predict(X[:10])
    q = sgpr.predict("Matern52", X, y, m.get_inducing_points(), m._theta, X[:10])
    np.testing.assert_allclose(p["f*"], q["f*"], rtol=1e-6, atol=1e-8)
