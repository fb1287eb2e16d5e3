"""The oracle against the reference's own known answers (CPU only).

Sources of the golden files: oracle/make_golden.py.
"""
import os

import numpy as np
import pytest

from oracle import gpr


def test_kat_localexperts(golden_dir):
    """/root/reference/tests/test_localexperts.py:204-227 (test_gpflow_gpr), tolerance 1e-6 there."""
    g = np.load(os.path.join(golden_dir, "kat_localexperts.npz"))
    X, y = g["x_train"], g["y_train"][:, 0]
    eps = float(g["eps"])
    tr = gpr.Transform.gpflow_default(1)
    tr.set_sigmoid(0, float(g["constraint_low"]), float(g["constraint_high"]))
    theta0 = np.array([1.0, 1.0, eps ** 2])
    th, l, ok, nev = gpr.fit("Matern32", X, y, theta0, tr, trainable=[True, False, False])
    assert ok
    assert abs(th[0] - float(g["ls"])) < 1e-6
    assert abs(l - float(g["ml"])) < 1e-6
    p = gpr.predict("Matern32", X, y, th, g["x_test"])
    assert abs(p["f*"][0] - g["pred_mean"][0]) < 1e-6
    assert abs(p["f*_var"][0] - g["pred_std"][0] ** 2) < 1e-6
    assert abs(p["y_var"][0] - (p["f*_var"][0] + eps ** 2)) < 1e-15


def test_pure_python_gpr(golden_dir):
    """GPSat/models/pure_python_gpr.py SMLII_mod (nlZ) and GPR (Rasmussen Alg. 2.1), Matern-3/2 ARD."""
    g = np.load(os.path.join(golden_dir, "pure_python_gpr.npz"))
    X, y, Xs = g["X"], g["y"], g["Xs"]
    for k, th in enumerate(g["thetas"]):
        nlz = float(np.ravel(g["nlZ"][k])[0])
        assert abs(-gpr.lml("Matern32", X, y, th) - nlz) <= 1e-10 * abs(nlz)
        p = gpr.predict("Matern32", X, y, th, Xs)
        np.testing.assert_allclose(p["f*"], g["fmean"][k], rtol=1e-8, atol=1e-10)
        np.testing.assert_allclose(p["f*_var"], g["fvar"][k], rtol=1e-7, atol=1e-12)


@pytest.mark.parametrize("kind", gpr.KERNELS)
def test_sklearn_ard(golden_dir, kind):
    g = np.load(os.path.join(golden_dir, "sklearn_ard.npz"))
    X, y, Xs, th = g["X"], g["y"], g["Xs"], g["theta"]
    v, grad = gpr.lml_and_grad(kind, X, y, th)
    assert abs(v - float(g[f"{kind}_lml"])) <= 1e-10 * abs(v)
    np.testing.assert_allclose(grad[:3], g[f"{kind}_grad_ls"], rtol=1e-8)
    np.testing.assert_allclose(grad[3], g[f"{kind}_grad_var"], rtol=1e-8)
    p = gpr.predict(kind, X, y, th, Xs)
    np.testing.assert_allclose(p["f*"], g[f"{kind}_fmean"], rtol=1e-8, atol=1e-10)
    np.testing.assert_allclose(p["f*_var"], g[f"{kind}_fvar"], rtol=1e-7, atol=1e-12)


@pytest.mark.parametrize("kind", gpr.KERNELS)
@pytest.mark.parametrize("D", [1, 2, 3])
def test_analytic_gradient_fd(kind, D):
    rng = np.random.default_rng(D)
    n = 40
    X = rng.uniform(-3, 3, (n, D))
    y = rng.standard_normal(n)
    th = np.concatenate([rng.uniform(0.5, 2.5, D), [0.9, 0.2]])
    v, g = gpr.lml_and_grad(kind, X, y, th)
    for i in range(D + 2):
        h = 1e-6 * th[i]
        tp, tm = th.copy(), th.copy()
        tp[i] += h
        tm[i] -= h
        fd = (gpr.lml(kind, X, y, tp) - gpr.lml(kind, X, y, tm)) / (2 * h)
        assert abs(g[i] - fd) <= 2e-6 * max(1.0, abs(fd))


def test_gradient_matches_torch_autograd():
    """Independent check of the analytic gradient: reverse-mode autodiff through Cholesky, the way
    TensorFlow computes it for gpflow.optimizers.Scipy (gpflow_models.py:317-321)."""
    import torch
    rng = np.random.default_rng(5)
    n, D = 60, 3
    X = rng.uniform(-3, 3, (n, D))
    y = rng.standard_normal(n)
    th = np.array([1.4, 0.8, 2.2, 0.7, 0.15])
    v, g = gpr.lml_and_grad("Matern32", X, y, th)
    t = torch.tensor(th, dtype=torch.float64, requires_grad=True)
    Xt, yt = torch.tensor(X), torch.tensor(y)
    A = Xt / t[:D]
    r2 = ((A[:, None, :] - A[None, :, :]) ** 2).sum(-1)
    r = torch.sqrt(torch.clamp(r2, min=1e-36))
    K = t[D] * (1 + np.sqrt(3.0) * r) * torch.exp(-np.sqrt(3.0) * r) + t[D + 1] * torch.eye(n, dtype=torch.float64)
    L = torch.linalg.cholesky(K)
    a = torch.linalg.solve_triangular(L, yt[:, None], upper=False)[:, 0]
    lml = -0.5 * (a @ a) - torch.log(torch.diagonal(L)).sum() - 0.5 * n * np.log(2 * np.pi)
    lml.backward()
    assert abs(float(lml.detach()) - v) < 1e-10
    np.testing.assert_allclose(t.grad.numpy(), g, rtol=1e-9)


def test_gpflow_distance_mode_agrees():
    """|a|^2+|b|^2-2ab (what GPflow evaluates) vs direct differences: same LML to 1e-8 for the
    smooth kernels on GPSat-scaled coordinates."""
    rng = np.random.default_rng(9)
    X = rng.uniform(-6, 6, (200, 3))
    y = rng.standard_normal(200)
    th = np.array([2.0, 3.0, 4.0, 0.5, 0.05])
    for kind in ("Matern32", "Matern52", "RBF"):
        a, b = gpr.lml(kind, X, y, th, "direct"), gpr.lml(kind, X, y, th, "gpflow")
        assert abs(a - b) <= 1e-10 * abs(a)


def test_transforms_roundtrip():
    tr = gpr.Transform.gpflow_default(3)
    tr.set_sigmoid([0, 1], [1e-8, 0.5], [12.0, 9.0])
    th = np.array([1.0, 3.0, 2.0, 0.4, 0.003])
    xi = tr.inverse(th)
    np.testing.assert_allclose(tr.forward(xi), th, rtol=1e-14)
    h = 1e-6
    fd = (tr.forward(xi + h) - tr.forward(xi - h)) / (2 * h)
    np.testing.assert_allclose(tr.dtheta_dxi(xi), fd, rtol=1e-8)
    # likelihood variance lower bound 1e-6 (gpflow Gaussian.variance_lower_bound)
    assert tr.forward(np.array([0, 0, 0, 0, -800.0]))[4] == pytest.approx(1e-6)


def test_not_pd_is_minus_inf():
    X = np.zeros((5, 2))
    y = np.ones(5)
    assert gpr.lml("RBF", X, y, np.array([1.0, 1.0, 1.0, 0.0])) == -np.inf


def test_sgpr_kat_localexperts(golden_dir):
    """/root/reference/tests/test_localexperts.py:229-251 (test_gpflow_sgpr, M = N = 50), tol 1e-4 there."""
    from oracle import sgpr
    g = np.load(os.path.join(golden_dir, "kat_localexperts.npz"))
    X, y = g["x_train"], g["y_train"][:, 0]
    eps = float(g["eps"])
    tr = gpr.Transform.gpflow_default(1)
    tr.set_sigmoid(0, float(g["constraint_low"]), float(g["constraint_high"]))
    th, l, ok, nev = sgpr.fit("Matern32", X, y, X.copy(), np.array([1.0, 1.0, eps ** 2]), tr,
                              trainable=[True, False, False])
    assert ok
    assert abs(th[0] - float(g["ls"])) < 1e-4
    p = sgpr.predict("Matern32", X, y, X.copy(), th, g["x_test"])
    assert abs(p["f*"][0] - g["pred_mean"][0]) < 1e-4
    assert abs(p["f*_var"][0] - g["pred_std"][0] ** 2) < 1e-4


@pytest.mark.parametrize("kind", gpr.KERNELS)
def test_sgpr_gradient_fd_and_tight_bound(kind):
    from oracle import sgpr
    rng = np.random.default_rng(4)
    n, M, D = 120, 25, 3
    X = rng.uniform(-3, 3, (n, D))
    y = np.sin(X[:, 0]) + 0.3 * np.cos(X[:, 1]) + 0.1 * rng.standard_normal(n)
    Z = X[rng.permutation(n)[:M]].copy()
    th = np.array([1.3, 0.9, 2.0, 0.8, 0.05])
    v, g = sgpr.elbo_and_grad(kind, X, y, Z, th)
    for i in range(5):
        e = np.zeros(5)
        e[i] = 1e-6 * th[i]
        fd = (sgpr.elbo(kind, X, y, Z, th + e) - sgpr.elbo(kind, X, y, Z, th - e)) / (2 * e[i])
        assert abs(fd - g[i]) <= 1e-6 * max(1.0, abs(g[i]))
    # the bound never exceeds the exact log marginal likelihood and is tight (up to the jitter) for Z = X
    exact = gpr.lml(kind, X, y, th)
    assert v <= exact + 1e-9
    assert abs(sgpr.elbo(kind, X, y, X.copy(), th) - exact) < 1e-2


@pytest.mark.parametrize("kind", gpr.KERNELS)
def test_sgpr_inducing_point_gradient_fd(kind):
    """d elbo / d Z of oracle/sgpr.py (the gradient train_inducing_points=True exposes, gpflow_models.py:896-897)
    against central differences of the ELBO."""
    from oracle import sgpr
    rng = np.random.default_rng(len(kind))
    n, M, D = 120, 15, 3
    X = rng.uniform(-2, 2, (n, D))
    y = np.sin(X[:, 0]) + 0.1 * rng.standard_normal(n)
    Z = X[:M] + 0.05 * rng.standard_normal((M, D))
    th = np.array([1.0, 1.3, 0.8, 0.9, 0.05])
    v, g, gz = sgpr.elbo_and_grad(kind, X, y, Z, th, with_z=True)
    v2, g2 = sgpr.elbo_and_grad(kind, X, y, Z, th)
    assert v == v2 and np.array_equal(g, g2)
    for k in range(0, M, 4):
        for d in range(D):
            Zp, Zm = Z.copy(), Z.copy()
            Zp[k, d] += 1e-6
            Zm[k, d] -= 1e-6
            fd = (sgpr.elbo(kind, X, y, Zp, th) - sgpr.elbo(kind, X, y, Zm, th)) / 2e-6
            assert abs(fd - gz[k, d]) <= 1e-6 * max(1.0, np.abs(gz).max()), (k, d, fd, gz[k, d])


# ---- VFF (oracle/vff.py) ---------------------------------------------------------------------
def test_vff_features_match_reference_numpy(golden_dir):
    """make_Kuf_np / make_kvs_np and the static quantities of GPR_kron.__init__
    (/root/reference/GPSat/vff.py:518-559, 600-610), executed from the reference source by
    oracle/make_golden.py."""
    from oracle import vff
    g = np.load(os.path.join(golden_dir, "vff_features.npz"))
    ms = [g["m0"], g["m1"], g["m2"]]
    np.testing.assert_allclose(vff.kuf_np(g["X"][:, 0:1], g["a"][0], g["b"][0], ms[0]), g["Kuf0"], rtol=0, atol=1e-15)
    pre = vff.Precomputed(g["X"], g["Y"][:, 0], g["a"], g["b"], ms)
    assert pre.Ms == [7, 5, 3]
    np.testing.assert_allclose(pre.Kuf, g["Kuf"], rtol=0, atol=1e-15)
    np.testing.assert_allclose(pre.KufY, g["KufY"][:, 0], rtol=1e-13, atol=1e-13)
    np.testing.assert_allclose(pre.KufKfu, g["KufKfu"], rtol=1e-13, atol=1e-13)
    assert abs(pre.trYY - float(g["tr_YTY"])) < 1e-12


@pytest.mark.parametrize("kind", ["Matern32", "Matern52"])
def test_vff_converges_to_exact_gpr(kind):
    """With a domain much wider than the data and many features the inducing features determine f on
    the data, so ELBO -> exact LML and predict_f -> exact GPR (1-D, where the product kernel is the
    Matern kernel itself).  Pins make_Kuu / elbo / predict_f (vff.py:381-454, 612-675)."""
    from oracle import vff
    rng = np.random.default_rng(0)
    n = 60
    X = rng.uniform(-1.8, 1.9, (n, 1))
    y = np.sin(2 * X[:, 0]) + 0.1 * rng.standard_normal(n)
    th = np.array([0.7, 1.1, 0.05])
    Xs = rng.uniform(-1.5, 1.5, (5, 1))
    pre = vff.Precomputed(X, y, [-8.0], [8.0], [np.arange(768)])
    exact = gpr.lml(kind, X, y, th)
    v = vff.elbo(kind, pre, th)
    assert v <= exact + 1e-9 and abs(v - exact) < 2e-4
    p, q = vff.predict(kind, pre, th, Xs), gpr.predict(kind, X, y, th, Xs)
    np.testing.assert_allclose(p["f*"], q["f*"], rtol=0, atol=1e-6)
    np.testing.assert_allclose(p["f*_var"], q["f*_var"], rtol=0, atol=1e-6)
    np.testing.assert_allclose(p["y_var"], p["f*_var"] + th[2])


@pytest.mark.parametrize("kind", ["Matern12", "Matern32", "Matern52"])
@pytest.mark.parametrize("D", [1, 2])
def test_vff_gradient_fd(kind, D):
    from oracle import vff
    rng = np.random.default_rng(3 + D)
    n = 80
    X = rng.uniform(-1.8, 1.9, (n, D))
    y = np.sin(2 * X[:, 0]) + 0.1 * rng.standard_normal(n)
    pre = vff.Precomputed(X, y, [-2.2] * D, [2.3] * D, [np.arange(m) for m in ([12] if D == 1 else [6, 5])])
    th = np.concatenate([rng.uniform(0.5, 1.5, D), rng.uniform(0.5, 1.5, D), [0.05]])
    v, g = vff.elbo_and_grad(kind, pre, th)
    for i in range(len(th)):
        h = np.zeros_like(th)
        h[i] = 1e-6 * th[i]
        fd = (vff.elbo(kind, pre, th + h) - vff.elbo(kind, pre, th - h)) / (2 * h[i])
        assert abs(fd - g[i]) <= 1e-6 * max(1.0, abs(g[i]))


def test_vff_oracle_against_extended_precision(golden_dir):
    """oracle/vff.py (float64) against the extended-precision restatement of the same formulas (oracle/vff_ld.py,
    goldens by oracle/make_vff_envelope_golden.py): ELBO, analytic gradient (vs central differences of the long-double
    ELBO) and predictions of the small sweep cases, at the error levels recorded in the goldens."""
    import envelope as env
    from oracle import vff
    gold = np.load(os.path.join(golden_dir, "vff_envelope.npz"))
    for c in env.VFF_CASES:
        if c["n"] > 1500:
            continue
        X, y, Xs, a, b, ms, theta = env.make_vff_case(c)
        D = X.shape[1]
        pre = vff.Precomputed(X, y, a, b, ms)
        v, g = vff.elbo_and_grad(c["kind"], pre, theta)
        p = vff.predict(c["kind"], pre, theta, Xs)
        ref = {k: gold[f"{c['name']}/{k}"] for k in ("elbo", "grad", "fmean", "fvar")}
        assert abs(v - float(ref["elbo"])) <= 1e-10 * abs(float(ref["elbo"])), c["name"]
        assert np.max(np.abs(g - ref["grad"])) <= 1e-8 * np.max(np.abs(ref["grad"])), c["name"]
        assert np.max(np.abs(p["f*"] - ref["fmean"])) <= 1e-10 * np.max(np.abs(ref["fmean"])), c["name"]
        assert np.max(np.abs(p["f*_var"] - ref["fvar"])) <= 1e-10 * np.prod(theta[D:2 * D]), c["name"]


def test_vff_edge_features_and_fit():
    """Prediction features outside [a, b] take the kernel's edge form (vff.py:472-497); Matern52
    refuses them (:498-511).  The optimiser increases the ELBO."""
    from oracle import vff
    ms = np.arange(4)
    X1 = np.array([[-3.0], [0.3], [2.9]])
    inside = vff.kuf_np(X1[1:2], -2.0, 2.0, ms)
    for kind in ("Matern12", "Matern32"):
        F = vff.kuf_pred(kind, 0.8, X1, -2.0, 2.0, ms)
        np.testing.assert_allclose(F[:, 1:2], inside)
        assert np.allclose(F[:4, 0], F[0, 0]) and np.allclose(F[:4, 2], F[0, 2])      # cos rows share the edge value
    np.testing.assert_allclose(vff.kuf_pred("Matern12", 0.8, X1, -2.0, 2.0, ms)[:4, 0], np.exp(-1.0 / 0.8))
    assert np.all(vff.kuf_pred("Matern12", 0.8, X1, -2.0, 2.0, ms)[4:, [0, 2]] == 0.0)
    with pytest.raises(AssertionError):
        vff.kuf_pred("Matern52", 0.8, X1, -2.0, 2.0, ms)
    rng = np.random.default_rng(9)
    X = rng.uniform(-1.5, 1.5, (50, 1))
    y = np.sin(3 * X[:, 0]) + 0.1 * rng.standard_normal(50)
    pre = vff.Precomputed(X, y, [-2.5], [2.5], [np.arange(10)])
    th0 = np.ones(3)
    th, val, ok, nfev = vff.fit("Matern32", pre, th0)
    assert ok and val > vff.elbo("Matern32", pre, th0) and nfev > 3


# ---- post-processing (oracle/postprocessing.py) ---------------------------------------------------
def test_postprocessing_oracle_matches_reference_functions(golden_dir):
    """gaussian_2d_weight / glue_local_predictions_1d / _2d of /root/reference/GPSat/postprocessing.py
    (:22-52, :462-593), executed from the reference source by oracle/make_golden.py."""
    import pandas as pd
    from oracle import postprocessing as pp
    g = np.load(os.path.join(golden_dir, "post_smooth.npz"))
    np.testing.assert_allclose(pp.gaussian_2d_weight(g["x"], g["y"], g["x"], g["y"], float(g["l_x"]), float(g["l_y"]),
                                                     g["vals"]), g["out"], rtol=1e-13)
    assert np.isnan(pp.gaussian_2d_weight([1e9], [1e9], g["x"], g["y"], 1.0, 1.0, g["vals"])[0]) and np.isnan(g["far"][0])
    q = np.load(os.path.join(golden_dir, "post_glue.npz"), allow_pickle=True)
    df = pd.DataFrame(q["df"], columns=list(q["cols"]))
    g2 = pp.glue_2d(df, ["pred_loc_x", "pred_loc_y"], ["x", "y"], ["f*", "f*_var"], 150.0, R=3)
    np.testing.assert_allclose(g2[list(q["g2cols"])].values, q["g2"], rtol=1e-13)
    radius = dict(zip(q["radius_keys"].tolist(), q["radius_vals"].tolist()))
    g1 = pp.glue_1d(df[df["y"] == 0.0], "pred_loc_x", "x", ["f*", "f*_var"], radius, R=2)
    np.testing.assert_allclose(g1[list(q["g1cols"])].values, q["g1"], rtol=1e-13)
