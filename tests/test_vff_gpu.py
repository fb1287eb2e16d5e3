"""Parity of the CUDA variational-Fourier-features path (C ABI: gpsat_vff_elbo / _predict / _fit) with
oracle/vff.py, the NumPy restatement of /root/reference/GPSat/vff.py (GPR_kron).

Tolerance at fixed hyper-parameters: 1e-8 relative on the ELBO and the predictions (absolute floor
scaled by the kernel variance), 1e-6 on the analytic gradient, 1e-4 on the optimum.
"""
import numpy as np
import pytest

from oracle import gpr, vff

pytestmark = pytest.mark.gpu

RTOL = 1e-8


def _expert(rng, n, D, nfreq, m, margin=0.4):
    X = rng.uniform(-2.0, 2.0, (n, D))
    y = np.sin(2 * X[:, 0]) + (0.5 * np.cos(X[:, 1]) if D > 1 else 0.0) + 0.1 * rng.standard_normal(n)
    a = X.min(axis=0) - margin * rng.uniform(0.5, 1.5, D)
    b = X.max(axis=0) + margin * rng.uniform(0.5, 1.5, D)
    Xs = rng.uniform(-1.9, 1.9, (m, D))
    return dict(X=X, y=y, Xs=Xs, a=a, b=b, nfreq=np.asarray(nfreq, dtype=np.int32))


def _theta(rng, D):
    return np.concatenate([rng.uniform(0.5, 1.5, D), rng.uniform(0.5, 1.5, D), [rng.uniform(0.02, 0.2)]])


def _batch(kind, experts):
    from gpsat_b200.engine import RaggedBatch
    return RaggedBatch.from_lists(kind, [e["X"] for e in experts], [e["y"] for e in experts],
                                  [e["Xs"] for e in experts])


def _dom(experts):
    return (np.stack([e["a"] for e in experts]), np.stack([e["b"] for e in experts]),
            np.stack([e["nfreq"] for e in experts]))


def _pre(e):
    return vff.Precomputed(e["X"], e["y"], e["a"], e["b"], [np.arange(k) for k in e["nfreq"]])


SHAPES = {1: [(200, [10], 30), (129, [64], 0), (500, [33], 140), (64, [3], 5), (400, [128], 20), (300, [100], 7)],
          2: [(300, [6, 5], 17), (130, [8, 8], 0), (700, [12, 6], 140), (257, [3, 20], 5), (64, [1, 2], 3)],
          3: [(300, [3, 3, 2], 17), (400, [5, 4, 3], 40), (150, [2, 2, 2], 0)]}


@pytest.mark.parametrize("kind", vff.KINDS)
@pytest.mark.parametrize("D", [1, 2, 3])
def test_elbo_grad_predict_ragged(engine, kind, D):
    rng = np.random.default_rng(13 * D + len(kind))
    experts = [_expert(rng, n, D, nf, m) for n, nf, m in SHAPES[D]]
    theta = np.stack([_theta(rng, D) for _ in experts])
    b = _batch(kind, experts)
    dom = _dom(experts)
    elbo, grad, status = engine.vff_elbo(b, dom, theta, with_grad=True)
    elbo0, none, _ = engine.vff_elbo(b, dom, theta)
    out = engine.vff_predict(b, dom, theta)
    assert none is None and np.all(status == 0) and np.all(out["status"] == 0)
    np.testing.assert_allclose(elbo0, elbo, rtol=1e-13)
    np.testing.assert_allclose(out["elbo"], elbo, rtol=1e-13)
    for e, ex in enumerate(experts):
        pre = _pre(ex)
        v, g = vff.elbo_and_grad(kind, pre, theta[e])
        assert abs(elbo[e] - v) <= RTOL * abs(v), (e, elbo[e], v)
        np.testing.assert_allclose(grad[e], g, rtol=1e-6, atol=1e-6 * np.abs(g).max())
        if len(ex["Xs"]):
            p = vff.predict(kind, pre, theta[e], ex["Xs"])
            sl = slice(int(b.off_m[e]), int(b.off_m[e + 1]))
            atol = 1e-8 * float(np.prod(theta[e][D:2 * D]))
            np.testing.assert_allclose(out["f*"][sl], p["f*"], rtol=1e-7, atol=atol * max(1.0, np.abs(ex["y"]).max()))
            np.testing.assert_allclose(out["f*_var"][sl], p["f*_var"], rtol=1e-7, atol=atol)
            np.testing.assert_allclose(out["y_var"][sl], p["y_var"], rtol=1e-7, atol=atol)


@pytest.mark.parametrize("kind", ["Matern12", "Matern32"])
def test_prediction_outside_domain_uses_edge_features(engine, kind):
    """vff.py:472-497: prediction points left of a / right of b."""
    rng = np.random.default_rng(2)
    ex = _expert(rng, 150, 2, [6, 6], 0)
    ex["Xs"] = np.array([[ex["a"][0] - 0.3, 0.1], [0.2, ex["b"][1] + 0.7], [ex["a"][0] - 0.1, ex["b"][1] + 0.2],
                         [0.0, 0.0]])
    theta = _theta(rng, 2)[None, :]
    b = _batch(kind, [ex])
    out = engine.vff_predict(b, _dom([ex]), theta)
    p = vff.predict(kind, _pre(ex), theta[0], ex["Xs"])
    np.testing.assert_allclose(out["f*"], p["f*"], rtol=1e-7, atol=1e-9)
    np.testing.assert_allclose(out["f*_var"], p["f*_var"], rtol=1e-7, atol=1e-9)


def test_converges_to_exact_gpr(engine):
    """1-D, wide domain, 64 frequencies per unit... : the ELBO approaches the exact log marginal
    likelihood and the predictions the exact GPR ones (the pin of the oracle, repeated on the GPU path)."""
    rng = np.random.default_rng(0)
    n = 60
    X = rng.uniform(-1.8, 1.9, (n, 1))
    y = np.sin(2 * X[:, 0]) + 0.1 * rng.standard_normal(n)
    ex = dict(X=X, y=y, Xs=rng.uniform(-1.5, 1.5, (5, 1)), a=np.array([-5.0]), b=np.array([5.0]),
              nfreq=np.array([64], dtype=np.int32))
    theta = np.array([[0.7, 1.1, 0.05]])
    out = engine.vff_predict(_batch("Matern52", [ex]), _dom([ex]), theta)
    exact = gpr.lml("Matern52", X, y, theta[0])
    assert out["elbo"][0] <= exact + 1e-9 and abs(out["elbo"][0] - exact) < 5e-3
    q = gpr.predict("Matern52", X, y, theta[0], ex["Xs"])
    np.testing.assert_allclose(out["f*"], q["f*"], rtol=0, atol=1e-4)
    np.testing.assert_allclose(out["f*_var"], q["f*_var"], rtol=0, atol=1e-4)


def test_fit_matches_oracle(engine):
    rng = np.random.default_rng(31)
    D = 2
    experts = [_expert(rng, n, D, nf, 0) for n, nf in [(300, [6, 5]), (200, [4, 7]), (500, [8, 8])]]
    E, P = len(experts), 2 * D + 1
    tr = vff.Transform.vff_default(D)
    theta0 = np.ones((E, P))
    b = _batch("Matern32", experts)
    res = engine.fit(b, theta0, np.tile(tr.kind, (E, 1)), np.tile(tr.a, (E, 1)), np.tile(tr.b, (E, 1)),
                     np.ones((E, P), np.uint8), vff=_dom(experts))
    assert np.all(res.status == 0)
    for e, ex in enumerate(experts):
        th, val, ok, nfev = vff.fit("Matern32", _pre(ex), theta0[e], tr)
        assert ok
        assert abs(res.lml[e] - val) <= 1e-4 * abs(val), (res.lml[e], val)
        # only the product of the per-dimension variances is identified; compare it and the rest
        np.testing.assert_allclose(res.theta[e][:D], th[:D], rtol=5e-3)
        np.testing.assert_allclose(np.prod(res.theta[e][D:2 * D]), np.prod(th[D:2 * D]), rtol=5e-3)
        np.testing.assert_allclose(res.theta[e][2 * D], th[2 * D], rtol=5e-3)


def test_bad_inputs(engine):
    from gpsat_b200 import _lib
    rng = np.random.default_rng(1)
    ex = _expert(rng, 50, 1, [5], 0)
    theta = _theta(rng, 1)[None, :]
    with pytest.raises(RuntimeError):
        engine.vff_elbo(_batch("RBF", [ex]), _dom([ex]), theta)                 # no VFF for RBF
    bad = dict(ex, a=np.array([ex["X"].min() + 0.1]))
    with pytest.raises(RuntimeError):
        engine.vff_elbo(_batch("Matern32", [bad]), _dom([bad]), theta)           # data outside (a, b)
    big = dict(ex, nfreq=np.array([129], dtype=np.int32))
    with pytest.raises(RuntimeError):
        engine.vff_elbo(_batch("Matern32", [big]), _dom([big]), theta)           # > 256 features / dim


def test_vff_model_driver_sequence_and_batched(engine):
    """B200VFFModel through the call order of LocalExpertOI.run (local_experts.py:1043-1164), against
    the oracle driven the same way; then the batched helper against the per-model loop."""
    import pandas as pd
    from gpsat_b200.models import B200VFFModel, fit_predict_models
    rng = np.random.default_rng(8)
    cs = np.array([5e4, 5e4])

    def make(seed):
        r = np.random.default_rng(seed)
        n = 250
        xy = r.uniform(-1.5e5, 1.5e5, (n, 2))
        obs = np.sin(xy[:, 0] / 6e4) + 0.4 * np.cos(xy[:, 1] / 5e4) + 0.1 * r.standard_normal(n) + 3.0
        return pd.DataFrame({"x": xy[:, 0], "y": xy[:, 1], "obs": obs})

    def build(df):
        m = B200VFFModel(data=df, obs_col="obs", coords_col=["x", "y"], coords_scale=list(cs), obs_mean="local",
                         num_inducing_features=[7, 6], domain_size=2.5e5, expert_loc=np.array([0.0, 0.0]),
                         verbose=False)
        m.set_parameter_constraints({"lengthscales": {"low": [1e-8, 1e-8], "high": [6e5, 6e5], "scale": True}},
                                    move_within_tol=True, tol=1e-2)
        return m

    dfs = [make(s) for s in (1, 2, 3)]
    pred = rng.uniform(-1.2e5, 1.2e5, (40, 2))
    m = build(dfs[0])
    assert m.optimise_parameters()
    th = m._theta.copy()
    # oracle on the same scaled data / domain / transforms
    pre = vff.Precomputed(m.coords, m.obs[:, 0], m.vff_a, m.vff_b, [np.arange(k) for k in m.vff_nfreq])
    tr = vff.Transform.vff_default(2)
    tr.set_sigmoid(0, 1e-8 / cs[0], 6e5 / cs[0])
    tr.set_sigmoid(1, 1e-8 / cs[1], 6e5 / cs[1])
    tho, val, ok, _ = vff.fit("Matern32", pre, np.ones(5), tr)
    assert ok
    assert abs(m.get_objective_function_value() - val) < 1e-4 * abs(val)
    np.testing.assert_allclose(m.get_lengthscales(), tho[:2], rtol=5e-3)
    np.testing.assert_allclose(m.get_kernel_variance(), np.prod(tho[2:4]), rtol=5e-3)
    p = m.predict(pred)
    po = vff.predict("Matern32", pre, th, pred / cs)
    np.testing.assert_allclose(p["f*"], po["f*"], rtol=1e-7, atol=1e-9)
    np.testing.assert_allclose(p["f*_var"], po["f*_var"], rtol=1e-7, atol=1e-9)
    np.testing.assert_allclose(p["f_bar"], np.full(40, dfs[0]["obs"].mean()))
    # batched helper == loop
    loop = [build(df) for df in dfs]
    loop_ok = [mm.optimise_parameters() for mm in loop]
    loop_obj = [mm.get_objective_function_value() for mm in loop]
    loop_pred = [mm.predict(pred) for mm in loop]
    batch = [build(df) for df in dfs]
    ok_b, obj_b, pred_b = fit_predict_models(batch, [pred] * 3)
    assert list(ok_b) == loop_ok
    for e in range(3):
        np.testing.assert_allclose(batch[e]._theta, loop[e]._theta, rtol=1e-12)
        assert abs(obj_b[e] - loop_obj[e]) <= 1e-12 * abs(loop_obj[e])
        np.testing.assert_allclose(pred_b[e]["f*"], loop_pred[e]["f*"], rtol=1e-12, atol=1e-14)


def test_vff_against_extended_precision(engine, golden_dir):
    """ELBO / gradient / predictions of the Fourier-feature model against extended precision (oracle/vff_ld.py, the
    formulas of GPR_kron.elbo / predict_f, vff.py:612-675, in numpy.longdouble; gradient = central differences of that
    ELBO; goldens: oracle/make_vff_envelope_golden.py).  The reference's own GPR_kron needs TensorFlow and its test is
    commented out (tests/test_localexperts.py:150-190), so this is the pin of the row: 1-D with 40 and 64 frequencies,
    2-D up to 20 x 20 (M = 1 521), 3-D, likelihood variance down to 1e-4, prediction points outside the domain box.
    Held to the north star's 1e-8 on all four quantities; the float64 NumPy oracle's own error is recorded in the
    goldens (f64_err) and written beside the CUDA path's to gpurun_out/vff_envelope.json."""
    import json
    import os
    import envelope as env
    from gpsat_b200.engine import RaggedBatch
    gold = np.load(os.path.join(golden_dir, "vff_envelope.npz"))
    rows, bad = [], []
    for c in env.VFF_CASES:
        X, y, Xs, a, b_, ms, theta = env.make_vff_case(c)
        D = X.shape[1]
        batch = RaggedBatch.from_lists(c["kind"], [X], [y], [Xs])
        dom = (a[None, :], b_[None, :], np.asarray([c["nfreq"]], dtype=np.int32))
        v, g, st = engine.vff_elbo(batch, dom, theta[None, :], with_grad=True)
        p = engine.vff_predict(batch, dom, theta[None, :])
        ref = {k: gold[f"{c['name']}/{k}"] for k in ("elbo", "grad", "fmean", "fvar", "f64_err", "cond_p")}
        err = np.array([abs(v[0] - float(ref["elbo"])) / abs(float(ref["elbo"])),
                        np.max(np.abs(g[0] - ref["grad"])) / np.max(np.abs(ref["grad"])),
                        np.max(np.abs(p["f*"] - ref["fmean"])) / np.max(np.abs(ref["fmean"])),
                        np.max(np.abs(p["f*_var"] - ref["fvar"])) / np.prod(theta[D:2 * D])])
        tol = np.full(4, 1e-8)
        rows.append(dict(case=c["name"], cond_p=float(ref["cond_p"]), gpu=err.tolist(), f64_oracle=ref["f64_err"].tolist(),
                         tol=tol.tolist(), status=int(st[0])))
        if st[0] != 0 or p["status"][0] != 0 or np.any(err > tol):
            bad.append(rows[-1])
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(rows, open(os.path.join("gpurun_out", "vff_envelope.json"), "w"), indent=1)
    assert not bad, bad
