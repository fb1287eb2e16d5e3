"""The plugin surface on a B200, checked like the reference checks GPflowGPRModel
(/root/reference/tests/test_localexperts.py:204-227) plus the batched entry point and priors."""
import os

import numpy as np
import pandas as pd
import pytest

from oracle import gpr

pytestmark = pytest.mark.gpu


def test_b200_gpr_reference_kat(built, golden_dir):
    """test_gpflow_gpr with B200GPRModel in place of GPflowGPRModel; tolerance 1e-6 as there."""
    from gpsat_b200.models import B200GPRModel
    g = np.load(os.path.join(golden_dir, "kat_localexperts.npz"))
    eps, tol = float(g["eps"]), 1e-6
    df = pd.DataFrame({"x": g["x_train"][:, 0], "y": g["y_train"][:, 0]})
    model = B200GPRModel(data=df, obs_col="y", coords_col="x", obs_mean=None)
    model.set_parameters(likelihood_variance=eps ** 2)
    model.set_parameter_constraints({"lengthscales": {"low": 1e-10, "high": 5.0}})
    result = model.optimise_parameters(fixed_params=["likelihood_variance", "kernel_variance"])
    out = model.predict(coords=g["x_test"])
    params = model.get_parameters()
    objfunc = -model.get_objective_function_value()
    assert result
    assert np.abs(params["lengthscales"][0] - float(g["ls"])) < tol
    assert np.abs(objfunc - float(g["ml"])) < tol
    assert np.abs(out["f*"] - g["pred_mean"]) < tol
    assert np.abs(out["f*_var"] - g["pred_std"] ** 2) < tol
    assert np.abs(out["y_var"] - (out["f*_var"] + eps ** 2)) < 1e-14
    assert set(out) == {"f*", "f*_var", "y_var", "f_bar"} and out["f_bar"].shape == out["f*"].shape
    assert model.gpu_name is None or "B200" in model.gpu_name


def _frames(rng, E, D=3):
    cols = ["x", "y", "t"][:D]
    out = []
    for e in range(E):
        n = int(rng.integers(40, 260))
        X = rng.uniform(-2e5, 2e5, (n, D))
        if D == 3:
            X[:, 2] = rng.integers(-4, 5, n)
        cs = np.array([5e4, 5e4, 1.0][:D])
        K = gpr.kernel_matrix("Matern32", X / cs, None, [2.0, 1.5, 3.0][:D], 0.7)
        y = np.linalg.cholesky(K + 1e-8 * np.eye(n)) @ rng.standard_normal(n) + 0.2 * rng.standard_normal(n) + 3.0
        df = pd.DataFrame(X, columns=cols)
        df["obs"] = y
        out.append(df)
    return out, cols


def test_driver_call_sequence_matches_oracle(built):
    """The call order of LocalExpertOI.run (local_experts.py:1043-1164) on one expert, against the
    oracle with the same transforms: constraints with scale=True and tol=1e-2, optimise, objective,
    parameters, predict on raw coordinates."""
    from gpsat_b200.models import B200GPRModel
    rng = np.random.default_rng(5)
    (df,), cols = _frames(rng, 1)
    init = {"coords_scale": [5e4, 5e4, 1], "obs_mean": "local", "kernel": "Matern52"}
    cons = {"lengthscales": {"low": [1e-8] * 3, "high": [6e5, 6e5, 9], "scale": True},
            "likelihood_variance": {"low": 1e-5, "high": 2.0}}
    model = B200GPRModel(data=df, obs_col="obs", coords_col=cols, expert_loc=np.zeros(3), **init)
    model.set_parameter_constraints(cons, move_within_tol=True, tol=1e-2)
    ok = model.optimise_parameters(max_iter=10_000)
    obj = model.get_objective_function_value()
    hyp = model.get_parameters("lengthscales", "kernel_variance", "likelihood_variance")
    pc = rng.uniform(-1e5, 1e5, (30, 3))
    pred = model.predict(coords=pc)

    cs = np.array([5e4, 5e4, 1.0])
    X = df[cols].values / cs
    y = df["obs"].values - df["obs"].mean()
    tr = gpr.Transform.gpflow_default(3)
    tr.set_sigmoid([0, 1, 2], np.array([1e-8] * 3) / cs, np.array([6e5, 6e5, 9]) / cs)
    tr.set_sigmoid([4], [1e-5], [2.0])
    th, l, ok_ref, nev = gpr.fit("Matern52", X, y, np.ones(5), tr)
    assert ok == ok_ref
    theta = np.concatenate([hyp["lengthscales"], [hyp["kernel_variance"], hyp["likelihood_variance"]]])
    np.testing.assert_allclose(theta, th, rtol=1e-4)
    assert abs(obj + l) <= 1e-4 * abs(l)
    p = gpr.predict("Matern52", X, y, theta, pc / cs)
    np.testing.assert_allclose(pred["f*"], p["f*"], rtol=1e-8, atol=1e-10)
    np.testing.assert_allclose(pred["f*_var"], p["f*_var"], rtol=1e-8, atol=1e-12)
    np.testing.assert_allclose(pred["f_bar"], np.repeat(df["obs"].mean(), 30))
    assert obj == pytest.approx(-gpr.lml("Matern52", X, y, theta), rel=1e-10)


def test_batched_equals_per_instance(built):
    """fit_predict_models (one engine call for all experts) gives bit-identical parameters and
    predictions to the per-instance calls the reference driver makes."""
    from gpsat_b200.models import B200GPRModel, fit_predict_models
    rng = np.random.default_rng(11)
    frames, cols = _frames(rng, 5)
    init = {"coords_scale": [5e4, 5e4, 1], "obs_mean": "local"}
    cons = {"lengthscales": {"low": [1e-8] * 3, "high": [6e5, 6e5, 9], "scale": True}}
    pcs = [rng.uniform(-1e5, 1e5, (int(rng.integers(1, 40)), 3)) for _ in frames]

    def make():
        ms = [B200GPRModel(data=df, obs_col="obs", coords_col=cols, **init) for df in frames]
        for m in ms:
            m.set_parameter_constraints(cons, move_within_tol=True, tol=1e-2)
        return ms

    a = make()
    succ, obj, preds = fit_predict_models(a, pcs, max_iter=25)
    b = make()
    for e, m in enumerate(b):
        s = m.optimise_parameters(max_iter=25)
        assert s == succ[e]
        np.testing.assert_array_equal(m._theta, a[e]._theta)
        assert m.get_objective_function_value() == obj[e]
        pr = m.predict(pcs[e])
        for k in ("f*", "f*_var", "y_var", "f_bar"):
            np.testing.assert_array_equal(pr[k], preds[e][k])


def test_priors_match_oracle(built):
    """Extension: log prior density (GPflow convention) added to the objective."""
    from gpsat_b200.models import B200GPRModel
    rng = np.random.default_rng(3)
    (df,), cols = _frames(rng, 1, D=2)
    model = B200GPRModel(data=df, obs_col="obs", coords_col=cols, coords_scale=[5e4, 5e4], obs_mean="local",
                         priors={"lengthscales": {"family": "gamma", "a": 2.0, "b": 1.5},
                                 "kernel_variance": {"family": "lognormal", "a": -0.5, "b": 0.8},
                                 "likelihood_variance": {"family": "normal", "a": 0.05, "b": 0.1}})
    ok = model.optimise_parameters()
    X = df[cols].values / np.array([5e4, 5e4])
    y = df["obs"].values - df["obs"].mean()
    tr = gpr.Transform.gpflow_default(2)
    pri = (np.array([3, 3, 2, 1]), np.array([2.0, 2.0, -0.5, 0.05]), np.array([1.5, 1.5, 0.8, 0.1]))
    th, l, ok_ref, nev, res = gpr.fit("Matern32", X, y, np.ones(4), tr, priors=pri, return_result=True)
    assert ok == ok_ref
    np.testing.assert_allclose(model._theta, th, rtol=1e-4)
    assert model.opt_info["objective"] == pytest.approx(res.fun, rel=1e-6)
    # the objective value reported to the driver stays the pure -LML (gpflow_models.py:337)
    assert model.get_objective_function_value() == pytest.approx(-l, rel=1e-6)


@pytest.mark.parametrize("n,m", [(60, 5), (127, 70), (200, 130), (300, 64)])
def test_full_cov(built, n, m):
    """predict(full_cov=True): f*_cov / y_cov as gpflow_models.py:245-263 builds them."""
    from gpsat_b200.models import B200GPRModel
    rng = np.random.default_rng(n + m)
    X = rng.uniform(-3, 3, (n, 2))
    y = np.sin(X[:, 0]) + 0.1 * rng.standard_normal(n)
    model = B200GPRModel(coords=X.copy(), obs=y.copy(), kernel="Matern52",
                         kernel_kwargs={"lengthscales": [1.3, 0.8], "variance": 0.9}, noise_variance=0.05)
    Xs = rng.uniform(-3, 3, (m, 2))
    out = model.predict(Xs, full_cov=True)
    ref = gpr.predict("Matern52", X, y, model._theta, Xs, full_cov=True)
    scale = np.abs(ref["f*_cov"]).max()
    np.testing.assert_allclose(out["f*_cov"], ref["f*_cov"], rtol=0, atol=1e-10 * scale)
    np.testing.assert_allclose(out["y_cov"], ref["y_cov"], rtol=0, atol=1e-10 * scale)
    np.testing.assert_allclose(out["f*_var"], np.diag(ref["f*_cov"]), rtol=1e-8, atol=1e-12)
    np.testing.assert_allclose(out["f*"], ref["f*"], rtol=1e-8, atol=1e-10)
    assert np.array_equal(out["f*_cov"], out["f*_cov"].T)
    assert set(out) == {"f*", "f*_var", "y_var", "f*_cov", "y_cov", "f_bar"}


def test_batched_driver_equals_reference_loop(built):
    """gpsat_b200.local_experts.LocalExpertOI.run == the reference's per-expert loop
    (local_experts.py:930-1260) executed with B200GPRModel instances, table by table."""
    from gpsat_b200.local_experts import LocalExpertOI, local_data_select, prediction_coords_for
    from gpsat_b200.models import B200GPRModel
    rng = np.random.default_rng(21)
    n = 2500
    data = pd.DataFrame({"x": rng.uniform(-4e5, 4e5, n), "y": rng.uniform(-4e5, 4e5, n),
                         "t": rng.integers(-6, 7, n).astype(float)})
    data["obs"] = np.sin(data.x / 1.5e5) + np.cos(data.y / 1e5) + 0.1 * data.t + 0.1 * rng.standard_normal(n)
    locs = pd.DataFrame({"x": [-1e5, 0.0, 1.5e5, 3.9e5, 9e5], "y": [0.0, 1e5, -1e5, 3.9e5, 9e5], "t": [0.0] * 5})
    grid = pd.DataFrame({"x": np.repeat(np.linspace(-4e5, 4e5, 33), 33), "y": np.tile(np.linspace(-4e5, 4e5, 33), 33)})
    data_cfg = {"obs_col": "obs", "coords_col": ["x", "y", "t"],
                "local_select": [{"col": "t", "comp": "<=", "val": 4}, {"col": "t", "comp": ">=", "val": -4},
                                 {"col": ["x", "y"], "comp": "<", "val": 1.5e5}]}
    model_cfg = {"oi_model": "B200GPRModel", "init_params": {"coords_scale": [5e4, 5e4, 1], "obs_mean": "local"},
                 "constraints": {"lengthscales": {"low": [1e-8] * 3, "high": [6e5, 6e5, 9]}},
                 "optim_kwargs": {"max_iter": 40}}
    oi = LocalExpertOI(locs, data, data_cfg, model_cfg, {"df": grid, "max_dist": 1e5})
    tabs = oi.run(min_obs=100)
    assert set(tabs) == {"run_details", "preds", "lengthscales", "kernel_variance", "likelihood_variance"}
    rd = tabs["run_details"]
    assert len(rd) == 4                      # the last expert has no prediction location in range: skipped
    assert rd["num_obs"].min() < 100 and np.isnan(rd.loc[rd["num_obs"] < 100, "objective_value"]).all()
    # the reference loop, one model at a time
    cons = {"lengthscales": {"low": [1e-8] * 3, "high": [6e5, 6e5, 9], "scale": True}}
    k = 0
    for _, rl in locs.iterrows():
        ref = {c: rl[c] for c in ["x", "y", "t"]}
        pc = prediction_coords_for(grid, ref, ["x", "y", "t"], 1e5)
        df_local = local_data_select(data, ref, data_cfg["local_select"])
        if len(pc) == 0 or len(df_local) < 100:
            continue
        m = B200GPRModel(data=df_local, obs_col="obs", coords_col=["x", "y", "t"], **model_cfg["init_params"])
        m.set_parameter_constraints(cons, move_within_tol=True, tol=1e-2)
        ok = m.optimise_parameters(max_iter=40)
        obj = m.get_objective_function_value()
        pr = m.predict(pc)
        key = (ref["x"], ref["y"], ref["t"])
        row = rd.loc[[key]]
        assert bool(row["optimise_success"].iloc[0]) == ok and row["objective_value"].iloc[0] == obj
        assert row["num_obs"].iloc[0] == len(df_local)
        p = tabs["preds"].loc[[key]]
        np.testing.assert_array_equal(p["f*"].values, pr["f*"])
        np.testing.assert_array_equal(p["f*_var"].values, pr["f*_var"])
        np.testing.assert_array_equal(p["pred_loc_x"].values, pc[:, 0])
        np.testing.assert_array_equal(tabs["lengthscales"].loc[[key]]["lengthscales"].values, m.get_lengthscales())
        assert tabs["kernel_variance"].loc[[key]]["kernel_variance"].iloc[0] == m.get_kernel_variance()
        k += 1
    assert k == 3


def _driver_case(seed=21, n=2500):
    rng = np.random.default_rng(seed)
    data = pd.DataFrame({"x": rng.uniform(-4e5, 4e5, n), "y": rng.uniform(-4e5, 4e5, n),
                         "t": rng.integers(-6, 7, n).astype(float)})
    data["obs"] = np.sin(data.x / 1.5e5) + np.cos(data.y / 1e5) + 0.1 * data.t + 0.1 * rng.standard_normal(n)
    locs = pd.DataFrame({"x": [-1e5, 0.0, 1.5e5, 3.9e5, 9e5, 2e5], "y": [0.0, 1e5, -1e5, 3.9e5, 9e5, 2e5],
                         "t": [0.0, 1.0, -1.0, 0.0, 0.0, 2.0]})
    grid = pd.DataFrame({"x": np.repeat(np.linspace(-4e5, 4e5, 33), 33), "y": np.tile(np.linspace(-4e5, 4e5, 33), 33)})
    data_cfg = {"obs_col": "obs", "coords_col": ["x", "y", "t"],
                "local_select": [{"col": "t", "comp": "<=", "val": 4}, {"col": "t", "comp": ">=", "val": -4},
                                 {"col": ["x", "y"], "comp": "<", "val": 1.5e5}]}
    model_cfg = {"oi_model": "B200GPRModel", "init_params": {"coords_scale": [5e4, 5e4, 1], "obs_mean": "local"},
                 "constraints": {"lengthscales": {"low": [1e-8] * 3, "high": [6e5, 6e5, 9]}},
                 "optim_kwargs": {"max_iter": 40}}
    return locs, data, data_cfg, model_cfg, grid


def test_device_selection_equals_host_selection(built):
    """gpsat_local_select (all experts in one launch) returns exactly the rows, in data order, that
    DataLoader.local_data_select (dataloader.py:2403-2447) and the prediction-location distance
    filter (prediction_locations.py:31-43) keep - including no-match experts and every comparison."""
    from gpsat_b200.local_experts import local_data_select, prediction_coords_for, select_all
    from gpsat_b200.models import get_engine
    locs, data, data_cfg, _, grid = _driver_case()
    cols = ["x", "y", "t"]
    table = data[cols + ["obs"]].values
    L = locs[cols].values
    conds = [(0, "<=", [2], [2], 4), (0, ">=", [2], [2], -4), (1, "<=", [0, 1], [0, 1], 1.5e5)]
    counts, off, idx = select_all(get_engine(0), table, L, conds)
    for e in range(len(L)):
        ref = dict(zip(cols, L[e]))
        want = local_data_select(data, ref, data_cfg["local_select"]).index.values
        np.testing.assert_array_equal(idx[off[e]:off[e + 1]], want)
        assert counts[e] == len(want)
    assert counts[4] == 0
    # every scalar comparison, against numpy
    for op, fn in ((">=", np.greater_equal), (">", np.greater), ("==", np.equal), ("<", np.less), ("<=", np.less_equal)):
        c, o, i = select_all(get_engine(0), table, L, [(0, op, [2], [2], 2.0)])
        for e in range(len(L)):
            np.testing.assert_array_equal(i[o[e]:o[e + 1]], np.flatnonzero(fn(table[:, 2], L[e, 2] + 2.0)))
    # prediction locations: strict d^2 < max_dist^2 over the shared columns
    c, o, i = select_all(get_engine(0), grid[["x", "y"]].values, L, [(1, "<", [0, 1], [0, 1], 1e5)])
    for e in range(len(L)):
        pc = prediction_coords_for(grid, dict(zip(cols, L[e])), cols, 1e5)
        np.testing.assert_array_equal(grid[["x", "y"]].values[i[o[e]:o[e + 1]]], pc[:, :2])
    # no condition at all selects everything; an empty table selects nothing
    c, o, i = select_all(get_engine(0), table[:7], L, [])
    assert np.all(c == 7) and np.array_equal(i, np.tile(np.arange(7), len(L)))
    c, o, i = select_all(get_engine(0), table[:0], L, conds)
    assert np.all(c == 0) and len(i) == 0


@pytest.mark.parametrize("optimise,predict", [(True, True), (False, True), (True, False)])
def test_vectorised_driver_equals_per_model_driver(built, optimise, predict):
    """LocalExpertOI.run(device_select=True) - GPU selection, array-built batch and tables - returns
    the same tables as the per-model path, value for value (run_time aside)."""
    from gpsat_b200.local_experts import LocalExpertOI
    locs, data, data_cfg, model_cfg, grid = _driver_case()
    a = LocalExpertOI(locs, data, data_cfg, model_cfg, {"df": grid, "max_dist": 1e5}).run(
        min_obs=100, optimise=optimise, predict=predict, table_suffix="_A", config_id=3, device_select=False)
    b = LocalExpertOI(locs, data, data_cfg, model_cfg, {"df": grid, "max_dist": 1e5}).run(
        min_obs=100, optimise=optimise, predict=predict, table_suffix="_A", config_id=3, device_select=True)
    assert set(a) == set(b)
    for k in a:
        ta, tb = a[k], b[k]
        if k.startswith("run_details"):
            ta, tb = ta.drop(columns="run_time"), tb.drop(columns="run_time")
        pd.testing.assert_frame_equal(ta, tb, check_exact=True, check_dtype=False)


def test_vectorised_driver_other_configs(built):
    """No prediction table (predict at the expert location), no local mean, fixed parameters, and the
    refusal for other model families."""
    from gpsat_b200.local_experts import LocalExpertOI
    locs, data, data_cfg, model_cfg, _ = _driver_case(seed=5, n=1500)
    model_cfg = dict(model_cfg, init_params={"coords_scale": [5e4, 5e4, 1]},
                     optim_kwargs={"max_iter": 25, "fixed_params": ["likelihood_variance"]},
                     params_to_store=["lengthscales", "likelihood_variance"])
    a = LocalExpertOI(locs, data, data_cfg, model_cfg).run(min_obs=30, device_select=False)
    b = LocalExpertOI(locs, data, data_cfg, model_cfg).run(min_obs=30, device_select=True)
    c = LocalExpertOI(locs, data, data_cfg, model_cfg).run(min_obs=30)                 # default: automatic = array-built
    assert set(a) == set(b) == {"run_details", "preds", "lengthscales", "likelihood_variance"}
    for k in a:
        ta, tb = a[k], b[k]
        if k == "run_details":
            ta, tb = ta.drop(columns="run_time"), tb.drop(columns="run_time")
        pd.testing.assert_frame_equal(ta, tb, check_exact=True, check_dtype=False)
    assert np.all(b["likelihood_variance"]["likelihood_variance"].values == 1.0)
    pd.testing.assert_frame_equal(c["preds"], b["preds"], check_exact=True)
    with pytest.raises(NotImplementedError):
        LocalExpertOI(locs, data, data_cfg, dict(model_cfg, oi_model="B200SGPRModel")).run(device_select=True)


def test_device_selection_equals_the_reference_functions(engine):
    """gpsat_local_select against vectors produced by the reference's own DataLoader.local_data_select
    (GPSat/dataloader.py:2354-2447) and _max_dist_bool (GPSat/prediction_locations.py:18-43) executed from source
    (oracle/make_golden.py --driver): including rows exactly on the ball and on the time-window edges."""
    import json
    from gpsat_b200 import local_experts as le
    gold = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    g = np.load(os.path.join(gold, "driver_select.npz"))
    cols = ["x", "y", "t"]
    table, locs = g["data"], g["locs"]
    for name, ls_list in json.loads(str(g["selects"])).items():
        conds = []
        for ls in ls_list:
            if isinstance(ls["col"], str):
                k = cols.index(ls["col"])
                conds.append((0, ls["comp"], [k], [k], ls["val"]))
            else:
                ids = [cols.index(c) for c in ls["col"]]
                conds.append((1, "<=", ids, ids, ls["val"]))      # the reference's ball query keeps d <= r for "<" too
        counts, off, idx = le.select_all(engine, table, locs, conds)
        np.testing.assert_array_equal(le.select_counts(engine, table, locs, conds), counts)
        for e in range(len(locs)):
            np.testing.assert_array_equal(idx[off[e]:off[e + 1]], g[f"{name}/{e}"], err_msg=f"{name}/{e}")
    m = np.load(os.path.join(gold, "driver_max_dist.npz"))
    for r in (100000, 200000):
        counts, off, idx = le.select_all(engine, m["grid"], m["refs"], [(1, "<", [0, 1], [0, 1], float(r))])
        for k in range(len(m["refs"])):
            np.testing.assert_array_equal(idx[off[k]:off[k + 1]], np.flatnonzero(m[f"{k}/{r}"]), err_msg=f"{k}/{r}")
