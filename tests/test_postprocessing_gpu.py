"""Post-processing reductions on the GPU (gpsat_gaussian_smooth, gpsat_glue_predictions) against the
reference's own outputs (tests/golden/post_*.npz) and the oracle; the smoothed-parameter, predict-only
pass of the batched driver (config c4 of SURVEY.md 8)."""
import os

import numpy as np
import pandas as pd
import pytest

from oracle import postprocessing as opp

pytestmark = pytest.mark.gpu


def test_gaussian_2d_weight_golden(engine, golden_dir):
    from gpsat_b200.postprocessing import gaussian_2d_weight
    g = np.load(os.path.join(golden_dir, "post_smooth.npz"))
    out = gaussian_2d_weight(g["x"], g["y"], g["x"], g["y"], float(g["l_x"]), float(g["l_y"]), g["vals"])
    np.testing.assert_allclose(out, g["out"], rtol=1e-12)
    assert np.isnan(gaussian_2d_weight([1e9], [1e9], g["x"], g["y"], 1.0, 1.0, g["vals"])[0])
    # several fields at once, all-NaN field, larger n
    rng = np.random.default_rng(0)
    n = 1500
    x, y = rng.uniform(-1, 1, n), rng.uniform(-1, 1, n)
    vals = rng.normal(size=(3, n))
    vals[1, rng.random(n) < 0.2] = np.nan
    vals[2] = np.nan
    out = gaussian_2d_weight(x, y, x, y, 0.3, 0.2, vals)
    for f in range(2):
        np.testing.assert_allclose(out[f], opp.gaussian_2d_weight(x, y, x, y, 0.3, 0.2, vals[f]), rtol=1e-11)
    assert np.all(np.isnan(out[2]))


def test_glue_golden(engine, golden_dir):
    from gpsat_b200.postprocessing import glue_local_predictions_1d, glue_local_predictions_2d
    q = np.load(os.path.join(golden_dir, "post_glue.npz"), allow_pickle=True)
    df = pd.DataFrame(q["df"], columns=list(q["cols"]))
    g2 = glue_local_predictions_2d(df, ["pred_loc_x", "pred_loc_y"], ["x", "y"], ["f*", "f*_var"], 150.0, R=3)
    assert list(g2.columns) == list(q["g2cols"])
    np.testing.assert_allclose(g2.values, q["g2"], rtol=1e-12)
    radius = dict(zip(q["radius_keys"].tolist(), q["radius_vals"].tolist()))
    g1 = glue_local_predictions_1d(df[df["y"] == 0.0], "pred_loc_x", "x", ["f*", "f*_var"], radius, R=2)
    assert list(g1.columns) == list(q["g1cols"])
    np.testing.assert_allclose(g1.values, q["g1"], rtol=1e-12)
    # shuffled rows give the same groups
    sh = df.sample(frac=1.0, random_state=3)
    g2s = glue_local_predictions_2d(sh, ["pred_loc_x", "pred_loc_y"], ["x", "y"], ["f*", "f*_var"], 150.0, R=3)
    np.testing.assert_allclose(g2s.values, q["g2"], rtol=1e-12)


def _oi(optimise_tables=None, suffix=""):
    from gpsat_b200.local_experts import LocalExpertOI
    rng = np.random.default_rng(12)
    n = 2500
    xy = rng.uniform(-2e5, 2e5, (n, 2))
    t = rng.integers(-2, 3, n).astype(float)
    obs = np.sin(xy[:, 0] / 8e4) + 0.5 * np.cos(xy[:, 1] / 6e4) + 0.05 * t + 0.1 * rng.standard_normal(n)
    data = pd.DataFrame({"x": xy[:, 0], "y": xy[:, 1], "t": t, "obs": obs})
    gx = np.arange(-1e5, 1.01e5, 1e5)
    locs = pd.DataFrame([(a, b, 0.0) for a in gx for b in gx], columns=["x", "y", "t"])
    pred = pd.DataFrame([(a, b) for a in np.arange(-1.5e5, 1.51e5, 5e4) for b in np.arange(-1.5e5, 1.51e5, 5e4)],
                        columns=["x", "y"])
    model = {"oi_model": "B200GPRModel",
             "init_params": {"coords_scale": [5e4, 5e4, 1], "obs_mean": "local"},
             "constraints": {"lengthscales": {"low": [1e-8, 1e-8, 1e-8], "high": [6e5, 6e5, 9]}}}
    if optimise_tables is not None:
        model["load_params"] = {"tables": optimise_tables, "table_suffix": suffix}
    return LocalExpertOI(locs, data,
                         {"obs_col": "obs", "coords_col": ["x", "y", "t"],
                          "local_select": [{"col": "t", "comp": "<=", "val": 2}, {"col": "t", "comp": ">=", "val": -2},
                                           {"col": ["x", "y"], "comp": "<", "val": 1.2e5}]},
                         model, {"df": pred, "max_dist": 1.0e5})


def test_smooth_then_predict_only_pass(engine):
    """fit pass -> smooth_hyperparameters -> predict-only pass with load_params from the _SMOOTHED
    tables (local_experts.py:553-689, postprocessing.py:96-395), then gluing of the overlapping
    predictions.  Checked against the oracle's smoothing and the per-expert model API."""
    from gpsat_b200.models import B200GPRModel
    from gpsat_b200.postprocessing import SmoothingConfig, glue_local_predictions_2d, smooth_hyperparameters
    oi = _oi()
    t1 = oi.run()
    assert len(t1["run_details"]) == 9 and t1["run_details"]["optimise_success"].all()
    coords = ["x", "y", "t"]
    cfg = {"lengthscales": SmoothingConfig(l_x=1e5, l_y=1e5, max=[10.0, 10.0, 8.0]),
           "kernel_variance": {"l_x": 1e5, "l_y": 1e5, "max": 0.8}}
    sm = smooth_hyperparameters(t1, ["lengthscales", "kernel_variance"], cfg, coords,
                                all_params=["lengthscales", "kernel_variance", "likelihood_variance"])
    assert set(sm) == {"lengthscales_SMOOTHED", "kernel_variance_SMOOTHED", "likelihood_variance_SMOOTHED"}
    ref = opp.smooth_table(t1["lengthscales"].reset_index(), "lengthscales", coords, ["x", "y"], 1e5, 1e5,
                           vmax=[10.0, 10.0, 8.0])
    a = sm["lengthscales_SMOOTHED"].sort_values("_dim_0", kind="stable").sort_index(kind="stable")
    b = ref.sort_values("_dim_0", kind="stable").sort_index(kind="stable")
    np.testing.assert_allclose(a["lengthscales"].values, b["lengthscales"].values, rtol=1e-12)
    refv = opp.smooth_table(t1["kernel_variance"].reset_index(), "kernel_variance", coords, ["x", "y"], 1e5, 1e5, vmax=0.8)
    np.testing.assert_allclose(sm["kernel_variance_SMOOTHED"].sort_index()["kernel_variance"].values,
                               refv.sort_index()["kernel_variance"].values, rtol=1e-12)
    pd.testing.assert_frame_equal(sm["likelihood_variance_SMOOTHED"], t1["likelihood_variance"])
    # predict-only pass
    oi2 = _oi(sm, "_SMOOTHED")
    t2 = oi2.run(optimise=False, table_suffix="_SMOOTHED")
    assert not t2["run_details_SMOOTHED"]["parameters_optimised"].any()
    ls2 = t2["lengthscales_SMOOTHED"].sort_values("_dim_0", kind="stable").sort_index(kind="stable")
    np.testing.assert_allclose(ls2["lengthscales"].values, a["lengthscales"].values, rtol=1e-14)
    # one expert by hand through the model API
    loc = (0.0, 0.0, 0.0)
    from gpsat_b200.local_experts import local_data_select, prediction_coords_for
    refloc = dict(zip(coords, loc))
    dfl = local_data_select(oi2.data, refloc, oi2.local_select)
    m = B200GPRModel(data=dfl, obs_col="obs", coords_col=coords, coords_scale=[5e4, 5e4, 1], obs_mean="local", verbose=False)
    m.set_parameters(lengthscales=sm["lengthscales_SMOOTHED"].loc[loc].sort_values("_dim_0")["lengthscales"].values,
                     kernel_variance=float(np.atleast_1d(sm["kernel_variance_SMOOTHED"].loc[loc]["kernel_variance"])[0]),
                     likelihood_variance=float(np.atleast_1d(sm["likelihood_variance_SMOOTHED"].loc[loc]["likelihood_variance"])[0]))
    pc = prediction_coords_for(oi2.pred_df, refloc, coords, 1.0e5)
    p = m.predict(pc)
    got = t2["preds_SMOOTHED"].loc[loc]
    np.testing.assert_allclose(got["f*"].values, p["f*"], rtol=1e-10, atol=1e-12)
    np.testing.assert_allclose(got["f*_var"].values, p["f*_var"], rtol=1e-10, atol=1e-12)
    # glue the overlapping predictions
    preds = t2["preds_SMOOTHED"].reset_index()
    glued = glue_local_predictions_2d(preds, ["pred_loc_x", "pred_loc_y"], ["x", "y"], ["f*", "f*_var"], 1.0e5, R=3)
    refg = opp.glue_2d(preds, ["pred_loc_x", "pred_loc_y"], ["x", "y"], ["f*", "f*_var"], 1.0e5, R=3)
    np.testing.assert_allclose(glued.values, refg.values, rtol=1e-11)
    assert len(glued) == preds[["pred_loc_x", "pred_loc_y"]].drop_duplicates().shape[0]


@pytest.mark.parametrize("oi_model,init", [("B200SGPRModel", {"num_inducing_points": 40}),
                                           ("GPflowVFFModel", {"num_inducing_features": [4, 4, 2],
                                                               "domain_size": [2.5e5, 2.5e5, 4.0]})])
def test_driver_runs_sparse_and_vff_models(engine, oi_model, init):
    """The batched driver with the other model families (oi_model by this package's or the reference's name)."""
    oi = _oi()
    from gpsat_b200.local_experts import resolve_model
    oi.model_cls = resolve_model(oi_model)
    oi.init_params.update(init)
    oi.constraints = None
    np.random.seed(0)
    t = oi.run()
    rd = t["run_details"]
    assert len(rd) == 9 and np.all(np.isfinite(rd["objective_value"])) and rd["model"].str.contains("B200").all()
    assert np.all(np.isfinite(t["preds"]["f*"])) and np.all(t["preds"]["f*_var"] > -1e-9)
    assert len(t["lengthscales"]) == 27
    if oi_model == "B200SGPRModel":
        ip = t["inducing_points"]
        assert list(ip.columns) == ["_dim_0", "_dim_1", "inducing_points"] and len(ip) == 9 * 40 * 3


def test_replacement_model_for_small_experts(engine):
    """local_experts.py:1021-1041: experts with fewer observations than replacement_threshold run the
    replacement model; the two families are batched separately and the tables keep expert order."""
    oi = _oi()
    from gpsat_b200.local_experts import resolve_model
    counts = []
    from gpsat_b200.local_experts import local_data_select
    for _, rl in oi.expert_locs.iterrows():
        counts.append(len(local_data_select(oi.data, {c: rl[c] for c in oi.coords_col}, oi.local_select)))
    thr = int(np.median(counts))
    oi.model_cls = resolve_model("B200SGPRModel")
    oi.init_params = dict(oi.init_params, num_inducing_points=60)
    oi.constraints = None
    oi.replacement_threshold = thr
    oi.replacement_cls = resolve_model("GPflowGPRModel")
    oi.replacement_init_params = {"coords_scale": [5e4, 5e4, 1], "obs_mean": "local"}
    oi.replacement_constraints = {"lengthscales": {"low": [1e-8, 1e-8, 1e-8], "high": [6e5, 6e5, 9]}}
    oi.replacement_optim_kwargs = {}
    oi.replacement_pred_kwargs = {}
    np.random.seed(1)
    t = oi.run()
    rd = t["run_details"]
    small = np.array(counts) < thr
    assert list(rd["num_obs"]) == counts
    assert all(("B200GPRModel" in mname) == s for mname, s in zip(rd["model"], small))
    assert np.all(np.isfinite(rd["objective_value"])) and np.all(np.isfinite(t["preds"]["f*"]))
    assert "inducing_points" in t and len(t["lengthscales"]) == 27
