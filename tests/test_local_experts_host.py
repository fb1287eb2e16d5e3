"""Selection and table semantics of the batched driver (CPU only), against the reference's rules:
GPSat/dataloader.py:2403-2447 (local_data_select), prediction_locations.py:18-43, 208-273,
local_experts.py:691-747 (dict_of_array_to_table)."""
import numpy as np
import pandas as pd

from gpsat_b200.local_experts import dict_to_tables, local_data_select, prediction_coords_for


def _obs(n=400, seed=0):
    rng = np.random.default_rng(seed)
    return pd.DataFrame({"x": rng.uniform(-5, 5, n), "y": rng.uniform(-5, 5, n),
                         "t": rng.integers(-9, 10, n).astype(float), "obs": rng.normal(size=n)})


def test_local_data_select_matches_brute_force():
    df = _obs()
    ref = {"x": 0.5, "y": -0.5, "t": 1.0}
    ls = [{"col": "t", "comp": "<=", "val": 4}, {"col": "t", "comp": ">=", "val": -4},
          {"col": ["x", "y"], "comp": "<", "val": 3.0}]
    out = local_data_select(df, ref, ls)
    # comparisons are against reference + val; the ball query includes its boundary (KDTree semantics)
    keep = (df.t <= 5) & (df.t >= -3) & (((df.x - 0.5) ** 2 + (df.y + 0.5) ** 2) <= 9.0)
    assert out.index.equals(df[keep].index) and 0 < len(out) < len(df)
    assert local_data_select(df, ref, [{"col": "t", "comp": "==", "val": 0}]).t.eq(1.0).all()


def test_prediction_coords_radius_and_fill():
    ref = {"x": 0.5, "y": -0.5, "t": 3.0}
    grid = pd.DataFrame({"x": np.repeat(np.arange(-5, 6.0), 11), "y": np.tile(np.arange(-5, 6.0), 11), "junk": 1.0})
    pc = prediction_coords_for(grid, ref, ["x", "y", "t"], max_dist=2.0)
    d2 = (grid.x - 0.5) ** 2 + (grid.y + 0.5) ** 2
    exp = grid[d2 < 4.0]
    np.testing.assert_array_equal(pc[:, :2], exp[["x", "y"]].values)
    assert np.all(pc[:, 2] == 3.0)                       # missing 't' comes from the expert location
    # strict inequality: a point at exactly max_dist is excluded (prediction_locations.py:31-43)
    one = pd.DataFrame({"x": [2.5], "y": [-0.5]})
    assert len(prediction_coords_for(one, ref, ["x", "y", "t"], max_dist=2.0)) == 0
    np.testing.assert_array_equal(prediction_coords_for(None, ref, ["x", "y", "t"]), [[0.5, -0.5, 3.0]])
    assert len(prediction_coords_for(grid, ref, ["x", "y", "t"])) == len(grid)


def test_tables_shape_and_index():
    ref = {"x": 1.0, "y": 2.0, "t": 3.0}
    t = dict_to_tables({"f*": np.arange(4.0), "f*_var": np.ones(4), "pred_loc_x": np.zeros(4)}, ref, "preds")
    # the reference's layout (tests/test_driver_golden.py): a leading _dim_0 column in every table
    assert list(t) == ["preds"] and t["preds"].shape == (4, 4) and list(t["preds"]["_dim_0"]) == [0, 1, 2, 3]
    assert t["preds"].index.names == ["x", "y", "t"] and t["preds"].index[0] == (1.0, 2.0, 3.0)
    rd = dict_to_tables({"num_obs": 7, "optimise_success": True, "model": "m"}, ref, "run_details")["run_details"]
    assert rd.shape == (1, 4) and rd["num_obs"].iloc[0] == 7 and rd["_dim_0"].iloc[0] == 0
    h = dict_to_tables({"lengthscales": np.array([1.0, 2.0, 3.0]), "kernel_variance": 0.5}, ref)
    assert list(h["lengthscales"].columns) == ["_dim_0", "lengthscales"] and len(h["lengthscales"]) == 3
    assert list(h["kernel_variance"].columns) == ["_dim_0", "kernel_variance"] and len(h["kernel_variance"]) == 1
