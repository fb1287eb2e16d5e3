"""Driver-side code against the reference's own functions executed from source (oracle/make_golden.py --driver):
DataLoader.local_data_select (GPSat/dataloader.py:2354-2447), the _max_dist_bool radius filter
(GPSat/prediction_locations.py:18-43) and LocalExpertOI.dict_of_array_to_table (GPSat/local_experts.py:691-747).
CPU tests for the host restatements; the GPU selection kernel is held to the same vectors in test_models_gpu.py."""
import json
import os

import numpy as np
import pandas as pd

from gpsat_b200 import local_experts as le

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _frame_from_json(j):
    df = pd.DataFrame({c: pd.Series(v).astype(t) for c, v, t in zip(j["columns"], j["data"], j["dtypes"])})
    if len(j["index_names"]) == 1:
        df.index = pd.Index(j["index"][0], name=j["index_names"][0])
    else:
        df.index = pd.MultiIndex.from_arrays(j["index"], names=j["index_names"])
    return df


def load_select_golden():
    g = np.load(os.path.join(GOLD, "driver_select.npz"))
    df = pd.DataFrame(g["data"], columns=["x", "y", "t", "obs"])
    locs = pd.DataFrame(g["locs"], columns=["x", "y", "t"])
    return g, df, locs, json.loads(str(g["selects"]))


def test_local_data_select_equals_the_reference_function():
    g, df, locs, selects = load_select_golden()
    for name, ls in selects.items():
        for e in range(len(locs)):
            got = le.local_data_select(df, locs.iloc[e].to_dict(), ls)
            np.testing.assert_array_equal(got.index.values, g[f"{name}/{e}"], err_msg=f"{name}/{e}")
    assert len(g["example/0"]) > 0 and len(g["example/5"]) == 0
    # the rows placed exactly on the ball / on the time-window edges are decided as the reference decides them
    assert 0 in g["example/0"] and 1 in g["example/0"] and 2 not in g["example/0"]


def test_prediction_radius_filter_equals_the_reference_numba_kernel():
    g = np.load(os.path.join(GOLD, "driver_max_dist.npz"))
    grid = pd.DataFrame(g["grid"], columns=["x", "y"])
    for k, ref in enumerate(g["refs"]):
        for r in (100000, 200000):
            want = g["grid"][g[f"{k}/{r}"]]
            got = le.prediction_coords_for(grid, {"x": ref[0], "y": ref[1], "t": 5.0}, ["x", "y", "t"], max_dist=float(r))
            np.testing.assert_array_equal(got[:, :2], want)
            assert np.all(got[:, 2] == 5.0)                  # missing coordinate filled from the expert location


def test_dict_to_tables_equals_the_reference_function():
    gold = json.load(open(os.path.join(GOLD, "driver_tables.json")))
    ref_loc = gold["ref_loc"]
    inp = gold["inputs"]
    pred = {k: np.asarray(v) for k, v in inp["pred"].items()}
    hypes = {k: (np.asarray(v) if isinstance(v, list) else v) for k, v in inp["hypes"].items()}
    ours = {"preds": le.dict_to_tables(pred, ref_loc, "preds"),
            "run_details": le.dict_to_tables(inp["run_details"], ref_loc, "run_details"),
            "hypes": le.dict_to_tables(hypes, ref_loc, None),
            "one_d_ref": le.dict_to_tables({"kernel_variance": 0.5}, {"x": 2.0}, None)}
    for grp, tabs in gold["tables"].items():
        assert sorted(ours[grp]) == sorted(tabs), grp
        for k, j in tabs.items():
            want = _frame_from_json(j)
            got = ours[grp][k]
            assert list(got.columns) == list(want.columns), (grp, k, list(got.columns))
            pd.testing.assert_frame_equal(got, want, check_exact=True, check_index_type=False)
