"""Config dataclasses, the JSON layout of the reference's config files, the reference-form constructor of
LocalExpertOI and the rank-merge of result tables - host logic, no GPU."""
import json

import numpy as np
import pandas as pd
import pytest

from gpsat_b200 import config_dataclasses as cd
from gpsat_b200.local_experts import LocalExpertOI, merge_rank_tables

# the key layout of configs/example_local_expert_oi.json (values shortened)
EXAMPLE = {
    "comment": ["example configuration for running local expert on example data."],
    "run_kwargs": {"store_path": "out_dir", "table_suffix": "", "store_every": 10, "optimise": True,
                   "check_config_compatible": False},
    "locations": {"source": "locs.csv",
                  "add_data_to_col": {"date": ["2020-03-04", "2020-03-05", "2020-03-06"]},
                  "col_funcs": {"date": {"func": "lambda x: x.astype('datetime64[D]')", "col_args": "date"},
                                "t": {"func": "lambda x: x.astype('datetime64[D]').astype(float)", "col_args": "date"}},
                  "col_select": ["x", "y", "t", "date"],
                  "row_select": [{"col": "date", "comp": "==", "val": "2020-03-05"}, {"col": "y", "comp": ">=", "val": 0}],
                  "sort_by": "date"},
    "data": {"data_source": "obs.csv", "table": "data", "obs_col": "obs", "coords_col": ["x", "y", "t"],
             "local_select": [{"col": "t", "comp": "<=", "val": 4}, {"col": "t", "comp": ">=", "val": -4},
                              {"col": ["x", "y"], "comp": "<", "val": 300000}],
             "global_select": [{"col": "y", "comp": ">=", "val": -1e9},
                               {"loc_col": "t", "src_col": "date", "func": "lambda x,y: np.datetime64(pd.to_datetime(x+y, unit='D'))"}],
             "row_select": None, "col_select": None, "col_funcs": None},
    "model": {"oi_model": "GPflowGPRModel", "init_params": {"coords_scale": [50000, 50000, 1]},
              "constraints": {"lengthscales": {"low": [1e-08, 1e-08, 1e-08], "high": [600000, 600000, 9]}}},
    "pred_loc": {"method": "from_dataframe", "df_file": "grid.csv", "max_dist": 200000},
}


def test_experiment_config_round_trips_the_reference_json_layout():
    cfg = cd.ExperimentConfig.from_dict(EXAMPLE)
    assert isinstance(cfg.data_config, cd.DataConfig) and cfg.data_config.obs_col == "obs"
    assert cfg.model_config.oi_model == "GPflowGPRModel" and cfg.model_config.params_to_store == "all"
    assert cfg.expert_locs_config.sort_by == "date" and cfg.prediction_locs_config.max_dist == 200000
    assert cfg.run_config.store_every == 10 and cfg.run_config.min_obs == 3            # default filled in
    d = cfg.to_dict()
    assert set(d) == {"data", "model", "locations", "pred_loc", "run_kwargs", "comment"}   # config_dataclasses.py:591-595
    for key in ("data", "model", "locations", "pred_loc", "run_kwargs"):
        for k, v in EXAMPLE[key].items():
            assert d[key][k] == v, (key, k)
    again = cd.ExperimentConfig.from_json(cfg.to_json())
    assert again == cfg
    with pytest.raises(KeyError):
        cd.ModelConfig.from_dict({"oi_modle": "typo"})


def test_dataframes_are_stored_as_dicts_and_come_back():
    df = pd.DataFrame({"x": [0.0, 1.0], "y": [2.0, 3.0], "obs": [0.1, 0.2]})
    dc = cd.DataConfig(data_source=df, obs_col="obs", coords_col=["x", "y"])
    assert isinstance(dc.data_source, dict)
    json.dumps(dc.to_dict())                                                            # JSON-serialisable
    back = dc.to_dict_with_dataframe()["data_source"]
    pd.testing.assert_frame_equal(back, df)
    pd.testing.assert_frame_equal(dc.load(), df)
    el = cd.ExpertLocsConfig(source=df[["x", "y"]])
    assert isinstance(el.source, dict) and isinstance(el.to_dict_with_dataframe()["source"], pd.DataFrame)
    pl = cd.PredictionLocsConfig(method="from_dataframe", df=df[["x", "y"]], max_dist=10)
    assert isinstance(pl.df, dict) and isinstance(pl.to_dict_with_dataframe()["df"], pd.DataFrame)
    with pytest.raises(AssertionError):
        cd.PredictionLocsConfig(method="nonsense")


def test_load_table_follows_the_reference_order(tmp_path):
    locs = pd.DataFrame({"x": [0.0, 1.0, 2.0], "y": [-1.0, 1.0, 2.0]})
    path = tmp_path / "locs.csv"
    locs.to_csv(path, index=False)
    out = cd.load_table(str(path), **{k: v for k, v in EXAMPLE["locations"].items() if k not in ("source", "sort_by")})
    # 3 rows x 3 dates, then date == 2020-03-05 and y >= 0
    assert list(out.columns) == ["x", "y", "t", "date"] and len(out) == 2
    assert np.all(out["t"].values == float(np.datetime64("2020-03-05").astype("datetime64[D]").astype(int)))
    w = cd.load_table(locs, where={"col": "x", "comp": ">=", "val": 1})                 # DataLoader.load docstring example
    assert list(w["x"]) == [1.0, 2.0]
    with pytest.raises(NotImplementedError):
        cd.load_table("data.nc")


def _files(tmp_path, seed=0):
    rng = np.random.default_rng(seed)
    t0 = float(np.datetime64("2020-03-05").astype("datetime64[D]").astype(int))
    obs = pd.DataFrame({"x": rng.uniform(-4e5, 4e5, 4000), "y": rng.uniform(-4e5, 4e5, 4000),
                        "t": t0 + rng.integers(-6, 7, 4000).astype(float)})
    obs["obs"] = np.sin(obs["x"] / 1e5) + 0.1 * rng.standard_normal(len(obs))
    locs = pd.DataFrame({"x": [-1e5, 0.0, 1e5, 2e5], "y": [-1e5, 1e5, 0.0, -5e5]})
    g = np.linspace(-3e5, 3e5, 13)
    grid = pd.DataFrame({"x": np.repeat(g, 13), "y": np.tile(g, 13)})
    obs.to_csv(tmp_path / "obs.csv", index=False)
    locs.to_csv(tmp_path / "locs.csv", index=False)
    grid.to_csv(tmp_path / "grid.csv", index=False)
    cfg = json.loads(json.dumps(EXAMPLE))
    cfg["locations"]["source"] = str(tmp_path / "locs.csv")
    cfg["locations"]["row_select"][1]["val"] = -2e5
    cfg["data"]["data_source"] = str(tmp_path / "obs.csv")
    cfg["pred_loc"]["df_file"] = str(tmp_path / "grid.csv")
    cfg["run_kwargs"]["store_path"] = str(tmp_path / "out")
    return cfg, obs, locs, grid


def test_reference_form_constructor_loads_the_three_tables(tmp_path):
    cfg, obs, locs, grid = _files(tmp_path)
    oi = LocalExpertOI(expert_loc_config=cfg["locations"], data_config=cfg["data"], model_config=cfg["model"],
                       pred_loc_config=cfg["pred_loc"])
    assert len(oi.expert_locs) == 3 and list(oi.expert_locs.columns) == ["x", "y", "t", "date"]   # y >= -2e5
    assert len(oi.data) == len(obs) and oi.obs_col == "obs" and oi.coords_col == ["x", "y", "t"]
    assert len(oi.pred_df) == len(grid) and oi.max_dist == 200000
    assert oi.model_cls.__name__ == "B200GPRModel" and oi.vectorisable()
    assert set(oi.config) == {"locations", "data", "model", "pred_loc"}
    json.dumps(oi.config)
    # the dataclass form and the ExperimentConfig form build the same object
    ec = cd.ExperimentConfig.from_dict(cfg)
    oi2 = LocalExpertOI(local_expert_config=ec)
    pd.testing.assert_frame_equal(oi2.expert_locs, oi.expert_locs)
    oi3 = LocalExpertOI(cd.ExpertLocsConfig(**cfg["locations"]), cd.DataConfig(**cfg["data"]), cd.ModelConfig(**cfg["model"]),
                        cd.PredictionLocsConfig(**cfg["pred_loc"]))
    pd.testing.assert_frame_equal(oi3.data, oi.data)
    # expert_loc method: predict at the expert location
    oi4 = LocalExpertOI(cfg["locations"], cfg["data"], cfg["model"], None)
    assert oi4.pred_df is None


def test_merge_rank_tables_restores_single_gpu_order():
    idx = lambda ids: pd.Index(np.asarray(ids, float), name="x")
    # single-GPU tables of 6 experts: experts 1 and 4 have too few observations
    keep, small = np.array([0, 2, 3, 5]), np.array([1, 4])
    rd = pd.DataFrame({"num_obs": [1, 2, 10, 12, 13, 15], "run_time": [np.nan, np.nan, .1, .1, .1, .1]},
                      index=idx(np.concatenate([small, keep])))
    ls = pd.DataFrame({"_dim_0": np.tile([0, 1], 4), "lengthscales": np.arange(8.0)}, index=idx(np.repeat(keep, 2)))
    parts = [np.array([0, 3, 4]), np.array([1, 2, 5])]
    gathered = []
    for p in parts:
        rd_rows = [k for k, e in enumerate(np.concatenate([small, keep])) if e in p]
        ls_rows = [k for k, e in enumerate(np.repeat(keep, 2)) if e in p]
        gathered.append({"tables": {"run_details": rd.iloc[rd_rows], "lengthscales": ls.iloc[ls_rows]},
                         "row_expert": {"run_details": np.concatenate([small, keep])[rd_rows],
                                        "lengthscales": np.repeat(keep, 2)[ls_rows]}})
    out = merge_rank_tables(gathered)
    pd.testing.assert_frame_equal(out["run_details"], rd)
    pd.testing.assert_frame_equal(out["lengthscales"], ls)
