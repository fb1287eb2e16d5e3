"""Batched local-expert driver: the loop body of ``LocalExpertOI.run`` for all experts at once.

The reference processes experts serially (``GPSat/local_experts.py:930-1260``): select local data,
build a model, apply constraints, optimise, read the objective and the hyper-parameters, predict,
append rows to result tables.  Experts are independent, so this driver runs the same steps in
three phases:

* phase A (host): per expert, the reference's selection semantics -
  ``DataLoader.local_data_select`` (``GPSat/dataloader.py:2354-2447``: column-wise comparisons
  against ``reference_location[col] + val`` and KD-tree ball queries for multi-column entries) and
  the ``max_dist`` radius filter of ``PredictionLocations`` (``GPSat/prediction_locations.py:18-43,
  208-273``, missing coordinate columns filled from the expert location); experts with fewer than
  ``min_obs`` observations get a NaN ``run_details`` row (``local_experts.py:988-1012``), experts
  without prediction locations are skipped (``:962-965``);
* phase B (device): one ``gpsat_gpr_fit`` + one ``gpsat_gpr_predict`` call for all remaining
  experts (:func:`gpsat_b200.models.fit_predict_models`);
* phase C (host): the same tables the reference writes - ``preds``, ``run_details`` and one table
  per hyper-parameter, indexed by the expert location (``dict_of_array_to_table``,
  ``local_experts.py:691-747``; ``lengthscales`` carries a ``_dim_0`` column,
  ``GPSat/utils.py:1437-1497``), returned as pandas DataFrames and appended to an HDF5 store when
  ``store_path`` is given and PyTables is importable.

Only what sits on either side of the engine is restated here; config parsing, data loading from
files, resume logic and plotting stay in GPSat (out of scope).
"""
from __future__ import annotations

import time
from typing import Dict, List, Optional

import numpy as np
import pandas as pd
from scipy.spatial import cKDTree

import importlib
import warnings

from .models import B200GPRModel, B200SGPRModel, B200VFFModel, fit_predict_models

# model names accepted in ``model_config["oi_model"]``: this package's classes, and the names of the
# reference classes they stand in for (``GPSat/models/__init__.py:3-28``)
_MODEL_CLASSES = {"B200GPRModel": B200GPRModel, "B200SGPRModel": B200SGPRModel, "B200VFFModel": B200VFFModel,
                  "GPflowGPRModel": B200GPRModel, "GPflowSGPRModel": B200SGPRModel, "GPflowVFFModel": B200VFFModel}


def resolve_model(oi_model):
    """``LocalExpertOI.set_model`` (``local_experts.py:309-330``): a name, a class, or
    ``{"path_to_model": module, "model_name": class}``; the class must derive from ``B200GPRModel``
    (the batched engine calls need its parameter layout)."""
    if isinstance(oi_model, dict):
        cls = getattr(importlib.import_module(oi_model["path_to_model"]), oi_model["model_name"])
    elif isinstance(oi_model, str):
        if oi_model not in _MODEL_CLASSES:
            raise NotImplementedError(f"oi_model '{oi_model}' is not available in the batched driver; "
                                      f"choose from {sorted(_MODEL_CLASSES)}")
        cls = _MODEL_CLASSES[oi_model]
    else:
        cls = oi_model
    assert isinstance(cls, type) and issubclass(cls, B200GPRModel), "the batched driver runs B200GPRModel subclasses"
    return cls

_COMPS = {">=": np.greater_equal, ">": np.greater, "==": np.equal, "<": np.less, "<=": np.less_equal}


def local_data_select(df: pd.DataFrame, reference_location: dict, local_select: List[dict],
                      kdtree: Optional[list] = None) -> pd.DataFrame:
    """Rows of ``df`` that satisfy every entry of ``local_select`` relative to the expert location
    (same semantics as ``DataLoader.local_data_select``, ``GPSat/dataloader.py:2403-2447``)."""
    select = np.ones(len(df), dtype=bool)
    for idx, ls in enumerate(local_select):
        col, comp = ls["col"], ls["comp"]
        if isinstance(col, str):
            assert col in df, f"col: {col} is not in data - {df.columns}"
            assert col in reference_location, f"col: {col} is not in reference_location - {reference_location.keys()}"
            assert comp in _COMPS, f"comp: {comp} is not valid"
            select &= _COMPS[comp](df[col].values, reference_location[col] + ls["val"])
        else:
            assert comp in ["<", "<="], "for multi dimensional values only less than comparison handled"
            for c in col:
                assert c in df, f"column: {c} is not in df.columns: {df.columns}"
                assert c in reference_location, f"col: {col} is not in reference_location - {reference_location.keys()}"
            kdt = None
            if kdtree is not None:
                kdt = kdtree[idx] if isinstance(kdtree, list) else kdtree
            if kdt is None:
                kdt = cKDTree(df.loc[:, col].values)
            in_ids = kdt.query_ball_point(x=[reference_location[c] for c in col], r=ls["val"])
            tmp = np.zeros(len(df), dtype=bool)
            tmp[in_ids] = True
            select &= tmp
    return df.loc[select, :]


def prediction_coords_for(pred_df: Optional[pd.DataFrame], expert_loc: dict, coords_col: List[str],
                          max_dist: Optional[float] = None, tree: Optional[cKDTree] = None,
                          tree_cols: Optional[List[str]] = None) -> np.ndarray:
    """Prediction coordinates of one expert: ``pred_df`` rows within ``max_dist`` of the expert
    (Euclidean over the columns ``pred_df`` shares with ``coords_col``), columns missing from
    ``pred_df`` filled with the expert's own value; ``pred_df=None`` predicts at the expert
    location itself (``GPSat/prediction_locations.py:208-273``)."""
    if pred_df is None:
        return np.array([[expert_loc[c] for c in coords_col]], dtype=np.float64)
    shared = tree_cols if tree_cols is not None else [c for c in coords_col if c in pred_df]
    sel = pred_df
    if max_dist is not None:
        if tree is None:
            tree = cKDTree(pred_df[shared].values)
        ref = np.array([expert_loc[c] for c in shared], dtype=np.float64)
        ids = np.array(sorted(tree.query_ball_point(ref, r=max_dist)), dtype=np.int64)
        if len(ids):     # the reference keeps d^2 < max_dist^2, strictly (prediction_locations.py:31-43)
            d2 = np.sum((pred_df[shared].values[ids] - ref) ** 2, axis=1)
            ids = ids[d2 < max_dist * max_dist]
        sel = pred_df.iloc[ids]
    out = np.empty((len(sel), len(coords_col)), dtype=np.float64)
    for k, c in enumerate(coords_col):
        out[:, k] = sel[c].values if c in sel else expert_loc[c]
    return out


def dict_to_tables(x: dict, ref_loc: dict, table: Optional[str] = None) -> Dict[str, pd.DataFrame]:
    """``dict_of_array_to_table`` of the reference (``local_experts.py:691-747``) for the shapes this path produces
    (pinned on the reference function executed from source: ``tests/golden/driver_tables.json``).
    ``table`` given (the reference's ``concat=True``, used for ``preds`` and ``run_details``, ``:1226-1233``): scalars
    and 1-d arrays become the columns of one frame, outer-joined on their position, which appears as the leading
    ``_dim_0`` column.  ``table=None`` (``concat=False``, the hyper-parameters, ``:1234-1236``): one frame per entry
    with one ``_dim_i`` column per axis (scalars count as length-1 arrays: ``_dim_0`` = 0; ``array_to_dataframe``,
    ``GPSat/utils.py:1437-1497``).  Frames are indexed by the expert location."""
    names = list(ref_loc.keys())
    vals = tuple(ref_loc.values())

    def index(nrow):
        if len(names) == 1:
            return pd.Index([vals[0]] * nrow, name=names[0])
        return pd.MultiIndex.from_tuples([vals] * nrow, names=names)

    def frame(k, v):
        v = np.array([v]) if isinstance(v, (int, float, bool, str)) else np.asarray(v)
        if v.ndim == 0:
            v = v.reshape(1)
        midx = pd.MultiIndex.from_product([np.arange(i) for i in v.shape], names=[f"_dim_{i}" for i in range(v.ndim)])
        return pd.DataFrame(v.flat, index=midx, columns=[k])

    if len(x) == 0:
        return {}
    if table is not None:
        df = pd.concat([frame(k, v) for k, v in x.items()], join="outer", axis=1).reset_index()
        df.index = index(len(df))
        return {table: df}
    out = {}
    for k, v in x.items():
        df = frame(k, v).reset_index()
        df.index = index(len(df))
        out[k] = df
    return out


class LocalExpertOI:
    """Batched counterpart of ``GPSat.local_experts.LocalExpertOI`` for in-memory data.

    ``expert_locs``: DataFrame with (at least) the coordinate columns; ``data``: DataFrame of
    observations; the config dicts use the reference's keys - ``data_config``: ``obs_col``,
    ``coords_col``, ``local_select``; ``model_config``: ``init_params``, ``constraints``,
    ``optim_kwargs``, ``pred_kwargs``, ``params_to_store``, ``replacement_threshold`` /
    ``replacement_model`` / ``replacement_init_params`` / ``replacement_constraints`` /
    ``replacement_optim_kwargs`` / ``replacement_pred_kwargs`` (another model for experts with fewer
    observations than the threshold, ``local_experts.py:1021-1041``; each family is one batch); ``pred_loc_config``: ``df`` (or ``None``
    to predict at the expert location) and ``max_dist`` (``config_dataclasses.py:9-509``).
    """

    def __init__(self, expert_loc_config=None, data_config=None, model_config=None, pred_loc_config=None,
                 local_expert_config=None):
        """Two call forms.

        Reference form (``GPSat/local_experts.py:142-222``):
        ``LocalExpertOI(expert_loc_config, data_config, model_config, pred_loc_config)`` with dicts or the
        dataclasses of :mod:`gpsat_b200.config_dataclasses` (or ``local_expert_config=ExperimentConfig``);
        expert locations, observations and prediction locations are loaded through
        :func:`gpsat_b200.config_dataclasses.load_table` (DataFrame / dict / csv / tsv / parquet / h5 sources).

        In-memory form: ``LocalExpertOI(expert_locs_df, data_df, data_config, model_config, pred_loc_config)`` with
        two DataFrames first."""
        from . import config_dataclasses as cd
        if isinstance(expert_loc_config, pd.DataFrame) and isinstance(data_config, pd.DataFrame):
            expert_locs, data = expert_loc_config, data_config
            data_config, model_config, pred_loc_config = model_config, pred_loc_config, local_expert_config
            self.config = {}
        else:
            if local_expert_config is not None:
                expert_loc_config = local_expert_config.expert_locs_config
                data_config = local_expert_config.data_config
                model_config = local_expert_config.model_config
                pred_loc_config = local_expert_config.prediction_locs_config
            plain = lambda c: c.to_dict_with_dataframe() if hasattr(c, "to_dict_with_dataframe") else (
                c.to_dict() if hasattr(c, "to_dict") else dict(c or {}))
            lc, dc, mc_, pc_ = plain(expert_loc_config), plain(data_config), plain(model_config), plain(pred_loc_config)
            # what the reference stores as the "oi_config" of a run (local_experts.py:170-174, 873-884)
            self.config = {"locations": cd._plain(lc), "data": cd._plain(dc), "model": cd._plain(mc_), "pred_loc": cd._plain(pc_)}
            expert_locs = self._load_expert_locations(**lc)
            data = cd.DataConfig(**{k: v for k, v in dc.items() if k in cd.DataConfig.__dataclass_fields__}).load()
            data_config, model_config = dc, mc_
            pred_loc_config = self._load_pred_loc(**pc_)
        self.expert_locs = expert_locs.reset_index(drop=True)
        self.data = data
        self.obs_col = data_config["obs_col"]
        self.coords_col = list(data_config["coords_col"])
        self.local_select = data_config.get("local_select") or []
        mc = dict(model_config or {})
        self.model_cls = resolve_model(mc.get("oi_model") or "B200GPRModel")
        self.load_params = mc.get("load_params")
        # replacement model for experts with few observations (local_experts.py:1021-1041)
        self.replacement_threshold = mc.get("replacement_threshold")
        if self.replacement_threshold is not None:
            self.replacement_cls = resolve_model(mc.get("replacement_model") or "B200GPRModel")
            self.replacement_init_params = dict(mc.get("replacement_init_params") or {})
            self.replacement_constraints = mc.get("replacement_constraints")
            self.replacement_optim_kwargs = dict(mc.get("replacement_optim_kwargs") or {})
            self.replacement_pred_kwargs = dict(mc.get("replacement_pred_kwargs") or {})
        self.init_params = dict(mc.get("init_params") or {})
        self.constraints = mc.get("constraints")
        self.optim_kwargs = dict(mc.get("optim_kwargs") or {})
        self.pred_kwargs = dict(mc.get("pred_kwargs") or {})
        pts = mc.get("params_to_store") or "all"
        self.params_to_store = None if pts == "all" else pts
        pc = dict(pred_loc_config or {})
        self.pred_df = pc.get("df")
        self.max_dist = pc.get("max_dist")

    @staticmethod
    def _load_expert_locations(df=None, file=None, source=None, where=None, add_data_to_col=None, col_funcs=None,
                               keep_cols=None, col_select=None, row_select=None, sort_by=None, reset_index=False,
                               source_kwargs=None, verbose=False, **kwargs) -> pd.DataFrame:
        """``LocalExpertOI.set_expert_locations`` (``local_experts.py:349-422``), legacy inputs included."""
        from .config_dataclasses import load_table
        if col_select is None and keep_cols is not None:
            warnings.warn("\n'keep_cols' provided to set_expert_locations, use 'col_select' instead")
            col_select = keep_cols
        if source is None and df is not None:
            warnings.warn("\n'df' was provided to set_expert_locations, use 'source' instead")
            source = df
        if source is None and file is not None:
            warnings.warn("\n'file' was provided to set_expert_locations, use 'source' instead")
            source = file
        assert source is not None, "expert locations: no 'source' given"
        locs = load_table(source, where=where, source_kwargs=source_kwargs, col_funcs=col_funcs, row_select=row_select,
                          col_select=col_select, reset_index=reset_index, add_data_to_col=add_data_to_col, verbose=verbose)
        if sort_by:
            locs = locs.sort_values(sort_by)
        return locs

    @staticmethod
    def _load_pred_loc(method="expert_loc", coords_col=None, df=None, df_file=None, max_dist=None, load_kwargs=None,
                       X_out=None, **kwargs) -> dict:
        """``PredictionLocations`` set-up (``GPSat/prediction_locations.py:50-206``): ``expert_loc`` predicts at the
        expert location, ``from_dataframe`` / ``from_source`` at the rows of a table within ``max_dist``."""
        from .config_dataclasses import load_table
        if method == "expert_loc":
            return {"df": None, "max_dist": None}
        assert method in ("from_dataframe", "from_source"), f"pred_loc method '{method}' is not handled"
        lk = dict(load_kwargs or {})
        src = df if df is not None else (df_file if df_file is not None else lk.pop("source", None))
        assert src is not None, "pred_loc: method is from_dataframe / from_source but no df, df_file or load_kwargs.source"
        assert max_dist is not None, "pred_loc: max_dist is required with a prediction-location table"
        return {"df": load_table(src, **lk), "max_dist": max_dist}

    @staticmethod
    def _scaled_constraints(constraints, init_params):
        """The driver forces ``scale=True`` on the lengthscale bounds when ``coords_scale`` is given
        (``local_experts.py:1110-1114``)."""
        if constraints is None:
            return None
        cons = {k: dict(v) for k, v in constraints.items()}
        if "coords_scale" in init_params and "lengthscales" in cons:
            cons["lengthscales"]["scale"] = True
        return cons

    def _constraints_for_model(self):
        return self._scaled_constraints(self.constraints, self.init_params)

    # ---- load_params (local_experts.py:553-689) ---------------------------------------------------
    def _param_lookup(self, param_names):
        """{param: {expert location tuple: values ordered by _dim_*}} from the configured tables."""
        lp = self.load_params
        if not lp:
            return None
        if lp.get("index_adjust"):
            raise NotImplementedError("load_params.index_adjust is not implemented in the batched driver")
        suffix = lp.get("table_suffix", "")
        names = lp.get("param_names") or param_names
        for pn in names:
            assert pn in param_names, f"provide param name:{pn}\nis not in param_names:{param_names}"
        tables = lp.get("tables")
        if tables is None:
            assert isinstance(lp.get("file"), str), "load_params needs 'tables' or 'file'"
            with pd.HDFStore(lp["file"], mode="r") as store:              # needs PyTables, like the reference
                tables = {f"{k}{suffix}": store.select(f"{k}{suffix}") for k in names if f"/{k}{suffix}" in store.keys()}
        lookup = {}
        for k in names:
            df = tables.get(f"{k}{suffix}")
            if df is None:
                print(f"KeyError\n 'No object named {k}{suffix}'\nskipping param_name: {k}")
                continue
            df = df.reset_index(drop=False) if df.index.names != [None] else df
            keys = [c for c in self.coords_col if c in df.columns]
            dim_cols = sorted(c for c in df.columns if c.startswith("_dim_"))
            per = {}
            for loc, sub in df.groupby(keys, sort=False):
                loc = loc if isinstance(loc, tuple) else (loc,)
                sub = sub.sort_values(dim_cols) if dim_cols else sub
                v = sub[k].values.astype(np.float64)
                if len(dim_cols) > 1:                  # n-d parameter: back to its array shape
                    v = v.reshape([int(sub[c].max()) + 1 for c in dim_cols])
                per[tuple(float(x) for x in loc)] = v if dim_cols else float(v[0])
            lookup[k] = (keys, per)
        return lookup

    @staticmethod
    def _params_for(lookup, ref: dict) -> dict:
        out = {}
        for k, (keys, per) in lookup.items():
            v = per.get(tuple(float(ref[c]) for c in keys))
            if v is None:
                warnings.warn(f"\n******\nno parameters found in table:\n{k}\nfor location:\n{ref}\n******")
                continue
            if np.any(np.isnan(v)):
                warnings.warn(f"\n{k}: found some nans for ref location: {ref}, removing those parameters")
                continue
            out[k] = v
        return out

    def vectorisable(self) -> bool:
        """The array-built path covers the exact model without a replacement model."""
        return self.model_cls is B200GPRModel and self.replacement_threshold is None

    def run(self, store_path: Optional[str] = None, optimise: bool = True, predict: bool = True,
            min_obs: int = 3, table_suffix: str = "", config_id: int = 1,
            device_select: Optional[bool] = None, store_every: int = 10, check_config_compatible: bool = True,
            skip_valid_checks_on: Optional[list] = None, world: Optional[bool] = None) -> Dict[str, pd.DataFrame]:
        """All experts: select, fit, predict, tabulate.  Returns ``{table name: DataFrame}`` with the
        reference's tables (``preds``, ``run_details``, one per stored hyper-parameter), each name
        extended by ``table_suffix``.  ``device_select=True`` evaluates the selection rules of all
        experts in one GPU launch (``gpsat_local_select``) and builds the packed batch and the
        tables with array operations instead of one model instance and one DataFrame per expert
        (exact-GPR model only; ``None``, the default, = whenever the configuration allows it; ``False`` forces the
        per-model path); the tables of the two paths are identical.

        ``store_every`` / ``check_config_compatible`` / ``skip_valid_checks_on`` are the reference's remaining
        ``run`` arguments (``local_experts.py:761-790``): all experts are processed in one batch and stored once, so
        ``store_every`` has no effect, and resuming into an existing store is GPSat's job (accepted, not acted on).

        ``world``: under ``torchrun`` (``WORLD_SIZE`` > 1; ``None`` = detect) the experts are partitioned over the
        ranks by ``parallel.lpt_partition`` on the selection counts, every rank runs its shard on its own GPU with
        no communication, and one ``gather_object`` brings the tables to rank 0, which stores them and returns them
        in the single-GPU order (other ranks return ``{}``).  Experts never interact, so the tables are bit-identical
        to a single-GPU run."""
        from . import parallel
        rank, nranks, _ = parallel.dist_env()
        if world is None:
            world = nranks > 1
        if device_select is None:
            device_select = self.vectorisable()
        if world and nranks > 1:
            if not self.vectorisable():
                raise NotImplementedError("multi-GPU runs use the array-built path: exact model (B200GPRModel) without a replacement model")
            parallel.init_process_group()
            tables = self._run_world(optimise, predict, min_obs, table_suffix, config_id, rank, nranks)
            if rank == 0:
                self._store(tables, store_path)
            return tables
        if device_select:
            tables = self._run_vectorised(optimise, predict, min_obs, table_suffix, config_id)
            self._store(tables, store_path)
            return tables
        t0 = time.time()
        coords_col = self.coords_col
        # phase A: selection (KD-trees are built once, as the reference caches them per global select)
        kdtrees = [cKDTree(self.data.loc[:, ls["col"]].values) if not isinstance(ls["col"], str) else None
                   for ls in self.local_select]
        pred_tree, pred_cols = None, None
        if self.pred_df is not None and self.max_dist is not None:
            pred_cols = [c for c in coords_col if c in self.pred_df]
            pred_tree = cKDTree(self.pred_df[pred_cols].values)
        cons = self._constraints_for_model()
        # group 0: the configured model; group 1: the replacement model for small experts
        groups = [dict(cls=self.model_cls, init=self.init_params, cons=cons, optim=self.optim_kwargs,
                       pred=self.pred_kwargs)]
        if self.replacement_threshold is not None:
            groups.append(dict(cls=self.replacement_cls, init=self.replacement_init_params,
                               cons=self._scaled_constraints(self.replacement_constraints, self.replacement_init_params),
                               optim=self.replacement_optim_kwargs, pred=self.replacement_pred_kwargs))
        lookup = None
        models, pcs, refs, skipped, group_of = [], [], [], [], []
        for _, rl in self.expert_locs.iterrows():
            ref = {c: rl[c] for c in coords_col}
            pc = prediction_coords_for(self.pred_df, ref, coords_col, self.max_dist, pred_tree, pred_cols)
            if len(pc) == 0:
                continue                                   # local_experts.py:962-965
            df_local = local_data_select(self.data, ref, self.local_select, kdtrees)
            if len(df_local) < min_obs:
                skipped.append((ref, len(df_local)))       # local_experts.py:988-1012
                continue
            gi = 1 if (self.replacement_threshold is not None and len(df_local) < self.replacement_threshold) else 0
            grp = groups[gi]
            m = grp["cls"](data=df_local, obs_col=self.obs_col, coords_col=coords_col,
                           expert_loc=np.array([ref[c] for c in coords_col]), verbose=False, **grp["init"])
            if self.load_params:                            # before the constraints, local_experts.py:1065-1115
                if lookup is None:
                    lookup = self._param_lookup(m.param_names)
                loaded = self._params_for(lookup, ref)
                if loaded:
                    m.set_parameters(**loaded)
            if grp["cons"] is not None:
                m.set_parameter_constraints(grp["cons"], move_within_tol=True, tol=1e-2)
            models.append(m); pcs.append(pc); refs.append(ref); group_of.append(gi)
        # phase B: per model family one fit call + one predict call for all its experts
        t_sel = time.time()
        success = np.zeros(len(models), bool)
        objective = np.full(len(models), np.nan)
        preds = [None] * len(models)
        for gi, grp in enumerate(groups):
            idx = [k for k, g in enumerate(group_of) if g == gi]
            if not idx:
                continue
            opt_kwargs = dict(grp["optim"])
            ok_g, obj_g, pred_g = fit_predict_models(
                [models[k] for k in idx], [pcs[k] for k in idx] if predict else None, optimise=optimise,
                max_iter=opt_kwargs.pop("max_iter", 10_000), fixed_params=opt_kwargs.pop("fixed_params", None),
                apply_scale=grp["pred"].get("apply_scale", True), **opt_kwargs)
            for j, k in enumerate(idx):
                success[k], objective[k] = ok_g[j], obj_g[j]
                if predict:
                    preds[k] = pred_g[j]
        t_fit = time.time()
        # phase C: tables
        frames: Dict[str, List[pd.DataFrame]] = {}

        def add(tabs):
            for k, v in tabs.items():
                frames.setdefault(k + table_suffix, []).append(v)

        model_name = f"{self.model_cls.__module__}.{self.model_cls.__name__}"[:64]
        for ref, nobs in skipped:
            add(dict_to_tables({"num_obs": nobs, "run_time": np.nan, "objective_value": np.nan,
                                "parameters_optimised": optimise, "optimise_success": False, "model": model_name,
                                "device": "", "config_id": config_id}, ref, "run_details"))
        per_expert = (t_fit - t_sel) / max(len(models), 1)
        for e, m in enumerate(models):
            device = m.cpu_name if m.gpu_name is None else m.gpu_name
            add(dict_to_tables({"num_obs": len(m.obs), "run_time": per_expert, "objective_value": float(objective[e]),
                                "parameters_optimised": optimise, "optimise_success": bool(success[e]) if optimise else False,
                                "model": f"{type(m).__module__}.{type(m).__name__}"[:64], "device": str(device)[:64],
                                "config_id": config_id},
                               refs[e], "run_details"))
            if predict:
                pred = dict(preds[e])
                for ci, c in enumerate(coords_col):
                    pred[f"pred_loc_{c}"] = pcs[e][:, ci]
                add(dict_to_tables(pred, refs[e], "preds"))
            hypes = m.get_parameters(*(self.params_to_store or m.param_names))
            add(dict_to_tables(hypes, refs[e], None))
        tables = {k: pd.concat(v) for k, v in frames.items()}
        self.timings = {"select_s": t_sel - t0, "engine_s": t_fit - t_sel, "tables_s": time.time() - t_fit}
        self._store(tables, store_path)
        return tables

    def _store(self, tables, store_path):
        """``.h5`` paths: appended to an HDF5 store like the reference (``local_experts.py:499-550``; needs
        PyTables) together with the ``oi_config`` table (``:873-884``).  Any other path is a directory that gets one
        parquet file per table and ``oi_config.json``."""
        if store_path is None:
            return
        import json
        cfg = json.dumps(getattr(self, "config", {}) or {}, default=str)
        if str(store_path).endswith((".h5", ".hdf5")):
            try:
                with pd.HDFStore(store_path, mode="a") as store:     # needs PyTables, like the reference
                    for k, v in tables.items():
                        store.append(k, v, data_columns=True)
                    store.append("oi_config", pd.DataFrame({"config": [cfg]}), min_itemsize={"config": max(len(cfg), 1)})
            except ImportError as exc:
                raise ImportError("writing result tables to HDF5 needs PyTables; give a directory as store_path "
                                  "(parquet files) or call run(store_path=None) to get the tables as DataFrames") from exc
            return
        import os
        os.makedirs(store_path, exist_ok=True)
        for k, v in tables.items():
            v.to_parquet(os.path.join(store_path, f"{k}.parquet"))
        with open(os.path.join(store_path, "oi_config.json"), "w") as f:
            f.write(cfg)

    @staticmethod
    def read_store(store_path) -> Dict[str, pd.DataFrame]:
        """Tables of a directory store written by :meth:`run`."""
        import os
        return {f[:-8]: pd.read_parquet(os.path.join(store_path, f)) for f in sorted(os.listdir(store_path))
                if f.endswith(".parquet")}

    # ---- multi-GPU: partition, run the shard, one gather ---------------------------------------------------------
    def _run_world(self, optimise, predict, min_obs, table_suffix, config_id, rank, nranks):
        from . import parallel
        import torch.distributed as dist
        parts = self._partition_experts(nranks)
        shard = LocalExpertOI.__new__(LocalExpertOI)
        shard.__dict__.update(self.__dict__)
        shard.expert_locs = self.expert_locs.iloc[parts[rank]].reset_index(drop=True)
        tables = shard._run_vectorised(optimise, predict, min_obs, table_suffix, config_id)
        payload = {"tables": tables, "row_expert": {k: parts[rank][v] for k, v in shard._row_expert.items()},
                   "timings": shard.timings}
        gathered = [None] * nranks if rank == 0 else None
        dist.gather_object(payload, gathered, dst=0)
        if rank != 0:
            return {}
        self.timings = {f"rank{r}": g["timings"] for r, g in enumerate(gathered)}
        return merge_rank_tables(gathered, table_suffix)

    def _partition_experts(self, nranks):
        """LPT over the experts on n^3 evals + n^2 m with n, m from the counting pass of the selection."""
        from . import parallel
        eng, table, locs, conds, ptab, pconds = self._selection_inputs()
        n = select_counts(eng, table, locs, conds)
        m = np.ones(len(locs), np.int64) if ptab is None else select_counts(eng, ptab, locs, pconds)
        return parallel.lpt_partition(parallel.expert_cost(n, m), nranks)


def merge_rank_tables(gathered: list, table_suffix: str = "") -> Dict[str, pd.DataFrame]:
    """Tables of all ranks in the row order of a single-GPU run: every table ordered by expert index, except that
    ``run_details`` lists the experts with too few observations first (the order ``_run_vectorised`` writes).
    ``gathered[r]`` = ``{"tables": {...}, "row_expert": {table: global expert index per row}}``."""
    out = {}
    names = []
    for g in gathered:
        for k in g["tables"]:
            if k not in names:
                names.append(k)
    for k in names:
        frames = [g["tables"][k] for g in gathered if k in g["tables"]]
        ids = np.concatenate([np.asarray(g["row_expert"][k]) for g in gathered if k in g["tables"]])
        df = pd.concat(frames)
        if k == "run_details" + table_suffix:
            small = np.isnan(df["run_time"].values)          # the NaN rows of experts below min_obs
            order = np.lexsort((ids, ~small))
        else:
            order = np.argsort(ids, kind="stable")
        out[k] = df.iloc[order]
    return out


# ---- vectorised driver path: batched selection on the GPU, array-built batch and tables ----------
_OPS = {">=": 0, ">": 1, "==": 2, "<": 3, "<=": 4}


def select_all(engine, table: np.ndarray, locs: np.ndarray, conds: list):
    """``gpsat_local_select`` for every row of ``locs``: (counts[E], off[E+1], idx[off[E]]).
    ``conds``: tuples ``(kind, op, cols, loc_cols, val)`` (see ``include/gpsat_b200.h``)."""
    import ctypes as C
    from . import _lib
    table = np.ascontiguousarray(table, dtype=np.float64)
    locs = np.ascontiguousarray(locs, dtype=np.float64)
    E = len(locs)
    arr = (_lib.SelectCond * max(len(conds), 1))()
    for q, (kind, op, cols, lcols, val) in enumerate(conds):
        arr[q].kind, arr[q].op, arr[q].ncol, arr[q].val = kind, _OPS[op], len(cols), float(val)
        for k in range(len(cols)):
            arr[q].col[k], arr[q].loc_col[k] = int(cols[k]), int(lcols[k])
    counts = np.zeros(E, dtype=np.int64)
    args = (engine._h, len(table), table.shape[1], C.c_void_p(table.ctypes.data), E, locs.shape[1],
            C.c_void_p(locs.ctypes.data), len(conds), arr)
    _lib.check(engine._lib.gpsat_local_select(*args, C.c_void_p(counts.ctypes.data), None, None))
    off = np.zeros(E + 1, dtype=np.int64)
    off[1:] = np.cumsum(counts)
    idx = np.zeros(max(int(off[-1]), 1), dtype=np.int64)
    if off[-1] > 0:
        _lib.check(engine._lib.gpsat_local_select(*args, C.c_void_p(counts.ctypes.data), C.c_void_p(off.ctypes.data),
                                                  C.c_void_p(idx.ctypes.data)))
    return counts, off, idx[:int(off[-1])]


def select_counts(engine, table: np.ndarray, locs: np.ndarray, conds: list) -> np.ndarray:
    """Counting pass of :func:`select_all` alone (what the multi-GPU partitioner needs)."""
    import ctypes as C
    from . import _lib
    table = np.ascontiguousarray(table, dtype=np.float64)
    locs = np.ascontiguousarray(locs, dtype=np.float64)
    arr = (_lib.SelectCond * max(len(conds), 1))()
    for q, (kind, op, cols, lcols, val) in enumerate(conds):
        arr[q].kind, arr[q].op, arr[q].ncol, arr[q].val = kind, _OPS[op], len(cols), float(val)
        for k in range(len(cols)):
            arr[q].col[k], arr[q].loc_col[k] = int(cols[k]), int(lcols[k])
    counts = np.zeros(len(locs), dtype=np.int64)
    _lib.check(engine._lib.gpsat_local_select(engine._h, len(table), table.shape[1], C.c_void_p(table.ctypes.data),
                                              len(locs), locs.shape[1], C.c_void_p(locs.ctypes.data), len(conds), arr,
                                              C.c_void_p(counts.ctypes.data), None, None))
    return counts


def select_count_noalloc(engine, table, locs, conds: list, counts):
    """``gpsat_local_select_count_noalloc``: the counting pass of :func:`select_all` without any allocation inside
    the library (no counterpart in the reference; see ``include/gpsat_b200.h``).  ``table`` [N, C], ``locs`` [E, CL]
    (float64) and ``counts`` [E] (int64) are contiguous torch CUDA tensors on the engine's device, owned by the
    caller; ``E * ceil(N / 1024)`` must not exceed ``_lib.NOALLOC_MAX_CHUNK_COUNTS``.  Calls that overlap on one
    GPU race on the library's fixed scratch: one call at a time per device."""
    import ctypes as C
    import torch
    from . import _lib
    for t, dt in ((table, torch.float64), (locs, torch.float64), (counts, torch.int64)):
        if not (t.is_cuda and t.is_contiguous() and t.dtype == dt):
            raise ValueError("select_count_noalloc needs contiguous CUDA tensors (float64 table / locs, int64 counts)")
    if counts.numel() != locs.shape[0]:
        raise ValueError("counts needs one entry per expert location")
    arr = (_lib.SelectCond * max(len(conds), 1))()
    for q, (kind, op, cols, lcols, val) in enumerate(conds):
        arr[q].kind, arr[q].op, arr[q].ncol, arr[q].val = kind, _OPS[op], len(cols), float(val)
        for k in range(len(cols)):
            arr[q].col[k], arr[q].loc_col[k] = int(cols[k]), int(lcols[k])
    torch.cuda.current_stream(table.device).synchronize()     # inputs were produced on torch's stream
    _lib.check(engine._lib.gpsat_local_select_count_noalloc(
        engine._h, table.shape[0], table.shape[1], C.c_void_p(table.data_ptr()), locs.shape[0], locs.shape[1],
        C.c_void_p(locs.data_ptr()), len(conds), arr, C.c_void_p(counts.data_ptr())))
    return counts


def _template_model(self):
    """One model instance on dummy data: parameter values, transforms and trainable flags do not depend on the data."""
    D = len(self.coords_col)
    obs_col = self.obs_col if isinstance(self.obs_col, str) else self.obs_col[0]
    return B200GPRModel(coords=np.arange(2.0 * D).reshape(2, D), obs=np.array([[0.0], [1.0]]),
                        coords_col=self.coords_col, obs_col=[obs_col], verbose=False, **self.init_params)


def _selection_inputs(self):
    """(engine, point table [N, D + 1], expert locations [E, D], observation conditions, prediction table or None,
    prediction conditions) in the form ``gpsat_local_select`` takes them."""
    coords_col = self.coords_col
    obs_col = self.obs_col if isinstance(self.obs_col, str) else self.obs_col[0]
    eng = _template_model(self).engine
    locs = self.expert_locs[coords_col].values.astype(np.float64)
    table = np.ascontiguousarray(self.data[coords_col + [obs_col]].values, dtype=np.float64)
    conds = []
    for ls in self.local_select:
        col, comp = ls["col"], ls["comp"]
        if isinstance(col, str):
            assert col in coords_col and comp in _OPS, f"local_select entry not understood: {ls}"
            conds.append((0, comp, [coords_col.index(col)], [coords_col.index(col)], ls["val"]))
        else:
            assert comp in ["<", "<="], "for multi dimensional values only less than comparison handled"
            ids = [coords_col.index(c) for c in col]
            conds.append((1, "<=", ids, ids, ls["val"]))          # KD-tree ball query: d <= r
    ptab, pconds = None, []
    if self.pred_df is not None:
        shared = [c for c in coords_col if c in self.pred_df]
        ptab = np.ascontiguousarray(self.pred_df[shared].values, dtype=np.float64)
        sid = [coords_col.index(c) for c in shared]
        pconds = [] if self.max_dist is None else [(1, "<", list(range(len(shared))), sid, self.max_dist)]
    return eng, table, locs, conds, ptab, pconds


def _run_vectorised(self, optimise, predict, min_obs, table_suffix, config_id):
    from .engine import RaggedBatch
    from . import _lib
    t0 = time.time()
    if not self.vectorisable():
        raise NotImplementedError("device_select=True supports the exact model (B200GPRModel) without a replacement model")
    coords_col, D = self.coords_col, len(self.coords_col)
    template = lambda: _template_model(self)
    tmpl = template()
    cs = np.broadcast_to(np.atleast_2d(tmpl.coords_scale)[0, :], (D,)).astype(np.float64)
    obs_scale = float(np.atleast_2d(tmpl.obs_scale).reshape(-1)[0])
    local_mean = isinstance(self.init_params.get("obs_mean"), str) and self.init_params.get("obs_mean") == "local"
    # ---- phase A: selection of observations and prediction locations for all experts ----
    eng, table, locs, conds, ptab, pconds = _selection_inputs(self)
    eng = tmpl.engine
    E = len(locs)
    n_all, off_all, idx_all = select_all(eng, table, locs, conds)
    if ptab is None:
        m_all = np.ones(E, dtype=np.int64)
        poff_all = np.arange(E + 1, dtype=np.int64)
        pidx_all, shared = None, []
    else:
        shared = [c for c in coords_col if c in self.pred_df]
        m_all, poff_all, pidx_all = select_all(eng, ptab, locs, pconds)
    has_pred = m_all > 0                                       # local_experts.py:962-965
    small = has_pred & (n_all < min_obs)                       # local_experts.py:988-1012
    keep = np.flatnonzero(has_pred & ~small)
    K = len(keep)
    # packed observations of the kept experts
    seg = [idx_all[off_all[e]:off_all[e + 1]] for e in keep]
    rows = np.concatenate(seg) if K else np.zeros(0, np.int64)
    off_n = np.zeros(K + 1, dtype=np.int64)
    off_n[1:] = np.cumsum(n_all[keep])
    local = table[rows]
    assert not np.isnan(local[:, :D]).any(), "nans found in coords"
    assert not np.isnan(local[:, D]).any(), "nans found in obs"
    X = local[:, :D] / cs
    y = local[:, D].copy()
    means = np.zeros(K)
    if local_mean:
        for k in range(K):
            means[k] = np.mean(y[off_n[k]:off_n[k + 1]].reshape(-1, 1), axis=0)[0]    # base_model.py:195-198
    y -= np.repeat(means, n_all[keep])
    y /= obs_scale
    # packed prediction coordinates (unscaled for the tables, scaled for the engine)
    off_m = np.zeros(K + 1, dtype=np.int64)
    off_m[1:] = np.cumsum(m_all[keep])
    pc = np.repeat(locs[keep], m_all[keep], axis=0)
    if pidx_all is not None and K:
        prow = np.concatenate([pidx_all[poff_all[e]:poff_all[e + 1]] for e in keep])
        for j, c in enumerate(shared):
            pc[:, coords_col.index(c)] = ptab[prow, j]
    apply_scale = self.pred_kwargs.get("apply_scale", True)
    Xs = pc / cs if apply_scale else pc.copy()
    batch = RaggedBatch(tmpl.kernel, D, off_n, np.ascontiguousarray(X), np.ascontiguousarray(y), off_m,
                        np.ascontiguousarray(Xs))
    # hyper-parameters, transforms, trainable flags per expert
    cons = self._constraints_for_model()
    P = D + 2
    theta = np.empty((K, P)); tkind = np.empty((K, P), np.int32); ta = np.empty((K, P)); tb = np.empty((K, P))
    lookup = self._param_lookup(tmpl.param_names) if self.load_params else None
    base = None
    for k, e in enumerate(keep):
        m = None
        if lookup is not None:
            loaded = self._params_for(lookup, dict(zip(coords_col, locs[e])))
            if loaded:                                      # before the constraints, local_experts.py:1065-1115
                m = template()
                m.set_parameters(**loaded)
                if cons is not None:
                    m.set_parameter_constraints(cons, move_within_tol=True, tol=1e-2)
        if m is None:
            if base is None:
                base = template()
                if cons is not None:
                    base.set_parameter_constraints(cons, move_within_tol=True, tol=1e-2)
            m = base
        theta[k], tkind[k], ta[k], tb[k] = m._theta, m._tkind, m._ta, m._tb
    opt_kwargs = dict(self.optim_kwargs)
    fixed = opt_kwargs.pop("fixed_params", None) or []
    tmpl._fix_hyperparameters(fixed)
    trainable = np.tile(tmpl._trainable.astype(np.uint8), (K, 1))
    t_sel = time.time()
    # ---- phase B ----
    success = np.zeros(K, bool)
    mean_evals = 0.0
    if K and optimise:
        tol = opt_kwargs.pop("tol", None)
        kw = {} if tol is None else {"ftol": float(tol), "pgtol": float(tol)}
        pri = None
        if np.any(tmpl._pkind):
            pri = (np.tile(tmpl._pkind, (K, 1)), np.tile(tmpl._pa, (K, 1)), np.tile(tmpl._pb, (K, 1)))
        res = eng.fit(batch, theta, tkind, ta, tb, trainable, max_iter=int(opt_kwargs.pop("max_iter", 10_000)),
                      priors=pri, **kw)
        ok_theta = res.status != _lib.STATUS_NOT_PD
        theta = np.where(ok_theta[:, None], res.theta, theta)
        success = res.status == _lib.STATUS_OK
        mean_evals = float(np.mean(res.n_evals))
    if K and predict:
        out = eng.predict(batch, theta)
        objective = -np.asarray(out["lml"])
        # an expert whose K + s2 I is not positive definite at its final hyper-parameters has no prediction: the
        # per-model path raises FloatingPointError there (models.py), a batch reports it per expert instead - NaN
        # predictions and objective, optimise_success False
        bad = np.asarray(out["status"]) != _lib.STATUS_OK
        if bad.any():
            warnings.warn(f"{int(bad.sum())} expert(s): covariance not positive definite at the final hyper-parameters; "
                          "their predictions are NaN")
            badrows = np.repeat(bad, m_all[keep])
            for key in ("f*", "f*_var", "y_var"):
                out[key] = np.where(badrows, np.nan, np.asarray(out[key]))
            objective = np.where(bad, np.nan, objective)
            success = success & ~bad
    elif K:
        objective = -np.asarray(eng.lml_grad(batch, theta, with_grad=False)[0])
    else:
        objective = np.zeros(0)
    t_fit = time.time()
    # ---- phase C: tables, in the row order of the per-expert path ----
    def midx(rows_locs):
        if D == 1:
            return pd.Index(rows_locs[:, 0], name=coords_col[0])
        return pd.MultiIndex.from_arrays([rows_locs[:, j] for j in range(D)], names=coords_col)

    model_name = f"{B200GPRModel.__module__}.{B200GPRModel.__name__}"[:64]
    device = str(tmpl.cpu_name if tmpl.gpu_name is None else tmpl.gpu_name)[:64]
    sk = np.flatnonzero(small)
    per_expert = (t_fit - t_sel) / max(K, 1)
    rd = pd.DataFrame({
        "_dim_0": np.zeros(len(sk) + K, dtype=np.int64),
        "num_obs": np.concatenate([n_all[sk], n_all[keep]]),
        "run_time": np.concatenate([np.full(len(sk), np.nan), np.full(K, per_expert)]),
        "objective_value": np.concatenate([np.full(len(sk), np.nan), objective]),
        "parameters_optimised": optimise,
        "optimise_success": np.concatenate([np.zeros(len(sk), bool), success if optimise else np.zeros(K, bool)]),
        "model": model_name,
        "device": np.concatenate([np.full(len(sk), "", dtype=object), np.full(K, device, dtype=object)]),
        "config_id": config_id})
    rd.index = midx(np.concatenate([locs[sk], locs[keep]]))
    tables = {"run_details" + table_suffix: rd} if len(rd) else {}
    row_expert = {"run_details" + table_suffix: np.concatenate([sk, keep])} if len(rd) else {}
    if K and predict:
        # without obs_mean='local' the reference stores np.array([0]) - an integer - as the mean (base_model.py:199-200)
        f_bar = np.repeat(means, m_all[keep]) if local_mean else np.zeros(len(pc), dtype=np.int64)
        pos = np.arange(len(pc), dtype=np.int64) - np.repeat(off_m[:-1], m_all[keep])      # row number within the expert
        pr = pd.DataFrame({"_dim_0": pos, "f*": out["f*"], "f*_var": out["f*_var"], "y_var": out["y_var"], "f_bar": f_bar})
        for j, c in enumerate(coords_col):
            pr[f"pred_loc_{c}"] = pc[:, j]
        pr.index = midx(np.repeat(locs[keep], m_all[keep], axis=0))
        tables["preds" + table_suffix] = pr
        row_expert["preds" + table_suffix] = np.repeat(keep, m_all[keep])
    store = self.params_to_store or tmpl.param_names
    if K:
        if "lengthscales" in store:
            ls = pd.DataFrame({"_dim_0": np.tile(np.arange(D), K), "lengthscales": theta[:, :D].reshape(-1)})
            ls.index = midx(np.repeat(locs[keep], D, axis=0))
            tables["lengthscales" + table_suffix] = ls
            row_expert["lengthscales" + table_suffix] = np.repeat(keep, D)
        for name, col in (("kernel_variance", D), ("likelihood_variance", D + 1)):
            if name in store:
                df = pd.DataFrame({"_dim_0": np.zeros(K, dtype=np.int64), name: theta[:, col].copy()})
                df.index = midx(locs[keep])
                tables[name + table_suffix] = df
                row_expert[name + table_suffix] = keep.copy()
    self.timings = {"select_s": t_sel - t0, "engine_s": t_fit - t_sel, "tables_s": time.time() - t_fit,
                    "mean_evals_per_expert": mean_evals}
    self._row_expert = row_expert          # expert index (row of self.expert_locs) of every table row
    return tables


LocalExpertOI._run_vectorised = _run_vectorised
LocalExpertOI._selection_inputs = _selection_inputs
