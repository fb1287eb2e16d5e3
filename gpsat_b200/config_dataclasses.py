"""Configuration dataclasses of the local-expert driver, with the reference's field names and JSON keys.

Re-declaration of ``GPSat/config_dataclasses.py`` with the standard library only (the reference builds on
``dataclasses_json``, absent here): ``DataConfig`` (``:9-216``), ``ModelConfig`` (``:219-328``),
``ExpertLocsConfig`` (``:331-445``), ``PredictionLocsConfig`` (``:448-509``), ``RunConfig`` (``:512-547``) and
``ExperimentConfig`` (``:550-610``; JSON keys ``data`` / ``model`` / ``locations`` / ``pred_loc`` / ``run_kwargs``,
``:591-595``).  Every class offers the ``dataclasses_json`` calls the reference code uses - ``to_dict``,
``from_dict``, ``to_json``, ``from_json`` - and ``to_dict_with_dataframe``; DataFrames handed to a config are
stored as dicts so that the config stays JSON-serialisable (``__post_init__``, ``:154-160, 425-433, 493-499``).

``load_table`` is the small part of ``DataLoader.load`` (``GPSat/dataloader.py:1522-1678``) the configs of this path
need for tabular sources: DataFrame / dict / csv / tsv / parquet (HDF5 when PyTables is importable), then
``add_data_to_col``, ``col_funcs``, ``row_select`` / ``where``, ``col_select``, ``reset_index``
(``_modify_df``, ``:1680-1800``).  netCDF / zarr sources and where-pushdown into HDF5 stay in GPSat (out of scope,
SURVEY.md section 2 row 14).
"""
from __future__ import annotations

import dataclasses
import json
import os
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Union

import numpy as np
import pandas as pd

MODELS = ("GPflowGPRModel", "GPflowSGPRModel", "GPflowSVGPModel", "sklearnGPRModel", "GPflowVFFModel", "GPflowASVGPModel",
          "B200GPRModel", "B200SGPRModel", "B200VFFModel")
METHODS = ("expert_loc", "from_dataframe", "from_source")

_COMPS = {"==": np.equal, "!=": np.not_equal, ">=": np.greater_equal, ">": np.greater, "<=": np.less_equal, "<": np.less}


# ---- dataclasses_json-like surface -------------------------------------------------------------------------
def _plain(v):
    if isinstance(v, pd.DataFrame):
        return v.to_dict()
    if dataclasses.is_dataclass(v):
        return v.to_dict()
    if isinstance(v, np.ndarray):
        return v.tolist()
    if isinstance(v, (np.integer,)):
        return int(v)
    if isinstance(v, (np.floating,)):
        return float(v)
    if isinstance(v, dict):
        return {k: _plain(x) for k, x in v.items()}
    if isinstance(v, (list, tuple)):
        return [_plain(x) for x in v]
    return v


class _JsonMixin:
    """``to_dict`` / ``from_dict`` / ``to_json`` / ``from_json`` as ``dataclasses_json`` provides them.  A field may
    carry ``metadata={"field_name": key}`` to appear under another key (``config(field_name=...)`` in the reference)."""

    @classmethod
    def _key(cls, f):
        return f.metadata.get("field_name", f.name)

    def to_dict(self) -> dict:
        return {self._key(f): _plain(getattr(self, f.name)) for f in dataclasses.fields(self)}

    @classmethod
    def from_dict(cls, d: dict):
        known = {cls._key(f): f for f in dataclasses.fields(cls)}
        unknown = [k for k in d if k not in known]
        if unknown:
            raise KeyError(f"{cls.__name__}: unknown keys {unknown}; valid keys: {sorted(known)}")
        kw = {}
        for k, v in d.items():
            f = known[k]
            sub = _NESTED.get((cls.__name__, f.name))
            kw[f.name] = sub.from_dict(v) if (sub is not None and isinstance(v, dict)) else v
        return cls(**kw)

    def to_json(self, **kwargs) -> str:
        return json.dumps(self.to_dict(), **kwargs)

    @classmethod
    def from_json(cls, s: str):
        return cls.from_dict(json.loads(s))


def _frame(v):
    return pd.DataFrame.from_dict(v) if isinstance(v, dict) else v


# ---- table loading ------------------------------------------------------------------------------------------
def _eval_func(func):
    """``GPSat.utils.config_func`` (``utils.py:311-520``) evaluates strings from the config with ``eval``; so does
    this (numpy and pandas in scope).  Operator strings ("<=", "+", ...) become two-argument functions."""
    if callable(func):
        return func
    assert isinstance(func, str), f"func must be a string or a callable, got {type(func)}"
    if func in _COMPS:
        return _COMPS[func]
    if func.strip() in ("+", "-", "*", "/", "%", "|", "&"):
        op = func.strip()
        return eval(f"lambda arg1, arg2: arg1 {op} arg2")
    return eval(func, {"np": np, "pd": pd, "numpy": np, "pandas": pd})


def _col_values(df, col):
    """``get_col_values(..., return_numpy=True)``: a NumPy array (strings as an object array, as pandas < 3 gave)."""
    v = df[col]
    if isinstance(v, pd.DataFrame):
        return v.to_numpy()
    if pd.api.types.is_string_dtype(v.dtype) and not pd.api.types.is_object_dtype(v.dtype):
        return v.to_numpy(dtype=object)
    return v.to_numpy()


def _apply_func(df, spec):
    fn = _eval_func(spec["func"])
    col_args = spec.get("col_args") or []
    col_args = col_args if isinstance(col_args, list) else [col_args]
    args = spec.get("args") or []
    args = args if isinstance(args, list) else [args]
    kwargs = dict(spec.get("kwargs") or {})
    for k, c in (spec.get("col_kwargs") or {}).items():
        kwargs[k] = _col_values(df, c)
    return fn(*[_col_values(df, c) for c in col_args], *args, **kwargs)


def row_select_bool(df: pd.DataFrame, conds, combine: str = "AND") -> np.ndarray:
    """Boolean mask of the rows that satisfy a list of ``{"col", "comp", "val"}`` (or ``{"func", "col_args"}``)
    conditions (``DataLoader._bool_xarray_from_where`` / ``row_select_bool``, ``dataloader.py:2449-2620``)."""
    if conds is None:
        return np.ones(len(df), bool)
    conds = conds if isinstance(conds, list) else [conds]
    out = np.ones(len(df), bool) if combine == "AND" else np.zeros(len(df), bool)
    for c in conds:
        if "func" in c:
            m = np.asarray(_apply_func(df, c), dtype=bool)
        else:
            col, comp, val = c["col"], c["comp"], c["val"]
            assert comp in _COMPS, f"comp: {comp} is not valid, must be one of {list(_COMPS)}"
            v = _col_values(df, col)
            if np.issubdtype(v.dtype, np.datetime64):
                val = np.datetime64(val)
            m = _COMPS[comp](v, val)
        out = (out & m) if combine == "AND" else (out | m)
    return out


def _read_source(source, table=None, engine=None, read_kwargs=None):
    kw = dict(read_kwargs or {})
    if isinstance(source, pd.DataFrame):
        return source.copy()
    if isinstance(source, dict):
        return pd.DataFrame.from_dict(source)
    assert isinstance(source, str), f"data source of type {type(source)} is not handled"
    suffix = source.rsplit(".", 1)[-1].lower()
    engine = engine or {"csv": "read_csv", "tsv": "read_csv", "h5": "HDFStore", "parquet": "read_parquet"}.get(suffix)
    if engine == "read_csv":
        if suffix == "tsv":
            kw.setdefault("sep", "\t")
        return pd.read_csv(source, **kw)
    if engine == "read_parquet":
        return pd.read_parquet(source, **kw)
    if engine == "HDFStore":
        try:
            import tables  # noqa: F401
        except ImportError as exc:
            raise ImportError(f"reading {source} needs PyTables, as in the reference; csv / parquet sources do not") from exc
        with pd.HDFStore(source, mode="r") as store:
            return store.get(table)
    raise NotImplementedError(f"source '{source}' (engine {engine}): only tabular files (csv, tsv, parquet, h5) are read here; "
                              "load netCDF / zarr data with GPSat's DataLoader and pass the DataFrame")


def load_table(source, where=None, table=None, col_funcs=None, row_select=None, col_select=None, engine=None,
               source_kwargs=None, add_data_to_col=None, reset_index=False, verbose=False, **_ignored) -> pd.DataFrame:
    """``DataLoader.load`` for tabular sources: read, ``where``, ``add_data_to_col``, ``col_funcs``, ``row_select``,
    ``col_select`` - in the reference's order (``dataloader.py:1645-1678``, ``_modify_df``)."""
    df = _read_source(source, table=table, engine=engine, read_kwargs=source_kwargs)
    if where:
        df = df.loc[row_select_bool(df, where)]
    if reset_index:
        df = df.reset_index(drop=True)
    if add_data_to_col:                      # rows are replicated for every value of a list (dataloader.py:251-330)
        for col, vals in add_data_to_col.items():
            vals = vals if isinstance(vals, (list, tuple, np.ndarray)) else [vals]
            parts = []
            for v in vals:
                p = df.copy()
                p[col] = v
                parts.append(p)
            df = pd.concat(parts, ignore_index=True) if parts else df
    if col_funcs:
        for new_col, spec in col_funcs.items():
            df[new_col] = _apply_func(df, spec)
    if row_select:
        df = df.loc[row_select_bool(df, row_select)]
    if col_select:
        missing = [c for c in col_select if c not in df]
        assert not missing, f"col_select has columns not in the data: {missing}"
        df = df.loc[:, col_select]
    return df


# ---- the configs ----------------------------------------------------------------------------------------------
@dataclass
class DataConfig(_JsonMixin):
    """``GPSat/config_dataclasses.py:9-216``."""
    data_source: Union[str, pd.DataFrame, dict, None] = None
    table: Optional[str] = None
    obs_col: Optional[str] = None
    coords_col: Optional[List[str]] = None
    local_select: Optional[List[dict]] = None
    global_select: Optional[List[dict]] = None
    row_select: Optional[List[dict]] = None
    col_select: Optional[List[str]] = None
    col_funcs: Union[List[str], dict, None] = None
    engine: Optional[str] = None
    read_kwargs: Optional[dict] = None

    def __post_init__(self):
        if isinstance(self.data_source, pd.DataFrame):
            self.data_source = self.data_source.to_dict()

    def to_dict_with_dataframe(self) -> dict:
        d = self.to_dict()
        d["data_source"] = _frame(d["data_source"])
        return d

    def load(self, where=None, verbose=False, **kwargs) -> pd.DataFrame:
        """``DataConfig.load`` (``:187-216``).  Static ``global_select`` entries (``col`` / ``comp`` / ``val``) are
        applied as ``where`` conditions; dynamic ones (``loc_col`` / ``src_col`` / ``func``) only bound what the
        reference reads from disk per expert and are implied by ``local_select`` for in-memory tables."""
        static = [g for g in (self.global_select or []) if "col" in g]
        if where is not None:
            static = static + (where if isinstance(where, list) else [where])
        return load_table(_frame(self.data_source), where=static or None, table=self.table, col_funcs=self.col_funcs,
                          row_select=self.row_select, col_select=self.col_select, engine=self.engine,
                          source_kwargs=self.read_kwargs, verbose=verbose, **kwargs)


@dataclass
class ModelConfig(_JsonMixin):
    """``GPSat/config_dataclasses.py:219-328``."""
    oi_model: Union[str, dict, None] = None
    init_params: Optional[dict] = None
    constraints: Optional[Dict[str, dict]] = None
    load_params: Optional[dict] = None
    optim_kwargs: Optional[dict] = None
    pred_kwargs: Optional[dict] = None
    params_to_store: Union[str, List[str]] = "all"
    replacement_threshold: Optional[int] = None
    replacement_model: Optional[str] = None
    replacement_init_params: Optional[dict] = None
    replacement_constraints: Optional[dict] = None
    replacement_optim_kwargs: Optional[dict] = None
    replacement_pred_kwargs: Optional[dict] = None


@dataclass
class ExpertLocsConfig(_JsonMixin):
    """``GPSat/config_dataclasses.py:331-445``."""
    source: Union[str, pd.DataFrame, dict, None] = None
    where: Union[dict, List[dict], None] = None
    add_data_to_col: Optional[dict] = None
    col_funcs: Optional[Dict[str, dict]] = None
    col_select: Optional[List[str]] = None
    row_select: Optional[List[dict]] = None
    reset_index: bool = False
    source_kwargs: Optional[dict] = None
    verbose: bool = False
    sort_by: Union[str, List[str], None] = None
    # legacy inputs (:420-423)
    df: Union[pd.DataFrame, dict, None] = None
    file: Optional[str] = None
    keep_cols: Optional[bool] = None

    def __post_init__(self):
        if isinstance(self.df, pd.DataFrame):
            self.df = self.df.to_dict()
        if isinstance(self.source, pd.DataFrame):
            self.source = self.source.to_dict()

    def to_dict_with_dataframe(self) -> dict:
        d = self.to_dict()
        d["df"], d["source"] = _frame(d["df"]), _frame(d["source"])
        return d


@dataclass
class PredictionLocsConfig(_JsonMixin):
    """``GPSat/config_dataclasses.py:448-509``."""
    method: str = "expert_loc"
    coords_col: Optional[List[str]] = None
    df: Union[pd.DataFrame, dict, None] = None
    df_file: Optional[str] = None
    max_dist: Union[int, float, None] = None
    load_kwargs: Optional[dict] = None
    X_out: Optional[str] = None

    def __post_init__(self):
        assert self.method in METHODS, f"method: {self.method} must be one of {METHODS}"
        if isinstance(self.df, pd.DataFrame):
            self.df = self.df.to_dict()

    def to_dict_with_dataframe(self) -> dict:
        d = self.to_dict()
        d["df"] = _frame(d["df"])
        return d


@dataclass
class RunConfig(_JsonMixin):
    """``GPSat/config_dataclasses.py:512-547``: the keyword arguments of ``LocalExpertOI.run``."""
    store_path: str = ""
    store_every: int = 10
    check_config_compatible: bool = True
    skip_valid_checks_on: Optional[List] = field(default_factory=list)
    optimise: bool = True
    predict: bool = True
    min_obs: int = 3
    table_suffix: str = ""


@dataclass
class ExperimentConfig(_JsonMixin):
    """``GPSat/config_dataclasses.py:550-610``: the whole experiment; JSON keys as in the reference's config files."""
    data_config: DataConfig = field(default_factory=DataConfig, metadata={"field_name": "data"})
    model_config: ModelConfig = field(default_factory=ModelConfig, metadata={"field_name": "model"})
    expert_locs_config: ExpertLocsConfig = field(default_factory=ExpertLocsConfig, metadata={"field_name": "locations"})
    prediction_locs_config: PredictionLocsConfig = field(default_factory=PredictionLocsConfig, metadata={"field_name": "pred_loc"})
    run_config: RunConfig = field(default_factory=RunConfig, metadata={"field_name": "run_kwargs"})
    comment: Union[str, List[str], None] = None

    def to_dict_with_dataframe(self) -> dict:
        return {"data_config": self.data_config.to_dict_with_dataframe(), "model_config": self.model_config.to_dict(),
                "expert_locs_config": self.expert_locs_config.to_dict_with_dataframe(),
                "prediction_locs_config": self.prediction_locs_config.to_dict_with_dataframe(),
                "run_config": self.run_config.to_dict(), "comment": self.comment}


_NESTED = {("ExperimentConfig", "data_config"): DataConfig, ("ExperimentConfig", "model_config"): ModelConfig,
           ("ExperimentConfig", "expert_locs_config"): ExpertLocsConfig,
           ("ExperimentConfig", "prediction_locs_config"): PredictionLocsConfig,
           ("ExperimentConfig", "run_config"): RunConfig}
