"""Post-processing steps of the local-expert workflow on the GPU.

Mirrors ``GPSat/postprocessing.py``:

* :func:`gaussian_2d_weight`          ``postprocessing.py:22-52``   (numba kernel there, CUDA kernel here)
* :class:`SmoothingConfig`            ``postprocessing.py:55-93``
* :func:`smooth_hyperparameters`      ``postprocessing.py:96-395``  on in-memory result tables
  (``LocalExpertOI.run`` output) or an HDF5 results file when PyTables is installed
* :func:`glue_local_predictions_1d` / :func:`glue_local_predictions_2d`   ``postprocessing.py:462-593``

The arithmetic runs through the C ABI (``gpsat_gaussian_smooth``, ``gpsat_glue_predictions``); there is
no CPU fallback.
"""
from __future__ import annotations

import ctypes as C
import re
from dataclasses import dataclass
from typing import Dict, List, Optional, Union

import numpy as np
import pandas as pd

from . import _lib
from .models import get_engine


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _p(a):
    return C.c_void_p(a.ctypes.data)


def gaussian_2d_weight(x0, y0, x, y, l_x, l_y, vals, device: int = 0) -> np.ndarray:
    """Normalised Gaussian-weighted average of ``vals`` (NaNs skipped) at every reference position
    ``(x0[i], y0[i])``: weights ``exp(-d^2 / 2)``, ``d^2 = ((x - x0) / l_x)^2 + ((y - y0) / l_y)^2``;
    NaN where all weights vanish (``postprocessing.py:22-52``).  ``vals`` may be [n] or [F, n]."""
    x0, y0, x, y = (_f64(np.atleast_1d(a)) for a in (x0, y0, x, y))
    v = _f64(vals)
    single = v.ndim == 1
    v = np.ascontiguousarray(v.reshape(-1, len(x)))
    assert len(x0) == len(y0) and len(x) == len(y) and v.shape[1] == len(x)
    l_x = float(np.asarray(l_x).reshape(-1)[0])
    l_y = float(np.asarray(l_y).reshape(-1)[0])
    out = np.empty((v.shape[0], len(x0)))
    eng = get_engine(device)
    _lib.check(eng._lib.gpsat_gaussian_smooth(eng._h, _lib.MEM_HOST, len(x0), _p(x0), _p(y0), len(x), v.shape[0],
                                              _p(x), _p(y), l_x, l_y, _p(v), _p(out)))
    return out[0] if single else out


@dataclass
class SmoothingConfig:
    """``postprocessing.py:55-93``: smoothing length scales and optional clamps of one parameter."""
    l_x: Union[int, float] = 1
    l_y: Union[int, float] = 1
    max: Union[int, float, list, None] = None
    min: Union[int, float, list, None] = None

    def __getitem__(self, item):
        if hasattr(self, item):
            return getattr(self, item)
        raise AttributeError(f"{item} is not an attribute of {type(self)} object")

    def get(self, key, default=None):
        return self[key] if hasattr(self, key) else default


def smooth_hyperparameters(tables: Dict[str, pd.DataFrame], params_to_smooth: List[str],
                           smooth_config_dict: Dict[str, Union[SmoothingConfig, dict]], coords_col: List[str],
                           all_params: Optional[List[str]] = None, xy_dims: List[str] = ("x", "y"),
                           reference_table_suffix: str = "", table_suffix: str = "_SMOOTHED",
                           device: int = 0) -> Dict[str, pd.DataFrame]:
    """Smooth hyper-parameter tables over the expert locations (``postprocessing.py:96-395``).

    ``tables`` maps table names to DataFrames with the coordinate columns, optional ``_dim_*``
    columns and one value column named like the parameter (what :meth:`LocalExpertOI.run` returns).
    Every parameter in ``params_to_smooth`` is clamped to ``[min, max]``, averaged with Gaussian
    weights over the ``xy_dims`` plane separately for every combination of the other coordinates
    and ``_dim_*`` components, clamped again and returned as ``<param><reference_table_suffix><table_suffix>``
    indexed by ``coords_col``; the other parameters of ``all_params`` are copied.  Rows whose
    smoothed value is NaN are dropped, as in the reference."""
    assert table_suffix != reference_table_suffix
    xy_dims = list(xy_dims)
    assert len(xy_dims) == 2, "dimensions to smooth over must have length 2"
    x_col, y_col = xy_dims
    all_params = list(all_params) if all_params is not None else list(params_to_smooth)
    assert all(p in all_params for p in params_to_smooth), \
        f"some params_to_smooth:\n{params_to_smooth} are not in model.param_names:\n{all_params}"
    cfg = {}
    for k, v in smooth_config_dict.items():
        cfg[f"{k}{reference_table_suffix}"] = SmoothingConfig(**v) if isinstance(v, dict) else v
    out: Dict[str, pd.DataFrame] = {}
    for param in params_to_smooth:
        hp = f"{param}{reference_table_suffix}"
        if hp not in cfg or hp not in tables:
            raise NotImplementedError(f"parameter: {hp} is not in dfs: {list(tables.keys())}\ncheck your inputs")
        df = tables[hp].reset_index(drop=False) if tables[hp].index.names != [None] else tables[hp].copy(True)
        df = df[[c for c in df.columns if c != "index"]]
        col_order = df.columns.values.tolist()
        sc = cfg[hp]
        other_dims = [c for c in coords_col if c not in xy_dims]
        dim_cols = [c for c in df.columns if re.search(r"^_dim_\d", c)]
        other_dims += dim_cols
        groups = [((), df)] if not other_dims else list(df.groupby(other_dims, sort=False, dropna=False))
        pieces = []
        for key, sub in groups:
            key = key if isinstance(key, tuple) else (key,)
            vals = sub[param].values.astype(np.float64).copy()
            _min, _max = sc.get("min", None), sc.get("max", None)
            comp = int(key[-1]) if dim_cols else 0
            if isinstance(_min, (list, np.ndarray)) and dim_cols and len(_min) > 0:
                _min = _min[min(comp, len(_min) - 1)]
            if isinstance(_max, (list, np.ndarray)) and dim_cols and len(_max) > 0:
                _max = _max[min(comp, len(_max) - 1)]
            if _max is not None:
                vals[vals > _max] = _max
            if _min is not None:
                vals[vals < _min] = _min
            x, y = sub[x_col].values, sub[y_col].values
            sm = gaussian_2d_weight(x, y, x, y, sc["l_x"], sc["l_y"], vals, device=device)
            if _min is not None:
                sm = np.maximum(sm, _min)
            if _max is not None:
                sm = np.minimum(sm, _max)
            tmp = pd.DataFrame({param: sm, x_col: x, y_col: y})
            tmp.dropna(inplace=True)
            for od, kv in zip(other_dims, key):
                tmp[od] = kv
            pieces.append(tmp[col_order])
        smooth_df = pd.concat(pieces)
        smooth_df.set_index(list(coords_col), inplace=True)
        out[f"{hp}{table_suffix}"] = smooth_df
    for param in all_params:
        if param in params_to_smooth:
            continue
        name = f"{param}{reference_table_suffix}"
        if name in tables:
            cp = tables[name].copy(True)
            if cp.index.names == [None]:
                cp.set_index(list(coords_col), inplace=True)
            out[f"{name}{table_suffix}"] = cp
    return out


def smooth_hyperparameters_file(result_file: str, params_to_smooth, smooth_config_dict, coords_col, all_params,
                                output_file: Optional[str] = None, **kwargs):
    """HDF5 front end of :func:`smooth_hyperparameters` (needs PyTables, like the reference): reads
    the parameter tables of ``result_file`` and appends the smoothed ones to ``output_file``."""
    suffix = kwargs.get("reference_table_suffix", "")
    with pd.HDFStore(result_file, mode="r") as store:
        tables = {f"{p}{suffix}": store.select(f"{p}{suffix}") for p in all_params if f"/{p}{suffix}" in store.keys()}
    out = smooth_hyperparameters(tables, params_to_smooth, smooth_config_dict, coords_col, all_params, **kwargs)
    with pd.HDFStore(result_file if output_file is None else output_file, mode="a") as store:
        for k, v in out.items():
            store.put(k, v, format="table", append=False)
    return out


def _glue(preds_df: pd.DataFrame, pred_loc_cols: List[str], xprt_loc_cols: List[str], vars_to_glue, sigma,
          device: int = 0) -> pd.DataFrame:
    if isinstance(vars_to_glue, str):
        vars_to_glue = [vars_to_glue]
    n = len(preds_df)
    keys = preds_df[pred_loc_cols]
    # groups in sorted key order, as DataFrame.groupby(...).sum() returns them; rows keep their order inside a group
    order = np.lexsort([keys[c].values for c in reversed(pred_loc_cols)]) if n else np.zeros(0, np.int64)
    sk = keys.values[order]
    new = np.ones(n, bool)
    if n > 1:
        new[1:] = np.any(sk[1:] != sk[:-1], axis=1)
    seg = np.concatenate([np.flatnonzero(new), [n]]).astype(np.int64)
    G = len(seg) - 1
    pred = _f64(preds_df[pred_loc_cols].values[order].T)
    xprt = _f64(preds_df[xprt_loc_cols].values[order].T)
    sig = _f64(np.broadcast_to(np.asarray(sigma, dtype=np.float64), (n,))[order] if np.ndim(sigma) else
               np.full(n, float(sigma)))
    vals = _f64(preds_df[vars_to_glue].values[order].T)
    out = np.empty((len(vars_to_glue), G))
    eng = get_engine(device)
    _lib.check(eng._lib.gpsat_glue_predictions(eng._h, _lib.MEM_HOST, n, G, _p(seg), len(pred_loc_cols), _p(pred),
                                               _p(xprt), _p(sig), len(vars_to_glue), _p(vals), _p(out)))
    res = pd.DataFrame(sk[seg[:-1]], columns=pred_loc_cols)
    for k, v in enumerate(vars_to_glue):
        res[v] = out[k]
    return res


def glue_local_predictions_1d(preds_df: pd.DataFrame, pred_loc_col: str, xprt_loc_col: str, vars_to_glue,
                              inference_radius, R=3, device: int = 0) -> pd.DataFrame:
    """Gaussian-weighted blend of overlapping expert predictions in 1-D (``postprocessing.py:462-530``);
    ``inference_radius`` is a number or a dict keyed by expert location."""
    if isinstance(inference_radius, dict):
        assert len(inference_radius) == len(preds_df[xprt_loc_col].unique())
        radius = np.array([inference_radius[loc] for loc in preds_df[xprt_loc_col]], dtype=np.float64)
    else:
        radius = float(inference_radius)
    return _glue(preds_df, [pred_loc_col], [xprt_loc_col], vars_to_glue, radius / R, device)


def glue_local_predictions_2d(preds_df: pd.DataFrame, pred_loc_cols: List[str], xprt_loc_cols: List[str],
                              vars_to_glue, inference_radius, R=3, device: int = 0) -> pd.DataFrame:
    """The 2-D version (``postprocessing.py:533-593``): product of the per-axis Gaussian weights."""
    return _glue(preds_df, list(pred_loc_cols), list(xprt_loc_cols), vars_to_glue, float(inference_radius) / R, device)
