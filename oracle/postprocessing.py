"""NumPy / pandas restatement of the reference's post-processing arithmetic (oracle; test infrastructure only).

* ``gaussian_2d_weight``           /root/reference/GPSat/postprocessing.py:22-52 (numba guvectorize there)
* ``smooth_table``                 the per-parameter loop of ``smooth_hyperparameters``, :247-330
* ``glue_1d`` / ``glue_2d``        ``glue_local_predictions_1d`` / ``_2d``, :462-593

Pinning: numba is importable in the build container, so ``oracle/make_golden.py`` executes the
reference's own ``gaussian_2d_weight`` source and stores its output in
``tests/golden/post_smooth.npz``; the glue functions are plain pandas / scipy and are executed
from the reference source the same way.
"""
from __future__ import annotations

import re

import numpy as np
import pandas as pd
from scipy.stats import norm


def gaussian_2d_weight(x0, y0, x, y, l_x, l_y, vals):
    x0, y0, x, y, vals = (np.asarray(a, dtype=np.float64) for a in (x0, y0, x, y, vals))
    out = np.empty(len(x0))
    ok = ~np.isnan(vals)
    for i in range(len(x0)):
        d2 = ((x - x0[i]) / l_x) ** 2 + ((y - y0[i]) / l_y) ** 2
        w = np.exp(-d2 / 2)
        ws = np.sum(w[ok])
        out[i] = np.nan if ws == 0 else np.sum(w[ok] * vals[ok]) / ws
    return out


def smooth_table(df: pd.DataFrame, param: str, coords_col, xy_dims, l_x, l_y, vmin=None, vmax=None) -> pd.DataFrame:
    """One hyper-parameter table (coordinates as columns) -> smoothed table indexed by coords_col."""
    other = [c for c in coords_col if c not in xy_dims] + [c for c in df.columns if re.search(r"^_dim_\d", c)]
    dim_cols = [c for c in df.columns if re.search(r"^_dim_\d", c)]
    order = df.columns.tolist()
    parts = []
    uniq = df[other].drop_duplicates() if other else pd.DataFrame(index=[0])
    for _, row in uniq.iterrows():
        sub = df
        for c in other:
            sub = sub[sub[c] == row[c]]
        vals = sub[param].values.astype(float).copy()
        lo, hi = vmin, vmax
        if isinstance(lo, (list, np.ndarray)) and dim_cols:
            lo = lo[min(int(row[dim_cols[-1]]), len(lo) - 1)]
        if isinstance(hi, (list, np.ndarray)) and dim_cols:
            hi = hi[min(int(row[dim_cols[-1]]), len(hi) - 1)]
        if hi is not None:
            vals[vals > hi] = hi
        if lo is not None:
            vals[vals < lo] = lo
        x, y = sub[xy_dims[0]].values, sub[xy_dims[1]].values
        sm = gaussian_2d_weight(x, y, x, y, l_x, l_y, vals)
        if lo is not None:
            sm = np.maximum(sm, lo)
        if hi is not None:
            sm = np.minimum(sm, hi)
        tmp = pd.DataFrame({param: sm, xy_dims[0]: x, xy_dims[1]: y}).dropna()
        for c in other:
            tmp[c] = row[c]
        parts.append(tmp[order])
    return pd.concat(parts).set_index(list(coords_col))


def glue_1d(preds_df, pred_loc_col, xprt_loc_col, vars_to_glue, inference_radius, R=3):
    p = preds_df.copy(deep=True)
    if isinstance(inference_radius, dict):
        inference_radius = np.array([inference_radius[loc] for loc in p[xprt_loc_col]])
    p["total_weights"] = norm.pdf(p[pred_loc_col], p[xprt_loc_col], inference_radius / R)
    for v in vars_to_glue:
        p[v] = p[v] * p["total_weights"]
    g = p[[pred_loc_col, "total_weights"] + list(vars_to_glue)].groupby([pred_loc_col]).sum().reset_index()
    for v in vars_to_glue:
        g[v] = g[v] / g["total_weights"]
    return g.drop("total_weights", axis=1)


def glue_2d(preds_df, pred_loc_cols, xprt_loc_cols, vars_to_glue, inference_radius, R=3):
    p = preds_df.copy(deep=True)
    p["total_weights"] = 1.0
    for pc, xc in zip(pred_loc_cols, xprt_loc_cols):
        p["total_weights"] *= norm.pdf(p[pc], p[xc], inference_radius / R)
    for v in vars_to_glue:
        p[v] = p[v] * p["total_weights"]
    g = p[list(pred_loc_cols) + ["total_weights"] + list(vars_to_glue)].groupby(list(pred_loc_cols)).sum().reset_index()
    for v in vars_to_glue:
        g[v] = g[v] / g["total_weights"]
    return g.drop("total_weights", axis=1)
