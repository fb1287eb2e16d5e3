"""Generate tests/golden/*.npz  (run in the build container, where /root/reference is mounted).

Three independent sources pin the oracle (and through it the CUDA path):

1. ``kat_localexperts.npz`` - the reference's own known-answer vector: the module-level fixture
   of ``/root/reference/tests/test_localexperts.py:22-49`` re-executed verbatim (sklearn
   Matern-3/2 fit on 50 points) - optimised lengthscale, LML, predictive mean / std.
2. ``pure_python_gpr.npz`` - the reference's in-repo NumPy GPR
   (``/root/reference/GPSat/models/pure_python_gpr.py:378-553``: ``SGPkernel``, ``SMLII_mod``,
   ``GPR``) executed from the reference tree (functions extracted with ``ast`` because the
   module's own imports pull in TensorFlow) on seeded 3-D ARD inputs.
3. ``sklearn_ard.npz`` - scikit-learn ``GaussianProcessRegressor`` LML, gradient and predictions
   for Matern-1/2, 3/2, 5/2 and RBF with ARD lengthscales (same math, independent code).

4. ``vff_features.npz`` - the NumPy-only pieces of the reference's VFF implementation
   (``/root/reference/GPSat/vff.py:518-559``: ``make_Kuf_np``, ``make_kvs_np``), executed from the
   reference tree, and the static quantities ``GPR_kron.__init__`` derives from them (``:600-610``).
   The TensorFlow parts of vff.py (``make_Kuu``, ``elbo``, ``predict_f``) cannot run here.

5. ``post_smooth.npz`` / ``post_glue.npz`` - the reference's own ``gaussian_2d_weight`` (numba),
   ``glue_local_predictions_1d`` and ``_2d`` (``/root/reference/GPSat/postprocessing.py:22-52,
   462-593``) executed from the reference source on seeded inputs.

Usage:  python oracle/make_golden.py
"""
from __future__ import annotations

import ast
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "tests", "golden")
REF = "/root/reference"


def kat_localexperts():
    from sklearn.gaussian_process import GaussianProcessRegressor
    from sklearn.gaussian_process.kernels import Matern

    # --- verbatim flow of tests/test_localexperts.py:22-49 ---
    np.random.seed(23435)
    kernel = Matern(length_scale=0.8, nu=3 / 2)
    gp = GaussianProcessRegressor(kernel)
    x = np.linspace(0, 10, 100)[:, None]
    f = gp.sample_y(x, random_state=0)
    N = 50
    eps = 1e-2
    indices = np.arange(100)
    np.random.shuffle(indices)
    x_train = x[indices[:N]]
    y_train = f[indices[:N]] + eps * np.random.randn(N, 1)
    gp.alpha = eps ** 2
    gp.fit(x_train, y_train)
    ls = gp.kernel_.length_scale
    ml = gp.log_marginal_likelihood()
    test_index = np.random.randint(0, 99)
    x_test = x[[test_index]]
    pred_mean, pred_std = gp.predict(x_test, return_std=True)
    np.savez(os.path.join(OUT, "kat_localexperts.npz"), x_train=x_train, y_train=y_train, eps=eps,
             ls=ls, ml=ml, x_test=x_test, pred_mean=pred_mean, pred_std=pred_std,
             constraint_low=1e-10, constraint_high=5.0)
    print("kat:", ls, ml, test_index, pred_mean, pred_std ** 2)


def _load_pure_python_functions():
    """Extract SGPkernel / SMLII_mod / GPR / Nystroem from the reference without importing GPSat."""
    path = os.path.join(REF, "GPSat", "models", "pure_python_gpr.py")
    tree = ast.parse(open(path).read())
    wanted = {"SGPkernel", "SMLII_mod", "GPR", "Nystroem"}
    body = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in wanted]
    mod = ast.Module(body=body, type_ignores=[])
    import scipy
    from numpy.linalg import multi_dot as mdot
    from scipy.spatial.distance import cdist, pdist, squareform
    ns = {"np": np, "scipy": scipy, "mdot": mdot, "squareform": squareform, "pdist": pdist, "cdist": cdist}
    exec(compile(mod, path, "exec"), ns)
    return ns


def pure_python():
    ns = _load_pure_python_functions()
    rng = np.random.default_rng(20240517)
    n, m = 120, 40
    X = rng.uniform(-3, 3, (n, 3))
    Xs = rng.uniform(-3, 3, (m, 3))
    y = np.sin(X[:, 0]) + 0.5 * np.cos(2 * X[:, 1]) + 0.1 * X[:, 2] + 0.1 * rng.standard_normal(n)
    thetas = np.array([[1.0, 1.0, 1.0, 1.0, 1.0],
                       [1.7, 0.6, 2.5, 0.8, 0.05],
                       [0.4, 3.0, 1.2, 2.0, 0.3]])
    nlz, fs, fvar = [], [], []
    for th in thetas:
        nlz.append(ns["SMLII_mod"](hypers=th, x=X, y=y, approx=False, M=None, grad=False))
        mean, sd = ns["GPR"](x=X, y=y, xs=Xs, ell=th[:3], sf2=th[3], sn2=th[4], mean=0, approx=False, M=None)
        fs.append(np.asarray(mean).ravel())
        fvar.append(np.asarray(sd).ravel() ** 2)
    np.savez(os.path.join(OUT, "pure_python_gpr.npz"), X=X, y=y, Xs=Xs, thetas=thetas,
             nlZ=np.array(nlz), fmean=np.array(fs), fvar=np.array(fvar))
    print("pure_python nlZ:", nlz)


def sklearn_ard():
    from sklearn.gaussian_process import GaussianProcessRegressor
    from sklearn.gaussian_process.kernels import RBF, ConstantKernel, Matern

    rng = np.random.default_rng(77)
    n, m = 150, 30
    X = rng.uniform(-4, 4, (n, 3))
    Xs = rng.uniform(-4, 4, (m, 3))
    y = np.sin(X[:, 0]) * np.cos(0.5 * X[:, 1]) + 0.2 * X[:, 2] + 0.15 * rng.standard_normal(n)
    theta = np.array([1.3, 0.7, 2.0, 0.8, 0.1])
    out = dict(X=X, y=y, Xs=Xs, theta=theta)
    for name, base in (("Matern12", Matern(length_scale=theta[:3], nu=0.5)),
                       ("Matern32", Matern(length_scale=theta[:3], nu=1.5)),
                       ("Matern52", Matern(length_scale=theta[:3], nu=2.5)),
                       ("RBF", RBF(length_scale=theta[:3]))):
        k = ConstantKernel(theta[3]) * base
        gp = GaussianProcessRegressor(k, alpha=theta[4], optimizer=None).fit(X, y)
        # sklearn theta = log([constant, l1, l2, l3]); gradient is w.r.t. log-parameters
        lml, g = gp.log_marginal_likelihood(gp.kernel_.theta, eval_gradient=True)
        mean, sd = gp.predict(Xs, return_std=True)
        out[f"{name}_lml"] = lml
        # convert to d/d(constrained): dlml/dp = g / p ; order -> [l1, l2, l3, var]
        out[f"{name}_grad_ls"] = g[1:4] / theta[:3]
        out[f"{name}_grad_var"] = g[0] / theta[3]
        out[f"{name}_fmean"] = mean
        out[f"{name}_fvar"] = sd ** 2
        print(name, lml)
    np.savez(os.path.join(OUT, "sklearn_ard.npz"), **out)


def vff_features():
    path = os.path.join(REF, "GPSat", "vff.py")
    tree = ast.parse(open(path).read())
    wanted = {"make_Kuf_np", "make_kvs_two_np", "make_kvs_np"}
    body = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in wanted]
    from functools import reduce
    ns = {"np": np, "reduce": reduce}
    exec(compile(ast.Module(body=body, type_ignores=[]), path, "exec"), ns)
    rng = np.random.default_rng(515)
    n = 70
    X = rng.uniform(-1.0, 1.0, (n, 3)) * np.array([2.0, 3.0, 1.0])
    Y = rng.standard_normal((n, 1))
    a = np.array([-2.5, -3.5, -1.25])
    b = np.array([2.75, 3.25, 1.5])
    ms = [np.arange(4), np.arange(3), np.arange(2)]
    # GPR_kron.__init__, vff.py:600-610
    Kuf = ns["make_kvs_np"]([ns["make_Kuf_np"](X[:, i:i + 1], a[i], b[i], ms[i]) for i in range(3)])
    np.savez(os.path.join(OUT, "vff_features.npz"), X=X, Y=Y, a=a, b=b, m0=ms[0], m1=ms[1], m2=ms[2],
             Kuf0=ns["make_Kuf_np"](X[:, 0:1], a[0], b[0], ms[0]), Kuf=Kuf, KufY=np.dot(Kuf, Y),
             KufKfu=np.dot(Kuf, Kuf.T), tr_YTY=np.sum(np.square(Y)))
    print("vff Kuf:", Kuf.shape)


def postprocessing():
    import numba as nb
    import pandas as pd
    from typing import Dict, List, Union
    from scipy.stats import norm
    path = os.path.join(REF, "GPSat", "postprocessing.py")
    tree = ast.parse(open(path).read())
    wanted = {"gaussian_2d_weight", "glue_local_predictions_1d", "glue_local_predictions_2d"}
    body = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in wanted]
    ns = {"np": np, "nb": nb, "pd": pd, "norm": norm, "Union": Union, "List": List, "Dict": Dict}
    exec(compile(ast.Module(body=body, type_ignores=[]), path, "exec"), ns)
    rng = np.random.default_rng(99)
    n = 60
    x, y = rng.uniform(-5e5, 5e5, n), rng.uniform(-5e5, 5e5, n)
    vals = rng.uniform(0.5, 3.0, n)
    vals[[3, 17, 40]] = np.nan
    sm = ns["gaussian_2d_weight"](x, y, x, y, 2e5, 1e5, vals)
    far = ns["gaussian_2d_weight"](np.array([1e9]), np.array([1e9]), x, y, 1.0, 1.0, vals)     # all weights underflow
    np.savez(os.path.join(OUT, "post_smooth.npz"), x=x, y=y, vals=vals, l_x=2e5, l_y=1e5, out=sm, far=far)
    # overlapping experts on a grid, predictions on shared locations
    rows = []
    for ex in (0.0, 100.0, 200.0):
        for ey in (0.0, 100.0):
            for px in np.arange(ex - 100, ex + 101, 50.0):
                for py in np.arange(ey - 100, ey + 101, 50.0):
                    rows.append((ex, ey, px, py, rng.normal(), rng.uniform(0.1, 1.0)))
    df = pd.DataFrame(rows, columns=["x", "y", "pred_loc_x", "pred_loc_y", "f*", "f*_var"])
    g2 = ns["glue_local_predictions_2d"](df, ["pred_loc_x", "pred_loc_y"], ["x", "y"], ["f*", "f*_var"], 150.0, R=3)
    d1 = df[df["y"] == 0.0]
    radius = {0.0: 150.0, 100.0: 120.0, 200.0: 90.0}
    g1 = ns["glue_local_predictions_1d"](d1, "pred_loc_x", "x", ["f*", "f*_var"], radius, R=2)
    np.savez(os.path.join(OUT, "post_glue.npz"), df=df.values, cols=np.array(df.columns), g2=g2.values,
             g2cols=np.array(g2.columns), g1=g1.values, g1cols=np.array(g1.columns),
             radius_keys=np.array(list(radius.keys())), radius_vals=np.array(list(radius.values())))
    print("postprocessing:", sm[:3], far, g2.shape, g1.shape)


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    kat_localexperts()
    pure_python()
    sklearn_ard()
    vff_features()
    postprocessing()


# ---- 6. driver-side functions of the reference, executed from source ---------------------
def _extract(path, names, ns, in_class=None):
    """exec the named top-level functions (or methods of `in_class`, decorators stripped) of a reference file."""
    tree = ast.parse(open(path).read())
    body = tree.body
    if in_class is not None:
        body = [n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == in_class][0].body
    fns = [n for n in body if isinstance(n, ast.FunctionDef) and n.name in names]
    assert len(fns) == len(names), f"{path}: found {[f.name for f in fns]} of {names}"
    for f in fns:
        f.decorator_list = []
    exec(compile(ast.Module(body=fns, type_ignores=[]), path, "exec"), ns)
    return ns


def _frame_to_json(df):
    """Lossless JSON form of a result table (floats through repr): columns, dtypes, index names and values."""
    import pandas as pd
    idx = df.index
    names = list(idx.names)
    ivals = [list(map(float, idx.get_level_values(k))) for k in range(idx.nlevels)]
    return {"columns": list(map(str, df.columns)), "dtypes": [str(df[c].dtype) for c in df.columns],
            "index_names": names, "index": ivals,
            "data": [df[c].tolist() for c in df.columns]}


def driver_functions():
    import json
    import warnings
    import numba as nb
    import pandas as pd
    from scipy.spatial import KDTree
    util = os.path.join(REF, "GPSat", "utils.py")
    ns = {"np": np, "pd": pd, "warnings": warnings, "KDTree": KDTree, "nb": nb}
    _extract(util, {"pandas_to_dict", "array_to_dataframe", "dict_of_array_to_dict_of_dataframe"}, ns)
    _extract(os.path.join(REF, "GPSat", "local_experts.py"), {"dict_of_array_to_table"}, ns, in_class="LocalExpertOI")
    _extract(os.path.join(REF, "GPSat", "dataloader.py"), {"local_data_select"}, ns, in_class="DataLoader")
    # _max_dist_bool keeps its numba decorator: re-apply the reference's guvectorize signature (prediction_locations.py:18-19)
    _extract(os.path.join(REF, "GPSat", "prediction_locations.py"), {"_max_dist_bool"}, ns)
    max_dist_bool = nb.guvectorize([(nb.float64[:, :], nb.float64[:], nb.float64[:], nb.bool_[:])],
                                   '(n, d),(d), () -> (n)')(ns["_max_dist_bool"])
    out = {}

    # -- DataLoader.local_data_select (dataloader.py:2354-2447) --
    rng = np.random.default_rng(11)
    N = 3000
    df = pd.DataFrame({"x": rng.uniform(-5e5, 5e5, N), "y": rng.uniform(-5e5, 5e5, N),
                       "t": rng.integers(0, 20, N).astype(float), "obs": rng.standard_normal(N)})
    # points exactly on the ball boundary of the first expert: (3, 4, .) * 1e5 is at distance 5e5 * ... use r = 5e5
    df.loc[0, ["x", "y", "t"]] = [3e5, 4e5, 10.0]
    df.loc[1, ["x", "y", "t"]] = [-3e5, -4e5, 14.0]         # on the ball and on the t <= +4 edge
    df.loc[2, ["x", "y", "t"]] = [0.0, 0.0, 5.0]            # below the t >= -4 edge by one
    locs = pd.DataFrame({"x": [0.0, 1e5, -2e5, 4e5, 0.0, 9e5], "y": [0.0, -1e5, 2e5, 4e5, 0.0, 9e5],
                         "t": [10.0, 3.0, 18.0, 10.0, 0.0, 10.0]})
    selects = {"example": [{"col": "t", "comp": "<=", "val": 4}, {"col": "t", "comp": ">=", "val": -4},
                           {"col": ["x", "y"], "comp": "<", "val": 5e5}],
               "strict_t": [{"col": "t", "comp": "<", "val": 4}, {"col": "t", "comp": ">", "val": -4},
                            {"col": ["x", "y"], "comp": "<=", "val": 3e5}],
               "ball3d": [{"col": ["x", "y", "t"], "comp": "<", "val": 2.5e5}, {"col": "t", "comp": "==", "val": 0}],
               "empty": []}
    sel = {}
    for name, ls in selects.items():
        for e in range(len(locs)):
            got = ns["local_data_select"](None, df, locs.iloc[e], ls, kdtree=None, verbose=False)
            sel[f"{name}/{e}"] = got.index.values.astype(np.int64)
    np.savez_compressed(os.path.join(OUT, "driver_select.npz"), data=df.values, locs=locs.values,
                        selects=json.dumps(selects), **sel)
    print("local_data_select:", {k: len(v) for k, v in list(sel.items())[:6]})

    # -- _max_dist_bool (prediction_locations.py:18-43) --
    g = np.linspace(-3e5, 3e5, 25)
    grid = np.stack([np.repeat(g, 25), np.tile(g, 25)], axis=1)
    refs = np.array([[0.0, 0.0], [1e5, -5e4], [2.9e5, 2.9e5], [9e5, 9e5]])
    md = {}
    for k, ref in enumerate(refs):
        for r in (1e5, 2e5):                                   # grid spacing 25 km: points exactly at distance r exist
            md[f"{k}/{int(r)}"] = max_dist_bool(grid, ref, r)
    np.savez_compressed(os.path.join(OUT, "driver_max_dist.npz"), grid=grid, refs=refs, **md)
    print("_max_dist_bool:", {k: int(v.sum()) for k, v in md.items()})

    # -- LocalExpertOI.dict_of_array_to_table (local_experts.py:691-747) as run() calls it (:1226-1237) --
    rl = pd.Series({"x": 1.5e5, "y": -2.5e4, "t": 18326.0})
    m = 7
    pred = {"f*": rng.standard_normal(m), "f*_var": rng.uniform(0.1, 1, m), "y_var": rng.uniform(0.1, 1, m),
            "f_bar": np.repeat(np.array([0]), m), "pred_loc_x": rng.uniform(-1, 1, m), "pred_loc_y": rng.uniform(-1, 1, m),
            "pred_loc_t": np.full(m, 18326.0)}
    run_details = {"num_obs": 412, "run_time": 1.25, "objective_value": -321.5, "parameters_optimised": True,
                   "optimise_success": False, "model": "GPSat.models.gpflow_models.GPflowGPRModel", "device": "NVIDIA B200",
                   "config_id": 1}
    hypes = {"lengthscales": np.array([1.5, 2.5, 3.5]), "kernel_variance": 0.75, "likelihood_variance": 0.02,
             "inducing_points": rng.standard_normal((4, 3))}
    tabs = {}
    tabs["preds"] = ns["dict_of_array_to_table"](dict(pred), ref_loc=rl, concat=True, table="preds")
    tabs["run_details"] = ns["dict_of_array_to_table"](dict(run_details), ref_loc=rl, concat=True, table="run_details")
    tabs["hypes"] = ns["dict_of_array_to_table"](dict(hypes), ref_loc=rl, concat=False)
    tabs["one_d_ref"] = ns["dict_of_array_to_table"]({"kernel_variance": 0.5}, ref_loc=pd.Series({"x": 2.0}), concat=False)
    gold = {"ref_loc": {k: float(v) for k, v in rl.items()},
            "inputs": {"pred": {k: v.tolist() for k, v in pred.items()}, "run_details": run_details,
                       "hypes": {k: (v.tolist() if isinstance(v, np.ndarray) else v) for k, v in hypes.items()}},
            "tables": {grp: {k: _frame_to_json(v) for k, v in t.items()} for grp, t in tabs.items()}}
    with open(os.path.join(OUT, "driver_tables.json"), "w") as f:
        json.dump(gold, f)
    for grp, t in tabs.items():
        for k, v in t.items():
            print("table", grp, k, list(v.columns), v.shape)


if __name__ == "__main__" and "--driver" in __import__("sys").argv:
    driver_functions()
