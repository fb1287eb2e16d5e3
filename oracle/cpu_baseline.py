"""Timed CPU baseline: the oracle port of the reference path on the host cores.

Used only by ``bench.py`` (``cpu_baseline`` leg and ``--impl reference``).  One worker process per
core, each running whole experts with single-threaded BLAS - the most favourable way to use the
host for this embarrassingly parallel loop (the reference itself runs the experts serially,
``GPSat/local_experts.py:930``, with TensorFlow intra-op threads).
Per expert it does what the reference does: SciPy L-BFGS-B on -LML with the same transforms
(``gpflow_models.py:317-321``), then the objective value (``:337``) and ``predict`` (``:229-230``).
"""
from __future__ import annotations

import os
import time

import numpy as np


def _init_worker():
    try:
        from threadpoolctl import threadpool_limits
        threadpool_limits(1)
    except Exception:
        os.environ["OMP_NUM_THREADS"] = "1"


def fit_predict_one(args):
    kind, X, y, Xs, theta0, tkind, ta, tb, trainable, max_iter = args
    from oracle import gpr
    tfm = gpr.Transform(tkind, ta, tb)
    t0 = time.perf_counter()
    theta, lml, ok, nev = gpr.fit(kind, X, y, theta0, tfm, trainable=np.asarray(trainable, bool), max_iter=max_iter)
    pred = gpr.predict(kind, X, y, theta, Xs) if len(Xs) else {"f*": np.zeros(0), "f*_var": np.zeros(0)}
    return dict(theta=theta, lml=lml, ok=ok, nev=nev, fmean=pred["f*"], fvar=pred["f*_var"],
                seconds=time.perf_counter() - t0, n=len(X))


class CpuPool:
    def __init__(self, procs: int):
        import multiprocessing as mp
        self.procs = procs
        self.pool = mp.get_context("spawn").Pool(procs, initializer=_init_worker)

    def run(self, jobs):
        """jobs: list of fit_predict_one argument tuples.  Returns (results, wall_seconds)."""
        t0 = time.perf_counter()
        res = self.pool.map(fit_predict_one, jobs, chunksize=1)
        return res, time.perf_counter() - t0

    def close(self):
        self.pool.close()
        self.pool.join()
