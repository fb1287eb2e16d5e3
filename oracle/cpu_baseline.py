"""Timed CPU baseline: the oracle port of the reference path on the host cores.

Used only by ``bench.py`` (``cpu_baseline`` leg and ``--impl reference``).  One worker process per
core, each running whole experts with single-threaded BLAS - the most favourable way to use the
host for this embarrassingly parallel loop (the reference itself runs the experts serially,
``GPSat/local_experts.py:930``, with TensorFlow intra-op threads).
Per expert it does what the reference does: SciPy L-BFGS-B on -LML with the same transforms
(``gpflow_models.py:317-321``), then the objective value (``:337``) and ``predict`` (``:229-230``).

Threading: NumPy and SciPy each bundle their own OpenBLAS, and SciPy's is only loaded when
``scipy.linalg`` is imported.  The pool therefore (i) exports ``*_NUM_THREADS=1`` in the parent before the
workers are spawned, (ii) limits the thread pools again inside the worker after ``oracle.gpr`` (which imports
``scipy.linalg``) is loaded, and (iii) returns ``threadpool_info()`` so the bench line can state, and assert,
that every BLAS in every worker ran with one thread.
"""
from __future__ import annotations

import os
import time

import numpy as np

_THREAD_VARS = ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS", "NUMEXPR_NUM_THREADS")


def _limit_blas():
    """Single-threaded BLAS in this process, whichever libraries are loaded by now.  Returns
    [(library file name, threads)] for every BLAS / OpenMP runtime threadpoolctl can see."""
    try:
        from threadpoolctl import threadpool_info, threadpool_limits
        threadpool_limits(1)
        return [(os.path.basename(d.get("filepath", "?")), int(d.get("num_threads", -1))) for d in threadpool_info()]
    except Exception:
        return []


def _init_worker():
    for v in _THREAD_VARS:
        os.environ[v] = "1"
    from oracle import gpr  # noqa: F401  (loads scipy.linalg and with it SciPy's own OpenBLAS)
    _limit_blas()


def fit_predict_one(args):
    kind, X, y, Xs, theta0, tkind, ta, tb, trainable, max_iter = args
    from oracle import gpr
    blas = _limit_blas()
    tfm = gpr.Transform(tkind, ta, tb)
    t0 = time.perf_counter()
    if max_iter > 0:
        theta, lml, ok, nev = gpr.fit(kind, X, y, theta0, tfm, trainable=np.asarray(trainable, bool), max_iter=max_iter)
    else:   # predict-only pass at fixed hyper-parameters (config c4): objective value + prediction
        theta, ok, nev = np.asarray(theta0, float), True, 0
        lml = gpr.lml(kind, X, y, theta)
    pred = gpr.predict(kind, X, y, theta, Xs) if len(Xs) else {"f*": np.zeros(0), "f*_var": np.zeros(0)}
    return dict(theta=theta, lml=lml, ok=ok, nev=nev, fmean=pred["f*"], fvar=pred["f*_var"],
                seconds=time.perf_counter() - t0, n=len(X), blas=blas)


def sgpr_fit_predict_one(args):
    """Sparse-GPR twin (config c5): SciPy L-BFGS-B on -ELBO with Z fixed, then predict."""
    kind, X, y, Xs, Z, theta0, tkind, ta, tb, trainable, max_iter = args
    from oracle import gpr, sgpr
    blas = _limit_blas()
    tfm = gpr.Transform(tkind, ta, tb)
    t0 = time.perf_counter()
    out = sgpr.fit(kind, X, y, Z, theta0, tfm, trainable=np.asarray(trainable, bool), max_iter=max_iter)
    theta, elbo, ok, nev = out[:4]
    pred = sgpr.predict(kind, X, y, Z, theta, Xs) if len(Xs) else {"f*": np.zeros(0), "f*_var": np.zeros(0)}
    return dict(theta=theta, lml=elbo, ok=ok, nev=nev, fmean=pred["f*"], fvar=pred["f*_var"],
                seconds=time.perf_counter() - t0, n=len(X), blas=blas)


class CpuPool:
    def __init__(self, procs: int):
        import multiprocessing as mp
        self.procs = procs
        saved = {v: os.environ.get(v) for v in _THREAD_VARS}
        for v in _THREAD_VARS:        # spawned workers inherit the environment: set it before they start
            os.environ[v] = "1"
        try:
            self.pool = mp.get_context("spawn").Pool(procs, initializer=_init_worker)
            self.pool.map(_warm, range(procs))      # workers exist (and have read the environment) before it is restored
        finally:
            for v, old in saved.items():
                if old is None:
                    os.environ.pop(v, None)
                else:
                    os.environ[v] = old

    def run(self, jobs, fn=fit_predict_one):
        """jobs: list of argument tuples of ``fn``.  Returns (results, wall_seconds)."""
        t0 = time.perf_counter()
        res = self.pool.map(fn, jobs, chunksize=1)
        return res, time.perf_counter() - t0

    @staticmethod
    def blas_report(results):
        """Distinct (library, threads) pairs seen by the workers; raises if any BLAS ran multi-threaded."""
        seen = sorted({tuple(b) for r in results for b in r.get("blas", [])})
        multi = [b for b in seen if b[1] != 1]
        if multi:
            raise RuntimeError(f"CPU baseline workers ran a multi-threaded BLAS: {multi}")
        return [list(b) for b in seen]

    def close(self):
        self.pool.close()
        self.pool.join()


def _warm(_):
    return os.getpid()
