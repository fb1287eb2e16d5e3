"""Seeded synthetic local-expert batches of the shapes named in BASELINE.json / SURVEY.md 8(d).

The reference's example data blobs are not in the tree (``.MISSING_LARGE_BLOBS``), so benchmarks
and full-size property tests use synthetic experts with the same geometry as
``configs/example_local_expert_oi.json`` (300 km radius, +-4 days, ``coords_scale=[50000, 50000, 1]``)
and the fork's production script (``IS2_GPSat_train.py:833``: 500 km radius, 25 km scale).
Observations are a draw from a Matern-3/2 GP (random Fourier features, O(n)) plus white noise.
Every expert has its own seed, so a shard of the batch reproduces the full batch bit for bit.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, Optional

import numpy as np

from .engine import RaggedBatch


@dataclass(frozen=True)
class WorkloadSpec:
    name: str
    E: int
    n_lo: int
    n_hi: int
    m_lo: int
    m_hi: int
    radius: float          # metres
    t_half: float          # days
    coords_scale: tuple
    ls_high: tuple         # upper lengthscale bounds in raw units (constraints "high")
    kernel: str = "Matern32"


WORKLOADS: Dict[str, WorkloadSpec] = {
    # c1  examples/inline_example.py shape (docs/notebooks/inline_example.ipynb: 25 experts, n=402..591)
    "c1": WorkloadSpec("c1_inline_example", 25, 400, 600, 3000, 5000, 300e3, 4, (50e3, 50e3, 1.0), (600e3, 600e3, 9.0)),
    # c2  configs/example_local_expert_oi.json shape
    "c2": WorkloadSpec("c2_example_local_expert_oi", 500, 800, 1200, 200, 1300, 300e3, 4, (50e3, 50e3, 1.0), (600e3, 600e3, 9.0)),
    # c3 / c4  pan-Arctic 50 km freeboard
    "c3": WorkloadSpec("c3_pan_arctic", 3000, 2000, 4000, 200, 1300, 500e3, 15, (25e3, 25e3, 1.0), (600e3, 600e3, 30.0)),
    # north-star run: 2048 experts x 2000 obs
    "ns": WorkloadSpec("ns_2048x2000", 2048, 2000, 2000, 400, 400, 400e3, 8, (50e3, 50e3, 1.0), (600e3, 600e3, 16.0)),
    # c4  predict-only pass with smoothed hyper-parameters (GPSat/postprocessing.py:365-395): c3 experts at fixed theta
    "c4": WorkloadSpec("c4_predict_only_smoothed", 3000, 2000, 4000, 200, 1300, 500e3, 15, (25e3, 25e3, 1.0), (600e3, 600e3, 30.0)),
    # c5  GPflowSGPRModel large experts: ~20k observations, 500 inducing points (gpflow_models.py:704)
    "c5": WorkloadSpec("c5_sgpr_20k_M500", 16, 20000, 20000, 200, 1300, 500e3, 15, (25e3, 25e3, 1.0), (600e3, 600e3, 30.0)),
}
SGPR_M = 500                                   # num_inducing_points default, gpflow_models.py:704
FIXED_THETA = (4.0, 4.0, 5.0, 0.05, 0.005)     # the generating hyper-parameters: what c4 predicts at


def inducing_points(X, index: int, seed: int = 0, M: int = SGPR_M):
    """First M rows of a seeded shuffle of the coordinates (the reference shuffles unseeded,
    gpflow_models.py:809-819, so parity needs Z passed in explicitly: SURVEY 7.3-7)."""
    rng = np.random.default_rng([seed, int(index), 1])
    return X[rng.permutation(len(X))[:M]].copy()


def _rff_sample(rng, Xs, ls, var, nfeat=256):
    """Draw from a Matern-3/2 GP at scaled inputs Xs via random Fourier features."""
    D = Xs.shape[1]
    nu = 1.5
    z = rng.standard_normal((nfeat, D))
    chi = rng.chisquare(2 * nu, size=(nfeat, 1))
    omega = z * np.sqrt(2 * nu / chi) / np.asarray(ls)[None, :]
    phase = rng.uniform(0, 2 * np.pi, nfeat)
    w = rng.standard_normal(nfeat)
    return np.sqrt(2.0 * var / nfeat) * (np.cos(Xs @ omega.T + phase) @ w)


def make_expert(spec: WorkloadSpec, index: int, seed: int = 0, with_pred: bool = True):
    """(X [n,3] scaled coords, y [n] centred obs, Xs [m,3] scaled prediction coords) of expert `index`."""
    rng = np.random.default_rng([seed, index])
    n = int(rng.integers(spec.n_lo, spec.n_hi + 1))
    m = int(rng.integers(spec.m_lo, spec.m_hi + 1)) if with_pred else 0
    cs = np.asarray(spec.coords_scale)

    def disc(k, radius):
        r = radius * np.sqrt(rng.uniform(0, 1, k))
        a = rng.uniform(0, 2 * np.pi, k)
        return r * np.cos(a), r * np.sin(a)

    x, yv = disc(n, spec.radius)
    t = rng.integers(-int(spec.t_half), int(spec.t_half) + 1, n).astype(float)
    X = np.stack([x, yv, t], axis=1) / cs
    f = _rff_sample(rng, X, ls=(4.0, 4.0, 5.0), var=0.05)
    obs = f + np.sqrt(0.005) * rng.standard_normal(n)
    obs = obs - obs.mean()                      # obs_mean='local' (GPSat/models/base_model.py:195-198)
    xp, yp = disc(m, 0.66 * spec.radius)
    Xs = np.stack([xp, yp, np.zeros(m)], axis=1) / cs
    return X, obs, Xs


def expert_sizes(spec: WorkloadSpec, indices, seed: int = 0, with_pred: bool = True):
    """(n[], m[]) of the experts `indices` without generating their data (the first two draws of
    :func:`make_expert`'s generator) - what the partitioner needs on every rank."""
    n = np.empty(len(indices), np.int64)
    m = np.zeros(len(indices), np.int64)
    for k, e in enumerate(indices):
        rng = np.random.default_rng([seed, int(e)])
        n[k] = rng.integers(spec.n_lo, spec.n_hi + 1)
        if with_pred:
            m[k] = rng.integers(spec.m_lo, spec.m_hi + 1)
    return n, m


def make_batch(spec: WorkloadSpec, seed: int = 0, experts: Optional[range] = None, with_pred: bool = True) -> RaggedBatch:
    idx = range(spec.E) if experts is None else experts
    Xl, yl, Sl = [], [], []
    for e in idx:
        X, y, Xs = make_expert(spec, e, seed, with_pred)
        Xl.append(X); yl.append(y); Sl.append(Xs)
    return RaggedBatch.from_lists(spec.kernel, Xl, yl, Sl if with_pred else None)


def default_fit_setup(spec: WorkloadSpec, E: int):
    """GPflow defaults + the example config's constraints:
    theta0 = 1; lengthscales sigmoid-constrained to [1e-8, ls_high] / coords_scale
    (configs/example_local_expert_oi.json:113-125, scale=True forced at local_experts.py:1113-1114);
    kernel variance softplus; likelihood variance 1e-6 + softplus."""
    D, P = 3, 5
    cs = np.asarray(spec.coords_scale)
    tkind = np.zeros((E, P), np.int32)
    ta = np.zeros((E, P))
    tb = np.zeros((E, P))
    tkind[:, :D] = 1
    ta[:, :D] = 1e-8 / cs
    tb[:, :D] = np.asarray(spec.ls_high) / cs
    ta[:, P - 1] = 1e-6
    theta0 = np.ones((E, P))
    trainable = np.ones((E, P), np.uint8)
    return theta0, tkind, ta, tb, trainable
