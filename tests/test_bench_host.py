"""Host-side logic of bench.py that needs no GPU: the partition of a global expert set over ranks, the CPU arm's
job construction and its single-threaded BLAS guarantee (SciPy's own OpenBLAS used to stay threaded)."""
import os
import sys
from dataclasses import replace

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import bench                                   # noqa: E402
from gpsat_b200 import synthetic as syn        # noqa: E402


def test_partition_covers_every_expert_once_and_balances_cost():
    spec = syn.WORKLOADS["c2"]
    for world in (1, 2, 4, 8):
        parts, info = bench.partition(spec, 64 * world, world, "fit")
        allidx = np.sort(np.concatenate(parts))
        np.testing.assert_array_equal(allidx, np.arange(64 * world))
        assert info["experts_per_rank"] == [len(p) for p in parts]
        assert info["cost_imbalance_max_over_mean"] < 1.02
    # equal-size experts (north-star set): equal counts
    parts, info = bench.partition(syn.WORKLOADS["ns"], 2048, 8, "fit")
    assert info["experts_per_rank"] == [256] * 8


def test_expert_sizes_match_generated_experts():
    spec = replace(syn.WORKLOADS["c2"], n_lo=20, n_hi=60, m_lo=3, m_hi=9)
    n, m = syn.expert_sizes(spec, [0, 5, 17])
    for k, e in enumerate([0, 5, 17]):
        X, y, Xs = syn.make_expert(spec, e)
        assert (len(X), len(Xs)) == (n[k], m[k])


def test_cpu_pool_runs_single_threaded_blas_and_all_modes():
    from oracle.cpu_baseline import CpuPool
    saved = os.environ.pop("OMP_NUM_THREADS", None)          # the defect only showed with the variable unset
    try:
        spec = replace(syn.WORKLOADS["c2"], n_lo=120, n_hi=140, m_lo=5, m_hi=6)
        pool = CpuPool(2)
        try:
            res, _ = pool.run(bench.cpu_jobs(spec, [0, 1], "fit"), bench.cpu_fn("fit"))
            blas = CpuPool.blas_report(res)
            assert blas and all(t == 1 for _, t in blas)
            assert any("scipy_openblas" in name for name, _ in blas)
            assert all(r["nev"] > 0 and np.isfinite(r["lml"]) for r in res)
            res_p, _ = pool.run(bench.cpu_jobs(spec, [0, 1], "predict"), bench.cpu_fn("predict"))
            assert all(r["nev"] == 0 and np.allclose(r["theta"], syn.FIXED_THETA) for r in res_p)
            spec5 = replace(syn.WORKLOADS["c5"], n_lo=300, n_hi=300, m_lo=4, m_hi=4)
            jobs = bench.cpu_jobs(spec5, [0], "sgpr")
            assert jobs[0][4].shape == (300, 3)               # fewer observations than M: every point is inducing
            res_s, _ = pool.run(jobs, bench.cpu_fn("sgpr"))
            assert np.isfinite(res_s[0]["lml"]) and len(res_s[0]["fmean"]) == 4
        finally:
            pool.close()
        assert "OMP_NUM_THREADS" not in os.environ           # the parent's environment is restored
    finally:
        if saved is not None:
            os.environ["OMP_NUM_THREADS"] = saved
