"""Throughput of the sparse-GPR path on the c5 shape (n ~ 20 000 observations per expert,
M = 500 inducing points, gpflow_models.py:704 default) - a measurement aid, not the bench line.

  python tools/sgpr_bench.py [--experts 16] [--n 20000] [--M 500] [--reps 3] [--fit]

Prints one JSON object: ELBO-only and ELBO+gradient evaluations per second with inputs resident in
HBM (CUDA-event time inside the library, gpsat_engine_stats.gpu_ms), the algorithmic FP64 flop rate
(forward 2 M^2 n + 2 M^3 / 3, gradient + 3 M^2 n + 4 M^3 per evaluation) against the measured FP64
peak, and optionally a full fit + predict.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--experts", type=int, default=16)
    ap.add_argument("--n", type=int, default=20000)
    ap.add_argument("--M", type=int, default=500)
    ap.add_argument("--m", type=int, default=700)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--fit", action="store_true")
    ap.add_argument("--peak", type=float, default=37.13, help="measured FP64 peak, TFLOP/s (tools/fp64_peak.cu)")
    args = ap.parse_args()

    import torch
    from gpsat_b200 import synthetic as syn
    from gpsat_b200.engine import Engine, RaggedBatch

    spec = syn.WorkloadSpec("c5_sgpr", args.experts, args.n, args.n, args.m, args.m, 500e3, 15, (25e3, 25e3, 1.0),
                            (600e3, 600e3, 30.0))
    Xl, yl, Sl, Zl = [], [], [], []
    for e in range(args.experts):
        X, y, Xs = syn.make_expert(spec, e, 5)
        rng = np.random.default_rng([5, e, 1])
        Xl.append(X); yl.append(y); Sl.append(Xs)
        Zl.append(X[rng.permutation(len(X))[:args.M]].copy())
    batch = RaggedBatch.from_lists("Matern32", Xl, yl, Sl).to_device("cuda:0")
    eng = Engine(0)
    E = args.experts
    theta = np.tile(np.array([4.0, 4.0, 5.0, 0.05, 0.005]), (E, 1))
    out = {"workload": f"c5 shape: {E} experts x n={args.n}, M={args.M}, D=3, Matern32", "experts": E}
    n, M = float(args.n), float(args.M)
    f_fwd = 2 * M * M * n + 2 * M ** 3 / 3
    f_grad = f_fwd + 3 * M * M * n + 4 * M ** 3
    for name, with_grad, flops in (("elbo", False, f_fwd), ("elbo_grad", True, f_grad)):
        eng.sgpr_elbo(batch, Zl, theta, with_grad=with_grad)       # warm-up (allocations)
        ms = []
        for _ in range(args.reps):
            eng.sgpr_elbo(batch, Zl, theta, with_grad=with_grad)
            ms.append(eng.stats()["gpu_ms"])
        t = float(np.median(ms)) * 1e-3
        out[name] = {"evals_per_s": E / t, "ms": t * 1e3, "tflops": E * flops / t / 1e12,
                     "frac_of_fp64_peak": E * flops / t / 1e12 / args.peak, "slots": eng.stats()["slots"]}
    eng.sgpr_predict(batch, Zl, theta)
    eng.sgpr_predict(batch, Zl, theta)
    out["predict"] = {"experts_per_s": E / (eng.stats()["gpu_ms"] * 1e-3), "ms": eng.stats()["gpu_ms"]}
    if args.fit:
        theta0, tkind, ta, tb, trainable = syn.default_fit_setup(spec, E)
        t0 = time.perf_counter()
        res = eng.fit(batch, theta0, tkind, ta, tb, trainable, inducing=Zl)
        torch.cuda.synchronize()
        st = eng.stats()
        nev = res.n_evals.cpu().numpy()
        out["fit"] = {"experts_per_s": E / (st["gpu_ms"] * 1e-3), "ms": st["gpu_ms"], "wall_s": time.perf_counter() - t0,
                      "evals": int(nev.sum()), "evals_per_s": float(nev.sum()) / (st["gpu_ms"] * 1e-3),
                      "status": res.status.cpu().numpy().tolist(),
                      "tflops": float(nev.sum()) * f_grad / (st["gpu_ms"] * 1e-3) / 1e12}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
