"""Throughput of the VFF path and of the predict-only pass (config c4) - measurement aids, not the
bench line.

  python tools/vff_bench.py [--experts 32] [--n 5000] [--nfreq 20 20] [--c4-experts 96]

VFF: ELBO + gradient evaluations per second with inputs resident in HBM (library CUDA-event time);
per evaluation the algorithmic work is M^3 (Cholesky M^3/3 + inverse 2M^3/3), the static
Kuf Kuf^T costs M^2 n once per expert.  c4: gpsat_gpr_predict at fixed theta on c3-shaped experts
(n = 2000..4000), n^3/3 + n^2 m flops per expert.
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--experts", type=int, default=32)
    ap.add_argument("--n", type=int, default=5000)
    ap.add_argument("--nfreq", type=int, nargs="+", default=[20, 20])
    ap.add_argument("--c4-experts", type=int, default=96)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--peak", type=float, default=37.13)
    args = ap.parse_args()
    from gpsat_b200 import synthetic as syn
    from gpsat_b200.engine import Engine, RaggedBatch

    eng = Engine(0)
    out = {}
    # ---- VFF ----
    E, D = args.experts, len(args.nfreq)
    rng = np.random.default_rng(3)
    Xl, yl, Sl = [], [], []
    for e in range(E):
        X = rng.uniform(-4.0, 4.0, (args.n, D))
        y = np.sin(X[:, 0]) + 0.3 * np.cos(2 * X[:, -1]) + 0.1 * rng.standard_normal(args.n)
        Xl.append(X); yl.append(y - y.mean()); Sl.append(rng.uniform(-3.5, 3.5, (500, D)))
    b = RaggedBatch.from_lists("Matern32", Xl, yl, Sl).to_device("cuda:0")
    dom = (np.full((E, D), -5.0), np.full((E, D), 5.0), np.tile(np.asarray(args.nfreq, np.int32), (E, 1)))
    M = float(np.prod(2 * np.asarray(args.nfreq) - 1))
    theta = np.tile(np.concatenate([np.full(D, 1.0), np.full(D, 0.8), [0.02]]), (E, 1))
    for name, wg, flops in (("elbo", False, M ** 3 / 3), ("elbo_grad", True, M ** 3)):
        eng.vff_elbo(b, dom, theta, with_grad=wg)
        ms = []
        for _ in range(args.reps):
            eng.vff_elbo(b, dom, theta, with_grad=wg)
            ms.append(eng.stats()["gpu_ms"])
        t = float(np.median(ms)) * 1e-3
        out[f"vff_{name}"] = {"evals_per_s": E / t, "ms": t * 1e3,
                              "tflops_incl_static": E * (flops + M * M * args.n) / t / 1e12,
                              "frac_of_fp64_peak": E * (flops + M * M * args.n) / t / 1e12 / args.peak}
    th0 = np.ones((E, 2 * D + 1))
    tk = np.zeros((E, 2 * D + 1), np.int32); ta = np.zeros((E, 2 * D + 1)); tb = np.zeros((E, 2 * D + 1))
    ta[:, -1] = 1e-6
    res = eng.fit(b, th0, tk, ta, tb, np.ones((E, 2 * D + 1), np.uint8), vff=dom)
    st = eng.stats()
    nev = res.n_evals.cpu().numpy()
    out["vff_fit"] = {"experts_per_s": E / (st["gpu_ms"] * 1e-3), "ms": st["gpu_ms"], "evals": int(nev.sum()),
                      "evals_per_s": float(nev.sum()) / (st["gpu_ms"] * 1e-3),
                      "status_ok": int(np.sum(res.status.cpu().numpy() == 0))}
    out["vff_workload"] = f"{E} experts x n={args.n}, D={D}, nfreq={args.nfreq} (M={int(M)}), Matern32"
    # ---- c4: predict only ----
    E4 = args.c4_experts
    spec = syn.WORKLOADS["c3"]
    hb = syn.make_batch(spec, seed=0, experts=range(E4), with_pred=True)
    db = hb.to_device("cuda:0")
    th = np.tile(np.array([4.0, 4.0, 5.0, 0.05, 0.005]), (E4, 1))
    eng.predict(db, th)
    ms = []
    for _ in range(args.reps):
        eng.predict(db, th)
        ms.append(eng.stats()["gpu_ms"])
    t = float(np.median(ms)) * 1e-3
    n, m = hb.n.astype(float), hb.m.astype(float)
    fl = float(np.sum(n ** 3 / 3 + n * n * m))
    out["c4_predict_only"] = {"workload": f"{E4} experts of c3_pan_arctic (n 2000..4000), fixed theta",
                              "experts_per_s": E4 / t, "ms": t * 1e3, "tflops": fl / t / 1e12,
                              "frac_of_fp64_peak": fl / t / 1e12 / args.peak}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
