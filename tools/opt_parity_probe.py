"""Developer probe: optimised hyper-parameters of the first K c2 bench experts, GPU engine against the CPU oracle
(SciPy L-BFGS-B): per expert the evaluation counts and the relative deviation per component.  Run with
GPSAT_B200_STEP=dmma GPSAT_B200_LAUUM=dmma (and / or a -DGPSAT_LIBM_KERNELS build via GPSAT_B200_LIB) to see which
arithmetic choice moves an end point.  CPU side = checker role of the oracle.

  python tools/opt_parity_probe.py [K]"""
import json
import os
import sys

import numpy as np

sys.path.insert(0, ".")
from bench import cpu_jobs
from gpsat_b200 import synthetic as syn
from gpsat_b200.engine import Engine
from oracle.cpu_baseline import CpuPool


def main():
    K = int(sys.argv[1]) if len(sys.argv) > 1 else 16
    spec = syn.WORKLOADS["c2"]
    pool = CpuPool(min(os.cpu_count() or 1, K))
    cpu, _ = pool.run(cpu_jobs(spec, range(K)))
    pool.close()
    host = syn.make_batch(spec, seed=0, experts=range(K), with_pred=True)
    th0, tk, ta, tb, tr = syn.default_fit_setup(spec, K)
    res = Engine(0).fit(host, th0, tk, ta, tb, tr)
    rows = []
    for e, r in enumerate(cpu):
        rel = np.abs(res.theta[e] - r["theta"]) / np.abs(r["theta"])
        rows.append({"expert": e, "n": int(r["n"]), "nev_gpu": int(res.n_evals[e]), "nev_cpu": int(r["nev"]),
                     "rel_theta": [float(f"{v:.2e}") for v in rel], "rel_lml": float(f"{abs(res.lml[e] - r['lml']) / abs(r['lml']):.2e}"),
                     "theta_cpu": [float(f"{v:.6g}") for v in r["theta"]]})
    worst = max(rows, key=lambda q: max(q["rel_theta"]))
    print(json.dumps({"step": os.environ.get("GPSAT_B200_STEP", "i8"), "lauum": os.environ.get("GPSAT_B200_LAUUM", "i8"),
                      "lib": os.environ.get("GPSAT_B200_LIB", "default"), "max_rel_theta": max(max(q["rel_theta"]) for q in rows),
                      "max_rel_lml": max(q["rel_lml"] for q in rows), "worst": worst,
                      "nev_differs": [q["expert"] for q in rows if q["nev_gpu"] != q["nev_cpu"]]}))


if __name__ == "__main__":
    main()
