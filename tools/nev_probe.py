"""Evaluations per expert of a bench workload with the low-rho tier on / off (which experts take a different search path)."""
import dataclasses
import sys
import numpy as np
sys.path.insert(0, ".")
from gpsat_b200 import synthetic as syn
from gpsat_b200.engine import Engine

name = sys.argv[1] if len(sys.argv) > 1 else "c3"
E = int(sys.argv[2]) if len(sys.argv) > 2 else 96
spec = dataclasses.replace(syn.WORKLOADS[name], E=E)
batch = syn.make_batch(spec, seed=0)
th0, tk, ta, tb, tr = syn.default_fit_setup(spec, batch.E)
eng = Engine(0)
out = {}
for tier in (True, False):
    eng.set_precision_low_tier(tier)
    res = eng.fit(batch, th0, tk, ta, tb, tr)
    out[tier] = res
    print("tier", tier, "evals mean", res.n_evals.mean(), "max", res.n_evals.max(), "argmax", int(res.n_evals.argmax()),
          "status", np.bincount(res.status, minlength=4))
a, b = out[True], out[False]
diff = np.flatnonzero(a.n_evals != b.n_evals)
for e in diff:
    n = int(batch.off_n[e + 1] - batch.off_n[e])
    print("expert", int(e), "n", n, "nev on/off", int(a.n_evals[e]), int(b.n_evals[e]), "lml on/off", a.lml[e], b.lml[e],
          "rel theta diff", float(np.max(np.abs(a.theta[e] - b.theta[e]) / np.abs(b.theta[e]))), "theta", np.round(a.theta[e], 5))
