"""rho = (var + noise) / noise of the optimised c2 / north-star experts (which tier of the precision rule they live in)."""
import sys
import numpy as np
sys.path.insert(0, ".")
from gpsat_b200 import synthetic as syn
from gpsat_b200.engine import Engine

eng = Engine(0)
for name in ("c2", "ns"):
    spec = syn.WORKLOADS[name]
    E = 500 if name == "c2" else 128
    import dataclasses
    batch = syn.make_batch(dataclasses.replace(spec, E=E), seed=0)
    th0, tk, ta, tb, tr = syn.default_fit_setup(spec, batch.E)
    res = eng.fit(batch, th0, tk, ta, tb, tr)
    rho = (res.theta[:, 3] + res.theta[:, 4]) / res.theta[:, 4]
    q = np.quantile(rho, [0, 0.1, 0.25, 0.5, 0.75, 0.9, 1.0])
    print(name, "rho quantiles 0/10/25/50/75/90/100:", np.round(q, 1), "share <= 16/32/64/128:",
          [float(np.mean(rho <= t)) for t in (16, 32, 64, 128)], "evals", res.n_evals.mean())
