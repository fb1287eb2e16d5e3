"""Short single-GPU command for ncu captures of the sparse-GPR path: two ELBO + gradient evaluations
of c5-shaped experts (n = 20 000, M = 500).  Per evaluation sgpr_gemm_kernel is launched 8 times, in
the order A, SYRK, V, H, GC, BINV, TT, GK (see engine_sgpr.inl).

  python tools/ncu_target_sgpr.py [E] [n] [M]
"""
import sys

import numpy as np

sys.path.insert(0, ".")
from gpsat_b200 import synthetic as syn
from gpsat_b200.engine import Engine, RaggedBatch

E = int(sys.argv[1]) if len(sys.argv) > 1 else 8
n = int(sys.argv[2]) if len(sys.argv) > 2 else 20000
M = int(sys.argv[3]) if len(sys.argv) > 3 else 500
spec = syn.WorkloadSpec("ncu_sgpr", E, n, n, 100, 100, 500e3, 15, (25e3, 25e3, 1.0), (600e3, 600e3, 30.0))
Xl, yl, Zl = [], [], []
for e in range(E):
    X, y, _ = syn.make_expert(spec, e, 5, with_pred=False)
    Xl.append(X); yl.append(y); Zl.append(X[:M].copy())
b = RaggedBatch.from_lists("Matern32", Xl, yl).to_device("cuda:0")
eng = Engine(0)
theta = np.tile(np.array([4.0, 4.0, 5.0, 0.05, 0.005]), (E, 1))
for it in range(2):
    v, g, st = eng.sgpr_elbo(b, Zl, theta, with_grad=True)
    s = eng.stats()
    print("sgpr elbo+grad", s["gpu_ms"], "ms", s["sweep_flops"] / s["gpu_ms"] / 1e9, "TF/s", "launches", s["kernel_launches"])
assert int(np.sum(st.cpu().numpy() != 0)) == 0
print("ok")
