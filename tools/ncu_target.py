"""Short single-GPU command for ncu captures: a few batched lml+grad evaluations and one predict
pass over c2-shaped experts (one evaluation = T x (diag_kernel, step_kernel) + lauum_grad_kernel).

  python tools/ncu_target.py [E] [n_lo] [n_hi] [evals]
"""
import sys

import numpy as np

sys.path.insert(0, ".")
from gpsat_b200 import synthetic as syn
from gpsat_b200.engine import Engine

E = int(sys.argv[1]) if len(sys.argv) > 1 else 296
n_lo = int(sys.argv[2]) if len(sys.argv) > 2 else 800
n_hi = int(sys.argv[3]) if len(sys.argv) > 3 else 1200
evals = int(sys.argv[4]) if len(sys.argv) > 4 else 2
spec = syn.WorkloadSpec("ncu", E, n_lo, n_hi, 200, 1300, 300e3, 4, (50e3, 50e3, 1.0), (600e3, 600e3, 9.0))
b = syn.make_batch(spec, with_pred=True)
eng = Engine(0)
bd = b.to_device("cuda:0")
theta = np.tile(np.array([2.0, 2.0, 3.0, 0.1, 0.01]), (E, 1))
for it in range(evals):
    lml, grad, st = eng.lml_grad(bd, theta)
    s = eng.stats()
    print("lml_grad", s["gpu_ms"], "ms", s["sweep_flops"] / s["gpu_ms"] / 1e9, "TF/s", "launches", s["kernel_launches"])
for it in range(2):
    p = eng.predict(bd, theta)
    s = eng.stats()
    print("predict", s["gpu_ms"], "ms", s["sweep_flops"] / s["gpu_ms"] / 1e9, "TF/s", "launches", s["kernel_launches"])
assert int(np.sum(st.cpu().numpy() != 0)) == 0
print("ok")
