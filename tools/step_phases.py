"""Per-phase cycle split of step_kernel (library built with -DGPSAT_STEP_PHASES into build/phases/):
  nvcc <flags of __graft_entry__.NVCC_FLAGS> -DGPSAT_STEP_PHASES -o build/phases/libgpsat_b200.so gpsat_b200/csrc/engine.cu
  GPSAT_B200_LIB=build/phases/libgpsat_b200.so GPSAT_B200_STEP=i8 python tools/step_phases.py [E] [n_lo] [n_hi]
Prints microseconds per tile and phase (thread 0's clock, 1.9 GHz assumed) for potrf and trtri tiles."""
import json
import sys

import numpy as np

sys.path.insert(0, ".")
from gpsat_b200 import synthetic as syn
from gpsat_b200.engine import Engine

E = int(sys.argv[1]) if len(sys.argv) > 1 else 296
n_lo = int(sys.argv[2]) if len(sys.argv) > 2 else 800
n_hi = int(sys.argv[3]) if len(sys.argv) > 3 else 1200
spec = syn.WorkloadSpec("ncu", E, n_lo, n_hi, 200, 1300, 300e3, 4, (50e3, 50e3, 1.0), (600e3, 600e3, 9.0))
b = syn.make_batch(spec, with_pred=False).to_device("cuda:0")
eng = Engine(0)
theta = np.tile(np.array([2.0, 2.0, 3.0, 0.1, 0.01]), (E, 1))
eng.lml_grad(b, theta)
eng.step_phases(reset=True)
eng.lml_grad(b, theta)
s = eng.stats()
ph = eng.step_phases()
names = ["prologue", "main_or_pass0", "drain0", "pass1", "drain1", "c_store", "covariance", "c_wt", "store_slice",
         "syrk_schur", "diag_factor"]
out = {"gpu_ms": s["gpu_ms"]}
for kind, base in (("potrf", 0), ("trtri", 16)):
    tiles = max(ph[base + 15], 1)
    o = {k: round(v / tiles / 1.9e3, 2) for k, v in zip(names, ph[base:base + 11])}
    if base == 0:
        o.update({k: round(ph[j] / tiles / 1.9e3, 2) for k, j in (("diag_panels", 11), ("diag_u_sweep", 12), ("diag_store_slice", 13))})
    o["tiles"] = ph[base + 15]
    o["sum_us_per_tile"] = round(sum(ph[base:base + 11]) / tiles / 1.9e3, 2)
    out[kind] = o
print(json.dumps(out))
