"""Per-phase cycle split of lauum_grad_i8_kernel (library built with -DGPSAT_I8_PHASES):
  nvcc <flags of __graft_entry__.NVCC_FLAGS> -DGPSAT_I8_PHASES -o gpsat_b200/libgpsat_b200.so gpsat_b200/csrc/engine.cu
  python tools/i8_phases.py [E] [n_lo] [n_hi]
Prints microseconds per tile and phase (thread 0's clock, 1.9 GHz assumed)."""
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
eng.i8_phases(reset=True)
eng.lml_grad(b, theta)
ph = eng.i8_phases()
names = ["prologue", "pass0_mma", "drain0", "pass1_mma", "drain1", "q_store", "grad_reduce"]
tiles = max(ph[7], 1)
out = {k: round(v / tiles / 1.9e3, 2) for k, v in zip(names, ph[:7])}
out["tiles"] = ph[7]
out["sum_us_per_tile"] = round(sum(ph[:7]) / tiles / 1.9e3, 2)
print(json.dumps(out))
