"""Dump LML, gradient and predictions of a fixed ragged batch at two parameter sets (low-rho tier and full int8 tier) to an
.npz, for a bitwise comparison of two builds of the library:
  GPSAT_B200_LIB=<lib A> python tools/ab_bitwise.py a.npz;  GPSAT_B200_LIB=<lib B> python tools/ab_bitwise.py b.npz
  python tools/ab_bitwise.py --compare a.npz b.npz
A change that only reorders the exact int32 k-sums of the int8 loops must give identical bits."""
import sys

import numpy as np

if sys.argv[1] == "--compare":
    a, b = np.load(sys.argv[2]), np.load(sys.argv[3])
    bad = [k for k in a.files if not np.array_equal(a[k].view(np.uint8), b[k].view(np.uint8))]
    print({"arrays": len(a.files), "different": bad})
    sys.exit(1 if bad else 0)

sys.path.insert(0, ".")
import torch

from gpsat_b200 import synthetic as syn
from gpsat_b200.engine import Engine


def host(x):
    return x.detach().cpu().numpy() if isinstance(x, torch.Tensor) else np.asarray(x)


spec = syn.WorkloadSpec("ab", 96, 300, 1900, 100, 400, 300e3, 4, (50e3, 50e3, 1.0), (600e3, 600e3, 9.0))
b = syn.make_batch(spec, with_pred=True).to_device("cuda:0")
eng = Engine(0)
out = {}
for name, th in (("low", [2.0, 2.0, 3.0, 1.0, 0.1]), ("full", [2.0, 2.0, 3.0, 3.0, 0.01])):
    theta = np.tile(np.array(th), (spec.E, 1))
    lml, grad, status = eng.lml_grad(b, theta)
    p = eng.predict(b, theta)
    out.update({f"{name}_lml": host(lml), f"{name}_grad": host(grad), f"{name}_status": host(status),
                f"{name}_f": host(p["f*"]), f"{name}_fvar": host(p["f*_var"]), f"{name}_plml": host(p["lml"])})
assert eng.i8_timeouts() == 0
np.savez(sys.argv[1], **out)
print({k: float(np.abs(v).sum()) for k, v in out.items() if k.endswith("lml")})
