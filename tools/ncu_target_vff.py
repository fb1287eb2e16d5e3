"""Short single-GPU command for ncu captures of the VFF path: two ELBO + gradient evaluations (the first one builds
the static feature tables and Kuf: vff_feat_kernel, vff_kuf_build_kernel).

  python tools/ncu_target_vff.py [E] [n]
"""
import sys

import numpy as np

sys.path.insert(0, ".")
from gpsat_b200.engine import Engine, RaggedBatch

E = int(sys.argv[1]) if len(sys.argv) > 1 else 16
n = int(sys.argv[2]) if len(sys.argv) > 2 else 20000
D, nfreq = 2, [20, 20]
rng = np.random.default_rng(3)
Xl, yl = [], []
for e in range(E):
    X = rng.uniform(-4.0, 4.0, (n, D))
    y = np.sin(X[:, 0]) + 0.3 * np.cos(2 * X[:, -1]) + 0.1 * rng.standard_normal(n)
    Xl.append(X); yl.append(y - y.mean())
b = RaggedBatch.from_lists("Matern32", Xl, yl).to_device("cuda:0")
dom = (np.full((E, D), -5.0), np.full((E, D), 5.0), np.tile(np.asarray(nfreq, np.int32), (E, 1)))
theta = np.tile(np.concatenate([np.full(D, 1.0), np.full(D, 0.8), [0.02]]), (E, 1))
eng = Engine(0)
for it in range(2):
    v, g, st = eng.vff_elbo(b, dom, theta, with_grad=True)
    print("vff elbo+grad", eng.stats()["gpu_ms"], "ms")
assert int(np.sum(st.cpu().numpy() != 0)) == 0
print("ok")
