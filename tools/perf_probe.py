"""Developer perf probe: batched lml+grad evals/s and a small fit on the GPU."""
import sys, time, json
import numpy as np
sys.path.insert(0, ".")
import torch
from gpsat_b200.engine import Engine, RaggedBatch
from gpsat_b200 import synthetic as syn

E = int(sys.argv[1]) if len(sys.argv) > 1 else 256
n = int(sys.argv[2]) if len(sys.argv) > 2 else 2000
spec = syn.WorkloadSpec("probe", E, n, n, 100, 100, 400e3, 8, (50e3, 50e3, 1.0), (600e3, 600e3, 16.0))
t0 = time.time()
b = syn.make_batch(spec, with_pred=True)
print("gen", time.time() - t0)
eng = Engine(0)
bd = b.to_device("cuda:0")
theta = np.tile(np.array([2.0, 2.0, 3.0, 0.1, 0.01]), (E, 1))
for it in range(3):
    lml, grad, st = eng.lml_grad(bd, theta)
    s = eng.stats()
    print("lml_grad", s["gpu_ms"], "ms", s["sweep_flops"] / s["gpu_ms"] / 1e9, "TF/s", "launches", s["kernel_launches"])
for it in range(2):
    lml, _, st = eng.lml_grad(bd, theta, with_grad=False)
    s = eng.stats()
    print("lml only", s["gpu_ms"], "ms", s["sweep_flops"] / s["gpu_ms"] / 1e9, "TF/s")
for it in range(2):
    p = eng.predict(bd, theta)
    s = eng.stats()
    print("predict", s["gpu_ms"], "ms", s["sweep_flops"] / s["gpu_ms"] / 1e9, "TF/s")
th0, tk, ta, tb, tr = syn.default_fit_setup(spec, E)
t0 = time.time()
res = eng.fit(bd, th0, tk, ta, tb, tr)
torch.cuda.synchronize()
s = eng.stats()
nev = res.n_evals.cpu().numpy()
fl = float(np.sum(nev.astype(float) * b.n.astype(float) ** 3))
print("fit wall", time.time() - t0, "gpu_ms", s["gpu_ms"], "ticks", s["ticks"], "evals", s["evals"], "mean nev", nev.mean(), "max", nev.max(),
      "TF/s", fl / s["gpu_ms"] / 1e9, "status", np.bincount(res.status.cpu().numpy()))
print("theta[0]", res.theta[0].cpu().numpy(), "lml0", float(res.lml[0]))
