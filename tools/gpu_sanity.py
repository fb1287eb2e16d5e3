"""Quick GPU sanity run against the oracle (developer tool; the real checks live in tests/)."""
import sys, time
import numpy as np
sys.path.insert(0, ".")
from gpsat_b200.engine import Engine, RaggedBatch
from oracle import gpr

rng = np.random.default_rng(0)
eng = Engine(0)

def make(n, D=3, m=0):
    X = rng.uniform(-4, 4, (n, D))
    K = gpr.kernel_matrix("Matern32", X, None, [2.0, 1.5, 3.0][:D], 0.7)
    y = np.linalg.cholesky(K + 1e-8 * np.eye(n)) @ rng.standard_normal(n) + 0.2 * rng.standard_normal(n)
    Xs = rng.uniform(-4, 4, (m, D))
    return X, y, Xs

for n in (5, 100, 127, 128, 129, 300, 700):
    X, y, _ = make(n)
    th = np.array([1.3, 0.7, 2.0, 0.8, 0.1])
    L, U = eng.factor_debug("Matern32", X, y, th)
    Ks = gpr.kernel_matrix("Matern32", X, None, th[:3], th[3]) + th[4] * np.eye(n)
    Lr = np.linalg.cholesky(Ks)
    a = np.linalg.solve(Lr, y)
    eL = np.max(np.abs(L[:n, :n] - Lr)); ea = np.max(np.abs(L[n, :n] - a))
    Laug = np.zeros((n + 1, n + 1)); Laug[:n, :n] = Lr; Laug[n, :n] = a; Laug[n, n] = 1
    Ur = np.linalg.inv(Laug).T
    eU = np.max(np.abs(U - Ur)) / np.max(np.abs(Ur))
    print(f"n={n}: |L-Lref|={eL:.2e} |a-aref|={ea:.2e} relU={eU:.2e}")

for kind in gpr.KERNELS:
    Xl, yl, Xsl = [], [], []
    for n in (40, 130, 257, 600):
        X, y, Xs = make(n, m=77 + n)
        Xl.append(X); yl.append(y); Xsl.append(Xs)
    b = RaggedBatch.from_lists(kind, Xl, yl, Xsl)
    th = np.tile(np.array([1.3, 0.7, 2.0, 0.8, 0.1]), (b.E, 1))
    lml, grad, st = eng.lml_grad(b, th)
    pr = eng.predict(b, th)
    for e in range(b.E):
        v, g = gpr.lml_and_grad(kind, Xl[e], yl[e], th[e])
        p = gpr.predict(kind, Xl[e], yl[e], th[e], Xsl[e])
        sl = slice(b.off_m[e], b.off_m[e + 1])
        print(kind, e, "lml rel", abs(lml[e] - v) / abs(v), "grad rel", np.max(np.abs(grad[e] - g) / np.abs(g)),
              "f* ", np.max(np.abs(pr["f*"][sl] - p["f*"])), "fvar rel", np.max(np.abs(pr["f*_var"][sl] - p["f*_var"]) / p["f*_var"]),
              "lml2", abs(pr["lml"][e] - v) / abs(v), st[e])

# fit
Xl, yl = [], []
for n in (60, 150, 300, 420, 90, 200):
    X, y, _ = make(n); Xl.append(X); yl.append(y)
b = RaggedBatch.from_lists("Matern32", Xl, yl)
E = b.E
tr = gpr.Transform.gpflow_default(3); tr.set_sigmoid([0, 1, 2], [1e-8] * 3, [12, 12, 9.0])
th0 = np.ones((E, 5))
t0 = time.time()
res = eng.fit(b, th0, np.tile(tr.kind, (E, 1)), np.tile(tr.a, (E, 1)), np.tile(tr.b, (E, 1)), np.ones((E, 5), np.uint8), max_slots=4)
print("fit time", time.time() - t0, eng.stats())
for e in range(E):
    th, l, ok, nev = gpr.fit("Matern32", Xl[e], yl[e], th0[e], tr)
    print(e, "theta rel", np.max(np.abs(res.theta[e] - th) / th), "lml", abs(res.lml[e] - l), "nev", res.n_evals[e], nev, "status", res.status[e], ok)
