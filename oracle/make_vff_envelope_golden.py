"""Writes tests/golden/vff_envelope.npz: ELBO, its gradient (central differences of the extended-precision ELBO) and
predictions of the Fourier-feature sweep cases of tests/envelope.py in extended precision (oracle/vff_ld.py), with the
error the float64 NumPy oracle (oracle/vff.py) makes on the same inputs and the condition number of the factored matrix
P = Kuf Kfu / s2 + kron Kuu.  The reference's own GPR_kron needs TensorFlow / GPflow (absent here) and its test is
commented out (tests/test_localexperts.py:150-190), so this is the pin the ELBO and predict_f restatement has.
Test infrastructure.    python -m oracle.make_vff_envelope_golden   (a few minutes, CPU only)"""
from __future__ import annotations

import os
import sys
import time
from functools import reduce

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import vff, vff_ld   # noqa: E402
import envelope as env           # noqa: E402


def main():
    out = {}
    t0 = time.time()
    for c in env.VFF_CASES:
        X, y, Xs, a, b, ms, theta = env.make_vff_case(c)
        D = X.shape[1]
        pre_ld = vff_ld.Precomputed(c["kind"], X, y, a, b, ms)
        elbo = vff_ld.elbo(pre_ld, theta)
        grad = vff_ld.grad_fd(pre_ld, theta)
        fm, fv = vff_ld.predict(pre_ld, theta, Xs)
        pre = vff.Precomputed(X, y, a, b, ms)
        v, g = vff.elbo_and_grad(c["kind"], pre, theta)
        p = vff.predict(c["kind"], pre, theta, Xs)
        Kuu = [vff.kuu_dim(c["kind"], theta[i], theta[D + i], a[i], b[i], ms[i]) for i in range(D)]
        cond = float(np.linalg.cond(pre.KufKfu / theta[2 * D] + reduce(np.kron, Kuu)))
        f64 = np.array([abs(v - elbo) / abs(elbo), np.max(np.abs(g - grad)) / np.max(np.abs(grad)),
                        np.max(np.abs(p["f*"] - fm)) / np.max(np.abs(fm)),
                        np.max(np.abs(p["f*_var"] - fv)) / np.prod(theta[D:2 * D])])
        for k, val in (("elbo", elbo), ("grad", grad), ("fmean", fm), ("fvar", fv), ("f64_err", f64), ("cond_p", cond)):
            out[f"{c['name']}/{k}"] = np.asarray(val)
        print(f"{c['name']:34s} elbo {elbo:.12e} cond(P) {cond:.1e}  float64 oracle err: elbo {f64[0]:.1e} grad {f64[1]:.1e} "
              f"f* {f64[2]:.1e} f*_var {f64[3]:.1e}  [{time.time() - t0:.0f} s]", flush=True)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "vff_envelope.npz"), **out)


if __name__ == "__main__":
    main()
