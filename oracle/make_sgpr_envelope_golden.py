"""Writes tests/golden/sgpr_envelope.npz: ELBO, its gradient (central differences of the extended-precision ELBO)
and predictions of the sparse-GPR sweep cases of tests/envelope.py in extended precision (oracle/longdouble_gpr.c
ld_sgpr), with the error the float64 NumPy oracle (oracle/sgpr.py) makes on the same inputs and the condition number
of Kuu + 1e-6 I.  Test infrastructure.    python -m oracle.make_sgpr_envelope_golden   (a few minutes, CPU only)"""
from __future__ import annotations

import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import gpr, sgpr, longdouble as ldo   # noqa: E402
import envelope as env                             # noqa: E402


def main():
    out = {}
    t0 = time.time()
    for c in env.SGPR_CASES:
        X, y, Z, Xs, theta = env.make_sgpr_case(c)
        elbo, fm, fv = ldo.sgpr(c["kind"], X, y, Z, theta, Xs)
        grad = ldo.sgpr_grad_fd(c["kind"], X, y, Z, theta)
        v, g = sgpr.elbo_and_grad(c["kind"], X, y, Z, theta)
        p = sgpr.predict(c["kind"], X, y, Z, theta, Xs)
        Kuu = gpr.kernel_matrix(c["kind"], Z, None, theta[:3], theta[3]) + 1e-6 * np.eye(len(Z))
        cond = float(np.linalg.cond(Kuu))
        f64 = np.array([abs(v - elbo) / abs(elbo), np.max(np.abs(g - grad)) / np.max(np.abs(grad)),
                        np.max(np.abs(p["f*"] - fm)) / np.max(np.abs(fm)), np.max(np.abs(p["f*_var"] - fv)) / theta[3]])
        for k, val in (("elbo", elbo), ("grad", grad), ("fmean", fm), ("fvar", fv), ("f64_err", f64), ("cond_kuu", cond)):
            out[f"{c['name']}/{k}"] = np.asarray(val)
        print(f"{c['name']:34s} elbo {elbo:.12e} cond(Kuu) {cond:.1e}  float64 oracle err: elbo {f64[0]:.1e} grad {f64[1]:.1e} "
              f"f* {f64[2]:.1e} f*_var {f64[3]:.1e}  [{time.time() - t0:.0f} s]", flush=True)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "sgpr_envelope.npz"), **out)


if __name__ == "__main__":
    main()
