"""Writes tests/golden/precision_envelope.npz: the conditioning sweep of tests/envelope.py evaluated in extended
precision (oracle/longdouble_gpr.c), together with the error the float64 NumPy / LAPACK oracle (oracle/gpr.py) makes
on the same inputs - the "attainable by a float64 implementation" level the GPU paths are judged against.

  python -m oracle.make_envelope_golden            (about ten minutes on 8 cores; CPU only)

Test infrastructure: nothing under gpsat_b200/ imports this.
"""
from __future__ import annotations

import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import gpr, longdouble as ldo   # noqa: E402
import envelope as env                       # noqa: E402


def one(kind, X, y, Xs, theta):
    lml, grad, fm, fv = ldo.gpr(kind, X, y, theta, Xs)
    ref = dict(lml=lml, grad=grad, fmean=fm, fvar=fv)
    v, g = gpr.lml_and_grad(kind, X, y, theta)
    p = gpr.predict(kind, X, y, theta, Xs)
    e = env.rel_errors(v, g, p["f*"], p["f*_var"], ref, theta)
    return ref, np.array([e["lml"], e["grad"], e["fmean"], e["fvar"]])


def low_tier():
    """tests/golden/low_tier_envelope.npz: the cases at and below the threshold of the precision rule's low-rho tier."""
    out = {}
    t0 = time.time()
    for c in env.low_tier_cases():
        X, y, Xs, theta = env.make_case(c)
        ref, f64 = one(c["kind"], X, y, Xs, theta)
        for k, v in ref.items():
            out[f"{c['name']}/{k}"] = np.asarray(v)
        out[f"{c['name']}/f64_err"] = f64
        print(f"{c['name']:32s} lml {ref['lml']:.12e}  float64 oracle err: lml {f64[0]:.1e} grad {f64[1]:.1e} "
              f"f* {f64[2]:.1e} f*_var {f64[3]:.1e}   [{time.time() - t0:.0f} s]", flush=True)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "low_tier_envelope.npz"), **out)


def main():
    if "--low-tier" in sys.argv:
        return low_tier()
    out = {}
    t0 = time.time()
    for c in env.cases():
        X, y, Xs, theta = env.make_case(c)
        ref, f64 = one(c["kind"], X, y, Xs, theta)
        for k, v in ref.items():
            out[f"{c['name']}/{k}"] = np.asarray(v)
        out[f"{c['name']}/f64_err"] = f64
        print(f"{c['name']:28s} lml {ref['lml']:.12e}  float64 oracle err: lml {f64[0]:.1e} grad {f64[1]:.1e} "
              f"f* {f64[2]:.1e} f*_var {f64[3]:.1e}   [{time.time() - t0:.0f} s]", flush=True)
    X, y, Xs, theta = env.smoke_case()
    ref, f64 = one("RBF", X, y, Xs, theta)
    for k, v in ref.items():
        out[f"smoke/{k}"] = np.asarray(v)
    out["smoke/f64_err"] = f64
    print("smoke", ref["lml"], f64)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "precision_envelope.npz"), **out)


if __name__ == "__main__":
    main()
