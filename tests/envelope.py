"""Cases of the conditioning sweep: which noise / signal-variance / size / duplicate-row
regimes the default (int8-sliced) arithmetic, the FP64 mma.sync path and a LAPACK float64 implementation can hold
to the 1e-8 parity bar, measured against the extended-precision oracle (oracle/longdouble_gpr.c).

Shared by oracle/make_envelope_golden.py (writes tests/golden/precision_envelope.npz, CPU only),
tests/test_gpu_parity.py::test_i8_precision_envelope and tools/precision_envelope.py.
Everything is seeded with numpy.random.default_rng, so the GPU box regenerates the same inputs bit for bit.
"""
from __future__ import annotations

import numpy as np

# GPflow's floor of the Gaussian likelihood variance is 1e-6 (gpflow_models.py:404-411); the sweep sits just above
NOISES = (1e-2, 1e-4, 1e-6 + 1e-9)
VARS = (0.05, 1.0, 10.0)
SIZES = (1000, 2000, 4000)
KINDS = ("Matern32", "RBF")
M_PRED = 32
LS = (3.0, 3.5, 4.0)


def cases(sizes=SIZES):
    out = []
    for kind in KINDS:
        for n in sizes:
            for noise in NOISES:
                for var in VARS:
                    out.append(dict(name=f"{kind}-n{n}-s{noise:.0e}-v{var:g}", kind=kind, n=n, noise=noise, var=var,
                                    dup=0.10, D=3, seed=len(out)))
    return out


def _rff(rng, kind, X, ls, var, nfeat=384):
    z = rng.standard_normal((nfeat, X.shape[1]))
    if kind == "RBF":
        omega = z / np.asarray(ls)[None, :]
    else:
        nu = {"Matern12": 0.5, "Matern32": 1.5, "Matern52": 2.5}[kind]
        omega = z * np.sqrt(2 * nu / rng.chisquare(2 * nu, size=(nfeat, 1))) / np.asarray(ls)[None, :]
    phase = rng.uniform(0, 2 * np.pi, nfeat)
    w = rng.standard_normal(nfeat)
    return np.sqrt(2.0 * var / nfeat) * (np.cos(X @ omega.T + phase) @ w)


def make_case(c):
    """(X [n, 3], y [n], Xs [M_PRED, 3], theta [5]) of one sweep case: c2-like geometry in scaled units (x, y in
    [-6, 6], t whole days in [-4, 4]), observations drawn from the model at theta (so y^T Ks^-1 y ~ n), and `dup`
    of the rows exact coordinate duplicates of other rows (repeat passes over the same bin)."""
    rng = np.random.default_rng([2024, c["seed"]])
    n = c["n"]
    X = np.stack([rng.uniform(-6, 6, n), rng.uniform(-6, 6, n), rng.integers(-4, 5, n).astype(float)], axis=1)
    ndup = int(c["dup"] * n)
    if ndup:
        dst = rng.choice(n, ndup, replace=False)
        src = rng.choice(np.setdiff1d(np.arange(n), dst), ndup, replace=True)
        X[dst] = X[src]
    theta = np.array([*LS, c["var"], c["noise"]])
    y = _rff(rng, c["kind"], X, LS, c["var"]) + np.sqrt(c["noise"]) * rng.standard_normal(n)
    Xs = np.stack([rng.uniform(-5, 5, M_PRED), rng.uniform(-5, 5, M_PRED), np.zeros(M_PRED)], axis=1)
    return X, y, Xs, theta


def smoke_case():
    """The reference's own ill-conditioned smoke (test_numerical_stability.py:15-75): a 20 x 20 grid in unscaled
    metre coordinates (+-1000 km), RBF with lengthscales 50 000 m, noise variance 0.1, kernel variance 1."""
    xs = np.linspace(-1000000, 1000000, 20)
    xm, ym = np.meshgrid(xs, xs)
    rng = np.random.default_rng(42)
    th = 0.5 + 0.3 * np.sin(xm / 500000) * np.cos(ym / 500000) + rng.normal(0, 0.1, xm.shape)
    X = np.stack([xm.ravel() + rng.normal(0, 1e-6, xm.size), ym.ravel() + rng.normal(0, 1e-6, xm.size)], axis=1)
    theta = np.array([50000.0, 50000.0, 1.0, 0.1])
    Xs = np.stack([rng.uniform(-9e5, 9e5, M_PRED), rng.uniform(-9e5, 9e5, M_PRED)], axis=1)
    return X, th.ravel(), Xs, theta


def rel_errors(lml, grad, fmean, fvar, ref, theta):
    """Error measures of one result against the extended-precision reference `ref` (dict lml, grad, fmean, fvar):
    relative for the LML, relative to the largest component for the gradient, relative to max |f*| for the mean,
    relative to the signal variance for f*_var (the prior variance the posterior variance is subtracted from)."""
    D = len(theta) - 2
    return dict(lml=abs(lml - ref["lml"]) / abs(ref["lml"]),
                grad=float(np.max(np.abs(grad - ref["grad"])) / np.max(np.abs(ref["grad"]))),
                fmean=float(np.max(np.abs(fmean - ref["fmean"])) / max(np.max(np.abs(ref["fmean"])), 1e-300)),
                fvar=float(np.max(np.abs(fvar - ref["fvar"])) / theta[D]))


# ---- sparse GPR: what a float64 implementation attains with the 1e-6 jitter on Kuu -----------
SGPR_CASES = (
    dict(name="c5-shape-Matern32-n20000-M500", kind="Matern32", n=20000, M=500, theta=(4.0, 4.0, 5.0, 0.05, 0.005), c5=True),
    dict(name="RBF-n5000-M300", kind="RBF", n=5000, M=300, theta=(2.0, 2.5, 3.0, 1.0, 1e-2), c5=False),
    dict(name="Matern32-n2000-M200-rho1e4", kind="Matern32", n=2000, M=200, theta=(3.0, 3.5, 4.0, 1.0, 1e-4), c5=False),
    dict(name="Matern52-n300-Z=X", kind="Matern52", n=300, M=300, theta=(2.0, 2.0, 3.0, 0.5, 0.05), c5=False),
)


def make_sgpr_case(c):
    """(X, y, Z, Xs, theta) of one sparse sweep case; the c5 case is expert 0 of the bench's c5 workload."""
    theta = np.asarray(c["theta"], dtype=np.float64)
    if c["c5"]:
        from gpsat_b200 import synthetic as syn
        X, y, Xs = syn.make_expert(syn.WORKLOADS["c5"], 0, 0)
        return X, y, syn.inducing_points(X, 0, 0, c["M"]), Xs[:M_PRED], theta
    rng = np.random.default_rng([77, c["n"], c["M"]])
    n = c["n"]
    X = np.stack([rng.uniform(-6, 6, n), rng.uniform(-6, 6, n), rng.integers(-4, 5, n).astype(float)], axis=1)
    y = _rff(rng, c["kind"], X, theta[:3], theta[3]) + np.sqrt(theta[4]) * rng.standard_normal(n)
    Z = X[rng.permutation(n)[:c["M"]]].copy()
    Xs = np.stack([rng.uniform(-5, 5, M_PRED), rng.uniform(-5, 5, M_PRED), np.zeros(M_PRED)], axis=1)
    return X, y, Z, Xs, theta


# ---- variational Fourier features: the ELBO / predictions of GPR_kron in extended precision --
VFF_CASES = (
    dict(name="Matern32-1D-n2000-f40", kind="Matern32", n=2000, nfreq=(40,), ell=(0.8,), var=(1.2,), noise=0.05),
    dict(name="Matern12-1D-n500-f64", kind="Matern12", n=500, nfreq=(64,), ell=(1.5,), var=(0.7,), noise=0.1),
    dict(name="Matern52-2D-n1500-f12x9", kind="Matern52", n=1500, nfreq=(12, 9), ell=(0.9, 1.3), var=(1.1, 0.8), noise=0.02),
    dict(name="Matern32-2D-n3000-f20x20", kind="Matern32", n=3000, nfreq=(20, 20), ell=(0.7, 0.6), var=(1.0, 1.0), noise=0.01),
    dict(name="Matern32-2D-n1000-f10x10-s1e-4", kind="Matern32", n=1000, nfreq=(10, 10), ell=(1.0, 1.2), var=(1.0, 2.0),
         noise=1e-4),
    dict(name="Matern32-3D-n1000-f5x4x3", kind="Matern32", n=1000, nfreq=(5, 4, 3), ell=(1.0, 1.4, 2.0), var=(1.0, 0.9, 1.1),
         noise=0.05),
)


def make_vff_case(c):
    """(X, y, Xs, a, b, ms, theta) of one Fourier-feature sweep case; two of the M_PRED prediction points lie outside
    the domain box for the kinds whose edge features the reference implements (vff.py:472-497)."""
    D = len(c["nfreq"])
    rng = np.random.default_rng([911, c["n"], D, len(c["kind"])])
    X = rng.uniform(-2.0, 2.0, (c["n"], D))
    y = np.sin(2 * X[:, 0]) + (0.5 * np.cos(X[:, 1]) if D > 1 else 0.0) + np.sqrt(c["noise"]) * rng.standard_normal(c["n"])
    a = X.min(axis=0) - 0.4 * rng.uniform(0.5, 1.5, D)
    b = X.max(axis=0) + 0.4 * rng.uniform(0.5, 1.5, D)
    Xs = rng.uniform(-1.9, 1.9, (M_PRED, D))
    if c["kind"] != "Matern52":
        Xs[0, 0] = a[0] - 0.3
        Xs[1, D - 1] = b[D - 1] + 0.25
    ms = [np.arange(k) for k in c["nfreq"]]
    theta = np.array([*c["ell"], *c["var"], c["noise"]])
    return X, y, Xs, a, b, ms, theta


# ---- low-rho tier of the precision rule: rho up to the tier's threshold 32, and one case outside -------
LOW_TIER = ((0.1, 1.0), (0.5, 5.0), (0.033, 1.0), (0.0033, 0.1), (0.02, 1.0))      # (noise, var): rho 11, 11, 31.3, 31.3, 51


LOW_TIER_KINDS = ("Matern32", "RBF", "Matern12", "Matern52")


def low_tier_cases():
    out = []
    for kind in LOW_TIER_KINDS:
        for n in (1000, 2000) if kind in KINDS else (2000,):
            for noise, var in LOW_TIER:
                out.append(dict(name=f"low-{kind}-n{n}-s{noise:g}-v{var:g}", kind=kind, n=n, noise=noise, var=var, dup=0.10, D=3,
                                seed=1000 + len(out)))
    for noise, var in ((0.1, 1.0), (0.033, 1.0)):          # the largest c3 experts: 32 tile columns
        out.append(dict(name=f"low-Matern32-n4000-s{noise:g}-v{var:g}", kind="Matern32", n=4000, noise=noise, var=var, dup=0.10,
                        D=3, seed=1000 + len(out)))
    return out
