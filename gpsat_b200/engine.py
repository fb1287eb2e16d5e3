"""NumPy / torch facing wrapper around the C ABI (one :class:`Engine` per GPU rank).

``RaggedBatch`` is the on-wire format between the driver and the engine: experts packed back to
back, described by ``off_n`` / ``off_m`` (see ``include/gpsat_b200.h``).  Host arrays go through
the library's own host<->device copies (the end-to-end path); torch CUDA tensors are passed as
device pointers (inputs resident in HBM).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import Optional, Sequence

import numpy as np

from . import _lib

SCIPY_FTOL = 2.220446049250313e-09


def _is_torch(x) -> bool:
    return type(x).__module__.startswith("torch")


def _ptr(x):
    if x is None:
        return None
    if _is_torch(x):
        return C.c_void_p(x.data_ptr())
    return C.c_void_p(x.ctypes.data)


def _f64(x, shape=None):
    a = np.ascontiguousarray(x, dtype=np.float64)
    if shape is not None:
        a = a.reshape(shape)
    return a


@dataclass
class RaggedBatch:
    """Packed experts.  ``X`` [N, D], ``y`` [N], ``Xs`` [M, D] are either all NumPy (host) or all
    torch CUDA float64 tensors (device); ``off_n`` / ``off_m`` are host int64 arrays."""
    kernel: str
    D: int
    off_n: np.ndarray
    X: object
    y: object
    off_m: Optional[np.ndarray] = None
    Xs: object = None
    _keep: list = field(default_factory=list, repr=False)

    @property
    def E(self) -> int:
        return len(self.off_n) - 1

    @property
    def on_device(self) -> bool:
        return _is_torch(self.X)

    @property
    def n(self) -> np.ndarray:
        return np.diff(self.off_n)

    @property
    def m(self) -> np.ndarray:
        return np.diff(self.off_m) if self.off_m is not None else np.zeros(self.E, np.int64)

    @classmethod
    def from_lists(cls, kernel: str, Xs_obs: Sequence[np.ndarray], ys: Sequence[np.ndarray],
                   Xs_pred: Optional[Sequence[np.ndarray]] = None) -> "RaggedBatch":
        assert len(Xs_obs) == len(ys)
        E = len(Xs_obs)
        D = int(np.atleast_2d(Xs_obs[0]).shape[1]) if E else 1
        off_n = np.zeros(E + 1, np.int64)
        off_n[1:] = np.cumsum([len(a) for a in Xs_obs])
        X = _f64(np.concatenate([_f64(a, (-1, D)) for a in Xs_obs], axis=0)) if E else np.zeros((0, D))
        y = _f64(np.concatenate([_f64(a).ravel() for a in ys])) if E else np.zeros(0)
        assert len(y) == off_n[-1], "obs / coords length mismatch"
        off_m, XS = None, None
        if Xs_pred is not None:
            assert len(Xs_pred) == E
            off_m = np.zeros(E + 1, np.int64)
            off_m[1:] = np.cumsum([len(a) for a in Xs_pred])
            XS = _f64(np.concatenate([_f64(a, (-1, D)) for a in Xs_pred], axis=0)) if E else np.zeros((0, D))
        return cls(kernel=kernel, D=D, off_n=off_n, X=X, y=y, off_m=off_m, Xs=XS)

    def to_device(self, device) -> "RaggedBatch":
        import torch
        t = lambda a: None if a is None else torch.as_tensor(a, dtype=torch.float64, device=device).contiguous()
        return RaggedBatch(self.kernel, self.D, self.off_n, t(self.X), t(self.y), self.off_m, t(self.Xs))

    def select(self, idx: Sequence[int]) -> "RaggedBatch":
        """Host-side sub-batch (used to shard experts across ranks)."""
        assert not self.on_device
        idx = list(idx)
        Xo = [self.X[self.off_n[e]:self.off_n[e + 1]] for e in idx]
        yo = [self.y[self.off_n[e]:self.off_n[e + 1]] for e in idx]
        Xp = None
        if self.off_m is not None:
            Xp = [self.Xs[self.off_m[e]:self.off_m[e + 1]] for e in idx]
        if not idx:
            b = RaggedBatch(self.kernel, self.D, np.zeros(1, np.int64), np.zeros((0, self.D)), np.zeros(0))
            if self.off_m is not None:
                b.off_m, b.Xs = np.zeros(1, np.int64), np.zeros((0, self.D))
            return b
        return RaggedBatch.from_lists(self.kernel, Xo, yo, Xp)

    def _c(self, with_pred: bool) -> _lib.Batch:
        b = _lib.Batch()
        b.kernel = _lib.KERNEL_IDS[self.kernel]
        b.D = self.D
        b.E = self.E
        self.off_n = np.ascontiguousarray(self.off_n, dtype=np.int64)
        b.off_n = self.off_n.ctypes.data
        b.X = _ptr(self.X).value if self.X is not None else None
        b.y = _ptr(self.y).value if self.y is not None else None
        if with_pred:
            assert self.off_m is not None, "batch has no prediction locations"
            self.off_m = np.ascontiguousarray(self.off_m, dtype=np.int64)
            b.off_m = self.off_m.ctypes.data
            b.Xs = _ptr(self.Xs).value if self.Xs is not None else None
        return b


@dataclass
class FitResult:
    theta: object
    objective: object      # minimised value -(lml + log prior): gpflow's training_loss at the optimum
    lml: object
    n_evals: object
    n_iters: object
    status: object

    @property
    def success(self):
        return self.status == _lib.STATUS_OK


class Engine:
    """One engine handle = one CUDA device + one stream.  Not thread-safe."""

    def __init__(self, device: int = 0):
        self._lib = _lib.load()
        h = C.c_void_p()
        _lib.check(self._lib.gpsat_engine_create(int(device), C.byref(h)))
        self._h = h
        self.device = int(device)

    def close(self):
        if getattr(self, "_h", None):
            self._lib.gpsat_engine_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_profile(self, on: bool):
        """Bracket every kernel launch with CUDA events; per-kernel times appear in :meth:`stats`."""
        _lib.check(self._lib.gpsat_engine_set_profile(self._h, int(bool(on))))

    def set_lauum_path(self, path: str):
        """``"i8"`` (default, one CTA per tile) / ``"i8q"`` (persistent kernel; bitwise the same results): K^-1 tile
        products of the gradient by error-free int8 slicing on the tcgen05 tensor path; ``"dmma"``: FP64 ``mma.sync``.  The int8 slices are taken against a-priori
        bounds, so they match FP64 only while (var + noise) / noise stays moderate: see
        :meth:`set_precision_rule` (on by default) and ``profiles/r02_precision_envelope.md``."""
        _lib.check(self._lib.gpsat_engine_set_lauum_path(self._h, {"dmma": 0, "i8": 1, "i8q": 2}[path]))

    def set_step_path(self, path: str):
        """'i8' (int8 tensor path for the potrf main loop) or 'dmma' (FP64 mma.sync)."""
        _lib.check(self._lib.gpsat_engine_set_step_path(self._h, {"dmma": 0, "i8": 1}[path]))

    def i8_phases(self, reset: bool = True):
        """Cycles per phase of the int8-path gradient kernel (development builds with -DGPSAT_I8_PHASES)."""
        out = (C.c_uint64 * 8)()
        _lib.check(self._lib.gpsat_engine_i8_phases(self._h, out, int(reset)))
        return [int(v) for v in out]

    def step_phases(self, reset: bool = True):
        """Cycles per phase of step_kernel (development builds with -DGPSAT_STEP_PHASES): potrf tiles [0, 16),
        trtri tiles [16, 32)."""
        out = (C.c_uint64 * 32)()
        _lib.check(self._lib.gpsat_engine_step_phases(self._h, out, int(reset)))
        return [int(v) for v in out]

    def i8_timeouts(self) -> int:
        """Number of calls that failed (raised) because a bounded barrier wait of the int8 pipeline gave up."""
        out = C.c_int(0)
        _lib.check(self._lib.gpsat_engine_i8_timeouts(self._h, C.byref(out)))
        return int(out.value)

    def debug_set_i8_spin_limit(self, polls: int):
        """Test hook: polls after which a barrier wait of the int8 pipeline gives up (default 2**22)."""
        _lib.check(self._lib.gpsat_engine_debug_set_i8_spin_limit(self._h, int(polls)))

    def set_precision_rule(self, rho_max=True):
        """Per expert and evaluation: int8-sliced k-sums only while ``(var + noise) / noise <= rho_max``,
        FP64 ``mma.sync`` above.  ``True`` = the library default, ``False`` / 0 = rule off."""
        if rho_max is True:
            rho_max = float(self._lib.gpsat_engine_precision_rule_default())
        _lib.check(self._lib.gpsat_engine_set_precision_rule(self._h, float(rho_max or 0.0)))

    def set_precision_low_tier(self, rho_low=True):
        """Low-rho tier of the precision rule: while ``(var + noise) / noise <= rho_low`` the int8-sliced k-sums drop
        their last slice-pair group.  ``True`` = the library default, ``False`` / 0 = tier off."""
        if rho_low is True:
            rho_low = float(self._lib.gpsat_engine_precision_low_tier_default())
        _lib.check(self._lib.gpsat_engine_set_precision_low_tier(self._h, float(rho_low or 0.0)))

    @property
    def stream_ptr(self) -> int:
        """cudaStream_t of the engine (wrap with torch.cuda.ExternalStream to record events on it)."""
        return int(self._lib.gpsat_engine_stream(self._h))

    def stats(self) -> dict:
        s = _lib.Stats()
        _lib.check(self._lib.gpsat_engine_stats(self._h, C.byref(s)))
        return {k: getattr(s, k) for k, _ in s._fields_}

    # -- allocation helper: outputs live where the inputs live --------------------------------
    def _empty(self, batch: RaggedBatch, shape, dtype):
        if batch.on_device:
            import torch
            td = {np.float64: torch.float64, np.int32: torch.int32}[dtype]
            return torch.empty(shape, dtype=td, device=batch.X.device)
        return np.empty(shape, dtype=dtype)

    def _in(self, batch: RaggedBatch, a, dtype, shape):
        if batch.on_device:
            import torch
            td = {np.float64: torch.float64, np.int32: torch.int32, np.uint8: torch.uint8}[dtype]
            if _is_torch(a):
                return a.to(device=batch.X.device, dtype=td).reshape(shape).contiguous()
            return torch.as_tensor(np.ascontiguousarray(a, dtype=dtype).reshape(shape), device=batch.X.device)
        return np.ascontiguousarray(a, dtype=dtype).reshape(shape)

    def _pre(self, batch: RaggedBatch):
        """Device inputs were produced on torch's current stream; the engine runs on its own
        non-blocking stream, so order the two before handing pointers over."""
        if batch.on_device:
            import torch
            torch.cuda.current_stream(batch.X.device).synchronize()

    def lml_grad(self, batch: RaggedBatch, theta, with_grad: bool = True):
        E, P = batch.E, batch.D + 2
        mem = _lib.MEM_DEVICE if batch.on_device else _lib.MEM_HOST
        th = self._in(batch, theta, np.float64, (E, P))
        lml = self._empty(batch, (E,), np.float64)
        grad = self._empty(batch, (E, P), np.float64) if with_grad else None
        status = self._empty(batch, (E,), np.int32)
        cb = batch._c(False)
        self._pre(batch)
        _lib.check(self._lib.gpsat_gpr_lml_grad(self._h, C.byref(cb), mem, _ptr(th), _ptr(lml), _ptr(grad), _ptr(status)))
        return lml, grad, status

    def fit(self, batch: RaggedBatch, theta0, tkind, ta, tb, trainable, *, max_iter: int = 10_000,
            pgtol: float = 1e-5, ftol: float = SCIPY_FTOL, maxfun: int = 15_000, maxls: int = 20,
            max_slots: int = 0, priors=None, inducing=None, vff=None) -> FitResult:
        """Batched L-BFGS fit.  ``priors``: optional ``(kind[E,P], a[E,P], b[E,P])`` (GPSAT_PRIOR_*).
        ``inducing``: one [M_e, D] array per expert switches to the sparse model (``gpsat_sgpr_fit``:
        minimises -elbo with the inducing points fixed; ``FitResult.lml`` then holds the ELBO).
        ``vff``: ``(a[E,D], b[E,D], nfreq[E,D])`` switches to variational Fourier features
        (``gpsat_vff_fit``; theta has 2 D + 1 columns: lengthscales, per-dimension variances, noise)."""
        E = batch.E
        P = 2 * batch.D + 1 if vff is not None else batch.D + 2
        mem = _lib.MEM_DEVICE if batch.on_device else _lib.MEM_HOST
        th = self._in(batch, theta0, np.float64, (E, P))
        th = th.clone() if _is_torch(th) else th.copy()
        tk = self._in(batch, tkind, np.int32, (E, P))
        a = self._in(batch, ta, np.float64, (E, P))
        b = self._in(batch, tb, np.float64, (E, P))
        tr = self._in(batch, trainable, np.uint8, (E, P))
        lml = self._empty(batch, (E,), np.float64)
        obj = self._empty(batch, (E,), np.float64)
        pri, keep = None, None
        if priors is not None:
            keep = (self._in(batch, priors[0], np.int32, (E, P)), self._in(batch, priors[1], np.float64, (E, P)),
                    self._in(batch, priors[2], np.float64, (E, P)))
            pri = _lib.Priors(_ptr(keep[0]).value, _ptr(keep[1]).value, _ptr(keep[2]).value)
        nev = self._empty(batch, (E,), np.int32)
        nit = self._empty(batch, (E,), np.int32)
        status = self._empty(batch, (E,), np.int32)
        opt = _lib.FitOptions(pgtol, ftol, int(max_iter), int(maxfun), int(maxls), int(max_slots))
        cb = batch._c(False)
        self._pre(batch)
        tail = (_ptr(th), _ptr(tk), _ptr(a), _ptr(b), _ptr(tr), C.byref(opt),
                C.byref(pri) if pri is not None else None, _ptr(obj), _ptr(lml), _ptr(nev), _ptr(nit), _ptr(status))
        if vff is not None:
            dom, keep_dom = self._vff_domain(batch, vff)
            _lib.check(self._lib.gpsat_vff_fit(self._h, C.byref(cb), C.byref(dom), mem, *tail))
        elif inducing is None:
            _lib.check(self._lib.gpsat_gpr_fit(self._h, C.byref(cb), mem, *tail))
        else:
            off_z, Zp = self._inducing(batch, inducing)
            _lib.check(self._lib.gpsat_sgpr_fit(self._h, C.byref(cb), _ptr(off_z), _ptr(Zp), mem, *tail))
        return FitResult(th, obj, lml, nev, nit, status)

    def predict(self, batch: RaggedBatch, theta):
        E, P = batch.E, batch.D + 2
        mem = _lib.MEM_DEVICE if batch.on_device else _lib.MEM_HOST
        M = int(batch.off_m[-1])
        th = self._in(batch, theta, np.float64, (E, P))
        fm = self._empty(batch, (M,), np.float64)
        fv = self._empty(batch, (M,), np.float64)
        yv = self._empty(batch, (M,), np.float64)
        lml = self._empty(batch, (E,), np.float64)
        status = self._empty(batch, (E,), np.int32)
        cb = batch._c(True)
        self._pre(batch)
        _lib.check(self._lib.gpsat_gpr_predict(self._h, C.byref(cb), mem, _ptr(th), _ptr(fm), _ptr(fv), _ptr(yv),
                                               _ptr(lml), _ptr(status)))
        return {"f*": fm, "f*_var": fv, "y_var": yv, "lml": lml, "status": status}

    def predict_full_cov(self, batch: RaggedBatch, theta):
        """:meth:`predict` plus ``"f*_cov"``: a list with one dense [m_e, m_e] block per expert."""
        E, P = batch.E, batch.D + 2
        mem = _lib.MEM_DEVICE if batch.on_device else _lib.MEM_HOST
        M = int(batch.off_m[-1])
        msz = batch.m.astype(np.int64)
        off_c = np.concatenate([[0], np.cumsum(msz * msz)])
        th = self._in(batch, theta, np.float64, (E, P))
        fm, fv, yv = (self._empty(batch, (M,), np.float64) for _ in range(3))
        cov = self._empty(batch, (int(off_c[-1]),), np.float64)
        lml = self._empty(batch, (E,), np.float64)
        status = self._empty(batch, (E,), np.int32)
        cb = batch._c(True)
        self._pre(batch)
        _lib.check(self._lib.gpsat_gpr_predict_cov(self._h, C.byref(cb), mem, _ptr(th), _ptr(fm), _ptr(fv), _ptr(yv),
                                                   _ptr(cov), _ptr(lml), _ptr(status)))
        blocks = [cov[int(off_c[e]):int(off_c[e + 1])].reshape(int(msz[e]), int(msz[e])) for e in range(E)]
        return {"f*": fm, "f*_var": fv, "y_var": yv, "f*_cov": blocks, "lml": lml, "status": status}

    # ---- sparse GPR (GPflowSGPRModel, gpflow_models.py:666-901) ------------------------------------
    def _inducing(self, batch: RaggedBatch, Z):
        """Z: one [M_e, D] array per expert -> (host off_z, packed Z on the batch's memory)."""
        Zl = [np.asarray(z.cpu() if _is_torch(z) else z, dtype=np.float64).reshape(-1, batch.D) for z in Z]
        if len(Zl) != batch.E:
            raise ValueError(f"need one inducing-point array per expert: got {len(Zl)} for {batch.E} experts")
        off_z = np.zeros(batch.E + 1, dtype=np.int64)
        off_z[1:] = np.cumsum([len(z) for z in Zl])
        Zp = np.ascontiguousarray(np.concatenate(Zl, axis=0)) if Zl else np.zeros((1, batch.D))
        return off_z, self._in(batch, Zp, np.float64, Zp.shape)

    def sgpr_elbo(self, batch: RaggedBatch, Z, theta, with_grad: bool = False, with_z: bool = False):
        """ELBO of ``gpflow.models.SGPR`` per expert at fixed theta (and, ``with_grad``, its gradient
        w.r.t. theta with Z fixed).  Returns (elbo, grad or None, status); ``with_z`` (implies ``with_grad``) appends
        the gradient w.r.t. the inducing points, packed [sum M_e, D] like Z (``gpsat_sgpr_elbo_gradz``)."""
        E, P = batch.E, batch.D + 2
        mem = _lib.MEM_DEVICE if batch.on_device else _lib.MEM_HOST
        off_z, Zp = self._inducing(batch, Z)
        th = self._in(batch, theta, np.float64, (E, P))
        elbo = self._empty(batch, (E,), np.float64)
        grad = self._empty(batch, (E, P), np.float64) if (with_grad or with_z) else None
        status = self._empty(batch, (E,), np.int32)
        cb = batch._c(False)
        self._pre(batch)
        if with_z:
            gz = self._empty(batch, (max(int(off_z[-1]), 1), batch.D), np.float64)
            _lib.check(self._lib.gpsat_sgpr_elbo_gradz(self._h, C.byref(cb), _ptr(off_z), _ptr(Zp), mem, _ptr(th),
                                                       _ptr(elbo), _ptr(grad), _ptr(gz), _ptr(status)))
            return elbo, grad, gz[:int(off_z[-1])], status
        _lib.check(self._lib.gpsat_sgpr_elbo(self._h, C.byref(cb), _ptr(off_z), _ptr(Zp), mem, _ptr(th), _ptr(elbo),
                                             _ptr(grad), _ptr(status)))
        return elbo, grad, status

    def sgpr_predict(self, batch: RaggedBatch, Z, theta, full_cov: bool = False):
        """``SGPR.predict_f`` / ``predict_y`` per expert at fixed theta; ``full_cov`` adds ``"f*_cov"``: one dense
        [m_e, m_e] block per expert (``gpsat_sgpr_predict_cov``)."""
        E, P = batch.E, batch.D + 2
        mem = _lib.MEM_DEVICE if batch.on_device else _lib.MEM_HOST
        M = int(batch.off_m[-1])
        off_z, Zp = self._inducing(batch, Z)
        th = self._in(batch, theta, np.float64, (E, P))
        fm, fv, yv = (self._empty(batch, (M,), np.float64) for _ in range(3))
        elbo = self._empty(batch, (E,), np.float64)
        status = self._empty(batch, (E,), np.int32)
        cb = batch._c(True)
        self._pre(batch)
        if full_cov:
            msz = batch.m.astype(np.int64)
            off_c = np.concatenate([[0], np.cumsum(msz * msz)])
            cov = self._empty(batch, (int(off_c[-1]),), np.float64)
            _lib.check(self._lib.gpsat_sgpr_predict_cov(self._h, C.byref(cb), _ptr(off_z), _ptr(Zp), mem, _ptr(th), _ptr(fm),
                                                        _ptr(fv), _ptr(yv), _ptr(cov), _ptr(elbo), _ptr(status)))
            blocks = [cov[int(off_c[e]):int(off_c[e + 1])].reshape(int(msz[e]), int(msz[e])) for e in range(E)]
            return {"f*": fm, "f*_var": fv, "y_var": yv, "f*_cov": blocks, "elbo": elbo, "status": status}
        _lib.check(self._lib.gpsat_sgpr_predict(self._h, C.byref(cb), _ptr(off_z), _ptr(Zp), mem, _ptr(th), _ptr(fm),
                                                _ptr(fv), _ptr(yv), _ptr(elbo), _ptr(status)))
        return {"f*": fm, "f*_var": fv, "y_var": yv, "elbo": elbo, "status": status}

    # ---- variational Fourier features (GPflowVFFModel, vff_model.py:14-264) ----------------------------
    def _vff_domain(self, batch: RaggedBatch, vff):
        E, D = batch.E, batch.D
        a = np.ascontiguousarray(vff[0], dtype=np.float64).reshape(E, D)
        b = np.ascontiguousarray(vff[1], dtype=np.float64).reshape(E, D)
        nf = np.ascontiguousarray(vff[2], dtype=np.int32).reshape(E, D)
        return _lib.VffDomain(a.ctypes.data, b.ctypes.data, nf.ctypes.data), (a, b, nf)

    def vff_elbo(self, batch: RaggedBatch, vff, theta, with_grad: bool = False):
        """``GPR_kron.elbo`` per expert at fixed theta ([E, 2 D + 1]); returns (elbo, grad or None, status)."""
        E, P = batch.E, 2 * batch.D + 1
        mem = _lib.MEM_DEVICE if batch.on_device else _lib.MEM_HOST
        dom, keep = self._vff_domain(batch, vff)
        th = self._in(batch, theta, np.float64, (E, P))
        elbo = self._empty(batch, (E,), np.float64)
        grad = self._empty(batch, (E, P), np.float64) if with_grad else None
        status = self._empty(batch, (E,), np.int32)
        cb = batch._c(False)
        self._pre(batch)
        _lib.check(self._lib.gpsat_vff_elbo(self._h, C.byref(cb), C.byref(dom), mem, _ptr(th), _ptr(elbo), _ptr(grad),
                                            _ptr(status)))
        return elbo, grad, status

    def vff_predict(self, batch: RaggedBatch, vff, theta):
        """``GPR_kron.predict_f`` / ``predict_y`` (full_cov=False) per expert at fixed theta."""
        E, P = batch.E, 2 * batch.D + 1
        mem = _lib.MEM_DEVICE if batch.on_device else _lib.MEM_HOST
        M = int(batch.off_m[-1])
        dom, keep = self._vff_domain(batch, vff)
        th = self._in(batch, theta, np.float64, (E, P))
        fm, fv, yv = (self._empty(batch, (M,), np.float64) for _ in range(3))
        elbo = self._empty(batch, (E,), np.float64)
        status = self._empty(batch, (E,), np.int32)
        cb = batch._c(True)
        self._pre(batch)
        _lib.check(self._lib.gpsat_vff_predict(self._h, C.byref(cb), C.byref(dom), mem, _ptr(th), _ptr(fm), _ptr(fv),
                                               _ptr(yv), _ptr(elbo), _ptr(status)))
        return {"f*": fm, "f*_var": fv, "y_var": yv, "elbo": elbo, "status": status}

    def factor_debug(self, kernel: str, X, y, theta, want_U: bool = True):
        X = _f64(X)
        X = X.reshape(len(X), -1)
        b = RaggedBatch.from_lists(kernel, [X], [_f64(y)])
        N = len(X) + 1
        L = np.empty((N, N))
        U = np.empty((N, N)) if want_U else None
        th = _f64(theta)
        cb = b._c(False)
        _lib.check(self._lib.gpsat_gpr_factor_debug(self._h, C.byref(cb), _ptr(th), _ptr(L), _ptr(U)))
        return L, U
