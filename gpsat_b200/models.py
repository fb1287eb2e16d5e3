"""Host-side mirror of GPSat's model plugin surface, backed by the CUDA engine.

``BaseGPRModel`` restates the abstract base every GPSat local-expert model inherits from
(reference: ``GPSat/models/base_model.py:17-448``) without its TensorFlow import;
``B200GPRModel`` has the method surface of ``GPflowGPRModel``
(``GPSat/models/gpflow_models.py:26-663``) - same method names, argument meaning, return types
and error behaviour - but every number it returns comes from ``libgpsat_b200.so`` (batch of one
expert through the C ABI).  The reference driver can select it without any change to GPSat via
the custom-model hook (``GPSat/local_experts.py:319-325``)::

    "model": {"oi_model": {"path_to_model": "gpsat_b200.models", "model_name": "B200GPRModel"}, ...}

There is no CPU fallback: constructing a model is pure host bookkeeping, but the first call that
needs arithmetic (``optimise_parameters`` / ``get_objective_function_value`` / ``predict``)
creates the engine and raises if the library or a B200 is missing.

For throughput use the batched entry points (:func:`fit_predict_models`, or
``gpsat_b200.local_experts.LocalExpertOI``), which pack many model instances into one
``RaggedBatch`` and make one engine call per phase instead of one per expert.
"""
from __future__ import annotations

import re
import warnings
from abc import ABC, abstractmethod
from typing import Dict, List, Optional, Sequence, Union

import numpy as np

from . import _lib
from .engine import Engine, RaggedBatch

try:  # pandas is optional for the array form of the constructor
    import pandas as pd
except Exception:  # pragma: no cover
    pd = None

# gpflow.likelihoods.Gaussian(variance_lower_bound=1e-6): the default likelihood-variance transform
# is 1e-6 + softplus(xi) (the reference refers to the bound at gpflow_models.py:404-409)
VARIANCE_LOWER_BOUND = 1e-6

_ENGINES: Dict[int, Engine] = {}


def get_engine(device: int = 0) -> Engine:
    """Process-wide engine handle per CUDA device (created on first use)."""
    eng = _ENGINES.get(int(device))
    if eng is None:
        eng = Engine(int(device))
        _ENGINES[int(device)] = eng
    return eng


def default_device(device=None) -> int:
    """CUDA device of a model: the one given, else this process's ``LOCAL_RANK`` (one rank per GPU under
    ``torchrun``), else 0."""
    import os
    return int(device) if device is not None else int(os.environ.get("LOCAL_RANK", 0))


def _processor_name() -> str:
    """CPU model string, as BaseGPRModel._get_processor_name does on Linux (base_model.py:302-324)."""
    try:
        with open("/proc/cpuinfo") as fh:
            for line in fh:
                if "model name" in line:
                    return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


class BaseGPRModel(ABC):
    """Data handling common to all local-expert models (reference ``base_model.py:83-277``).

    * ``data`` (DataFrame) with ``coords_col`` / ``obs_col``, or ``coords`` / ``obs`` arrays
      (1-d arrays become column vectors) - ``base_model.py:134-185``
    * NaNs are rejected (``:188-189``)
    * ``obs_mean='local'`` subtracts the column mean; **any other value is replaced by 0**, as in
      the reference (``:195-200``)
    * ``coords /= coords_scale; obs -= obs_mean; obs /= obs_scale`` on the model's own copy
      (``:243-245``)
    * every name in ``param_names`` must have ``get_<name>`` / ``set_<name>`` (``:270-277``)
    """

    def __init__(self, data=None, coords_col=None, obs_col=None, coords=None, obs=None, coords_scale=None,
                 obs_scale=None, obs_mean=None, verbose=True, **kwargs):
        if data is not None:
            assert coords_col is not None, "data was provided, but coord_col was not"
            assert obs_col is not None, "data was provided, but obs_col was not"
            if isinstance(coords_col, str):
                coords_col = [coords_col]
            if isinstance(obs_col, str):
                obs_col = [obs_col]
            self.obs = np.array(data.loc[:, obs_col].values)        # own copy: scaled in place below
            self.coords = np.array(data.loc[:, coords_col].values)
        else:
            assert obs is not None, f"data is {data}, and so is obs: {obs}, provide either"
            assert coords is not None, f"data is {data}, and so is coords: {coords}, provide either"
            assert isinstance(obs, np.ndarray), "if obs is provided directly it must be an np.array"
            assert isinstance(coords, np.ndarray), "if obs is provided directly it must be an np.array"
            if obs.ndim == 1:
                obs = obs[:, None]
            if coords.ndim == 1:
                coords = coords[:, None]
            assert len(obs) == len(coords), "obs and coords lengths don't match "
            self.obs, self.coords = obs, coords
            if coords_col is None:
                coords_col = list(range(self.coords.shape[1]))
            if obs_col is None:
                obs_col = [0]
        self.obs_col, self.coords_col = obs_col, coords_col

        assert not np.isnan(self.coords).any(), "nans found in coords"
        assert not np.isnan(self.obs).any(), "nans found in obs"

        if isinstance(obs_mean, str) and obs_mean == "local":
            obs_mean = np.mean(self.obs, axis=0)
        else:
            obs_mean = np.array([0])[None, :]
        self.obs_mean = obs_mean

        def as_row(v):
            if v is None:
                return np.atleast_2d(1)
            if isinstance(v, list):
                return np.array(v)[None, :]
            if isinstance(v, (int, float)):
                return np.array([v])[None, :]
            return v

        self.obs_scale = as_row(obs_scale)
        self.coords_scale = as_row(coords_scale)
        if self.coords.dtype.kind in "iu":
            self.coords = self.coords.astype(float)
        if self.obs.dtype.kind in "iu":
            self.obs = self.obs.astype(float)
        self.coords /= self.coords_scale
        self.obs -= self.obs_mean
        self.obs /= self.obs_scale

        self.gpu_name, self.cpu_name = self._get_device_names()

        for pn in self.param_names:
            assert not bool(re.search(" ", pn)), f"param_name: '{pn}' has a space (' ') in it, which is prohibited"
            getattr(self, f"set_{pn}")
            getattr(self, f"get_{pn}")

    def _get_device_names(self):
        """(gpu_name, cpu_name) for ``run_details.device`` (``local_experts.py:1180``).  The
        reference asks TensorFlow (``base_model.py:279-300``); here the CUDA runtime is asked
        through the C ABI, and only if the library is already loaded (no side effects in a ctor)."""
        gpu = None
        try:
            gpu = _lib.device_name(getattr(self, "device", 0))
        except Exception:
            gpu = None
        return gpu, _processor_name()

    @abstractmethod
    def predict(self, coords) -> Dict[str, np.ndarray]:
        ...

    @abstractmethod
    def optimise_parameters(self):
        ...

    @property
    @abstractmethod
    def param_names(self) -> List[str]:
        ...

    @abstractmethod
    def get_objective_function_value(self):
        ...

    def get_parameters(self, *args, return_dict=True) -> Union[dict, list]:
        """``get_<name>()`` for the given names, or all ``param_names`` (``base_model.py:370-403``)."""
        if len(args) == 0:
            args = self.param_names
        for a in args:
            assert a in self.param_names, f"cannot get parameters for: {a}, it's not in param_names: {self.param_names}"
        if return_dict:
            return {a: getattr(self, f"get_{a}")() for a in args}
        return [getattr(self, f"get_{a}")() for a in args]

    def set_parameters(self, **kwargs):
        """``set_<name>(value)`` for every keyword (``base_model.py:405-422``)."""
        for k, v in kwargs.items():
            assert k in self.param_names, f"cannot get parameters for: {k}, it's not in param_names: {self.param_names}"
            getattr(self, f"set_{k}")(v)

    def set_parameter_constraints(self, constraints_dict, **kwargs):
        """``set_<name>_constraints(**spec, **kwargs)`` per entry (``base_model.py:424-439``)."""
        for k, v in constraints_dict.items():
            assert k in self.param_names, f"cannot get parameters for: {k}, it's not in param_names: {self.param_names}"
            getattr(self, f"set_{k}_constraints")(**v, **kwargs)


# prior families of the extension `priors=` (GPflow convention: the log prior density of the
# constrained value plus the log-Jacobian of the transform is added to the log marginal likelihood)
PRIOR_IDS = {"none": 0, "normal": 1, "lognormal": 2, "gamma": 3}


class B200GPRModel(BaseGPRModel):
    """Exact GP regression, ARD stationary kernel, Gaussian likelihood, float64, on one B200.

    Mirrors ``GPflowGPRModel`` (``gpflow_models.py:33-157``): ``kernel`` is one of "Matern12"
    (= "Exponential"), "Matern32" (default), "Matern52", "RBF" (= "SquaredExponential");
    ``kernel_kwargs`` may hold ``lengthscales`` (default ``ones(D)``, ``:129-131``) and ``variance``
    (default 1.0); ``noise_variance`` defaults to GPflow's 1.0.  Unknown keyword arguments are
    swallowed like the reference's ``**kwargs``.  Mean functions other than zero and
    custom likelihood objects are not part of the engine and raise ``NotImplementedError``.
    """

    _P_NAMES = ["lengthscales", "kernel_variance", "likelihood_variance"]

    def __init__(self, data=None, coords_col=None, obs_col=None, coords=None, obs=None, coords_scale=None,
                 obs_scale=None, obs_mean=None, verbose=True, *, kernel="Matern32", kernel_kwargs=None,
                 mean_function=None, mean_func_kwargs=None, noise_variance=None, likelihood=None,
                 device: Optional[int] = None, priors: Optional[dict] = None, **kwargs):
        self.device = default_device(device)
        super().__init__(data=data, coords_col=coords_col, obs_col=obs_col, coords=coords, obs=obs,
                         coords_scale=coords_scale, obs_scale=obs_scale, obs_mean=obs_mean, verbose=verbose)
        assert kernel is not None, "kernel was not provided"
        if not isinstance(kernel, str):
            raise NotImplementedError("B200GPRModel takes the kernel by name (a gpflow.kernels object has no meaning here)")
        if kernel not in _lib.KERNEL_IDS:
            raise NotImplementedError(f"kernel '{kernel}' is not implemented; choose from {sorted(_lib.KERNEL_IDS)}")
        if mean_function is not None:
            raise NotImplementedError("only the zero mean function is implemented (use obs_mean='local')")
        if likelihood is not None:
            raise NotImplementedError("custom likelihood objects are not supported; pass noise_variance")
        assert self.obs.shape[1] == 1, "B200GPRModel takes one observation column"
        self.kernel = kernel
        D = self.coords.shape[1]
        assert 1 <= D <= 6, "B200GPRModel supports 1..6 input dimensions"
        kernel_kwargs = dict(kernel_kwargs or {})
        ls = np.asarray(kernel_kwargs.pop("lengthscales", np.ones(D)), dtype=np.float64)
        ls = np.full(D, float(ls)) if ls.ndim == 0 else ls.reshape(D).copy()
        var = float(kernel_kwargs.pop("variance", 1.0))
        kernel_kwargs.pop("active_dims", None)
        kernel_kwargs.pop("name", None)
        if kernel_kwargs:
            raise TypeError(f"unexpected kernel_kwargs: {sorted(kernel_kwargs)}")
        noise = 1.0 if noise_variance is None else float(noise_variance)
        assert noise > VARIANCE_LOWER_BOUND, "noise_variance must exceed the likelihood's variance_lower_bound (1e-6)"
        # constrained parameter vector and its bijections: [l_1..l_D, kernel variance, likelihood variance]
        self._theta = np.concatenate([ls, [var, noise]])
        P = D + 2
        self._tkind = np.zeros(P, np.int32)             # softplus
        self._ta = np.zeros(P)
        self._tb = np.zeros(P)
        self._ta[P - 1] = VARIANCE_LOWER_BOUND          # shifted softplus on the likelihood variance
        self._trainable = np.ones(P, bool)
        self._pkind = np.zeros(P, np.int32)
        self._pa = np.zeros(P)
        self._pb = np.zeros(P)
        if priors:
            for name, spec in priors.items():
                self.set_prior(name, **spec)
        self.opt_info: dict = {}

    # ---- shape helpers ------------------------------------------------------------------------
    @property
    def D(self) -> int:
        return self.coords.shape[1]

    def _slice(self, name):
        D = self.D
        return {"lengthscales": slice(0, D), "kernel_variance": slice(D, D + 1),
                "likelihood_variance": slice(D + 1, D + 2)}[name]

    def _batch(self, pred_coords=None) -> RaggedBatch:
        X = np.ascontiguousarray(self.coords, dtype=np.float64)
        y = np.ascontiguousarray(self.obs[:, 0], dtype=np.float64)
        return RaggedBatch.from_lists(self.kernel, [X], [y], None if pred_coords is None else [pred_coords])

    @property
    def engine(self) -> Engine:
        return get_engine(self.device)

    def _inducing_list(self):
        """``None`` for the exact model; the sparse subclass returns ``[Z]``."""
        return None

    def _fit_kwargs(self) -> dict:
        """Extra ``Engine.fit`` arguments that select the model family."""
        return {"inducing": self._inducing_list()}

    def update_obs_data(self, data=None, coords_col=None, obs_col=None, coords=None, obs=None, coords_scale=None,
                        obs_scale=None):
        """Swap the training data, keeping hyper-parameters (``gpflow_models.py:159-176``)."""
        BaseGPRModel.__init__(self, data=data, coords_col=coords_col, obs_col=obs_col, coords=coords, obs=obs,
                              coords_scale=coords_scale, obs_scale=obs_scale)

    @property
    def param_names(self) -> list:
        return list(self._P_NAMES)

    # ---- predict --------------------------------------------------------------------------------
    def _prep_pred_coords(self, coords, apply_scale=True) -> np.ndarray:
        """Input handling of ``GPflowGPRModel.predict`` (``gpflow_models.py:212-227``): DataFrame /
        Series / list / ndarray; a 1-d array is one point with D coordinates."""
        if pd is not None and isinstance(coords, (pd.Series, pd.DataFrame)):
            coords = coords[self.coords_col].values if self.coords_col is not None else coords.values
        if isinstance(coords, list):
            coords = np.array(coords)
        assert isinstance(coords, np.ndarray), "coords should be an ndarray (one can be converted from)"
        if coords.ndim == 1:
            coords = coords[None, :]
        coords = coords.astype(self.coords.dtype)
        if apply_scale:
            coords = coords / self.coords_scale
        assert coords.shape[1] == self.D, f"prediction coords have {coords.shape[1]} columns, model has {self.D}"
        return np.ascontiguousarray(coords, dtype=np.float64)

    def _f_bar(self, m: int) -> np.ndarray:
        f_bar = np.asarray(self.obs_mean)[:, 0] if np.asarray(self.obs_mean).ndim == 2 else np.atleast_1d(self.obs_mean)
        if len(f_bar) != m:
            assert len(f_bar) == 1, f"'f_bar' did not match the length of 'f*' and f_bar len is not, got: {len(f_bar)}"
            return np.repeat(f_bar, m)
        return f_bar

    def predict(self, coords, full_cov=False, apply_scale=True) -> Dict[str, np.ndarray]:
        """Posterior mean "f*", marginal variance "f*_var", predictive variance "y_var" and the
        stored mean "f_bar" at ``coords`` (``gpflow_models.py:186-273``); ``full_cov=True`` adds
        "f*_cov" and "y_cov".  Values are in scaled-observation units, as in the reference."""
        Xs = self._prep_pred_coords(coords, apply_scale)
        b = self._batch(Xs)
        if full_cov:
            out = self.engine.predict_full_cov(b, self._theta[None, :])
        else:
            out = self.engine.predict(b, self._theta[None, :])
        if int(out["status"][0]) != _lib.STATUS_OK:
            raise FloatingPointError("K + likelihood_variance * I is not positive definite at the current parameters")
        res = {"f*": out["f*"], "f*_var": out["f*_var"], "y_var": out["y_var"]}
        if full_cov:
            f_cov = out["f*_cov"][0]
            y_cov = f_cov.copy()
            y_cov[np.diag_indices(len(y_cov))] += res["y_var"] - res["f*_var"]
            res["f*_var"] = np.diag(f_cov).copy()
            res["f*_cov"] = f_cov
            res["y_cov"] = y_cov
        res["f_bar"] = self._f_bar(len(res["f*"]))
        return res

    # ---- optimise ---------------------------------------------------------------------------------
    def _fix_hyperparameters(self, params_list):
        """``gpflow.set_trainable(param, False)`` (``gpflow_models.py:275-288``); like there, the
        flag persists on the model."""
        for param in params_list:
            if param in self._P_NAMES:
                self._trainable[self._slice(param)] = False
            else:
                print(f"{param} is not detected as a hyperparameter. Skipping...")

    def optimise_parameters(self, max_iter=10_000, fixed_params=None, **opt_kwargs):
        """L-BFGS on -(log marginal likelihood + log prior) in unconstrained space, SciPy
        ``L-BFGS-B`` defaults (``gpflow_models.py:290-329``).  Returns the optimiser's ``success``.

        ``opt_kwargs`` understood: ``method`` (must be "L-BFGS-B"), ``tol`` (SciPy semantics: sets
        both ``ftol`` and ``gtol``), and the gpflow ``Scipy.minimize`` flags ``compile``,
        ``allow_unused_variables``, ``tf_fun_args``, ``track_loss_history`` (accepted, no effect).
        """
        if fixed_params is None:
            fixed_params = []
        self._fix_hyperparameters(fixed_params)
        method = opt_kwargs.pop("method", "L-BFGS-B")
        if method != "L-BFGS-B":
            raise NotImplementedError(f"optimiser method '{method}' is not implemented on the device (L-BFGS-B only)")
        tol = opt_kwargs.pop("tol", None)
        for k in ("compile", "allow_unused_variables", "tf_fun_args", "track_loss_history"):
            opt_kwargs.pop(k, None)
        if opt_kwargs.pop("step_callback", None) is not None:
            raise NotImplementedError("step_callback is not available: the optimiser loop runs on the device")
        if opt_kwargs:
            raise TypeError(f"unexpected optimiser arguments: {sorted(opt_kwargs)}")
        kw = {}
        if tol is not None:
            kw = {"ftol": float(tol), "pgtol": float(tol)}
        res = self.engine.fit(self._batch(), self._theta[None, :], self._tkind[None, :], self._ta[None, :],
                              self._tb[None, :], self._trainable[None, :].astype(np.uint8), max_iter=int(max_iter),
                              priors=self._prior_arrays(1), **self._fit_kwargs(), **kw)
        status = int(res.status[0])
        self.opt_info = {"status": status, "n_evals": int(res.n_evals[0]), "n_iters": int(res.n_iters[0]),
                         "objective": float(res.objective[0])}
        if status != _lib.STATUS_NOT_PD:
            self._theta = np.array(res.theta[0], dtype=np.float64)
        success = status == _lib.STATUS_OK
        if not success:
            print("*" * 10)
            print("optimization failed!")
        return bool(success)

    def get_objective_function_value(self):
        """Negative log marginal likelihood at the current parameters (``gpflow_models.py:334-337``)."""
        lml, _, status = self.engine.lml_grad(self._batch(), self._theta[None, :], with_grad=False)
        return -float(lml[0])

    # ---- getters / setters ----------------------------------------------------------------------
    def get_lengthscales(self) -> np.ndarray:
        return self._theta[: self.D].copy()

    def get_kernel_variance(self) -> float:
        return float(self._theta[self.D])

    def get_likelihood_variance(self) -> float:
        return float(self._theta[self.D + 1])

    def _assign(self, name, values):
        """gpflow ``Parameter.assign``: the value must map to a finite unconstrained value."""
        sl = self._slice(name)
        values = np.asarray(values, dtype=np.float64)
        k, a, b = self._tkind[sl], self._ta[sl], self._tb[sl]
        ok = np.where(k == _lib.TRANSFORM_SIGMOID, (values > a) & (values < b), values > a)
        if not np.all(ok) or not np.all(np.isfinite(values)):
            raise ValueError(f"cannot assign {values} to '{name}': outside the parameter's transform domain "
                             f"(low={a}, high={np.where(k == _lib.TRANSFORM_SIGMOID, b, np.inf)})")
        self._theta[sl] = values

    def set_lengthscales(self, lengthscales):
        """Array-like of size D, or a scalar applied to every dimension (``gpflow_models.py:351-361``)."""
        v = np.asarray(lengthscales, dtype=np.float64)
        v = np.full(self.D, float(v)) if v.ndim == 0 else v.reshape(-1)
        assert len(v) == self.D, f"lengthscales must have {self.D} entries, got {len(v)}"
        self._assign("lengthscales", v)

    @staticmethod
    def _scalar(value, what):
        if isinstance(value, np.ndarray):
            assert (len(value) == 1) & (len(value.shape) == 1), \
                f"set_{what} expected to receive float, or np.array with len(1), shape:(1,), got" \
                f"len: {len(value)}, shape: {value.shape}"
            value = value[0]
        return float(value)

    def set_kernel_variance(self, kernel_variance):
        """float or shape-(1,) array (``gpflow_models.py:363-381``)."""
        self._assign("kernel_variance", [self._scalar(kernel_variance, "kernel_variance")])

    def set_likelihood_variance(self, likelihood_variance):
        """float or shape-(1,) array; values under the likelihood's lower bound (1e-6) are raised to
        it with a warning (``gpflow_models.py:383-411``)."""
        v = self._scalar(likelihood_variance, "likelihood_variance")
        if v < VARIANCE_LOWER_BOUND:
            warnings.warn("\n***\ntrying to set likelihood_variance to value less than "
                          "model.likelihood.variance_lower_bound\nwill set to variance_lower_bound\n***\n")
            v = VARIANCE_LOWER_BOUND
        sl = self._slice("likelihood_variance")
        if self._tkind[sl][0] == _lib.TRANSFORM_SOFTPLUS and v <= self._ta[sl][0]:
            v = float(np.nextafter(self._ta[sl][0], np.inf))     # boundary of 1e-6 + softplus: xi = -inf in GPflow
        self._assign("likelihood_variance", [v])

    # ---- constraints ----------------------------------------------------------------------------
    def _set_param_constraints(self, name, low, high, move_within_tol=True, tol=1e-8, scale=False,
                               scale_magnitude=None):
        """Sigmoid(low, high) re-parametrisation (``gpflow_models.py:416-494``): optional division of
        the bounds by ``coords_scale`` (``scale=True``), then the current value is moved inside
        ``[low + tol, high - tol]`` with ``tol`` capped at half the narrowest interval."""
        sl = self._slice(name)
        if isinstance(low, (list, tuple)):
            low = np.array(low, dtype=np.float64)
        elif isinstance(low, (int, np.integer, float)):
            low = np.array([low], dtype=np.float64)
        if isinstance(high, (list, tuple)):
            high = np.array(high, dtype=np.float64)
        elif isinstance(high, (int, np.integer, float)):
            high = np.array([high], dtype=np.float64)
        low, high = np.asarray(low, np.float64), np.asarray(high, np.float64)
        assert len(low.shape) == 1
        assert len(high.shape) == 1
        vals = self._theta[sl].copy()
        assert len(vals) == len(low), "len of low constraint does not match param length"
        assert len(vals) == len(high), "len of high constraint does not match param length"
        assert np.all(low <= high), "all values in high constraint must be greater than low"
        if scale:
            if scale_magnitude is None:
                low = low / self.coords_scale[0, :]
                high = high / self.coords_scale[0, :]
            else:
                low = low / scale_magnitude
                high = high / scale_magnitude
        if move_within_tol:
            half_min_width = np.min(high - low) / 2
            if tol > half_min_width:
                tol = half_min_width
            hi_mask = vals > (high - tol)
            vals[hi_mask] = high[hi_mask] - tol
            lo_mask = vals < (low + tol)
            vals[lo_mask] = low[lo_mask] + tol
        self._tkind[sl] = _lib.TRANSFORM_SIGMOID
        self._ta[sl] = low
        self._tb[sl] = high
        self._assign(name, vals)

    def set_lengthscales_constraints(self, low, high, move_within_tol=True, tol=1e-8, scale=False,
                                     scale_magnitude=None):
        self._set_param_constraints("lengthscales", low, high, move_within_tol, tol, scale, scale_magnitude)

    def set_kernel_variance_constraints(self, low, high, move_within_tol=True, tol=1e-8, scale=False,
                                        scale_magnitude=None):
        self._set_param_constraints("kernel_variance", low, high, move_within_tol, tol, scale, scale_magnitude)

    def set_likelihood_variance_constraints(self, low, high, move_within_tol=True, tol=1e-8, scale=False,
                                            scale_magnitude=None):
        self._set_param_constraints("likelihood_variance", low, high, move_within_tol, tol, scale, scale_magnitude)

    # ---- priors (extension asked for by the north star; the reference never sets one) -------------
    def set_prior(self, name, family="none", a=0.0, b=1.0):
        """Prior density on the constrained value of parameter ``name``: "normal"(mean a, std b),
        "lognormal"(mu a, sigma b), "gamma"(shape a, rate b) or "none".  Follows GPflow's
        ``log_prior_density`` (prior on the constrained value plus the transform's log-Jacobian),
        which ``training_loss`` subtracts."""
        assert name in self._P_NAMES, f"unknown parameter '{name}'"
        assert family in PRIOR_IDS, f"unknown prior family '{family}', choose from {sorted(PRIOR_IDS)}"
        sl = self._slice(name)
        self._pkind[sl] = PRIOR_IDS[family]
        self._pa[sl] = a
        self._pb[sl] = b

    def _prior_arrays(self, E: int):
        if not np.any(self._pkind):
            return None
        return (np.tile(self._pkind, (E, 1)), np.tile(self._pa, (E, 1)), np.tile(self._pb, (E, 1)))


class B200SGPRModel(B200GPRModel):
    """Sparse GP regression (Titsias 2009) with fixed inducing points, mirroring ``GPflowSGPRModel``
    (``gpflow_models.py:666-901``): the ``num_inducing_points`` (default 500) inducing points are
    the first rows of a shuffled copy of the scaled training coordinates (``:816-819``; all of them
    when there are fewer data points, ``:810-814``), the objective value is **+ELBO**
    (``:860-862``), ``predict`` is the parent's (``predict_f`` / ``predict_y`` of the sparse model)
    and ``param_names`` gains "inducing_points".  ``optimise_parameters(train_inducing_points=True)``
    (``:864-899``) optimises Z together with the hyper-parameters (host L-BFGS-B around CUDA evaluations).
    """

    def __init__(self, data=None, coords_col=None, obs_col=None, coords=None, obs=None, coords_scale=None,
                 obs_scale=None, obs_mean=None, verbose=True, *, kernel="Matern32", num_inducing_points=500,
                 kernel_kwargs=None, mean_function=None, mean_func_kwargs=None, noise_variance=None,
                 likelihood=None, device: Optional[int] = None, priors: Optional[dict] = None, **kwargs):
        super().__init__(data=data, coords_col=coords_col, obs_col=obs_col, coords=coords, obs=obs,
                         coords_scale=coords_scale, obs_scale=obs_scale, obs_mean=obs_mean, verbose=verbose,
                         kernel=kernel, kernel_kwargs=kernel_kwargs, mean_function=mean_function,
                         mean_func_kwargs=mean_func_kwargs, noise_variance=noise_variance, likelihood=likelihood,
                         device=device, priors=priors, **kwargs)
        assert num_inducing_points is not None, "num_inducing_points is None, must be specified for SGPR"
        if len(self.coords) < num_inducing_points:
            print("number of inducing points is more than number of data points, "
                  "setting inducing points to data points...")
            self.inducing_points = self.coords
        else:
            X = self.coords.copy()
            np.random.shuffle(X)          # the reference draws from numpy's global generator too
            self.inducing_points = X[:int(num_inducing_points)]

    def _inducing_list(self):
        return [np.ascontiguousarray(self.inducing_points, dtype=np.float64)]

    @property
    def param_names(self) -> list:
        return super().param_names + ["inducing_points"]

    def get_inducing_points(self) -> np.ndarray:
        return np.array(self.inducing_points, dtype=np.float64)

    def set_inducing_points(self, inducing_points):
        Z = np.asarray(inducing_points, dtype=np.float64)
        assert Z.ndim == 2 and Z.shape[1] == self.D, f"inducing_points must have shape [M, {self.D}]"
        assert np.all(np.isfinite(Z)), "inducing_points contains non-finite values"
        self.inducing_points = Z.copy()

    def predict(self, coords, full_cov=False, apply_scale=True) -> Dict[str, np.ndarray]:
        Xs = self._prep_pred_coords(coords, apply_scale)
        out = self.engine.sgpr_predict(self._batch(Xs), self._inducing_list(), self._theta[None, :], full_cov=full_cov)
        if int(out["status"][0]) != _lib.STATUS_OK:
            raise FloatingPointError("Kuu + jitter * I or B is not positive definite at the current parameters")
        res = {"f*": out["f*"], "f*_var": out["f*_var"], "y_var": out["y_var"]}
        if full_cov:                                  # same keys as the exact model (gpflow_models.py:245-263)
            f_cov = out["f*_cov"][0]
            y_cov = f_cov.copy()
            y_cov[np.diag_indices(len(y_cov))] += res["y_var"] - res["f*_var"]
            res["f*_var"] = np.diag(f_cov).copy()
            res["f*_cov"] = f_cov
            res["y_cov"] = y_cov
        res["f_bar"] = self._f_bar(len(res["f*"]))
        return res

    def optimise_parameters(self, train_inducing_points=False, max_iter=10_000, fixed_params=None, **opt_kwargs):
        """``gpflow_models.py:864-899``.  With the inducing points fixed (the reference's default, ``:896-897``) this is
        the parent's device-side batched L-BFGS.  ``train_inducing_points=True`` makes Z [M, D] (unconstrained, as in
        GPflow) part of the search: D + 2 + M D variables are beyond the device optimiser's state, so the loop is the
        host routine GPflow's ``Scipy`` optimiser calls itself - ``scipy.optimize.minimize(method="L-BFGS-B", jac=True)``
        - around one CUDA evaluation per step (``gpsat_sgpr_elbo_gradz``: ELBO, d/d theta and d/d Z in one pass)."""
        if not train_inducing_points:
            return super().optimise_parameters(max_iter=max_iter, fixed_params=fixed_params, **opt_kwargs)
        import scipy.optimize as sopt
        if np.any(self._pkind):
            raise NotImplementedError("priors are not supported together with train_inducing_points=True")
        self._fix_hyperparameters([] if fixed_params is None else fixed_params)
        method = opt_kwargs.pop("method", "L-BFGS-B")
        if method != "L-BFGS-B":
            raise NotImplementedError(f"optimiser method '{method}' is not implemented (L-BFGS-B only)")
        tol = opt_kwargs.pop("tol", None)
        for k in ("compile", "allow_unused_variables", "tf_fun_args", "track_loss_history"):
            opt_kwargs.pop(k, None)
        if opt_kwargs:
            raise TypeError(f"unexpected optimiser arguments: {sorted(opt_kwargs)}")
        free = self._trainable.astype(bool)
        nfree = int(free.sum())
        sig = self._tkind == _lib.TRANSFORM_SIGMOID
        a, b = self._ta, self._tb
        theta0 = self._theta.copy()
        Z0 = self.get_inducing_points()
        y0 = theta0 - a
        with np.errstate(divide="ignore", invalid="ignore"):
            xi0 = np.where(sig, np.log(theta0 - a) - np.log(b - theta0), y0 + np.log(-np.expm1(-y0)))
        batch = self._batch().to_device(f"cuda:{self.device}")      # X, y stay resident across the evaluations
        eng = self.engine

        def unpack(v):
            xi = xi0.copy()
            xi[free] = v[:nfree]
            s = np.where(xi >= 0, 1.0 / (1.0 + np.exp(-np.abs(xi))), np.exp(-np.abs(xi)) / (1.0 + np.exp(-np.abs(xi))))
            sp = np.log1p(np.exp(-np.abs(xi))) + np.maximum(xi, 0.0)
            theta = np.where(free, np.where(sig, b * s + a * (1.0 - s), a + sp), theta0)
            dth = np.where(sig, (b - a) * s * (1.0 - s), s)
            return theta, dth, v[nfree:].reshape(Z0.shape)

        def fun(v):
            theta, dth, Z = unpack(v)
            elbo, g, gz, st = eng.sgpr_elbo(batch, [Z], theta[None, :], with_z=True)
            elbo, g, gz, st = (np.asarray(t.cpu() if hasattr(t, "cpu") else t) for t in (elbo, g, gz, st))
            if int(st[0]) != _lib.STATUS_OK or not np.isfinite(elbo[0]):
                return np.inf, np.zeros(len(v))
            return -float(elbo[0]), np.concatenate([(-g[0] * dth)[free], -gz.ravel()])

        options = {"maxiter": int(max_iter)}
        res = sopt.minimize(fun, np.concatenate([xi0[free], Z0.ravel()]), jac=True, method="L-BFGS-B", tol=tol, options=options)
        theta, _, Z = unpack(res.x)
        if np.isfinite(res.fun):
            self._theta = np.array(theta, dtype=np.float64)
            self.inducing_points = np.array(Z, dtype=np.float64)
        self.opt_info = {"status": 0 if res.success else 2, "n_evals": int(res.nfev), "n_iters": int(res.nit),
                         "objective": float(res.fun)}
        if not res.success:
            print("*" * 10)
            print("optimization failed!")
        return bool(res.success)

    def get_objective_function_value(self):
        """ELBO at the current parameters (``gpflow_models.py:860-862``; note the sign: not negated)."""
        elbo, _, _ = self.engine.sgpr_elbo(self._batch(), self._inducing_list(), self._theta[None, :])
        return float(elbo[0])


class B200VFFModel(B200GPRModel):
    """Variational Fourier features (Hensman et al. 2017), mirroring ``GPflowVFFModel``
    (``GPSat/models/vff_model.py:14-264`` -> ``GPR_kron``, ``GPSat/vff.py:564-675``): one 1-D Matern
    kernel per input dimension, ``num_inducing_features`` frequencies per dimension (int or list;
    ``2 k - 1`` cos / sin features each, their product over the dimensions in total), on the box
    ``expert_loc +- domain_size`` (unscaled units) widened to contain the data (``:180-199``).

    Parameters: "lengthscales" has one entry per dimension; every 1-D kernel has its own variance,
    ``get_kernel_variance`` returns their product and ``set_kernel_variance`` gives each the D-th
    root (``:213-233``); bounds passed to ``set_kernel_variance_constraints`` are raised to 1/D
    (``:249-258``).  The likelihood variance starts at GPflow's 1.0 (no ``noise_variance`` argument
    in the reference).  ``get_objective_function_value`` is **+ELBO** (``:209-211``).
    Reference quirk kept: ``fixed_params=["lengthscales"]`` / ``["kernel_variance"]`` freeze only the
    *last* dimension's kernel, because ``GPR_kron`` registers that one as ``model.kernel``
    (``vff.py:572``, ``gpflow_models.py:281-284``).  ``domain_size=None`` (which raises ``TypeError``
    in the reference, ``vff_model.py:177``) selects the data range +- 1e-8, the branch the reference
    intended (``:181-184``).  At most 3 input dimensions and 128 frequencies per dimension.
    """

    def __init__(self, data=None, coords_col=None, obs_col=None, coords=None, coords_scale=None, obs=None,
                 obs_scale=None, obs_mean=None, *, kernel="Matern32", num_inducing_features=None, kernel_kwargs=None,
                 mean_function=None, mean_func_kwargs=None, domain_size=None, expert_loc=None, device: Optional[int] = None,
                 verbose=True, **kwargs):
        self.device = default_device(device)
        BaseGPRModel.__init__(self, data=data, coords_col=coords_col, obs_col=obs_col, coords=coords, obs=obs,
                              coords_scale=coords_scale, obs_scale=obs_scale, obs_mean=obs_mean, verbose=verbose)
        assert kernel is not None, "kernel was not provided"
        assert num_inducing_features is not None, "Number of inducing points per dimension not specified"
        if kernel not in ("Matern12", "Matern32", "Matern52"):
            raise NotImplementedError(f"VFF is defined for Matern12 / Matern32 / Matern52, got '{kernel}'")
        if mean_function is not None:
            raise NotImplementedError("only the zero mean function is implemented")
        assert self.obs.shape[1] == 1, "B200VFFModel takes one observation column"
        self.kernel = kernel
        D = self.coords.shape[1]
        assert 1 <= D <= 3, "B200VFFModel supports 1..3 input dimensions"
        if kernel_kwargs is None:
            kernel_kwargs = [dict(lengthscales=1.0) for _ in range(D)]
        elif isinstance(kernel_kwargs, dict):
            kernel_kwargs = [dict(kernel_kwargs) for _ in range(D)]
        assert len(kernel_kwargs) == D
        ls = np.array([float(np.asarray(k.get("lengthscales", 1.0))) for k in kernel_kwargs])
        var = np.array([float(k.get("variance", 1.0)) for k in kernel_kwargs])
        # domain of the Fourier features, scaled coordinates (vff_model.py:174-199)
        if isinstance(domain_size, (int, float)):
            domain_size = [domain_size for _ in range(D)]
        cmin, cmax = self.coords.min(axis=0), self.coords.max(axis=0)
        if domain_size is None:
            a, b = cmin - 1e-8, cmax + 1e-8
        else:
            assert len(domain_size) == D, "length of margin list must match number of coordinate dimensions"
            if expert_loc is None:
                expert_loc = np.mean(self.coords, axis=0) * self.coords_scale.squeeze()
            expert_loc = np.asarray(expert_loc, dtype=np.float64).reshape(-1)
            scale = np.broadcast_to(np.atleast_2d(self.coords_scale)[0, :], (D,))
            a = (expert_loc[:D] - np.asarray(domain_size, dtype=np.float64)) / scale
            b = (expert_loc[:D] + np.asarray(domain_size, dtype=np.float64)) / scale
            a = np.where(a < cmin, a, cmin - 1e-8)
            b = np.where(b > cmax, b, cmax + 1e-8)
        self.vff_a, self.vff_b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
        if isinstance(num_inducing_features, (int, np.integer)):
            nf = [int(num_inducing_features)] * D
        else:
            nf = [int(k) for k in num_inducing_features]
            nf = nf * D if len(nf) == 1 else nf
        assert len(nf) == D, "Number of ms must match number of kernels."
        assert all(1 <= k <= 128 for k in nf), "num_inducing_features must be in [1, 128] per dimension"
        self.vff_nfreq = np.asarray(nf, dtype=np.int32)
        P = 2 * D + 1
        self._theta = np.concatenate([ls, var, [1.0]])
        self._tkind = np.zeros(P, np.int32)
        self._ta = np.zeros(P)
        self._tb = np.zeros(P)
        self._ta[P - 1] = VARIANCE_LOWER_BOUND
        self._trainable = np.ones(P, bool)
        self._pkind = np.zeros(P, np.int32)
        self._pa = np.zeros(P)
        self._pb = np.zeros(P)
        self.opt_info = {}

    def _slice(self, name):
        D = self.D
        return {"lengthscales": slice(0, D), "kernel_variance": slice(D, 2 * D),
                "likelihood_variance": slice(2 * D, 2 * D + 1)}[name]

    def _vff(self):
        return (self.vff_a[None, :], self.vff_b[None, :], self.vff_nfreq[None, :])

    def update_obs_data(self, data=None, coords_col=None, obs_col=None, coords=None, obs=None, coords_scale=None,
                        obs_scale=None):
        """Swap the training data (``gpflow_models.py:159-176`` inherited by the reference's VFF model, where the
        static ``Kuf`` products of ``GPR_kron.__init__`` are NOT rebuilt; here they are rebuilt on the next call).  The
        Fourier domain keeps its box and is widened, by the constructor's rule (``vff_model.py:186-199``), wherever the
        new data would fall outside it."""
        super().update_obs_data(data=data, coords_col=coords_col, obs_col=obs_col, coords=coords, obs=obs,
                                coords_scale=coords_scale, obs_scale=obs_scale)
        cmin, cmax = self.coords.min(axis=0), self.coords.max(axis=0)
        self.vff_a = np.where(self.vff_a < cmin, self.vff_a, cmin - 1e-8)
        self.vff_b = np.where(self.vff_b > cmax, self.vff_b, cmax + 1e-8)

    def _fit_kwargs(self) -> dict:
        return {"vff": self._vff()}

    @property
    def num_features(self) -> int:
        return int(np.prod(2 * self.vff_nfreq - 1))

    def _fix_hyperparameters(self, params_list):
        D = self.D
        for param in params_list:
            if param == "likelihood_variance":
                self._trainable[2 * D] = False
            elif param == "kernel_variance":
                self._trainable[2 * D - 1] = False          # model.kernel is the last dimension's kernel
            elif param == "lengthscales":
                self._trainable[D - 1] = False
            else:
                print(f"{param} is not detected as a hyperparameter. Skipping...")

    def predict(self, coords, full_cov=False, apply_scale=True) -> Dict[str, np.ndarray]:
        if full_cov:
            raise NotImplementedError("full_cov=True is not implemented (GPR_kron.predict_f raises too, vff.py:664-665)")
        Xs = self._prep_pred_coords(coords, apply_scale)
        if self.kernel == "Matern52":
            assert np.all(Xs >= self.vff_a) and np.all(Xs <= self.vff_b), "Edges not implemented for Matern52"
        out = self.engine.vff_predict(self._batch(Xs), self._vff(), self._theta[None, :])
        if int(out["status"][0]) != _lib.STATUS_OK:
            raise FloatingPointError("Kuf Kfu / s2 + Kuu is not positive definite at the current parameters")
        return {"f*": out["f*"], "f*_var": out["f*_var"], "y_var": out["y_var"], "f_bar": self._f_bar(len(out["f*"]))}

    def get_objective_function_value(self):
        """ELBO at the current parameters (``vff_model.py:209-211``)."""
        elbo, _, _ = self.engine.vff_elbo(self._batch(), self._vff(), self._theta[None, :])
        return float(elbo[0])

    def get_kernel_variance(self) -> float:
        return float(np.prod(self._theta[self._slice("kernel_variance")]))

    def set_kernel_variance(self, kernel_variance):
        v = self._scalar(kernel_variance, "kernel_variance")
        self._assign("kernel_variance", np.full(self.D, v ** (1.0 / self.D)))

    def get_likelihood_variance(self) -> float:
        return float(self._theta[2 * self.D])

    def set_kernel_variance_constraints(self, low, high, move_within_tol=True, tol=1e-8, scale=False,
                                        scale_magnitude=None):
        alpha = 1.0 / self.D
        low = float(np.asarray(low, dtype=np.float64).reshape(-1)[0]) ** alpha
        high = float(np.asarray(high, dtype=np.float64).reshape(-1)[0]) ** alpha
        self._set_param_constraints("kernel_variance", [low] * self.D, [high] * self.D, move_within_tol, tol, scale,
                                    scale_magnitude)


def fit_predict_models(models: Sequence[B200GPRModel], pred_coords: Optional[Sequence] = None, *,
                       optimise: bool = True, max_iter: int = 10_000, fixed_params=None, apply_scale: bool = True,
                       **opt_kwargs):
    """Batched equivalent (exact or sparse models) of the reference loop body (``local_experts.py:1126-1164``) for a list of
    model instances that share kernel and input dimension: one ``gpsat_gpr_fit`` call for all
    experts, then one ``gpsat_gpr_predict`` call.  Updates every model's parameters in place and
    returns ``(success[E], objective[E], predictions)`` where ``predictions`` is a list of the
    dicts ``model.predict`` would return (or ``None`` without ``pred_coords``)."""
    E = len(models)
    if E == 0:
        return np.zeros(0, bool), np.zeros(0), ([] if pred_coords is not None else None)
    m0 = models[0]
    assert all(m.kernel == m0.kernel and m.D == m0.D and m.device == m0.device for m in models), \
        "fit_predict_models needs one kernel / dimension / device per batch"
    sparse, fourier = isinstance(m0, B200SGPRModel), isinstance(m0, B200VFFModel)
    assert all(type(m) is type(m0) for m in models), \
        "fit_predict_models needs one model class per batch (partition by model type, local_experts.py:1021-1041)"
    inducing = [m._inducing_list()[0] for m in models] if sparse else None
    vff = ((np.stack([m.vff_a for m in models]), np.stack([m.vff_b for m in models]),
            np.stack([m.vff_nfreq for m in models])) if fourier else None)
    family = {"vff": vff} if fourier else {"inducing": inducing}
    for m in models:
        m._fix_hyperparameters(fixed_params or [])
    Xs_list = None
    if pred_coords is not None:
        Xs_list = [m._prep_pred_coords(c, apply_scale) for m, c in zip(models, pred_coords)]
    batch = RaggedBatch.from_lists(m0.kernel, [np.ascontiguousarray(m.coords, dtype=np.float64) for m in models],
                                   [np.ascontiguousarray(m.obs[:, 0], dtype=np.float64) for m in models], Xs_list)
    eng = m0.engine
    theta = np.stack([m._theta for m in models])
    success = np.ones(E, bool)
    if optimise:
        tol = opt_kwargs.pop("tol", None)
        kw = {} if tol is None else {"ftol": float(tol), "pgtol": float(tol)}
        pri = None
        if any(np.any(m._pkind) for m in models):
            pri = (np.stack([m._pkind for m in models]), np.stack([m._pa for m in models]),
                   np.stack([m._pb for m in models]))
        res = eng.fit(batch, theta, np.stack([m._tkind for m in models]), np.stack([m._ta for m in models]),
                      np.stack([m._tb for m in models]), np.stack([m._trainable for m in models]).astype(np.uint8),
                      max_iter=int(max_iter), priors=pri, **family, **kw)
        for e, m in enumerate(models):
            st = int(res.status[e])
            m.opt_info = {"status": st, "n_evals": int(res.n_evals[e]), "n_iters": int(res.n_iters[e]),
                          "objective": float(res.objective[e])}
            if st != _lib.STATUS_NOT_PD:
                m._theta = np.array(res.theta[e], dtype=np.float64)
            success[e] = st == _lib.STATUS_OK
        theta = np.stack([m._theta for m in models])
    preds = None
    if Xs_list is not None:
        if fourier:
            out = eng.vff_predict(batch, vff, theta)
            objective = np.asarray(out["elbo"])
        elif sparse:
            out = eng.sgpr_predict(batch, inducing, theta)
            objective = np.asarray(out["elbo"])           # the sparse model reports +ELBO
        else:
            out = eng.predict(batch, theta)
            objective = -np.asarray(out["lml"])
        preds = []
        for e, m in enumerate(models):
            sl = slice(int(batch.off_m[e]), int(batch.off_m[e + 1]))
            preds.append({"f*": out["f*"][sl].copy(), "f*_var": out["f*_var"][sl].copy(),
                          "y_var": out["y_var"][sl].copy(), "f_bar": m._f_bar(sl.stop - sl.start)})
    else:
        if fourier:
            objective = np.asarray(eng.vff_elbo(batch, vff, theta)[0])
        elif sparse:
            objective = np.asarray(eng.sgpr_elbo(batch, inducing, theta)[0])
        else:
            lml, _, _ = eng.lml_grad(batch, theta, with_grad=False)
            objective = -np.asarray(lml)
    return success, objective, preds
