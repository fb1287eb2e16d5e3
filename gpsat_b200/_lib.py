"""ctypes binding of ``libgpsat_b200.so`` (the C ABI declared in ``include/gpsat_b200.h``).

There is no CPU fallback: if the shared library is missing, or no B200 is visible when an engine
is created, an exception is raised.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# GPSAT_B200_LIB: an instrumented development build (tools/step_phases.py); the default is the in-tree library
LIB_PATH = os.environ.get("GPSAT_B200_LIB") or os.path.join(_HERE, "libgpsat_b200.so")

MEM_HOST = 0
MEM_DEVICE = 1

KERNEL_IDS = {"Matern12": 0, "Exponential": 0, "Matern32": 1, "Matern52": 2, "RBF": 3,
              "SquaredExponential": 3}
TRANSFORM_SOFTPLUS = 0
TRANSFORM_SIGMOID = 1

STATUS_OK, STATUS_NOT_PD, STATUS_LINESEARCH, STATUS_MAXITER, STATUS_BAD_INPUT = range(5)

EXPORTED_SYMBOLS = (
    "gpsat_version", "gpsat_last_error", "gpsat_device_name", "gpsat_engine_create", "gpsat_engine_destroy",
    "gpsat_engine_stats", "gpsat_engine_stream", "gpsat_engine_set_profile", "gpsat_engine_set_lauum_path", "gpsat_engine_set_step_path",
    "gpsat_engine_i8_timeouts", "gpsat_engine_debug_set_i8_spin_limit", "gpsat_engine_set_precision_rule",
    "gpsat_engine_precision_rule_default", "gpsat_engine_set_precision_low_tier", "gpsat_engine_precision_low_tier_default", "gpsat_engine_i8_phases", "gpsat_engine_step_phases", "gpsat_fit_options_default",
    "gpsat_gpr_lml_grad", "gpsat_gpr_fit", "gpsat_gpr_predict", "gpsat_gpr_predict_cov", "gpsat_gpr_factor_debug",
    "gpsat_sgpr_elbo", "gpsat_sgpr_elbo_gradz", "gpsat_sgpr_predict", "gpsat_sgpr_predict_cov", "gpsat_sgpr_fit",
    "gpsat_vff_elbo", "gpsat_vff_predict", "gpsat_vff_fit",
    "gpsat_gaussian_smooth", "gpsat_glue_predictions", "gpsat_local_select", "gpsat_local_select_count_noalloc",
)
NOALLOC_MAX_CHUNK_COUNTS = 1 << 20


class Batch(C.Structure):
    _fields_ = [("kernel", C.c_int32), ("D", C.c_int32), ("E", C.c_int64),
                ("off_n", C.c_void_p), ("off_m", C.c_void_p),
                ("X", C.c_void_p), ("y", C.c_void_p), ("Xs", C.c_void_p)]


class FitOptions(C.Structure):
    _fields_ = [("pgtol", C.c_double), ("ftol", C.c_double), ("maxiter", C.c_int32),
                ("maxfun", C.c_int32), ("maxls", C.c_int32), ("max_slots", C.c_int32)]


class Priors(C.Structure):
    _fields_ = [("kind", C.c_void_p), ("a", C.c_void_p), ("b", C.c_void_p)]


class VffDomain(C.Structure):
    _fields_ = [("a", C.c_void_p), ("b", C.c_void_p), ("nfreq", C.c_void_p)]


class SelectCond(C.Structure):
    _fields_ = [("kind", C.c_int32), ("op", C.c_int32), ("ncol", C.c_int32), ("col", C.c_int32 * 3),
                ("loc_col", C.c_int32 * 3), ("val", C.c_double)]


class Stats(C.Structure):
    _fields_ = [("kernel_launches", C.c_int64), ("ticks", C.c_int64), ("evals", C.c_int64),
                ("slots", C.c_int64), ("h2d_bytes", C.c_double), ("d2h_bytes", C.c_double),
                ("gpu_ms", C.c_double), ("sweep_flops", C.c_double),
                ("ms_diag", C.c_double), ("ms_step", C.c_double), ("ms_lauum", C.c_double),
                ("ms_opt", C.c_double), ("ms_reduce", C.c_double),
                ("n_diag", C.c_int64), ("n_step", C.c_int64), ("n_lauum", C.c_int64),
                ("n_opt", C.c_int64), ("n_reduce", C.c_int64)]


_lib = None


def load():
    """Load the shared library (once).  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). gpsat_b200 has no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    vp = C.c_void_p
    lib.gpsat_version.restype = C.c_int
    lib.gpsat_last_error.restype = C.c_char_p
    lib.gpsat_device_name.argtypes = [C.c_int, C.c_char_p, C.c_size_t]
    lib.gpsat_engine_create.argtypes = [C.c_int, C.POINTER(vp)]
    lib.gpsat_engine_destroy.argtypes = [vp]
    lib.gpsat_engine_stats.argtypes = [vp, C.POINTER(Stats)]
    lib.gpsat_engine_stream.argtypes = [vp]
    lib.gpsat_engine_stream.restype = vp
    lib.gpsat_engine_set_profile.argtypes = [vp, C.c_int]
    lib.gpsat_engine_set_lauum_path.argtypes = [vp, C.c_int]
    lib.gpsat_engine_set_step_path.argtypes = [vp, C.c_int]
    lib.gpsat_engine_i8_timeouts.argtypes = [vp, C.POINTER(C.c_int)]
    lib.gpsat_engine_debug_set_i8_spin_limit.argtypes = [vp, C.c_int]
    lib.gpsat_engine_set_precision_rule.argtypes = [vp, C.c_double]
    lib.gpsat_engine_precision_rule_default.argtypes = []
    lib.gpsat_engine_precision_rule_default.restype = C.c_double
    lib.gpsat_engine_set_precision_low_tier.argtypes = [vp, C.c_double]
    lib.gpsat_engine_precision_low_tier_default.argtypes = []
    lib.gpsat_engine_precision_low_tier_default.restype = C.c_double
    lib.gpsat_engine_i8_phases.argtypes = [vp, C.POINTER(C.c_uint64), C.c_int]
    lib.gpsat_engine_step_phases.argtypes = [vp, C.POINTER(C.c_uint64), C.c_int]
    lib.gpsat_fit_options_default.argtypes = [C.POINTER(FitOptions)]
    lib.gpsat_fit_options_default.restype = None
    lib.gpsat_gpr_lml_grad.argtypes = [vp, C.POINTER(Batch), C.c_int, vp, vp, vp, vp]
    lib.gpsat_gpr_fit.argtypes = [vp, C.POINTER(Batch), C.c_int, vp, vp, vp, vp, vp,
                                  C.POINTER(FitOptions), C.POINTER(Priors), vp, vp, vp, vp, vp]
    lib.gpsat_gpr_predict.argtypes = [vp, C.POINTER(Batch), C.c_int, vp, vp, vp, vp, vp, vp]
    lib.gpsat_gpr_predict_cov.argtypes = [vp, C.POINTER(Batch), C.c_int, vp, vp, vp, vp, vp, vp, vp]
    lib.gpsat_gpr_factor_debug.argtypes = [vp, C.POINTER(Batch), vp, vp, vp]
    lib.gpsat_sgpr_elbo.argtypes = [vp, C.POINTER(Batch), vp, vp, C.c_int, vp, vp, vp, vp]
    lib.gpsat_sgpr_elbo_gradz.argtypes = [vp, C.POINTER(Batch), vp, vp, C.c_int, vp, vp, vp, vp, vp]
    lib.gpsat_sgpr_fit.argtypes = [vp, C.POINTER(Batch), vp, vp, C.c_int, vp, vp, vp, vp, vp,
                                   C.POINTER(FitOptions), C.POINTER(Priors), vp, vp, vp, vp, vp]
    lib.gpsat_sgpr_predict.argtypes = [vp, C.POINTER(Batch), vp, vp, C.c_int, vp, vp, vp, vp, vp, vp]
    lib.gpsat_sgpr_predict_cov.argtypes = [vp, C.POINTER(Batch), vp, vp, C.c_int, vp, vp, vp, vp, vp, vp, vp]
    lib.gpsat_vff_elbo.argtypes = [vp, C.POINTER(Batch), C.POINTER(VffDomain), C.c_int, vp, vp, vp, vp]
    lib.gpsat_vff_predict.argtypes = [vp, C.POINTER(Batch), C.POINTER(VffDomain), C.c_int, vp, vp, vp, vp, vp, vp]
    lib.gpsat_vff_fit.argtypes = [vp, C.POINTER(Batch), C.POINTER(VffDomain), C.c_int, vp, vp, vp, vp, vp,
                                  C.POINTER(FitOptions), C.POINTER(Priors), vp, vp, vp, vp, vp]
    lib.gpsat_gaussian_smooth.argtypes = [vp, C.c_int, C.c_int64, vp, vp, C.c_int64, C.c_int64, vp, vp,
                                          C.c_double, C.c_double, vp, vp]
    lib.gpsat_glue_predictions.argtypes = [vp, C.c_int, C.c_int64, C.c_int64, vp, C.c_int, vp, vp, vp, C.c_int64, vp, vp]
    lib.gpsat_local_select.argtypes = [vp, C.c_int64, C.c_int32, vp, C.c_int64, C.c_int32, vp, C.c_int32,
                                       C.POINTER(SelectCond), vp, vp, vp]
    lib.gpsat_local_select_count_noalloc.argtypes = [vp, C.c_int64, C.c_int32, vp, C.c_int64, C.c_int32, vp, C.c_int32,
                                                     C.POINTER(SelectCond), vp]
    _lib = lib
    return lib


def device_name(device: int = 0) -> str:
    """Name of CUDA device `device` (raises if the library or the device is missing)."""
    buf = C.create_string_buffer(256)
    check(load().gpsat_device_name(int(device), buf, 256))
    return buf.value.decode()


def last_error() -> str:
    return load().gpsat_last_error().decode()


def check(rc: int):
    if rc != 0:
        raise RuntimeError(f"libgpsat_b200 call failed (rc={rc}): {last_error()}")
