"""The C-ABI library loads and exports every symbol include/gpsat_b200.h declares (no GPU calls)."""
import ctypes as C
import os
import re

from gpsat_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "gpsat_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gpsat_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported(built):
    lib = _lib.load()
    names = _declared()
    assert len(names) >= 9
    for n in names:
        assert hasattr(lib, n), f"{n} declared in the header but not exported"
    assert set(_lib.EXPORTED_SYMBOLS) <= set(names)


def test_version_and_defaults(built):
    lib = _lib.load()
    assert lib.gpsat_version() >= 100
    o = _lib.FitOptions()
    lib.gpsat_fit_options_default(C.byref(o))
    # SciPy L-BFGS-B defaults + the reference's max_iter (gpflow_models.py:291)
    assert (o.pgtol, o.ftol, o.maxiter, o.maxfun, o.maxls) == (1e-5, 2.220446049250313e-09, 10000, 15000, 20)


def test_struct_layouts_match_header():
    assert C.sizeof(_lib.Batch) == 56
    assert C.sizeof(_lib.FitOptions) == 32
    assert C.sizeof(_lib.Stats) == 144


def test_no_cpu_fallback(built):
    """Without a GPU the engine must refuse to start rather than compute on the host."""
    import torch
    if torch.cuda.is_available():
        return
    from gpsat_b200.engine import Engine
    try:
        Engine(0)
    except RuntimeError as e:
        assert "libgpsat_b200" in str(e)
    else:
        raise AssertionError("Engine(0) succeeded without a GPU")


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "gpsat_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(".py"):
                txt = open(os.path.join(dp, f)).read()
                assert "import oracle" not in txt and "from oracle" not in txt, f
