"""Loader for the compiled UNMODIFIED reference (oracle/_ref/, built by oracle/build_ref.py).

Test infrastructure only: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  Never imported by hddm_b200/.

Harness-side shims (reference untouched): scipy >= 1.14 removed `cumtrapz` / `simps`,
which src/wfpt.pyx:409 and hddm/tests/test_likelihoods.py:176 still call.
"""
import glob
import importlib.util
import os

_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")


def available():
    return bool(glob.glob(os.path.join(_DIR, "wfpt*.so")) and
                glob.glob(os.path.join(_DIR, "cdfdif_wrapper*.so")))


def _shim_scipy():
    import scipy.integrate as si
    if not hasattr(si, "cumtrapz"):
        si.cumtrapz = si.cumulative_trapezoid
    if not hasattr(si, "simps"):
        si.simps = si.simpson


def _load(name):
    so = sorted(glob.glob(os.path.join(_DIR, name + "*.so")))
    if not so:
        raise ImportError("oracle/_ref/%s*.so missing: run `python oracle/build_ref.py` where "
                          "/root/reference exists" % name)
    spec = importlib.util.spec_from_file_location(name, so[0])
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


_cache = {}


def wfpt():
    """The reference `wfpt` extension module (src/wfpt.pyx)."""
    if "wfpt" not in _cache:
        _shim_scipy()
        _cache["wfpt"] = _load("wfpt")
    return _cache["wfpt"]


def cdfdif_wrapper():
    """The reference `cdfdif_wrapper` extension module (src/cdfdif_wrapper.pyx + src/cdfdif.c)."""
    if "cdf" not in _cache:
        _cache["cdf"] = _load("cdfdif_wrapper")
    return _cache["cdf"]
