"""CPU-side checks of the drop-in boundary: the C-ABI library loads and exports every symbol
include/hddm_b200.h declares, the Python mirror of the reference's module surface keeps the
reference's signatures and argument errors, the host-built series-selection table reproduces the
oracle's (= the reference's) branch/K decisions, and the sharding logic is correct over gloo with
world_size 2.  No compute call is made: there is no GPU in this container."""
import ctypes as C
import inspect
import os
import re
import socket
import sys

import numpy as np
import pytest

from hddm_b200 import _lib, batch, cdfdif_wrapper, wfpt
from oracle import oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    text = open(os.path.join(ROOT, "include", "hddm_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(hddm_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    L = _lib.lib()
    declared = _header_symbols()
    assert len(declared) >= 25
    for name in declared:
        assert hasattr(L, name), "libhddm_b200.so does not export %s" % name
        assert name in _lib.SIGNATURES, "%s declared in the header but not bound in _lib.SIGNATURES" % name
    assert sorted(_lib.SIGNATURES) == declared


def test_library_has_no_oracle_or_torch_dependency():
    import subprocess
    out = subprocess.run(["ldd", _lib.library_path()], stdout=subprocess.PIPE, text=True).stdout
    assert "oracle" not in out and "torch" not in out and "python" not in out, out


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "hddm_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", src, flags=re.M), f
                assert "liboracle" not in src and not re.search(r"\bwo_[a-z]", src), f


# reference signatures: src/wfpt.pyx:32,54,244,357,389,323,417; src/pdf.pxi:104; src/cdfdif_wrapper.pyx:16
REFERENCE_SIGNATURES = {
    wfpt.full_pdf: "(x, v, sv, a, z, sz, t, st, err, n_st=2, n_sz=2, use_adaptive=1, simps_err=0.001)",
    wfpt.pdf_array: "(x, v, sv, a, z, sz, t, st, err=0.0001, logp=0, n_st=2, n_sz=2, use_adaptive=1, "
                    "simps_err=0.001, p_outlier=0, w_outlier=0)",
    wfpt.wiener_like: "(x, v, sv, a, z, sz, t, st, err, n_st=10, n_sz=10, use_adaptive=1, simps_err=1e-08, "
                      "p_outlier=0, w_outlier=0.1)",
    wfpt.wiener_like_multi: "(x, v, sv, a, z, sz, t, st, err, multi=None, n_st=10, n_sz=10, use_adaptive=1, "
                            "simps_err=0.001, p_outlier=0, w_outlier=0)",
    wfpt.wiener_like_contaminant: "(x, cont_x, v, sv, a, z, sz, t, st, t_min, t_max, err, n_st=10, n_sz=10, "
                                  "use_adaptive=1, simps_err=1e-08)",
    wfpt.gen_cdf_using_pdf: "(v, sv, a, z, sz, t, st, err, N=500, time=5.0, n_st=2, n_sz=2, use_adaptive=1, "
                            "simps_err=0.001, p_outlier=0, w_outlier=0)",
    wfpt.split_cdf: "(x, data)",
    cdfdif_wrapper.dmat_cdf_array: "(x, v, sv, a, z, sz, t, st, p_outlier, w_outlier)",
}


def test_module_surface_matches_reference_signatures():
    for fn, want in REFERENCE_SIGNATURES.items():
        assert str(inspect.signature(fn)) == want, fn.__name__
    # gen_rts_from_cdf: the reference's arguments first, then the two NEW optional uniform buffers
    assert str(inspect.signature(wfpt.gen_rts_from_cdf)).startswith(
        "(v, sv, a, z, sz, t, st, samples=1000, cdf_lb=-6, cdf_ub=6, dt=0.01")
    import hddm_b200
    assert hddm_b200.cdfdif is cdfdif_wrapper  # hddm/__init__.py:16-21
    # the reference installs TOP-LEVEL modules `wfpt` and `cdfdif_wrapper` (setup.py:6-7, imported bare at
    # hddm/__init__.py:11-21): the repo root carries shims of those names
    import importlib
    top_w, top_c = importlib.import_module("wfpt"), importlib.import_module("cdfdif_wrapper")
    for fn in REFERENCE_SIGNATURES:
        owner = top_c if fn.__name__ == "dmat_cdf_array" else top_w
        assert getattr(owner, fn.__name__) is fn
    assert top_w.gen_rts_from_cdf is wfpt.gen_rts_from_cdf and top_w.wiener_like_rlddm is wfpt.wiener_like_rlddm


def test_debug_knobs_are_gated_by_the_environment(monkeypatch):
    """hddm_b200_set_kernel / _set_sort act on every caller in the process: refused unless HDDM_B200_DEBUG_KNOBS=1
    (tests/conftest.py sets it for the A/B parity tests)."""
    L = _lib.lib()
    monkeypatch.delenv("HDDM_B200_DEBUG_KNOBS", raising=False)
    assert L.hddm_b200_set_kernel(0) == _lib.ERR_ARG and b"HDDM_B200_DEBUG_KNOBS" in L.hddm_b200_last_error()
    assert L.hddm_b200_set_sort(1) == _lib.ERR_ARG
    monkeypatch.setenv("HDDM_B200_DEBUG_KNOBS", "1")
    assert L.hddm_b200_set_kernel(0) == _lib.OK and L.hddm_b200_set_sort(1) == _lib.OK
    assert L.hddm_b200_set_kernel(9) == _lib.ERR_ARG


def test_buffer_contract_errors_match_cython():
    # SURVEY.md 8(b) probes of np.ndarray[double, ndim=1]
    with pytest.raises(TypeError):
        wfpt.wiener_like([0.5, 0.6], 1, 0, 2, .5, 0, .3, 0, 1e-4)
    with pytest.raises(ValueError, match="dtype mismatch"):
        wfpt.wiener_like(np.array([0.5, 0.6], dtype=np.float32), 1, 0, 2, .5, 0, .3, 0, 1e-4)
    with pytest.raises(ValueError, match="dimensions"):
        wfpt.pdf_array(np.zeros((2, 2)), 1, 0, 2, .5, 0, .3, 0)
    with pytest.raises(ValueError, match="dtype mismatch"):
        wfpt.wiener_like_contaminant(np.array([0.5]), np.array([0], dtype=np.int64), 1, 0, 2, .5, 0, .3, 0,
                                     0.1, 5., 1e-4)
    with pytest.raises(ZeroDivisionError):
        cdfdif_wrapper.dmat_cdf_array(np.array([0.5]), 1, 0, 2, .5, 0, .3, 0, 0.05, 0)


def test_no_gpu_means_loud_failure_not_fallback():
    n = C.c_int(0)
    rc = _lib.lib().hddm_b200_device_count(C.byref(n))
    if rc == 0 and n.value > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(RuntimeError, match="hddm_b200"):
        wfpt.wiener_like(np.array([0.5, 0.6]), 1, 0, 2, .5, 0, .3, 0, 1e-4)
    with pytest.raises(RuntimeError, match="hddm_b200"):
        wfpt.wiener_like_batch(np.array([0.5, 0.6]), np.ones((9, 3)))
    with pytest.raises(RuntimeError):
        batch.DeviceSweep(np.array([0.5, 0.6]))


def _table(err):
    thr = np.zeros(31)
    code = np.zeros(32, dtype=np.int32)
    n = C.c_int()
    _lib.check(_lib.lib().hddm_series_table(err, _lib.dptr(thr), code.ctypes.data_as(_lib.c_ip), C.byref(n)))
    return thr[:max(n.value, 0)].copy(), code[:n.value + 1].copy(), n.value


def _lookup(thr, code, tt):
    return code[np.searchsorted(thr, tt, side="right")]


@pytest.mark.parametrize("err", [1e-4, 1e-3, 1e-10, 1e-8, 1e-2, 0.3, 1e-15, 3.3e-6])
def test_series_table_reproduces_reference_decisions(err):
    """src/pdf.pxi:36-54,60: (branch, K) from the table == the formulas, on random normalised times
    and on the adjacent doubles either side of every break point."""
    thr, code, n = _table(err)
    assert n >= 2 and np.all(np.diff(thr) > 0)
    rng = np.random.default_rng(5)
    tts = np.concatenate([10.0 ** rng.uniform(-6, 3, 20000), rng.uniform(1e-3, 8, 20000),
                          thr, np.nextafter(thr, 0), np.nextafter(thr, np.inf),
                          [5e-324, 1e-300, 1e300, 1.7e308]])
    got = _lookup(thr, code, tts)
    want = np.array([O.series_code(float(t), err) for t in tts])
    assert np.array_equal(got, want), np.where(got != want)[0][:10]


def test_series_table_survey_breakpoints():
    # SURVEY.md appendix B, err=1e-4: small-time 3 terms below 0.0555, 4 to 0.154; large-time 4 to
    # 0.217, 3 to 0.451, 2 to 1.548, 1 above
    thr, code, n = _table(1e-4)
    assert list(code[1:]) == [3, 4, -4, -3, -2, -1]
    np.testing.assert_allclose(thr[1:], [0.0555, 0.154, 0.217, 0.451, 1.548], rtol=6e-3)


def test_shard_bounds_partition():
    for P in (0, 1, 7, 8, 10000, 10001):
        for world in (1, 2, 3, 8):
            spans = [batch.shard_bounds(P, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == P
            assert all(spans[r][1] == spans[r + 1][0] for r in range(world - 1))
            sizes = batch.shard_sizes(P, world)
            assert max(sizes) - min(sizes) <= 1 and sum(sizes) == P


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _gloo_worker(rank, world, port, P, q):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sys.path.insert(0, ROOT)
    from tests.util import sweep_inputs
    x, params = sweep_inputs(64, P)

    def evaluate(xs, block):  # stand-in for the CUDA launch: the CPU oracle, in the tests only
        return torch.from_numpy(O.wiener_like_batch(xs, block, **batch.HDDM_DEFAULTS))

    full = batch.wiener_like_batch_sharded(x, params, evaluate=evaluate)
    q.put((rank, full.numpy().copy()))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("P", [7, 1, 8])  # ragged shards, an empty shard, equal shards (the no-pad path)
def test_sharded_batch_over_gloo_world2(P):
    """N>1 host path: sets split over 2 ranks, all_gather, every rank sees the single-process vector
    bit for bit (each per-set sum is computed on one rank only)."""
    import torch.multiprocessing as mp
    from tests.util import sweep_inputs
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, P, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    x, params = sweep_inputs(64, P)
    want = O.wiener_like_batch(x, params, **batch.HDDM_DEFAULTS)
    for r in range(2):
        assert got[r].shape == (P,)
        assert np.array_equal(got[r], want)


# ---- host logic of the plugin boundary that needs no GPU (the C call is replaced by a recorder) -------------

def test_regression_closure_reindexes_like_the_reference(monkeypatch):
    """hddm/models/hddm_regression.py:26-36: every regression outcome is a Series over the WHOLE data set and is
    re-indexed by the node's own trials (`.loc[value['rt'].index].values`); p_outlier defaults to 0.05 and
    w_outlier comes from wiener_params.  A full-length plain vector must be refused, not truncated."""
    import pandas as pd
    from hddm_b200 import likelihoods as L
    seen = {}

    def fake_multi(x, v, sv, a, z, sz, t, st, err, multi=None, n_st=10, n_sz=10, use_adaptive=1, simps_err=1e-3,
                   p_outlier=0, w_outlier=0):
        seen.update(x=x, v=v, a=a, err=err, multi=multi, p_outlier=p_outlier, w_outlier=w_outlier)
        return -1.0

    monkeypatch.setattr(L.wfpt, "wiener_like_multi", fake_multi)
    rng = np.random.default_rng(3)
    data = pd.DataFrame({"rt": rng.uniform(0.4, 2.0, 60) * np.where(rng.random(60) < .7, 1, -1), "subj": np.arange(60) % 3})
    v_all = pd.Series(rng.normal(1.0, 0.3, 60), index=data.index)
    a_all = pd.Series(rng.uniform(1.0, 2.0, 60), index=data.index)
    node = data[data.subj == 1]                     # second subject: rows 1, 4, 7, ...
    like = L.generate_wfpt_reg_stochastic_class(wiener_params=dict(L.DEFAULT_WIENER_PARAMS, w_outlier=0.2))
    inst = like(value=node, v=v_all, sv=0.1, a=a_all, z=0.5, sz=0.1, t=0.3, st=0.1, reg_outcomes=["v", "a"])
    assert inst.logp == -1.0
    assert np.array_equal(seen["x"], node["rt"].values)
    assert np.array_equal(seen["v"], v_all.values[1::3]) and np.array_equal(seen["a"], a_all.values[1::3])
    assert seen["p_outlier"] == 0.05 and seen["w_outlier"] == 0.2 and seen["err"] == 1e-4
    assert list(seen["multi"]) == ["v", "a"]
    with pytest.raises(ValueError):                 # un-indexed full-length vector: cannot be aligned
        L.wiener_multi_like(node, v_all.values, 0.1, 1.5, 0.5, 0.1, 0.3, 0.1, ["v"])
    assert L.wiener_multi_like(node["rt"].values, v_all.values[1::3], 0.1, 1.5, 0.5, 0.1, 0.3, 0.1, ["v"]) == -1.0


def test_trial_cache_identity_content_and_lru(monkeypatch):
    import gc
    import pandas as pd
    from hddm_b200 import likelihoods as L

    class FakeTrials(object):
        made, closed = 0, 0

        def __init__(self, x):
            FakeTrials.made += 1
            self.x = x.copy()

        def close(self):
            FakeTrials.closed += 1

    monkeypatch.setattr(L.wfpt, "Trials", FakeTrials)
    cache = L.DeviceTrialCache(max_entries=3)
    df = pd.DataFrame({"rt": np.random.default_rng(1).random(100)})
    handles = [cache.get(L._rt_array(df)) for _ in range(4)]
    assert FakeTrials.made == 1 and all(h is handles[0] for h in handles)
    a = np.random.default_rng(2).random(50)
    h1 = cache.get(a)
    a[0] = 5.0                                       # writeable data changed in place: must not hit
    h2 = cache.get(a)
    assert h1 is not h2 and h2.x[0] == 5.0
    assert cache.get(a.copy()) is h2                 # same content, other array: hit by content
    for k in range(4):
        cache.get(np.full(7, float(k)))
    assert len(cache._handles) == 3 and FakeTrials.closed == FakeTrials.made - 3   # evicted one at a time
    del df, handles
    gc.collect()
    assert not cache._by_id                          # identity entries die with the memory they describe
