"""hddm_wiener_like_batch_dev_static: the allocation-free variant of hddm_wiener_like_batch_dev (scratch in a
static __device__ array of the library; include/hddm_b200.h).  It has no counterpart in the reference, so the
checker is the existing entry point, on the same device buffers, within the summed-log-likelihood tolerance
of the other GPU tests (tests/util.py RTOL_SUM), on two calls in a row with different shapes -- the second
call finds the first one's partial sums in the array and must not be affected by them.

Collected last on purpose: a failure here hides no other test."""
import ctypes as C

import numpy as np
import pytest

from tests.util import RTOL_SUM, assert_parity, draw_rts, sweep_inputs

STATIC_DOUBLES = 1 << 18  # kStaticPartialDoubles, hddm_b200/csrc/kernels.cuh


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def _both(L, torch, x, seg, max_seg, params, numerics=(1e-4, 2, 2, 1, 1e-3)):
    """(static entry point, existing entry point) -> logp[P] and pdf_sv counts, buffers made up front."""
    from hddm_b200 import _lib
    dev = x.device
    P = int(params.shape[1])
    n = int(x.numel())
    need = int(L.hddm_wiener_like_batch_scratch(max_seg if seg is not None else n, P))
    scratch = torch.empty(max(need, 1), dtype=torch.float64, device=dev)
    out_s = torch.full((P,), 7.0, dtype=torch.float64, device=dev)
    out_d = torch.full((P,), 9.0, dtype=torch.float64, device=dev)
    ev_s = torch.zeros(1, dtype=torch.int64, device=dev)
    ev_d = torch.zeros(1, dtype=torch.int64, device=dev)
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    torch.cuda.synchronize(dev)
    free0 = torch.cuda.mem_get_info(dev)[0]
    _lib.check(L.hddm_wiener_like_batch_dev_static(_ptr(x), n, _ptr(seg), max_seg, _ptr(params), P, *numerics,
                                                   _ptr(out_s), _ptr(ev_s), stream))
    torch.cuda.synchronize(dev)
    free1 = torch.cuda.mem_get_info(dev)[0]
    _lib.check(L.hddm_wiener_like_batch_dev(_ptr(x), n, _ptr(seg), max_seg, _ptr(params), P, *numerics,
                                            _ptr(out_d), _ptr(scratch), _ptr(ev_d), stream))
    torch.cuda.synchronize(dev)
    return out_s.cpu().numpy(), out_d.cpu().numpy(), int(ev_s.item()), int(ev_d.item()), free0 - free1


@pytest.mark.gpu
def test_static_scratch_entry_point_matches_the_allocating_one(gpu):
    import torch
    from hddm_b200 import _lib
    L = _lib.lib()
    _lib.check(L.hddm_b200_init(0))          # set-up: the static entry point itself binds nothing
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)

    # warm-up call (module load, symbol address, exponential table): outside the no-new-memory check below
    xw, pw = sweep_inputs(4096, 8, onset=0.70, seed=5)
    _both(L, torch, torch.from_numpy(xw).to(dev), None, 0, torch.from_numpy(pw).to(dev))

    # call 1: one trial vector shared by 96 sets, 20000 trials -> 10 blocks of 2048 per set in the array
    x1, p1 = sweep_inputs(20000, 96, onset=0.70, seed=11)
    s1, d1, es1, ed1, grew1 = _both(L, torch, torch.from_numpy(x1).to(dev), None, 0, torch.from_numpy(p1).to(dev))
    assert np.isfinite(d1).all()
    assert_parity(s1, d1, RTOL_SUM, what="static vs allocating entry point, shared trials")
    assert es1 == ed1 and es1 > 0, (es1, ed1)   # same pdf_sv evaluations: same decisions
    assert grew1 <= 0, "device memory in use grew by %d bytes across the static call" % grew1

    # call 2, straight after, other shapes: ragged segments (1 .. 3 blocks), onset-mixed RTs, half of the sets
    # without outlier mass (log-likelihood -inf where a trial lies before the onset: exact class)
    rng = np.random.default_rng(12)
    P2 = 300
    lens = rng.integers(50, 6000, P2)
    lens[:4] = (1, 2048, 2049, 5999)
    seg = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    x2 = draw_rts(rng, int(seg[-1]), 0.30)
    _, p2 = sweep_inputs(8, P2, seed=13)
    p2[7, ::2] = 0.0
    s2, d2, es2, ed2, grew2 = _both(L, torch, torch.from_numpy(x2).to(dev), torch.from_numpy(seg).to(dev),
                                    int(lens.max()), torch.from_numpy(np.ascontiguousarray(p2)).to(dev))
    assert np.isneginf(d2).any() and np.isfinite(d2).any()
    assert_parity(s2, d2, RTOL_SUM, what="static vs allocating entry point, ragged segments")
    assert es2 == ed2 and es2 > 0, (es2, ed2)
    assert grew2 <= 0, "device memory in use grew by %d bytes across the static call" % grew2

    # call 1 again after call 2: the same numbers as before, bit for bit (nothing of call 2 leaks in)
    s1b, _, _, _, _ = _both(L, torch, torch.from_numpy(x1).to(dev), None, 0, torch.from_numpy(p1).to(dev))
    assert np.array_equal(s1b, s1)


@pytest.mark.gpu
def test_static_scratch_entry_point_rejects_what_does_not_fit(gpu):
    import torch
    from hddm_b200 import _lib
    L = _lib.lib()
    _lib.check(L.hddm_b200_init(0))
    dev = torch.device("cuda", 0)
    P, n = 1000, 2048 * 263                  # 263000 partial sums > 262144
    assert P * 263 > STATIC_DOUBLES
    x, p = sweep_inputs(n, P, seed=3)
    xd, pd = torch.from_numpy(x).to(dev), torch.from_numpy(p).to(dev)
    out = torch.full((P,), 5.0, dtype=torch.float64, device=dev)
    rc = L.hddm_wiener_like_batch_dev_static(_ptr(xd), n, None, 0, _ptr(pd), P, 1e-4, 2, 2, 1, 1e-3, _ptr(out), None,
                                             C.c_void_p(torch.cuda.current_stream(dev).cuda_stream))
    assert rc == _lib.ERR_ARG
    msg = L.hddm_b200_last_error().decode()
    assert "263000" in msg and str(STATIC_DOUBLES) in msg, msg
    torch.cuda.synchronize(dev)
    assert (out.cpu().numpy() == 5.0).all()  # refused before anything was launched
    # just below the limit it runs: 1000 sets x 262 blocks = 262000 partial sums
    n = 2048 * 262
    _lib.check(L.hddm_wiener_like_batch_dev_static(_ptr(xd), n, None, 0, _ptr(pd), P, 1e-4, 2, 2, 1, 1e-3, _ptr(out),
                                                   None, C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)))
    torch.cuda.synchronize(dev)
    assert np.isfinite(out.cpu().numpy()).all()
