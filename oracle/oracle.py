"""ctypes front end of oracle/liboracle.so (wfpt_oracle.c + cdfdif_oracle.c).

TEST INFRASTRUCTURE ONLY.  Mirrors the reference's module-function names
(src/wfpt.pyx, src/cdfdif_wrapper.pyx) so parity tests read like the reference's own tests.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liboracle.so")
_lib = None

c_dp = C.POINTER(C.c_double)
c_lp = C.POINTER(C.c_int64)


def build(force=False):
    srcs = [os.path.join(_HERE, f) for f in ("wfpt_oracle.c", "cdfdif_oracle.c")]
    if (not force and os.path.exists(_SO) and
            os.path.getmtime(_SO) >= max(os.path.getmtime(s) for s in srcs)):
        return _SO
    subprocess.check_call(["make", "-s", "-B", "-C", _HERE, "liboracle.so"])
    return _SO


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        L = C.CDLL(_SO)
        d, i, l = C.c_double, C.c_int, C.c_int64
        L.wo_ftt_01w.restype = d
        L.wo_ftt_01w.argtypes = [d, d, d]
        L.wo_series_code.restype = i
        L.wo_series_code.argtypes = [d, d]
        L.wo_prob_ub.restype = d
        L.wo_prob_ub.argtypes = [d, d, d]
        L.wo_pdf_sv.restype = d
        L.wo_pdf_sv.argtypes = [d] * 6
        L.wo_full_pdf.restype = d
        L.wo_full_pdf.argtypes = [d] * 9 + [i, i, i, d]
        L.wo_pdf_array.restype = None
        L.wo_pdf_array.argtypes = [c_dp, l] + [d] * 8 + [i, i, i, i, d, d, d, c_dp]
        L.wo_wiener_like.restype = d
        L.wo_wiener_like.argtypes = [c_dp, l] + [d] * 8 + [i, i, i, d, d, d]
        L.wo_wiener_like_multi.restype = d
        L.wo_wiener_like_multi.argtypes = [c_dp, l] + [c_dp, l] * 7 + [d, i, i, i, d, d, d]
        L.wo_wiener_like_contaminant.restype = d
        L.wo_wiener_like_contaminant.argtypes = ([c_dp, C.POINTER(C.c_int32), l] + [d] * 10 +
                                                 [i, i, i, d])
        L.wo_wiener_like_batch.restype = None
        L.wo_wiener_like_batch.argtypes = [c_dp, l, c_lp, c_dp, l, d, i, i, i, d, c_dp]
        L.wo_gen_cdf_using_pdf.restype = i
        L.wo_gen_cdf_using_pdf.argtypes = [c_dp, l] + [d] * 8 + [i, i, i, d, d, d, c_dp]
        L.wo_gen_rts_from_cdf.restype = None
        L.wo_gen_rts_from_cdf.argtypes = [c_dp, l] + [d] * 7 + [c_dp, c_dp, l, c_dp, c_dp]
        L.wo_wiener_like_rlddm.restype = d
        L.wo_wiener_like_rlddm.argtypes = [c_dp, c_lp, c_dp, c_lp, l] + [d] * 11 + [i, i, i, d, d, d]
        L.wo_wiener_like_rl.restype = d
        L.wo_wiener_like_rl.argtypes = [c_lp, c_dp, c_lp, l] + [d] * 7
        L.wo_wiener_like_multi_rlddm.restype = d
        L.wo_wiener_like_multi_rlddm.argtypes = [c_dp, c_lp, c_dp, c_lp, l, d, C.POINTER(c_dp), c_lp,
                                                 d, i, i, i, d, d, d]
        L.wo_cdfdif.restype = d
        L.wo_cdfdif.argtypes = [d, i, c_dp, c_dp]
        L.wo_dmat_cdf_array.restype = i
        L.wo_dmat_cdf_array.argtypes = [c_dp, l] + [d] * 9 + [c_dp]
        L.wo_pdfsv_calls.restype = C.c_longlong
        L.wo_pdfsv_calls_reset.restype = None
        L.wo_cdf_series_terms.restype = C.c_longlong
        L.wo_cdf_series_terms_reset.restype = None
        _lib = L
    return _lib


def _f64(x):
    x = np.asarray(x)
    if x.dtype != np.float64 or x.ndim != 1:
        raise ValueError("Buffer dtype mismatch / wrong number of dimensions")
    return np.ascontiguousarray(x)


def _p(a):
    return a.ctypes.data_as(c_dp)


def ftt_01w(tt, w, err):
    return lib().wo_ftt_01w(tt, w, err)


def series_code(tt, err):
    return lib().wo_series_code(tt, err)


def prob_ub(v, a, z):
    return lib().wo_prob_ub(v, a, z)


def pdf_sv(x, v, sv, a, z, err):
    return lib().wo_pdf_sv(x, v, sv, a, z, err)


def full_pdf(x, v, sv, a, z, sz, t, st, err, n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3):
    return lib().wo_full_pdf(x, v, sv, a, z, sz, t, st, err, n_st, n_sz, int(use_adaptive), simps_err)


def pdf_array(x, v, sv, a, z, sz, t, st, err=1e-4, logp=0, n_st=2, n_sz=2, use_adaptive=1,
              simps_err=1e-3, p_outlier=0, w_outlier=0):
    x = _f64(x)
    y = np.empty_like(x)
    lib().wo_pdf_array(_p(x), x.size, v, sv, a, z, sz, t, st, err, int(logp), n_st, n_sz,
                       int(use_adaptive), simps_err, p_outlier, w_outlier, _p(y))
    return y


def wiener_like(x, v, sv, a, z, sz, t, st, err, n_st=10, n_sz=10, use_adaptive=1, simps_err=1e-8,
                p_outlier=0, w_outlier=0.1):
    x = _f64(x)
    return lib().wo_wiener_like(_p(x), x.size, v, sv, a, z, sz, t, st, err, n_st, n_sz,
                                int(use_adaptive), simps_err, p_outlier, w_outlier)


_NAMES = ("v", "sv", "a", "z", "sz", "t", "st")


def wiener_like_multi(x, v, sv, a, z, sz, t, st, err, multi=None, n_st=10, n_sz=10, use_adaptive=1,
                      simps_err=1e-3, p_outlier=0, w_outlier=0):
    x = _f64(x)
    vals = dict(zip(_NAMES, (v, sv, a, z, sz, t, st)))
    multi = set(multi or ())
    args, keep = [], []
    for name in _NAMES:
        if name in multi:
            arr = np.ascontiguousarray(np.asarray(vals[name], dtype=np.float64))
            assert arr.shape == x.shape
            stride = 1
        else:
            arr = np.array([vals[name]], dtype=np.float64)
            stride = 0
        keep.append(arr)
        args += [_p(arr), stride]
    return lib().wo_wiener_like_multi(_p(x), x.size, *args, err, n_st, n_sz, int(use_adaptive),
                                      simps_err, p_outlier, w_outlier)


def _i64(a):
    a = np.asarray(a)
    if a.dtype.kind not in "iu" or a.ndim != 1:
        raise ValueError("Buffer dtype mismatch / wrong number of dimensions")
    return np.ascontiguousarray(a, dtype=np.int64)


def wiener_like_rlddm(x, response, feedback, split_by, q, alpha, pos_alpha, v, sv, a, z, sz, t, st, err,
                      n_st=10, n_sz=10, use_adaptive=1, simps_err=1e-8, p_outlier=0, w_outlier=0):
    x, feedback = _f64(x), _f64(feedback)
    response, split_by = _i64(response), _i64(split_by)
    return lib().wo_wiener_like_rlddm(_p(x), response.ctypes.data_as(c_lp), _p(feedback),
                                      split_by.ctypes.data_as(c_lp), x.size, q, alpha, pos_alpha, v, sv, a,
                                      z, sz, t, st, err, n_st, n_sz, int(use_adaptive), simps_err, p_outlier,
                                      w_outlier)


def wiener_like_rl(response, feedback, split_by, q, alpha, pos_alpha, v, z, err=1e-4, n_st=10, n_sz=10,
                   use_adaptive=1, simps_err=1e-8, p_outlier=0, w_outlier=0):
    feedback = _f64(feedback)
    response, split_by = _i64(response), _i64(split_by)
    return lib().wo_wiener_like_rl(response.ctypes.data_as(c_lp), _p(feedback), split_by.ctypes.data_as(c_lp),
                                   response.size, q, alpha, pos_alpha, v, z, p_outlier, w_outlier)


_RL_NAMES = _NAMES + ("alpha",)


def wiener_like_multi_rlddm(x, response, feedback, split_by, q, v, sv, a, z, sz, t, st, alpha, err,
                            multi=None, n_st=10, n_sz=10, use_adaptive=1, simps_err=1e-3, p_outlier=0,
                            w_outlier=0):
    x, feedback = _f64(x), _f64(feedback)
    response, split_by = _i64(response), _i64(split_by)
    vals = dict(zip(_RL_NAMES, (v, sv, a, z, sz, t, st, alpha)))
    multi = set(multi or ())
    keep, strides = [], []
    for name in _RL_NAMES:
        if name in multi:
            arr = np.ascontiguousarray(np.asarray(vals[name], dtype=np.float64))
            assert arr.shape == x.shape
            strides.append(1)
        else:
            arr = np.array([vals[name]], dtype=np.float64)
            strides.append(0)
        keep.append(arr)
    ptrs = (c_dp * 8)(*[_p(k) for k in keep])
    st_arr = np.asarray(strides, dtype=np.int64)
    return lib().wo_wiener_like_multi_rlddm(_p(x), response.ctypes.data_as(c_lp), _p(feedback),
                                            split_by.ctypes.data_as(c_lp), x.size, q, ptrs,
                                            st_arr.ctypes.data_as(c_lp), err, n_st, n_sz, int(use_adaptive),
                                            simps_err, p_outlier, w_outlier)


def wiener_like_contaminant(x, cont_x, v, sv, a, z, sz, t, st, t_min, t_max, err, n_st=10, n_sz=10,
                            use_adaptive=1, simps_err=1e-8):
    x = _f64(x)
    cont_x = np.ascontiguousarray(cont_x, dtype=np.int32)
    return lib().wo_wiener_like_contaminant(_p(x), cont_x.ctypes.data_as(C.POINTER(C.c_int32)),
                                            x.size, v, sv, a, z, sz, t, st, t_min, t_max, err,
                                            n_st, n_sz, int(use_adaptive), simps_err)


def wiener_like_batch(x, params, seg=None, err=1e-4, n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3):
    """params: float64 [9, P] rows v, sv, a, z, sz, t, st, p_outlier, w_outlier."""
    x = _f64(x)
    params = np.ascontiguousarray(params, dtype=np.float64)
    assert params.ndim == 2 and params.shape[0] == 9
    P = params.shape[1]
    out = np.empty(P)
    segp = None
    if seg is not None:
        seg = np.ascontiguousarray(seg, dtype=np.int64)
        assert seg.size == P + 1
        segp = seg.ctypes.data_as(c_lp)
    lib().wo_wiener_like_batch(_p(x), x.size, segp, _p(params), P, err, n_st, n_sz,
                               int(use_adaptive), simps_err, _p(out))
    return out


def gen_cdf_using_pdf(v, sv, a, z, sz, t, st, err, N=500, time=5., n_st=2, n_sz=2, use_adaptive=1,
                      simps_err=1e-3, p_outlier=0, w_outlier=0):
    x = np.linspace(-time, time, 2 * N + 1)
    cdf = np.empty_like(x)
    rc = lib().wo_gen_cdf_using_pdf(_p(x), x.size, v, sv, a, z, sz, t, st, err, n_st, n_sz,
                                    int(use_adaptive), simps_err, p_outlier, w_outlier, _p(cdf))
    if rc:
        raise ValueError("at least one of the parameters is out of the support")
    return x, cdf


def gen_rts_from_cdf(v, sv, a, z, sz, t, st, samples=1000, cdf_lb=-6, cdf_ub=6, dt=1e-2,
                     f=None, u_delay=None, return_table=False):
    """As src/wfpt.pyx:323-354 but the uniforms are passed in (default: numpy global RNG in the
    reference's draw order: f first, then the delay uniforms when st != 0)."""
    x = np.arange(cdf_lb, cdf_ub, dt)
    if f is None:
        f = np.random.rand(samples)
    if st != 0 and u_delay is None:
        u_delay = np.random.rand(samples)
    f = _f64(f)
    ud = _f64(u_delay) if st != 0 else np.zeros(1)
    table = np.empty_like(x)
    rts = np.empty(samples)
    lib().wo_gen_rts_from_cdf(_p(x), x.size, v, sv, a, z, sz, t, st, _p(f), _p(ud), samples,
                              _p(table), _p(rts))
    return (rts, x, table) if return_table else rts


def cdfdif(t, x, par):
    par = np.ascontiguousarray(par, dtype=np.float64)
    prob = C.c_double()
    F = lib().wo_cdfdif(t, int(x), _p(par), C.byref(prob))
    return F, prob.value


def dmat_cdf_array(x, v, sv, a, z, sz, t, st, p_outlier, w_outlier):
    x = _f64(x)
    y = np.empty_like(x)
    rc = lib().wo_dmat_cdf_array(_p(x), x.size, v, sv, a, z, sz, t, st, p_outlier, w_outlier, _p(y))
    if rc == 2:
        raise AssertionError("1. / (2*w_outlier) must be smaller than RT")
    if rc == 1:
        raise ValueError("at least one of the parameters is out of the support")
    return y


def pdfsv_calls(reset=False):
    n = lib().wo_pdfsv_calls()
    if reset:
        lib().wo_pdfsv_calls_reset()
    return n
