"""Drop-in for the reference's `wfpt` extension module (hddm.wfpt, /root/reference/src/wfpt.pyx).

Same function names, positional order, keyword names and defaults; same return types and error
behaviour (-inf for zero probability, ValueError out of support).  Every function forwards to the
CUDA library through the C ABI of include/hddm_b200.h -- nothing is computed in Python except
what the reference itself does in numpy (grid construction with np.linspace / np.arange, the
uniforms of np.random, split_cdf's pure-numpy reshuffle).
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import as_rt_buffer, check, dptr, lib

__all__ = ["Regression", "full_pdf", "pdf_array", "wiener_like", "wiener_like_multi", "wiener_like_contaminant",
           "wiener_like_rlddm", "wiener_like_rl", "wiener_like_multi_rlddm", "wiener_like_rlddm_batch",
           "wiener_like_rl_batch",
           "gen_cdf_using_pdf", "gen_rts_from_cdf", "split_cdf", "wiener_like_batch",
           "wiener_like_multi_batch", "wiener_like_regression_batch", "wiener_like_nodes", "Trials"]

_PARAM_NAMES = ("v", "sv", "a", "z", "sz", "t", "st")


def full_pdf(x, v, sv, a, z, sz, t, st, err, n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3):
    """src/pdf.pxi:104-146 (cpdef full_pdf)."""
    out = C.c_double()
    check(lib().hddm_full_pdf(x, v, sv, a, z, sz, t, st, err, n_st, n_sz, int(bool(use_adaptive)),
                              simps_err, C.byref(out)))
    return out.value


def pdf_array(x, v, sv, a, z, sz, t, st, err=1e-4, logp=0, n_st=2, n_sz=2, use_adaptive=1,
              simps_err=1e-3, p_outlier=0, w_outlier=0):
    """src/wfpt.pyx:32-48."""
    x = as_rt_buffer(x)
    y = np.empty(x.shape[0], dtype=np.double)
    check(lib().hddm_pdf_array(dptr(x), x.shape[0], v, sv, a, z, sz, t, st, err, int(bool(logp)), n_st,
                               n_sz, int(bool(use_adaptive)), simps_err, p_outlier, w_outlier, dptr(y)))
    return y


def wiener_like(x, v, sv, a, z, sz, t, st, err, n_st=10, n_sz=10, use_adaptive=1, simps_err=1e-8,
                p_outlier=0, w_outlier=0.1):
    """src/wfpt.pyx:54-76."""
    x = as_rt_buffer(x)
    out = C.c_double()
    check(lib().hddm_wiener_like(dptr(x), x.shape[0], v, sv, a, z, sz, t, st, err, n_st, n_sz,
                                 int(bool(use_adaptive)), simps_err, p_outlier, w_outlier, C.byref(out)))
    return out.value


def wiener_like_multi(x, v, sv, a, z, sz, t, st, err, multi=None, n_st=10, n_sz=10, use_adaptive=1,
                      simps_err=1e-3, p_outlier=0, w_outlier=0):
    """src/wfpt.pyx:244-274.  Parameters named in `multi` are per-trial float64 arrays."""
    x = as_rt_buffer(x)
    if multi is None:
        # src/wfpt.pyx:253-254 passes the whole array to the scalar full_pdf and cannot succeed
        raise TypeError("only size-1 arrays can be converted to Python scalars")
    values = dict(zip(_PARAM_NAMES, (v, sv, a, z, sz, t, st)))
    multi = set(multi)
    keep, ptrs, strides = [], (C.POINTER(C.c_double) * 7)(), (C.c_int64 * 7)()
    for k, name in enumerate(_PARAM_NAMES):
        if name in multi:
            arr = np.ascontiguousarray(np.asarray(values[name], dtype=np.float64))
            if arr.ndim != 1 or arr.shape[0] < x.shape[0]:
                raise IndexError("per-trial parameter '%s' is shorter than x" % name)
            strides[k] = 1
        else:
            arr = np.array([values[name]], dtype=np.float64)
            strides[k] = 0
        keep.append(arr)
        ptrs[k] = dptr(arr)
    out = C.c_double()
    check(lib().hddm_wiener_like_multi(dptr(x), x.shape[0], ptrs, strides, err, n_st, n_sz,
                                       int(bool(use_adaptive)), simps_err, p_outlier, w_outlier,
                                       C.byref(out)))
    return out.value


def wiener_like_contaminant(x, cont_x, v, sv, a, z, sz, t, st, t_min, t_max, err, n_st=10, n_sz=10,
                            use_adaptive=1, simps_err=1e-8):
    """src/wfpt.pyx:357-386.  cont_x must be an int32 array (np.ndarray[int, ndim=1])."""
    x = as_rt_buffer(x)
    if not isinstance(cont_x, np.ndarray):
        raise TypeError("Argument 'cont_x' has incorrect type")
    if cont_x.dtype != np.int32 or cont_x.ndim != 1:
        raise ValueError("Buffer dtype mismatch, expected 'int' but got '%s'" % cont_x.dtype)
    cont_x = np.ascontiguousarray(cont_x)
    out = C.c_double()
    check(lib().hddm_wiener_like_contaminant(dptr(x), cont_x.ctypes.data_as(_lib.c_ip), x.shape[0], v, sv,
                                             a, z, sz, t, st, t_min, t_max, err, n_st, n_sz,
                                             int(bool(use_adaptive)), simps_err, C.byref(out)))
    return out.value


def _as_long_buffer(a, name):
    """`np.ndarray[long, ndim=1]` (src/wfpt.pyx:80,82): int64 1-D ndarray, any stride."""
    if not isinstance(a, np.ndarray):
        raise TypeError("Argument '%s' has incorrect type (expected numpy.ndarray, got %s)"
                        % (name, type(a).__name__))
    if a.ndim != 1:
        raise ValueError("Buffer has wrong number of dimensions (expected 1, got %d)" % a.ndim)
    if a.dtype != np.int64:
        raise ValueError("Buffer dtype mismatch, expected 'long' but got '%s'" % a.dtype)
    return np.ascontiguousarray(a)


def _rl_task(x, response, feedback, split_by):
    response = _as_long_buffer(response, "response")
    feedback = as_rt_buffer(feedback, "feedback")
    split_by = _as_long_buffer(split_by, "split_by")
    n = response.shape[0] if x is None else x.shape[0]
    if feedback.shape[0] < n or split_by.shape[0] < n or response.shape[0] < n:
        raise IndexError("response / feedback / split_by shorter than the trial vector")
    lp = _lib.c_lp
    return response, feedback, split_by, n, response.ctypes.data_as(lp), dptr(feedback), split_by.ctypes.data_as(lp)


def wiener_like_rlddm(x, response, feedback, split_by, q, alpha, pos_alpha, v, sv, a, z, sz, t, st, err,
                      n_st=10, n_sz=10, use_adaptive=1, simps_err=1e-8, p_outlier=0, w_outlier=0):
    """src/wfpt.pyx:79-155."""
    x = as_rt_buffer(x)
    response, feedback, split_by, n, rp, fp, sp = _rl_task(x, response, feedback, split_by)
    out = C.c_double()
    check(lib().hddm_wiener_like_rlddm(dptr(x), rp, fp, sp, n, q, alpha, pos_alpha, v, sv, a, z, sz, t, st, err,
                                       n_st, n_sz, int(bool(use_adaptive)), simps_err, p_outlier, w_outlier,
                                       C.byref(out)))
    return out.value


def wiener_like_rl(response, feedback, split_by, q, alpha, pos_alpha, v, z, err=1e-4, n_st=10, n_sz=10,
                   use_adaptive=1, simps_err=1e-8, p_outlier=0, w_outlier=0):
    """src/wfpt.pyx:158-241 (err ... simps_err are accepted and unused, as in the reference)."""
    response, feedback, split_by, n, rp, fp, sp = _rl_task(None, response, feedback, split_by)
    out = C.c_double()
    check(lib().hddm_wiener_like_rl(rp, fp, sp, n, q, alpha, pos_alpha, v, z, p_outlier, w_outlier, C.byref(out)))
    return out.value


_RL_PARAM_NAMES = _PARAM_NAMES + ("alpha",)


def wiener_like_multi_rlddm(x, response, feedback, split_by, q, v, sv, a, z, sz, t, st, alpha, err, multi=None,
                            n_st=10, n_sz=10, use_adaptive=1, simps_err=1e-3, p_outlier=0, w_outlier=0):
    """src/wfpt.pyx:277-320.  Parameters named in `multi` (v, sv, a, z, sz, t, st, alpha) are per-trial."""
    x = as_rt_buffer(x)
    if multi is None:
        # src/wfpt.pyx:293-294 passes the whole array to the scalar full_pdf and cannot succeed
        raise TypeError("only size-1 arrays can be converted to Python scalars")
    response, feedback, split_by, n, rp, fp, sp = _rl_task(x, response, feedback, split_by)
    values = dict(zip(_RL_PARAM_NAMES, (v, sv, a, z, sz, t, st, alpha)))
    multi = set(multi)
    keep, ptrs, strides = [], (C.POINTER(C.c_double) * 8)(), (C.c_int64 * 8)()
    for k, name in enumerate(_RL_PARAM_NAMES):
        if name in multi:
            arr = np.ascontiguousarray(np.asarray(values[name], dtype=np.float64))
            if arr.ndim != 1 or arr.shape[0] < n:
                raise IndexError("per-trial parameter '%s' is shorter than x" % name)
            strides[k] = 1
        else:
            arr = np.array([values[name]], dtype=np.float64)
            strides[k] = 0
        keep.append(arr)
        ptrs[k] = dptr(arr)
    out = C.c_double()
    check(lib().hddm_wiener_like_multi_rlddm(dptr(x), rp, fp, sp, n, q, ptrs, strides, err, n_st, n_sz,
                                             int(bool(use_adaptive)), simps_err, p_outlier, w_outlier,
                                             C.byref(out)))
    return out.value


def wiener_like_rlddm_batch(x, response, feedback, split_by, params, err=1e-4, n_st=2, n_sz=2, use_adaptive=1,
                            simps_err=1e-3, return_evals=False):
    """NEW: P parameter sets against one subject's task.  params float64 [12, P]: rows v, sv, a, z, sz, t,
    st, p_outlier, w_outlier, q, alpha, pos_alpha; returns float64[P] of wiener_like_rlddm values."""
    x = as_rt_buffer(x)
    response, feedback, split_by, n, rp, fp, sp = _rl_task(x, response, feedback, split_by)
    params = np.ascontiguousarray(params, dtype=np.float64)
    if params.ndim != 2 or params.shape[0] != 12:
        raise ValueError("params must be float64 [12, P]")
    P = params.shape[1]
    out = np.empty(P, dtype=np.double)
    evals = C.c_ulonglong(0)
    check(lib().hddm_wiener_like_rlddm_batch(dptr(x), rp, fp, sp, n, dptr(params), P, err, n_st, n_sz,
                                             int(bool(use_adaptive)), simps_err, dptr(out), C.byref(evals) if return_evals else None))
    return (out, evals.value) if return_evals else out


def wiener_like_rl_batch(response, feedback, split_by, params):
    """NEW: params float64 [7, P]: rows q, alpha, pos_alpha, v, z, p_outlier, w_outlier."""
    response, feedback, split_by, n, rp, fp, sp = _rl_task(None, response, feedback, split_by)
    params = np.ascontiguousarray(params, dtype=np.float64)
    if params.ndim != 2 or params.shape[0] != 7:
        raise ValueError("params must be float64 [7, P]")
    P = params.shape[1]
    out = np.empty(P, dtype=np.double)
    check(lib().hddm_wiener_like_rl_batch(rp, fp, sp, n, dptr(params), P, dptr(out)))
    return out


def gen_cdf_using_pdf(v, sv, a, z, sz, t, st, err, N=500, time=5., n_st=2, n_sz=2, use_adaptive=1,
                      simps_err=1e-3, p_outlier=0, w_outlier=0):
    """src/wfpt.pyx:389-414.  Returns (x, cdf_array); ValueError out of support."""
    x = np.linspace(-time, time, 2 * N + 1)
    cdf = np.empty(x.shape[0], dtype=np.double)
    check(lib().hddm_gen_cdf_using_pdf(dptr(x), x.shape[0], v, sv, a, z, sz, t, st, err, n_st, n_sz,
                                       int(bool(use_adaptive)), simps_err, p_outlier, w_outlier, dptr(cdf)))
    return x, cdf


def gen_rts_from_cdf(v, sv, a, z, sz, t, st, samples=1000, cdf_lb=-6, cdf_ub=6, dt=1e-2,
                     f=None, u_delay=None):
    """src/wfpt.pyx:323-354.  The uniforms come from numpy's global RNG in the reference's draw
    order (f, then the delay uniforms when st != 0) unless passed explicitly (NEW, for parity
    tests and device-side pipelines)."""
    x = np.arange(cdf_lb, cdf_ub, dt)
    if f is None:
        f = np.random.rand(samples)
    f = np.ascontiguousarray(f, dtype=np.float64)
    if st != 0:
        if u_delay is None:
            u_delay = np.random.rand(samples)
        u_delay = np.ascontiguousarray(u_delay, dtype=np.float64)
        up = dptr(u_delay)
    else:
        up = None
    rts = np.empty(samples, dtype=np.double)
    check(lib().hddm_gen_rts_from_cdf(dptr(x), x.shape[0], v, sv, a, z, sz, t, st, dptr(f), up, samples,
                                      dptr(rts), None))
    return rts


def split_cdf(x, data):
    """src/wfpt.pyx:417-433 (pure numpy in the reference too; `/` is Python-2 integer division)."""
    N = (len(data) - 1) // 2
    x_lb = -x[:N][::-1]
    lb = data[:N][::-1]
    lb = np.cumsum(np.concatenate([np.array([0]), -np.diff(lb)]))
    x_ub = x[N + 1:]
    ub = data[N + 1:]
    ub -= ub[0]
    return (x_lb, lb, x_ub, ub)


# ---- NEW: batched entry points -------------------------------------------------------------------

def _pack_params(params, P=None):
    """Accept float64 [9, P] (rows v, sv, a, z, sz, t, st, p_outlier, w_outlier) or a dict of
    scalars/arrays with those names (p_outlier default 0, w_outlier default 0.1)."""
    if isinstance(params, dict):
        names = _PARAM_NAMES + ("p_outlier", "w_outlier")
        defaults = {"p_outlier": 0.0, "w_outlier": 0.1}
        cols = [np.atleast_1d(np.asarray(params.get(n, defaults.get(n)), dtype=np.float64)) for n in names]
        P = max(c.shape[0] for c in cols)
        params = np.stack([np.broadcast_to(c, (P,)) for c in cols])
    params = np.ascontiguousarray(params, dtype=np.float64)
    if params.ndim != 2 or params.shape[0] != 9:
        raise ValueError("params must be float64 [9, P]: v, sv, a, z, sz, t, st, p_outlier, w_outlier")
    return params


def wiener_like_batch(x, params, seg=None, err=1e-4, n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3,
                      return_evals=False):
    """P parameter sets against their trials in ONE launch; returns float64[P] of wiener_like values
    (src/wfpt.pyx:54-76 semantics per set, including -inf).  seg: int64[P+1] trial offsets per set
    (ragged: subjects x proposals), or None when every set sees all of x."""
    x = as_rt_buffer(x)
    params = _pack_params(params)
    P = params.shape[1]
    segp = None
    if seg is not None:
        seg = np.ascontiguousarray(seg, dtype=np.int64)
        if seg.shape != (P + 1,):
            raise ValueError("seg must have P+1 entries")
        segp = seg.ctypes.data_as(_lib.c_lp)
    out = np.empty(P, dtype=np.double)
    evals = C.c_ulonglong(0)
    check(lib().hddm_wiener_like_batch(dptr(x), x.shape[0], segp, dptr(params), P, err, n_st, n_sz,
                                       int(bool(use_adaptive)), simps_err, dptr(out), C.byref(evals) if return_evals else None))
    return (out, evals.value) if return_evals else out


def wiener_like_multi_batch(x, v_trial, params, err=1e-4, n_st=10, n_sz=10, use_adaptive=1, simps_err=1e-3,
                            return_evals=False):
    """Regression layout: v_trial float64 [P, n] per-trial drift for each of P sets, the other eight
    parameters per set in params [9, P] (row 0 ignored); wiener_like_multi semantics per set."""
    x = as_rt_buffer(x)
    params = _pack_params(params)
    P = params.shape[1]
    v_trial = np.ascontiguousarray(v_trial, dtype=np.float64)
    if v_trial.shape != (P, x.shape[0]):
        raise ValueError("v_trial must be [P, n]")
    out = np.empty(P, dtype=np.double)
    evals = C.c_ulonglong(0)
    check(lib().hddm_wiener_like_multi_batch(dptr(x), x.shape[0], dptr(v_trial), dptr(params), P, err, n_st,
                                             n_sz, int(bool(use_adaptive)), simps_err, dptr(out),
                                             C.byref(evals) if return_evals else None))
    return (out, evals.value) if return_evals else out


_OUTCOMES = {"v": 0, "sv": 1, "a": 2, "z": 3, "sz": 4, "t": 5, "st": 6}
_LINKS = {"identity": 0, None: 0, "invlogit": 1, "logistic": 1, "exp": 2}


class Regression(object):
    """Trials and design matrices of an HDDMRegressor node resident on the GPU (hddm_regdata_*): the reference
    regresses any of v, sv, a, z, sz, t, st (hddm/models/hddm_regression.py:26-36) and calls wiener_like_multi with one
    per-trial column per outcome; here every coefficient proposal is scored in one launch and the linear predictors
    are evaluated inside the kernel.

        reg = Regression(x, {"v": design_v, "a": design_a})        # design_*: float64 [K, n], intercept included
        logp = reg.like_batch({"v": coef_v, "a": coef_a}, params)   # coef_*: [K, P]; params [9, P]
    `links`: outcome -> "identity" (default) | "invlogit" | "exp", applied to the linear predictor."""

    def __init__(self, x, designs, links=None):
        x = as_rt_buffer(x)
        self.n = x.shape[0]
        self.names = list(designs)
        if len(self.names) > 4:
            raise ValueError("at most 4 regressed outcomes")
        links = links or {}
        self.outcome = (C.c_int * len(self.names))(*[_OUTCOMES[k] for k in self.names])
        self.link = (C.c_int * len(self.names))(*[_LINKS[links.get(k)] for k in self.names])
        mats = []
        for k in self.names:
            d = np.ascontiguousarray(designs[k], dtype=np.float64)
            if d.ndim != 2 or d.shape[1] != self.n or not 1 <= d.shape[0] <= 16:
                raise ValueError("design of %r must be float64 [K, n] with 1 <= K <= 16" % k)
            mats.append(d)
        self.K = [d.shape[0] for d in mats]
        Ks = (C.c_int64 * len(mats))(*self.K)
        ptrs = (_lib.c_dp * len(mats))(*[dptr(d) for d in mats])
        self._h = C.c_void_p()
        check(lib().hddm_regdata_upload(dptr(x), self.n, len(mats), Ks, ptrs, C.byref(self._h)))

    def like_batch(self, coefs, params, err=1e-4, n_st=10, n_sz=10, use_adaptive=1, simps_err=1e-3, return_evals=False):
        """wiener_like_multi (src/wfpt.pyx:244-274) per coefficient proposal; numerics default to
        wiener_like_multi's own, as hddm/models/hddm_regression.py:31-36 leaves them."""
        params = _pack_params(params)
        P = params.shape[1]
        mats = []
        for k, K in zip(self.names, self.K):
            c = np.ascontiguousarray(coefs[k], dtype=np.float64)
            if c.shape != (K, P):
                raise ValueError("coefficients of %r must be [K, P] = %r" % (k, (K, P)))
            mats.append(c)
        ptrs = (_lib.c_dp * len(mats))(*[dptr(c) for c in mats])
        out = np.empty(P, dtype=np.double)
        evals = C.c_ulonglong(0)
        check(lib().hddm_regdata_like_batch(self._h, self.outcome, self.link, ptrs, dptr(params), P, err, n_st, n_sz,
                                            int(bool(use_adaptive)), simps_err, dptr(out),
                                            C.byref(evals) if return_evals else None))
        return (out, evals.value) if return_evals else out

    def close(self):
        if self._h:
            lib().hddm_regdata_free(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def wiener_like_regression_batch(x, design, coef, params, err=1e-4, n_st=10, n_sz=10, use_adaptive=1,
                                 simps_err=1e-3, return_evals=False):
    """Regression on the drift with the identity link, one shot: design float64 [K, n] (rows = columns of the
    design matrix of `v ~ ...`, intercept included), coef float64 [K, P] (columns = coefficient proposals), params
    [9, P] (row 0 ignored).  Equivalent to wiener_like_multi_batch(x, (design.T @ coef).T, params) without ever
    forming the [P, n] drift table.  (Regression keeps the data on the GPU across calls and takes other outcomes.)"""
    x = as_rt_buffer(x)
    params = _pack_params(params)
    P = params.shape[1]
    design = np.ascontiguousarray(design, dtype=np.float64)
    coef = np.ascontiguousarray(coef, dtype=np.float64)
    if design.ndim != 2 or design.shape[1] != x.shape[0] or coef.shape != (design.shape[0], P):
        raise ValueError("design must be [K, n] and coef [K, P]")
    out = np.empty(P, dtype=np.double)
    evals = C.c_ulonglong(0)
    check(lib().hddm_wiener_like_regression_batch(dptr(x), x.shape[0], dptr(design), design.shape[0], dptr(coef),
                                                  dptr(params), P, err, n_st, n_sz, int(bool(use_adaptive)),
                                                  simps_err, dptr(out), C.byref(evals) if return_evals else None))
    return (out, evals.value) if return_evals else out


def _node_tables(node_off, set_node, P):
    node_off = np.ascontiguousarray(node_off, dtype=np.int64)
    set_node = np.ascontiguousarray(set_node, dtype=np.int64)
    if node_off.ndim != 1 or node_off.shape[0] < 1:
        raise ValueError("node_off must hold n_nodes + 1 offsets")
    if set_node.shape != (P,):
        raise ValueError("set_node must have one entry per parameter set")
    return node_off, set_node


def wiener_like_nodes(x, node_off, set_node, params, err=1e-4, n_st=2, n_sz=2, use_adaptive=1,
                      simps_err=1e-3, return_evals=False):
    """Hierarchical layout: observed node s holds trials x[node_off[s]:node_off[s+1]]; parameter set p
    (column p of params [9, P]) is evaluated against node set_node[p].  One launch for all proposals /
    chains of all nodes; returns float64[P] of wiener_like values (src/wfpt.pyx:54-76 per set)."""
    x = as_rt_buffer(x)
    params = _pack_params(params)
    P = params.shape[1]
    node_off, set_node = _node_tables(node_off, set_node, P)
    out = np.empty(P, dtype=np.double)
    evals = C.c_ulonglong(0)
    check(lib().hddm_wiener_like_nodes(dptr(x), x.shape[0], node_off.ctypes.data_as(_lib.c_lp),
                                       node_off.shape[0] - 1, set_node.ctypes.data_as(_lib.c_lp), dptr(params),
                                       P, err, n_st, n_sz, int(bool(use_adaptive)), simps_err, dptr(out),
                                       C.byref(evals) if return_evals else None))
    return (out, evals.value) if return_evals else out


class Trials(object):
    """Device-resident signed RTs of one observed node: uploaded once, reused by every logp call of
    the MCMC loop (the reference re-reads `x['rt'].values` on every call, hddm/likelihoods.py:54)."""

    def __init__(self, x):
        x = as_rt_buffer(x)
        self.n = x.shape[0]
        self._h = C.c_void_p()
        check(lib().hddm_trials_upload(dptr(x), self.n, C.byref(self._h)))

    def wiener_like(self, v, sv, a, z, sz, t, st, err, n_st=10, n_sz=10, use_adaptive=1, simps_err=1e-8,
                    p_outlier=0, w_outlier=0.1):
        out = C.c_double()
        check(lib().hddm_trials_wiener_like(self._h, v, sv, a, z, sz, t, st, err, n_st, n_sz,
                                            int(bool(use_adaptive)), simps_err, p_outlier, w_outlier,
                                            C.byref(out)))
        return out.value

    def wiener_like_batch(self, params, seg=None, err=1e-4, n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3,
                          return_evals=False):
        params = _pack_params(params)
        P = params.shape[1]
        segp = None
        if seg is not None:
            seg = np.ascontiguousarray(seg, dtype=np.int64)
            if seg.shape != (P + 1,):  # the library reads seg[0..P]
                raise ValueError("seg must have P+1 entries")
            segp = seg.ctypes.data_as(_lib.c_lp)
        out = np.empty(P, dtype=np.double)
        evals = C.c_ulonglong(0)
        check(lib().hddm_trials_wiener_like_batch(self._h, segp, dptr(params), P, err, n_st, n_sz,
                                                  int(bool(use_adaptive)), simps_err, dptr(out),
                                                  C.byref(evals) if return_evals else None))
        return (out, evals.value) if return_evals else out

    def wiener_like_nodes(self, node_off, set_node, params, err=1e-4, n_st=2, n_sz=2, use_adaptive=1,
                          simps_err=1e-3, return_evals=False):
        """As wfpt.wiener_like_nodes on the resident trials (sorted once per node table)."""
        params = _pack_params(params)
        P = params.shape[1]
        node_off, set_node = _node_tables(node_off, set_node, P)
        out = np.empty(P, dtype=np.double)
        evals = C.c_ulonglong(0)
        check(lib().hddm_trials_wiener_like_nodes(self._h, node_off.ctypes.data_as(_lib.c_lp),
                                                  node_off.shape[0] - 1, set_node.ctypes.data_as(_lib.c_lp),
                                                  dptr(params), P, err, n_st, n_sz, int(bool(use_adaptive)),
                                                  simps_err, dptr(out), C.byref(evals) if return_evals else None))
        return (out, evals.value) if return_evals else out

    def close(self):
        if self._h:
            lib().hddm_trials_free(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
