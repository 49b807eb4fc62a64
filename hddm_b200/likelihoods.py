"""The WFPT stochastic of the reference's plugin boundary (hddm/likelihoods.py) on top of the CUDA path.

`generate_wfpt_stochastic_class(wiener_params, sampling_method, cdf_range, sampling_dt)` has the
reference's signature and defaults (hddm/likelihoods.py:30-48) and returns a class with the same
members: `logp` = `wfpt_like(x, v, sv, a, z, sz, t, st, p_outlier)` (hddm/likelihoods.py:52-73), `.pdf`,
`.cdf`, `.cdf_vec`, `.random` (:76-99) and the quantile / chi-square / G-square methods (:107-283).
When kabuki is installed the class is built by `kabuki.utils.stochastic_from_dist` exactly as the
reference does (:94); otherwise by a small stand-in with the same `value` / `parents` / `logp`
protocol, which is what the tests and the batched sampler of hddm_b200/sampler.py use (pymc 2.3 and
kabuki are not installable here).

Extensions (the reason this file exists instead of the three-line import swap of INTEGRATION.md):
  * every observed node keeps its signed RTs resident on the GPU (`wfpt.Trials`), uploaded on first
    use and re-created lazily after unpickling -- no device handle is ever pickled
    (hddm/models/base.py:731-739 drops and rebuilds the stochastic class on load);
  * `wfpt_like_batch(nodes, set_node, params)` evaluates many parameter sets (proposals, chains,
    all children of a group-level parameter) against many observed nodes in ONE launch.
Nothing here computes a likelihood on the CPU.
"""
import zlib

import numpy as np

from . import cdfdif_wrapper, wfpt

DEFAULT_WIENER_PARAMS = {'err': 1e-4, 'n_st': 2, 'n_sz': 2, 'use_adaptive': 1, 'simps_err': 1e-3,
                         'w_outlier': 0.1}  # hddm/likelihoods.py:43-48 == hddm/models/base.py:712-718
_PARENTS = ('v', 'sv', 'a', 'z', 'sz', 't', 'st')


def _rt_array(x):
    """Signed RTs of an observed node's value: DataFrame / structured array with an 'rt' column, or a
    plain float64 vector."""
    if isinstance(x, np.ndarray) and x.dtype.names is None:
        return np.ascontiguousarray(x, dtype=np.float64)
    col = x['rt']
    return np.ascontiguousarray(getattr(col, 'values', col), dtype=np.float64)


class DeviceTrialCache(object):
    """Signed-RT vectors resident on the GPU, so that the thousands of logp calls an MCMC run makes on one
    observed node upload its data once.  Read-only data (what pandas' copy-on-write `.values` returns) is
    recognised by identity -- address, shape, strides, plus a weak reference to the owning array that drops
    the entry when the memory is released -- so the latency path does not re-hash the vector; anything else
    is looked up by content (length + two checksums, confirmed byte for byte on a hit).
    Least-recently-used entries are evicted one at a time."""

    def __init__(self, max_entries=4096):
        from collections import OrderedDict
        self.max_entries = max_entries
        self._handles = OrderedDict()   # content key -> (Trials, host copy of the bytes)
        self._by_id = {}                # (address, shape, strides) -> (weakref to the owning array, content key)

    @staticmethod
    def _key(rt):
        buf = memoryview(rt).cast('B')
        return (rt.shape[0], zlib.adler32(buf), zlib.crc32(buf))

    def get(self, rt):
        import weakref
        owner = rt  # the array that owns the memory: pandas hands out a fresh read-only view per `.values`
        while isinstance(owner.base, np.ndarray):
            owner = owner.base
        layout = (rt.__array_interface__['data'][0], rt.shape, rt.strides)
        hit = self._by_id.get(layout)
        # identity is trusted only for read-only data (pandas copy-on-write views): nothing can change
        # the bytes behind our back, and the weak reference drops the entry before the address is reused
        if hit is not None and hit[0]() is owner and not rt.flags.writeable:
            ent = self._handles.get(hit[1])
            if ent is not None:
                self._handles.move_to_end(hit[1])
                return ent[0]
        key = self._key(rt)
        ent = self._handles.get(key)
        if ent is not None and not np.array_equal(ent[1], rt, equal_nan=True):
            key = key + (len(self._handles),)  # checksum collision: keep both, under distinct keys
            ent = None
        if ent is None:
            while len(self._handles) >= self.max_entries:
                _, (old, _) = self._handles.popitem(last=False)
                old.close()
            ent = self._handles[key] = (wfpt.Trials(rt), rt.copy())
        else:
            self._handles.move_to_end(key)
        if not rt.flags.writeable:
            self._by_id[layout] = (weakref.ref(owner, lambda _, k=layout: self._by_id.pop(k, None)), key)
        return ent[0]

    def clear(self):
        for h, _ in self._handles.values():
            h.close()
        self._handles.clear()
        self._by_id.clear()

    def __getstate__(self):  # never pickle device handles
        return {'max_entries': self.max_entries}

    def __setstate__(self, state):
        self.__init__(state['max_entries'])


_cache = DeviceTrialCache()


def _prob_upper(v, a, z):
    """hddm/likelihoods.py:66-69 (== src/pdf.pxi:67-72)."""
    if v == 0:
        return z
    return (np.exp(-2 * a * z * v) - 1) / (np.exp(-2 * a * v) - 1)


def make_wfpt_like(wiener_params=None, cache=None):
    """The logp closure of hddm/likelihoods.py:52-73 on device-resident trials."""
    wp = dict(DEFAULT_WIENER_PARAMS if wiener_params is None else wiener_params)
    cache = _cache if cache is None else cache

    def wfpt_like(x, v, sv, a, z, sz, t, st, p_outlier=0):
        rt = _rt_array(x)
        if rt.size == 0 or np.abs(rt).max() < 998:
            return cache.get(rt).wiener_like(v, sv, a, z, sz, t, st, p_outlier=p_outlier, **wp)
        # missing responses coded +-999 (hddm/likelihoods.py:56-73): RT trials as usual, the rest binomial
        from scipy import stats
        noresponse = np.abs(rt) >= 999
        logp_resp = cache.get(np.ascontiguousarray(rt[~noresponse])).wiener_like(
            v, sv, a, z, sz, t, st, p_outlier=p_outlier, **wp)
        n_noresponse = int(noresponse.sum())
        k_upper = int((rt[noresponse] > 0).sum())
        return logp_resp + stats.binom.logpmf(k_upper, n_noresponse, _prob_upper(v, a, z))

    return wfpt_like


class BatchedNodes(object):
    """All observed nodes of a model as ONE device-resident trial vector with a node table.

        nodes = BatchedNodes([rt_subj0, rt_subj1, ...])
        logp = nodes.logp(set_node, params)     # params float64 [9, P]; set p <-> node set_node[p]

    One launch evaluates every (node, proposal / chain) pair: this is what turns kernel throughput
    into sampling speed for 200-trial nodes, where a launch per node is latency-bound."""

    def __init__(self, rts, wiener_params=None):
        rts = [_rt_array(r) for r in rts]
        self.node_off = np.concatenate([[0], np.cumsum([r.shape[0] for r in rts])]).astype(np.int64)
        self.n_nodes = len(rts)
        self._x = np.concatenate(rts) if rts else np.zeros(0)
        self._trials = None
        wp = dict(DEFAULT_WIENER_PARAMS if wiener_params is None else wiener_params)
        self.w_outlier = wp.pop('w_outlier')
        self.numerics = wp

    def _handle(self):
        if self._trials is None:  # lazily (re-)created: survives pickling of the owner
            self._trials = wfpt.Trials(self._x)
        return self._trials

    def pack(self, v, sv, a, z, sz, t, st, p_outlier=0.0):
        """Broadcast scalars / vectors to the [9, P] layout of the batched entry points."""
        cols = [np.atleast_1d(np.asarray(c, dtype=np.float64)) for c in (v, sv, a, z, sz, t, st, p_outlier)]
        P = max(c.shape[0] for c in cols)
        out = np.empty((9, P))
        for k, c in enumerate(cols):
            out[k] = c
        out[8] = self.w_outlier
        return out

    def logp(self, set_node, params, return_evals=False):
        return self._handle().wiener_like_nodes(self.node_off, set_node, params, return_evals=return_evals,
                                                **self.numerics)

    def __getstate__(self):
        d = dict(self.__dict__)
        d['_trials'] = None
        return d


def wfpt_like_batch(nodes, set_node, params):
    """Batched counterpart of wfpt_like: `nodes` a BatchedNodes, returns float64[P] (-inf where the
    reference returns -inf)."""
    return nodes.logp(set_node, params)


# ---- the stochastic class ------------------------------------------------------------------------

def _standin_stochastic_from_dist(name, logp):
    """Minimal stand-in for kabuki.utils.stochastic_from_dist (-> pymc.stochastic_from_dist): a class whose
    instances hold `value` (the observed data) and `parents` (a dict) and whose `logp` calls
    logp(value, **parents) -- the protocol hddm/likelihoods.py relies on (:52, :85-99)."""

    class Node(object):
        def __init__(self, name_=name, value=None, observed=True, **parents):
            self.__name__ = name_
            self.value = value
            self.observed = observed
            self.parents = dict(parents)

        @property
        def shape(self):
            return np.shape(_rt_array(self.value))

        @property
        def logp(self):
            return logp(self.value, **self.parents)

    Node.__name__ = name
    Node._logp_fun = staticmethod(logp)
    return Node


def _stochastic_from_dist(name, logp):
    try:
        from kabuki.utils import stochastic_from_dist  # hddm/likelihoods.py:5,94
    except ImportError:
        return _standin_stochastic_from_dist(name, logp)
    return stochastic_from_dist(name, logp)


def data_quantiles(rt, quantiles=(0.1, 0.3, 0.5, 0.7, 0.9)):
    """hddm/utils.py:646-666 on signed RTs: (q_lower, q_upper, p_upper)."""
    from scipy.stats.mstats import mquantiles
    rt = _rt_array(rt)
    quantiles = np.asarray(quantiles)
    return (np.asarray(mquantiles(-rt[rt < 0], quantiles)), np.asarray(mquantiles(rt[rt > 0], quantiles)),
            float(np.mean(rt > 0)))


def generate_wfpt_stochastic_class(wiener_params=None, sampling_method='cdf', cdf_range=(-5, 5),
                                   sampling_dt=1e-4):
    """hddm/likelihoods.py:30-105.  sampling_method 'cdf' (src/wfpt.pyx:323-354) or 'drift' (the random-walk
    simulator of hddm/generate.py:208-319), both on the GPU."""
    wp = dict(DEFAULT_WIENER_PARAMS if wiener_params is None else wiener_params)
    wfpt_like = make_wfpt_like(wp)

    def random(self):  # :76-81 -> hddm/generate.py:190-194 -> src/wfpt.pyx:323-354
        par = {k: self.parents[k] for k in _PARENTS}
        size = int(np.prod(self.shape))
        if sampling_method == 'drift':  # hddm/generate.py:187-188
            from . import posterior
            return posterior.gen_rts_from_simulated_drift(par, size, sampling_dt)
        if sampling_method != 'cdf':
            raise TypeError("Sampling method %s not found." % sampling_method)  # hddm/generate.py:195-196
        return wfpt.gen_rts_from_cdf(par['v'], par['sv'], par['a'], par['z'], par['sz'], par['t'], par['st'],
                                     size, cdf_range[0], cdf_range[1], sampling_dt)

    def pdf(self, x):  # :85-87
        return wfpt.pdf_array(_rt_array(x), **self.parents)

    def cdf(self, x):  # :90-91
        return cdfdif_wrapper.dmat_cdf_array(_rt_array(x), w_outlier=wp['w_outlier'], **self.parents)

    def cdf_vec(self):  # :98
        return wfpt.gen_cdf_using_pdf(time=cdf_range[1], **dict(list(self.parents.items()) + list(wp.items())))

    cls = _stochastic_from_dist('wfpt', wfpt_like)
    cls.pdf = pdf
    cls.cdf_vec = cdf_vec
    cls.cdf = cdf
    cls.random = random
    add_quantiles_functions_to_pymc_class(cls)
    return cls


def add_quantiles_functions_to_pymc_class(pymc_class):
    """Quantile statistics of hddm/likelihoods.py:107-283 (chi-square / G-square objectives of
    hddm/models/base.py:250-264); the CDF evaluations go through cdfdif on the GPU."""

    def compute_quantiles_stats(self, quantiles=(0.1, 0.3, 0.5, 0.7, 0.9)):  # :115-173
        q = np.asarray(quantiles)
        if getattr(self, '_quantiles_edges', None) is not None and np.array_equal(self._quantiles_edges, q):
            return
        if hasattr(self, '_is_average_node'):
            raise AttributeError('cannot recompute stats of average model')
        self._quantiles_edges = q
        rt = _rt_array(self.value)
        pos = np.diff(np.concatenate(([0.], q, [1.])))
        self._n_samples = rt.shape[0]
        self._empirical_quantiles = data_quantiles(rt, q)
        q_lower, q_upper, _ = self._empirical_quantiles
        if not np.isnan(rt).any():
            neg = pos[::-1]
            self._emp_rt = np.concatenate(((-q_lower)[::-1], [0.], q_upper))
            self._freq_obs = np.concatenate((np.sum(rt < 0) * neg, np.sum(rt > 0) * pos))
        else:  # go/no-go data: NaN = no response (:152-173)
            self._emp_rt = np.concatenate(([0.], q_upper))
            self._freq_obs = np.concatenate(([float(np.isnan(rt).sum())], np.sum(rt > 0) * pos))

    def set_quantiles_stats(self, quantiles, n_samples, emp_rt, freq_obs, p_upper):  # :175-187
        self._quantiles_edges = np.asarray(quantiles)
        self._n_samples = n_samples
        self._emp_rt = emp_rt
        self._freq_obs = freq_obs
        nq = len(quantiles)
        self._empirical_quantiles = (-emp_rt[:nq][::-1], emp_rt[nq + 1:], p_upper)

    def get_quantiles_stats(self, quantiles=(0.1, 0.3, 0.5, 0.7, 0.9)):  # :189-196
        self.compute_quantiles_stats(quantiles)
        return {'n_samples': self._n_samples, 'emp_rt': self._emp_rt, 'freq_obs': self._freq_obs}

    def _get_theoretical_proportion(self):  # :198-212
        theo_cdf = np.concatenate(([0.], self.cdf(np.ascontiguousarray(self._emp_rt, dtype=np.float64)), [1.]))
        proportion = np.diff(theo_cdf)
        proportion[proportion <= 1e-6] = 1e-6
        return proportion

    def chisquare(self):  # :214-226
        from scipy import stats
        try:
            theo = self._get_theoretical_proportion()
        except (ValueError, FloatingPointError):
            return np.inf
        score, _ = stats.chisquare(self._freq_obs, theo * self._n_samples)
        return score

    def gsquare(self):  # :228-239
        try:
            theo = self._get_theoretical_proportion()
        except ValueError:
            return -np.inf
        return 2 * np.sum(self._freq_obs * np.log(theo))

    def empirical_quantiles(self, quantiles=(.1, .3, .5, .7, .9)):  # :241-251
        self.compute_quantiles_stats(quantiles)
        return self._empirical_quantiles

    def theoretical_quantiles(self, quantiles=(.1, .3, .5, .7, .9)):  # :253-274
        quantiles = np.asarray(quantiles)
        x_lower, cdf_lower, x_upper, cdf_upper = wfpt.split_cdf(*self.cdf_vec())
        lower_idx = np.searchsorted(cdf_lower, quantiles * cdf_lower[-1])
        upper_idx = np.searchsorted(cdf_upper, quantiles * cdf_upper[-1])
        return (x_lower[lower_idx], x_upper[upper_idx], cdf_upper[-1])

    for f in (compute_quantiles_stats, set_quantiles_stats, get_quantiles_stats, chisquare, gsquare,
              _get_theoretical_proportion, empirical_quantiles, theoretical_quantiles):
        setattr(pymc_class, f.__name__, f)


def wiener_like_contaminant(value, cont_x, v, sv, a, z, sz, t, st, t_min, t_max, err, n_st, n_sz,
                            use_adaptive, simps_err):
    """hddm/likelihoods.py:12-16."""
    return wfpt.wiener_like_contaminant(_rt_array(value), np.asarray(cont_x).astype(np.int32), v, sv, a, z, sz,
                                        t, st, t_min, t_max, err, n_st, n_sz, use_adaptive, simps_err)


def general_WienerCont(err=1e-4, n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3):
    """hddm/likelihoods.py:21-28."""
    def _like(value, cont_x, v, sv, a, z, sz, t, st, t_min, t_max, err=err, n_st=n_st, n_sz=n_sz,
              use_adaptive=use_adaptive, simps_err=simps_err):
        return wiener_like_contaminant(value, cont_x, v, sv, a, z, sz, t, st, t_min, t_max, err=err, n_st=n_st,
                                       n_sz=n_sz, use_adaptive=use_adaptive, simps_err=simps_err)
    return _stochastic_from_dist("Wiener Diffusion Contaminant Process", _like)


def _reg_column(col, value, n):
    """One regression outcome aligned with the node's trials, as hddm/models/hddm_regression.py:29-30 does it:
    a pandas Series (the regressor node's value spans the whole data set) is re-indexed by the node's own
    trial index, `.loc[value['rt'].index].values`.  Anything without an index must already have one entry
    per trial of the node; a longer vector is refused rather than silently truncated."""
    rt = value['rt'] if not (isinstance(value, np.ndarray) and value.dtype.names is None) else None
    if hasattr(col, 'loc') and rt is not None and hasattr(rt, 'index'):
        col = col.loc[rt.index].values
    col = np.ascontiguousarray(getattr(col, 'values', col), dtype=np.float64)
    if col.ndim != 1 or col.shape[0] != n:
        raise ValueError("regression outcome has %s entries for a node of %d trials" % (col.shape, n))
    return col


def make_wiener_multi_like(wiener_params=None):
    """The logp closure of hddm/models/hddm_regression.py:26-36."""
    wp = dict(DEFAULT_WIENER_PARAMS if wiener_params is None else wiener_params)

    def wiener_multi_like(value, v, sv, a, z, sz, t, st, reg_outcomes, p_outlier=0.05):
        x = _rt_array(value)
        params = {'v': v, 'sv': sv, 'a': a, 'z': z, 'sz': sz, 't': t, 'st': st}
        for name in reg_outcomes:
            params[name] = _reg_column(params[name], value, x.shape[0])
        return wfpt.wiener_like_multi(x, params['v'], params['sv'], params['a'], params['z'], params['sz'],
                                      params['t'], params['st'], 1e-4, reg_outcomes, w_outlier=wp['w_outlier'],
                                      p_outlier=p_outlier)

    return wiener_multi_like


def generate_wfpt_reg_stochastic_class(wiener_params=None, sampling_method='cdf', cdf_range=(-5, 5),
                                       sampling_dt=1e-4):
    """hddm/models/hddm_regression.py:15-60: the stochastic of HDDMRegressor.  Same closure signature and
    defaults (p_outlier=0.05, w_outlier from wiener_params, err fixed at 1e-4, numerics left at
    wiener_like_multi's own defaults as the reference leaves them)."""
    wp = dict(DEFAULT_WIENER_PARAMS if wiener_params is None else wiener_params)
    wiener_multi_like = make_wiener_multi_like(wp)

    def random(self):  # :39-54: one simulated trial per row, each with its own regressed parameters
        from . import posterior
        parents = dict(self.parents)
        reg = list(parents.pop('reg_outcomes'))
        parents.pop('p_outlier', None)
        x = _rt_array(self.value)
        n = x.shape[0]
        cols = np.empty((7, n))
        for k, name in enumerate(_PARENTS):
            cols[k] = _reg_column(parents[name], self.value, n) if name in reg else parents[name]
        seed = np.random.randint(0, 2 ** 31 - 1)
        if sampling_method == 'drift':  # hddm/generate.py:187-188
            rts = posterior.simulate_drift_batch(cols, 1, sampling_dt, seed=seed)[:, 0]
        elif sampling_method == 'cdf':
            rts = posterior.gen_rts_from_cdf_batch(cols, 1, cdf_range[0], cdf_range[1], sampling_dt, seed=seed)[:, 0]
        else:
            raise TypeError("Sampling method %s not found." % sampling_method)  # hddm/generate.py:195-196
        out = self.value.copy() if hasattr(self.value, 'copy') else np.array(self.value)
        if isinstance(out, np.ndarray) and out.dtype.names is None:
            return rts
        out['rt'] = rts  # signed, as hddm.utils.flip_errors leaves them (:52)
        return out

    cls = _stochastic_from_dist('wfpt_reg', wiener_multi_like)
    cls.random = random
    return cls


wfpt_reg_like = generate_wfpt_reg_stochastic_class(sampling_method='drift')  # hddm/models/hddm_regression.py:62
wiener_multi_like = make_wiener_multi_like()


# ---- reinforcement-learning stochastics ------------------------------------------------------------

def _rl_columns(x):
    """The columns hddm/models/hddm_rl.py:68-72 pulls out of an observed node's DataFrame (or a dict /
    structured array with the same names)."""
    col = lambda name: np.asarray(getattr(x[name], 'values', x[name]))
    q_init = x['q_init']
    q = float(q_init.iloc[0] if hasattr(q_init, 'iloc') else np.asarray(q_init).ravel()[0])
    return (np.ascontiguousarray(col('response').astype(int), dtype=np.int64), q,
            np.ascontiguousarray(col('feedback').astype(float)), np.ascontiguousarray(col('split_by').astype(int), dtype=np.int64))


def wienerRL_like(x, v, alpha, pos_alpha, sv, a, z, sz, t, st, p_outlier=0):
    """hddm/models/hddm_rl.py:61-73."""
    wp = DEFAULT_WIENER_PARAMS
    response, q, feedback, split_by = _rl_columns(x)
    return wfpt.wiener_like_rlddm(_rt_array(x), response, feedback, split_by, q, alpha, pos_alpha, v, sv, a, z, sz,
                                  t, st, p_outlier=p_outlier, **wp)


def RL_like(x, v, alpha, pos_alpha, z=0.5, p_outlier=0):
    """hddm/models/rl.py:72-85."""
    wp = DEFAULT_WIENER_PARAMS
    response, q, feedback, split_by = _rl_columns(x)
    return wfpt.wiener_like_rl(response, feedback, split_by, q, alpha, pos_alpha, v, z, p_outlier=p_outlier, **wp)


def wienerRL_multi_like(value, v, sv, a, z, sz, t, st, alpha, reg_outcomes, p_outlier=0):
    """hddm/models/hddm_rl_regression.py:25-38: parameters named in reg_outcomes are per-trial Series,
    re-indexed by the node's trials (:32-33) as in wiener_multi_like."""
    response, q, feedback, split_by = _rl_columns(value)
    params = {'v': v, 'sv': sv, 'a': a, 'z': z, 'sz': sz, 't': t, 'st': st, 'alpha': alpha}
    n = response.shape[0]
    for name in reg_outcomes:
        params[name] = _reg_column(params[name], value, n)
    return wfpt.wiener_like_multi_rlddm(_rt_array(value), response, feedback, split_by, q, params['v'], params['sv'],
                                        params['a'], params['z'], params['sz'], params['t'], params['st'],
                                        params['alpha'], 1e-4, reg_outcomes, p_outlier=p_outlier)


WienerRL = _stochastic_from_dist('wienerRL', wienerRL_like)  # hddm/models/hddm_rl.py:73
RL = _stochastic_from_dist('RL', RL_like)                    # hddm/models/rl.py:85
