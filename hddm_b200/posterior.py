"""Posterior-predictive batch on the GPU (BASELINE config 5).

The reference's posterior predictive check loops in Python: for every posterior draw and every observed
node, `node.random()` (hddm/likelihoods.py:76-81 -> hddm/generate.py:190-194 -> wfpt.gen_rts_from_cdf,
src/wfpt.pyx:323-354) simulates one data set, and hddm/utils.py:241-298 reduces it to summary statistics
(accuracy, mean / std / quantiles per boundary).  Here all draws of a node go through the GPU together and
only the statistics need to come back; the chi-square / G-square quantile objectives
(hddm/likelihoods.py:198-239) likewise evaluate `cdfdif` for many parameter sets in one launch.

Draws are independent, so under torch.distributed they are sharded in contiguous blocks over the ranks
(`shard_bounds` of hddm_b200/batch.py) and only the [D, 15] statistics / [D, n] probabilities are gathered.
"""
import ctypes as C

import functools

import numpy as np

from . import _lib
from ._lib import check, dptr, lib

STAT_NAMES = (("accuracy", "mean_ub", "std_ub") + tuple("%dq_ub" % q for q in (10, 30, 50, 70, 90)) +
              ("mean_lb", "std_lb") + tuple("%dq_lb" % q for q in (10, 30, 50, 70, 90)))  # hddm/utils.py:250-266
MAX_SAMPLES = 16384


def _params7(params):
    params = np.ascontiguousarray(params, dtype=np.float64)
    if params.ndim != 2 or params.shape[0] != 7:
        raise ValueError("params must be float64 [7, D]: v, sv, a, z, sz, t, st")
    return params


@functools.lru_cache(maxsize=8)
def _rt_grid(cdf_lb, cdf_ub, dt):
    """np.arange(cdf_lb, cdf_ub, dt) of src/wfpt.pyx:331, kept between calls (read-only)."""
    x = np.arange(cdf_lb, cdf_ub, dt)
    x.setflags(write=False)
    return x


def gen_rts_from_cdf_batch(params, samples=1000, cdf_lb=-6, cdf_ub=6, dt=1e-2, f=None, u_delay=None, seed=None,
                           return_rts=True, return_stats=False, first_draw=0):
    """gen_rts_from_cdf (src/wfpt.pyx:323-354) for D posterior draws in one pass.

    params: float64 [7, D] (rows v, sv, a, z, sz, t, st).  Uniforms: `f` / `u_delay` float64 [D, samples]
    (the reference's draw order per call is f, then the delays when st != 0), or `seed` for device-side
    Philox uniforms (keyed by the global draw index first_draw + column, so any split of the draws over
    calls or GPUs gives the same data sets), or neither -- then numpy's global RNG is used draw by draw exactly as D reference
    calls would.  Returns rts [D, samples] and/or stats [D, 15] (columns STAT_NAMES)."""
    params = _params7(params)
    D = params.shape[1]
    x = _rt_grid(float(cdf_lb), float(cdf_ub), float(dt))
    rts = np.empty((D, samples), dtype=np.double) if return_rts else None
    stats = np.empty((D, len(STAT_NAMES)), dtype=np.double) if return_stats else None
    rp = dptr(rts) if rts is not None else None
    sp = dptr(stats) if stats is not None else None
    if seed is not None:
        check(lib().hddm_gen_rts_from_cdf_batch_seeded(dptr(x), x.shape[0], dptr(params), D, int(seed), int(first_draw), samples,
                                                       rp, sp))
    else:
        if f is None:
            f = np.empty((D, samples))
            u_delay = np.zeros((D, samples))
            for d in range(D):  # numpy's global stream in the order of D successive reference calls
                f[d] = np.random.rand(samples)
                if params[6, d] != 0:
                    u_delay[d] = np.random.rand(samples)
        f = np.ascontiguousarray(f, dtype=np.float64)
        if f.shape != (D, samples):
            raise ValueError("f must be [D, samples]")
        up = None
        if u_delay is not None:
            u_delay = np.ascontiguousarray(u_delay, dtype=np.float64)
            if u_delay.shape != (D, samples):
                raise ValueError("u_delay must be [D, samples]")
            up = dptr(u_delay)
        check(lib().hddm_gen_rts_from_cdf_batch(dptr(x), x.shape[0], dptr(params), D, dptr(f), up, samples, rp, sp))
    if return_rts and return_stats:
        return rts, stats
    return rts if return_rts else stats


def ppc_stats(rts):
    """Statistics of hddm/utils.py:241-267 for D data sets: rts float64 [D, samples] of signed RTs (or a
    single vector) -> [D, 15] (columns STAT_NAMES)."""
    rts = np.ascontiguousarray(np.atleast_2d(np.asarray(rts, dtype=np.float64)))
    D, S = rts.shape
    out = np.empty((D, len(STAT_NAMES)), dtype=np.double)
    check(lib().hddm_ppc_stats(dptr(rts), D, S, dptr(out)))
    return out


def dmat_cdf_batch(x, params, return_status=False):
    """dmat_cdf_array (src/cdfdif_wrapper.pyx:16-52) for D parameter sets: x float64 [D, n] signed RTs (set
    d's quantile edges), params float64 [9, D] (v, sv, a, z, sz, t, st, p_outlier, w_outlier).  Returns y
    [D, n]; rows where the reference would raise (out of support, RT range) are NaN and flagged in status."""
    x = np.ascontiguousarray(x, dtype=np.float64)
    params = np.ascontiguousarray(params, dtype=np.float64)
    if x.ndim != 2 or params.ndim != 2 or params.shape != (9, x.shape[0]):
        raise ValueError("x must be [D, n] and params [9, D]")
    if (params[8] == 0).any():
        raise ZeroDivisionError("float division")  # src/cdfdif_wrapper.pyx:12
    D, n = x.shape
    y = np.empty((D, n), dtype=np.double)
    status = np.zeros(D, dtype=np.int32)
    check(lib().hddm_dmat_cdf_batch(dptr(x), n, dptr(params), D, dptr(y), status.ctypes.data_as(_lib.c_ip)))
    return (y, status) if return_status else y


def quantile_objectives(emp_rt, freq_obs, n_samples, params):
    """chi-square and G-square of hddm/likelihoods.py:198-239 for D parameter sets against D nodes'
    quantile statistics: emp_rt [D, n_edges] signed quantile edges, freq_obs [D, n_edges + 1] observed
    counts per bin, n_samples [D].  Returns (chisq [D], gsq [D]); sets out of support give (inf, -inf)
    as the reference's except-branches do (:221-223, :236-238)."""
    from scipy import stats
    emp_rt = np.ascontiguousarray(emp_rt, dtype=np.float64)
    freq_obs = np.asarray(freq_obs, dtype=np.float64)
    n_samples = np.asarray(n_samples, dtype=np.float64)
    cdf, status = dmat_cdf_batch(emp_rt, params, return_status=True)
    D = emp_rt.shape[0]
    chisq, gsq = np.full(D, np.inf), np.full(D, -np.inf)
    for d in range(D):
        if status[d] == _lib.ERR_RT_RANGE:
            raise AssertionError("1. / (2*w_outlier) must be smaller than RT")
        if status[d] != _lib.OK:
            continue
        prop = np.diff(np.concatenate(([0.], cdf[d], [1.])))  # :207-212
        prop[prop <= 1e-6] = 1e-6
        chisq[d] = stats.chisquare(freq_obs[d], prop * n_samples[d])[0]
        gsq[d] = 2 * np.sum(freq_obs[d] * np.log(prop))
    return chisq, gsq


def simulate_drift_batch(params, samples=1000, dt=1e-4, intra_sv=1.0, seed=0, u_init=None, u_steps=None,
                         max_steps=10000000, switch=None):
    """Random-walk simulator of hddm/generate.py:208-319 (_gen_rts_from_simulated_drift) for D parameter
    sets x `samples` trials on the GPU: params float64 [7, D] -> signed RTs [D, samples].  Device Philox
    uniforms keyed by `seed`, or -- parity mode -- the caller's u_init [D, samples, 3] (delay uniform,
    start-point uniform, standard-normal drift deviate) and u_steps [D, samples, max_steps].
    `switch`: float64 [3, D] = v_switch, V_switch, t_switch, the variant of :234-241,272-276,283-285 (the
    drift becomes normal(v_switch, V_switch) from step round(t_switch / dt) on); u_init then has a fourth
    column, the standard-normal deviate of the switched drift."""
    params = _params7(params)
    D = params.shape[1]
    cols = 3
    if switch is not None:
        switch = np.ascontiguousarray(switch, dtype=np.float64)
        if switch.shape != (3, D):
            raise ValueError("switch must be float64 [3, D]: v_switch, V_switch, t_switch")
        cols = 4
    rts = np.empty((D, samples), dtype=np.double)
    ip = up = None
    if u_init is not None or u_steps is not None:
        u_init = np.ascontiguousarray(u_init, dtype=np.float64)
        u_steps = np.ascontiguousarray(u_steps, dtype=np.float64)
        if u_init.shape != (D, samples, cols) or u_steps.ndim != 3 or u_steps.shape[:2] != (D, samples):
            raise ValueError("u_init must be [D, samples, %d] and u_steps [D, samples, max_steps]" % cols)
        max_steps = u_steps.shape[2]
        ip, up = dptr(u_init), dptr(u_steps)
    if switch is None:
        check(lib().hddm_simulate_drift_batch(dptr(params), D, samples, dt, intra_sv, int(seed), ip, up,
                                              int(max_steps), dptr(rts)))
    else:
        check(lib().hddm_simulate_drift_switch_batch(dptr(params), dptr(switch), D, samples, dt, intra_sv, int(seed),
                                                     ip, up, int(max_steps), dptr(rts)))
    return rts


def gen_rts_from_simulated_drift(params, samples=1000, dt=1e-4, intra_sv=1., seed=None):
    """hddm/generate.py:208 signature for one parameter dict (keys v, a, z, t and optionally sv, sz, st, and
    v_switch, t_switch[, V_switch] for the switched-drift variant); returns the RT vector (the reference
    also returns the drift paths, which only its plotting code uses)."""
    col = np.array([[params['v']], [params.get('sv', 0.)], [params['a']], [params.get('z', .5)],
                    [params.get('sz', 0.)], [params['t']], [params.get('st', 0.)]], dtype=np.float64)
    switch = None
    if 'v_switch' in params:
        switch = np.array([[params['v_switch']], [params.get('V_switch', 0.)], [params['t_switch']]], dtype=np.float64)
    if seed is None:
        seed = np.random.randint(0, 2 ** 31 - 1)
    return simulate_drift_batch(col, 1 if samples is None else samples, dt, intra_sv, seed, switch=switch)[0]


def post_pred_stats_sharded(params, samples, cdf_lb=-6, cdf_ub=6, dt=1e-4, seed=0, evaluate=None, group=None,
                            timings=None):
    """Posterior predictive statistics for D draws with the draws sharded over the ranks of `group` (one
    process per GPU, contiguous blocks as hddm_b200.batch.shard_bounds); every rank returns the full
    float64 [D, 15] table.  The only collective is one all_gather of the per-draw statistics (120 bytes per
    draw); the simulated data sets never leave their GPU.  Device uniforms are keyed by the global draw
    index, so the table does not depend on the world size.

    `evaluate(params_block, first_draw) -> float64 [block, 15]` defaults to the CUDA path; tests pass a
    stand-in to exercise the partition / gather logic over gloo on CPU.  `timings` (a dict, optional) receives
    this rank's wall-clock seconds per stage: "evaluate" (host staging + kernels + statistics back on the host),
    "gather" (to the device, all_gather, full table back to the host)."""
    import time
    import torch
    import torch.distributed as dist
    from .batch import gather_rows, shard_bounds
    params = _params7(params)
    D = params.shape[1]
    if dist.is_available() and dist.is_initialized():
        world, rank = dist.get_world_size(group), dist.get_rank(group)
    else:
        world, rank = 1, 0
    lo, hi = shard_bounds(D, world, rank)
    block = np.ascontiguousarray(params[:, lo:hi])
    if evaluate is None:
        evaluate = lambda b, first: gen_rts_from_cdf_batch(b, samples, cdf_lb, cdf_ub, dt, seed=seed, return_rts=False,
                                                           return_stats=True, first_draw=first)
    t0 = time.perf_counter()
    local = evaluate(block, lo) if hi > lo else np.empty((0, len(STAT_NAMES)))
    t1 = time.perf_counter()
    local = torch.from_numpy(np.ascontiguousarray(local, dtype=np.float64))
    if world > 1 and dist.get_backend(group) == "nccl":
        local = local.cuda()
    table = gather_rows(local, D, group).cpu().numpy()
    if timings is not None:
        timings["evaluate"] = t1 - t0
        timings["gather"] = time.perf_counter() - t1
    return table


def dmat_cdf_sharded(x, params, evaluate=None, group=None):
    """dmat_cdf_batch (cdfdif quantile probabilities of D parameter sets, src/cdfdif_wrapper.pyx:16-52) with the sets
    sharded over the ranks of `group` like post_pred_stats_sharded: every rank evaluates its contiguous block of
    rows of x [D, n] / columns of params [9, D] and returns the full [D, n] table after one all_gather (8 n bytes
    per set).  Rows the reference would refuse are NaN.  `evaluate(x_block, params_block) -> [block, n]` defaults
    to the CUDA path; tests pass a stand-in to exercise the partition / gather logic over gloo on CPU."""
    import torch
    import torch.distributed as dist
    from .batch import gather_rows, shard_bounds
    x = np.ascontiguousarray(x, dtype=np.float64)
    params = np.ascontiguousarray(params, dtype=np.float64)
    if x.ndim != 2 or params.ndim != 2 or params.shape != (9, x.shape[0]):
        raise ValueError("x must be [D, n] and params [9, D]")
    D, n = x.shape
    if dist.is_available() and dist.is_initialized():
        world, rank = dist.get_world_size(group), dist.get_rank(group)
    else:
        world, rank = 1, 0
    lo, hi = shard_bounds(D, world, rank)
    if evaluate is None:
        evaluate = dmat_cdf_batch
    local = evaluate(x[lo:hi], np.ascontiguousarray(params[:, lo:hi])) if hi > lo else np.empty((0, n))
    local = torch.from_numpy(np.ascontiguousarray(local, dtype=np.float64).reshape(hi - lo, n))
    if world > 1 and dist.get_backend(group) == "nccl":
        local = local.cuda()
    return gather_rows(local, D, group).cpu().numpy()
