#!/usr/bin/env python
"""bench_paths.py -- secondary measurements of the hot path's siblings (bench.py holds the headline).

One JSON line per workload, each timed END TO END through the public host-buffer API (H2D of the inputs
and D2H of the results inside the timed region, wall clock around the blocking call, best of `--reps`
after a warm-up call), next to the compiled reference (oracle/_ref; the C port if absent) timed on ONE
host core on a bounded sample of the same inputs:

  simple     C1-shaped sweep: v/a/t only (sv = sz = st = 0), T trials x P sets          evals/s
  regression C3: per-trial drift (wiener_like_multi semantics) + p_outlier, T x P       evals/s, GB/s
  ppc        C5: D posterior draws x S simulated trials (dt = 1e-4 table, 15 statistics)  draws/s
  cdf        C5: cdfdif quantile probabilities, D parameter sets x 11 edges             cdf evals/s
  rlddm      RL-DDM likelihood, P proposals against one subject's task of n trials      evals/s
  simulator  random-walk simulator, D sets x S trials at dt = 1e-4                      trials/s
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
HDDM = dict(err=1e-4, n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3)


def best_of(fn, reps):
    fn()
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        out = fn()
        ts.append(time.perf_counter() - t0)
    return min(ts), out


def cpu_modules():
    from oracle import ref
    if ref.available():
        return "reference", ref.wfpt(), ref.cdfdif_wrapper()
    from oracle import oracle as O
    O.build()
    return "port", O, O


def draw(rng, n, sv=True, sz=True, st=True):
    svv = rng.uniform(0, 2.5, n) if sv else np.zeros(n)
    szv = rng.uniform(0, 0.4, n) if sz else np.zeros(n)
    stv = rng.uniform(0, 0.35, n) if st else np.zeros(n)
    z = rng.uniform(0.4, 0.6, n); v = rng.uniform(-4, 4, n); t = rng.uniform(0.2, 0.5, n); a = rng.uniform(0.5, 2.0, n)
    return np.stack([v, svv, a, z, szv, t, stv])


def rts(rng, n, onset=0.70):
    return (onset + rng.gamma(2.0, 0.25, n)) * np.where(rng.random(n) < 0.6, 1.0, -1.0)


def emit(name, value, unit, cpu_value, kind, sample, **extra):
    line = {"workload": name, "value": value, "unit": unit, "n_gpus": 1, "dtype": "f64", "data": "synthetic",
            "timing": "end to end through the host-buffer API, best of reps",
            "cpu_baseline": {"value": cpu_value, "unit": unit, "cores": 1, "kind": kind, "sample": sample},
            "speedup_vs_one_core": value / cpu_value if cpu_value else None}
    line.update(extra)
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--only", default="")
    args = ap.parse_args()
    from hddm_b200 import _lib, posterior, wfpt
    _lib.check(_lib.lib().hddm_b200_init(0))
    kind, W, CW = cpu_modules()
    rng = np.random.default_rng(20261104)
    want = set(args.only.split(",")) if args.only else None
    on = lambda k: want is None or k in want

    if on("simple"):
        T, P = 100000, 10000
        x = rts(rng, T)
        params = np.concatenate([draw(rng, P, False, False, False), np.full((1, P), 0.05), np.full((1, P), 0.1)])
        dt, lp = best_of(lambda: wfpt.wiener_like_batch(x, params, **HDDM), args.reps)
        xs = x[:20000]
        t0 = time.perf_counter()
        ref = [W.wiener_like(xs, *params[:7, p], 1e-4, 2, 2, 1, 1e-3, 0.05, 0.1) for p in range(64)]
        cdt = time.perf_counter() - t0
        emit("simple DDM sweep (v,a,t; sv=sz=st=0): %d trials x %d sets" % (T, P), T * P / dt, "evals/s",
             xs.size * 64 / cdt, kind, "first 20000 trials x first 64 sets", ms=dt * 1e3)

    if on("regression"):
        T, P = 100000, 512
        x = rts(rng, T)
        params = np.concatenate([draw(rng, P), np.full((1, P), 0.05), np.full((1, P), 0.1)])
        cov = rng.standard_normal(T)
        v_trial = np.ascontiguousarray(params[0][:, None] + 0.5 * cov[None, :])  # v ~ 1 + covariate per set
        nm = dict(err=1e-4, n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3)
        dt, lp = best_of(lambda: wfpt.wiener_like_multi_batch(x, v_trial, params, **nm), args.reps)
        xs, vs = x[:5000], v_trial[:4, :5000]
        t0 = time.perf_counter()
        for p in range(4):
            W.wiener_like_multi(xs, np.ascontiguousarray(vs[p]), *params[1:7, p], 1e-4, ['v'], 2, 2, 1, 1e-3, 0.05, 0.1)
        cdt = time.perf_counter() - t0
        emit("regression sweep (per-trial drift, full DDM, p_outlier): %d trials x %d sets" % (T, P),
             T * P / dt, "evals/s", xs.size * 4 / cdt, kind, "first 5000 trials x first 4 sets", ms=dt * 1e3,
             h2d_bytes=int(v_trial.nbytes + x.nbytes + params.nbytes),
             algorithmic_GBps=16.0 * T * P / dt / 1e9)

    if on("regression_design"):
        T, P, K = 100000, 512, 3
        x = rts(rng, T)
        params = np.concatenate([draw(rng, P), np.full((1, P), 0.05), np.full((1, P), 0.1)])
        design = np.concatenate([np.ones((1, T)), rng.standard_normal((K - 1, T))])
        coef = np.concatenate([params[0][None, :], rng.uniform(-0.5, 0.5, (K - 1, P))])
        nm = dict(err=1e-4, n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3)
        dt, lp = best_of(lambda: wfpt.wiener_like_regression_batch(x, design, coef, params, **nm), args.reps)
        vs = np.ascontiguousarray((design[:, :5000].T @ coef[:, :4]).T)
        t0 = time.perf_counter()
        for p in range(4):
            W.wiener_like_multi(x[:5000], vs[p], *params[1:7, p], 1e-4, ['v'], 2, 2, 1, 1e-3, 0.05, 0.1)
        cdt = time.perf_counter() - t0
        emit("regression sweep, drift built on the device from a [%d, n] design: %d trials x %d coefficient sets"
             % (K, T, P), T * P / dt, "evals/s", 5000 * 4 / cdt, kind, "first 5000 trials x first 4 sets",
             ms=dt * 1e3, h2d_bytes=int(design.nbytes + x.nbytes + params.nbytes + coef.nbytes),
             algorithmic_GBps=16.0 * T * P / dt / 1e9)

    if on("regression_resident"):
        # C3 as an MCMC over coefficients sees it: trials + design resident (wfpt.Regression), one call per batch of
        # proposals; outcome v (lean evaluators with a drift per trial) and outcome a (table-driven walk)
        T, P, K = 100000, 512, 3
        x = rts(rng, T)
        params = np.concatenate([draw(rng, P), np.full((1, P), 0.05), np.full((1, P), 0.1)])
        design = np.concatenate([np.ones((1, T)), rng.standard_normal((K - 1, T))])
        nm = dict(err=1e-4, n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3)
        for outcome, centre, spread in (("v", params[0], 0.5), ("a", params[2], 0.05)):
            coef = np.concatenate([centre[None, :], rng.uniform(-spread, spread, (K - 1, P))])
            reg = wfpt.Regression(x, {outcome: design})
            dt, lp = best_of(lambda: reg.like_batch({outcome: coef}, params, **nm), args.reps)
            cols = np.ascontiguousarray((design[:, :5000].T @ coef[:, :4]).T)
            t0 = time.perf_counter()
            for p in range(4):
                a7 = [cols[p] if n == outcome else params[k, p] for k, n in enumerate(("v", "sv", "a", "z", "sz", "t", "st"))]
                W.wiener_like_multi(x[:5000], *a7, 1e-4, [outcome], 2, 2, 1, 1e-3, 0.05, 0.1)
            cdt = time.perf_counter() - t0
            emit("regression %s ~ 1 + x1 + x2, data resident: %d trials x %d coefficient proposals" % (outcome, T, P),
                 T * P / dt, "evals/s", 5000 * 4 / cdt, kind, "first 5000 trials x first 4 proposals", ms=dt * 1e3,
                 h2d_bytes=int(params.nbytes + coef.nbytes), design_read_GBps=8.0 * (K + 1) * T * P / dt / 1e9)
            reg.close()

    if on("nodes"):
        # C2-shaped node batches: 20 observed nodes x 200 trials, P proposals per launch (what the sampler issues)
        from hddm_b200 import likelihoods as L
        S, Tn = 20, 200
        nodes = L.BatchedNodes([rts(rng, Tn, 0.45) for _ in range(S)])
        for per_node in (1, 16, 205, 4096):
            P = S * per_node
            params = np.concatenate([draw(rng, P), np.full((1, P), 0.05), np.full((1, P), 0.1)])
            set_node = np.tile(np.arange(S, dtype=np.int64), per_node)
            dt, lp = best_of(lambda: nodes.logp(set_node, params), max(args.reps, 5))
            emit("node batch: %d nodes x %d trials x %d proposals each (%d sets)" % (S, Tn, per_node, P), Tn * P / dt,
                 "evals/s", None, kind, "-", ms=dt * 1e3)
        for T2, P2 in ((1000, 10000),):
            x = rts(rng, T2)
            params = np.concatenate([draw(rng, P2), np.full((1, P2), 0.05), np.full((1, P2), 0.1)])
            dt, lp = best_of(lambda: wfpt.wiener_like_batch(x, params, **HDDM), args.reps)
            emit("short segments: %d trials x %d sets" % (T2, P2), T2 * P2 / dt, "evals/s", None, kind, "-", ms=dt * 1e3)

    if on("ppc"):
        D, S = 10000, 200
        params = draw(rng, D)
        dt, st = best_of(lambda: posterior.gen_rts_from_cdf_batch(params, S, -6, 6, 1e-4, seed=1, return_rts=False,
                                                                  return_stats=True), args.reps)
        nref = 8
        np.random.seed(0)
        t0 = time.perf_counter()
        for d in range(nref):
            W.gen_rts_from_cdf(*params[:, d], S, -6, 6, 1e-4)
        cdt = time.perf_counter() - t0
        emit("posterior predictive: %d draws x %d trials, cdf table dt=1e-4 on [-6,6), 15 statistics per draw" % (D, S),
             D / dt, "draws/s", nref / cdt, kind, "first %d draws (gen_rts_from_cdf only, no statistics)" % nref,
             ms=dt * 1e3, table_points=int(120000 * D), d2h_bytes=int(D * 15 * 8))

    if on("cdf"):
        D, n = 10000, 11
        par7 = draw(rng, D)
        par7[4] = np.maximum(par7[4], 0.02); par7[6] = np.maximum(par7[6], 0.02)
        params = np.concatenate([par7, np.full((1, D), 0.05), np.full((1, D), 0.1)])
        x = np.ascontiguousarray(((par7[5] + par7[6] / 2 + rng.gamma(2.0, 0.3, (n, D))).T *
                                  np.where(rng.random((D, n)) < 0.5, 1.0, -1.0)).clip(-4.9, 4.9))
        dt, y = best_of(lambda: posterior.dmat_cdf_batch(x, params), args.reps)
        nref = 200
        t0 = time.perf_counter()
        for d in range(nref):
            CW.dmat_cdf_array(x[d], *params[:, d])
        cdt = time.perf_counter() - t0
        emit("cdfdif quantile probabilities: %d parameter sets x %d edges" % (D, n), D * n / dt, "cdf evals/s",
             nref * n / cdt, kind, "first %d sets" % nref, ms=dt * 1e3)

    if on("rlddm"):
        n, P = 2000, 4096
        split_by = rng.integers(0, 4, n).astype(np.int64)
        response = rng.integers(0, 2, n).astype(np.int64)
        feedback = (rng.random(n) < np.where(response == 1, 0.8, 0.3)).astype(np.float64)
        x = (0.45 + rng.gamma(2.0, 0.25, n)) * np.where(response == 1, 1.0, -1.0)
        p7 = draw(rng, P); p7[0] = np.abs(p7[0]) + 0.5; p7[5] = rng.uniform(0.1, 0.3, P); p7[6] = rng.uniform(0, 0.2, P)
        params = np.concatenate([p7, np.full((1, P), 0.05), np.full((1, P), 0.1), np.full((1, P), 0.5),
                                 rng.normal(0, 1.5, (1, P)), rng.normal(0, 1.5, (1, P))])
        dt, lp = best_of(lambda: wfpt.wiener_like_rlddm_batch(x, response, feedback, split_by, params, **HDDM), args.reps)
        nref = 8
        t0 = time.perf_counter()
        for p in range(nref):
            W.wiener_like_rlddm(x, response, feedback, split_by, params[9, p], params[10, p], params[11, p],
                                *params[:7, p], 1e-4, 2, 2, 1, 1e-3, 0.05, 0.1)
        cdt = time.perf_counter() - t0
        emit("RL-DDM likelihood: %d trials (4 conditions) x %d proposals" % (n, P), n * P / dt, "evals/s",
             n * nref / cdt, kind, "first %d proposals" % nref, ms=dt * 1e3)

    if on("simulator"):
        D, S = 1000, 200
        params = draw(rng, D)
        dt, r = best_of(lambda: posterior.simulate_drift_batch(params, S, dt=1e-4, seed=2), args.reps)
        steps = float(np.nansum(np.abs(r) - np.repeat(params[5][:, None], S, 1)) / 1e-4)
        emit("random-walk simulator: %d sets x %d trials at dt=1e-4" % (D, S), D * S / dt, "trials/s",
             None, "unavailable", "hddm/generate.py needs kabuki/pymc to import; its pure-Python loop makes "
             "~1000-step numpy blocks per trial (order 1e3 trials/s/core)", ms=dt * 1e3,
             approx_steps_per_s=steps / dt)


if __name__ == "__main__":
    main()
