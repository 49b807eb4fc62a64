#!/usr/bin/env python
"""bench_mcmc.py -- MCMC iterations per second of the hierarchical full DDM (BASELINE.json configs[1]:
20 subjects x 200 trials, include=('sv','sz','st'), HDDM numerics), the second half of the metric.

The SAME batched slice-within-Gibbs sampler (hddm_b200/sampler.py) is run with
  (a) the CUDA likelihood (hddm_b200.likelihoods.BatchedNodes, one launch per proposal round), and
  (b) the reference's own compiled Cython wiener_like (oracle/_ref; the C port if absent), one call per
      observed node exactly as PyMC would make them (hddm/likelihoods.py:54-55), on one host core --
      PyMC 2.3 / kabuki are not installable here, so this is the closest measurable stand-in for the
      reference's sampling loop and it EXCLUDES PyMC's own per-step overhead (favouring the reference).
Prints one JSON line per arm.  An iteration updates every stochastic of the model once.
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


def make_data(S, T, seed):
    """Synthetic subjects from the generative model (parameters in the ranges of hddm/generate.py:38-46),
    RTs drawn with the GPU CDF sampler (src/wfpt.pyx:323-354 semantics)."""
    from hddm_b200 import wfpt
    rng = np.random.default_rng(seed)
    rts = []
    for s in range(S):
        v, a, t = rng.normal(1.0, 0.5), rng.uniform(1.2, 2.0), rng.uniform(0.25, 0.4)
        f, u = rng.random(T), rng.random(T)
        rts.append(wfpt.gen_rts_from_cdf(v, 0.8, a, 0.5, 0.2, t, 0.15, T, -6, 6, 1e-3, f=f, u_delay=u))
    return rts


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--subjects", type=int, default=20)
    ap.add_argument("--trials", type=int, default=200)
    ap.add_argument("--iters", type=int, default=200)
    ap.add_argument("--cpu-iters", type=int, default=5)
    ap.add_argument("--chains", type=int, nargs="+", default=[1, 16, 256])
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    from hddm_b200 import likelihoods as L, sampler
    S, T = args.subjects, args.trials
    rts = make_data(S, T, 20261104)
    nodes = L.BatchedNodes(rts)
    cfg = {"workload": "hierarchical HDDM %d subjects x %d trials, include=(sv,sz,st), p_outlier=.05, HDDM numerics; "
                       "slice-within-Gibbs, 15 stochastic families per iteration" % (S, T)}
    for C in args.chains:
        m = sampler.HierarchicalDDM(S, nodes.logp, chains=C, seed=1)
        m.sample(5, trace=False)  # warm-up (uploads + sorts the trials once)
        calls0, sets0 = m.loglik_calls, m.sets_evaluated
        _, dt = m.sample(args.iters, trace=False)
        print(json.dumps({"arm": "hddm_b200", "metric": "MCMC it/s", "chains": C, "iters": args.iters,
                          "it_per_s_per_chain": args.iters / dt, "chain_it_per_s": C * args.iters / dt,
                          "launches_per_iter": (m.loglik_calls - calls0) / args.iters,
                          "evals_per_s": (m.sets_evaluated - sets0) * T / dt, "config": cfg}), flush=True)
    if args.no_cpu:
        return 0
    from oracle import ref
    if ref.available():
        Wref, kind = ref.wfpt(), "reference"
    else:
        from oracle import oracle as Wref
        Wref.build()
        kind = "port"

    def cpu_loglik(set_node, params):
        return np.array([Wref.wiener_like(rts[set_node[p]], *params[:7, p], HDDM["err"], HDDM["n_st"], HDDM["n_sz"],
                                          HDDM["use_adaptive"], HDDM["simps_err"], params[7, p], params[8, p])
                         for p in range(params.shape[1])])

    m = sampler.HierarchicalDDM(S, cpu_loglik, chains=1, seed=1)
    m.sample(1, trace=False)
    _, dt = m.sample(args.cpu_iters, trace=False)
    print(json.dumps({"arm": "cpu_" + kind, "metric": "MCMC it/s", "chains": 1, "iters": args.cpu_iters,
                      "it_per_s_per_chain": args.cpu_iters / dt, "chain_it_per_s": args.cpu_iters / dt, "cores": 1,
                      "config": cfg}), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
