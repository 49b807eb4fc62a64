#!/usr/bin/env python
"""bench.py -- full-DDM WFPT log-density throughput (evals/s, 1 eval = one trial x parameter set).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--trials T] [--sets P] [--sweep A|B] [--no-cpu] [--no-sub]

Workload (BASELINE.json configs[3]): batched likelihood sweep, T trials x P
parameter sets per GPU, full DDM (sv, sz, st), HDDM numerics err=1e-4 n_st=n_sz=2 adaptive
simps_err=1e-3 p_outlier=.05 w_outlier=.1, synthetic RTs (sweep S-A: |rt| = 0.70 + Gamma(2,.25)).
A step is one pass of the hot path over the whole T x P batch.  Weak scaling: every rank owns P
sets (global batch = N*P sets) and the per-set log-likelihoods are all-gathered over NCCL inside
the timed region.

Prints ONE JSON line (rank 0):
  value        device-resident throughput (CUDA events, L2 flushed between steps, max over ranks)
  e2e          the same through the public host-buffer API (hddm_b200.wfpt.wiener_like_batch: H2D of
               RTs + parameters and D2H of the log-likelihoods inside the timed region)
  roofline     EXECUTED work of the likelihood kernel against the FP64-FMA issue peak measured in this
               run: frac = FP64-pipe thread-instructions per second / (peak FMA per second), <= 1.  The
               executed instruction count per pdf_sv comes from the ncu capture of THIS kernel source
               (profiles/ncu_fp64.json, keyed by a hash of hddm_b200/csrc); with a stale capture frac is
               null.  The SURVEY.md 8(d) figure (the reference's operation sequence, 300 FLOP per pdf_sv)
               is kept as achieved_reference_accounting.
  parity       after timing: 16 random sets of the headline launch itself, at full T, against the
               reference's own compiled code (oracle/_ref; the C port if absent) on the host
  cpu_baseline the reference's own Cython path timed on the host cores on a bounded sample
  strong       BASELINE config C4 as worded: P sets GLOBAL, sharded over the N GPUs (strong scaling)
  mcmc         "MCMC it/s" of the metric: BASELINE config C2 (20 subjects x 200 trials, sv sz st), 1 and
               16 chains, same sampler on the CUDA likelihood and on the reference's CPU likelihood
  ppc          BASELINE config C5: posterior predictive draws/s (1e4 draws x 200 trials) over the N GPUs
--impl reference times only the reference's CPU path (and its MCMC arm).
"""
import argparse
import hashlib
import json
import math
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

HDDM = dict(err=1e-4, n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3)
P_OUT, W_OUT = 0.05, 0.1
FLOP_PER_PDFSV = 300.0   # fixed accounting of the REFERENCE's operation sequence
FLOP_PER_EVAL_TAIL = 28.0
SETS_PER_CORE = 64       # CPU sample: 20000 trials x 64 sets per core ~ 8-10 s of work per core
METRIC = "full-DDM wfpt log-density evals/sec (trial x param)"
UNIT = "evals/s"
PARITY_SETS = 16


def sweep_inputs(T, P, onset, seed):
    """SURVEY.md 8(d) draw order: sv, sz, st, z, v, t, a, then the gamma RTs and their signs."""
    rng = np.random.default_rng(seed)
    sv = rng.uniform(0, 2.5, P); sz = rng.uniform(0, 0.4, P); st = rng.uniform(0, 0.35, P)
    z = rng.uniform(0.4, 0.6, P); v = rng.uniform(-4, 4, P); t = rng.uniform(0.2, 0.5, P)
    a = rng.uniform(0.5, 2.0, P)
    g = rng.gamma(2.0, 0.25, T)
    sgn = np.where(rng.random(T) < 0.6, 1.0, -1.0)
    x = (onset + g) * sgn
    params = np.stack([v, sv, a, z, sz, t, st, np.full(P, P_OUT), np.full(P, W_OUT)])
    return np.ascontiguousarray(x), np.ascontiguousarray(params)


def csrc_sha():
    """Hash of the kernel sources: ties an ncu capture under profiles/ to the code it measured."""
    h = hashlib.sha1()
    d = os.path.join(ROOT, "hddm_b200", "csrc")
    for name in sorted(os.listdir(d)):
        if name.endswith((".cu", ".cuh")):
            h.update(name.encode())
            h.update(open(os.path.join(d, name), "rb").read())
    return h.hexdigest()[:16]


# ---------------------------------------------------------------------------------------------
# CPU arm: the reference's own implementation on the host cores (oracle/_ref), else the C port.

def _cpu_like(kind):
    if kind == "reference":
        from oracle import ref
        W = ref.wfpt()
    else:
        from oracle import oracle as W
    return lambda x, params, p: W.wiener_like(x, *params[:7, p], HDDM["err"], HDDM["n_st"], HDDM["n_sz"],
                                              HDDM["use_adaptive"], HDDM["simps_err"], params[7, p], params[8, p])


def _cpu_worker(job):
    kind, x, params, lo, hi = job
    sys.path.insert(0, ROOT)
    fn = _cpu_like(kind)
    t0 = time.perf_counter()
    out = [fn(x, params, p) for p in range(lo, hi)]
    return time.perf_counter() - t0, out


def _cpu_count_worker(job):
    """pdf_sv evaluations the reference's recursion makes (counted by the C port) for sets lo..hi."""
    x, params, lo, hi = job
    sys.path.insert(0, ROOT)
    from oracle import oracle as O
    O.pdfsv_calls(reset=True)
    for p in range(lo, hi):
        O.wiener_like(x, *params[:7, p], HDDM["err"], HDDM["n_st"], HDDM["n_sz"], HDDM["use_adaptive"],
                      HDDM["simps_err"], params[7, p], params[8, p])
    return int(O.pdfsv_calls(reset=True))


def cpu_kind():
    from oracle import ref
    if ref.available():
        try:
            ref.wfpt()
            return "reference"
        except Exception:
            pass
    from oracle import oracle as O
    O.build()
    return "port"


def host_cores():
    return len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else os.cpu_count()


class CpuArm(object):
    """Times wiener_like of the CPU implementation on `cores` processes over a bounded sample:
    the first t_sample trials x (sets_per_core * cores) parameter sets of the bench workload."""

    def __init__(self, x, params, t_sample, sets_per_core, cores=None):
        import multiprocessing as mp
        self.kind = cpu_kind()
        self.cores = cores or host_cores()
        self.x = np.ascontiguousarray(x[:t_sample])
        self.sets = min(params.shape[1], sets_per_core * self.cores)
        self.params = np.ascontiguousarray(params[:, :self.sets])
        self.pool = mp.get_context("spawn").Pool(self.cores)
        self.sample = "first %d trials x first %d parameter sets of the workload, %d sets per process" % (
            self.x.size, self.sets, int(math.ceil(self.sets / self.cores)))

    def step(self, sets=None):
        sets = self.sets if sets is None else min(sets, self.sets)
        per = int(math.ceil(sets / self.cores))
        jobs = [(self.kind, self.x, self.params, lo, min(lo + per, sets)) for lo in range(0, sets, per)]
        t0 = time.perf_counter()
        res = self.pool.map(_cpu_worker, jobs)
        wall = time.perf_counter() - t0
        return wall, self.x.size * sets, res

    def loglik_sets(self, x, params):
        """wiener_like of every column of params against x, one set per job (the in-bench parity check)."""
        P = params.shape[1]
        res = self.pool.map(_cpu_worker, [(self.kind, x, params, p, p + 1) for p in range(P)])
        return np.array([r[1][0] for r in res])

    def pdfsv_counts(self, x, params):
        return sum(self.pool.map(_cpu_count_worker, [(x, params, p, p + 1) for p in range(params.shape[1])]))

    def close(self):
        self.pool.close()
        self.pool.join()


# ---------------------------------------------------------------------------------------------
# MCMC it/s (BASELINE config C2): the same batched slice-within-Gibbs sampler on both likelihoods.

MCMC_CFG = {"workload": "C2 hierarchical HDDM 20 subjects x 200 trials, include=(sv,sz,st), p_outlier=.05, HDDM "
                        "numerics; slice-within-Gibbs, 15 stochastic families per iteration (hddm_b200/sampler.py; "
                        "PyMC 2.3 / kabuki are not installable: reference it/s through PyMC unavailable on this box)"}


def mcmc_data(S=20, T=200, seed=20261104, dt=1e-3):
    """Synthetic subjects drawn from the full DDM itself by an Euler-Maruyama walk in numpy (per trial: drift
    ~ N(v, sv), start point ~ U(z +- sz/2) a, non-decision time ~ U(t +- st/2)), so that neither the GPU nor
    the oracle is needed to make the data and both arms see the same vectors."""
    rng = np.random.default_rng(seed)
    rts = []
    for _ in range(S):
        v, a, t = rng.normal(1.0, 0.5), rng.uniform(1.2, 2.0), rng.uniform(0.25, 0.4)
        sv, sz, st = 0.8, 0.2, 0.15
        drift = rng.normal(v, sv, T)
        y = (0.5 + sz * (rng.random(T) - 0.5)) * a
        rt = np.zeros(T)
        alive = np.ones(T, dtype=bool)
        k = 0
        while alive.any() and k < 20000:
            k += 1
            y = np.where(alive, y + drift * dt + math.sqrt(dt) * rng.standard_normal(T), y)
            hit = alive & ((y <= 0) | (y >= a))
            rt[hit] = k * dt
            alive &= ~hit
        sign = np.where(y >= a, 1.0, -1.0)
        rts.append(np.ascontiguousarray((rt + t + st * (rng.random(T) - 0.5)) * sign))
    return rts


def _mcmc_cpu_worker(job):
    kind, rts, set_node, params, lo, hi = job
    sys.path.insert(0, ROOT)
    fn = _cpu_like(kind)
    return [fn(rts[set_node[p]], params, p) for p in range(lo, hi)]


def mcmc_cpu(pool, kind, cores, rts, chains, iters):
    """Same sampler, reference likelihood: one wiener_like call per observed node as PyMC makes them
    (hddm/likelihoods.py:54-55), the sets of a proposal round spread over all host cores."""
    from hddm_b200 import sampler

    def loglik(set_node, params):
        P = params.shape[1]
        per = int(math.ceil(P / cores))
        out = pool.map(_mcmc_cpu_worker, [(kind, rts, set_node, params, lo, min(lo + per, P)) for lo in range(0, P, per)])
        return np.concatenate([np.asarray(o, dtype=np.float64) for o in out])

    m = sampler.HierarchicalDDM(len(rts), loglik, chains=chains, seed=1)
    m.sample(1, trace=False)
    _, dt = m.sample(iters, trace=False)
    return {"chains": chains, "iters": iters, "it_per_s_per_chain": iters / dt, "chain_it_per_s": chains * iters / dt,
            "cores": cores, "kind": kind}


def mcmc_gpu(rts, chains, iters, native=True):
    """native: the C++ driver (csrc/abi_mcmc.cu) -- the same chain as the Python sampler draws, without Python
    between the launches; else hddm_b200/sampler.py with speculative batches."""
    from hddm_b200 import likelihoods as L, sampler
    nodes = L.BatchedNodes(rts)
    T = rts[0].shape[0]
    m = sampler.HierarchicalDDM(len(rts), nodes.logp, chains=chains, seed=1)
    run = (lambda n: m.sample_native(nodes, n, trace=False)) if native else (lambda n: m.sample(n, trace=False))
    run(5)  # warm-up (uploads + sorts the trials once)
    calls0 = m.loglik_calls
    _, dt = run(iters)
    return {"chains": chains, "iters": iters, "it_per_s_per_chain": iters / dt, "chain_it_per_s": chains * iters / dt,
            "launches_per_iter": (m.loglik_calls - calls0) / iters, "driver": "native" if native else "python"}


def mcmc_reference_record(pool, kind, cores, iters=(10, 2)):
    rts = mcmc_data()
    return {"config": MCMC_CFG,
            "cpu_1_chain": mcmc_cpu(pool, kind, cores, rts, 1, iters[0]),
            "cpu_16_chains": mcmc_cpu(pool, kind, cores, rts, 16, iters[1])}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    onset = 0.70 if args.sweep == "A" else 0.30
    x, params = sweep_inputs(args.trials, args.sets, onset, 20261104)
    arm = CpuArm(x, params, t_sample=min(args.trials, 20000), sets_per_core=SETS_PER_CORE)
    for _ in range(max(args.warmup, 1)):
        arm.step(sets=arm.cores)  # spins the worker processes up; one set each
    total_t, total_e = 0.0, 0
    for _ in range(args.steps):
        w, e, _ = arm.step()
        total_t += w
        total_e += e
    mcmc = None
    if not args.no_sub:
        try:
            mcmc = mcmc_reference_record(arm.pool, arm.kind, arm.cores)
        except Exception as ex:
            mcmc = {"error": repr(ex)}
    arm.close()
    value = total_e / total_t
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total_t / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args, 1),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": arm.cores, "kind": arm.kind, "sample": arm.sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "mcmc": mcmc,
    }
    print(json.dumps(line))
    return 0


def workload_config(args, n):
    return {"workload": "C4 batched likelihood sweep S-%s: %d trials x %d parameter sets per GPU, full DDM "
                        "(sv,sz,st), HDDM numerics err=1e-4 n_st=n_sz=2 adaptive simps_err=1e-3 p_outlier=.05 "
                        "w_outlier=.1" % (args.sweep, args.trials, args.sets),
            "trials": args.trials, "sets_per_gpu": args.sets, "global_sets": args.sets * n, "sweep": args.sweep,
            "sharding": "parameter sets over %d GPU(s), all_gather of logp" % n,
            "l2": "L2 flushed (256 MiB write) between timed steps; path is FP64-bound, inputs 0.9 MB"}


# ---------------------------------------------------------------------------------------------

class ClockSampler(threading.Thread):
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.rows = []
        self._stop_evt = threading.Event()

    def run(self):
        while not self._stop_evt.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                for ln in out.strip().splitlines():
                    self.rows.append([c.strip() for c in ln.split(",")])
            except Exception:
                pass
            self._stop_evt.wait(0.2)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=6)
        sm, mx, reasons = [], 0.0, set()
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx = max(mx, float(r[2]))
            except Exception:
                continue
            for name, col in (("hw_slowdown", 5), ("hw_thermal_slowdown", 6), ("sw_thermal_slowdown", 7), ("sw_power_cap", 8)):
                if len(r) > col and r[col].lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None,
                "reasons": sorted(reasons), "samples": len(sm)}


def split_sweep(T, P):
    """launch_like's choice (wfpt_kernels.cu): big sweeps run the two passes as two launches."""
    return T > 6144 and P * math.ceil(T / 8192.0) >= 2368


def roofline_record(args, T, P, k_ms, pdfsv_per_step, peak_tflops):
    """Executed-work roofline of the likelihood kernel (see the module docstring)."""
    alg_flop = float(T) * P * FLOP_PER_EVAL_TAIL + pdfsv_per_step * FLOP_PER_PDFSV
    ref_acc = alg_flop / (k_ms * 1e-3) / 1e12
    sha = csrc_sha()
    cap, stale = None, None
    try:
        caps = json.load(open(os.path.join(ROOT, "profiles", "ncu_fp64.json")))
        cap = caps.get("%s:%dx%d" % (args.sweep, T, P))
    except Exception:
        caps = None
    rec = {"bound": "fp64", "unit": "TFLOP/s", "peak": peak_tflops, "achieved": None, "frac": None, "traffic": None,
           "peak_source": "FP64 FMA probe (k_fp64_peak, 8 independent chains per thread) measured in this run; "
                          "MEASURED_PEAKS.json has no FP64 entry",
           "kernel": "k_like_partial first pass + deferred-trial pass (+ k_build_tables, k_like_final)",
           "kernel_ms": k_ms, "pdf_sv_per_eval": pdfsv_per_step / (float(T) * P),
           "achieved_reference_accounting": ref_acc,
           "reference_accounting": "pdf_sv evals x 300 + 28 FLOP per eval (SURVEY.md 8d: the REFERENCE's operation "
                                   "sequence, not what the kernel executes; not a roofline fraction)",
           "algorithmic_bytes": int(8 * T + 72 * P),
           "scratch_bytes": int(8 * P * math.ceil(T / 2048.0) +
                                ((2 * 1024 + 3072) * P * math.ceil(T / 8192.0) +
                                 3072 * P * (1 + math.ceil(math.ceil(T / 8192.0) / 8.0)) if split_sweep(T, P) else 0)),
           "csrc_sha": sha}
    if cap:
        stale = cap.get("csrc_sha") != sha
        inst = float(cap["fp64_thread_inst"])                      # per step of this workload, from ncu
        per_pdfsv = inst / float(cap["pdf_sv_evals"])
        flop = float(cap["fp64_flop"]) / float(cap["pdf_sv_evals"])   # FMA = 2, DMUL / DADD / DSETP .. = 1
        issue_rate = per_pdfsv * pdfsv_per_step / (k_ms * 1e-3)       # FP64-pipe thread-instructions per second
        frac = issue_rate / (peak_tflops * 1e12 / 2.0)
        rec.update({
            "executed_fp64_inst_per_pdf_sv": per_pdfsv, "fused_share": cap.get("fused_share"),
            "executed_flop_per_pdf_sv": flop,
            "achieved_executed_flops": flop * pdfsv_per_step / (k_ms * 1e-3) / 1e12,
            "traffic": cap.get("dram_bytes"), "capture": cap.get("source"), "capture_csrc_sha": cap.get("csrc_sha"),
            "capture_stale": stale,
            "fp64_pipe_active_pct_ncu": cap.get("fp64_pipe_pct"),
            "accounting": "achieved = executed FP64-pipe thread-instructions/s x 2 (every FP64 instruction holds an "
                          "FMA issue slot) against the FMA peak; instruction count per pdf_sv from the ncu capture "
                          "named in `capture`, pdf_sv evals counted by the kernel in this run"})
        if not stale:
            rec["achieved"] = issue_rate * 2.0 / 1e12
            rec["frac"] = frac
        else:
            rec["frac_from_stale_capture"] = frac
    else:
        rec["accounting"] = "no ncu capture of this workload under profiles/ncu_fp64.json: executed fraction unknown"
    return rec


def run_ours(args):
    import torch
    import torch.distributed as dist
    from hddm_b200 import _lib, wfpt
    from hddm_b200.batch import DeviceSweep, gather_loglik, shard_bounds

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the hddm_b200 path has no CPU fallback")
    # stdout carries exactly ONE JSON line: anything native libraries print there meanwhile (NCCL's version
    # banner at NCCL_DEBUG=VERSION/WARN) is sent to stderr; fd 1 is restored for the final print
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    _lib.check(_lib.lib().hddm_b200_init(local))

    onset = 0.70 if args.sweep == "A" else 0.30
    T, P = args.trials, args.sets
    # every rank sees the same trials; rank r owns the r-th block of P parameter sets
    x, params0 = sweep_inputs(T, P, onset, 20261104)
    params = params0 if rank == 0 else sweep_inputs(1, P, onset, 20261104 + rank)[1]
    n_global = world * P

    # FP64 roofline denominator, measured now on this GPU
    import ctypes as C
    fl, ms = C.c_double(), C.c_float()
    _lib.check(_lib.lib().hddm_b200_fp64_peak(8192, C.byref(fl), C.byref(ms)))
    peak_tflops = fl.value / 1e12

    sweep = DeviceSweep(x, device=dev)
    d_params = torch.from_numpy(params).to(dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(step_fn, steps, warmup):
        """W warm-up steps, then `steps` timed ones: L2 flush + barrier outside, CUDA events on the launching
        stream inside; returns (sum of step ms as the max over ranks, per-step kernel ms, last result)."""
        for _ in range(warmup):
            out = step_fn(None)
        barrier()
        step_ms, kern_ms = [], []
        for _ in range(steps):
            flush.fill_(1)  # L2 flush between timed iterations (outside the timed region)
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            out = step_fn((k0, k1))
            e1.record()
            barrier()
            step_ms.append(e0.elapsed_time(e1))
            kern_ms.append(k0.elapsed_time(k1))
        t = torch.tensor([sum(step_ms)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()), kern_ms, out

    def one_step(record):
        if record is not None:
            record[0].record()
        lp = sweep.loglik(d_params, **HDDM)
        if record is not None:
            record[1].record()
        return gather_loglik(lp, n_global)

    warm = max(args.warmup, 3)
    for _ in range(warm):
        one_step(None)
    barrier()
    sweep.evals.zero_()
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    total_ms, kern_ms, out = timed(one_step, args.steps, 0)
    pdfsv = int(sweep.evals.item())
    clocks = sampler.stop() if sampler else None

    evals_per_step = float(T) * P * world
    value = evals_per_step * args.steps / (total_ms * 1e-3)
    # every rank's own kernel time per step (weak scaling: each rank has its own 1e4 sets; the step ends with the
    # slowest rank's all-gather contribution)
    k_rank = torch.tensor([float(np.mean(kern_ms))], dtype=torch.float64, device=dev)
    if world > 1:
        k_ranks = [torch.zeros_like(k_rank) for _ in range(world)]
        dist.all_gather(k_ranks, k_rank)
        k_rank = torch.cat(k_ranks)
    kernel_ms_per_rank = [round(float(v), 4) for v in k_rank.cpu().tolist()]

    # ---- e2e: public host-buffer API, pinned host inputs, H2D + D2H inside the timed region
    xh = torch.from_numpy(x).pin_memory().numpy()
    ph = torch.from_numpy(params).pin_memory().numpy()
    e2e_steps = max(1, min(args.steps, 3))
    wfpt.wiener_like_batch(xh, ph, **HDDM)
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        lp_host = wfpt.wiener_like_batch(xh, ph, **HDDM)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = evals_per_step * e2e_steps / float(t.item())
    headline_local = out[rank * P:(rank + 1) * P].cpu().numpy()
    same = bool(np.array_equal(lp_host, headline_local))

    # ---- strong scaling of C4 as BASELINE.json words it: P sets GLOBAL, sharded over the ranks
    strong = None
    if not args.no_sub:
        lo, hi = shard_bounds(P, world, rank)
        d_block = torch.from_numpy(np.ascontiguousarray(params0[:, lo:hi])).to(dev)

        def strong_step(record):
            if record is not None:
                record[0].record()
            lp = sweep.loglik(d_block, **HDDM)
            if record is not None:
                record[1].record()
            return gather_loglik(lp, P)

        s_ms, s_kern, s_out = timed(strong_step, args.steps, 3)
        # every rank's own kernel time: the collective ends with the slowest rank, so the spread between ranks
        # (sets differ in cost) is part of the step
        k_all = torch.tensor([float(np.mean(s_kern))], dtype=torch.float64, device=dev)
        if world > 1:
            k_list = [torch.zeros_like(k_all) for _ in range(world)]
            dist.all_gather(k_list, k_all)
            k_all = torch.cat(k_list)
        strong = {"value": float(T) * P * args.steps / (s_ms * 1e-3), "unit": UNIT, "scaling": "strong",
                  "global_sets": P, "sets_per_gpu": hi - lo, "ms_per_step": s_ms / args.steps,
                  "kernel_ms_rank0": float(np.mean(s_kern)),
                  "kernel_ms_per_rank": [round(float(v), 4) for v in k_all.cpu().tolist()],
                  # the sharded result is rank 0's own single-GPU vector bit for bit, sets computed on other
                  # GPUs included (each set is summed on one GPU in a fixed order)
                  "equals_single_gpu_result": bool(np.array_equal(s_out.cpu().numpy(), headline_local))}

    # ---- C5 posterior predictive over the ranks (draws sharded, one all_gather of [D, 15])
    ppc = None
    if not args.no_sub:
        from hddm_b200 import posterior
        rng = np.random.default_rng(20261104)
        D, S = 10000, 200
        pp = np.stack([rng.uniform(-4, 4, D), rng.uniform(0, 2.5, D), rng.uniform(0.5, 2.0, D), rng.uniform(0.4, 0.6, D),
                       rng.uniform(0, 0.4, D), rng.uniform(0.2, 0.5, D), rng.uniform(0, 0.35, D)])
        posterior.post_pred_stats_sharded(pp, S, -6, 6, 1e-4, seed=1)
        best, stages = None, None
        for _ in range(3):
            barrier()
            tm = {}
            t0 = time.perf_counter()
            table = posterior.post_pred_stats_sharded(pp, S, -6, 6, 1e-4, seed=1, timings=tm)
            t_call = time.perf_counter() - t0
            barrier()
            dt = time.perf_counter() - t0
            if best is None or dt < best:
                best = dt
                stages = {"evaluate_ms_rank0": 1e3 * tm["evaluate"], "gather_ms_rank0": 1e3 * tm["gather"],
                          "closing_barrier_ms_rank0": 1e3 * (dt - t_call)}
        # the other half of C5: cdfdif quantile probabilities of the same draws (11 edges each), sharded the same way
        n_edges = 11
        p9 = np.concatenate([pp, np.full((1, D), P_OUT), np.full((1, D), W_OUT)])
        p9[4] = np.maximum(p9[4], 0.02)
        p9[6] = np.maximum(p9[6], 0.02)
        edges = np.ascontiguousarray(((p9[5] + p9[6] / 2 + rng.gamma(2.0, 0.3, (n_edges, D))).T *
                                      np.where(rng.random((D, n_edges)) < 0.5, 1.0, -1.0)).clip(-4.9, 4.9))
        posterior.dmat_cdf_sharded(edges, p9)
        cdf_best = None
        for _ in range(3):
            barrier()
            t0 = time.perf_counter()
            probs = posterior.dmat_cdf_sharded(edges, p9)
            barrier()
            dtc = time.perf_counter() - t0
            cdf_best = dtc if cdf_best is None else min(cdf_best, dtc)
        t = torch.tensor([best, cdf_best], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        cdf = {"value": D * n_edges / float(t[1].item()), "unit": "cdf evals/s", "scaling": "strong",
               "ms": 1e3 * float(t[1].item()), "sets": D, "edges_per_set": n_edges,
               "table_sha1": hashlib.sha1(np.nan_to_num(probs).tobytes()).hexdigest()}
        t = t[:1]
        ppc = {"value": D / float(t.item()), "unit": "draws/s", "scaling": "strong", "ms": 1e3 * float(t.item()),
               "config": {"workload": "C5: %d posterior draws x %d trials, gen_rts_from_cdf table dt=1e-4 on [-6,6) + "
                                      "15 gen_ppc_stats statistics per draw, host parameters in, host table out" % (D, S)},
               "stages": stages, "table_sha1": hashlib.sha1(np.nan_to_num(table).tobytes()).hexdigest(),
               "cdf_quantiles": cdf}

    line = None
    if rank == 0:
        k_ms = float(np.mean(kern_ms))
        pdfsv_per_step = pdfsv / float(args.steps)
        cpu, parity, mcmc = None, None, None
        try:
            if args.no_cpu:
                raise RuntimeError("skipped (--no-cpu)")
            arm = CpuArm(x, params, t_sample=min(T, 20000), sets_per_core=SETS_PER_CORE)
            arm.step(sets=arm.cores)
            w, e, res = arm.step()
            cpu = {"value": e / w, "unit": UNIT, "cores": arm.cores, "kind": arm.kind, "sample": arm.sample}
            # ---- parity of the HEADLINE launch itself: random sets at full T against the reference's code
            try:
                idx = np.sort(np.random.default_rng(5).choice(P, size=min(PARITY_SETS, P), replace=False))
                sub = np.ascontiguousarray(params[:, idx])
                want = arm.loglik_sets(x, sub)
                got = headline_local[idx]
                fin = np.isfinite(want)
                rel = float(np.max(np.abs(got[fin] - want[fin]) / np.abs(want[fin]))) if fin.any() else 0.0
                _, ev_gpu = wfpt.wiener_like_batch(x, sub, return_evals=True, **HDDM)
                ev_cpu = arm.pdfsv_counts(x, sub)
                parity = {"sets": int(idx.size), "trials": T, "against": arm.kind, "max_rel_sum": rel,
                          "tolerance": 1e-9, "ok": bool(rel <= 1e-9 and np.array_equal(got[~fin], want[~fin])),
                          "pdf_sv_evals_gpu": int(ev_gpu), "pdf_sv_evals_oracle": int(ev_cpu),
                          "evals_equal": bool(int(ev_gpu) == int(ev_cpu))}
            except Exception as ex:
                parity = {"error": repr(ex)}
            if not args.no_sub:
                try:
                    rts = mcmc_data()
                    mcmc = {"config": MCMC_CFG,
                            "gpu_1_chain": mcmc_gpu(rts, 1, 4 * args.mcmc_iters),
                            "gpu_16_chains": mcmc_gpu(rts, 16, args.mcmc_iters),
                            "gpu_1_chain_python_driver": mcmc_gpu(rts, 1, args.mcmc_iters, native=False),
                            "cpu_1_chain": mcmc_cpu(arm.pool, arm.kind, arm.cores, rts, 1, 10),
                            "cpu_16_chains": mcmc_cpu(arm.pool, arm.kind, arm.cores, rts, 16, 2)}
                except Exception as ex:
                    mcmc = {"error": repr(ex)}
            arm.close()
        except Exception as ex:  # the CPU arm is a reported baseline; never let it sink the GPU line
            cpu = {"value": None, "unit": UNIT, "cores": 0, "kind": "unavailable", "sample": repr(ex)}
            if not args.no_sub and mcmc is None:
                try:
                    rts = mcmc_data()
                    mcmc = {"config": MCMC_CFG, "gpu_1_chain": mcmc_gpu(rts, 1, 4 * args.mcmc_iters),
                            "gpu_16_chains": mcmc_gpu(rts, 16, args.mcmc_iters)}
                except Exception as ex2:
                    mcmc = {"error": repr(ex2)}
        launches = (4 if split_sweep(T, P) else 2) * args.steps
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": warm, "ms_per_step": total_ms / args.steps, "kernel_ms_per_rank": kernel_ms_per_rank,
            "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(args, world),
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(x.nbytes + params.nbytes),
                    "d2h_bytes_per_step": int(8 * P + 8), "steps": e2e_steps, "bitwise_equal_to_device_path": same},
            # big sweeps: per-set tables, first pass, deferred-trial pass, final sum; else 2
            "gpu_launches": launches,
            "roofline": roofline_record(args, T, P, k_ms, pdfsv_per_step, peak_tflops),
            "parity": parity,
            "cpu_baseline": cpu,
            "strong": strong, "mcmc": mcmc, "ppc": ppc,
            "clocks": clocks,
        }
        sys.stdout.flush()
        os.dup2(saved_stdout, 1)
        print(json.dumps(line))
        sys.stdout.flush()
        os.dup2(2, 1)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--trials", type=int, default=100000)
    ap.add_argument("--sets", type=int, default=10000)
    ap.add_argument("--sweep", default="A", choices=["A", "B"])
    ap.add_argument("--mcmc-iters", type=int, default=100)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline / parity / CPU-MCMC legs (profiling runs)")
    ap.add_argument("--no-sub", action="store_true", help="skip the strong / mcmc / ppc sub-records")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.gpus > 1 and world == 1:
        # convenience: re-launch under torchrun when called plainly with --gpus N
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(args.gpus),
               "--master-addr", "127.0.0.1", "--master-port", "29531", os.path.abspath(__file__)] + sys.argv[1:]
        return subprocess.call(cmd)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
