#!/usr/bin/env python
"""bench.py -- full-DDM WFPT log-density throughput (evals/s, 1 eval = one trial x parameter set).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--trials T] [--sets P] [--sweep A|B]

Workload (BASELINE.json configs[3]): batched likelihood sweep, T trials x P
parameter sets per GPU, full DDM (sv, sz, st), HDDM numerics err=1e-4 n_st=n_sz=2 adaptive
simps_err=1e-3 p_outlier=.05 w_outlier=.1, synthetic RTs (sweep S-A: |rt| = 0.70 + Gamma(2,.25)).
A step is one pass of the hot path over the whole T x P batch.  Weak scaling: every rank owns P
sets (global batch = N*P sets) and the per-set log-likelihoods are all-gathered over NCCL inside
the timed region.

Prints ONE JSON line (rank 0).  `value` = device-resident throughput; `e2e` = the same through the
public host-buffer API (hddm_b200.wfpt.wiener_like_batch: H2D of RTs+parameters and D2H of the
log-likelihoods inside the timed region); `roofline` = algorithmic FP64 FLOP/s of the likelihood
kernel against the FP64-FMA peak measured in this run; `cpu_baseline` = the reference's own Cython
path (oracle/_ref) or its C restatement timed on the host cores on a bounded sample.
--impl reference times only that CPU path.
"""
import argparse
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
FLOP_PER_PDFSV = 300.0   # fixed accounting of the reference's operation sequence
FLOP_PER_EVAL_TAIL = 28.0
SETS_PER_CORE = 64       # CPU sample: 20000 trials x 64 sets per core ~ 8-10 s of work per core
METRIC = "full-DDM wfpt log-density evals/sec (trial x param)"
UNIT = "evals/s"


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


# ---------------------------------------------------------------------------------------------
# CPU arm: the reference's own implementation on the host cores (oracle/_ref), else the C port.

def _cpu_worker(job):
    kind, x, params, lo, hi = job
    sys.path.insert(0, ROOT)
    if kind == "reference":
        from oracle import ref
        W = ref.wfpt()
        fn = lambda p: W.wiener_like(x, *params[:7, p], HDDM["err"], HDDM["n_st"], HDDM["n_sz"],
                                     HDDM["use_adaptive"], HDDM["simps_err"], params[7, p], params[8, p])
    else:
        from oracle import oracle as O
        fn = lambda p: O.wiener_like(x, *params[:7, p], HDDM["err"], HDDM["n_st"], HDDM["n_sz"],
                                     HDDM["use_adaptive"], HDDM["simps_err"], params[7, p], params[8, p])
    t0 = time.perf_counter()
    out = [fn(p) for p in range(lo, hi)]
    return time.perf_counter() - t0, out


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


class CpuArm(object):
    """Times wiener_like of the CPU implementation on `cores` processes over a bounded sample:
    the first t_sample trials x (sets_per_core * cores) parameter sets of the bench workload."""

    def __init__(self, x, params, t_sample, sets_per_core, cores=None):
        import multiprocessing as mp
        self.kind = cpu_kind()
        self.cores = cores or (len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else os.cpu_count())
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

    def close(self):
        self.pool.close()
        self.pool.join()


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
    arm.close()
    value = total_e / total_t
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total_t / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args, 1),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": arm.cores, "kind": arm.kind, "sample": arm.sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
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


def run_ours(args):
    import torch
    import torch.distributed as dist
    from hddm_b200 import _lib, wfpt
    from hddm_b200.batch import DeviceSweep, gather_loglik

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

    def one_step(record=None):
        if record is not None:
            record[0].record()
        lp = sweep.loglik(d_params, **HDDM)
        if record is not None:
            record[1].record()
        return gather_loglik(lp, n_global)

    for _ in range(max(args.warmup, 3)):
        out = one_step()
    barrier()
    sweep.evals.zero_()
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()

    step_ms, kern_ms = [], []
    total_ms = 0.0
    for _ in range(args.steps):
        flush.fill_(1)  # L2 flush between timed iterations (outside the timed region)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out = one_step((k0, k1))
        e1.record()
        barrier()
        step_ms.append(e0.elapsed_time(e1))
        kern_ms.append(k0.elapsed_time(k1))
    t = torch.tensor([sum(step_ms)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    pdfsv = int(sweep.evals.item())
    clocks = sampler.stop() if sampler else None

    evals_per_step = float(T) * P * world
    value = evals_per_step * args.steps / (total_ms * 1e-3)

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
    same = bool(np.array_equal(lp_host, out[rank * P:(rank + 1) * P].cpu().numpy()))

    line = None
    if rank == 0:
        k_ms = float(np.mean(kern_ms))
        mean_pdfsv = pdfsv / (float(T) * P * args.steps)
        alg_flop = float(T) * P * (mean_pdfsv * FLOP_PER_PDFSV + FLOP_PER_EVAL_TAIL)
        achieved = alg_flop / (k_ms * 1e-3) / 1e12
        traffic, traffic_src = None, None
        try:  # measured DRAM bytes per launch of this exact workload, if an ncu capture of it is on file
            tj = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
            ent = tj.get("%s:%dx%d" % (args.sweep, T, P))
            if ent:
                traffic, traffic_src = ent["bytes"], ent["source"]
        except Exception:
            pass
        cpu = None
        try:
            if args.no_cpu:
                raise RuntimeError("skipped (--no-cpu)")
            arm = CpuArm(x, params, t_sample=min(T, 20000), sets_per_core=SETS_PER_CORE)
            arm.step(sets=arm.cores)
            w, e, res = arm.step()
            arm.close()
            cpu = {"value": e / w, "unit": UNIT, "cores": arm.cores, "kind": arm.kind, "sample": arm.sample}
        except Exception as ex:  # the CPU arm is a reported baseline; never let it sink the GPU line
            cpu = {"value": None, "unit": UNIT, "cores": 0, "kind": "unavailable", "sample": repr(ex)}
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": total_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(args, world),
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(x.nbytes + params.nbytes),
                    "d2h_bytes_per_step": int(8 * P + 8), "steps": e2e_steps, "bitwise_equal_to_device_path": same},
            # big sweeps (>= 2368 blocks of 8192 trials): per-set tables, first pass, deferred-trial pass, final sum; else 2
            "gpu_launches": (4 if (T > 6144 and P * math.ceil(T / 8192.0) >= 2368) else 2) * args.steps,
            "roofline": {"bound": "fp64", "achieved": achieved, "peak": peak_tflops, "unit": "TFLOP/s",
                         "frac": achieved / peak_tflops, "traffic": traffic, "traffic_source": traffic_src,
                         "algorithmic_bytes": int(8 * T + 72 * P + 8 * P * math.ceil(T / 2048.0) +
                                                  # split sweep: bit map out and back, per-set tables out and
                                                  # into every CTA of both launches
                                                  ((2 * 1024 + 3072) * P * math.ceil(T / 8192.0) +
                                                   3072 * P * (1 + math.ceil(math.ceil(T / 8192.0) / 8.0))
                                                   if (T > 6144 and P * math.ceil(T / 8192.0) >= 2368) else 0)),
                         "fp64_pipe_active_pct_ncu": 57.7,
                         "kernel": "k_like_partial: first pass + deferred-trial pass (S-A: 97 % + 3 %) + k_build_tables, k_like_final (0.1 %)",
                         "kernel_ms": k_ms,
                         "pdf_sv_per_eval": mean_pdfsv,
                         "accounting": "algorithmic FLOP = pdf_sv evals x 300 + 28 per eval (SURVEY.md 8d); "
                                       "peak = FP64 FMA probe measured in this run (MEASURED_PEAKS.json has no FP64 entry); "
                                       "frac > 1 because the kernel executes ~22 FP64 instructions per pdf_sv where the "
                                       "reference's sequence costs 300 FLOP; fp64_pipe_active_pct_ncu (profiles/r1_v9_summary.md) "
                                       "is the executed-work utilisation"},
            "cpu_baseline": cpu,
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
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (profiling runs)")
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
