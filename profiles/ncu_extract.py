#!/usr/bin/env python
"""Executed-work figures of the likelihood kernel from .ncu-rep captures -> profiles/ncu_fp64.json.

    python profiles/ncu_extract.py --key A:100000x10000 --pdf-sv-evals N --sha <csrc sha> --source <text> \
        --out profiles/ncu_fp64.json first.ncu-rep [second.ncu-rep ...]

Each report holds ONE launch of k_like_partial of one step of the workload named by --key (a split sweep
has two: first pass + deferred-trial pass); their counts are added.  Per launch, from the source page
(needs -lineinfo + --import-source on): executed warp-level and thread-level instructions per SASS opcode;
from the raw page: duration, FP64 pipe utilisation, DRAM bytes.  bench.py divides the FP64-pipe
thread-instruction total by the pdf_sv evaluations of the step (counted by the kernel, identical run to
run) and refuses the entry when --sha differs from the hash of the current hddm_b200/csrc.
"""
import argparse
import collections
import csv
import io
import json
import os
import subprocess
import sys

FP64_OPS = ("DFMA", "DMUL", "DADD", "DSETP", "DMNMX")   # what the FP64 pipe executes in these kernels


def pages(rep):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = [r for r in csv.reader(io.StringIO(raw)) if len(r) > 20]
    scale = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12,      # -> bytes
             "ns": 1e-6, "us": 1e-3, "usecond": 1e-3, "ms": 1.0, "msecond": 1.0, "s": 1e3, "second": 1e3}  # -> ms
    metrics = {}
    for name, unit, val in zip(rows[0], rows[1], rows[2]):
        try:
            metrics[name] = repr(float(val.replace(",", "")) * scale.get(unit, 1.0)) if unit in scale else val
        except ValueError:
            metrics[name] = val
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    hdr, data = rows[1], rows[2:]
    iI, iT = hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed")
    warp, thread = collections.Counter(), collections.Counter()
    for r in data:
        if len(r) <= iT:
            continue
        op = r[1].split()
        if not op:
            continue
        o = (op[0] if not op[0].startswith('@') else op[1]).split('.')[0]
        warp[o] += int(r[iI] or 0)
        thread[o] += int(r[iT] or 0)
    return metrics, warp, thread


def num(metrics, name):
    try:
        return float(metrics[name].replace(",", ""))
    except Exception:
        return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--key", required=True)
    ap.add_argument("--pdf-sv-evals", type=float, required=True)
    ap.add_argument("--sha", required=True)
    ap.add_argument("--source", default="")
    ap.add_argument("--out", default=os.path.join(os.path.dirname(os.path.abspath(__file__)), "ncu_fp64.json"))
    ap.add_argument("reps", nargs="+")
    a = ap.parse_args()
    tot_w, tot_t = collections.Counter(), collections.Counter()
    launches = []
    for rep in a.reps:
        m, w, t = pages(rep)
        tot_w.update(w)
        tot_t.update(t)
        launches.append({
            "report": os.path.basename(rep), "kernel": m.get("Kernel Name"),
            "duration_ms": num(m, "gpu__time_duration.sum"),
            "fp64_pipe_pct": num(m, "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active"),
            "issue_active_pct": num(m, "sm__issue_active.avg.pct_of_peak_sustained_elapsed"),
            "warps_eligible": num(m, "smsp__warps_eligible.avg.per_cycle_active"),
            "registers": num(m, "launch__registers_per_thread"),
            "dram_read": num(m, "dram__bytes_read.sum"), "dram_write": num(m, "dram__bytes_write.sum"),
            "warp_inst": sum(w.values()), "fp64_warp_inst": sum(w[o] for o in FP64_OPS),
            "fp64_thread_inst": sum(t[o] for o in FP64_OPS)})
    fp64_t = sum(tot_t[o] for o in FP64_OPS)
    fused, other = tot_t["DFMA"], fp64_t - tot_t["DFMA"]
    entry = {
        "csrc_sha": a.sha, "source": a.source, "pdf_sv_evals": a.pdf_sv_evals,
        "fp64_thread_inst": fp64_t, "fp64_warp_inst": sum(tot_w[o] for o in FP64_OPS),
        "fp64_flop": 2 * fused + other, "fused_share": fused / float(fp64_t) if fp64_t else None,
        "thread_inst_by_opcode": {o: tot_t[o] for o in FP64_OPS},
        "fp64_share_of_warp_inst": sum(tot_w[o] for o in FP64_OPS) / float(sum(tot_w.values()) or 1),
        "warp_inst_top": dict(tot_w.most_common(14)),
        "fp64_pipe_pct": launches[0]["fp64_pipe_pct"],
        "dram_bytes": sum((l["dram_read"] or 0) + (l["dram_write"] or 0) for l in launches),
        "launches": launches}
    try:
        out = json.load(open(a.out))
    except Exception:
        out = {}
    out["_comment"] = ("Executed FP64 work of k_like_partial per bench step, from ncu --set full captures "
                       "(profiles/ncu_extract.py).  bench.py uses an entry only when csrc_sha matches the kernel sources.")
    out[a.key] = entry
    json.dump(out, open(a.out, "w"), indent=1, sort_keys=True)
    print(json.dumps({k: entry[k] for k in ("fp64_thread_inst", "fp64_flop", "fused_share", "fp64_share_of_warp_inst",
                                            "fp64_pipe_pct", "dram_bytes")}))


if __name__ == "__main__":
    sys.exit(main())
