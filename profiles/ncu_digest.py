#!/usr/bin/env python
"""Digest of an .ncu-rep: headline metrics + SASS opcode mix.  usage: ncu_digest.py file.ncu-rep"""
import collections, csv, io, re, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = [r for r in csv.reader(io.StringIO(raw)) if len(r) > 20]
hdr, units, vals = rows[0], rows[1], rows[2]
want = ["gpu__time_duration.sum", "launch__registers_per_thread", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__issue_active.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "smsp__warps_eligible.avg.per_cycle_active", "sass__inst_executed_local_loads", "sass__inst_executed_local_stores",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]
for h, u, v in zip(hdr, units, vals):
    if h in want or re.match(r"smsp__average_warps_issue_stalled_.*_per_issue_active.ratio", h) and float(v or 0) > 0.15:
        print("%-90s %-12s %s" % (h, u, v))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hdr = rows[1]; data = rows[2:]
iS, iI, iT = hdr.index("# Samples"), hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed")
ops, samp, thr = collections.Counter(), collections.Counter(), collections.Counter()
tot = ts = 0
for r in data:
    if len(r) <= iT: continue
    op = r[1].split()
    if not op: continue
    o = (op[0] if not op[0].startswith('@') else op[1]).split('.')[0]
    n, s = int(r[iI] or 0), int(r[iS] or 0)
    ops[o] += n; samp[o] += s; thr[o] += int(r[iT] or 0); tot += n; ts += s
print("SASS instructions", len(data), "executed warp-inst", tot, "samples", ts)
for o, n in ops.most_common(22):
    print("%-10s inst %6.2f%%  samples %6.2f%%  thr/inst %5.1f" % (o, 100 * n / tot, 100 * samp[o] / ts, thr[o] / max(n, 1)))
