#!/usr/bin/env python
"""Per-source-line instruction / stall-sample totals of an .ncu-rep (needs -lineinfo).  usage: ncu_lines.py rep [topN]"""
import csv, io, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
cur = None; hdr = None; out = []
for r in csv.reader(io.StringIO(txt)):
    if len(r) >= 2 and r[0] == "File Path": cur = r[1].split('/')[-1]; continue
    if len(r) >= 2 and r[0] == "Line No": hdr = r; continue
    if hdr and len(r) == len(hdr) and r[0] not in ("", "Line No"):
        iI = hdr.index("Instructions Executed"); iS = hdr.index("# Samples")
        try: n = int(r[iI] or 0); s = int(r[iS] or 0)
        except ValueError: continue
        if n: out.append((n, s, cur, r[0], r[1].strip()[:100]))
tot = sum(o[0] for o in out); ts = sum(o[1] for o in out)
print("warp-inst", tot, "samples", ts)
for n, s, f, l, src in sorted(out, reverse=True)[:top]:
    print("%5.2f%% inst %5.2f%% smp  %s:%s  %s" % (100 * n / tot, 100 * s / max(ts, 1), f, l, src))
