#!/usr/bin/env python
"""Per-source-line instruction / stall-sample shares from an ncu report.
usage: tools/ncu_lines.py report.ncu-rep [min_pct]"""
import csv, subprocess, sys, io, collections
rep = sys.argv[1]; thr = float(sys.argv[2]) if len(sys.argv) > 2 else 0.5
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
cur = None; hdr = None; agg = collections.OrderedDict()
for r in rows:
    if r and r[0] == "File Path": cur = r[1].split("/")[-1]; continue
    if r and r[0] == "Line No": hdr = r; continue
    if not hdr or len(r) < len(hdr) or not r[0]: continue
    ci = {n: j for j, n in enumerate(hdr)}
    try:
        inst = int(r[ci["Instructions Executed"]]); smp = int(r[ci["# Samples"]])
    except ValueError:
        continue
    long_sb = int(r[ci["stall_long_sb"]] or 0); short_sb = int(r[ci["stall_short_sb"]] or 0)
    bar = int(r[ci["stall_barrier"]] or 0)
    k = (cur, int(r[0]))
    a = agg.setdefault(k, [0, 0, 0, 0, 0, r[1]])
    a[0] += inst; a[1] += smp; a[2] += long_sb; a[3] += short_sb; a[4] += bar
ti = sum(a[0] for a in agg.values()) or 1; ts = sum(a[1] for a in agg.values()) or 1
print("total warp-instructions %d, samples %d" % (ti, ts))
print("%-18s %5s %7s %7s %6s %6s %6s  source" % ("file", "line", "inst%", "samp%", "longsb", "shortsb", "bar"))
for (f, l), a in sorted(agg.items()):
    if 100 * a[0] / ti >= thr or 100 * a[1] / ts >= thr:
        print("%-18s %5d %6.2f%% %6.2f%% %6d %6d %6d  %s" % (f[:18], l, 100 * a[0] / ti, 100 * a[1] / ts, a[2], a[3], a[4], a[5].strip()[:100]))
