#!/usr/bin/env python
"""Per-source-line instruction and stall-sample shares of one kernel launch in an ncu report.
usage: python profiles/srclines.py REPORT.ncu-rep [launch_index] [min_pct]"""
import csv, io, subprocess, sys
rep = sys.argv[1]; li = sys.argv[2] if len(sys.argv) > 2 else "0"; minp = float(sys.argv[3]) if len(sys.argv) > 3 else 0.5
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--launch-skip", li, "--launch-count", "1"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
fname = None; hdr = None; agg = []
for r in rows:
    if r and r[0] == "File Path": fname = r[1].split("/")[-1]; continue
    if r and r[0] == "Function Name": print("#", r[1][:100]); continue
    if r and r[0] == "Line No": hdr = r; ie = hdr.index("Instructions Executed"); ss = hdr.index("# Samples"); continue
    if hdr and r and r[0] not in ("",):
        try: agg.append((fname, int(r[0]), r[1], float(r[ie] or 0), float(r[ss] or 0)))
        except (ValueError, IndexError): pass   # source lines with embedded quotes split badly
ti = sum(a[3] for a in agg); ts = sum(a[4] for a in agg)
print(f"# total warp instructions {ti:.0f}, samples {ts:.0f}")
for f, ln, src, i, s in agg:
    if i >= ti * minp / 100 or s >= ts * minp / 100:
        print(f"{i / ti * 100:5.1f}% inst {s / ts * 100:5.1f}% samp  {f}:{ln:<4d} {src.strip()[:120]}")
