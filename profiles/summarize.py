#!/usr/bin/env python
"""Turn the ncu outputs of profiles/run_profile.sh (in gpurun_out/) into the text summaries committed under profiles/.
usage: python profiles/summarize.py TAG   (reads gpurun_out/TAG_launches.csv and gpurun_out/TAG_full.ncu-rep)"""
import collections
import csv
import io
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum",
        "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_sectors.sum",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__inst_executed.sum", "sm__cycles_elapsed.max"]


def launches(tag):
    path = os.path.join(ROOT, "gpurun_out", f"{tag}_launches.csv")
    rows = list(csv.reader(l for l in open(path) if l.startswith('"')))
    hdr = rows[0]
    ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        a = agg.setdefault(re.sub(r"\(.*", "", r[ki]), [0, 0.0, 0.0])
        a[0] += 1
        a[1] += float(r[vi].replace(",", ""))
        a[2] = max(a[2], float(r[vi].replace(",", "")))
    tot = sum(a[1] for a in agg.values())
    out = io.StringIO()
    out.write(f"# {tag}: ncu --metrics gpu__time_duration.sum --clock-control none over `python bench.py --steps 1 --warmup 3 --e2e-steps 1 --no-cpu --parts '' --min-seconds 0`; max_us = the full-size (10 M-region) launch, the e2e arm's chunked launches pull avg_us down\n")
    out.write("# cold-cache, serialised launches: compare SHARES, not absolutes\n")
    out.write(f"{'kernel':72s} {'n':>5s} {'total_us':>11s} {'avg_us':>9s} {'max_us':>9s} {'share':>6s}\n")
    for k, (n, t, mx) in sorted(agg.items(), key=lambda x: -x[1][1]):
        out.write(f"{k[:72]:72s} {n:5d} {t / 1e3:11.1f} {t / 1e3 / n:9.1f} {mx / 1e3:9.1f} {t / tot:6.3f}\n")
    return out.getvalue()


def full(tag):
    rep = os.path.join(ROOT, "gpurun_out", f"{tag}_full.ncu-rep")
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    cols = [hdr.index(w) for w in WANT if w in hdr]
    out = io.StringIO()
    out.write(f"# {tag}: ncu --set full --clock-control none --import-source on (per launch; dram bytes are `traffic`)\n")
    for r in rows[2:]:
        out.write(f"\n## {r[hdr.index('Kernel Name')]}\n")
        for c in cols:
            out.write(f"{hdr[c]:62s} {units[c]:14s} {r[c]}\n")
    return out.getvalue()


if __name__ == "__main__":
    tag = sys.argv[1]
    p = os.path.join(ROOT, "gpurun_out", f"{tag}_launches.csv")
    if os.path.exists(p):
        open(os.path.join(ROOT, "profiles", f"{tag}_launches_summary.txt"), "w").write(launches(tag))
    p = os.path.join(ROOT, "gpurun_out", f"{tag}_full.ncu-rep")
    if os.path.exists(p):
        open(os.path.join(ROOT, "profiles", f"{tag}_ncu_full_summary.txt"), "w").write(full(tag))
