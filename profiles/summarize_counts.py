#!/usr/bin/env python
"""Summaries of the counting kernels' --set full captures (profiles/prof_counts.sh): python profiles/summarize_counts.py TAG > profiles/TAG_counts_ncu_full_summary.txt"""
import csv
import io
import os
import subprocess
import sys

from summarize import WANT

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXTRA = ["lts__t_sectors_op_atom.sum", "lts__t_sectors_op_red.sum", "lts__t_sectors_op_read.sum", "lts__t_sectors_op_write.sum", "lts__t_requests_srcunit_tex.sum",
         "smsp__issue_active.avg.pct_of_peak_sustained_active", "l1tex__t_requests_pipe_lsu_mem_global_op_atom.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_red.sum"]


def one(rep, title):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    out = []
    for r in rows[2:]:
        out.append(f"\n## {title}: {r[hdr.index('Kernel Name')][:110]}")
        for w in WANT + EXTRA:
            if w in hdr:
                c = hdr.index(w)
                out.append(f"{w:62s} {units[c]:14s} {r[c]}")
        st = []
        for c, h in enumerate(hdr):
            if h.startswith("smsp__pcsamp_warps_issue_stalled_") and not h.endswith("_not_issued"):
                try:
                    st.append((float(r[c].replace(",", "")), h[len("smsp__pcsamp_warps_issue_stalled_"):]))
                except ValueError:
                    pass
        tot = sum(v for v, _ in st) or 1.0
        for v, h in sorted(st, reverse=True)[:7]:
            out.append(f"   stall {100 * v / tot:5.1f}% {h}")
    return "\n".join(out)


if __name__ == "__main__":
    tag = sys.argv[1]
    print(f"# {tag}: ncu --set full --clock-control none --import-source on, one launch each (100 M 150-base reads) over\n"
          "# `python bench.py --steps 1 --warmup 3 --e2e-steps 1 --no-cpu --parts=counts --min-seconds 0` (profiles/prof_counts.sh)")
    print(one(os.path.join(ROOT, "gpurun_out", f"{tag}_dep_full.ncu-rep"), "depth"))
    print(one(os.path.join(ROOT, "gpurun_out", f"{tag}_cov_full.ncu-rep"), "coverage"))
