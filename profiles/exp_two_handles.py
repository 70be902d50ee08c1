#!/usr/bin/env python
"""Experiment (r02): how much do two intersect jobs gain from running side by side?  Two handles on the same index (own
streams, own job state), passes of the bench sweep alternating between them, against one handle taking all of them."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
import gffx_b200 as g
from gffx_b200 import _lib

gff = bench.ensure_workload(60000, lambda p, s, n: g.synth_gff(p, seed=s, n_genes=n, quirks=False), g.build_index)
H = [g.TreeIndexData.load_tree_index(gff, device=0) for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 2)]
Q = 10_000_000
sets = [H[0].synth_regions_device(bench.BED_SEED + k, Q, bench.REGION_LEN, first=0, device=torch.device("cuda", 0)) for k in range(4)]
torch.cuda.synchronize()
F = _lib.F_PAIRS_ON_DEVICE | _lib.F_NO_HOST_COPY | _lib.F_ASYNC


def run(handles, steps):
    pend = []
    t0 = time.perf_counter()
    k = 0
    for _ in range(steps):
        new = []
        for (name, mode, inv), regs in zip(bench.PASSES, sets):
            new.append(handles[k % len(handles)].intersect(regs, mode, inv, False, None, flags=F))
            k += 1
        for p in pend:
            p.wait().close()
        pend = new
    for p in pend:
        p.wait().close()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / steps * 1e3


for hs in (H[:1], H):
    run(hs, 5)
    print(len(hs), "handle(s): %.3f ms/step" % run(hs, 200), flush=True)
