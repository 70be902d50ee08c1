"""Per-call cost of gffx_intersect at a tiny and at the bench size: wall time, device span, the gap the host adds, per-stage device ms."""
import time, sys, os
sys.path.insert(0, os.getcwd())
import numpy as np, torch
import gffx_b200 as g
from gffx_b200 import _lib
from bench import ensure_workload, workload_dir
gff = ensure_workload(60000, g.build_index)
idx = g.TreeIndexData.load_tree_index(gff, device=0)
dev = torch.device("cuda", 0)
F = _lib.F_PAIRS_ON_DEVICE | _lib.F_NO_HOST_COPY | (_lib.F_STAGE_TIMES if os.environ.get("STAGES") else 0)
for n in (1000, 10_000_000):
    regs = idx.synth_regions_device(5, n, 20000, device=dev)
    for _ in range(5):
        idx.intersect(regs, 2, False, False, None, flags=F, copy=False).close()
    torch.cuda.synchronize()
    t0 = time.perf_counter(); K = 50
    kms = 0
    for _ in range(K):
        r = idx.intersect(regs, 2, False, False, None, flags=F, copy=False); kms += r.kernel_ms; r.close()
    torch.cuda.synchronize()
    w = (time.perf_counter() - t0) / K * 1e3
    print(n, "stage_ms", [round(x, 3) for x in r.stage_ms])
    print(n, "wall ms/call", round(w, 4), "kernel span ms", round(kms / K, 4), "host gap", round(w - kms / K, 4))
