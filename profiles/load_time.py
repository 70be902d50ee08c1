"""Wall time of TreeIndexData.load_tree_index on the bench GFF (host line scan + index build + uploads), cold and warm."""
import os, sys, time
sys.path.insert(0, os.getcwd())
import gffx_b200 as g
from bench import ensure_workload
gff = ensure_workload(60000, g.build_index)
for k in range(3):
    t0 = time.perf_counter()
    idx = g.TreeIndexData.load_tree_index(gff, device=0)
    t1 = time.perf_counter()
    print("load", k, round(t1 - t0, 3), "s;", os.path.getsize(gff) >> 20, "MiB GFF,", idx.n_lines, "lines")
    idx.close()
