"""Wall time of TreeIndexData.load_tree_index on the bench GFF: the line table from the host scanner, the device scanner
(linescan.cu) and the side-car cache (linecache.cpp); GFFX_TIMING=1 prints the phases."""
import os, sys, time
sys.path.insert(0, os.getcwd())
import gffx_b200 as g
from bench import ensure_workload
gff = ensure_workload(60000, lambda p, s, n: g.synth_gff(p, seed=s, n_genes=n), g.build_index)
cache = gff + ".b200cache"
def run(tag, **env):
    for k, v in env.items():
        os.environ[k] = v
    for k in range(3):
        t0 = time.perf_counter()
        idx = g.TreeIndexData.load_tree_index(gff, device=0)
        t1 = time.perf_counter()
        print(tag, k, round(t1 - t0, 3), "s; source", idx.lines_info()[1], ";", os.path.getsize(gff) >> 20, "MiB GFF,", idx.n_lines, "lines", flush=True)
        idx.close()
run("host scan, no cache  ", GFFX_CACHE="0", GFFX_LINE_SCAN="host")
run("device scan, no cache", GFFX_CACHE="0", GFFX_LINE_SCAN="device")
if os.path.exists(cache):
    os.remove(cache)
run("device scan + cache  ", GFFX_CACHE="1", GFFX_LINE_SCAN="device")
