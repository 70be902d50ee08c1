"""bench.py's two arms: the CPU arm (--impl reference) runs the oracle on the host cores WITHOUT mapping the engine's library and
with an explicit thread count; the workload it makes for itself is the one the GPU arm joins (same GFF text, same region stream)."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def test_reference_arm_runs_without_the_engine_library(tmp_path):
    code = r"""
import json, os, sys
sys.argv = ["bench.py", "--impl", "reference", "--genes", "400", "--regions", "20000", "--steps", "2", "--warmup", "1", "--ref-seconds", "2"]
os.environ["OMP_NUM_THREADS"] = "1"   # what torchrun exports; the arm must not inherit it
import bench
rc = bench.main()
maps = open("/proc/self/maps").read()
print(json.dumps({"rc": rc, "engine": "libgffx_b200" in maps, "cuda": "libcudart" in maps or "libcuda.so" in maps, "oracle": "liboracle" in maps,
                  "synth": "libgffx_synth" in maps, "torch": "torch" in sys.modules}))
"""
    env = dict(os.environ, GFFX_BENCH_DIR=str(tmp_path / "w"))
    p = subprocess.run([sys.executable, "-c", code], cwd=ROOT, env=env, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [json.loads(x) for x in p.stdout.splitlines() if x.startswith("{")]
    line, probe = lines[0], lines[-1]
    assert probe == {"rc": 0, "engine": False, "cuda": False, "oracle": True, "synth": True, "torch": False}
    assert line["impl"] == "reference" and line["steps"] == 2 and line["warmup"] == 1
    assert line["cpu_baseline"]["cores"] == (os.cpu_count() or 1) and line["cpu_baseline"]["kind"] == "port"
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["value"] > 0
    for k in ("metric", "unit", "n_gpus", "ms_per_step", "higher_is_better", "scaling", "vs_baseline", "dtype", "data", "config"):
        assert k in line


def test_reference_arm_other_ranks_exit_without_work():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    p = subprocess.run([sys.executable, "bench.py", "--impl", "reference", "--gpus", "2"], cwd=ROOT, env=env, capture_output=True, text=True, timeout=120)
    assert p.returncode == 0 and p.stdout.strip() == ""


def test_config_is_the_same_in_both_arms():
    import argparse

    import bench

    a = argparse.Namespace(genes=60000)
    assert bench.workload_config(a, 60000, 10_000_000) == bench.workload_config(a, 60000, 10_000_000)
    src = open(os.path.join(ROOT, "bench.py")).read()
    assert src.count('"config": workload_config(args, ') == 2   # one call per arm, nothing arm-specific merged in


def test_synth_text_of_the_standalone_generator_is_the_engines(tmp_path):
    """libgffx_synth.so and libgffx_b200.so are built from the same synth.cpp: byte-identical GFF text"""
    import bench
    import gffx_b200 as g

    a, b = str(tmp_path / "a.gff"), str(tmp_path / "b.gff")
    assert bench.synth_lib().gffx_synth_gff_file(a.encode(), 77, 300, 0, 0) == 0
    g.synth_gff(b, seed=77, n_genes=300, quirks=False)
    assert open(a, "rb").read() == open(b, "rb").read()


@pytest.mark.gpu
def test_region_stream_of_the_cpu_arm_is_the_engines(tmp_path):
    import bench
    import gffx_b200 as g
    from tests import _oracle

    gff = str(tmp_path / "s.gff")
    g.synth_gff(gff, seed=5, n_genes=2000, quirks=False)
    g.build_index(gff)
    with g.TreeIndexData.load_tree_index(gff) as idx, _oracle.Oracle(gff) as orc:
        ext = bench.oracle_extents(orc)
        for seed, first, n, length in ((1, 0, 50000, 20000), (2, 123456789012, 30000, 150), (3, 7, 1000, 400_000_000)):
            want = idx.synth_regions(seed, n, length, first=first)
            got = bench.synth_regions_cpu(ext, seed, first, n, length)
            dev = idx.synth_regions_device(seed, n, length, first=first)
            for w, x, d in zip(want, got, dev):
                assert np.array_equal(w, x) and np.array_equal(w, d.cpu().numpy().view(np.uint32))
