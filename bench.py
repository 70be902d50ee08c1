#!/usr/bin/env python
"""bench.py — the hot path of GFFx on B200: `intersect --bed` (+ depth / coverage) over synthetic human-scale GFF3.

Headline workload (BASELINE.json configs[1]): 60k-gene human-like GFF3 (~2.7 M lines, seed 20260001) x
10 M synthetic 20 kb BED regions per GPU.  One STEP = the configuration's mode sweep, four full intersect passes (root-level
join with the (root_fid,start,end) pairs materialised in HBM, root -> block selection, line-level filter, ordered compaction
into output byte ranges):
    overlap | --contained | --contains-region | --contained --invert
each pass over its OWN 10 M-region set (4 x 120 MB of inputs rotate through the 126 MB L2, so no pass finds its inputs
cached).  value = query intervals/s = n_gpus * 4 * Q * steps / time.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

The same JSON line carries the other BASELINE configs as sub-objects, each with its own roofline / e2e / cpu_baseline:
    entire_group_types   configs[2]: -e -T gene,mRNA,exon on 10 M regions
    depth, coverage      configs[3]: 100 M 150-base read intervals (N > 1: the sharded job over N x 100 M)
    sharded_1b           configs[4], N > 1: 1 B regions split over the ranks, intersect (overlap) + depth
and, N > 1, "parity_n": the sharded job's byte ranges on a 200 k-regions-per-rank job equal the single-GPU answer over the
union and the oracle's.

N > 1 is launched by torchrun, one rank per GPU; every rank joins its own regions against a replicated index (weak scaling;
shards are independent, the only exchange is the element-wise combination of the job's byte flags — roots hit | lines matched
— by one NCCL all-reduce(MAX) per pass; gffx_b200/shard.py).

--impl reference times the CPU implementation of the path (the C++ restatement under oracle/: the Rust reference cannot be
built in this image) on the box's host cores, all of them, on a bounded sample of the workload.  That arm loads oracle/ and
the workload generator (libgffx_synth.so) only — never the engine's library.
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

GFF_SEED = 20260001
BED_SEED = 20260002
READ_SEED = 20260010
REGION_LEN = 20000
READ_LEN = 150
# (name, mode, invert) — OverlapMode: 0 contained, 1 contains-region, 2 overlap
PASSES = [("overlap", 2, False), ("contained", 0, False), ("contains_region", 1, False), ("contained_invert", 0, True)]
EG_TYPES = "gene,mRNA,exon"
METRIC = "query_intervals_per_sec"
UNIT = "intervals/s"


def env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


def workload_dir(genes):
    return os.environ.get("GFFX_BENCH_DIR", f"/tmp/gffx_b200_bench_{GFF_SEED}_{genes}")


def ensure_workload(genes, synth, builder):
    """Synthetic GFF + its index files, written once per box.  synth(path, seed, genes) writes the text, builder(path) the index."""
    d = workload_dir(genes)
    gff = os.path.join(d, "hs_synth.gff")
    done = os.path.join(d, ".done")
    if not os.path.exists(done):
        os.makedirs(d, exist_ok=True)
        synth(gff, GFF_SEED, genes)
        builder(gff)
        open(done, "w").write("ok\n")
    return gff


def load_peaks():
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except (OSError, ValueError):
        peaks = {}
    peak = float(peaks.get("hbm_gbs", 6650.0))
    return peak, ("measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback (B200_PROFILING.md 6.65 TB/s)")


class ClockSampler:
    """nvidia-smi clocks + throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.rows = []
        self.proc = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(device), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), [x.strip() for x in line.split(",")]))

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        rows = [r for t, r in self.rows if t0 <= t <= t1 and len(r) >= 6] or [r for _, r in self.rows if len(r) >= 6]
        mhz = sorted(int(r[0]) for r in rows if r[0].isdigit())
        mx = max((int(r[1]) for r in rows if r[1].isdigit()), default=None)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(r[2 + k].lower().startswith("active") for r in rows)]
        return {"sm_mhz": mhz[len(mhz) // 2] if mhz else None, "sm_max_mhz": mx, "reasons": reasons, "samples": len(rows)}


# ---------------------------------------------------------------------------------------------------
# CPU arm: the C++ restatement of the reference (oracle/), timed on the host cores.  This section and
# tests/ are the only places that may execute oracle/ code; nothing below it is on the product path.
# ---------------------------------------------------------------------------------------------------
def synth_lib():
    """the workload generator alone (gffx_b200/lib/libgffx_synth.so = synth.cpp + util.cpp): no engine, no CUDA"""
    lib = ctypes.CDLL(os.path.join(ROOT, "gffx_b200", "lib", "libgffx_synth.so"))
    lib.gffx_synth_gff_file.restype = ctypes.c_int
    lib.gffx_synth_gff_file.argtypes = [ctypes.c_char_p, ctypes.c_uint64, ctypes.c_uint32, ctypes.c_int, ctypes.c_uint32]
    lib.gffx_synth_regions_ext.restype = ctypes.c_int
    lib.gffx_synth_regions_ext.argtypes = [ctypes.c_void_p, ctypes.c_uint32] + [ctypes.c_uint64] * 3 + [ctypes.c_uint32] + [ctypes.c_void_p] * 3
    return lib


_SYNTH = None


def synth_regions_cpu(extents, seed, first, n, length):
    """regions [first, first+n) of stream `seed` without the engine: libgffx_synth.so, the same counter-based function
    (gffx_b200/csrc/synth_common.h) that gffx_synth_regions_host / _device evaluate; extents[c] = largest root end of sequence c"""
    import numpy as np

    global _SYNTH
    _SYNTH = _SYNTH or synth_lib()
    ext = np.ascontiguousarray(extents, dtype=np.uint32)
    out = [np.empty(n, dtype=np.uint32) for _ in range(3)]
    if _SYNTH.gffx_synth_regions_ext(ext.ctypes.data, len(ext), seed, first, n, length, *(a.ctypes.data for a in out)) != 0:
        raise SystemExit("synthetic regions: the index has no root intervals")
    return tuple(out)


def oracle_extents(orc):
    import numpy as np

    ext = []
    for c in range(len(orc.seq_names)):
        t = orc.seq_intervals(c)
        ext.append(int(t[:, 1].max()) if len(t) else 0)
    return np.array(ext, dtype=np.uint32)


def cpu_intersect_step(orc, region_sets, passes):
    """one step of the mode sweep through the oracle's intersect (intersect.rs:586-654 restated)"""
    n = 0
    for (name, mode, inv, entire, types), regs in zip(passes, region_sets):
        orc.intersect(regs, mode, inv, entire, types)
        n += regs[0].shape[0]
    return n


SWEEP = [(n, m, i, False, None) for n, m, i in PASSES]


def calibrate(fn, probe_n, full_n, budget_s):
    """largest sample (multiple of probe_n, at most full_n) whose run is expected to fit budget_s, from one probe run"""
    t = time.perf_counter()
    fn(probe_n)
    dt = max(time.perf_counter() - t, 1e-3)
    q = int(probe_n * max(1.0, budget_s / dt))
    return min(full_n, max(probe_n, (q // probe_n) * probe_n)), dt


def cpu_baseline_intersect(orc, ext, q_full, budget_s, passes, cores):
    def run(q):
        sets = [synth_regions_cpu(ext, BED_SEED + k, 0, q, REGION_LEN) for k in range(len(passes))]
        t = time.perf_counter()
        n = cpu_intersect_step(orc, sets, passes)
        return n, time.perf_counter() - t

    q_s, _ = calibrate(lambda q: run(q), 50000, q_full, budget_s)
    n, dt = run(q_s)
    return {"value": n / dt, "unit": UNIT, "cores": cores, "kind": "port", "seconds": dt,
            "sample": f"one step of the same passes on {q_s} of {q_full} regions per pass ({cores} OpenMP threads)",
            "note": "C++ restatement of GFFx v0.4.0 (oracle/); the Rust reference cannot be compiled in this image"}


def cpu_baseline_counts(orc, ext, kind, q_full, budget_s, cores):
    fn = orc.depth if kind == 0 else orc.coverage

    def run(q):
        regs = synth_regions_cpu(ext, READ_SEED, 0, q, READ_LEN)
        t = time.perf_counter()
        fn(regs)
        return time.perf_counter() - t

    q_s, _ = calibrate(lambda q: run(q), 200000, q_full, budget_s)
    dt = run(q_s)
    return {"value": q_s / dt, "unit": UNIT, "cores": cores, "kind": "port", "seconds": dt,
            "sample": f"one {'depth' if kind == 0 else 'coverage'} job on {q_s} of {q_full} read intervals ({cores} OpenMP threads)",
            "note": "C++ restatement of GFFx v0.4.0 (oracle/)"}


def open_oracle(gff):
    from tests import _oracle

    cores = os.cpu_count() or 1
    _oracle.set_threads(cores)      # explicit: torchrun exports OMP_NUM_THREADS=1
    return _oracle.Oracle(gff), cores


def run_reference(args, rank, world):
    if rank != 0:
        return 0
    from tests import _oracle

    sl = synth_lib()

    def synth(path, seed, genes):
        if sl.gffx_synth_gff_file(path.encode(), seed, genes, 0, 0) != 0:
            raise SystemExit("synthetic GFF generation failed")

    gff = ensure_workload(args.genes, synth, _oracle.build_index)
    orc, cores = open_oracle(gff)
    with orc:
        ext = oracle_extents(orc)
        budget = args.ref_seconds / max(1, args.steps + args.warmup)

        def probe(q):
            cpu_intersect_step(orc, [synth_regions_cpu(ext, BED_SEED + k, 0, q, REGION_LEN) for k in range(len(PASSES))], SWEEP)

        q_s, _ = calibrate(probe, 50000, args.regions, budget)
        sets = [synth_regions_cpu(ext, BED_SEED + k, 0, q_s, REGION_LEN) for k in range(len(PASSES))]
        for _ in range(args.warmup):
            cpu_intersect_step(orc, sets, SWEEP)
        t0 = time.perf_counter()
        n = 0
        for _ in range(args.steps):
            n += cpu_intersect_step(orc, sets, SWEEP)
        dt = time.perf_counter() - t0
        n_roots = orc.n_roots
    val = n / dt
    sample = f"{q_s} of {args.regions} regions per pass x {len(PASSES)} passes per step (host RAM resident, all {cores} threads)"
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": workload_config(args, n_roots, args.regions),
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample,
                         "note": "C++ restatement of GFFx v0.4.0 (oracle/); the Rust reference cannot be compiled in this image"},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "libraries": "oracle/_build/liboracle.so + gffx_b200/lib/libgffx_synth.so (workload text only); the engine's library is not loaded",
    }
    print(json.dumps(line), flush=True)
    return 0


def workload_config(args, n_roots, q):
    """identical in both arms (the driver compares them): nothing here depends on which implementation ran"""
    return {
        "workload": f"intersect --bed: synthetic human-like GFF3 ({args.genes} genes, seed {GFF_SEED}) x {q} "
                    f"20 kb BED regions per pass per GPU; step = overlap | --contained | --contains-region | --contained --invert",
        "regions_per_pass_per_gpu": q, "passes_per_step": len(PASSES), "region_len": REGION_LEN, "n_roots": n_roots,
        "pairs": "materialised in HBM",
        "timing": "value / ms_per_step: the K steps are repeated until the timed region (two CUDA events on the engine's stream) lasts "
                  ">= 1 s (timing.loops x K steps, ms_per_step = their mean); the passes are queued asynchronously (GFFX_F_ASYNC), results "
                  "waited one step behind; roofline / per_pass: K more steps, synchronous, with GFFX_F_STAGE_TIMES (per-stage events)",
        "l2": "inputs larger than L2: 4 x 12 B x Q region sets rotate (no flush needed)",
        "sharding": "regions split across ranks (counter-based stream, rank r owns [r*Q,(r+1)*Q)); index replicated; N>1: the job's "
                    "byte flags (roots hit | lines matched, one array) are combined by one NCCL all-reduce(MAX) per pass "
                    "and every rank emits the byte ranges of the whole N*Q-region job",
    }


# ---------------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------------
JOIN_NAME = "root join + rank-space scatter (wj_root_kernel, offsets scan, wj_expand_kernel)"


def stage_table(q, p, l_sel, e_out):
    """what each timed stage of one intersect call is, and its ALGORITHMIC bytes (every input read once, every
    output written once; re-reads and scan passes the implementation adds are NOT credited)"""
    return {
        8: ("window partition (hist + scan + scatter)", 0),
        # regions (3 x u32) read once — by the join and, in the same kernel, by the line filter's scatter (a7's 12*Q is NOT
        # credited a second time) — and pairs (3 x u32) written once; stage 7, the expansion, is added to stage 0
        0: (JOIN_NAME, 12 * q + 12 * p),
        3: ("select_blocks + merge_blocks", 0),
        5: ("rank-space scans", 0),
        6: ("eval_lines + emit_ranges", 12 * l_sel + 16 * e_out),
    }


def bind_to_gpu_cpus(local_rank):
    """Host threads and first-touch pinned allocations next to the GPU: the CPU set NVML reports for the device (its NUMA node /
    PCIe root).  Without it every rank's page-locked buffers land wherever the launcher started the process."""
    info = {"bound": False}
    try:
        import pynvml

        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local_rank)
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = [64 * w + b for w, m in enumerate(mask) for b in range(64) if (m >> b) & 1]
        allowed = sorted(set(cpus) & os.sched_getaffinity(0))
        if allowed:
            os.sched_setaffinity(0, allowed)
            info = {"bound": True, "cpus": f"{allowed[0]}-{allowed[-1]} ({len(allowed)})"}
            try:
                info["numa_node"] = int(pynvml.nvmlDeviceGetNumaNodeId(h))
            except Exception:
                pass
    except Exception as e:   # no NVML, or a container without the right to change its affinity
        info["error"] = str(e)[:100]
    return info


class Ctx:
    pass


def make_ctx(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist

    import gffx_b200 as g

    c = Ctx()
    c.args, c.rank, c.world, c.local_rank = args, rank, world, local_rank
    c.affinity = bind_to_gpu_cpus(local_rank) if not args.no_bind else {"bound": False}
    torch.cuda.set_device(local_rank)
    c.dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=c.dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    c.barrier = barrier
    if rank == 0:
        ensure_workload(args.genes, lambda p, s, n: g.synth_gff(p, seed=s, n_genes=n, quirks=False), g.build_index)
    barrier()
    c.gff = os.path.join(workload_dir(args.genes), "hs_synth.gff")
    c.idx = g.TreeIndexData.load_tree_index(c.gff, device=local_rank)
    c.stream = torch.cuda.ExternalStream(c.idx.cuda_stream, device=c.dev)
    c.combiner = None
    if world > 1:
        from gffx_b200.shard import Combiner

        c.combiner = Combiner(c.idx)
    c.peak, c.peak_src = load_peaks()
    return c


def allmax(c, x):
    import torch
    import torch.distributed as dist

    t = torch.tensor([x], device=c.dev, dtype=torch.float64)
    if c.world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def allsum(c, xs):
    import torch
    import torch.distributed as dist

    t = torch.tensor(list(xs), device=c.dev, dtype=torch.float64)
    if c.world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return [float(x) for x in t.tolist()]


def queue_pass(c, regs, mode, inv, entire, types, flags):
    if c.world == 1:
        return c.idx.intersect(regs, mode, inv, entire, types, flags=flags, copy=False)
    return c.idx.shard_intersect(regs, mode, inv, entire, types, combine=c.combiner, flags=flags, copy=False)


def timed_async(c, passes, sets, steps, min_seconds):
    """`steps` steps of `passes`, repeated until the event-timed region lasts min_seconds; jobs are queued asynchronously and the
    results of one step are waited for after the next step has been queued, so the stream never drains.
    -> (ms_total over loops*steps, loops, tally per pass name)"""
    import torch

    from gffx_b200 import _lib

    F = _lib.F_PAIRS_ON_DEVICE | _lib.F_NO_HOST_COPY | _lib.F_ASYNC
    tally = {}

    def settle(pend):
        for name, p in pend:
            r = p.wait()
            a = tally.setdefault(name, {"calls": 0, "pairs": 0, "ranges": 0, "roots": 0, "launches": 0, "kernel_ms": 0.0})
            a["calls"] += 1
            a["pairs"] += r.n_pairs
            a["ranges"] += r.n_ranges
            a["roots"] += r.n_roots
            a["launches"] += r.launches
            a["kernel_ms"] += r.kernel_ms
            r.close()

    def one_step(prev):
        pend = [(name, queue_pass(c, regs, mode, inv, entire, types, F)) for (name, mode, inv, entire, types), regs in zip(passes, sets)]
        if prev:
            settle(prev)
        return pend

    # one untimed step to estimate the step time, then loops x steps between two events
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    settle(one_step(None))   # first use of the second lane allocates its buffers
    ev0.record(c.stream)
    settle(one_step(None))
    ev1.record(c.stream)
    torch.cuda.synchronize()
    est = max(allmax(c, ev0.elapsed_time(ev1)), 1e-3)
    loops = max(1, int(1.15 * min_seconds * 1e3 / (est * steps) + 0.999))   # the first step runs cold: margin
    tally.clear()
    c.barrier()
    t0 = time.perf_counter()
    ev0.record(c.stream)
    prev = None
    for _ in range(loops * steps):
        prev = one_step(prev)
    settle(prev)
    ev1.record(c.stream)
    c.barrier()
    t1 = time.perf_counter()
    return allmax(c, ev0.elapsed_time(ev1)), loops, tally, (t0, t1)


def staged_sync(c, passes, sets, steps):
    """the same steps, synchronous, with per-stage CUDA events: what `roofline` and `per_pass` are computed from"""
    from gffx_b200 import _lib

    F = _lib.F_PAIRS_ON_DEVICE | _lib.F_NO_HOST_COPY | _lib.F_STAGE_TIMES
    acc = {}
    for _ in range(steps):
        for (name, mode, inv, entire, types), regs in zip(passes, sets):
            r = queue_pass(c, regs, mode, inv, entire, types, F)
            a = acc.setdefault(name, {"calls": 0, "pairs": 0, "ranges": 0, "roots": 0, "launches": 0, "kernel_ms": 0.0, "stage_ms": [0.0] * 12})
            a["calls"] += 1
            a["pairs"] += r.n_pairs
            a["ranges"] += r.n_ranges
            a["roots"] += r.n_roots
            a["launches"] += r.launches
            a["kernel_ms"] += r.kernel_ms
            for s in range(12):
                a["stage_ms"][s] += r.stage_ms[s]
            r.close()
    return acc


def e2e_sync(c, passes, host_sets, steps):
    """the public call with pinned HOST arrays in and byte ranges out (copies inside the timed region) -> (seconds, ranges)"""
    import torch

    from gffx_b200 import _lib

    def step():
        n = 0
        for (name, mode, inv, entire, types), regs in zip(passes, host_sets):
            r = queue_pass(c, regs, mode, inv, entire, types, _lib.F_PAIRS_ON_DEVICE)
            n += r.n_ranges
            r.close()
        return n

    step()
    c.barrier()
    t0 = time.perf_counter()
    n_ranges = 0
    for _ in range(steps):
        n_ranges += step()
    torch.cuda.synchronize()
    return allmax(c, time.perf_counter() - t0), n_ranges


def e2e_queued(c, passes, host_sets, steps):
    """the same host-in / host-out passes QUEUED (GFFX_F_ASYNC without GFFX_F_DEVICE_INPUT / GFFX_F_NO_HOST_COPY): a pass is
    waited for after the next one has been queued, so one pass's ranges travel to the host while the next pass's regions travel
    to the device (the two directions of the link).  Every result is waited for and its host ranges looked at inside the
    timed region.  -> (seconds, ranges)"""
    import torch

    from gffx_b200 import _lib

    F = _lib.F_PAIRS_ON_DEVICE | _lib.F_ASYNC

    def settle(p):
        r = p.wait(copy=False)
        n = r.n_ranges
        if n:
            assert int(r.ranges[n - 1, 1]) > 0   # the host copy is there
        r.close()
        return n

    def run(k):
        n, prev = 0, None
        for _ in range(k):
            for (name, mode, inv, entire, types), regs in zip(passes, host_sets):
                p = queue_pass(c, regs, mode, inv, entire, types, F)
                if prev is not None:
                    n += settle(prev)
                prev = p
        return n + settle(prev)

    run(1)
    c.barrier()
    t0 = time.perf_counter()
    n_ranges = run(steps)
    torch.cuda.synchronize()
    return allmax(c, time.perf_counter() - t0), n_ranges


def h2d_ceiling(c, host_sets):
    """what the box can move: every rank copies its pinned region sets to its GPU at the same time, nothing else running"""
    import torch

    dst = [tuple(torch.empty_like(t, device=c.dev) for t in hs) for hs in host_sets]
    nbytes = sum(t.numel() * 4 for hs in host_sets for t in hs)
    best = None
    for _ in range(3):
        c.barrier()
        t0 = time.perf_counter()
        for hs, ds in zip(host_sets, dst):
            for h, d in zip(hs, ds):
                d.copy_(h, non_blocking=True)
        torch.cuda.synchronize()
        dt = allmax(c, time.perf_counter() - t0)
        best = dt if best is None else min(best, dt)
    del dst
    return {"gbs_per_rank": nbytes / best / 1e9, "gbs_aggregate": c.world * nbytes / best / 1e9, "bytes_per_rank": nbytes}


def roofline_of(c, acc, q, steps):
    """dominant kernel group over the step and the whole pass, from the stage-timed run"""
    per_kernel = {}
    for name, a in acc.items():
        p, e_out = a["pairs"] / a["calls"], a["ranges"] / a["calls"]
        for s, (kname, nbytes) in stage_table(q, p, c.idx.n_lines, e_out).items():
            k = per_kernel.setdefault(kname, {"ms": 0.0, "bytes": 0.0, "launches": 0})
            if a["stage_ms"][s] > 0:
                k["ms"] += a["stage_ms"][s] + (a["stage_ms"][7] if s == 0 else 0.0)
                k["bytes"] += nbytes * a["calls"]
                k["launches"] += a["calls"]
    stage_sum = sum(k["ms"] for k in per_kernel.values()) or 1.0
    dom_name, dom = max(((n, k) for n, k in per_kernel.items() if k["bytes"] > 0), key=lambda x: x[1]["ms"])
    achieved = dom["bytes"] / (dom["ms"] * 1e-3) / 1e9 if dom["ms"] > 0 else 0.0
    roofline = {"bound": "hbm", "kernel": dom_name, "achieved": achieved, "peak": c.peak, "unit": "GB/s", "frac": achieved / c.peak,
                "traffic": None, "peak_source": c.peak_src, "avg_launch_ms": dom["ms"] / max(1, dom["launches"]),
                "algorithmic_bytes_per_launch": dom["bytes"] / max(1, dom["launches"]), "share_of_step": dom["ms"] / stage_sum,
                "kernels": {n: {"ms_per_step": k["ms"] / steps, "share": k["ms"] / stage_sum,
                                "GBps": (k["bytes"] / (k["ms"] * 1e-3) / 1e9 if k["ms"] > 0 else 0.0)} for n, k in per_kernel.items()}}
    pass_bytes = sum(k["bytes"] for k in per_kernel.values())
    pass_ms = sum(a["kernel_ms"] for a in acc.values())
    if pass_ms > 0:
        roofline["whole_pass"] = {"algorithmic_bytes_per_step": pass_bytes / steps, "ms_per_step": pass_ms / steps,
                                  "achieved": pass_bytes / (pass_ms * 1e-3) / 1e9, "frac": pass_bytes / (pass_ms * 1e-3) / 1e9 / c.peak,
                                  "note": "every algorithmic byte of a5 + a7 over the synchronous, stage-timed pass (kernel span incl. partition and scans)"}
    traffic_file = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(traffic_file):
        try:
            roofline["traffic"] = json.load(open(traffic_file)).get(dom_name)
        except (OSError, ValueError):
            pass
    return roofline, pass_bytes


def pinned_sets(c, seeds, q, length, first):
    import torch

    out = []
    for seed in seeds:
        t3 = tuple(torch.empty(q, dtype=torch.int32, pin_memory=True) for _ in range(3))
        c.idx.synth_regions(seed, q, length, first=first, out=t3)
        out.append(t3)
    return out


def bench_intersect(c):
    """the headline: configs[1]"""
    args, idx, Q = c.args, c.idx, c.args.regions
    first = c.rank * Q
    dev_sets = [idx.synth_regions_device(BED_SEED + k, Q, REGION_LEN, first=first, device=c.dev) for k in range(len(PASSES))]
    host_sets = pinned_sets(c, [BED_SEED + k for k in range(len(PASSES))], Q, REGION_LEN, first)
    for _ in range(args.warmup):
        for (name, mode, inv, entire, types), regs in zip(SWEEP, dev_sets):
            from gffx_b200 import _lib
            queue_pass(c, regs, mode, inv, entire, types, _lib.F_PAIRS_ON_DEVICE | _lib.F_NO_HOST_COPY).close()
    sampler = ClockSampler(c.local_rank) if c.rank == 0 else None
    ms_total, loops, tally, (t0, t1) = timed_async(c, SWEEP, dev_sets, args.steps, args.min_seconds)
    clocks = sampler.stop(t0, t1) if sampler else None
    n_steps = loops * args.steps
    acc = staged_sync(c, SWEEP, dev_sets, args.steps)
    e2e_steps = max(1, min(args.steps, args.e2e_steps))
    e2e_sync_s, ranges_e2e = e2e_sync(c, SWEEP, host_sets, e2e_steps)
    e2e_s, e2e_api = e2e_sync_s, "gffx_intersect (C ABI) with pinned host region arrays; byte ranges copied back"
    if True:   # the same calls queued, two jobs in flight (a sharded job's flag exchange rides on its lane's stream)
        e2e_s, ranges_q = e2e_queued(c, SWEEP, host_sets, e2e_steps)
        assert ranges_q == ranges_e2e
        e2e_api = ("gffx_intersect" if c.world == 1 else "gffx_shard_begin / add / finish") + " (C ABI) with GFFX_F_ASYNC, pinned host region arrays in, byte ranges copied back to page-locked host memory, two jobs in flight; every result waited for and read inside the timed region"
    ceiling = h2d_ceiling(c, host_sets)
    pairs_total, launches_total, ranges_sum = allsum(c, [sum(a["pairs"] for a in tally.values()), sum(a["launches"] for a in tally.values()), ranges_e2e])
    c.dev_sets, c.host_sets = dev_sets, host_sets
    if c.rank != 0:
        return None
    roofline, _ = roofline_of(c, acc, Q, args.steps)
    value = c.world * len(PASSES) * Q * n_steps / (ms_total * 1e-3)
    h2d = len(PASSES) * 12 * Q
    d2h = int(16 * ranges_sum / c.world / e2e_steps) + 64 * len(PASSES)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": c.world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_total / n_steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u32",
        "data": "synthetic", "config": workload_config(args, idx.n_roots, Q),
        "timing": {"loops": loops, "steps_timed": n_steps, "timed_region_s": ms_total * 1e-3, "min_seconds": args.min_seconds,
                   "synchronous_stage_timed_ms_per_step": sum(a["kernel_ms"] for a in acc.values()) / args.steps},
        "index": {"n_lines": idx.n_lines, "n_roots": idx.n_roots, "n_seqs": idx.n_seqs, "virtual_sequences": idx.vseq(0)[3], "device_bytes": idx.device_bytes},
        "pairs_per_sec": pairs_total / (ms_total * 1e-3), "pairs_per_step": pairs_total / n_steps,
        "per_pass": {n: {"ms": a["kernel_ms"] / a["calls"], "pairs": a["pairs"] // a["calls"], "ranges": a["ranges"] // a["calls"],
                         "roots": a["roots"] // a["calls"], "launches": a["launches"] // a["calls"],
                         "ms_queued": tally[n]["kernel_ms"] / tally[n]["calls"]} for n, a in acc.items()},
        "roofline": roofline, "clocks": clocks,
        "e2e": {"value": c.world * len(PASSES) * Q * e2e_steps / e2e_s, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "steps": e2e_steps, "ms_per_step": e2e_s / e2e_steps * 1e3, "h2d_ceiling_gbs": ceiling["gbs_aggregate"], "h2d_ceiling": ceiling,
                "host_affinity": c.affinity, "api": e2e_api,
                "synchronous": {"value": c.world * len(PASSES) * Q * e2e_steps / e2e_sync_s, "ms_per_step": e2e_sync_s / e2e_steps * 1e3,
                                "d2h_bytes_per_step": int(16 * ranges_sum / c.world / e2e_steps) + 64 * len(PASSES),
                                "note": "the same passes as plain synchronous gffx_intersect calls, one after the other"}},
        "gpu_launches": int(launches_total),
    }
    return line


def bench_entire_group_types(c):
    """configs[2]: -e -T gene,mRNA,exon on the first region set (entire-group + type filter = the match-only path, intersect.rs:619)"""
    args, idx, Q = c.args, c.idx, c.args.regions
    passes = [("entire_group_types", 2, False, True, EG_TYPES)]
    ms_total, loops, tally, _ = timed_async(c, passes, c.dev_sets[:1], max(2, args.steps), args.min_seconds / 4)
    n_steps = loops * max(2, args.steps)
    acc = staged_sync(c, passes, c.dev_sets[:1], 3)
    e2e_s, ranges = e2e_sync(c, passes, c.host_sets[:1], 2)
    if c.rank != 0:
        return None
    a = acc["entire_group_types"]
    p, e_out = a["pairs"] / a["calls"], a["ranges"] / a["calls"]
    nbytes = 12 * Q + 12 * p + 12 * idx.n_lines + 16 * e_out   # a5 + a7: regions once, pairs, every line of the selected blocks, the ranges
    ms = a["kernel_ms"] / a["calls"]
    return {"workload": f"intersect -e -T {EG_TYPES}: {Q} regions per GPU, overlap mode", "ms": ms_total / n_steps, "value": c.world * Q * n_steps / (ms_total * 1e-3),
            "unit": UNIT, "pairs": int(p), "ranges": int(e_out), "roots": a["roots"] // a["calls"], "launches": a["launches"] // a["calls"],
            "roofline": {"bound": "hbm", "kernel": "whole pass (partition, join + scatter, select, scans, line filter, expansion)", "algorithmic_bytes": nbytes,
                         "ms": ms, "achieved": nbytes / (ms * 1e-3) / 1e9, "peak": c.peak, "unit": "GB/s", "frac": nbytes / (ms * 1e-3) / 1e9 / c.peak,
                         "note": "L_sel = every line (all blocks are selected at this density)"},
            "e2e": {"value": c.world * Q * 2 / e2e_s, "unit": UNIT, "h2d_bytes_per_step": 12 * Q, "d2h_bytes_per_step": int(16 * ranges / 2) + 64, "ms_per_step": e2e_s / 2 * 1e3}}


def bench_counts(c, kind):
    """configs[3]: depth / coverage from Q read intervals of 150 bases per GPU (N > 1: the sharded job, accumulators combined over NCCL)"""
    import torch

    args, idx, Q = c.args, c.idx, c.args.reads
    name = "depth" if kind == 0 else "coverage"
    reads = idx.synth_regions_device(READ_SEED, Q, READ_LEN, first=c.rank * Q, device=c.dev)
    fn = idx.depth if kind == 0 else idx.coverage
    fn(reads, combine=c.combiner, copy=False).close()
    c.barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    steps = 3
    ev0.record(c.stream)
    kms, launches, n_rows = 0.0, 0, 0
    for _ in range(steps):
        rows = fn(reads, combine=c.combiner, copy=False)
        kms += rows.kernel_ms
        launches += rows.launches
        n_rows = int(rows.id_idx.shape[0])
        rows.close()
    ev1.record(c.stream)
    c.barrier()
    ms_total = allmax(c, ev0.elapsed_time(ev1))
    # end to end: pinned host reads in, rows out
    host = tuple(torch.empty(Q, dtype=torch.int32, pin_memory=True) for _ in range(3))
    for h, d in zip(host, reads):
        h.copy_(d)
    del reads
    torch.cuda.synchronize()
    fn(host, combine=c.combiner, copy=False).close()
    c.barrier()
    te0 = time.perf_counter()
    for _ in range(2):
        fn(host, combine=c.combiner, copy=False).close()
    torch.cuda.synchronize()
    e2e_s = allmax(c, time.perf_counter() - te0)
    del host
    if c.rank != 0:
        return None
    nbytes = 12 * Q + 16 * idx.n_lines + 24 * n_rows
    ms_k = kms / steps
    return {"workload": f"{name}: {Q} read intervals of {READ_LEN} bases per GPU", "ms": ms_total / steps, "value": c.world * Q * steps / (ms_total * 1e-3), "unit": UNIT,
            "rows": n_rows, "launches": launches // steps,
            "roofline": {"bound": "hbm", "kernel": f"{name} job (partition + scatter + scans + rows)", "algorithmic_bytes": nbytes, "ms": ms_k,
                         "achieved": nbytes / (ms_k * 1e-3) / 1e9, "peak": c.peak, "unit": "GB/s", "frac": nbytes / (ms_k * 1e-3) / 1e9 / c.peak},
            "e2e": {"value": c.world * Q * 2 / e2e_s, "unit": UNIT, "h2d_bytes_per_step": 12 * Q, "d2h_bytes_per_step": 24 * n_rows, "ms_per_step": e2e_s / 2 * 1e3}}


def parity_n(c):
    """N > 1, outside every timed region: the sharded job on 200 k regions per rank gives every rank the byte ranges that ONE
    GPU computes over the union — and (rank 0) the oracle's"""
    import numpy as np

    from gffx_b200 import _lib

    idx, q = c.idx, 200_000
    ok = True
    for k, (name, mode, inv) in enumerate(PASSES):
        mine = idx.synth_regions_device(BED_SEED + 100 + k, q, REGION_LEN, first=c.rank * q, device=c.dev)
        union = idx.synth_regions_device(BED_SEED + 100 + k, q * c.world, REGION_LEN, first=0, device=c.dev)
        a = idx.shard_intersect(mine, mode, inv, False, None, combine=c.combiner)
        b = idx.intersect(union, mode, inv, False, None)
        ok = ok and a.n_ranges == b.n_ranges and np.array_equal(a.ranges, b.ranges)
        if c.rank == 0 and k == 0 and not c.args.no_cpu:
            orc, _ = open_oracle(c.gff)
            with orc:
                want, _, _ = orc.intersect(idx.synth_regions(BED_SEED + 100 + k, q * c.world, REGION_LEN, first=0), mode, inv, False, None)
            ok = ok and np.array_equal(b.ranges, want)
    return bool(allsum(c, [0.0 if ok else 1.0])[0] == 0.0)


def bench_sharded_1b(c):
    """configs[4], N > 1: 1 B regions in total, split over the ranks; one overlap intersect pass and one depth job, end to end
    on device-resident regions generated per shard (never materialised as text)"""
    import torch

    from gffx_b200 import _lib

    idx = c.idx
    q = c.args.big // c.world
    F = _lib.F_NO_HOST_COPY
    out = {"regions_total": q * c.world, "regions_per_gpu": q}
    regs = idx.synth_regions_device(BED_SEED + 7, q, REGION_LEN, first=c.rank * q, device=c.dev)
    for rep in range(2):
        c.barrier()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record(c.stream)
        r = c.idx.shard_intersect(regs, 2, False, False, None, combine=c.combiner, flags=F, batches=max(1, q // 50_000_000), copy=False)
        ev1.record(c.stream)
        n_ranges = r.n_ranges
        r.close()
        c.barrier()
        ms = allmax(c, ev0.elapsed_time(ev1))
    out["intersect"] = {"ms": ms, "value": q * c.world / (ms * 1e-3), "unit": UNIT, "ranges": int(n_ranges),
                        "roofline_frac": (12 * q + 12 * idx.n_lines + 16 * n_ranges) / (ms * 1e-3) / 1e9 / c.peak}
    del regs
    reads = idx.synth_regions_device(READ_SEED + 7, q, READ_LEN, first=c.rank * q, device=c.dev)
    for rep in range(2):
        c.barrier()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record(c.stream)
        rows = idx.depth(reads, combine=c.combiner, copy=False, batches=max(1, q // 100_000_000))
        ev1.record(c.stream)
        n_rows = int(rows.id_idx.shape[0])
        rows.close()
        c.barrier()
        ms = allmax(c, ev0.elapsed_time(ev1))
    out["depth"] = {"ms": ms, "value": q * c.world / (ms * 1e-3), "unit": UNIT, "rows": n_rows,
                    "roofline_frac": (12 * q + 16 * idx.n_lines + 24 * n_rows) / (ms * 1e-3) / 1e9 / c.peak}
    del reads
    return out if c.rank == 0 else None


def extras(c, line):
    """record emission (K8) and BED text parsing (K9) on the device: not part of the timed step"""
    import numpy as np

    from gffx_b200 import _lib

    idx, Q = c.idx, c.args.regions
    F_DEV = _lib.F_PAIRS_ON_DEVICE | _lib.F_NO_HOST_COPY
    name, mode, inv = PASSES[0]
    r = idx.intersect(c.dev_sets[0], mode, inv, False, None, flags=F_DEV, copy=False)
    nr = r.n_ranges
    best, nbytes = None, 0
    for _ in range(4):
        _, nbytes = idx.gather_text(None, nr, device=True)
        best = idx.gather_ms if best is None else min(best, idx.gather_ms)
    r.close()
    if best and nbytes:
        gbs = 2.0 * nbytes / (best * 1e-3) / 1e9
        line["emit"] = {"kernel": "gather_text_kernel", "pass": name, "ranges": int(nr), "bytes_out": int(nbytes), "ms": best, "scan_ms": idx.gather_scan_ms,
                        "algorithmic_bytes": 2 * int(nbytes), "achieved": gbs, "peak": c.peak, "unit": "GB/s", "frac": gbs / c.peak}
    import pandas as pd

    n_bed = min(Q, 10_000_000)
    names = np.array(idx.num_to_seqid, dtype=object)
    h = c.host_sets[0]
    bed_path = os.path.join(workload_dir(c.args.genes), f"bench_{n_bed}.bed")
    if not os.path.exists(bed_path):
        pd.DataFrame({"c": names[h[0][:n_bed].numpy()], "s": h[1][:n_bed].numpy(), "e": h[2][:n_bed].numpy()}).to_csv(bed_path, sep="\t", header=False, index=False)
    best = None
    for _ in range(3):
        t = idx.parse_bed_device(bed_path, 0, device=c.local_rank)
        best = idx.bed_parse_ms if best is None else min(best, idx.bed_parse_ms)
    n_parsed = int(t[0].numel())
    t0 = time.perf_counter()
    hp = idx.parse_bed_file(bed_path)
    host_ms = (time.perf_counter() - t0) * 1e3
    nbytes = idx.bed_bytes + 12 * n_parsed
    line["bed_parse"] = {"kernels": "bed_count_nl + bed_line_starts + bed_parse_lines + scan + bed_compact", "lines": n_parsed, "text_bytes": idx.bed_bytes,
                         "host_decided_lines": idx.bed_host_lines, "ms": best, "lines_per_sec": n_parsed / (best * 1e-3),
                         "algorithmic_bytes": nbytes, "achieved": nbytes / (best * 1e-3) / 1e9, "peak": c.peak, "unit": "GB/s",
                         "frac": nbytes / (best * 1e-3) / 1e9 / c.peak, "host_parser_ms_1_thread": host_ms, "same_as_host": bool(
                             len(hp[0]) == n_parsed and np.array_equal(hp[1], t[1].cpu().numpy().view(np.uint32)))}


def run_gpu(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the engine has no CPU fallback (use --impl reference for the CPU arm)")
    c = make_ctx(args, rank, world, local_rank)
    line = bench_intersect(c)
    subs = {}
    if "eg" in args.parts:
        subs["entire_group_types"] = bench_entire_group_types(c)
    if world == 1 and "extras" in args.parts:
        extras(c, line)
    c.dev_sets = c.dev_sets[:1]
    c.host_sets = None
    torch.cuda.empty_cache()
    if "counts" in args.parts:
        subs["depth"] = bench_counts(c, 0)
        subs["coverage"] = bench_counts(c, 1)
    par = None
    if world > 1:
        par = parity_n(c)
        if "big" in args.parts:
            c.dev_sets = None
            torch.cuda.empty_cache()
            subs["sharded_1b"] = bench_sharded_1b(c)
    if rank == 0:
        line.update({k: v for k, v in subs.items() if v is not None})
        if par is not None:
            line["parity_n"] = par
        if world == 1 and not args.no_cpu:   # the oracle on bounded samples of the same workloads, all host threads (rank 0, N = 1 only)
            orc, cores = open_oracle(c.gff)
            with orc:
                ext = oracle_extents(orc)
                line["cpu_baseline"] = cpu_baseline_intersect(orc, ext, args.regions, args.cpu_seconds, SWEEP, cores)
                if "entire_group_types" in line:
                    line["entire_group_types"]["cpu_baseline"] = cpu_baseline_intersect(orc, ext, args.regions, args.cpu_seconds / 2,
                                                                                       [("entire_group_types", 2, False, True, EG_TYPES)], cores)
                for k, name in ((0, "depth"), (1, "coverage")):
                    if name in line:
                        line[name]["cpu_baseline"] = cpu_baseline_counts(orc, ext, k, args.reads, args.cpu_seconds / 2, cores)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--regions", type=int, default=10_000_000, help="BED regions per pass per GPU")
    ap.add_argument("--reads", type=int, default=100_000_000, help="read intervals per GPU of the depth / coverage jobs")
    ap.add_argument("--big", type=int, default=1_000_000_000, help="regions of the sharded job of configs[4] (N > 1), in total")
    ap.add_argument("--genes", type=int, default=60000)
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--min-seconds", type=float, default=1.0, help="the timed region is repeated until it lasts this long")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="target CPU time of the in-line cpu_baseline sample")
    ap.add_argument("--ref-seconds", type=float, default=150.0, help="target total CPU time of --impl reference")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-bind", action="store_true", help="do not bind the process to the GPU's CPU set")
    ap.add_argument("--parts", default="eg,counts,extras,big", help="sub-benchmarks to run besides the headline (comma list of eg,counts,extras,big)")
    args = ap.parse_args()
    args.parts = set(p for p in args.parts.split(",") if p)
    if args.warmup < 3 and args.impl == "b200":
        args.warmup = 3  # timing rule: W >= 3
    rank, world, local_rank = env_int("RANK", 0), env_int("WORLD_SIZE", 1), env_int("LOCAL_RANK", 0)
    if args.impl == "reference":
        return run_reference(args, rank, world)
    return run_gpu(args, rank, world, local_rank)


if __name__ == "__main__":
    sys.exit(main())
