#!/usr/bin/env python
"""bench.py — the hot path of GFFx on B200: `intersect --bed` over synthetic human-scale GFF3.

Workload (BASELINE.json configs[1]): 60k-gene human-like GFF3 (~2.7 M lines, seed
20260001) x 10 M synthetic 20 kb BED regions per GPU.  One STEP = the configuration's mode sweep, four
full intersect passes (root-level join with the (root_fid,start,end) pairs materialised in HBM, root ->
block selection, line-level filter, ordered compaction into output byte ranges):
    overlap | --contained | --contains-region | --contained --invert
each pass over its OWN 10 M-region set (4 x 120 MB of inputs rotate through the 126 MB L2, so no pass
finds its inputs cached).  value = query intervals/s = n_gpus * 4 * Q * steps / time.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

N > 1 is launched by torchrun, one rank per GPU; every rank joins its own regions against a replicated
index (weak scaling; shards are independent, the only exchange is the element-wise combination of the job's
accumulators — root hit flags, line rank-space keys — before the final line filter; gffx_b200/shard.py).

--impl reference times the CPU implementation of the path (the C++ restatement under oracle/: the Rust
reference cannot be built in this image) on the box's host cores, on a bounded sample of the workload.
"""
import argparse
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
REGION_LEN = 20000
# (name, mode, invert) — OverlapMode: 0 contained, 1 contains-region, 2 overlap
PASSES = [("overlap", 2, False), ("contained", 0, False), ("contains_region", 1, False), ("contained_invert", 0, True)]
METRIC = "query_intervals_per_sec"
UNIT = "intervals/s"


def env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


def workload_dir(genes):
    return os.environ.get("GFFX_BENCH_DIR", f"/tmp/gffx_b200_bench_{GFF_SEED}_{genes}")


def ensure_workload(genes, builder):
    """Synthetic GFF + its index files, written once per box.  `builder(path)` writes the index."""
    import gffx_b200 as g

    d = workload_dir(genes)
    gff = os.path.join(d, "hs_synth.gff")
    done = os.path.join(d, ".done")
    if not os.path.exists(done):
        os.makedirs(d, exist_ok=True)
        g.synth_gff(gff, seed=GFF_SEED, n_genes=genes, quirks=False)
        builder(gff)
        open(done, "w").write("ok\n")
    return gff


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
# CPU arm: the C++ restatement of the reference (oracle/), timed on the host cores.  This function and
# tests/ are the only places that may execute oracle/ code; nothing below it is on the product path.
# ---------------------------------------------------------------------------------------------------
def cpu_reference_step(orc, region_sets):
    """one step of the same mode sweep through the oracle's intersect (intersect.rs:586-654 restated)"""
    n = 0
    for (name, mode, inv), regs in zip(PASSES, region_sets):
        orc.intersect(regs, mode, inv, False, None)
        n += regs[0].shape[0]
    return n


def cpu_sample_sets(idx, q_sample, rank=0):
    return [idx.synth_regions(BED_SEED + k, q_sample, REGION_LEN, first=0) for k in range(len(PASSES))]


def calibrate_cpu_sample(orc, idx, q_full, budget_s):
    """largest sample (regions per pass) whose step is expected to fit `budget_s`, from a 50k-region probe.
    The reference's line scan saturates with Q (it stops at the first matching region), so a linear
    extrapolation of the probe over-estimates the time: the sample errs on the small side."""
    probe = 50000
    sets = cpu_sample_sets(idx, probe)
    t = time.perf_counter()
    cpu_reference_step(orc, sets)
    dt = max(time.perf_counter() - t, 1e-3)
    q = int(probe * max(1.0, budget_s / dt))
    q = min(q_full, max(probe, (q // probe) * probe))
    return q, dt


def run_reference(args, rank, world):
    if rank != 0:
        return 0
    import gffx_b200 as g
    from tests import _oracle

    gff = ensure_workload(args.genes, _oracle.build_index)
    idx = g.TreeIndexData.load_tree_index(gff, device=-1)  # host-only: workload generators, no GPU
    cores = _oracle.max_threads()
    with _oracle.Oracle(gff) as orc:
        budget = args.ref_seconds / max(1, args.steps + args.warmup)
        q_s, _ = calibrate_cpu_sample(orc, idx, args.regions, budget)
        sets = cpu_sample_sets(idx, q_s)
        for _ in range(args.warmup):
            cpu_reference_step(orc, sets)
        t0 = time.perf_counter()
        n = 0
        for _ in range(args.steps):
            n += cpu_reference_step(orc, sets)
        dt = time.perf_counter() - t0
    val = n / dt
    sample = f"{q_s} of {args.regions} regions per pass x {len(PASSES)} passes per step (host RAM resident, all {cores} threads)"
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": workload_config(args, idx, args.regions),
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample,
                         "note": "C++ restatement of GFFx v0.4.0 (oracle/); the Rust reference cannot be compiled in this image"},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


def workload_config(args, idx, q):
    return {
        "workload": f"intersect --bed: synthetic human-like GFF3 ({args.genes} genes, {idx.n_lines} lines, seed {GFF_SEED}) x {q} "
                    f"20 kb BED regions per pass per GPU; step = overlap | --contained | --contains-region | --contained --invert",
        "regions_per_pass_per_gpu": q, "passes_per_step": len(PASSES), "region_len": REGION_LEN, "n_roots": idx.n_roots,
        "n_lines": idx.n_lines, "pairs": "materialised in HBM",
        "timing": "value / ms_per_step: K steps between two CUDA events on the engine's stream, calls without per-stage events; "
                  "roofline / per_pass: the same K steps repeated with GFFX_F_STAGE_TIMES (per-stage CUDA events inside every call, "
                  "~40 us per call; roofline.staged_run_ms_per_step is that run's step time)",
        "l2": "inputs larger than L2: 4 x 12 B x Q region sets rotate (no flush needed)",
        "sharding": "regions split across ranks (counter-based stream, rank r owns [r*Q,(r+1)*Q)); index replicated; N>1: the job's "
                    "byte flags (roots hit | lines matched, one array) are combined by one NCCL all-reduce(MAX) per pass "
                    "and every rank emits the byte ranges of the whole N*Q-region job",
    }


# ---------------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------------
# what each timed stage of one intersect call is, and its ALGORITHMIC bytes (every input
# read once, every output written once; re-reads and scan passes the implementation adds are NOT credited)
JOIN_NAME = "root join + rank-space scatter (wj_count_kernel, offsets scan, wj_expand_kernel)"


def stage_table(q, p, l_sel, e_out):
    return {
        8: ("window partition (hist + scan + scatter)", 0),  # an implementation pass: no algorithmic bytes credited
        # regions (3 x u32) read once — by the join and, in the same kernel, by the line filter's scatter (a7's 12*Q is NOT
        # credited a second time) — and pairs (3 x u32) written once; stage 7, the expansion, is added to stage 0
        0: (JOIN_NAME, 12 * q + 12 * p),
        3: ("select_blocks + merge_blocks", 0),
        5: ("rank-space scans", 0),
        6: ("eval_lines + emit_ranges", 12 * l_sel + 16 * e_out),
    }


def run_gpu(args, rank, world, local_rank):
    import numpy as np
    import torch
    import torch.distributed as dist

    import gffx_b200 as g
    from gffx_b200 import _lib

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the engine has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    if rank == 0:
        gff = ensure_workload(args.genes, g.build_index)
    barrier()
    gff = os.path.join(workload_dir(args.genes), "hs_synth.gff")
    idx = g.TreeIndexData.load_tree_index(gff, device=local_rank)
    Q = args.regions
    first = rank * Q
    dev_sets = [idx.synth_regions_device(BED_SEED + k, Q, REGION_LEN, first=first, device=dev) for k in range(len(PASSES))]
    host_sets = []
    for k in range(len(PASSES)):
        t3 = tuple(torch.empty(Q, dtype=torch.int32, pin_memory=True) for _ in range(3))
        idx.synth_regions(BED_SEED + k, Q, REGION_LEN, first=first, out=t3)
        host_sets.append(t3)
    stream = torch.cuda.ExternalStream(idx.cuda_stream, device=dev)
    F_DEV = _lib.F_PAIRS_ON_DEVICE | _lib.F_NO_HOST_COPY

    combiner = None
    if world > 1:
        from gffx_b200.shard import Combiner

        combiner = Combiner(idx)  # exchanges the job's accumulators over NCCL before the finish step (gffx_b200/shard.py)

    def step(sets, flags, acc=None):
        for (name, mode, inv), regs in zip(PASSES, sets):
            # copy=False: the byte ranges stay in the result's page-locked buffer (what a Rust caller of the C ABI reads too)
            if world == 1:
                r = idx.intersect(regs, mode, inv, False, None, flags=flags, copy=False)
            else:  # every rank joins its own Q regions; after the exchange each rank holds the ranges of the world's N*Q regions
                r = idx.shard_intersect(regs, mode, inv, False, None, combine=combiner, flags=flags, copy=False)
            if acc is not None:
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

    # ---- device-resident arm: K steps between two events on the engine's stream --------------------
    for _ in range(args.warmup):
        step(dev_sets, F_DEV)
    tally = {}   # pairs / ranges / launches of the timed steps (no per-stage events inside them: what a caller of the ABI runs)
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    ev0.record(stream)
    for _ in range(args.steps):
        step(dev_sets, F_DEV, tally)
    ev1.record(stream)
    barrier()
    t1 = time.perf_counter()
    clocks = sampler.stop(t0, t1) if sampler else None
    ms = torch.tensor([ev0.elapsed_time(ev1)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_total = float(ms.item())
    # the same K steps once more with GFFX_F_STAGE_TIMES: the per-stage CUDA events behind `roofline` (a dozen events and
    # their read-back cost ~40 us per call, which is why the timed steps above run without them)
    acc = {}
    evs0, evs1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    evs0.record(stream)
    for _ in range(args.steps):
        step(dev_sets, F_DEV | _lib.F_STAGE_TIMES, acc)
    evs1.record(stream)
    torch.cuda.synchronize()
    ms_staged = evs0.elapsed_time(evs1)

    # ---- end-to-end arm: the same calls with pinned HOST inputs and the byte ranges read back -------
    e2e_steps = max(1, min(args.steps, args.e2e_steps))
    step(host_sets, _lib.F_PAIRS_ON_DEVICE)
    acc_e = {}
    barrier()
    te0 = time.perf_counter()
    for _ in range(e2e_steps):
        step(host_sets, _lib.F_PAIRS_ON_DEVICE, acc_e)
    torch.cuda.synchronize()
    te = torch.tensor([time.perf_counter() - te0], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_s = float(te.item())

    # ---- totals over ranks --------------------------------------------------------------------------
    tot = torch.tensor([sum(a["pairs"] for a in tally.values()), sum(a["launches"] for a in tally.values()),
                        sum(a["ranges"] for a in acc_e.values())], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    pairs_total, launches_total, ranges_e2e = (float(x) for x in tot.tolist())
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    n_intervals = world * len(PASSES) * Q * args.steps
    value = n_intervals / (ms_total * 1e-3)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except (OSError, ValueError):
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback (B200_PROFILING.md 6.65 TB/s)"

    # dominant kernel over the step: largest share of the event-timed stage durations (rank 0)
    per_kernel = {}
    for name, a in acc.items():
        q, p = Q, a["pairs"] / a["calls"]
        e_out = a["ranges"] / a["calls"]
        tab = stage_table(q, p, idx.n_lines, e_out)
        for s, (kname, nbytes) in tab.items():
            k = per_kernel.setdefault(kname, {"ms": 0.0, "bytes": 0.0, "launches": 0})
            if a["stage_ms"][s] > 0:
                k["ms"] += a["stage_ms"][s] + (a["stage_ms"][7] if s == 0 else 0.0)
                k["bytes"] += nbytes * a["calls"]
                k["launches"] += a["calls"]
    stage_sum = sum(k["ms"] for k in per_kernel.values()) or 1.0
    dom_name, dom = max(((n, k) for n, k in per_kernel.items() if k["bytes"] > 0), key=lambda x: x[1]["ms"])
    achieved = dom["bytes"] / (dom["ms"] * 1e-3) / 1e9 if dom["ms"] > 0 else 0.0
    roofline = {"bound": "hbm", "kernel": dom_name, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": None, "peak_source": peak_src, "avg_launch_ms": dom["ms"] / max(1, dom["launches"]),
                "algorithmic_bytes_per_launch": dom["bytes"] / max(1, dom["launches"]), "share_of_step": dom["ms"] / stage_sum,
                "kernels": {n: {"ms_per_step": k["ms"] / args.steps, "share": k["ms"] / stage_sum,
                                "GBps": (k["bytes"] / (k["ms"] * 1e-3) / 1e9 if k["ms"] > 0 else 0.0)} for n, k in per_kernel.items()}}
    # the whole pass against the same roof: every algorithmic byte of SURVEY.md 8(d) for a5 + a7 (each credited once)
    # over the pass's whole kernel span, partition and scans included
    pass_bytes = sum(k["bytes"] for k in per_kernel.values())
    pass_ms = sum(a["kernel_ms"] for a in acc.values())
    if pass_ms > 0:
        roofline["staged_run_ms_per_step"] = ms_staged / args.steps
        roofline["whole_pass"] = {"algorithmic_bytes_per_step": pass_bytes / args.steps, "ms_per_step": pass_ms / args.steps,
                                  "achieved": pass_bytes / (pass_ms * 1e-3) / 1e9, "frac": pass_bytes / (pass_ms * 1e-3) / 1e9 / peak}
    traffic_file = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(traffic_file):
        try:
            roofline["traffic"] = json.load(open(traffic_file)).get(dom_name)
        except (OSError, ValueError):
            pass

    # ---- record emission (K8, not part of the timed step): the text of the overlap pass's ranges gathered on the device ----
    emit = None
    if world == 1 and not getattr(args, "no_emit", False):
        name, mode, inv = PASSES[0]
        r = idx.intersect(dev_sets[0], mode, inv, False, None, flags=F_DEV, copy=False)
        nr = r.n_ranges
        best, nbytes = None, 0
        for _ in range(4):
            _, nbytes = idx.gather_text(None, nr, device=True)
            best = idx.gather_ms if best is None else min(best, idx.gather_ms)
        r.close()
        if best and nbytes:
            gbs = 2.0 * nbytes / (best * 1e-3) / 1e9
            emit = {"kernel": "gather_text_kernel", "pass": name, "ranges": int(nr), "bytes_out": int(nbytes), "ms": best, "scan_ms": idx.gather_scan_ms,
                    "algorithmic_bytes": 2 * int(nbytes), "achieved": gbs, "peak": peak, "unit": "GB/s", "frac": gbs / peak}

    # ---- BED text parsing on the device (K9, not part of the timed step): a 3-column BED of the first pass's regions ----
    bed_parse = None
    if world == 1 and not getattr(args, "no_emit", False):
        import pandas as pd
        n_bed = min(Q, 10_000_000)
        names = np.array(idx.num_to_seqid, dtype=object)
        h = host_sets[0]
        bed_path = os.path.join(workload_dir(args.genes), f"bench_{n_bed}.bed")
        if not os.path.exists(bed_path):
            pd.DataFrame({"c": names[h[0][:n_bed].numpy()], "s": h[1][:n_bed].numpy(), "e": h[2][:n_bed].numpy()}).to_csv(
                bed_path, sep="\t", header=False, index=False)
        best = None
        for _ in range(3):
            t = idx.parse_bed_device(bed_path, 0, device=local_rank)
            best = idx.bed_parse_ms if best is None else min(best, idx.bed_parse_ms)
        n_parsed = int(t[0].numel())
        t0 = time.perf_counter()
        hp = idx.parse_bed_file(bed_path)
        host_ms = (time.perf_counter() - t0) * 1e3
        nbytes = idx.bed_bytes + 12 * n_parsed
        bed_parse = {"kernels": "bed_count_nl + bed_line_starts + bed_parse_lines + scan + bed_compact", "lines": n_parsed, "text_bytes": idx.bed_bytes,
                     "host_decided_lines": idx.bed_host_lines, "ms": best, "lines_per_sec": n_parsed / (best * 1e-3),
                     "algorithmic_bytes": nbytes, "achieved": nbytes / (best * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                     "frac": nbytes / (best * 1e-3) / 1e9 / peak, "host_parser_ms_1_thread": host_ms, "same_as_host": bool(
                         len(hp[0]) == n_parsed and np.array_equal(hp[1], t[1].cpu().numpy().view(np.uint32)))}

    h2d = len(PASSES) * 12 * Q
    d2h = int(16 * ranges_e2e / world / e2e_steps) + 64 * len(PASSES)
    e2e_val = world * len(PASSES) * Q * e2e_steps / e2e_s
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u32",
        "data": "synthetic", "config": workload_config(args, idx, Q),
        "pairs_per_sec": pairs_total / (ms_total * 1e-3), "pairs_per_step": pairs_total / args.steps,
        "per_pass": {n: {"ms": a["kernel_ms"] / a["calls"], "pairs": a["pairs"] // a["calls"], "ranges": a["ranges"] // a["calls"],
                         "roots": a["roots"] // a["calls"], "launches": a["launches"] // a["calls"],
                         "ms_without_stage_events": tally[n]["kernel_ms"] / tally[n]["calls"]} for n, a in acc.items()},
        "roofline": roofline, "clocks": clocks,
        "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "steps": e2e_steps,
                "ms_per_step": e2e_s / e2e_steps * 1e3, "api": "gffx_intersect (C ABI) with pinned host region arrays; byte ranges copied back"},
        "gpu_launches": int(launches_total),
    }
    if emit:
        line["emit"] = emit
    if bed_parse:
        line["bed_parse"] = bed_parse
    if world == 1 and not args.no_cpu:
        line["cpu_baseline"] = cpu_baseline(args, idx, gff)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


# ---------------------------------------------------------------------------------------------------
# depth / coverage (BASELINE.json configs[3]): --workload depth|coverage.  One step = one whole job over Q read
# intervals of 150 bases (regions -> histograms / bitmap, scans, per-ID rows).  Not the default line of the contract
# (that is the intersect sweep); run by hand and recorded under profiles/.
# ---------------------------------------------------------------------------------------------------
def run_counts(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist

    import gffx_b200 as g

    kind = 0 if args.workload == "depth" else 1
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    if rank == 0:
        ensure_workload(args.genes, g.build_index)
    barrier()
    gff = os.path.join(workload_dir(args.genes), "hs_synth.gff")
    idx = g.TreeIndexData.load_tree_index(gff, device=local_rank)
    Q = args.regions
    reads = idx.synth_regions_device(20260010, Q, 150, first=rank * Q, device=dev)
    combiner = None
    if world > 1:
        from gffx_b200.shard import Combiner

        combiner = Combiner(idx)
    fn = idx.depth if kind == 0 else idx.coverage
    stream = torch.cuda.ExternalStream(idx.cuda_stream, device=dev)
    # copy=False: the rows stay in the job's page-locked buffers (what a caller of the C ABI reads too)
    for _ in range(args.warmup):
        fn(reads, combine=combiner, copy=False).close()
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    ev0.record(stream)
    kms, launches = 0.0, 0
    for _ in range(args.steps):
        rows = fn(reads, combine=combiner, copy=False)
        kms += rows.kernel_ms
        launches += rows.launches
        n_rows = int(rows.id_idx.shape[0])
        rows.close()
    ev1.record(stream)
    barrier()
    t1 = time.perf_counter()
    clocks = sampler.stop(t0, t1) if sampler else None
    ms = torch.tensor([ev0.elapsed_time(ev1)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_total = float(ms.item())
    # end to end: pinned host reads in, rows out
    e2e = None
    if args.e2e_steps > 0:
        host = tuple(torch.empty(Q, dtype=torch.int32, pin_memory=True) for _ in range(3))
        idx.synth_regions(20260010, Q, 150, first=rank * Q, out=host)
        fn(host, combine=combiner, copy=False).close()
        barrier()
        te0 = time.perf_counter()
        for _ in range(args.e2e_steps):
            r2 = fn(host, combine=combiner, copy=False)
            n_rows_e2e = int(r2.id_idx.shape[0])
            r2.close()
        torch.cuda.synchronize()
        te = torch.tensor([time.perf_counter() - te0], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        e2e_s = float(te.item())
        e2e = {"value": world * Q * args.e2e_steps / e2e_s, "unit": UNIT, "h2d_bytes_per_step": 12 * Q,
               "d2h_bytes_per_step": n_rows_e2e * 24, "steps": args.e2e_steps, "ms_per_step": e2e_s / args.e2e_steps * 1e3}
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except (OSError, ValueError):
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    # B = 12*Q + 16*L + 24*F_out per job, over the job's kernels (event-timed inside the library)
    nbytes = 12 * Q + 16 * idx.n_lines + 24 * n_rows
    ach = nbytes * args.steps / (kms * 1e-3) / 1e9 if kms > 0 else 0.0
    line = {
        "metric": METRIC, "value": world * Q * args.steps / (ms_total * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u32", "data": "synthetic",
        "config": {"workload": f"{args.workload}: synthetic human-like GFF3 ({args.genes} genes, {idx.n_lines} lines) x {Q} 150-base read intervals per GPU per step",
                   "rows": n_rows, "l2": "inputs larger than L2 (12 B x Q)"},
        "roofline": {"bound": "hbm", "kernel": f"{args.workload} job (scatter + scans + rows)", "achieved": ach, "peak": peak, "unit": "GB/s",
                     "frac": ach / peak, "traffic": None, "kernel_ms_per_step": kms / args.steps, "algorithmic_bytes_per_step": nbytes},
        "clocks": clocks, "e2e": e2e, "gpu_launches": launches,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


def cpu_baseline(args, idx, gff):
    """rank 0, N=1 only: the oracle (C++ restatement) on a bounded sample of the same workload, all host threads"""
    from tests import _oracle

    cores = _oracle.max_threads()
    with _oracle.Oracle(gff) as orc:
        q_s, _ = calibrate_cpu_sample(orc, idx, args.regions, args.cpu_seconds)
        sets = cpu_sample_sets(idx, q_s)
        t0 = time.perf_counter()
        n = cpu_reference_step(orc, sets)
        dt = time.perf_counter() - t0
    return {"value": n / dt, "unit": UNIT, "cores": cores, "kind": "port", "seconds": dt,
            "sample": f"one step of the same mode sweep on {q_s} of {args.regions} regions per pass ({cores} OpenMP threads)",
            "note": "C++ restatement of GFFx v0.4.0 (oracle/); the Rust reference cannot be compiled in this image"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--regions", type=int, default=10_000_000, help="BED regions per pass per GPU")
    ap.add_argument("--genes", type=int, default=60000)
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="target CPU time of the in-line cpu_baseline sample")
    ap.add_argument("--ref-seconds", type=float, default=150.0, help="target total CPU time of --impl reference")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--workload", default="intersect", choices=["intersect", "depth", "coverage"],
                    help="intersect = the contract's line (configs[1]); depth / coverage = configs[3], reported by hand")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "b200":
        args.warmup = 3  # timing rule: W >= 3
    rank, world, local_rank = env_int("RANK", 0), env_int("WORLD_SIZE", 1), env_int("LOCAL_RANK", 0)
    if args.impl == "reference":
        return run_reference(args, rank, world)
    if args.workload != "intersect":
        return run_counts(args, rank, world, local_rank)
    return run_gpu(args, rank, world, local_rank)


if __name__ == "__main__":
    sys.exit(main())
