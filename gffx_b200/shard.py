"""Multi-GPU sharding of the hot path: one process per GPU, every rank joins ITS share of the regions
against a replicated index, and the only exchange is the element-wise combination of the job's accumulators before
the finish step — after which every rank holds the complete result.

Why accumulators and not results: the reference tests every line against ALL regions of the line's sequence
(src/commands/intersect.rs:621-633) inside blocks selected by ANY region (:598-617), so the union of per-shard outputs is
not the output; but "was this root hit by any region" and "does any region match this line" are ORs over the regions,
hence over the shards: a sealed intersect job (gffx_job_buffer) exposes exactly these two byte-flag arrays, ~1 byte per
root and per line.  depth's histograms add, coverage's bitmap ORs and its word difference array adds (counts.cuh).

Byte flags (MAX) and 32-bit counters (SUM) go through NCCL's all-reduce.  The other accumulators (unsigned 64-bit MAX /
MIN keys, bitmaps combined by OR) have no NCCL reduction: all-to-all of the W slices (each rank receives slice r of every
rank), one fold kernel over the W copies (gffx_fold), all-gather of the folded slices: 2 (W-1)/W n elements per rank over
NVLink instead of the (W-1) n of a plain all-gather; NCCL has no bitwise-OR / unsigned-64 MIN reduction, which is why the
reduction is our own kernel.  With the gloo backend (CPU tests) the same protocol runs on host arrays with a numpy fold.
"""
from __future__ import annotations

import ctypes as C
from typing import Callable, List, Optional, Tuple

import numpy as np

# gffx_reduce_op (include/gffx_b200.h) -> (numpy dtype, fold over axis 0)
OPS = {
    0: (np.uint8, lambda a: a.max(axis=0)),
    1: (np.uint64, lambda a: a.max(axis=0)),
    2: (np.uint64, lambda a: a.min(axis=0)),
    3: (np.uint32, lambda a: a.sum(axis=0, dtype=np.uint32)),
    4: (np.int32, lambda a: a.sum(axis=0, dtype=np.int32)),
    5: (np.uint64, lambda a: np.bitwise_or.reduce(a, axis=0)),
}


def shard_bounds(n: int, world: int, rank: int, align: int = 4096) -> Tuple[int, int]:
    """[lo, hi) of rank's share of n regions: equal chunks rounded up to `align` (the partition tile), the tail to the last"""
    if world <= 1:
        return 0, n
    chunk = -(-n // world)
    chunk = -(-chunk // align) * align
    lo = min(n, rank * chunk)
    hi = min(n, lo + chunk)
    return lo, hi


class _DevView:
    """a raw device pointer as a __cuda_array_interface__ producer (uint8)"""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (int(ptr), False), "version": 2}


def combine_array(host: np.ndarray, op: int, group=None) -> np.ndarray:
    """The exchange protocol on a host array (gloo): returns the element-wise combination over the group's ranks."""
    import torch
    import torch.distributed as dist

    dt, fold = OPS[op]
    a = np.ascontiguousarray(host, dtype=dt)
    W, r = dist.get_world_size(group), dist.get_rank(group)
    n = a.shape[0]
    if W == 1 or n == 0:
        return a.copy()
    seg = -(-n // W)
    send = np.zeros(W * seg, dtype=dt)
    send[:n] = a
    recv = np.empty_like(send)
    ts, tr = torch.from_numpy(send.view(np.uint8)), torch.from_numpy(recv.view(np.uint8))
    dist.all_to_all_single(tr, ts, group=group)
    mine = np.ascontiguousarray(fold(recv.reshape(W, seg)))
    out = np.empty(W * seg, dtype=dt)
    parts = [torch.empty(seg * a.itemsize, dtype=torch.uint8) for _ in range(W)]
    dist.all_gather(parts, torch.from_numpy(mine.view(np.uint8)), group=group)
    for k, p in enumerate(parts):
        out.view(np.uint8)[k * seg * a.itemsize:(k + 1) * seg * a.itemsize] = p.numpy()
    return out[:n]


class Combiner:
    """Callable handed to TreeIndexData.shard_intersect / depth / coverage: combines every accumulator of the open job over
    the ranks of a torch.distributed group (NCCL), in place, on the engine's stream."""

    def __init__(self, idx, group=None):
        import torch
        import torch.distributed as dist

        self.idx = idx
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self._streams = {}
        self.device = torch.device("cuda", torch.cuda.current_device())
        self.bytes_exchanged = 0
        self._scratch = {}
        self._views = {}

    def _buf(self, name, nbytes, device):
        import torch

        t = self._scratch.get(name)
        if t is None or t.numel() < nbytes:
            t = torch.empty(int(nbytes * 1.25) + 256, dtype=torch.uint8, device=device)
            self._scratch[name] = t
        return t[:nbytes]

    def __call__(self, buffers: List[Tuple[int, int, int]]):
        """buffers: (device pointer, elements, gffx_reduce_op) as reported by gffx_job_buffer"""
        import torch
        import torch.distributed as dist

        from . import _lib

        if self.world == 1:
            return
        W, r = self.world, self.rank
        sp = self.idx.cuda_stream   # the open job's lane (queued jobs alternate between two streams)
        stream = self._streams.get(sp)
        if stream is None:
            stream = self._streams[sp] = torch.cuda.ExternalStream(sp, device=self.device)
        with torch.cuda.stream(stream):
            for ptr, n, op in buffers:
                if n == 0:
                    continue
                es = np.dtype(OPS[op][0]).itemsize
                seg = -(-n // W)
                t = self._views.get((ptr, n * es))   # the engine's accumulators keep their addresses from job to job
                if t is None:
                    t = self._views[(ptr, n * es)] = torch.as_tensor(_DevView(ptr, n * es), device=self.device)
                if op == 0:      # byte flags: NCCL's own MAX
                    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
                    self.bytes_exchanged += 2 * (W - 1) * n // W
                    continue
                if op in (3, 4):  # 32-bit sums (wrap-around addition is the same for signed and unsigned)
                    dist.all_reduce(t.view(torch.int32), op=dist.ReduceOp.SUM, group=self.group)
                    self.bytes_exchanged += 2 * (W - 1) * n * 4 // W
                    continue
                # unsigned 64-bit MAX / MIN and bitwise OR: no NCCL reduction exists -> slices, our fold kernel, gather
                send = self._buf("send", W * seg * es, t.device)
                recv = self._buf("recv", W * seg * es, t.device)
                send[: n * es].copy_(t)
                if W * seg > n:
                    send[n * es:].zero_()
                dist.all_to_all_single(recv, send, group=self.group)
                mine = self._buf("mine", seg * es, t.device)
                rc = _lib.lib.gffx_fold(self.idx._h, mine.data_ptr(), recv.data_ptr(), W, seg, seg, op)
                if rc != 0:
                    raise RuntimeError(_lib.last_error())
                dist.all_gather_into_tensor(send, mine, group=self.group)
                t.copy_(send[: n * es])
                self.bytes_exchanged += 2 * (W - 1) * seg * es
        # no host synchronisation: the finish step is queued on the same (engine) stream


def merge_count_rows(rows_per_shard: List[dict], kind: int) -> dict:
    """Host-side alternative for depth when only rank 0 needs the table ("gather to 0"): per-ID rows of the
    shards merged as the reference merges batches (src/commands/depth.rs:501-508): depth summed, start = min, end = max.
    rows: {id: (chrom, start, end, value)}.  Not valid for coverage (a union is not additive over region chunks)."""
    if kind != 0:
        raise ValueError("coverage rows of region-chunk shards cannot be merged; combine the accumulators instead")
    out: dict = {}
    for rows in rows_per_shard:
        for k, (chrom, s, e, v) in rows.items():
            if k in out:
                c0, s0, e0, v0 = out[k]
                out[k] = (c0, min(s0, s), max(e0, e), v0 + v)
            else:
                out[k] = (chrom, s, e, v)
    return out
