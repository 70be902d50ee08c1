"""Host-side mirror of the reference's public interface for the hot path, over the C ABI.

Names follow the reference's Rust API (``src/lib.rs:7-22``): ``TreeIndexData.load_tree_index``,
``parse_bed_file``, ``parse_region``, ``query_features``, ``OverlapMode``, ``IntersectArgs`` /
``run`` (``src/commands/intersect.rs``), ``DepthArgs`` / ``CoverageArgs`` and their ``run``
(``src/commands/depth.rs``, ``src/commands/coverage.rs``).  Errors surface as ``GffxError`` with
the message the engine reports (the reference's ``anyhow::Error`` reaching ``main``).
"""
import ctypes as C
import enum
from dataclasses import dataclass, field
from typing import List, Optional, Sequence, Tuple

import numpy as np

from . import _lib
from ._lib import lib


class GffxError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"[{code}] {msg}")
        self.code = code
        self.msg = msg


def _check(rc: int):
    if rc != 0:
        raise GffxError(rc, _lib.last_error())


class OverlapMode(enum.IntEnum):
    """src/commands/intersect.rs:74-78"""

    Contained = 0
    ContainsRegion = 1
    Overlap = 2


def _is_torch_cuda(x) -> bool:
    return hasattr(x, "is_cuda") and bool(x.is_cuda)


def _as_region_ptrs(regions, keep: list):
    """regions = (chr, start, end) as numpy uint32 arrays (host) or torch cuda int32/uint32 tensors."""
    chr_, s, e = regions
    if _is_torch_cuda(chr_):
        for t in (chr_, s, e):
            if not t.is_contiguous() or t.element_size() != 4:
                raise ValueError("device regions must be contiguous 4-byte tensors")
        n = int(chr_.numel())
        if int(s.numel()) != n or int(e.numel()) != n:
            raise ValueError("region arrays differ in length")
        keep.extend([chr_, s, e])
        return chr_.data_ptr(), s.data_ptr(), e.data_ptr(), n, _lib.F_DEVICE_INPUT
    if hasattr(chr_, "data_ptr") and hasattr(chr_, "numel"):  # torch CPU tensor (e.g. pinned)
        n = int(chr_.numel())
        keep.extend([chr_, s, e])
        return chr_.data_ptr(), s.data_ptr(), e.data_ptr(), n, 0
    arrs = [np.ascontiguousarray(a, dtype=np.uint32) for a in (chr_, s, e)]
    n = arrs[0].shape[0]
    if arrs[1].shape[0] != n or arrs[2].shape[0] != n:
        raise ValueError("region arrays differ in length")
    keep.extend(arrs)
    return arrs[0].ctypes.data, arrs[1].ctypes.data, arrs[2].ctypes.data, n, 0


@dataclass
class Result:
    """What one join / intersect call produced."""

    n_queries: int = 0
    n_pairs: int = 0
    pairs: Optional[np.ndarray] = None  # (n_pairs, 3) uint32: root_fid, start, end
    pairs_device_ptr: int = 0
    n_roots: int = 0
    n_ranges: int = 0
    ranges: Optional[np.ndarray] = None  # (n_ranges, 2) uint64 byte ranges, output order
    bytes_out: int = 0
    warn_fids: List[int] = field(default_factory=list)
    kernel_ms: float = 0.0
    stage_ms: Tuple[float, ...] = ()
    launches: int = 0
    _h: object = None  # the C result handle when `pairs` / `ranges` are zero-copy views of library-owned (pinned) memory

    def close(self):
        """release the library-owned arrays behind zero-copy `pairs` / `ranges` views (they must not be used afterwards)"""
        if self._h is not None:
            h, self._h = self._h, None
            self.pairs = self.ranges = None
            lib.gffx_result_free(h)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class PendingResult:
    """A GFFX_F_ASYNC job that has been queued on the engine's stream: wait() blocks until its counters (and, unless the job
    was queued with GFFX_F_NO_HOST_COPY, the host copy of its ranges) have arrived and returns the Result.  The region arrays
    are kept alive until then."""

    def __init__(self, h, keep):
        self._h, self._keep = h, keep

    def wait(self, copy: bool = True) -> "Result":
        """copy=False: `ranges` (a job queued without GFFX_F_NO_HOST_COPY) is a view of the result's page-locked buffer,
        valid until Result.close()"""
        h, self._h = self._h, None
        rc = lib.gffx_result_wait(h)
        if rc != 0:
            lib.gffx_result_free(h)
            raise GffxError(rc, _lib.last_error())
        self._keep = None
        return _take_result(h, copy)

    def close(self):
        if self._h is not None:
            h, self._h = self._h, None
            lib.gffx_result_free(h)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def _take_result(h, copy: bool = True) -> Result:
    """copy=False: `pairs` / `ranges` are views of the result's page-locked buffers, valid until Result.close() (or its
    collection); the default copies them into ordinary numpy arrays and frees the handle at once."""
    done = False
    try:
        r = Result()
        info = _lib.ResultInfo()
        _check(lib.gffx_result_get_info(h, C.byref(info)))   # one call for the whole summary
        r.n_queries, r.n_pairs, r.n_roots, r.n_ranges, r.bytes_out = info.n_queries, info.n_pairs, info.n_roots, info.n_ranges, info.bytes_out
        if info.pairs and r.n_pairs:
            r.pairs = np.ctypeslib.as_array(info.pairs, shape=(r.n_pairs, 3))
            if copy:
                r.pairs = r.pairs.copy()
        r.pairs_device_ptr = info.pairs_device or 0
        if info.ranges and r.n_ranges:
            r.ranges = np.ctypeslib.as_array(info.ranges, shape=(r.n_ranges, 2))
            if copy:
                r.ranges = r.ranges.copy()
        elif r.n_ranges == 0:
            r.ranges = np.zeros((0, 2), dtype=np.uint64)
        if info.n_warn:
            r.warn_fids = list(np.ctypeslib.as_array(info.warn_fids, shape=(info.n_warn,)))
        r.kernel_ms = info.kernel_ms
        r.stage_ms = tuple(info.stage_ms)
        r.launches = info.launches
        if not copy:
            r._h = h
            done = True
        return r
    finally:
        if not done:
            lib.gffx_result_free(h)


class TreeIndexData:
    """Device-resident index: ``TreeIndexData`` (src/utils/tree_index.rs:12-16) + ``GofMap``
    (src/index_loader/gof.rs:20-27) + the GFF text's line table."""

    def __init__(self, handle):
        self._h = handle

    # -- construction ---------------------------------------------------------------------
    @classmethod
    def load_tree_index(cls, gff_path: str, device: int = 0) -> "TreeIndexData":
        h = C.c_void_p()
        _check(lib.gffx_index_open(str(gff_path).encode(), int(device), C.byref(h)))
        return cls(h)

    @classmethod
    def create(cls, seq_names: Sequence[str], root_seq, root_start, root_end, root_fid, gof_fid, gof_start, gof_end,
               gff_text: Optional[bytes] = None, device: int = 0) -> "TreeIndexData":
        names = (C.c_char_p * len(seq_names))(*[s.encode() for s in seq_names])
        arrs32 = [np.ascontiguousarray(a, dtype=np.uint32) for a in (root_seq, root_start, root_end, root_fid, gof_fid)]
        arrs64 = [np.ascontiguousarray(a, dtype=np.uint64) for a in (gof_start, gof_end)]
        d = _lib.IndexDesc()
        d.n_seqs = len(seq_names)
        d.seq_names = names
        d.n_roots = arrs32[0].shape[0]
        d.root_seq, d.root_start, d.root_end, d.root_fid = (a.ctypes.data for a in arrs32[:4])
        d.n_gof = arrs32[4].shape[0]
        d.gof_fid = arrs32[4].ctypes.data
        d.gof_start, d.gof_end = (a.ctypes.data for a in arrs64)
        buf = None
        if gff_text:
            buf = C.create_string_buffer(gff_text, len(gff_text))
            d.gff_text = C.addressof(buf)
            d.gff_len = len(gff_text)
        d.device = int(device)
        h = C.c_void_p()
        _check(lib.gffx_index_create(C.byref(d), C.byref(h)))
        return cls(h)   # the library has copied the text (include/gffx_b200.h: inputs are borrowed for the call only)

    def close(self):
        if self._h:
            lib.gffx_index_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # -- accessors ------------------------------------------------------------------------
    @property
    def num_to_seqid(self) -> List[str]:
        return [lib.gffx_index_seq_name(self._h, i).decode() for i in range(lib.gffx_index_n_seqs(self._h))]

    @property
    def seqid_to_num(self) -> dict:
        return {n: lib.gffx_index_seq_num(self._h, n.encode()) for n in self.num_to_seqid}

    n_seqs = property(lambda self: lib.gffx_index_n_seqs(self._h))

    def vseq(self, seq: int):
        """(virtual sequence, base, clamp, number of virtual sequences) of sequence `seq` (gffx_index_vseq)"""
        v, b, c, n = C.c_uint32(), C.c_uint32(), C.c_uint32(), C.c_uint32()
        _check(lib.gffx_index_vseq(self._h, int(seq), C.byref(v), C.byref(b), C.byref(c), C.byref(n)))
        return v.value, b.value, c.value, n.value
    n_roots = property(lambda self: lib.gffx_index_n_roots(self._h))
    n_blocks = property(lambda self: lib.gffx_index_n_blocks(self._h))
    n_lines = property(lambda self: lib.gffx_index_n_lines(self._h))
    n_ids = property(lambda self: lib.gffx_index_n_ids(self._h))

    def lines_info(self):
        """(digest of the GFF line table, its source: 0 host scanner, 1 device scanner, 2 side-car cache)"""
        d, s = C.c_uint64(), C.c_int()
        _check(lib.gffx_index_lines_info(self._h, C.byref(d), C.byref(s)))
        return d.value, s.value
    device_bytes = property(lambda self: lib.gffx_index_device_bytes(self._h))
    cuda_stream = property(lambda self: lib.gffx_index_cuda_stream(self._h) or 0)

    def seq_roots(self, seq: int) -> np.ndarray:
        p = _lib.u32p()
        n = C.c_uint64()
        _check(lib.gffx_index_seq_roots(self._h, seq, C.byref(p), C.byref(n)))
        if n.value == 0:
            return np.zeros((0, 3), dtype=np.uint32)
        return np.ctypeslib.as_array(p, shape=(n.value, 3)).copy()

    # -- BED parsing ----------------------------------------------------------------------
    def _parse(self, fn, path):
        ps = [_lib.u32p() for _ in range(3)]
        n = C.c_uint64()
        _check(fn(self._h, str(path).encode(), C.byref(ps[0]), C.byref(ps[1]), C.byref(ps[2]), C.byref(n)))
        try:
            out = tuple(np.ctypeslib.as_array(p, shape=(max(n.value, 1),))[: n.value].copy() for p in ps)
        finally:
            for p in ps:
                lib.gffx_free(p)
        return out

    def parse_bed_file(self, bed_path):
        """src/commands/intersect.rs:201-230"""
        return self._parse(lib.gffx_parse_bed_intersect, bed_path)

    def parse_bed_counts(self, bed_path):
        """BED dialect of depth/coverage: src/commands/depth.rs:449-495, src/commands/coverage.rs:222-255"""
        return self._parse(lib.gffx_parse_bed_counts, bed_path)

    def parse_reads(self, path, stats: Optional[dict] = None):
        """Read intervals of a SAM / BAM file: what process_bam (src/commands/depth.rs:297-423) and collect_by_root_from_bam
        (src/commands/coverage.rs:125-205) take from rust-htslib — (sequence number, pos, reference_end) per kept record."""
        st = (C.c_uint64 * 5)()
        out = self._parse(lambda h, p, a, b, c, n: lib.gffx_parse_reads(h, p, a, b, c, n, st), path)
        if stats is not None:
            stats.update(records=int(st[0]), unmapped=int(st[1]), unknown_ref=int(st[2]), skipped=int(st[3]), container=int(st[4]))
        return out

    def parse_bed_device(self, bed_path, dialect: int = 0, device: int = 0):
        """The same dialects parsed on the device (K9, bedparse.cuh): dialect 0 = parse_bed_file, 1 = depth / coverage.  Returns three
        torch int32 CUDA tensors (views of engine-owned memory, valid until the next parse on this handle) — pass them as regions."""
        import torch

        ps = [C.c_void_p() for _ in range(3)]
        n = C.c_uint64()
        _check(lib.gffx_parse_bed_device(self._h, str(bed_path).encode(), int(dialect), C.byref(ps[0]), C.byref(ps[1]), C.byref(ps[2]), C.byref(n)))
        self.bed_parse_ms = lib.gffx_bed_parse_stat(self._h, 0)
        self.bed_host_lines = int(lib.gffx_bed_parse_stat(self._h, 2))
        self.bed_bytes = int(lib.gffx_bed_parse_stat(self._h, 3))
        dev = torch.device("cuda", int(device))   # the device this handle was opened on
        if n.value == 0:
            return tuple(torch.empty(0, dtype=torch.int32, device=dev) for _ in range(3))
        out = []
        for p in ps:
            class _Mem:   # __cuda_array_interface__ view of engine-owned device memory
                pass
            m = _Mem()
            m.__cuda_array_interface__ = {"shape": (n.value,), "typestr": "<i4", "data": (p.value, False), "version": 2}
            out.append(torch.as_tensor(m, device=dev))
        return tuple(out)

    def parse_region(self, region: str):
        """src/commands/intersect.rs:172-198"""
        c, s, e = C.c_uint32(), C.c_uint32(), C.c_uint32()
        _check(lib.gffx_parse_region(self._h, region.encode(), C.byref(c), C.byref(s), C.byref(e)))
        return c.value, s.value, e.value

    # -- compute --------------------------------------------------------------------------
    def query_features(self, regions, mode: OverlapMode = OverlapMode.Overlap, invert: bool = False, *, want_pairs: bool = True,
                       flags: int = 0, copy: bool = True) -> Result:
        """src/commands/intersect.rs:105-169"""
        keep: list = []
        pc, ps, pe, n, f = _as_region_ptrs(regions, keep)
        f |= flags | (_lib.F_WANT_PAIRS if want_pairs else 0)
        h = C.c_void_p()
        _check(lib.gffx_query_features(self._h, pc, ps, pe, n, int(mode), int(bool(invert)), f, C.byref(h)))
        return _take_result(h, copy)

    def intersect(self, regions, mode: OverlapMode = OverlapMode.Overlap, invert: bool = False, entire_group: bool = False,
                  types: Optional[str] = None, *, want_pairs: bool = False, flags: int = 0, copy: bool = True) -> Result:
        """run() after BED parsing, src/commands/intersect.rs:586-654"""
        keep: list = []
        pc, ps, pe, n, f = _as_region_ptrs(regions, keep)
        f |= flags | (_lib.F_WANT_PAIRS if want_pairs else 0)
        h = C.c_void_p()
        _check(lib.gffx_intersect(self._h, pc, ps, pe, n, int(mode), int(bool(invert)), int(bool(entire_group)),
                                  types.encode() if types is not None else None, f, C.byref(h)))
        if f & _lib.F_ASYNC:   # queued, not finished: the caller waits when it wants the numbers (gffx_result_wait)
            return PendingResult(h, keep)
        return _take_result(h, copy)

    # -- batched / sharded jobs (gffx_shard_*, gffx_job_buffer) ----------------------------------
    def job_buffers(self, counts_handle=None):
        """accumulators of the open intersect job (or of the depth / coverage job `counts_handle`): (device ptr, elements, op)"""
        out = []
        k = 0
        while True:
            p, n, op = C.c_void_p(), C.c_uint64(), C.c_int()
            rc = lib.gffx_job_buffer(self._h, counts_handle, k, C.byref(p), C.byref(n), C.byref(op))
            if rc == 1:
                return out
            _check(rc)
            out.append((p.value or 0, n.value, op.value))
            k += 1

    def shard_intersect(self, regions, mode: OverlapMode = OverlapMode.Overlap, invert: bool = False, entire_group: bool = False,
                        types: Optional[str] = None, *, combine=None, batches: int = 1, want_pairs: bool = False, flags: int = 0,
                        copy: bool = True) -> Result:
        """The same result as intersect() over the union of every shard's regions: begin, add this shard's regions (in
        `batches` pieces), `combine(buffers)` exchanges the accumulators (gffx_b200.shard.Combiner; None = single shard),
        finish.  Result.n_pairs / pairs are THIS shard's."""
        keep: list = []
        pc, ps, pe, n, f = _as_region_ptrs(regions, keep)
        flags |= _lib.F_WANT_PAIRS if want_pairs else 0
        announce = n if flags & (_lib.F_WANT_PAIRS | _lib.F_PAIRS_ON_DEVICE) else 0
        _check(lib.gffx_shard_begin(self._h, int(mode), int(bool(invert)), int(bool(entire_group)),
                                    types.encode() if types is not None else None, flags, announce, max(1, batches) + 1))
        step = max(1, (n + max(1, batches) - 1) // max(1, batches))
        for b0 in range(0, n, step):
            m = min(step, n - b0)
            _check(lib.gffx_shard_add(self._h, pc + 4 * b0, ps + 4 * b0, pe + 4 * b0, m, f, None))
        if combine is not None:
            combine(self.job_buffers())
        h = C.c_void_p()
        _check(lib.gffx_shard_finish(self._h, C.byref(h)))
        if flags & _lib.F_ASYNC:
            return PendingResult(h, keep)
        return _take_result(h, copy)

    def sync(self):
        _check(lib.gffx_index_sync(self._h))

    def count_pairs_noalloc(self, d_chr: int, d_start: int, d_end: int, n: int, mode: int, invert: bool, d_out: int, stream: int = 0):
        """gffx_count_pairs_noalloc: device pointers in (n sequence numbers, starts, ends), two uint64 out on the device
        ([0] pairs, [1] bad sequence number seen); nothing is allocated, the call is asynchronous on `stream` (0: the
        index's own, wait with sync()).  No counterpart in the reference (include/gffx_b200.h)."""
        _check(lib.gffx_count_pairs_noalloc(self._h, d_chr, d_start, d_end, n, mode, 1 if invert else 0, d_out, stream or None))

    def write_ranges(self, ranges: np.ndarray, out_path: Optional[str]):
        r = np.ascontiguousarray(ranges, dtype=np.uint64)
        _check(lib.gffx_write_ranges(self._h, r.ctypes.data, r.shape[0], str(out_path).encode() if out_path else None))

    def gather_text(self, ranges: Optional[np.ndarray] = None, n_ranges: Optional[int] = None, device: bool = False):
        """Record emission on the device (K8, gather.cuh): the bytes write_ranges would write, gathered from the HBM-resident
        GFF into one buffer.  ranges=None: the ranges of the last intersect on this handle (pass its n_ranges).
        Returns bytes (device=False) or (device_pointer, n_bytes); `gather_ms` / `gather_scan_ms` hold the device times."""
        text, nb = C.c_void_p(), C.c_uint64()
        if ranges is None:
            _check(lib.gffx_gather_text(self._h, None, int(n_ranges or 0), _lib.F_NO_HOST_COPY if device else 0, C.byref(text), C.byref(nb)))
        else:
            r = np.ascontiguousarray(ranges, dtype=np.uint64)
            _check(lib.gffx_gather_text(self._h, r.ctypes.data, r.shape[0], _lib.F_NO_HOST_COPY if device else 0, C.byref(text), C.byref(nb)))
        self.gather_ms = lib.gffx_gather_kernel_ms(self._h, 0)
        self.gather_scan_ms = lib.gffx_gather_kernel_ms(self._h, 1)
        if device:
            return text.value, nb.value
        return C.string_at(text.value, nb.value) if nb.value else b""

    def _counts(self, kind: int, regions, batches: int = 1, combine=None, copy: bool = True):
        st = C.c_void_p()
        _check(lib.gffx_counts_begin(self._h, kind, C.byref(st)))
        done = False
        try:
            keep: list = []
            pc, ps, pe, n, f = _as_region_ptrs(regions, keep)
            if batches <= 1 or n == 0:
                _check(lib.gffx_counts_add(st, pc, ps, pe, n, f))
            else:
                step = (n + batches - 1) // batches
                for b0 in range(0, n, step):
                    m = min(step, n - b0)
                    _check(lib.gffx_counts_add(st, pc + 4 * b0, ps + 4 * b0, pe + 4 * b0, m, f))
            if combine is not None:  # multi-GPU: exchange the accumulators, then every rank finishes with the global state
                combine(self.job_buffers(st))
            nr = C.c_uint64()
            pid, pch, pst, pen = (_lib.u32p() for _ in range(4))
            pv = _lib.u64p()
            _check(lib.gffx_counts_finish(st, C.byref(nr), C.byref(pid), C.byref(pch), C.byref(pst), C.byref(pen), C.byref(pv)))
            k = nr.value

            def arr(p, dt):
                if not k:
                    return np.zeros(0, dtype=dt)
                a = np.ctypeslib.as_array(p, shape=(k,))
                return a.copy() if copy else a

            rows = CountRows(arr(pid, np.uint32), arr(pch, np.uint32), arr(pst, np.uint32), arr(pen, np.uint32), arr(pv, np.uint64),
                             lib.gffx_counts_kernel_ms(st), lib.gffx_counts_launches(st), kind, self)
            if not copy:   # the arrays are views of the job's page-locked row buffers: the job lives until rows.close()
                rows._st = st
                done = True
            return rows
        finally:
            if not done:
                lib.gffx_counts_free(st)

    def depth(self, regions, batches: int = 1, combine=None, copy: bool = True) -> "CountRows":
        """compute_hit_depth over all batches, src/commands/depth.rs:222-293,429-512.  copy=False: the row arrays are views
        of library-owned page-locked memory, valid until CountRows.close()."""
        return self._counts(0, regions, batches, combine, copy)

    def coverage(self, regions, batches: int = 1, combine=None, copy: bool = True) -> "CountRows":
        """collect_by_root + finalize_compute_breadth, src/commands/coverage.rs:208-450"""
        return self._counts(1, regions, batches, combine, copy)

    def id_string(self, i: int) -> bytes:
        n = C.c_uint32()
        p = lib.gffx_index_id_string(self._h, int(i), C.byref(n))
        return C.string_at(p, n.value) if p else b""

    def chrom_string(self, i: int) -> bytes:
        n = C.c_uint32()
        p = lib.gffx_index_chrom_string(self._h, int(i), C.byref(n))
        return C.string_at(p, n.value) if p else b""

    # -- synthetic workloads --------------------------------------------------------------
    def synth_regions(self, seed: int, n: int, length: int, first: int = 0, out=None):
        """regions [first, first+n) of the counter-based stream `seed`; `out` = three 4-byte host tensors
        (e.g. pinned torch tensors) to fill in place instead of fresh numpy arrays"""
        if out is not None:
            if any(int(t.numel()) != n or t.element_size() != 4 or t.is_cuda for t in out):
                raise ValueError("out must be three host tensors of n 4-byte elements")
            _check(lib.gffx_synth_regions_host(self._h, seed, first, n, length, *(t.data_ptr() for t in out)))
            return tuple(out)
        out = [np.empty(n, dtype=np.uint32) for _ in range(3)]
        _check(lib.gffx_synth_regions_host(self._h, seed, first, n, length, *(a.ctypes.data for a in out)))
        return tuple(out)

    def synth_regions_device(self, seed: int, n: int, length: int, first: int = 0, device=None):
        import torch

        dev = device if device is not None else torch.device("cuda", torch.cuda.current_device())
        out = [torch.empty(n, dtype=torch.int32, device=dev) for _ in range(3)]
        _check(lib.gffx_synth_regions_device(self._h, seed, first, n, length, *(t.data_ptr() for t in out)))
        return tuple(out)


@dataclass
class CountRows:
    id_idx: np.ndarray
    chrom_idx: np.ndarray
    start: np.ndarray
    end: np.ndarray
    value: np.ndarray
    kernel_ms: float
    launches: int
    kind: int
    index: "TreeIndexData"
    _st: object = None   # the C job handle while the arrays are zero-copy views (copy=False)

    def close(self):
        """release the library-owned row buffers behind zero-copy arrays (they must not be used afterwards)"""
        if self._st is not None:
            st, self._st = self._st, None
            self.id_idx = self.chrom_idx = self.start = self.end = self.value = None
            lib.gffx_counts_free(st)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def to_tsv(self, sort_rows: bool = True) -> bytes:
        """The reference's TSV (write_depth_results / write_breadth_results); rows sorted bytewise when asked,
        because the reference's own row order is hash order."""
        rows = []
        for i in range(self.id_idx.shape[0]):
            ident = self.index.id_string(self.id_idx[i])
            chrom = self.index.chrom_string(self.chrom_idx[i])
            s, e, v = int(self.start[i]), int(self.end[i]), int(self.value[i])
            if self.kind == 0:
                rows.append(ident + b"\t" + chrom + b"\t%d\t%d\t%d\n" % (s, e, v))
            else:
                length = e - s if e > s else 0
                frac = (v / length) if length > 0 else 0.0
                rows.append(ident + b"\t" + chrom + b"\t%d\t%d\t%d\t%s\n" % (s, e, v, ("%.6f" % frac).encode()))
        if sort_rows:
            rows.sort()
        head = b"id\tchr\tstart\tend\tdepth\n" if self.kind == 0 else b"id\tchr\tstart\tend\tbreadth\tfraction\n"
        return head + b"".join(rows)


def build_index(gff_path: str, attribute: str = "gene_name", skip_types: Optional[str] = None):
    """`gffx index` (src/commands/index.rs:25-37): writes the eight index files beside the GFF."""
    _check(lib.gffx_index_build(str(gff_path).encode(), attribute.encode(), skip_types.encode() if skip_types is not None else None))


def synth_gff(path: str, seed: int = 20260001, n_genes: int = 60000, quirks: bool = False, n_scaffolds: int = 0) -> int:
    """human-like synthetic GFF3 (24 chromosomes); n_scaffolds > 0 appends that many small unplaced scaffolds"""
    n = C.c_uint64()
    if n_scaffolds:
        _check(lib.gffx_synth_gff_scaffolds(str(path).encode(), seed, n_genes, int(quirks), int(n_scaffolds), C.byref(n)))
    else:
        _check(lib.gffx_synth_gff(str(path).encode(), seed, n_genes, int(quirks), C.byref(n)))
    return n.value


# ---- command level: the reference's Args structs and run() functions ------------------------------
@dataclass
class IntersectArgs:
    """src/commands/intersect.rs:43-70 + CommonArgs src/utils/common.rs:17-52"""

    input: str
    output: Optional[str] = None
    bed: Optional[str] = None
    region: Optional[str] = None
    contained: bool = False
    contains_region: bool = False
    overlap: bool = False
    invert: bool = False
    entire_group: bool = False
    types: Optional[str] = None
    threads: int = 12
    verbose: bool = False
    device: int = 0


def run_intersect(args: IntersectArgs):
    """src/commands/intersect.rs:541-655"""
    if sum([args.contained, args.contains_region, args.overlap]) > 1:
        raise GffxError(-1, "--contained, --contains-region and --overlap cannot be used together")
    mode = OverlapMode.Contained if args.contained else OverlapMode.ContainsRegion if args.contains_region else OverlapMode.Overlap
    a = _lib.IntersectArgs(
        str(args.input).encode(), str(args.output).encode() if args.output else None, str(args.bed).encode() if args.bed else None,
        args.region.encode() if args.region else None, int(mode), int(args.invert), int(args.entire_group),
        args.types.encode() if args.types is not None else None, int(args.verbose), int(args.device))
    _check(lib.gffx_intersect_run(C.byref(a)))


@dataclass
class DepthArgs:
    """src/commands/depth.rs:40-74"""

    input: str
    source: str
    output: Optional[str] = None
    bin_shift: int = 12
    threads: int = 12
    verbose: bool = False
    device: int = 0


CoverageArgs = DepthArgs  # src/commands/coverage.rs:37-57 has the same fields minus bin_shift


def _counts_args(args):
    return _lib.CountsArgs(str(args.input).encode(), str(args.source).encode(), str(args.output).encode() if args.output else None,
                           int(args.bin_shift), int(args.threads), int(args.verbose), int(args.device))


def run_depth(args: DepthArgs):
    """src/commands/depth.rs:548-635"""
    a = _counts_args(args)
    _check(lib.gffx_depth_run(C.byref(a)))


def run_coverage(args: DepthArgs):
    """src/commands/coverage.rs:493-582"""
    a = _counts_args(args)
    _check(lib.gffx_coverage_run(C.byref(a)))
