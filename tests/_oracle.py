"""ctypes wrapper of the CPU oracle (oracle/_build/liboracle.so).  TEST INFRASTRUCTURE ONLY."""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.path.join(ROOT, "oracle", "_build", "liboracle.so")
_lib = C.CDLL(LIB_PATH)
_lib.orc_last_error.restype = C.c_char_p
_lib.orc_seq_name.restype = C.c_char_p
_lib.orc_n_roots.restype = C.c_uint64
_lib.orc_n_seqs.restype = C.c_uint32
_lib.orc_n_roots.argtypes = [C.c_void_p]
_lib.orc_n_seqs.argtypes = [C.c_void_p]
_lib.orc_seq_name.argtypes = [C.c_void_p, C.c_uint32]
_lib.orc_close.argtypes = [C.c_void_p]
_lib.orc_free.argtypes = [C.c_void_p]
_u32p = C.POINTER(C.c_uint32)
_u64p = C.POINTER(C.c_uint64)
_lib.orc_open.argtypes = [C.c_char_p, C.POINTER(C.c_void_p)]
_lib.orc_seq_intervals.argtypes = [C.c_void_p, C.c_uint32, C.POINTER(_u32p), _u64p]
_lib.orc_parse_bed_intersect.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(_u32p), _u64p]
_lib.orc_parse_bed_dc.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(_u32p), _u64p]
_lib.orc_parse_reads.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(_u32p), _u64p]
_lib.orc_query_features.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.c_int, C.POINTER(_u32p), _u64p]
_lib.orc_intersect.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.c_int, C.c_int, C.c_char_p,
                               C.POINTER(_u64p), _u64p, _u64p, _u64p]
_lib.orc_reencode_rit.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), _u64p]
_lib.orc_depth.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint32, C.POINTER(C.c_void_p), _u64p]
_lib.orc_coverage.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(C.c_void_p), _u64p]


class OracleError(RuntimeError):
    pass


def _check(rc):
    if rc != 0:
        raise OracleError((_lib.orc_last_error() or b"").decode("utf-8", "replace"))


def build_index(gff_path: str):
    """`gffx index` restated (src/index_builder/core.rs:41-242): writes the index files beside the GFF."""
    _check(_lib.orc_build_index(str(gff_path).encode()))


def set_threads(n: int):
    _lib.orc_set_threads(int(n))


def max_threads() -> int:
    return int(_lib.orc_max_threads())


def _regs(regions):
    arrs = [np.ascontiguousarray(a, dtype=np.uint32) for a in regions]
    return arrs, arrs[0].shape[0]


def _take_triples(p, n):
    try:
        if n == 0:
            return np.zeros((0, 3), dtype=np.uint32)
        return np.ctypeslib.as_array(p, shape=(n, 3)).copy()
    finally:
        _lib.orc_free(p)


class Oracle:
    def __init__(self, gff_path: str):
        self._h = C.c_void_p()
        _check(_lib.orc_open(str(gff_path).encode(), C.byref(self._h)))

    def close(self):
        if self._h:
            _lib.orc_close(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    @property
    def seq_names(self):
        return [_lib.orc_seq_name(self._h, i).decode() for i in range(_lib.orc_n_seqs(self._h))]

    @property
    def n_roots(self):
        return _lib.orc_n_roots(self._h)

    def seq_intervals(self, seq):
        p, n = _u32p(), C.c_uint64()
        _check(_lib.orc_seq_intervals(self._h, seq, C.byref(p), C.byref(n)))
        return _take_triples(p, n.value)

    def reencode_rit(self) -> bytes:
        """the index's trees rebuilt by the restated IntervalTree::new and re-serialised (must equal the .rit file)"""
        p, n = C.c_void_p(), C.c_uint64()
        _check(_lib.orc_reencode_rit(self._h, C.byref(p), C.byref(n)))
        try:
            return C.string_at(p, n.value)
        finally:
            _lib.orc_free(p)

    def parse_bed_intersect(self, bed):
        p, n = _u32p(), C.c_uint64()
        _check(_lib.orc_parse_bed_intersect(self._h, str(bed).encode(), C.byref(p), C.byref(n)))
        t = _take_triples(p, n.value)
        return t[:, 0].copy(), t[:, 1].copy(), t[:, 2].copy()

    def parse_bed_counts(self, bed):
        p, n = _u32p(), C.c_uint64()
        _check(_lib.orc_parse_bed_dc(self._h, str(bed).encode(), C.byref(p), C.byref(n)))
        t = _take_triples(p, n.value)
        return t[:, 0].copy(), t[:, 1].copy(), t[:, 2].copy()

    def parse_reads(self, path):
        p, n = _u32p(), C.c_uint64()
        _check(_lib.orc_parse_reads(self._h, str(path).encode(), C.byref(p), C.byref(n)))
        t = _take_triples(p, n.value)
        return t[:, 0].copy(), t[:, 1].copy(), t[:, 2].copy()

    def query_features(self, regions, mode, invert=False, want_pairs=True):
        a, n = _regs(regions)
        p, k = _u32p(), C.c_uint64()
        _check(_lib.orc_query_features(self._h, a[0].ctypes.data, a[1].ctypes.data, a[2].ctypes.data, n, int(mode), int(bool(invert)),
                                       C.byref(p) if want_pairs else None, C.byref(k)))
        if not want_pairs:
            return k.value
        return _take_triples(p, k.value)

    def intersect(self, regions, mode, invert=False, entire_group=False, types=None):
        """-> (ranges (n,2) uint64 in output order, n_pairs, n_warn)"""
        a, n = _regs(regions)
        p = _u64p()
        nr, npairs, nwarn = C.c_uint64(), C.c_uint64(), C.c_uint64()
        _check(_lib.orc_intersect(self._h, a[0].ctypes.data, a[1].ctypes.data, a[2].ctypes.data, n, int(mode), int(bool(invert)),
                                  int(bool(entire_group)), types.encode() if types is not None else None, C.byref(p), C.byref(nr),
                                  C.byref(npairs), C.byref(nwarn)))
        try:
            r = np.ctypeslib.as_array(p, shape=(nr.value, 2)).copy() if nr.value else np.zeros((0, 2), dtype=np.uint64)
        finally:
            _lib.orc_free(p)
        return r, npairs.value, nwarn.value

    def _tsv(self, fn, regions, *extra):
        a, n = _regs(regions)
        p, ln = C.c_void_p(), C.c_uint64()
        _check(fn(self._h, a[0].ctypes.data, a[1].ctypes.data, a[2].ctypes.data, n, *extra, C.byref(p), C.byref(ln)))
        try:
            return C.string_at(p, ln.value)
        finally:
            _lib.orc_free(p)

    def depth(self, regions, bin_shift=12) -> bytes:
        """header + bytewise-sorted rows"""
        return self._tsv(_lib.orc_depth, regions, C.c_uint32(bin_shift))

    def coverage(self, regions) -> bytes:
        return self._tsv(_lib.orc_coverage, regions)


def sorted_pairs(p: np.ndarray) -> np.ndarray:
    """canonical order of a pair multiset (the reference's pair order is unspecified)"""
    if p.shape[0] == 0:
        return p.reshape(0, 3)
    idx = np.lexsort((p[:, 2], p[:, 1], p[:, 0]))
    return p[idx]
