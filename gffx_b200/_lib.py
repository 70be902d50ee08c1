"""ctypes binding of the C ABI declared in include/gffx_b200.h.

The shared library is built in-tree (``make`` or ``__graft_entry__.build()``) as
``gffx_b200/lib/libgffx_b200.so``.  There is no Python or CPU fallback: importing this module
without the built CUDA library raises, and every compute call fails when no B200-class GPU
is usable.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("GFFX_B200_LIB") or os.path.join(_HERE, "lib", "libgffx_b200.so")  # the override is for A/B builds of the same ABI

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} is missing: build the CUDA extension first (run `make` at the repo root or "
        "`python -c 'import __graft_entry__ as g; g.build()'`). gffx_b200 has no CPU fallback."
    )

lib = C.CDLL(LIB_PATH)

u32p = C.POINTER(C.c_uint32)
u64p = C.POINTER(C.c_uint64)
vp = C.c_void_p


class IndexDesc(C.Structure):
    _fields_ = [
        ("n_seqs", C.c_uint32),
        ("seq_names", C.POINTER(C.c_char_p)),
        ("n_roots", C.c_uint64),
        ("root_seq", vp),
        ("root_start", vp),
        ("root_end", vp),
        ("root_fid", vp),
        ("n_gof", C.c_uint64),
        ("gof_fid", vp),
        ("gof_start", vp),
        ("gof_end", vp),
        ("gff_text", vp),
        ("gff_len", C.c_uint64),
        ("device", C.c_int),
    ]


class ResultInfo(C.Structure):   # gffx_result_info
    _fields_ = [(n, C.c_uint64) for n in ("n_queries", "n_pairs", "n_roots", "n_ranges", "bytes_out", "n_warn")] + [
        ("pairs", u32p), ("pairs_device", vp), ("ranges", u64p), ("warn_fids", u32p),
        ("kernel_ms", C.c_double), ("stage_ms", C.c_double * 12), ("launches", C.c_uint32), ("reserved", C.c_uint32)]


class IntersectArgs(C.Structure):
    _fields_ = [
        ("input", C.c_char_p),
        ("output", C.c_char_p),
        ("bed", C.c_char_p),
        ("region", C.c_char_p),
        ("mode", C.c_int),
        ("invert", C.c_int),
        ("entire_group", C.c_int),
        ("types", C.c_char_p),
        ("verbose", C.c_int),
        ("device", C.c_int),
    ]


class CountsArgs(C.Structure):
    _fields_ = [
        ("input", C.c_char_p),
        ("source", C.c_char_p),
        ("output", C.c_char_p),
        ("bin_shift", C.c_uint32),
        ("threads", C.c_int),
        ("verbose", C.c_int),
        ("device", C.c_int),
    ]


def _sig(name, restype, *argtypes):
    f = getattr(lib, name)
    f.restype = restype
    f.argtypes = list(argtypes)
    return f


# every symbol of include/gffx_b200.h (tests/test_abi.py checks the header against this table)
SIGNATURES = {
    "gffx_last_error": (C.c_char_p,),
    "gffx_version": (C.c_char_p,),
    "gffx_device_count": (C.c_int,),
    "gffx_index_open": (C.c_int, C.c_char_p, C.c_int, C.POINTER(vp)),
    "gffx_index_create": (C.c_int, C.POINTER(IndexDesc), C.POINTER(vp)),
    "gffx_index_destroy": (None, vp),
    "gffx_index_cuda_stream": (vp, vp),
    "gffx_index_build": (C.c_int, C.c_char_p, C.c_char_p, C.c_char_p),
    "gffx_index_n_seqs": (C.c_uint32, vp),
    "gffx_index_seq_name": (C.c_char_p, vp, C.c_uint32),
    "gffx_index_seq_num": (C.c_int64, vp, C.c_char_p),
    "gffx_index_n_roots": (C.c_uint64, vp),
    "gffx_index_n_blocks": (C.c_uint64, vp),
    "gffx_index_n_lines": (C.c_uint64, vp),
    "gffx_index_n_ids": (C.c_uint64, vp),
    "gffx_index_lines_info": (C.c_int, vp, u64p, C.POINTER(C.c_int)),
    "gffx_index_device_bytes": (C.c_uint64, vp),
    "gffx_index_seq_roots": (C.c_int, vp, C.c_uint32, C.POINTER(u32p), u64p),
    "gffx_parse_bed_intersect": (C.c_int, vp, C.c_char_p, C.POINTER(u32p), C.POINTER(u32p), C.POINTER(u32p), u64p),
    "gffx_parse_bed_counts": (C.c_int, vp, C.c_char_p, C.POINTER(u32p), C.POINTER(u32p), C.POINTER(u32p), u64p),
    "gffx_parse_reads": (C.c_int, vp, C.c_char_p, C.POINTER(u32p), C.POINTER(u32p), C.POINTER(u32p), u64p, u64p),
    "gffx_parse_bed_device": (C.c_int, vp, C.c_char_p, C.c_int, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), u64p),
    "gffx_bed_parse_stat": (C.c_double, vp, C.c_int),
    "gffx_parse_region": (C.c_int, vp, C.c_char_p, u32p, u32p, u32p),
    "gffx_free": (None, vp),
    "gffx_query_features": (C.c_int, vp, vp, vp, vp, C.c_uint64, C.c_int, C.c_int, C.c_uint32, C.POINTER(vp)),
    "gffx_intersect": (C.c_int, vp, vp, vp, vp, C.c_uint64, C.c_int, C.c_int, C.c_int, C.c_char_p, C.c_uint32, C.POINTER(vp)),
    "gffx_result_wait": (C.c_int, vp),
    "gffx_result_pending": (C.c_int, vp),
    "gffx_result_n_queries": (C.c_uint64, vp),
    "gffx_result_n_pairs": (C.c_uint64, vp),
    "gffx_result_pairs": (u32p, vp),
    "gffx_result_pairs_device": (vp, vp),
    "gffx_result_n_roots": (C.c_uint64, vp),
    "gffx_result_n_ranges": (C.c_uint64, vp),
    "gffx_result_ranges": (u64p, vp),
    "gffx_result_bytes_out": (C.c_uint64, vp),
    "gffx_result_n_warn": (C.c_uint64, vp),
    "gffx_result_warn_fids": (u32p, vp),
    "gffx_result_kernel_ms": (C.c_double, vp),
    "gffx_result_launches": (C.c_uint32, vp),
    "gffx_result_stage_ms": (C.c_double, vp, C.c_int),
    "gffx_result_get_info": (C.c_int, vp, vp),
    "gffx_result_free": (None, vp),
    "gffx_write_ranges": (C.c_int, vp, vp, C.c_uint64, C.c_char_p),
    "gffx_gather_text": (C.c_int, vp, vp, C.c_uint64, C.c_uint32, C.POINTER(vp), C.POINTER(C.c_uint64)),
    "gffx_gather_kernel_ms": (C.c_double, vp, C.c_int),
    "gffx_write_text": (C.c_int, vp, C.c_uint64, C.c_char_p),
    "gffx_counts_begin": (C.c_int, vp, C.c_int, C.POINTER(vp)),
    "gffx_counts_add": (C.c_int, vp, vp, vp, vp, C.c_uint64, C.c_uint32),
    "gffx_counts_finish": (C.c_int, vp, u64p, C.POINTER(u32p), C.POINTER(u32p), C.POINTER(u32p), C.POINTER(u32p), C.POINTER(u64p)),
    "gffx_counts_kernel_ms": (C.c_double, vp),
    "gffx_counts_launches": (C.c_uint32, vp),
    "gffx_counts_free": (None, vp),
    "gffx_index_id_string": (vp, vp, C.c_uint32, u32p),
    "gffx_index_chrom_string": (vp, vp, C.c_uint32, u32p),
    "gffx_counts_write": (C.c_int, vp, C.c_char_p),
    "gffx_shard_begin": (C.c_int, vp, C.c_int, C.c_int, C.c_int, C.c_char_p, C.c_uint32, C.c_uint64, C.c_uint32),
    "gffx_shard_add": (C.c_int, vp, vp, vp, vp, C.c_uint64, C.c_uint32, u64p),
    "gffx_shard_finish": (C.c_int, vp, C.POINTER(vp)),
    "gffx_job_buffer": (C.c_int, vp, vp, C.c_uint32, C.POINTER(vp), u64p, C.POINTER(C.c_int)),
    "gffx_fold": (C.c_int, vp, vp, vp, C.c_uint32, C.c_uint64, C.c_uint64, C.c_int),
    "gffx_index_sync": (C.c_int, vp),
    "gffx_inflate_raw": (C.c_int64, vp, C.c_uint64, vp, C.c_uint64),
    "gffx_crc32": (C.c_uint32, vp, C.c_uint64),
    "gffx_count_pairs_noalloc": (C.c_int, vp, vp, vp, vp, C.c_uint64, C.c_int, C.c_int, vp, vp),
    "gffx_noalloc_max_regions": (C.c_uint64,),
    "gffx_noalloc_max_roots": (C.c_uint64,),
    "gffx_intersect_run": (C.c_int, C.POINTER(IntersectArgs)),
    "gffx_depth_run": (C.c_int, C.POINTER(CountsArgs)),
    "gffx_coverage_run": (C.c_int, C.POINTER(CountsArgs)),
    "gffx_synth_gff": (C.c_int, C.c_char_p, C.c_uint64, C.c_uint32, C.c_int, u64p),
    "gffx_synth_gff_scaffolds": (C.c_int, C.c_char_p, C.c_uint64, C.c_uint32, C.c_int, C.c_uint32, u64p),
    "gffx_index_vseq": (C.c_int, vp, C.c_uint32, u32p, u32p, u32p, u32p),
    "gffx_synth_regions_host": (C.c_int, vp, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint32, vp, vp, vp),
    "gffx_synth_regions_device": (C.c_int, vp, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint32, vp, vp, vp),
}

for _name, (_res, *_args) in SIGNATURES.items():
    _sig(_name, _res, *_args)

F_DEVICE_INPUT = 1
F_WANT_PAIRS = 2
F_NO_HOST_COPY = 4
F_PAIRS_ON_DEVICE = 8
F_STAGE_TIMES = 16
F_ASYNC = 32


def last_error() -> str:
    return (lib.gffx_last_error() or b"").decode("utf-8", "replace")
