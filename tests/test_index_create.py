"""gffx_index_create: the entry a Rust caller uses when it already holds TreeIndexData / GofMap / &Mmap
(src/utils/tree_index.rs:12-16, src/index_loader/gof.rs:20-27) — SURVEY.md 8(b)'s first bullet.  Built from the arrays the
oracle decodes out of the index files, it must behave exactly like gffx_index_open on the same files, and it must COPY the
GFF text (include/gffx_b200.h: inputs are borrowed for the duration of the call only)."""
import ctypes as C
import os

import numpy as np
import pytest

from tests._oracle import Oracle
from tests.test_intersect_gpu import MODES, mixed_regions


def arrays_of(path):
    """what a caller holding the reference's structs has: names, per-tree intervals, .gof records in file order"""
    with Oracle(path) as o:
        names = o.seq_names
        seq, s, e, fid = [], [], [], []
        for c in range(len(names)):
            t = o.seq_intervals(c)   # (start, end, root_fid) per interval of tree c
            seq.append(np.full(len(t), c, dtype=np.uint32))
            s.append(t[:, 0]); e.append(t[:, 1]); fid.append(t[:, 2])
    gof = np.frombuffer(open(path + ".gof", "rb").read(), dtype=np.dtype([("fid", "<u4"), ("seq", "<u4"), ("s", "<u8"), ("e", "<u8")]))
    return names, np.concatenate(seq), np.concatenate(s), np.concatenate(e), np.concatenate(fid), gof["fid"].copy(), gof["s"].copy(), gof["e"].copy()


def create_from(gffx, path, device):
    names, seq, s, e, fid, gfid, gs, ge = arrays_of(path)
    text = bytearray(open(path, "rb").read())
    idx = gffx.TreeIndexData.create(names, seq, s, e, fid, gfid, gs, ge, gff_text=bytes(text), device=device)
    for k in range(0, len(text), 4096):   # the caller's buffer is gone / reused after the call
        text[k] = 0x23
    del text
    return idx


def test_create_host_only_equals_open(gffx, synth_quirk_gff):
    a = create_from(gffx, synth_quirk_gff, -1)
    b = gffx.TreeIndexData.load_tree_index(synth_quirk_gff, device=-1)
    assert (a.n_seqs, a.n_roots, a.n_lines, a.n_blocks) == (b.n_seqs, b.n_roots, b.n_lines, b.n_blocks)
    assert a.num_to_seqid == b.num_to_seqid
    assert [a.vseq(c) for c in range(a.n_seqs)] == [b.vseq(c) for c in range(b.n_seqs)]
    a.close(); b.close()


def test_create_rejects_bad_descriptors(gffx):
    from gffx_b200 import _lib

    h = C.c_void_p()
    assert _lib.lib.gffx_index_create(None, C.byref(h)) == -1
    d = _lib.IndexDesc()
    d.n_roots = 3                       # arrays missing
    d.device = -1
    assert _lib.lib.gffx_index_create(C.byref(d), C.byref(h)) == -1 and "root arrays" in _lib.last_error()
    with pytest.raises(gffx.GffxError):   # a tree sequence number beyond the name list
        gffx.TreeIndexData.create(["chr1"], [1], [0], [10], [0], [0], [0], [5], gff_text=b"x\ty\n", device=-1)


@pytest.mark.gpu
def test_create_gives_the_results_of_open(gffx, synth_quirk_gff):
    a = create_from(gffx, synth_quirk_gff, 0)
    b = gffx.TreeIndexData.load_tree_index(synth_quirk_gff, device=0)
    try:
        regs = mixed_regions(b, 4711, 3000)
        for mode, invert in MODES:
            for entire, types in ((False, None), (True, None), (False, "gene,exon")):
                ra = a.intersect(regs, mode, invert, entire, types, want_pairs=True)
                rb = b.intersect(regs, mode, invert, entire, types, want_pairs=True)
                assert ra.n_pairs == rb.n_pairs and np.array_equal(ra.ranges, rb.ranges)
                if ra.n_pairs:
                    assert np.array_equal(ra.pairs, rb.pairs)
        r = a.intersect(regs, 2, False, False, None)
        text = open(synth_quirk_gff, "rb").read()
        got = a.gather_text(r.ranges, r.n_ranges)        # record emission reads the library's own copy of the text
        assert got == b"".join(text[s:e] for s, e in r.ranges.tolist())
        reads = b.synth_regions(5, 20000, 150)
        assert a.depth(reads).to_tsv() == b.depth(reads).to_tsv()      # ID strings come out of the copied text
        assert a.coverage(reads).to_tsv() == b.coverage(reads).to_tsv()
    finally:
        a.close(); b.close()
