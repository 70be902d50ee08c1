"""Host-side checks of the virtual-sequence layout (VSeqMap, gffx_b200/csrc/host_index.h) that lifts the engine past a few
hundred sequences: TreeIndexData has one tree per sequence and no limit (src/utils/tree_index.rs:36-82), the reference's own
benchmark/hs.gff.sqs lists 583.  No GPU needed: the layout is built by the host loaders (device = -1)."""
import os

import numpy as np
import pytest

from tests._hs import HS, write_hs_like


def layout(idx):
    v = np.array([idx.vseq(c) for c in range(idx.n_seqs)], dtype=np.uint64)
    return v[:, 0], v[:, 1], v[:, 2], int(v[0, 3])


def check_disjoint(vseq, base, clamp):
    for v in np.unique(vseq):
        m = np.flatnonzero(vseq == v)
        if len(m) == 1:
            assert base[m[0]] == 0 and clamp[m[0]] == 0xFFFFFFFF      # alone: the sequence keeps its own coordinates
            continue
        o = m[np.argsort(base[m])]
        lo, hi = base[o], base[o] + clamp[o]
        assert np.all(hi < 2**32)
        assert np.all(lo[1:] > hi[:-1])                                # [base, base + clamp] never touch


def test_reference_hs_header_files_open(gffx, tmp_path):
    """the reference's own 583-sequence .sqs / .rix (tests/golden/hs, byte copies) go through the loaders"""
    assert sum(1 for _ in open(os.path.join(HS, "hs.gff.sqs"))) == 583
    base, trees = write_hs_like(str(tmp_path))
    idx = gffx.TreeIndexData.load_tree_index(base, device=-1)
    assert idx.n_seqs == 583 and idx.n_roots == sum(len(t) for t in trees)
    assert idx.num_to_seqid[:2] == [l.rstrip("\n") for l in open(os.path.join(HS, "hs.gff.sqs"))][:2]
    vseq, b, c, n_v = layout(idx)
    assert n_v <= 192 and n_v == len(np.unique(vseq))
    check_disjoint(vseq, b, c)
    for s, t in enumerate(trees):                                      # the clamp sits above every root coordinate
        if len(t) and c[s] != 0xFFFFFFFF:
            assert int(t[:, 1].max()) < c[s]
    idx.close()


@pytest.mark.parametrize("cap", [None, "65536", "4000000"])
def test_scaffold_assembly_layout(gffx, oracle, tmp_path, monkeypatch, cap):
    if cap:
        monkeypatch.setenv("GFFX_VSEQ_CAP", cap)
    path = str(tmp_path / "scaf.gff")
    gffx.synth_gff(path, seed=5, n_genes=600, quirks=True, n_scaffolds=1500)
    oracle.build_index(path)
    idx = gffx.TreeIndexData.load_tree_index(path, device=-1)
    assert idx.n_seqs > 1500
    vseq, b, c, n_v = layout(idx)
    assert n_v <= 192
    check_disjoint(vseq, b, c)
    names = idx.num_to_seqid
    top = [s for s, n in enumerate(names) if n.startswith("scaf") and (int(n[4:]) % 401) == 5]
    assert top and all(c[s] == 0xFFFFFFFF for s in top)               # a coordinate near 2^32: that sequence stays alone
    idx.close()
