"""More sequences than partition windows: a scaffold-level assembly (24 chromosomes + 1200 scaffolds, quirks on) and an
index with the 583 sequences of the reference's benchmark/hs.gff.sqs.  The reference loads one tree per sequence with
no limit (src/utils/tree_index.rs:36-82); the engine lays them on virtual sequences (host_index.h: VSeqMap) and every
result must stay bit-exact: pairs, byte ranges for all mode x invert x {plain, -e, -T, -e -T}, depth, coverage."""
import numpy as np
import pytest

from tests._hs import write_hs_like
from tests._oracle import Oracle, sorted_pairs
from tests.test_counts_gpu import first_diff
from tests.test_oracle_kat import pair_checksum

pytestmark = pytest.mark.gpu

MODES = [(0, False), (0, True), (1, False), (1, True), (2, False), (2, True)]


def scaffold_regions(idx, seed, n, valid_only=False):
    """regions on every kind of sequence: proportional to extent (mostly chromosomes), uniform over the sequences (mostly
    scaffolds), reaching past the sequence's end, up to u32::MAX, and (intersect only) inverted / empty ones"""
    rng = np.random.default_rng(seed)
    a = idx.synth_regions(seed, n, 20000)
    b = idx.synth_regions(seed + 1, n, 150)
    n_seq = idx.n_seqs
    c_chr = rng.integers(0, n_seq, size=2 * n).astype(np.uint32)
    c_s = rng.integers(0, 160000, size=2 * n).astype(np.uint32)
    c_e = (c_s + rng.choice(np.array([1, 50, 150, 2000, 40000, 300000, 6_000_000], dtype=np.uint32), size=2 * n)).astype(np.uint32)
    d_chr = rng.integers(0, n_seq, size=n // 4).astype(np.uint32)
    d_s = rng.choice(np.array([0, 100, 4294950000, 4999000, 5000100], dtype=np.uint32), size=n // 4)
    d_e = rng.choice(np.array([4294967295, 4294960001, 5000120, 5100000], dtype=np.uint32), size=n // 4)
    chr_ = np.concatenate([a[0], b[0], c_chr, d_chr])
    s = np.concatenate([a[1], b[1], c_s, d_s]).astype(np.uint32)
    e = np.concatenate([a[2], b[2], c_e, d_e]).astype(np.uint32)
    if not valid_only:
        k = len(chr_) // 40
        pick = rng.choice(len(chr_), size=k, replace=False)
        s[pick[: k // 2]], e[pick[: k // 2]] = e[pick[: k // 2]].copy(), s[pick[: k // 2]].copy()
        e[pick[k // 2:]] = s[pick[k // 2:]]
    perm = rng.permutation(len(chr_))
    return chr_[perm].copy(), s[perm].copy(), e[perm].copy()


@pytest.fixture(scope="module")
def scaf_gff(tmp_path_factory):
    import gffx_b200
    from tests import _oracle

    path = str(tmp_path_factory.mktemp("scaf") / "scaf.gff")
    gffx_b200.synth_gff(path, seed=9, n_genes=1500, quirks=True, n_scaffolds=1200)
    _oracle.build_index(path)
    return path


@pytest.fixture(scope="module", params=[None, "3000000"])
def scaf(request, gffx, scaf_gff):
    """default packing, and a small capacity that spreads the scaffolds over many virtual sequences"""
    import os

    old = os.environ.get("GFFX_VSEQ_CAP")
    if request.param:
        os.environ["GFFX_VSEQ_CAP"] = request.param
    try:
        idx = gffx.TreeIndexData.load_tree_index(scaf_gff, device=0)
    finally:
        if request.param:
            if old is None:
                del os.environ["GFFX_VSEQ_CAP"]
            else:
                os.environ["GFFX_VSEQ_CAP"] = old
    o = Oracle(scaf_gff)
    yield idx, o
    o.close()
    idx.close()


def test_more_sequences_than_windows(scaf):
    idx, _ = scaf
    assert idx.n_seqs > 1200
    assert idx.vseq(0)[3] <= 192


@pytest.mark.parametrize("mode,invert", MODES)
def test_pairs(scaf, mode, invert):
    idx, o = scaf
    regs = scaffold_regions(idx, 300 + mode, 4000)
    r = idx.query_features(regs, mode, invert)
    want = o.query_features(regs, mode, invert)
    assert r.n_pairs == want.shape[0]
    got = r.pairs if r.pairs is not None else np.zeros((0, 3), dtype=np.uint32)
    assert np.array_equal(sorted_pairs(got), sorted_pairs(want))   # coordinates come back in the sequence's own system


@pytest.mark.parametrize("mode,invert", MODES)
@pytest.mark.parametrize("entire,types", [(False, None), (True, None), (False, "gene,mRNA,exon"), (True, "exon, CDS ,,nosuchtype")])
def test_ranges(scaf, mode, invert, entire, types):
    idx, o = scaf
    regs = scaffold_regions(idx, 17 + mode, 3000)
    r = idx.intersect(regs, mode, invert, entire, types)
    want, n_pairs, n_warn = o.intersect(regs, mode, invert, entire, types)
    assert r.n_pairs == n_pairs
    assert r.n_ranges == want.shape[0] and np.array_equal(r.ranges, want)
    assert len(r.warn_fids) == n_warn


@pytest.mark.parametrize("seed", [1, 2])
def test_depth_and_coverage(scaf, seed):
    idx, o = scaf
    regs = scaffold_regions(idx, 50 + seed, 6000, valid_only=seed == 2)
    got, want = idx.depth(regs).to_tsv(), o.depth(regs)
    assert got == want, first_diff(got, want)
    got, want = idx.coverage(regs).to_tsv(), o.coverage(regs)
    assert got == want, first_diff(got, want)
    assert got.count(b"\n") > 500


def test_device_resident_batches(scaf):
    """the same through the batched job API with device-resident regions (what the sharded bench runs)"""
    import torch

    idx, o = scaf
    regs = scaffold_regions(idx, 77, 5000)
    dev = tuple(torch.from_numpy(a.view(np.int32)).cuda() for a in regs)
    r = idx.intersect(dev, 1, False, False, "exon", want_pairs=True)
    want, n_pairs, _ = o.intersect(regs, 1, False, False, "exon")
    assert r.n_pairs == n_pairs and np.array_equal(r.ranges, want)


def brute_pairs(trees, regs, mode, invert):
    """numpy restatement of tree.rs:110 + intersect.rs:145-163 per region, straight from the interval arrays"""
    out = []
    for c, rs, re in zip(*regs):
        t = trees[c]
        if not len(t):
            continue
        s, e = t[:, 0].astype(np.int64), t[:, 1].astype(np.int64)
        hit = (s < int(re)) & (e > int(rs))
        keep = (s >= int(rs)) & (e <= int(re)) if mode == 0 else (s <= int(rs)) & (e >= int(re)) if mode == 1 else np.ones(len(t), bool)
        sel = hit & (keep != invert)
        if sel.any():
            out.append(t[sel][:, [2, 0, 1]])
    return np.concatenate(out) if out else np.zeros((0, 3), dtype=np.uint32)


def test_hs_shaped_index_root_join(gffx, tmp_path):
    """583 sequences named and sized as in the reference's benchmark/hs.gff.{sqs,rix}: root join == brute force"""
    base, trees = write_hs_like(str(tmp_path))
    idx = gffx.TreeIndexData.load_tree_index(base, device=0)
    assert idx.n_seqs == 583
    rng = np.random.default_rng(583)
    n = 3000
    chr_ = np.concatenate([rng.integers(0, 583, size=n), rng.integers(0, 24, size=n)]).astype(np.uint32)
    ext = np.array([int(t[:, 1].max()) if len(t) else 1000 for t in trees], dtype=np.int64)
    s = (rng.random(2 * n) * ext[chr_]).astype(np.uint32)
    e = (s + rng.choice(np.array([1, 150, 20000, 1_000_000], dtype=np.uint32), size=2 * n)).astype(np.uint32)
    flip = rng.random(2 * n) < 0.02
    s, e = np.where(flip, e, s).astype(np.uint32), np.where(flip, s, e).astype(np.uint32)
    regs = (chr_, s, e)
    for mode, invert in MODES:
        r = idx.query_features(regs, mode, invert)
        want = brute_pairs(trees, regs, mode, invert)
        assert r.n_pairs == len(want), (mode, invert)
        got = r.pairs if r.pairs is not None else np.zeros((0, 3), dtype=np.uint32)
        assert pair_checksum(got) == pair_checksum(want), (mode, invert)
        assert np.array_equal(sorted_pairs(got), sorted_pairs(want)), (mode, invert)
    idx.close()
