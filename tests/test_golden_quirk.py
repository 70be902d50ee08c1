"""Holds the CPU oracle — and, on the GPU box, the CUDA engine — to answers that NEITHER of them produced:
tests/golden/quirk/ was written by tests/golden/naive_eval.py, a pure-Python O(L x Q) restatement of
intersect.rs:105-169,232-523, common.rs:188-245, depth.rs:121-293,449-495 and coverage.rs:92-450 that shares no code with
oracle/ (tests/golden/make_golden.py quirk).  The fixture: ~2.5 k GFF lines with every quirk of SURVEY.md App. D + 12
scaffolds, 421 BED lines placed on every boundary of every root interval (offsets -1 / 0 / +1, inverted, empty, beyond the
sequence, unknown chromosomes, headers, CRLF, '+' signs, overflow).  24 intersect cases (mode x invert x plain/-e/-T/-e -T),
the depth table and the coverage table, byte for byte."""
import hashlib
import json
import lzma
import os
import shutil
import struct

import numpy as np
import pytest

from tests._oracle import Oracle
from tests.conftest import GOLDEN, has_cuda
from tests.test_oracle_kat import pair_checksum

Q = os.path.join(GOLDEN, "quirk")
GOLD = json.load(open(os.path.join(Q, "golden.json")))
CASES = sorted(GOLD["intersect"])
TYPES = {"plain": (False, None), "e": (True, None), "T": (False, "gene,mRNA,exon"), "eT": (True, "exon, CDS ,,nosuchtype")}


@pytest.fixture(scope="module")
def quirk_dir(tmp_path_factory):
    d = str(tmp_path_factory.mktemp("gquirk"))
    with lzma.open(os.path.join(Q, "quirk.gff.xz"), "rb") as f, open(os.path.join(d, "quirk.gff"), "wb") as g:
        g.write(f.read())
    for ext in ("sqs", "rix", "rit", "gof"):
        shutil.copy(os.path.join(Q, "quirk.gff." + ext), os.path.join(d, "quirk.gff." + ext))
    for b in ("intersect.bed", "counts.bed"):
        shutil.copy(os.path.join(Q, b), os.path.join(d, b))
    return d


def split_case(case):
    m, t = case.split("/")
    return {"contained": 0, "contains_region": 1, "overlap": 2}[m.split("+")[0]], m.endswith("+invert"), TYPES[t][0], TYPES[t][1]


def digests(text, ranges):
    h, t = hashlib.sha256(), hashlib.sha256()
    for s, e in np.asarray(ranges, dtype=np.uint64).reshape(-1, 2).tolist():
        h.update(struct.pack("<QQ", s, e))
        t.update(text[s:e])
    return h.hexdigest(), t.hexdigest()


def check_intersect(want, n_pairs, pairs, n_roots, ranges, warn, text):
    assert n_pairs == want["pairs"]
    if pairs is not None:
        assert pair_checksum(pairs) == want["pair_checksum"]
    if n_roots is not None:
        assert n_roots == want["roots"]
    r = np.asarray(ranges, dtype=np.uint64).reshape(-1, 2)
    assert len(r) == want["ranges"]
    assert int((r[:, 1] - r[:, 0]).sum()) == want["bytes"]
    assert [list(map(int, x)) for x in r[:4]] == want["first_ranges"]
    assert digests(text, r) == (want["ranges_sha256"], want["text_sha256"])
    assert sorted(int(w) for w in warn) == want["warn"]


def test_the_index_writer_reproduces_the_committed_index_files(oracle, quirk_dir, tmp_path):
    """the fixture's .sqs/.rix/.rit/.gof are what `gffx index` (restated) writes for the fixture text today as well"""
    p = str(tmp_path / "quirk.gff")
    shutil.copy(os.path.join(quirk_dir, "quirk.gff"), p)
    oracle.build_index(p)
    for ext in ("sqs", "rix", "rit", "gof"):
        assert open(p + "." + ext, "rb").read() == open(os.path.join(quirk_dir, "quirk.gff." + ext), "rb").read(), ext


@pytest.mark.parametrize("case", CASES)
def test_oracle_intersect_equals_the_naive_evaluator(quirk_dir, case):
    mode, invert, entire, types = split_case(case)
    gff = os.path.join(quirk_dir, "quirk.gff")
    text = open(gff, "rb").read()
    with Oracle(gff) as o:
        regs = o.parse_bed_intersect(os.path.join(quirk_dir, "intersect.bed"))
        assert regs[0].shape[0] == GOLD["n_regions_intersect"]
        ranges, n_pairs, n_warn = o.intersect(regs, mode, invert, entire, types)
        pairs = o.query_features(regs, mode, invert)
        want = GOLD["intersect"][case]
        assert n_warn == len(want["warn"])              # the oracle reports how many fids it warned about, not which
        check_intersect(want, n_pairs, pairs, len(np.unique(pairs[:, 0])), ranges, want["warn"], text)


def test_oracle_depth_and_coverage_equal_the_naive_evaluator(quirk_dir):
    gff = os.path.join(quirk_dir, "quirk.gff")
    with Oracle(gff) as o:
        regs = o.parse_bed_counts(os.path.join(quirk_dir, "counts.bed"))
        assert regs[0].shape[0] == GOLD["n_regions_counts"]
        d, c = o.depth(regs), o.coverage(regs)
    assert d == lzma.open(os.path.join(Q, "depth.tsv.xz")).read()
    assert c == lzma.open(os.path.join(Q, "coverage.tsv.xz")).read()
    assert d.count(b"\n") - 1 == GOLD["depth_rows"] and c.count(b"\n") - 1 == GOLD["coverage_rows"]
    assert GOLD["chrom_order_dependent_ids"] == []   # no row of the fixture depends on the reference's hash-map order


def test_naive_evaluator_reproduces_its_committed_answers(quirk_dir):
    """the committed goldens are what naive_eval.py says today (a changed evaluator must regenerate them on purpose)"""
    from tests.golden import naive_eval as nv

    ix = nv.Index(os.path.join(quirk_dir, "quirk.gff"))
    regions = nv.parse_bed_file(open(os.path.join(quirk_dir, "intersect.bed"), "rb").read(), ix.seqid_to_num)
    for case in ("overlap/plain", "contained+invert/eT", "contains_region/e"):
        mode, invert, entire, types = split_case(case)
        r = nv.intersect_run(ix, regions, mode, invert, entire, types)
        want = GOLD["intersect"][case]
        assert (len(r["pairs"]), nv.pair_checksum(r["pairs"]), len(r["ranges"])) == (want["pairs"], want["pair_checksum"], want["ranges"])
        assert nv.ranges_digest(ix, r["ranges"]) == (want["ranges_sha256"], want["text_sha256"])
    cregs = nv.parse_bed_counts(open(os.path.join(quirk_dir, "counts.bed"), "rb").read(), ix.seqid_to_num)
    assert nv.depth_tsv(nv.depth_run(ix, cregs)[0]) == lzma.open(os.path.join(Q, "depth.tsv.xz")).read()
    assert nv.coverage_tsv(nv.coverage_run(ix, cregs)[0]) == lzma.open(os.path.join(Q, "coverage.tsv.xz")).read()


# ---- the CUDA engine against the same bytes ----------------------------------------------------------------------------
@pytest.fixture(scope="module")
def gpu_idx(gffx, quirk_dir):
    idx = gffx.TreeIndexData.load_tree_index(os.path.join(quirk_dir, "quirk.gff"), device=0)
    yield idx
    idx.close()


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES)
def test_gpu_intersect_equals_the_naive_evaluator(gpu_idx, quirk_dir, case):
    mode, invert, entire, types = split_case(case)
    text = open(os.path.join(quirk_dir, "quirk.gff"), "rb").read()
    regs = gpu_idx.parse_bed_file(os.path.join(quirk_dir, "intersect.bed"))
    assert len(regs[0]) == GOLD["n_regions_intersect"]
    r = gpu_idx.intersect(regs, mode, invert, entire, types, want_pairs=True)
    pairs = r.pairs if r.pairs is not None else np.zeros((0, 3), dtype=np.uint32)
    check_intersect(GOLD["intersect"][case], r.n_pairs, pairs, r.n_roots, r.ranges, r.warn_fids, text)


@pytest.mark.gpu
def test_gpu_depth_and_coverage_equal_the_naive_evaluator(gpu_idx, quirk_dir):
    bed = os.path.join(quirk_dir, "counts.bed")
    host = gpu_idx.parse_bed_counts(bed)
    assert len(host[0]) == GOLD["n_regions_counts"]
    assert gpu_idx.depth(host).to_tsv() == lzma.open(os.path.join(Q, "depth.tsv.xz")).read()
    assert gpu_idx.coverage(host).to_tsv() == lzma.open(os.path.join(Q, "coverage.tsv.xz")).read()
    dev = gpu_idx.parse_bed_device(bed, 1, device=0)     # the device BED parser (K9) reads the same 395 regions out of it
    assert int(dev[0].numel()) == GOLD["n_regions_counts"]
    assert gpu_idx.depth(dev).to_tsv() == lzma.open(os.path.join(Q, "depth.tsv.xz")).read()
