"""SAM / BAM as the read source of depth / coverage: the product's reader (own BGZF inflate, own
record walk) == the oracle's (zlib + std::string tools) == the triples the test derives from the records it wrote.
PARITY UNPINNED BY THE REFERENCE: the reference reads alignments through rust-htslib, which cannot be built or run here
and ships no fixture; the three statements below agree on the SAM specification and htslib's published record rules."""
import gzip
import os
import random

import numpy as np
import pytest

from tests import _alnfiles as A
from tests._oracle import Oracle


@pytest.fixture(scope="module")
def host_index(gffx, synth_plain_gff):
    idx = gffx.TreeIndexData.load_tree_index(synth_plain_gff, device=-1)   # host side only: parsing needs no GPU
    o = Oracle(synth_plain_gff)
    yield idx, o
    o.close()
    idx.close()


def _targets(idx):
    names = idx.num_to_seqid
    # the header lists the index's sequences in another order, one name the index lacks, and only some of them
    t = [(names[3], 5_000_000), ("chrUn_not_in_index", 1000), (names[0], 9_000_000), (names[1], 7_000_000)]
    return t, {0: 3, 2: 0, 3: 1}


def _expected(recs, tid2num):
    out = []
    for r in recs:
        e = r.expected()
        if e is not None and e[0] in tid2num:
            out.append((tid2num[e[0]], e[1], e[2]))
    return np.array(out, dtype=np.uint32).reshape(-1, 3)


def _check(idx, o, path, want):
    st = {}
    got = np.stack(idx.parse_reads(path, st), axis=1) if True else None
    orc = np.stack(o.parse_reads(path), axis=1)
    assert np.array_equal(got, want), path
    assert np.array_equal(orc, want), path
    return st


@pytest.mark.parametrize("kind", ["sam", "sam_crlf", "bam", "bam_small_blocks", "bam_stored", "sam_bgzf", "sam_gz"])
def test_readers_agree(host_index, tmp_path, kind):
    idx, o = host_index
    targets, tid2num = _targets(idx)
    rng = random.Random(hash(kind) & 0xFFFF)
    recs = A.random_records(rng, targets, 20000, 4_000_000)
    want = _expected(recs, tid2num)
    p = str(tmp_path / ("x." + ("bam" if kind.startswith("bam") else "sam")))
    if kind == "sam":
        data = A.sam_text(targets, recs)
    elif kind == "sam_crlf":
        data = A.sam_text(targets, recs, crlf=True)
    elif kind == "bam":
        data = A.bgzf(A.bam_bytes(targets, recs), levels=(6, 1, 9))
    elif kind == "bam_small_blocks":   # records and the header straddle block boundaries at every offset
        data = A.bgzf(A.bam_bytes(targets, recs), block=977, rng=rng)
    elif kind == "bam_stored":
        data = A.bgzf(A.bam_bytes(targets, recs), levels=(0,))
    elif kind == "sam_bgzf":
        data = A.bgzf(A.sam_text(targets, recs), block=4000)
    else:
        data = gzip.compress(A.sam_text(targets, recs), 6)
    open(p, "wb").write(data)
    st = _check(idx, o, p, want)
    assert st["records"] == len(recs)
    assert st["container"] == {"sam": 0, "sam_crlf": 0, "sam_gz": 2}.get(kind, 1)
    assert st["records"] - st["unmapped"] - st["unknown_ref"] - st["skipped"] == want.shape[0]


@pytest.mark.parametrize("chunk", [1, 300, 5000, 70000])
def test_chunked_pipeline(host_index, tmp_path, monkeypatch, chunk):
    """the reader walks one chunk of inflated bytes while the next one inflates (reads.cpp): tiny chunks put a chunk boundary
    inside the header, inside records (carried over in the buffers' head room) and inside a record larger than the head room
    (the 320 KB CIGAR of the CG tag); BAM and bgzip-ed SAM"""
    idx, o = host_index
    targets, tid2num = _targets(idx)
    rng = random.Random(chunk)
    recs = A.random_records(rng, targets, 3000, 4_000_000)
    recs.insert(1500, A.Rec("long", 0, 0, 1000, [(1, "M"), (1, "D")] * 40000, tags=b"NMC\x05"))
    want = _expected(recs, tid2num)
    monkeypatch.setenv("GFFX_BGZF_CHUNK", str(chunk))
    p = str(tmp_path / "c.bam")
    open(p, "wb").write(A.bgzf(A.bam_bytes(targets, recs), block=977, rng=rng))
    assert _check(idx, o, p, want)["records"] == len(recs)
    p = str(tmp_path / "c.sam")
    open(p, "wb").write(A.bgzf(A.sam_text(targets, recs[:1500] + recs[1501:]), block=3000))
    _check(idx, o, p, _expected(recs[:1500] + recs[1501:], tid2num))


def test_sam_text_on_all_threads(gffx, host_index, tmp_path, monkeypatch):
    """plain SAM text is cut at line ends and parsed on the host threads (reads.cpp: sam_text_threads): same intervals in file
    order, same counters, and a bad record still fails the call"""
    idx, o = host_index
    targets, tid2num = _targets(idx)
    rng = random.Random(11)
    recs = A.random_records(rng, targets, 5000, 4_000_000)
    monkeypatch.setenv("GFFX_SAM_PAR_MIN", "0")
    for crlf in (False, True):
        p = str(tmp_path / "par.sam")
        open(p, "wb").write(A.sam_text(targets, recs, crlf=crlf))
        assert _check(idx, o, p, _expected(recs, tid2num))["records"] == len(recs)
    text = A.sam_text(targets, recs)
    cutat = text.index(b"\n", len(text) * 3 // 4) + 1
    open(p, "wb").write(text[:cutat] + b"broken\tline\n" + text[cutat:])
    with pytest.raises(gffx.GffxError):
        idx.parse_reads(p)
    open(p, "wb").write(b"@HD\tVN:1.6\n@SQ\tSN:%s\tLN:10\n" % targets[0][0].encode())   # a header and nothing else
    assert idx.parse_reads(p)[0].size == 0


def test_long_cigar_in_cg_tag(host_index, tmp_path):
    idx, o = host_index
    targets, tid2num = _targets(idx)
    long_cigar = [(1, "M"), (1, "D")] * 40000   # 80000 operations: more than the 16-bit count holds
    recs = [A.Rec("long", 0, 0, 1000, long_cigar, tags=b"NMC\x05XZZabc\0"), A.Rec("plain", 16, 2, 5, [(50, "M")]),
            # a soft clip of the read's whole length that is NOT a placeholder (no CG tag): taken as written
            A.Rec("clip", 0, 3, 77, [(30, "S"), (12, "N")], l_seq=30)]
    p = str(tmp_path / "cg.bam")
    open(p, "wb").write(A.bgzf(A.bam_bytes(targets, recs)))
    want = np.array([(3, 1000, 1000 + 80000), (0, 5, 55), (1, 77, 89)], dtype=np.uint32)
    _check(idx, o, p, want)


def test_edge_records(host_index, tmp_path):
    idx, o = host_index
    targets, tid2num = _targets(idx)
    recs = [A.Rec("a", 0, 0, 0, [(10, "M")]),                      # POS 1 -> pos 0
            A.Rec("b", 0, 0, -1, [(10, "M")]),                     # POS 0: no coordinate -> dropped
            A.Rec("c", 4, 0, 100, [(10, "M")]),                    # unmapped flag with a position -> dropped
            A.Rec("d", 0, -1, 100, [(10, "M")]),                   # RNAME '*'
            A.Rec("e", 0, 1, 100, [(10, "M")]),                    # target the index lacks
            A.Rec("f", 0, 2, 2_000_000_000, [(2_000_000_00, "N"), (5, "M")]),
            A.Rec("g", 0, 3, 7, []),                               # CIGAR '*' -> one base
            A.Rec("h", 0, 3, 9, [(5, "I"), (3, "S")])]             # no reference-consuming operation -> one base
    want = np.array([(3, 0, 10), (0, 2_000_000_000, 2_200_000_005), (1, 7, 8), (1, 9, 10)], dtype=np.uint32)
    for name, data in [("e.sam", A.sam_text(targets, recs)), ("e.bam", A.bgzf(A.bam_bytes(targets, recs)))]:
        p = str(tmp_path / name)
        open(p, "wb").write(data)
        _check(idx, o, p, want)
    # SAM only: an RNAME that no @SQ line declares is treated as unmapped, hex / octal FLAGs are numbers
    sam = A.sam_text(targets, [A.Rec("x", 0, 0, 10, [(4, "M")])]).decode().rstrip("\n").split("\n")
    body = sam[-1].split("\t")
    extra = ["\t".join([body[0], "0x10", "nowhere"] + body[3:]), "\t".join([body[0], "020", body[2]] + body[3:]), "\t".join([body[0], "0x4"] + body[2:])]
    p = str(tmp_path / "f.sam")
    open(p, "w").write("\n".join(sam + extra) + "\n")
    _check(idx, o, p, np.array([(3, 10, 14), (3, 10, 14)], dtype=np.uint32))


def test_malformed_inputs_fail_loudly(gffx, host_index, tmp_path):
    idx, o = host_index
    targets, _ = _targets(idx)
    good = A.bgzf(A.bam_bytes(targets, [A.Rec("a", 0, 0, 5, [(10, "M")])] * 50))
    cases = {"crc.bam": good[:40] + bytes([good[40] ^ 0x55]) + good[41:], "trunc.bam": good[: len(good) // 2], "tid.bam": A.bgzf(A.bam_bytes(targets, [A.Rec("a", 0, 9, 5, [(10, "M")])])),
             "short.sam": b"@SQ\tSN:x\tLN:5\nr1\t0\tx\t1\t30\t5M\n", "cigar.sam": A.sam_text(targets, [A.Rec("a", 0, 0, 5, [(10, "M")])]).replace(b"10M", b"10Q"),
             "x.cram": b"CRAM\x03\x00" + bytes(30)}
    for name, data in cases.items():
        p = str(tmp_path / name)
        open(p, "wb").write(data)
        with pytest.raises(gffx.GffxError) as ei:
            idx.parse_reads(p)
        assert ei.value.code in ((-1,) if name.endswith("cram") else (-3,)), name
    with pytest.raises(gffx.GffxError) as ei:
        idx.parse_reads(str(tmp_path / "missing.bam"))
    assert ei.value.code == -2


@pytest.mark.gpu
@pytest.mark.parametrize("kind", ["bam", "sam"])
def test_depth_and_coverage_from_alignments(gffx, synth_plain_gff, tmp_path, kind):
    """gffx depth / coverage -s reads.bam: rows == the oracle's on the triples of the same file == the BED of the same reads"""
    idx = gffx.TreeIndexData.load_tree_index(synth_plain_gff, device=0)
    o = Oracle(synth_plain_gff)
    try:
        names = idx.num_to_seqid
        targets = [(n, 50_000_000) for n in names[:6]] + [("other", 5)]
        regs = idx.synth_regions(77, 30000, 150)
        rng = random.Random(5)
        recs = []
        for c, s, e in zip(*regs):
            if int(c) >= 6:
                continue
            a = rng.randint(1, int(e - s) - 1)
            recs.append(A.Rec(f"r{len(recs)}", rng.choice([0, 16]), int(c), int(s), [(a, "M"), (3, "I"), (int(e - s) - a, "M")]))
        p = str(tmp_path / ("r." + kind))
        open(p, "wb").write(A.bgzf(A.bam_bytes(targets, recs)) if kind == "bam" else A.sam_text(targets, recs))
        triples = o.parse_reads(p)
        assert len(triples[0]) == len(recs)
        for cmd, ofn in [(gffx.run_depth, o.depth), (gffx.run_coverage, o.coverage)]:
            out = str(tmp_path / "out.tsv")
            cmd(gffx.DepthArgs(input=synth_plain_gff, source=p, output=out))
            got = open(out, "rb").read().split(b"\n")
            want = ofn(triples).split(b"\n")
            assert got[0] == want[0] and sorted(got[1:]) == sorted(want[1:])
    finally:
        o.close()
        idx.close()
