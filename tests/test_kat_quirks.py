"""Hand-derived known answers, one per item of SURVEY.md App. D (the quirk checklist).  Every expectation below was worked
out BY HAND from the cited reference lines (paths relative to the reference root) — not produced by any implementation —
and is checked against the CPU oracle here and, with `-m gpu`, against the CUDA engine through the C ABI.

Coordinates: a GFF line `s e` (1-based, closed) makes the root interval [min(s,e)-1, max(s,e)) (src/index_builder/core.rs:
102-109); a BED line `qs qe` is used as is: half-open against root intervals (src/utils/tree.rs:110), CLOSED against the raw
columns 4/5 of a line (src/commands/intersect.rs:500-517)."""
import os

import numpy as np
import pytest

from tests._oracle import Oracle

C, CR, O = 0, 1, 2   # OverlapMode::Contained, ContainsRegion, Overlap


def L(seq, typ, s, e, attrs, strand="+"):
    return f"{seq}\tsrc\t{typ}\t{s}\t{e}\t.\t{strand}\t.\t{attrs}".encode()


class Kat:
    """a tiny GFF on disk + its index (built by the oracle's restated `gffx index`)"""

    def __init__(self, tmp, name, lines, last_newline=True):
        from tests import _oracle

        self.lines = lines
        self.path = os.path.join(str(tmp), name + ".gff")
        text = b"\n".join(lines) + (b"\n" if last_newline else b"")
        open(self.path, "wb").write(text)
        self.text = text
        _oracle.build_index(self.path)
        self.bed_dir = str(tmp)

    def bed(self, name, content):
        p = os.path.join(self.bed_dir, name)
        open(p, "wb").write(content)
        return p

    def out(self, *idx):
        """the bytes of the given GFF lines as they sit in the file (with their newline when they have one)"""
        pieces = []
        for i in idx:
            pieces.append(self.lines[i] + (b"\n" if i + 1 < len(self.lines) or self.text.endswith(b"\n") else b""))
        return b"".join(pieces)


class OracleImpl:
    name = "oracle"

    def __init__(self, k):
        self.k, self.o = k, Oracle(k.path)

    def close(self):
        self.o.close()

    def intersect(self, bed, mode, invert=False, entire=False, types=None):
        regs = self.o.parse_bed_intersect(bed)
        r, n_pairs, n_warn = self.o.intersect(regs, mode, invert, entire, types)
        return b"".join(self.k.text[s:e] for s, e in r.tolist()), n_pairs

    def depth(self, bed):
        return self.o.depth(self.o.parse_bed_counts(bed))

    def coverage(self, bed):
        return self.o.coverage(self.o.parse_bed_counts(bed))


class GpuImpl:
    name = "gpu"

    def __init__(self, k):
        import gffx_b200

        self.k, self.idx = k, gffx_b200.TreeIndexData.load_tree_index(k.path, device=0)

    def close(self):
        self.idx.close()

    def intersect(self, bed, mode, invert=False, entire=False, types=None):
        regs = self.idx.parse_bed_file(bed)
        r = self.idx.intersect(regs, mode, invert, entire, types)
        return b"".join(self.k.text[s:e] for s, e in r.ranges.tolist()), r.n_pairs

    def depth(self, bed):
        return self.idx.depth(self.idx.parse_bed_counts(bed)).to_tsv()

    def coverage(self, bed):
        return self.idx.coverage(self.idx.parse_bed_counts(bed)).to_tsv()


@pytest.fixture(params=["oracle", pytest.param("gpu", marks=pytest.mark.gpu)])
def impl(request):
    made = []

    def make(k):
        i = (OracleImpl if request.param == "oracle" else GpuImpl)(k)
        made.append(i)
        return i

    yield make
    for i in made:
        i.close()


DEPTH_H = b"id\tchr\tstart\tend\tdepth\n"
COV_H = b"id\tchr\tstart\tend\tbreadth\tfraction\n"


def rows(header, *r):
    return header + b"".join(sorted(r))


# ---- D.1: root test half-open 0-based (tree.rs:110) vs line test closed raw 1-based (intersect.rs:500-517) -------------------
def test_d1_half_open_roots_closed_lines(tmp_path, impl):
    k = Kat(tmp_path, "d1", [L("chr1", "gene", 101, 200, "ID=g1"), L("chr1", "exon", 101, 150, "ID=e1;Parent=g1")])
    i = impl(k)
    # root g1 = [100,200).  (200,300): 200 > 200 is false -> no hit at all
    assert i.intersect(k.bed("a.bed", b"chr1\t200\t300\n"), O) == (b"", 0)
    # (199,300): hit.  gene 101..200: 199<=200<=300 (clause 2) -> emitted; exon 101..150: no clause holds
    assert i.intersect(k.bed("b.bed", b"chr1\t199\t300\n"), O) == (k.out(0), 1)
    # (50,100): 100 < 100 is false -> no hit, although the closed line test would not match either
    assert i.intersect(k.bed("c.bed", b"chr1\t50\t100\n"), O) == (b"", 0)
    # (50,101): hit; both lines start at 101 and 50<=101<=101 (clause 1)
    assert i.intersect(k.bed("d.bed", b"chr1\t50\t101\n"), O) == (k.out(0, 1), 1)
    # (150,190): exon 101..150 is emitted through clause 2 (150<=150<=190) although [150,190) shares no base with it
    assert i.intersect(k.bed("e.bed", b"chr1\t150\t190\n"), O) == (k.out(0, 1), 1)


# ---- D.2: --invert acts on the root test only (intersect.rs:161); the line test never inverts --------------------------------
def test_d2_invert_is_root_level_only(tmp_path, impl):
    k = Kat(tmp_path, "d2", [L("chr1", "gene", 101, 200, "ID=g1"), L("chr1", "exon", 101, 150, "ID=e1;Parent=g1")])
    i = impl(k)
    assert i.intersect(k.bed("a.bed", b"chr1\t150\t190\n"), O, invert=True) == (b"", 0)          # keep = true for Overlap: invert drops everything
    # -c -I, (150,190): hit, contained is false (100 >= 150 fails) -> pair; plain contained line test: 101>=150 fails for both lines
    assert i.intersect(k.bed("a.bed", b"chr1\t150\t190\n"), C, invert=True) == (b"", 1)
    # -c -I, (90,250): contained holds -> inverted away -> no pair, no output
    assert i.intersect(k.bed("b.bed", b"chr1\t90\t250\n"), C, invert=True) == (b"", 0)
    # -c, (90,250): pair; both lines lie inside [90,250] -> both emitted
    assert i.intersect(k.bed("b.bed", b"chr1\t90\t250\n"), C) == (k.out(0, 1), 1)
    # -C -I, (90,250): contains-region fails (100 <= 90 is false) -> pair; line test start<=90 fails -> nothing emitted
    assert i.intersect(k.bed("b.bed", b"chr1\t90\t250\n"), CR, invert=True) == (b"", 1)


# ---- D.3: -e with -T takes the match-only path (intersect.rs:619) --------------------------------------------------------------
def test_d3_entire_group_with_types_is_match_only(tmp_path, impl):
    k = Kat(tmp_path, "d3", [L("chr1", "gene", 101, 200, "ID=g1"), L("chr1", "mRNA", 101, 200, "ID=m1;Parent=g1"),
                             L("chr1", "exon", 101, 150, "ID=e1;Parent=m1"), L("chr1", "exon", 181, 200, "ID=e2;Parent=m1")])
    i = impl(k)
    bed = k.bed("b.bed", b"chr1\t185\t190\n")
    assert i.intersect(bed, O)[0] == k.out(0, 1, 3)                       # e1 101..150 matches no clause with (185,190)
    assert i.intersect(bed, O, entire=True)[0] == k.out(0, 1, 2, 3)       # the whole block, verbatim
    assert i.intersect(bed, O, types="exon")[0] == k.out(3)
    assert i.intersect(bed, O, entire=True, types="exon")[0] == k.out(3)  # NOT the block's exons: the line test still applies
    assert i.intersect(bed, O, types=" exon ,, gene")[0] == k.out(0, 3)   # tokens are trimmed, empty ones dropped (:252-259)


# ---- D.4 / D.5: a line is tested against ALL regions of ITS OWN sequence name; block membership is physical -------------------
def test_d4_d5_line_test_by_name_and_physical_blocks(tmp_path, impl):
    k = Kat(tmp_path, "d45", [
        L("chr1", "gene", 101, 200, "ID=g1"),                     # 0  root [100,200)
        L("chr1", "exon", 101, 120, "ID=e1;Parent=g1"),           # 1
        L("chr1", "exon", 2001, 2010, "ID=far;Parent=g1"),        # 2  far outside its root
        L("chr2", "exon", 301, 310, "ID=x2;Parent=g1"),           # 3  names another sequence
        L("chr1", "region", 1, 1000, "ID=reg1"),                  # 4  skipped type: not indexed, physically inside g1's block
        L("chr1", "gene", 1001, 1100, "ID=g2"),                   # 5  root [1000,1100)
        L("chr2", "gene", 5001, 5100, "ID=g3"),                   # 6  root on chr2 (so that chr2 is a known sequence)
    ])
    i = impl(k)
    # R1 (150,160) hits g1 only; R2 (2000,2020) and R3 chr2 (300,320) hit no root but take part in the line tests (:621-633)
    bed = k.bed("a.bed", b"chr1\t150\t160\nchr1\t2000\t2020\nchr2\t300\t320\n")
    # g1: 101<=150<=200 (clause 3); e1: nothing; far: 2000<=2001<=2020; x2 (chr2): 300<=301<=320; reg1: 1<=150<=1000
    assert i.intersect(bed, O) == (k.out(0, 2, 3, 4), 1)
    # without R2 / R3 the same block keeps only the lines R1 reaches
    assert i.intersect(k.bed("b.bed", b"chr1\t150\t160\n"), O) == (k.out(0, 4), 1)
    # depth: every line of a routed block that has 9 columns and an ID counts, whatever its type or sequence name
    # (depth.rs:131-157); chr = column 1 of the line.  (140,170) is routed to g1: g1 [100,200) hit; e1 [100,120) no;
    # far no; x2 [300,310) no; reg1 [0,1000) hit
    assert i.depth(k.bed("c.bed", b"chr1\t140\t170\n")) == rows(DEPTH_H, b"g1\tchr1\t100\t200\t1\n", b"reg1\tchr1\t0\t1000\t1\n")
    # a region on chr1 that overlaps x2's coordinates: x2 counts although its line says chr2 (seqid ignored), and reports chr2
    assert i.depth(k.bed("d.bed", b"chr1\t190\t305\n")) == rows(DEPTH_H, b"g1\tchr1\t100\t200\t1\n", b"reg1\tchr1\t0\t1000\t1\n",
                                                               b"x2\tchr2\t300\t310\t1\n")


# ---- D.6: intersect keeps start >= end regions (intersect.rs:223-225), depth / coverage drop them (depth.rs:488) ---------------
def test_d6_inverted_regions(tmp_path, impl):
    k = Kat(tmp_path, "d6", [L("chr1", "gene", 101, 200, "ID=g1"), L("chr1", "exon", 101, 150, "ID=e1;Parent=g1"),
                             L("chr1", "exon", 160, 170, "ID=e2;Parent=g1")])
    i = impl(k)
    bed = k.bed("a.bed", b"chr1\t190\t150\n")
    # root: 100 < 150 && 200 > 190 -> hit.  gene 101..200: 101<=190<=200 (clause 3).  e1 101..150: 101<=150<=150 (clause 4, qe inside).
    # e2 160..170: 190<=160 no; 190<=170 no; 160<=190<=170 no; 160<=150 no -> not emitted
    assert i.intersect(bed, O) == (k.out(0, 1), 1)
    assert i.depth(bed) == DEPTH_H
    assert i.coverage(bed) == COV_H
    # an empty region (150,150): 100 < 150 && 200 > 150 -> hit; gene: clause 3; e1: 150<=150<=150 (clause 2)
    assert i.intersect(k.bed("b.bed", b"chr1\t150\t150\n"), O) == (k.out(0, 1), 1)


# ---- D.7: duplicate root IDs: two tree intervals, one fid, block = LAST .gof record (index_builder/core.rs:139-142, gof.rs:32-37)
def test_d7_duplicate_root_ids(tmp_path, impl):
    k = Kat(tmp_path, "d7", [
        L("chr1", "gene", 101, 200, "ID=dup"),                    # 0  feature 0; fid = 2 (the LAST feature named dup)
        L("chr1", "exon", 101, 150, "ID=e1;Parent=dup"),          # 1
        L("chr1", "gene", 1001, 1100, "ID=dup"),                  # 2  feature 2; fid = 2
        L("chr1", "exon", 1001, 1050, "ID=e2;Parent=dup"),        # 3
    ])
    i = impl(k)
    bed = k.bed("a.bed", b"chr1\t150\t160\n")                      # hits the FIRST interval [100,200) only
    # -e: the block of fid 2 is its last .gof record = lines 2..3, not the lines the region lies on
    assert i.intersect(bed, O, entire=True) == (k.out(2, 3), 1)
    # match-only: that block's lines (1001.., 1001..) match nothing of (150,160)
    assert i.intersect(bed, O) == (b"", 1)
    # a region over both intervals: two pairs (one per tree interval), one root, one block
    assert i.intersect(k.bed("b.bed", b"chr1\t150\t1010\n"), O, entire=True) == (k.out(2, 3), 2)
    # depth routes once per (region, fid) (depth.rs:239-241) into the LAST record's slice: dup's second line [1000,1100) hit, e2 [1000,1050) hit
    assert i.depth(k.bed("b.bed", b"chr1\t150\t1010\n")) == rows(DEPTH_H, b"dup\tchr1\t1000\t1100\t1\n", b"e2\tchr1\t1000\t1050\t1\n")


# ---- D.8: the line test reads columns 4/5 raw: no swap (intersect.rs:526-538); the indexer and depth / coverage swap ------------
def test_d8_reversed_columns(tmp_path, impl):
    k = Kat(tmp_path, "d8", [L("chr1", "gene", 200, 101, "ID=g1"), L("chr1", "exon", 120, 130, "ID=e1;Parent=g1")])
    i = impl(k)
    # root [100,200) (swapped by the indexer).  Overlap with (150,160): raw start 200, end 101: 150<=200<=160 no; 150<=101 no;
    # 200<=150 no; 200<=160 no -> the gene line is NOT emitted although the region lies inside it; e1 120..130: no clause
    assert i.intersect(k.bed("a.bed", b"chr1\t150\t160\n"), O) == (b"", 1)
    # Contained with (90,250): start>=qs && end<=qe reads 200>=90 && 101<=250 -> emitted; e1 too
    assert i.intersect(k.bed("b.bed", b"chr1\t90\t250\n"), C) == (k.out(0, 1), 1)
    # ContainsRegion with (150,160): 200<=150 fails for the gene; e1 120<=150 but 130>=160 fails
    assert i.intersect(k.bed("a.bed", b"chr1\t150\t160\n"), CR) == (b"", 1)
    assert i.depth(k.bed("a.bed", b"chr1\t150\t160\n")) == rows(DEPTH_H, b"g1\tchr1\t100\t200\t1\n")


# ---- D.9: depth reports the hull of the HIT lines (depth.rs:198-199), coverage the hull of ALL lines (coverage.rs:345-346) ------
def test_d9_hulls(tmp_path, impl):
    k = Kat(tmp_path, "d9", [L("chr1", "gene", 101, 200, "ID=g1"), L("chr1", "CDS", 101, 120, "ID=c1;Parent=g1"),
                             L("chr1", "CDS", 181, 200, "ID=c1;Parent=g1")])
    i = impl(k)
    bed = k.bed("a.bed", b"chr1\t100\t110\n")
    assert i.depth(bed) == rows(DEPTH_H, b"g1\tchr1\t100\t200\t1\n", b"c1\tchr1\t100\t120\t1\n")
    assert i.coverage(bed) == rows(COV_H, b"g1\tchr1\t100\t200\t10\t0.100000\n", b"c1\tchr1\t100\t200\t10\t0.100000\n")
    # a region in the gap between the two CDS lines: c1 is not hit (no depth row) but coverage still reports it with breadth 0
    gap = k.bed("b.bed", b"chr1\t130\t170\n")
    assert i.depth(gap) == rows(DEPTH_H, b"g1\tchr1\t100\t200\t1\n")
    assert i.coverage(gap) == rows(COV_H, b"g1\tchr1\t100\t200\t40\t0.400000\n", b"c1\tchr1\t100\t200\t0\t0.000000\n")
    # two regions on the two lines: ONE region each, depth counts regions per ID -> 2; breadth = 5 + 10
    two = k.bed("c.bed", b"chr1\t115\t125\nchr1\t170\t190\n")
    assert i.depth(two) == rows(DEPTH_H, b"g1\tchr1\t100\t200\t2\n", b"c1\tchr1\t100\t200\t2\n")
    assert i.coverage(two) == rows(COV_H, b"g1\tchr1\t100\t200\t30\t0.300000\n", b"c1\tchr1\t100\t200\t15\t0.150000\n")
    # one region over both lines counts ONCE for c1 (deduplicated per region, depth.rs:202-206)
    assert i.depth(k.bed("d.bed", b"chr1\t110\t190\n")) == rows(DEPTH_H, b"g1\tchr1\t100\t200\t1\n", b"c1\tchr1\t100\t200\t1\n")


# ---- D.10: fast_id takes the first "ID=" anywhere in column 9 (depth.rs:104-118) ------------------------------------------------
def test_d10_fast_id(tmp_path, impl):
    k = Kat(tmp_path, "d10", [L("chr1", "gene", 101, 200, "geneID=abc;ID=real1"),
                              L("chr1", "exon", 101, 150, "Note=x;ID=spaced id;Parent=abc"),
                              L("chr1", "region", 160, 170, "Parent=abc")])   # no ID: only a skipped type may lack one (index_builder/core.rs:113-116)
    i = impl(k)
    # "geneID=abc": the match is the ID= inside geneID= -> "abc"; "ID=spaced id" ends at the space; a line without ID= is ignored
    assert i.depth(k.bed("a.bed", b"chr1\t100\t200\n")) == rows(DEPTH_H, b"abc\tchr1\t100\t200\t1\n", b"spaced\tchr1\t100\t150\t1\n")


# ---- D.11: touching ranges merge: s <= current end (common.rs:216, coverage.rs:98) ----------------------------------------------
def test_d11_touching_ranges_merge(tmp_path, impl):
    k = Kat(tmp_path, "d11", [L("chr1", "gene", 101, 200, "ID=g1"), L("chr1", "gene", 301, 400, "ID=g2"), L("chr1", "gene", 501, 600, "ID=g3")])
    i = impl(k)
    import ctypes

    # g1+g2 selected: their blocks are contiguous (end == next start) -> ONE merged range; g1+g3: two ranges, g2's bytes left out
    o = Oracle(k.path) if i.name == "oracle" else None
    for bed, want, n_ranges in ((b"chr1\t150\t350\n", k.out(0, 1), 1), (b"chr1\t150\t160\nchr1\t550\t560\n", k.out(0, 2), 2)):
        p = k.bed("x.bed", bed)
        assert i.intersect(p, O, entire=True)[0] == want
        if o:
            assert len(o.intersect(o.parse_bed_intersect(p), O, False, True, None)[0]) == n_ranges
        else:
            assert i.idx.intersect(i.idx.parse_bed_file(p), O, False, True, None).n_ranges == n_ranges
    if o:
        o.close()
    # coverage: [100,150) and [150,200) touch -> one merged interval; breadth is the union either way
    assert i.coverage(k.bed("c.bed", b"chr1\t100\t150\nchr1\t150\t200\n")) == rows(COV_H, b"g1\tchr1\t100\t200\t100\t1.000000\n")


# ---- D.12: '\r' is never stripped; the last line may lack its '\n' (intersect.rs:286-297) ---------------------------------------
def test_d12_crlf_and_missing_final_newline(tmp_path, impl):
    k = Kat(tmp_path, "d12", [L("chr1", "gene", 101, 200, "ID=g1"), L("chr1", "exon", 101, 150, "ID=e1;Parent=g1") + b"\r",
                              L("chr1", "exon", 160, 170, "Parent=g1;ID=e2") + b"\r", L("chr1", "exon", 181, 200, "ID=e3;Parent=g1")],
            last_newline=False)
    i = impl(k)
    out, _ = i.intersect(k.bed("a.bed", b"chr1\t90\t250\n"), C)
    assert out == k.text                                     # every line, '\r' included, the last one without a newline
    assert out.endswith(b"Parent=g1") and b"\r\n" in out
    # depth: the ID runs to ';', space or tab — a trailing '\r' is part of it (depth.rs:110-113): "e2\r"
    assert i.depth(k.bed("a.bed", b"chr1\t90\t250\n")) == rows(DEPTH_H, b"g1\tchr1\t100\t200\t1\n", b"e1\tchr1\t100\t150\t1\n",
                                                              b"e2\r\tchr1\t159\t170\t1\n", b"e3\tchr1\t180\t200\t1\n")


# ---- App. A.1: the BED dialects -------------------------------------------------------------------------------------------------
def test_bed_dialects(tmp_path, impl):
    k = Kat(tmp_path, "bed", [L("chr1", "gene", 101, 200, "ID=g1")])
    i = impl(k)
    # intersect: header / unknown-chromosome lines are skipped BEFORE their numbers are looked at (intersect.rs:220-224);
    # any ASCII whitespace separates fields; extra columns are ignored; short lines are skipped
    ok = k.bed("ok.bed", b"track name=x\n#c\n\nchrNope\tabc\t-1\nchr1 150\x0c160 extra\t.\nchr1\t5\n")
    assert i.intersect(ok, O) == (k.out(0), 1)
    # one leading '+' is part of lexical_core's u32 language (lexical-parse-integer 1.0, parse_sign): (+300,+310) misses g1
    assert i.intersect(k.bed("plus.bed", b"chr1\t+150\t+160\nchr1\t+300\t310\n"), O) == (k.out(0), 1)
    with pytest.raises(Exception):   # a known chromosome with a bad number aborts the command (intersect.rs:223-224)
        i.intersect(k.bed("bad.bed", b"chr1\t150\t16x\n"), O)
    with pytest.raises(Exception):
        i.intersect(k.bed("neg.bed", b"chr1\t-5\t160\n"), O)
    # depth / coverage: only tab and space split; numbers parse first, '+' is fine, trailing '\r' is trimmed off field 3,
    # bad lines are dropped one by one (depth.rs:470-492)
    cb = k.bed("c.bed", b"chr1\t+150\t160\r\nchr1\tx\t170\nchr1\t150\x0c160\t170\nchr1\t199\t99999999999\nchr1 100  101\n")
    assert i.depth(cb) == rows(DEPTH_H, b"g1\tchr1\t100\t200\t2\n")
    # 1/128 = 0.0078125 exactly: `{:.6}` rounds the binary value half to even -> 0.007812 (coverage.rs:472)
    k2 = Kat(tmp_path, "frac", [L("chr1", "gene", 1, 128, "ID=w")])
    i2 = impl(k2)
    assert i2.coverage(k2.bed("f.bed", b"chr1\t5\t6\n")) == rows(COV_H, b"w\tchr1\t0\t128\t1\t0.007812\n")
