"""Pins the CPU oracle on the hand-derived known-answer fixture of SURVEY.md App. C and on the
reference's own index files (tests/golden/ps, answers by numpy brute force: make_golden.py)."""
import json
import os

import numpy as np
import pytest

from tests._oracle import Oracle, sorted_pairs
from tests.conftest import GOLDEN

CONTAINED, CONTAINS, OVERLAP = 0, 1, 2
LINES = open(os.path.join(GOLDEN, "kat.gff"), "rb").read().split(b"\n")[:-1]
NAME = {b"g1": 1, b"m1": 2, b"e1": 3, b"e2": 4, b"g2": 5, b"e3": 6, b"g3": 7}  # line numbers in kat.gff


def emitted(path, ranges):
    text = open(path, "rb").read()
    return b"".join(text[s:e] for s, e in ranges.tolist())


def lines_of(*names):
    return b"".join(LINES[NAME[n]] + b"\n" for n in names)


def bed_a(o):
    num = {n: i for i, n in enumerate(o.seq_names)}
    # chr3 is unknown and dropped by the BED parser before the engine is reached
    return (np.array([num["chr1"]] * 3, dtype=np.uint32), np.array([150, 200, 90], dtype=np.uint32), np.array([190, 310, 250], dtype=np.uint32))


def bed_b(o):
    num = {n: i for i, n in enumerate(o.seq_names)}
    return (np.array([num["chr1"]], dtype=np.uint32), np.array([185], dtype=np.uint32), np.array([190], dtype=np.uint32))


@pytest.mark.parametrize(
    "mode,invert,entire,types,n_pairs,want",
    [
        (OVERLAP, False, False, None, 3, (b"g1", b"m1", b"e1", b"e2", b"g2", b"e3")),
        (OVERLAP, False, True, None, 3, (b"g1", b"m1", b"e1", b"e2", b"g2", b"e3")),
        (CONTAINED, False, False, None, 1, (b"g1", b"m1", b"e1", b"e2")),
        (CONTAINED, True, False, None, 2, (b"g1", b"m1", b"e1", b"e2")),
        (CONTAINS, False, False, None, 1, (b"g1", b"m1")),
        (OVERLAP, True, False, None, 0, ()),
    ],
)
def test_kat_intersect_bed_a(kat_gff, mode, invert, entire, types, n_pairs, want):
    with Oracle(kat_gff) as o:
        ranges, npairs, nwarn = o.intersect(bed_a(o), mode, invert, entire, types)
        assert npairs == n_pairs and nwarn == 0
        assert emitted(kat_gff, ranges) == lines_of(*want)


@pytest.mark.parametrize(
    "entire,types,want",
    [(False, None, (b"g1", b"m1", b"e2")), (True, None, (b"g1", b"m1", b"e1", b"e2")), (False, "exon", (b"e2",)), (True, "exon", (b"e2",))],
)
def test_kat_intersect_bed_b(kat_gff, entire, types, want):
    with Oracle(kat_gff) as o:
        ranges, _, _ = o.intersect(bed_b(o), OVERLAP, False, entire, types)
        assert emitted(kat_gff, ranges) == lines_of(*want)


def test_kat_depth_coverage(kat_gff):
    with Oracle(kat_gff) as o:
        d = o.depth(bed_a(o))
        assert d == (b"id\tchr\tstart\tend\tdepth\n" + b"".join(sorted([
            b"g1\tchr1\t100\t200\t2\n", b"m1\tchr1\t100\t200\t2\n", b"e1\tchr1\t100\t150\t1\n", b"e2\tchr1\t180\t200\t2\n",
            b"g2\tchr1\t300\t400\t1\n", b"e3\tchr1\t300\t400\t1\n"])))
        c = o.coverage(bed_a(o))
        assert c == (b"id\tchr\tstart\tend\tbreadth\tfraction\n" + b"".join(sorted([
            b"g1\tchr1\t100\t200\t100\t1.000000\n", b"m1\tchr1\t100\t200\t100\t1.000000\n", b"e1\tchr1\t100\t150\t50\t1.000000\n",
            b"e2\tchr1\t180\t200\t20\t1.000000\n", b"g2\tchr1\t300\t400\t10\t0.100000\n", b"e3\tchr1\t300\t400\t10\t0.100000\n"])))


def pair_checksum(p):
    mix = np.uint64(0x9E3779B97F4A7C15)
    with np.errstate(over="ignore"):
        x = p[:, 0].astype(np.uint64) * mix + p[:, 1].astype(np.uint64) * np.uint64(0xBF58476D1CE4E5B9) + p[:, 2].astype(np.uint64) * np.uint64(0x94D049BB133111EB)
        x ^= x >> np.uint64(31)
        x *= mix
        return int(x.sum(dtype=np.uint64))


GOLD = json.load(open(os.path.join(GOLDEN, "ps", "golden.json")))


@pytest.mark.parametrize("case", sorted(GOLD["cases"]))
def test_oracle_on_reference_index_files(ps_index, case):
    gff, bed = ps_index
    mode = {"contained": 0, "contains_region": 1, "overlap": 2}[case.split("+")[0]]
    invert = case.endswith("+invert")
    with Oracle(gff) as o:
        assert o.n_roots == GOLD["n_roots_total"]
        regs = o.parse_bed_intersect(bed)
        assert regs[0].shape[0] == GOLD["n_regions"]
        p = o.query_features(regs, mode, invert)
        want = GOLD["cases"][case]
        assert p.shape[0] == want["pairs"]
        assert len(np.unique(p[:, 0])) == want["roots"]
        assert pair_checksum(p) == want["checksum"]


def test_oracle_tree_build_reproduces_reference_rit_bytes(ps_index):
    """Reference-written bytes as the golden vector: the restated IntervalTree::new (src/utils/tree.rs:30-64) + bincode2
    encoder (src/utils/tree_io.rs:37-54), fed the decoded intervals of the reference's own ps.gff.rit, must write the
    very same file (also true for benchmark/at and benchmark/dm, checked when /root/reference is present)."""
    gff, _ = ps_index
    with Oracle(gff) as o:
        assert o.reencode_rit() == open(gff + ".rit", "rb").read()
    for ds in ("at", "dm"):
        ref = f"/root/reference/benchmark/{ds}.gff"
        if os.path.exists(ref + ".rit") and os.path.exists(ref + ".gof"):
            with Oracle(ref) as o:
                assert o.reencode_rit() == open(ref + ".rit", "rb").read()
