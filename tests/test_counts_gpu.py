"""Parity of the CUDA depth / coverage paths (through the C ABI) with the CPU oracle.
The reference's row order is hash order: header exact + bytewise-sorted rows exact."""
import numpy as np
import pytest

from tests._oracle import Oracle

pytestmark = pytest.mark.gpu


def region_mix(idx, seed, n):
    """short reads + 20 kb windows + a few multi-Mb regions + invalid (start >= end) ones, shuffled"""
    a = idx.synth_regions(seed, n, 150)
    b = idx.synth_regions(seed + 1, n // 4, 20000)
    c = idx.synth_regions(seed + 2, max(n // 100, 3), 2_500_000)
    chr_ = np.concatenate([a[0], b[0], c[0]])
    s = np.concatenate([a[1], b[1], c[1]])
    e = np.concatenate([a[2], b[2], c[2]])
    rng = np.random.default_rng(seed)
    bad = rng.choice(len(chr_), size=len(chr_) // 40, replace=False)
    e[bad[::2]] = s[bad[::2]]
    s[bad[1::2]], e[bad[1::2]] = e[bad[1::2]].copy(), s[bad[1::2]].copy()
    # reads of varying length so that some fall inside introns, some span several exons
    ln = rng.integers(1, 4000, size=len(chr_)).astype(np.uint32)
    short = rng.random(len(chr_)) < 0.3
    e = np.where(short & (e > s), s + ln, e).astype(np.uint32)
    perm = rng.permutation(len(chr_))
    return chr_[perm].copy(), s[perm].copy(), e[perm].copy()


@pytest.fixture(scope="module", params=["quirk", "plain"])
def pair(request, gffx, synth_quirk_gff, synth_plain_gff):
    path = synth_quirk_gff if request.param == "quirk" else synth_plain_gff
    idx = gffx.TreeIndexData.load_tree_index(path, device=0)
    o = Oracle(path)
    yield idx, o, path
    o.close()
    idx.close()


def first_diff(a: bytes, b: bytes):
    la, lb = a.split(b"\n"), b.split(b"\n")
    for i, (x, y) in enumerate(zip(la, lb)):
        if x != y:
            return i, x, y
    return min(len(la), len(lb)), None, None


@pytest.mark.parametrize("seed,n", [(1, 20000), (2, 3000)])
def test_depth(pair, seed, n):
    idx, o, _ = pair
    regs = region_mix(idx, seed, n)
    got = idx.depth(regs).to_tsv()
    want = o.depth(regs)
    assert got == want, first_diff(got, want)
    assert got.count(b"\n") > 1000


@pytest.mark.parametrize("seed,n", [(3, 20000), (4, 3000)])
def test_coverage(pair, seed, n):
    idx, o, _ = pair
    regs = region_mix(idx, seed, n)
    got = idx.coverage(regs).to_tsv()
    want = o.coverage(regs)
    assert got == want, first_diff(got, want)
    assert got.count(b"\n") > 1000


def test_batches_accumulate(pair):
    """The reference feeds depth in batches of 100k BED lines (depth.rs:24,462); any batching gives the same rows."""
    idx, o, _ = pair
    regs = region_mix(idx, 9, 8000)
    one = idx.depth(regs).to_tsv()
    many = idx.depth(regs, batches=7).to_tsv()
    assert one == many == o.depth(regs)
    assert idx.coverage(regs, batches=5).to_tsv() == o.coverage(regs)


def test_empty_and_all_invalid(pair):
    idx, o, _ = pair
    z = np.zeros(0, dtype=np.uint32)
    assert idx.depth((z, z, z)).to_tsv() == b"id\tchr\tstart\tend\tdepth\n"
    assert idx.coverage((z, z, z)).to_tsv() == b"id\tchr\tstart\tend\tbreadth\tfraction\n"
    c = np.zeros(5, dtype=np.uint32)
    s = np.full(5, 1000, dtype=np.uint32)
    assert idx.depth((c, s, s)).to_tsv() == o.depth((c, s, s)) == b"id\tchr\tstart\tend\tdepth\n"


def test_whole_chromosome_regions(pair):
    """Regions that cover everything: every ID of every routed root is hit exactly once per region."""
    idx, o, _ = pair
    n = idx.n_seqs
    c = np.arange(n, dtype=np.uint32).repeat(2)
    s = np.zeros(2 * n, dtype=np.uint32)
    e = np.full(2 * n, 0xFFFFFFFF, dtype=np.uint32)
    assert idx.depth((c, s, e)).to_tsv() == o.depth((c, s, e))
    assert idx.coverage((c, s, e)).to_tsv() == o.coverage((c, s, e))


def test_kat(gffx, kat_gff):
    idx = gffx.TreeIndexData.load_tree_index(kat_gff, device=0)
    num = idx.seqid_to_num
    regs = (np.array([num["chr1"]] * 3, dtype=np.uint32), np.array([150, 200, 90], dtype=np.uint32), np.array([190, 310, 250], dtype=np.uint32))
    d = idx.depth(regs).to_tsv(sort_rows=False)
    assert d == (b"id\tchr\tstart\tend\tdepth\n" b"g1\tchr1\t100\t200\t2\n" b"m1\tchr1\t100\t200\t2\n" b"e1\tchr1\t100\t150\t1\n"
                 b"e2\tchr1\t180\t200\t2\n" b"g2\tchr1\t300\t400\t1\n" b"e3\tchr1\t300\t400\t1\n")
    c = idx.coverage(regs).to_tsv(sort_rows=False)
    assert c == (b"id\tchr\tstart\tend\tbreadth\tfraction\n" b"g1\tchr1\t100\t200\t100\t1.000000\n" b"m1\tchr1\t100\t200\t100\t1.000000\n"
                 b"e1\tchr1\t100\t150\t50\t1.000000\n" b"e2\tchr1\t180\t200\t20\t1.000000\n" b"g2\tchr1\t300\t400\t10\t0.100000\n"
                 b"e3\tchr1\t300\t400\t10\t0.100000\n")
    idx.close()


def test_cli_depth_and_coverage_files(gffx, pair, tmp_path):
    """gffx_depth_run / gffx_coverage_run (the reference's run(), depth.rs:548-635, coverage.rs:493-582): BED in, TSV out."""
    idx, o, path = pair
    regs = region_mix(idx, 21, 3000)
    names = idx.num_to_seqid
    bed = tmp_path / "reads.bed"
    with open(bed, "w") as f:
        f.write("# reads\n")
        for c, s, e in zip(*regs):
            f.write(f"{names[c]}\t{s}\t{e}\n")
        f.write("chrNope\t1\t2\nchr1\tx\t9\n")
    parsed = o.parse_bed_counts(str(bed))
    for fn, ofn in ((gffx.run_depth, o.depth), (gffx.run_coverage, o.coverage)):
        out = tmp_path / "out.tsv"
        fn(gffx.DepthArgs(input=path, source=str(bed), output=str(out)))
        rows = out.read_bytes().split(b"\n")
        got = rows[0] + b"\n" + b"".join(r + b"\n" for r in sorted(rows[1:-1]))
        assert got == ofn(parsed)
