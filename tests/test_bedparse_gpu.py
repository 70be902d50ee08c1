"""BED text parsed on the device (K9, bedparse.cuh) must give exactly the arrays of the host parsers,
which restate parse_bed_file (src/commands/intersect.rs:201-230) and the depth / coverage dialect
(src/commands/depth.rs:449-495, src/commands/coverage.rs:222-255) and are themselves checked against the oracle's
parsers in tests/test_host.py.  Same order, same skips, same abort."""
import numpy as np
import pytest

from tests._oracle import Oracle

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def quirk(gffx, synth_quirk_gff):
    idx = gffx.TreeIndexData.load_tree_index(synth_quirk_gff, device=0)
    o = Oracle(synth_quirk_gff)
    yield idx, o
    o.close()
    idx.close()


def dev_arrays(t3):
    return tuple(t.cpu().numpy().view(np.uint32) for t in t3)


def same(a3, b3):
    return all(np.array_equal(np.asarray(a, dtype=np.uint32), np.asarray(b, dtype=np.uint32)) for a, b in zip(a3, b3))


def plain_bed(idx, path, n, seed, cols6=False, crlf=False):
    regs = idx.synth_regions(seed, n, 20000)
    names = idx.num_to_seqid
    eol = "\r\n" if crlf else "\n"
    with open(path, "w", newline="") as f:
        for c, s, e in zip(*regs):
            f.write(f"{names[c]}\t{s}\t{e}" + ("\tr\t0\t+" if cols6 else "") + eol)
    return regs


QUIRK_LINES = [
    "track name=x description=\"y\"",            # header: unknown chromosome in dialect 0, bad numbers in 1
    "browser position chr1:1-2",
    "# comment\t1\t2",
    "",
    "   ",
    "chr1",                                        # < 3 fields
    "chr1\t5",
    "chr1 7 9",                                    # spaces as separators
    "chr1\t\t11\t\t13",                            # empty fields dropped
    "  chr1\t15\t17",                              # leading white space
    "chr1\t30\t20",                                # start >= end: kept by intersect, dropped by counts
    "chr1\t40\t40",
    "chrNope\tzz\tyy",                             # unknown chromosome with garbage numbers
    "chrNope\t1\t2",
    "chr2\t0\t4294967295",                         # u32::MAX
    "chr2\t007\t0009",                             # leading zeros
    "chr2\t1\t2\textra\tcolumns\there",
    "chr1\t50\t60\x0c",                            # form feed: white space for dialect 0, trimmed from field 3 in dialect 1
    "chr1\t70\t80 ",                               # trailing blank
    "chr1\x0c90\x0c100",                           # form feed as the only separator: fields in dialect 0, one field in dialect 1
    "chr1\t110\t120\r",
]
COUNTS_ONLY_LINES = [                              # dialect 0 would abort on these; dialect 1 drops or accepts them
    "chr1\t+5\t+9",                                # Rust's str::parse::<u32> accepts a leading '+'
    "chr1\t-5\t9",
    "chr1\t5x\t9",
    "chr1\t5\t4294967296",                         # overflow
    "chr1\t00000000005\t00000000009",              # 11 digits, value fits
    "chr1\t5\t9 ",                            # non-ASCII white space after field 3: trim_end strips it
    "chré\t5\t9",                             # non-ASCII name
    "chr1\t1.5\t9",
]


@pytest.mark.parametrize("dialect", [0, 1])
@pytest.mark.parametrize("variant", ["3col", "6col", "crlf", "no_final_newline", "empty", "only_newlines"])
def test_plain_files(quirk, tmp_path, dialect, variant):
    idx, _ = quirk
    bed = tmp_path / "a.bed"
    if variant == "empty":
        bed.write_bytes(b"")
    elif variant == "only_newlines":
        bed.write_bytes(b"\n\n\n")
    else:
        plain_bed(idx, bed, 20011, 3, cols6=variant == "6col", crlf=variant == "crlf")
        if variant == "no_final_newline":
            bed.write_bytes(bed.read_bytes().rstrip(b"\n"))
    host = (idx.parse_bed_file if dialect == 0 else idx.parse_bed_counts)(str(bed))
    got = dev_arrays(idx.parse_bed_device(str(bed), dialect))
    assert same(got, host)
    if variant in ("3col", "6col"):
        assert idx.bed_host_lines == 0          # every line decided on the device
        assert len(got[0]) == 20011 and idx.bed_parse_ms > 0


@pytest.mark.parametrize("dialect", [0, 1])
def test_quirk_lines_match_the_host_parser_and_the_oracle(quirk, tmp_path, dialect):
    idx, o = quirk
    names = idx.num_to_seqid
    lines = [l.replace("chr1", names[0]).replace("chr2", names[min(1, len(names) - 1)]) for l in QUIRK_LINES]
    if dialect == 1:
        lines += [l.replace("chr1", names[0]) for l in COUNTS_ONLY_LINES]
    rng = np.random.default_rng(1)
    body = []
    for rep in range(40):                        # quirk lines scattered between plain ones, several tiles of text
        for l in lines:
            body.append(l)
            for _ in range(int(rng.integers(0, 60))):
                body.append(f"{names[int(rng.integers(0, len(names)))]}\t{int(rng.integers(0, 1 << 27))}\t{int(rng.integers(0, 1 << 27))}")
    bed = tmp_path / "q.bed"
    bed.write_bytes(("\n".join(body) + "\n").encode("utf-8"))
    host = (idx.parse_bed_file if dialect == 0 else idx.parse_bed_counts)(str(bed))
    want = (o.parse_bed_intersect if dialect == 0 else o.parse_bed_counts)(str(bed))
    got = dev_arrays(idx.parse_bed_device(str(bed), dialect))
    assert same(host, want)
    assert same(got, want)
    assert len(got[0]) > 1000


@pytest.mark.parametrize("bad", ["chr1\t5x\t9", "chr1\t-1\t9", "chr1\t5\t4294967296", "chr1\t\xff\t9"])
def test_intersect_dialect_aborts_like_the_host_parser(gffx, quirk, tmp_path, bad):
    idx, _ = quirk
    names = idx.num_to_seqid
    bed = tmp_path / "bad.bed"
    good = f"{names[0]}\t1\t2\n"
    bed.write_bytes((good * 300 + bad.replace("chr1", names[0]) + "\n" + good * 300).encode("latin-1"))
    with pytest.raises(gffx.GffxError) as eh:
        idx.parse_bed_file(str(bed))
    with pytest.raises(gffx.GffxError) as ed:
        idx.parse_bed_device(str(bed), 0)
    assert eh.value.code == ed.value.code and eh.value.msg == ed.value.msg
    # on an unknown chromosome the same garbage is skipped silently — except invalid UTF-8, which aborts before the look-up
    bed.write_bytes((good + bad.replace("chr1", "chrNope") + "\n" + good).encode("latin-1"))
    if "\xff" in bad:
        for fn in (lambda: idx.parse_bed_file(str(bed)), lambda: idx.parse_bed_device(str(bed), 0)):
            with pytest.raises(gffx.GffxError) as ex:
                fn()
            assert "utf-8" in ex.value.msg
    else:
        assert same(dev_arrays(idx.parse_bed_device(str(bed), 0)), idx.parse_bed_file(str(bed)))


def test_parsed_regions_feed_intersect_and_depth_directly(quirk, tmp_path):
    idx, o = quirk
    bed = tmp_path / "r.bed"
    plain_bed(idx, bed, 30000, 9)
    host = idx.parse_bed_file(str(bed))
    dev = idx.parse_bed_device(str(bed), 0)
    a = idx.intersect(host, 2, False, False, "exon")
    b = idx.intersect(dev, 2, False, False, "exon")
    assert a.n_pairs == b.n_pairs and np.array_equal(a.ranges, b.ranges)
    devc = idx.parse_bed_device(str(bed), 1)
    assert idx.depth(devc).to_tsv() == o.depth(o.parse_bed_counts(str(bed)))


def test_cli_runs_give_the_same_files_with_device_and_host_parsing(gffx, quirk, synth_quirk_gff, tmp_path, monkeypatch):
    idx, _ = quirk
    bed = tmp_path / "c.bed"
    plain_bed(idx, bed, 5000, 21, cols6=True)
    outs = {}
    for host in (False, True):
        if host:
            monkeypatch.setenv("GFFX_HOST_BED", "1")
        o1, o2 = tmp_path / f"i{int(host)}.gff", tmp_path / f"d{int(host)}.tsv"
        gffx.run_intersect(gffx.IntersectArgs(input=synth_quirk_gff, bed=str(bed), output=str(o1)))
        gffx.run_depth(gffx.DepthArgs(input=synth_quirk_gff, source=str(bed), output=str(o2)))
        outs[host] = (o1.read_bytes(), sorted(o2.read_bytes().splitlines()))
    assert outs[False] == outs[True] and len(outs[False][0]) > 0
