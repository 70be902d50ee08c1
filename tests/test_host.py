"""Host-side pieces of the product (index loaders, BED dialects, region parsing, emission) against
the oracle's restatement of the reference loaders — CPU only (index opened with device=-1)."""
import os

import numpy as np
import pytest

from tests._oracle import Oracle, OracleError


def test_loaders_on_reference_index_files(gffx, ps_index):
    """.sqs/.rix/.rit/.gof written by the reference's own `gffx index` (old pretty-printed .rix, zero .gof padding)."""
    gff, bed = ps_index
    idx = gffx.TreeIndexData.load_tree_index(gff, device=-1)
    with Oracle(gff) as o:
        assert idx.num_to_seqid == o.seq_names
        assert idx.n_roots == o.n_roots == 20213
        a = idx.parse_bed_file(bed)
        b = o.parse_bed_intersect(bed)
        for x, y in zip(a, b):
            assert np.array_equal(x, y)
        assert a[0].shape[0] == 100000
    assert idx.n_lines == 0  # the GFF text is absent: join-only use


def test_index_of_synthetic_gff(gffx, synth_quirk_gff):
    idx = gffx.TreeIndexData.load_tree_index(synth_quirk_gff, device=-1)
    with Oracle(synth_quirk_gff) as o:
        assert idx.num_to_seqid == o.seq_names
        assert idx.n_roots == o.n_roots
    # duplicate root ids: fewer reachable blocks than tree intervals
    assert idx.n_blocks < idx.n_roots
    assert idx.n_lines > 80000 and idx.n_ids > 50000


BED_QUIRKS = (
    b"track name=foo\n"
    b"# comment\n"
    b"\n"
    b"chr1\t100\t200\n"
    b"chr1 300 250 extra cols\n"          # start >= end: kept by intersect, dropped by depth/coverage
    b"chr1\t5\t5\n"
    b"chrUnknown\tabc\tdef\n"             # unknown chromosome: skipped before numbers are parsed
    b"chr2\t7\t90\tname\t0\t+\n"
    b"chr2\t10\n"                          # < 3 fields
    b"chr1\t  12\t 40\r\n"                # CRLF + repeated separators
    b"chrX\t1000\t2000"                    # no trailing newline
)


def test_bed_dialects(gffx, synth_plain_gff, tmp_path):
    bed = tmp_path / "q.bed"
    bed.write_bytes(BED_QUIRKS)
    idx = gffx.TreeIndexData.load_tree_index(synth_plain_gff, device=-1)
    with Oracle(synth_plain_gff) as o:
        for mine, theirs in ((idx.parse_bed_file, o.parse_bed_intersect), (idx.parse_bed_counts, o.parse_bed_counts)):
            a, b = mine(str(bed)), theirs(str(bed))
            for x, y in zip(a, b):
                assert np.array_equal(x, y)
        assert idx.parse_bed_file(str(bed))[0].shape[0] == 6
        assert idx.parse_bed_counts(str(bed))[0].shape[0] == 4


@pytest.mark.parametrize("bad", [b"chr1\t-5\t10\n", b"chr1\t1e3\t2000\n", b"chr1\t1\t99999999999\n", b"chr1\tx\t7\n"])
def test_bed_number_error_aborts_intersect_dialect(gffx, synth_plain_gff, tmp_path, bad):
    """src/commands/intersect.rs:223-224: a bad number on a KNOWN chromosome aborts; depth/coverage drop the line."""
    bed = tmp_path / "bad.bed"
    bed.write_bytes(b"chr1\t1\t2\n" + bad)
    idx = gffx.TreeIndexData.load_tree_index(synth_plain_gff, device=-1)
    with pytest.raises(gffx.GffxError) as ei:
        idx.parse_bed_file(str(bed))
    assert ei.value.code == -3
    with Oracle(synth_plain_gff) as o:
        with pytest.raises(OracleError):
            o.parse_bed_intersect(str(bed))
        a, b = idx.parse_bed_counts(str(bed)), o.parse_bed_counts(str(bed))
        assert a[0].shape[0] == b[0].shape[0] == 1


def test_parse_region(gffx, synth_plain_gff):
    idx = gffx.TreeIndexData.load_tree_index(synth_plain_gff, device=-1)
    num = idx.seqid_to_num
    assert idx.parse_region("chr2:100-2000") == (num["chr2"], 100, 2000)
    for bad in ("chr2", "chr2:100", "chr2:200-100", "nochr:1-2", "chr2:a-b"):
        with pytest.raises(gffx.GffxError):
            idx.parse_region(bad)


def test_write_ranges(gffx, kat_gff, tmp_path):
    idx = gffx.TreeIndexData.load_tree_index(kat_gff, device=-1)
    text = open(kat_gff, "rb").read()
    out = tmp_path / "o.gff"
    r = np.array([[16, 40], [40, 60], [100, 100], [200, len(text)], [5, len(text) + 9]], dtype=np.uint64)
    idx.write_ranges(r, str(out))
    assert out.read_bytes() == text[16:60] + text[200:]
    idx.write_ranges(np.zeros((0, 2), dtype=np.uint64), str(out))
    assert out.read_bytes() == b""


def test_synth_regions_are_a_pure_function_of_seed_and_index(gffx, synth_plain_gff):
    idx = gffx.TreeIndexData.load_tree_index(synth_plain_gff, device=-1)
    a = idx.synth_regions(42, 1000, 20000)
    b = idx.synth_regions(42, 400, 20000, first=600)
    for x, y in zip(a, b):
        assert np.array_equal(x[600:], y)
    assert np.all(a[2] - a[1] == 20000) and a[0].max() < idx.n_seqs
    assert len(np.unique(a[0])) > 10


@pytest.mark.parametrize("fixture", ["synth_quirk_gff", "synth_plain_gff", "kat_gff"])
def test_index_builder_is_byte_compatible(gffx, request, fixture, tmp_path):
    """gffx_index_build (src/index_builder/core.rs:41-242) writes the same bytes as the oracle's restatement."""
    import shutil

    src = request.getfixturevalue(fixture)
    dst = str(tmp_path / "copy.gff")
    shutil.copy(src, dst)
    gffx.build_index(dst)
    for ext in ("sqs", "gof", "rit", "rix", "fts", "prt"):
        assert open(dst + "." + ext, "rb").read() == open(src + "." + ext, "rb").read(), ext
    # attribute table: header + distinct values in first-seen order; a2f = one u32 per indexed line
    gffx.build_index(dst, attribute="gene")
    atn = open(dst + ".atn", "rb").read().split(b"\n")
    assert atn[0] == b"#attribute=gene"
    a2f = np.fromfile(dst + ".a2f", dtype="<u4")
    assert a2f.shape[0] == np.fromfile(dst + ".prt", dtype="<u4").shape[0]
    used = a2f[a2f != 0xFFFFFFFF]
    assert (len(atn) - 2 == (used.max() + 1 if len(used) else 0))


def test_index_builder_errors(gffx, tmp_path):
    p = tmp_path / "bad.gff"
    p.write_bytes(b"chr1\tsrc\tgene\t1\t9\t.\t+\t.\n")  # 8 columns
    with pytest.raises(gffx.GffxError):
        gffx.build_index(str(p))
    p.write_bytes(b"chr1\tsrc\tgene\t1\t9\t.\t+\t.\tName=x\n")  # no ID
    with pytest.raises(gffx.GffxError):
        gffx.build_index(str(p))
    with pytest.raises(gffx.GffxError):
        gffx.build_index(str(tmp_path / "missing.gff"))


def test_id_dictionary_is_numbered_by_first_appearance(gffx, synth_plain_gff):
    """The depth / coverage view numbers the ID attribute values in order of first appearance in the GFF (the order a serial
    scan interns them in); the loader builds that dictionary on all host threads from per-line hashes."""
    idx = gffx.TreeIndexData.load_tree_index(synth_plain_gff, device=-1)
    seen, order = set(), []
    with open(synth_plain_gff, "rb") as f:
        for line in f:
            if line.startswith(b"#"):
                continue
            cols = line.rstrip(b"\r\n").split(b"\t")
            if len(cols) < 9:
                continue
            # `region` lines are skipped at index build: the one in front of the first root lies in no block; the later ones
            # sit in the tail of the preceding root's block and count
            if not seen and cols[2] == b"region":
                continue
            for kv in cols[8].split(b";"):
                if kv.startswith(b"ID="):
                    v = kv[3:]
                    if v not in seen:
                        seen.add(v)
                        order.append(v)
                    break
    assert idx.n_ids == len(order)
    assert len(order) < idx.n_lines   # multi-line CDS features share one ID: the dictionary really dedups
    got = [idx.id_string(i) for i in range(idx.n_ids)]
    assert got == order
