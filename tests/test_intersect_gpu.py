"""Parity of the CUDA intersect path (through the C ABI) with the CPU oracle: root-level join pairs,
selected byte ranges for every mode x invert x -e x -T combination, quirk inputs, edge cases."""
import json
import os

import numpy as np
import pytest

from tests._oracle import Oracle, sorted_pairs
from tests.conftest import GOLDEN
from tests.test_oracle_kat import pair_checksum

pytestmark = pytest.mark.gpu

MODES = [(0, False), (0, True), (1, False), (1, True), (2, False), (2, True)]


def mixed_regions(idx, seed, n):
    """20 kb windows + short reads + a few degenerate regions (start >= end, beyond every root, huge)."""
    a = idx.synth_regions(seed, n, 20000)
    b = idx.synth_regions(seed + 1, n, 150)
    c = idx.synth_regions(seed + 2, max(n // 10, 4), 3_000_000)
    chr_ = np.concatenate([a[0], b[0], c[0]])
    s = np.concatenate([a[1], b[1], c[1]])
    e = np.concatenate([a[2], b[2], c[2]])
    k = len(chr_) // 50
    rng = np.random.default_rng(seed)
    pick = rng.choice(len(chr_), size=k, replace=False)
    s[pick[: k // 2]], e[pick[: k // 2]] = e[pick[: k // 2]].copy(), s[pick[: k // 2]].copy()  # inverted (BED keeps them)
    e[pick[k // 2:]] = s[pick[k // 2:]]                                                     # empty
    s[pick[:2]] = 0xFFFFFF00
    e[pick[:2]] = 0xFFFFFFFF
    perm = rng.permutation(len(chr_))
    return chr_[perm].copy(), s[perm].copy(), e[perm].copy()


@pytest.fixture(scope="module")
def quirk(gffx, synth_quirk_gff):
    idx = gffx.TreeIndexData.load_tree_index(synth_quirk_gff, device=0)
    o = Oracle(synth_quirk_gff)
    yield idx, o, synth_quirk_gff
    o.close()
    idx.close()


@pytest.mark.parametrize("mode,invert", MODES)
def test_query_features_pairs(quirk, mode, invert):
    idx, o, _ = quirk
    regs = mixed_regions(idx, 100 + mode, 5000)
    r = idx.query_features(regs, mode, invert)
    want = o.query_features(regs, mode, invert)
    assert r.n_pairs == want.shape[0]
    got = r.pairs if r.pairs is not None else np.zeros((0, 3), dtype=np.uint32)
    assert np.array_equal(sorted_pairs(got), sorted_pairs(want))
    assert r.n_roots == len(np.unique(want[:, 0]))
    assert r.launches > 0 or (mode == 2 and invert)


@pytest.mark.parametrize("mode,invert", MODES)
@pytest.mark.parametrize("entire,types", [(False, None), (True, None), (False, "gene,mRNA,exon"), (True, "exon, CDS ,,nosuchtype")])
def test_intersect_ranges(quirk, mode, invert, entire, types):
    idx, o, path = quirk
    regs = mixed_regions(idx, 7 + mode, 3000)
    r = idx.intersect(regs, mode, invert, entire, types)
    want, n_pairs, n_warn = o.intersect(regs, mode, invert, entire, types)
    assert r.n_pairs == n_pairs
    assert r.n_ranges == want.shape[0]
    assert np.array_equal(r.ranges, want)
    assert r.bytes_out == int((want[:, 1] - want[:, 0]).sum())
    assert len(r.warn_fids) == n_warn


def test_few_regions_and_single_chromosome(quirk):
    idx, o, _ = quirk
    for regs in [(np.array([3], dtype=np.uint32), np.array([0], dtype=np.uint32), np.array([0xFFFFFFFF], dtype=np.uint32)),
                 (np.array([0, 0], dtype=np.uint32), np.array([5, 5], dtype=np.uint32), np.array([5, 4], dtype=np.uint32))]:
        for mode, invert in MODES:
            r = idx.intersect(regs, mode, invert, False, None, want_pairs=True)
            want, n_pairs, _ = o.intersect(regs, mode, invert, False, None)
            assert r.n_pairs == n_pairs and np.array_equal(r.ranges, want)


def test_empty_input(quirk):
    idx, _, _ = quirk
    z = np.zeros(0, dtype=np.uint32)
    r = idx.intersect((z, z, z), 2, False, False, None, want_pairs=True)
    assert r.n_pairs == 0 and r.n_ranges == 0 and r.n_roots == 0 and r.bytes_out == 0


def test_bad_chromosome_number_is_an_error(gffx, quirk):
    idx, _, _ = quirk
    with pytest.raises(gffx.GffxError) as ei:
        idx.query_features((np.array([999], dtype=np.uint32), np.array([1], dtype=np.uint32), np.array([2], dtype=np.uint32)))
    assert ei.value.code == -1


def test_device_resident_regions_give_the_same_answer(quirk):
    import torch

    idx, o, _ = quirk
    regs = idx.synth_regions(5, 40001, 20000)  # odd length: exercises the unaligned tail
    dev = tuple(torch.from_numpy(a.view(np.int32)).cuda() for a in regs)
    a = idx.intersect(regs, 2, False, False, "exon", want_pairs=True)
    b = idx.intersect(dev, 2, False, False, "exon", want_pairs=True)
    assert a.n_pairs == b.n_pairs and np.array_equal(a.ranges, b.ranges)
    assert np.array_equal(sorted_pairs(a.pairs), sorted_pairs(b.pairs))
    # unaligned device views (offset by one element)
    dev1 = tuple(t[1:] for t in dev)
    c = idx.intersect(dev1, 2, False, False, "exon")
    want, _, _ = o.intersect(tuple(x[1:] for x in regs), 2, False, False, "exon")
    assert np.array_equal(c.ranges, want)
    # the device generator is the same function of (seed, index) as the host generator
    g = idx.synth_regions_device(5, 40001, 20000)
    for x, y in zip(g, regs):
        assert np.array_equal(x.cpu().numpy().view(np.uint32), y)


def test_reference_index_files_root_join(gffx, ps_index):
    """Root-level join on the reference's own ps index files against the brute-force golden numbers."""
    gold = json.load(open(os.path.join(GOLDEN, "ps", "golden.json")))
    gff, bed = ps_index
    idx = gffx.TreeIndexData.load_tree_index(gff, device=0)
    regs = idx.parse_bed_file(bed)
    for case, want in gold["cases"].items():
        mode = {"contained": 0, "contains_region": 1, "overlap": 2}[case.split("+")[0]]
        r = idx.query_features(regs, mode, case.endswith("+invert"))
        assert r.n_pairs == want["pairs"], case
        assert r.n_roots == want["roots"], case
        p = r.pairs if r.pairs is not None else np.zeros((0, 3), dtype=np.uint32)
        assert pair_checksum(p) == want["checksum"], case
    idx.close()


def test_cli_run_writes_the_same_file_as_the_oracle(gffx, quirk, tmp_path):
    """gffx_intersect_run (the reference's run(), src/commands/intersect.rs:541-655) end to end: BED file in, GFF out."""
    idx, o, path = quirk
    regs = idx.synth_regions(77, 2000, 20000)
    names = idx.num_to_seqid
    bed = tmp_path / "r.bed"
    with open(bed, "w") as f:
        f.write("track name=x\n")
        for c, s, e in zip(*regs):
            f.write(f"{names[c]}\t{s}\t{e}\tr\t0\t+\n")
        f.write("chrNope\tzz\tyy\n")
    text = open(path, "rb").read()
    for kw in ({}, {"contained": True}, {"contains_region": True, "invert": True}, {"entire_group": True}, {"types": "CDS"}):
        out = tmp_path / "out.gff"
        gffx.run_intersect(gffx.IntersectArgs(input=path, bed=str(bed), output=str(out), **kw))
        mode = 0 if kw.get("contained") else 1 if kw.get("contains_region") else 2
        want, _, _ = o.intersect(o.parse_bed_intersect(str(bed)), mode, kw.get("invert", False), kw.get("entire_group", False), kw.get("types"))
        assert out.read_bytes() == b"".join(text[s:e] for s, e in want.tolist())


@pytest.mark.parametrize("sub", ["0", "3"])
def test_join_coarse_bins_overflow_the_slice(gffx, synth_quirk_gff, sub, monkeypatch):
    """GFFX_WJ_SUB caps the bins per window of the K1 index (winjoin_build.h): with one or eight bins per window the
    lists outgrow both the 16 position bits of the hit masks and the shared-memory slice, so the global-memory tails
    and the second walk of the emit pass are exercised.  Same pairs, same ranges."""
    monkeypatch.setenv("GFFX_WJ_SUB", sub)
    idx = gffx.TreeIndexData.load_tree_index(synth_quirk_gff, device=0)
    try:
        with Oracle(synth_quirk_gff) as o:
            regs = mixed_regions(idx, 777, 4000)
            for mode, invert in MODES:
                r = idx.query_features(regs, mode, invert)
                want = o.query_features(regs, mode, invert)
                got = r.pairs if r.pairs is not None else np.zeros((0, 3), dtype=np.uint32)
                assert np.array_equal(sorted_pairs(got), sorted_pairs(want))
                ri = idx.intersect(regs, mode, invert, False, None)
                want_ranges, n_pairs, _ = o.intersect(regs, mode, invert, False, None)
                assert ri.n_pairs == n_pairs
                assert np.array_equal(ri.ranges, want_ranges)
    finally:
        idx.close()


def test_pair_buffer_grows_when_the_result_outruns_the_first_guess(quirk):
    """Chromosome-wide regions hit every root of their sequence: far more than the 2 pairs per region the pair buffer is
    first sized for, so the expansion (K1b) runs again into a buffer of the exact size.  Same multiset as the oracle."""
    idx, o, _ = quirk
    n_seq = len(idx.num_to_seqid)
    chr_ = np.arange(200, dtype=np.uint32) % n_seq
    regs = (chr_, np.zeros(200, dtype=np.uint32), np.full(200, 0xFFFFFFF0, dtype=np.uint32))
    for mode, invert in [(2, False), (0, False), (1, True)]:
        r = idx.query_features(regs, mode, invert)
        want = o.query_features(regs, mode, invert)
        assert r.n_pairs == want.shape[0]
        if mode != 1:
            assert r.n_pairs > 20 * 200
        got = r.pairs if r.pairs is not None else np.zeros((0, 3), dtype=np.uint32)
        assert np.array_equal(sorted_pairs(got), sorted_pairs(want))


def test_pairs_come_back_grouped_by_root_in_start_order(quirk):
    """The documented order of the pair array (include/gffx_b200.h): grouped by root, roots by (sequence, start, end)."""
    idx, _, _ = quirk
    regs = mixed_regions(idx, 4242, 3000)
    r = idx.query_features(regs, 2, False)
    p = r.pairs
    assert p is not None and len(p) > 0
    change = np.flatnonzero(np.any(p[1:] != p[:-1], axis=1)) + 1
    heads = p[np.concatenate([[0], change])]
    assert len(np.unique(heads, axis=0)) == len(heads)          # every root's copies are contiguous
    r2 = idx.query_features(regs, 2, False)
    assert np.array_equal(p, r2.pairs)                            # and the order is reproducible


def test_result_summary_in_one_call_equals_the_accessors(quirk):
    """gffx_result_get_info fills one struct with what the gffx_result_* accessors return one by one."""
    import ctypes as C

    from gffx_b200 import _lib

    idx, _, _ = quirk
    regs = mixed_regions(idx, 99, 2000)
    keep = []
    h = C.c_void_p()
    arrs = [np.ascontiguousarray(a, dtype=np.uint32) for a in regs]
    rc = _lib.lib.gffx_intersect(idx._h, arrs[0].ctypes.data, arrs[1].ctypes.data, arrs[2].ctypes.data, len(arrs[0]), 2, 0, 0, None,
                                 _lib.F_WANT_PAIRS | _lib.F_STAGE_TIMES, C.byref(h))
    assert rc == 0, _lib.last_error()
    try:
        info = _lib.ResultInfo()
        assert _lib.lib.gffx_result_get_info(h, C.byref(info)) == 0
        L = _lib.lib
        assert info.n_queries == L.gffx_result_n_queries(h) == len(arrs[0])
        assert info.n_pairs == L.gffx_result_n_pairs(h) > 0
        assert info.n_roots == L.gffx_result_n_roots(h)
        assert info.n_ranges == L.gffx_result_n_ranges(h) > 0
        assert info.bytes_out == L.gffx_result_bytes_out(h)
        assert info.n_warn == L.gffx_result_n_warn(h)
        assert C.cast(info.pairs, C.c_void_p).value == C.cast(L.gffx_result_pairs(h), C.c_void_p).value
        assert C.cast(info.ranges, C.c_void_p).value == C.cast(L.gffx_result_ranges(h), C.c_void_p).value
        assert (info.pairs_device or 0) == (L.gffx_result_pairs_device(h) or 0)
        assert info.kernel_ms == L.gffx_result_kernel_ms(h) > 0
        assert info.launches == L.gffx_result_launches(h) > 0
        assert [info.stage_ms[k] for k in range(12)] == [L.gffx_result_stage_ms(h, k) for k in range(12)]
        assert info.stage_ms[0] > 0 and info.stage_ms[8] > 0   # GFFX_F_STAGE_TIMES was asked for
        plain = idx.intersect(regs, 2, False, False, None)
        assert plain.kernel_ms > 0 and not any(plain.stage_ms)   # without the flag only the call's whole span is timed
        assert _lib.lib.gffx_result_get_info(h, None) != 0
    finally:
        _lib.lib.gffx_result_free(h)


def test_compaction_offsets_by_scan_for_very_large_line_tables(gffx, synth_quirk_gff):
    """Past 16.8 M lines the compaction kernel takes its tile offsets from a scan pass instead of adding up the tile counts
    itself; GFFX_EMIT_SCAN forces that path on a small table (read once per process: run in a child).  Same ranges."""
    import subprocess
    import sys

    code = (
        "import os, sys, numpy as np\n"
        "sys.path.insert(0, os.getcwd())\n"
        "import gffx_b200 as g\n"
        "from tests._oracle import Oracle\n"
        "from tests.test_intersect_gpu import mixed_regions\n"
        f"gff = {synth_quirk_gff!r}\n"
        "idx = g.TreeIndexData.load_tree_index(gff, device=0)\n"
        "regs = mixed_regions(idx, 31, 3000)\n"
        "with Oracle(gff) as o:\n"
        "    for mode, inv in [(2, False), (0, False), (1, True)]:\n"
        "        r = idx.intersect(regs, mode, inv, False, None)\n"
        "        want, n_pairs, _ = o.intersect(regs, mode, inv, False, None)\n"
        "        assert r.n_pairs == n_pairs and np.array_equal(r.ranges, want), (mode, inv)\n"
        "print('ok')\n"
    )
    env = dict(os.environ, GFFX_EMIT_SCAN="1")
    out = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=300,
                         cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    assert out.returncode == 0 and out.stdout.strip().endswith("ok"), out.stderr[-2000:]
