"""Record emission on the device (K8, gather.cuh): the gathered text must be byte for byte what the
reference's write loops produce from the same ranges (src/commands/intersect.rs:385-429, src/utils/common.rs:231-278):
the concatenation of text[start:end] in range order, skipping start >= end and end > file length."""
import zlib

import numpy as np
import pytest

from tests._oracle import Oracle
from tests.test_intersect_gpu import MODES, mixed_regions

pytestmark = pytest.mark.gpu


def expected_text(gff: bytes, ranges: np.ndarray) -> bytes:
    out = []
    for s, e in np.asarray(ranges, dtype=np.uint64).reshape(-1, 2).tolist():
        if s < e <= len(gff):
            out.append(gff[s:e])
    return b"".join(out)


@pytest.fixture(scope="module")
def quirk(gffx, synth_quirk_gff):
    idx = gffx.TreeIndexData.load_tree_index(synth_quirk_gff, device=0)
    o = Oracle(synth_quirk_gff)
    gff = open(synth_quirk_gff, "rb").read()
    yield idx, o, gff
    o.close()
    idx.close()


@pytest.mark.parametrize("mode,invert", MODES)
@pytest.mark.parametrize("entire,types", [(False, None), (True, None), (False, "gene,mRNA,exon")])
def test_gather_last_result_matches_oracle_ranges(quirk, mode, invert, entire, types):
    """ranges still on the device (ranges=None) -> text == slices of the ORACLE's ranges"""
    idx, o, gff = quirk
    regs = mixed_regions(idx, 300 + mode, 3000)
    r = idx.intersect(regs, mode, invert, entire, types)
    want_ranges, _, _ = o.intersect(regs, mode, invert, entire, types)
    got = idx.gather_text(None, r.n_ranges)
    assert got == expected_text(gff, want_ranges)
    assert len(got) == r.bytes_out


def test_gather_explicit_ranges_edge_cases(quirk):
    idx, _, gff = quirk
    n = len(gff)
    rng = np.random.default_rng(5)
    cases = [
        np.zeros((0, 2), dtype=np.uint64),                                   # nothing
        np.array([[0, n]], dtype=np.uint64),                                 # the whole file as one range
        np.array([[5, 5], [9, 3], [0, n + 1], [n - 1, n]], dtype=np.uint64),  # empty, inverted, past the end, last byte
        np.array([[0, 1], [1, 2], [2, 3], [3, 7], [7, 8], [100, 101], [101, 4000]], dtype=np.uint64),  # tiny contiguous runs
    ]
    # 5000 random ranges, all lengths and alignments, overlapping and out of order (the kernel must not assume either)
    s = rng.integers(0, n, size=5000)
    ln = rng.integers(0, 700, size=5000)
    cases.append(np.stack([s, np.minimum(s + ln, n + 3)], axis=1).astype(np.uint64))
    # every line once, in file order (source-contiguous: fused into long runs), then reversed (no two contiguous)
    nl = np.flatnonzero(np.frombuffer(gff, dtype=np.uint8) == 10).astype(np.uint64) + 1
    starts = np.concatenate([[0], nl[:-1]]).astype(np.uint64)
    lines = np.stack([starts, nl], axis=1)
    cases += [lines, lines[::-1].copy()]
    for ranges in cases:
        got = idx.gather_text(ranges)
        assert got == expected_text(gff, ranges), ranges[:4]


def test_gather_equals_host_writer(quirk, tmp_path):
    """same bytes as gffx_write_ranges (the host writev path) on a real result, via the device pointer variant too"""
    import torch

    idx, _, gff = quirk
    regs = mixed_regions(idx, 77, 4000)
    r = idx.intersect(regs, 2, False, False, None)
    out = tmp_path / "host.gff"
    idx.write_ranges(r.ranges, str(out))
    want = out.read_bytes()
    assert idx.gather_text(None, r.n_ranges) == want
    ptr, nb = idx.gather_text(None, r.n_ranges, device=True)
    assert nb == len(want)
    buf = torch.empty(nb, dtype=torch.uint8, device="cuda:0")
    torch.cuda.synchronize()
    import ctypes

    cudart = ctypes.CDLL("libcudart.so.12")
    assert cudart.cudaMemcpy(ctypes.c_void_p(buf.data_ptr()), ctypes.c_void_p(ptr), ctypes.c_size_t(nb), 3) == 0
    assert bytes(buf.cpu().numpy()) == want
    assert idx.gather_ms > 0


def test_cli_device_emission_matches_host_emission(gffx, quirk, synth_quirk_gff, tmp_path, monkeypatch):
    """gffx_intersect_run writes the same file whether records are gathered on the device (default) or sliced on the host"""
    idx, _, _ = quirk
    bed = tmp_path / "q.bed"
    regs = idx.synth_regions(9, 2000, 20000)
    names = idx.num_to_seqid
    with open(bed, "w") as f:
        for c, s, e in zip(*regs):
            f.write(f"{names[c]}\t{s}\t{e}\n")
    outs = []
    for host in (False, True):
        if host:
            monkeypatch.setenv("GFFX_HOST_EMIT", "1")
        out = tmp_path / f"o{int(host)}.gff"
        gffx.run_intersect(gffx.IntersectArgs(input=synth_quirk_gff, bed=str(bed), output=str(out), types="exon,CDS"))
        outs.append(out.read_bytes())
    assert outs[0] == outs[1] and len(outs[0]) > 0


def test_gather_full_size_properties(gffx, tmp_path_factory):
    """60k-gene GFF (0.4 GB of text): overlap of 1 M regions emits nearly the whole file; length and CRC against the
    slices of the same ranges"""
    from tests.conftest import _make_synth

    path = _make_synth(tmp_path_factory.mktemp("big"), "big.gff", 20260001, 60000, False)
    idx = gffx.TreeIndexData.load_tree_index(path, device=0)
    try:
        gff = open(path, "rb").read()
        regs = idx.synth_regions(11, 1_000_000, 20000)
        r = idx.intersect(regs, 2, False, False, None)
        got = idx.gather_text(None, r.n_ranges)
        assert len(got) == r.bytes_out
        assert zlib.crc32(got) == zlib.crc32(expected_text(gff, r.ranges))
        # -e: merged block ranges (few, long)
        r = idx.intersect(regs, 2, False, True, None)
        got = idx.gather_text(None, r.n_ranges)
        assert len(got) == r.bytes_out and zlib.crc32(got) == zlib.crc32(expected_text(gff, r.ranges))
    finally:
        idx.close()
