"""The root-parallel join kernel (gffx_b200/csrc/rootjoin.cuh) against the oracle, forced on at every job size, with the
tail threshold moved so that each of its parts carries the job in turn: prefix-sum bins + exact boundary bins (normal
part), the region-parallel tail (inverted / empty / long regions), and roots with end <= start (every region tested).
The region-parallel kernel (winjoin.cuh) is forced for the same inputs: both must equal the oracle, hence each other."""
import os

import numpy as np
import pytest

from tests._oracle import Oracle, sorted_pairs
from tests.test_intersect_gpu import MODES, mixed_regions

pytestmark = pytest.mark.gpu


def _open(gffx, path, kernel, tail_bins):
    old = {k: os.environ.get(k) for k in ("GFFX_WJ_KERNEL", "GFFX_WR_TAIL_BINS")}
    os.environ["GFFX_WJ_KERNEL"] = kernel
    if tail_bins is not None:
        os.environ["GFFX_WR_TAIL_BINS"] = str(tail_bins)
    try:
        return gffx.TreeIndexData.load_tree_index(path, device=0)   # the engine reads both when it is built
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


@pytest.fixture(scope="module")
def oracle_quirk(synth_quirk_gff):
    o = Oracle(synth_quirk_gff)
    yield o
    o.close()


@pytest.mark.parametrize("kernel,tail_bins", [("roots", None), ("roots", 1024), ("roots", 0), ("roots", 1 << 20), ("rounds", None)])
def test_forced_kernel_equals_oracle(gffx, synth_quirk_gff, oracle_quirk, kernel, tail_bins):
    idx = _open(gffx, synth_quirk_gff, kernel, tail_bins)
    try:
        for mode, invert in MODES:
            regs = mixed_regions(idx, 300 + mode, 6000)
            r = idx.query_features(regs, mode, invert)
            want = oracle_quirk.query_features(regs, mode, invert)
            assert r.n_pairs == want.shape[0], (kernel, tail_bins, mode, invert)
            got = r.pairs if r.pairs is not None else np.zeros((0, 3), dtype=np.uint32)
            assert np.array_equal(sorted_pairs(got), sorted_pairs(want)), (kernel, tail_bins, mode, invert)
        for mode, invert, entire, types in [(2, False, False, None), (0, False, False, "gene,mRNA,exon"), (1, True, True, None), (0, True, False, None)]:
            regs = mixed_regions(idx, 17 + mode, 3000)
            r = idx.intersect(regs, mode, invert, entire, types)
            want, n_pairs, n_warn = oracle_quirk.intersect(regs, mode, invert, entire, types)
            assert r.n_pairs == n_pairs and np.array_equal(r.ranges, want) and len(r.warn_fids) == n_warn
    finally:
        idx.close()


@pytest.mark.parametrize("length", [150, 2000, 20000, 400000])
def test_single_window_tiles(gffx, synth_plain_gff, length):
    """enough regions for whole tiles inside one window (the bench's shape), each length class on its own"""
    o = Oracle(synth_plain_gff)
    idx = _open(gffx, synth_plain_gff, "roots", None)
    try:
        regs = idx.synth_regions(900 + length, 400_000, length)
        sample = tuple(a[:20000] for a in regs)   # the oracle walks its tree per region: keep its share small
        for mode, invert in MODES:
            r = idx.query_features(regs, mode, invert, want_pairs=False)
            # additivity pins the large job to the oracle-checked small ones: pairs(all) = sum over chunks
            tot = 0
            for k in range(0, 400_000, 100_000):
                tot += idx.query_features(tuple(a[k:k + 100_000] for a in regs), mode, invert, want_pairs=False).n_pairs
            assert r.n_pairs == tot
            rs = idx.query_features(sample, mode, invert)
            want = o.query_features(sample, mode, invert)
            got = rs.pairs if rs.pairs is not None else np.zeros((0, 3), dtype=np.uint32)
            assert np.array_equal(sorted_pairs(got), sorted_pairs(want)), (length, mode, invert)
    finally:
        idx.close()
        o.close()


def test_roots_and_rounds_agree_at_scale(gffx, synth_plain_gff):
    a = _open(gffx, synth_plain_gff, "roots", None)
    b = _open(gffx, synth_plain_gff, "rounds", None)
    try:
        regs = mixed_regions(a, 4242, 300_000)
        for mode, invert in MODES:
            ra = a.intersect(regs, mode, invert, False, None, want_pairs=True)
            rb = b.intersect(regs, mode, invert, False, None, want_pairs=True)
            assert ra.n_pairs == rb.n_pairs and ra.n_roots == rb.n_roots
            assert np.array_equal(ra.ranges, rb.ranges)
            pa = ra.pairs if ra.pairs is not None else np.zeros((0, 3), dtype=np.uint32)
            pb = rb.pairs if rb.pairs is not None else np.zeros((0, 3), dtype=np.uint32)
            assert np.array_equal(sorted_pairs(pa), sorted_pairs(pb))
    finally:
        a.close()
        b.close()
