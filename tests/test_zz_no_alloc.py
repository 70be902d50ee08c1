"""gffx_count_pairs_noalloc (include/gffx_b200.h): the allocation-free join-pair count, through the C ABI, against
gffx_query_features on the same regions — exact (integer work), on two calls in a row (the second call finds the static
scratch array full of the first call's contents), for every mode, at a size that takes the region-parallel kernel and at
one that takes the root-parallel kernel; the size limit is an error, not a fault.  Collected last on purpose
(tests/test_zz_*)."""
import numpy as np
import pytest

from tests.test_intersect_gpu import MODES, mixed_regions

pytestmark = pytest.mark.gpu


def _dev(a, torch):
    return torch.from_numpy(np.ascontiguousarray(a).view(np.int32)).cuda()


def _noalloc(idx, torch, regs, mode, invert):
    c, s, e = (_dev(a, torch) for a in regs)
    out = torch.full((2,), -1, dtype=torch.int64, device="cuda")   # the call clears it itself
    torch.cuda.synchronize()
    idx.count_pairs_noalloc(c.data_ptr(), s.data_ptr(), e.data_ptr(), len(regs[0]), mode, invert, out.data_ptr())
    idx.sync()
    o = out.cpu().numpy()
    return int(o[0]), int(o[1])


@pytest.mark.parametrize("n", [3000, 150_000])
def test_noalloc_equals_query_features_twice(gffx, synth_plain_gff, n):
    import torch
    idx = gffx.TreeIndexData.load_tree_index(synth_plain_gff, device=0)
    try:
        for mode, invert in MODES:
            for rep in range(2):   # two calls in a row: different regions, the array keeps the first call's contents
                regs = mixed_regions(idx, 40 + 7 * mode + rep, n)
                want = idx.query_features(regs, mode, invert, want_pairs=False).n_pairs
                got, bad = _noalloc(idx, torch, regs, mode, invert)
                assert bad == 0
                assert got == want, (n, mode, invert, rep)
    finally:
        idx.close()


def test_noalloc_edges(gffx, synth_plain_gff):
    import torch
    from gffx_b200 import _lib
    idx = gffx.TreeIndexData.load_tree_index(synth_plain_gff, device=0)
    try:
        lim = _lib.lib.gffx_noalloc_max_regions()
        assert lim == 1 << 20 and _lib.lib.gffx_noalloc_max_roots() == 1 << 18
        regs = idx.synth_regions(5, 1000, 5000)
        empty = tuple(a[:0] for a in regs)
        assert _noalloc(idx, torch, empty, 0, False) == (0, 0)
        # a sequence number beyond the index's list is reported, the other regions are counted
        bad = (regs[0].copy(), regs[1], regs[2])
        bad[0][17] = 0x7FFFFFF0
        keep = np.ones(1000, dtype=bool)
        keep[17] = False
        want = idx.query_features(tuple(a[keep] for a in regs), 0, False, want_pairs=False).n_pairs
        got, flag = _noalloc(idx, torch, bad, 0, False)
        assert flag != 0 and got == want
        # beyond the maximum: a clear error and no launch
        out = torch.zeros(2, dtype=torch.int64, device="cuda")
        buf = torch.zeros(8, dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        with pytest.raises(gffx.GffxError, match="exceed the fixed scratch"):
            idx.count_pairs_noalloc(buf.data_ptr(), buf.data_ptr(), buf.data_ptr(), lim + 1, 0, False, out.data_ptr())
        # the largest batch allowed
        big = idx.synth_regions(6, lim, 2000)
        want = idx.query_features(big, 1, False, want_pairs=False).n_pairs
        assert _noalloc(idx, torch, big, 1, False) == (want, 0)
    finally:
        idx.close()
