"""Properties of the hot path at BASELINE.json's full size (configs[1]: 60 000-gene human-like GFF3, 10 M 20 kb regions per
pass), where the oracle cannot finish: identities between the modes, additivity over region chunks, determinism, the shape
of the range list — plus a bit-exact oracle comparison on a bounded sample of the same stream against the same full index."""
import numpy as np
import pytest

from tests._oracle import Oracle

pytestmark = pytest.mark.gpu

GENES, Q, LEN, SEED = 60000, 10_000_000, 20000, 20260002


@pytest.fixture(scope="module")
def full():
    import torch

    import gffx_b200 as g
    from bench import ensure_workload

    gff = ensure_workload(GENES, lambda p, s, n: g.synth_gff(p, seed=s, n_genes=n, quirks=False), g.build_index)
    idx = g.TreeIndexData.load_tree_index(gff, device=0)
    regs = idx.synth_regions_device(SEED, Q, LEN, device=torch.device("cuda", 0))
    yield idx, gff, regs
    idx.close()


def run(idx, regs, mode, invert, **kw):
    from gffx_b200 import _lib

    r = idx.intersect(regs, mode, invert, False, None, flags=_lib.F_PAIRS_ON_DEVICE, copy=True, **kw)
    return r


def test_mode_identities_and_range_shape(full):
    idx, _, regs = full
    res = {(m, inv): run(idx, regs, m, inv) for m, inv in [(2, False), (2, True), (0, False), (0, True), (1, False), (1, True)]}
    ov = res[(2, False)]
    assert ov.n_queries == Q and ov.n_pairs > Q   # ~1.4 hits per 20 kb region on this annotation
    # --invert keeps the overlapping intervals that FAIL the mode's extra condition: the two halves partition the overlap set
    assert res[(0, False)].n_pairs + res[(0, True)].n_pairs == ov.n_pairs
    assert res[(1, False)].n_pairs + res[(1, True)].n_pairs == ov.n_pairs
    assert res[(2, True)].n_pairs == 0 and res[(2, True)].n_ranges == 0   # overlap has no extra condition to fail
    for r in res.values():
        assert r.n_roots <= idx.n_roots and r.n_ranges <= idx.n_lines
        if r.n_ranges:
            a = r.ranges
            assert np.all(a[:, 0] < a[:, 1])
            assert np.all(a[1:, 0] >= a[:-1, 1])                      # file order, no line twice
            assert int((a[:, 1] - a[:, 0]).sum()) == r.bytes_out
    # 10 M uniformly placed 20 kb regions cover the genome many times over: every line of every hit root overlaps one
    assert ov.n_roots == idx.n_roots and ov.n_ranges >= idx.n_lines - 1


def test_same_call_twice_gives_the_same_result(full):
    idx, _, regs = full
    a, b = run(idx, regs, 0, False), run(idx, regs, 0, False)
    assert (a.n_pairs, a.n_roots, a.n_ranges, a.bytes_out) == (b.n_pairs, b.n_roots, b.n_ranges, b.bytes_out)
    assert np.array_equal(a.ranges, b.ranges)


def test_job_in_batches_equals_one_call_and_pairs_add_up(full):
    import torch

    idx, _, regs = full
    from gffx_b200 import _lib

    whole = run(idx, regs, 1, False)
    batched = idx.shard_intersect(regs, 1, False, False, None, batches=4, flags=_lib.F_PAIRS_ON_DEVICE)
    assert batched.n_pairs == whole.n_pairs and batched.n_roots == whole.n_roots
    assert np.array_equal(batched.ranges, whole.ranges)
    # the join is additive over region chunks: two halves as two calls
    h = Q // 2
    lo = tuple(t[:h].contiguous() for t in regs)
    hi = tuple(t[h:].contiguous() for t in regs)
    assert run(idx, lo, 1, False).n_pairs + run(idx, hi, 1, False).n_pairs == whole.n_pairs
    torch.cuda.synchronize()


def test_pair_array_on_the_device_names_roots_of_the_index(full):
    """Every materialised pair is a (root_fid, start, end) triple of the index, the array is grouped by root in (sequence,
    start, end) order, and its length is n_pairs (checked on the device: 13.9 M pairs)."""
    import torch

    idx, _, regs = full
    from gffx_b200 import _lib
    from gffx_b200.shard import _DevView

    r = idx.intersect(regs, 2, False, False, None, flags=_lib.F_PAIRS_ON_DEVICE | _lib.F_NO_HOST_COPY, copy=False)
    try:
        assert r.pairs_device_ptr and r.n_pairs > 0
        p = torch.as_tensor(_DevView(r.pairs_device_ptr, r.n_pairs * 12), device="cuda:0").view(torch.int32).view(-1, 3)
        assert p.shape[0] == r.n_pairs
        uniq, counts = torch.unique_consecutive(p, dim=0, return_counts=True)
        assert uniq.shape[0] == torch.unique(p, dim=0).shape[0]          # a root's pairs are one run
        assert int(counts.sum()) == r.n_pairs
        roots = set()
        for s in range(len(idx.num_to_seqid)):
            t = np.asarray(idx.seq_roots(s), dtype=np.uint32).reshape(-1, 3)   # (start, end, root_fid)
            roots.update(map(tuple, t[:, [2, 0, 1]].tolist()))
        got = set(map(tuple, uniq.cpu().numpy().view(np.uint32).tolist()))
        assert got <= roots and len(got) == r.n_roots
    finally:
        r.close()


def test_sample_of_the_full_stream_matches_the_oracle_on_the_full_index(full):
    idx, gff, _ = full
    n = 20000
    regs = idx.synth_regions(SEED, n, LEN)   # the first n regions of the same counter-based stream
    with Oracle(gff) as o:
        for mode, invert in [(2, False), (0, False), (1, True)]:
            r = idx.intersect(regs, mode, invert, False, None, want_pairs=True)
            want_pairs = o.query_features(regs, mode, invert)
            got = r.pairs if r.pairs is not None else np.zeros((0, 3), dtype=np.uint32)
            order = lambda a: a[np.lexsort((a[:, 2], a[:, 1], a[:, 0]))] if len(a) else a
            assert np.array_equal(order(got), order(want_pairs))
            w_ranges, w_pairs, _ = o.intersect(regs, mode, invert, False, None)
            assert w_pairs == r.n_pairs
            assert np.array_equal(r.ranges, w_ranges)


def _by_id(rows):
    return dict(zip(rows.id_idx.tolist(), rows.value.tolist()))


def test_depth_adds_over_read_chunks_and_coverage_is_idempotent(full):
    """depth is a sum over the reads (linearity), breadth a union (reading the same reads twice changes nothing); a job fed in
    batches equals the job fed at once — 10 M 150-base reads against the full index."""
    import torch

    idx, _, _ = full
    reads = idx.synth_regions_device(20260010, Q, 150, device=torch.device("cuda", 0))
    h = Q // 2
    lo = tuple(t[:h].contiguous() for t in reads)
    hi = tuple(t[h:].contiguous() for t in reads)
    whole, a, b = idx.depth(reads), idx.depth(lo), idx.depth(hi)
    dw, da, db = _by_id(whole), _by_id(a), _by_id(b)
    assert len(dw) > 100000 and sum(dw.values()) > Q // 2
    assert set(dw) == set(da) | set(db)
    assert all(dw[k] == da.get(k, 0) + db.get(k, 0) for k in dw)
    batched = idx.depth(reads, batches=4)
    for f in ("id_idx", "chrom_idx", "start", "end", "value"):
        assert np.array_equal(getattr(batched, f), getattr(whole, f))
    cov = idx.coverage(reads)
    twice = idx.coverage(tuple(torch.cat([t, t]) for t in reads))
    for f in ("id_idx", "chrom_idx", "start", "end", "value"):
        assert np.array_equal(getattr(cov, f), getattr(twice, f))
    length = np.maximum(cov.end.astype(np.int64) - cov.start.astype(np.int64), 0)
    assert np.all(cov.value.astype(np.int64) <= length)
    ca, cw = _by_id(idx.coverage(lo)), _by_id(cov)
    assert all(ca[k] <= cw.get(k, 0) for k in ca)   # more reads never cover less
