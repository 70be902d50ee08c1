"""Batched and sharded jobs through the C ABI (gffx_shard_*, gffx_job_buffer, gffx_fold) against the oracle run on the
UNION of the regions.  Two shards are simulated on one GPU: shard A's accumulators are saved, shard B's job folds them in
with the engine's own fold kernel before its finish step — exactly what gffx_b200.shard.Combiner does across ranks."""
import os
import subprocess
import sys

import numpy as np
import pytest

from tests._oracle import Oracle
from tests.test_intersect_gpu import MODES, mixed_regions

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def quirk(gffx, synth_quirk_gff):
    idx = gffx.TreeIndexData.load_tree_index(synth_quirk_gff, device=0)
    o = Oracle(synth_quirk_gff)
    yield idx, o
    o.close()
    idx.close()


def _split(regs, k):
    return tuple(a[:k].copy() for a in regs), tuple(a[k:].copy() for a in regs)


class TwoShardsOnOneGpu:
    """combine callback: first call (shard A) saves the accumulators, second call (shard B) folds A's copies in"""

    def __init__(self, idx):
        self.idx = idx
        self.saved = None

    def __call__(self, buffers):
        import torch

        from gffx_b200 import _lib
        from gffx_b200.shard import OPS, _DevView

        self.idx.sync()
        views = []
        for ptr, n, op in buffers:
            es = np.dtype(OPS[op][0]).itemsize
            views.append((torch.as_tensor(_DevView(ptr, n * es), device="cuda:0") if n else None, n, op, es))
        if self.saved is None:
            self.saved = [v.clone() if v is not None else None for v, _, _, _ in views]
            return
        assert len(self.saved) == len(views)
        for (v, n, op, es), a in zip(views, self.saved):
            if n == 0:
                continue
            two = torch.cat([a, v])   # rows: shard A, shard B
            torch.cuda.synchronize()
            rc = _lib.lib.gffx_fold(self.idx._h, v.data_ptr(), two.data_ptr(), 2, n, n, op)
            assert rc == 0, _lib.last_error()
            self.idx.sync()
        self.saved = None


@pytest.mark.parametrize("mode,invert", MODES)
@pytest.mark.parametrize("entire,types", [(False, None), (True, None), (False, "gene,exon")])
def test_batched_job_equals_one_call(quirk, mode, invert, entire, types):
    idx, o = quirk
    regs = mixed_regions(idx, 31 + mode, 3000)
    r = idx.shard_intersect(regs, mode, invert, entire, types, batches=3)
    want_ranges, n_pairs, _ = o.intersect(regs, mode, invert, entire, types)
    assert r.n_pairs == n_pairs
    assert np.array_equal(r.ranges, want_ranges)


@pytest.mark.parametrize("mode,invert", MODES)
@pytest.mark.parametrize("entire,types", [(False, None), (True, None), (False, "gene,exon")])
def test_two_shards_combined_equal_the_union(quirk, mode, invert, entire, types):
    idx, o = quirk
    regs = mixed_regions(idx, 57 + mode, 3000)
    a, b = _split(regs, len(regs[0]) // 3)
    comb = TwoShardsOnOneGpu(idx)
    ra = idx.shard_intersect(a, mode, invert, entire, types, combine=comb)   # shard A alone (its accumulators are saved)
    rb = idx.shard_intersect(b, mode, invert, entire, types, combine=comb)   # shard B + A's accumulators
    want_ranges, n_pairs, _ = o.intersect(regs, mode, invert, entire, types)
    assert ra.n_pairs + rb.n_pairs == n_pairs
    assert np.array_equal(rb.ranges, want_ranges)


@pytest.mark.parametrize("kind", [0, 1])
def test_two_shards_depth_coverage(quirk, kind):
    idx, o = quirk
    a3 = idx.synth_regions(91, 20000, 150)
    b3 = idx.synth_regions(92, 3000, 20000)
    regs = tuple(np.concatenate([x, y]) for x, y in zip(a3, b3))
    a, b = _split(regs, 9000)
    comb = TwoShardsOnOneGpu(idx)
    fn = idx.depth if kind == 0 else idx.coverage
    fn(a, combine=comb)
    rows = fn(b, batches=2, combine=comb)
    want = o.depth(regs) if kind == 0 else o.coverage(regs)
    assert rows.to_tsv() == want


def test_batched_job_with_pairs(quirk):
    """a job that announces its size keeps the bucketed batches and materialises the pairs like one call does"""
    from tests._oracle import sorted_pairs

    idx, o = quirk
    regs = mixed_regions(idx, 77, 3000)
    for mode, invert in MODES:
        r = idx.shard_intersect(regs, mode, invert, batches=3, want_pairs=True)
        want = o.query_features(regs, mode, invert)
        got = r.pairs if r.pairs is not None else np.zeros((0, 3), dtype=np.uint32)
        assert r.n_pairs == want.shape[0]
        assert np.array_equal(sorted_pairs(got), sorted_pairs(want))


def test_streaming_job_refuses_pairs(quirk, gffx):
    idx, _ = quirk
    from gffx_b200 import _lib

    with pytest.raises(gffx.GffxError):
        gffx.api._check(_lib.lib.gffx_shard_begin(idx._h, 2, 0, 0, None, _lib.F_WANT_PAIRS, 0, 1))


def test_two_gpus_nccl(synth_quirk_gff):
    """real exchange over NCCL when the box has two GPUs (tests/run_shard_nccl.py under torchrun)"""
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29517", os.path.join(ROOT, "tests", "run_shard_nccl.py"), synth_quirk_gff]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and "shard nccl ok" in out.stdout, out.stdout[-2000:] + out.stderr[-2000:]
