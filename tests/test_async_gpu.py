"""GFFX_F_ASYNC: jobs queued back to back on the engine's stream, counters read later (include/gffx_b200.h: gffx_result_wait).
Same numbers as the synchronous calls, in any wait order; errors surface at the wait; the rank-space epochs (filter.cuh)
survive the wrap of both counters."""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from gffx_b200 import _lib
from tests._oracle import Oracle
from tests.test_intersect_gpu import MODES, mixed_regions

pytestmark = pytest.mark.gpu
F = _lib.F_ASYNC | _lib.F_NO_HOST_COPY | _lib.F_PAIRS_ON_DEVICE


@pytest.fixture(scope="module")
def quirk(gffx, synth_quirk_gff):
    idx = gffx.TreeIndexData.load_tree_index(synth_quirk_gff, device=0)
    yield idx
    idx.close()


def dev(regs):
    return tuple(torch.from_numpy(a.view(np.int32)).cuda() for a in regs)


def test_queued_jobs_equal_synchronous_calls(quirk):
    idx = quirk
    sets = [mixed_regions(idx, 900 + k, 2500) for k in range(4)]
    dsets = [dev(s) for s in sets]
    cases = [(m, i, e, t) for (m, i) in MODES for (e, t) in ((False, None), (True, None), (False, "gene,mRNA,exon"))]
    want = [idx.intersect(sets[k % 4], m, i, e, t) for k, (m, i, e, t) in enumerate(cases)]
    pend = [idx.intersect(dsets[k % 4], m, i, e, t, flags=F) for k, (m, i, e, t) in enumerate(cases)]
    order = list(range(len(pend)))
    np.random.default_rng(1).shuffle(order)
    for k in order:                       # waiting out of order is fine: every job has its own slot
        r = pend[k].wait()
        w = want[k]
        assert (r.n_pairs, r.n_roots, r.n_ranges, r.bytes_out, sorted(r.warn_fids)) == (w.n_pairs, w.n_roots, w.n_ranges, w.bytes_out, sorted(w.warn_fids)), cases[k]
        assert r.kernel_ms > 0 and r.launches > 0


def test_queued_jobs_with_host_regions_and_host_ranges(gffx, quirk):
    """GFFX_F_ASYNC without GFFX_F_NO_HOST_COPY and without GFFX_F_DEVICE_INPUT: host arrays in, the host copy of the ranges
    behind the job; two jobs in flight, so one job's ranges travel back while the next job's regions travel in"""
    idx = quirk
    FH = _lib.F_ASYNC | _lib.F_PAIRS_ON_DEVICE
    sets = [mixed_regions(idx, 300 + k, 4000) for k in range(4)]
    cases = [(m, i, e, t) for (m, i) in MODES for (e, t) in ((False, None), (True, None), (False, "gene,mRNA,exon"))]
    want = [idx.intersect(sets[k % 4], m, i, e, t) for k, (m, i, e, t) in enumerate(cases)]

    def same(r, w, what):
        assert (r.n_pairs, r.n_roots, r.n_ranges, r.bytes_out) == (w.n_pairs, w.n_roots, w.n_ranges, w.bytes_out), what
        if w.n_ranges:
            assert r.ranges is not None and np.array_equal(np.asarray(r.ranges), w.ranges), what
        r.close()

    for copy in (True, False):
        pend = []
        for k, (m, i, e, t) in enumerate(cases):
            pend.append((k, idx.intersect(sets[k % 4], m, i, e, t, flags=FH)))
            if len(pend) > 2:
                j, p = pend.pop(0)
                same(p.wait(copy=copy), want[j], cases[j])
        for j, p in pend:
            same(p.wait(copy=copy), want[j], cases[j])
    # a result dropped while its ranges are in flight, then the lanes are used again
    for _ in range(3):
        idx.intersect(sets[0], 0, False, False, None, flags=FH).close()
    p = idx.intersect(sets[1], 0, False, False, None, flags=FH)
    w = idx.intersect(sets[1], 0, False, False, None)       # a synchronous call between queue and wait
    same(p.wait(), w, "after abandoned results")
    with pytest.raises(gffx.GffxError):                      # pairs are not copied to the host behind a queued job
        idx.intersect(sets[0], 0, False, False, None, flags=FH | _lib.F_WANT_PAIRS)
    # the batched form (gffx_shard_begin / add / finish), three batches of host regions per job, two jobs in flight
    pend = [(k, idx.shard_intersect(sets[k % 4], m, i, e, t, batches=3, flags=FH)) for k, (m, i, e, t) in enumerate(cases[:6])]
    for j, p in pend:
        same(p.wait(copy=False), want[j], ("batched", cases[j]))


def test_last_jobs_device_outputs_are_readable(quirk):
    idx = quirk
    regs = mixed_regions(idx, 77, 3000)
    w = idx.intersect(regs, 2, False, False, None, want_pairs=True)
    p = idx.intersect(dev(regs), 2, False, False, None, flags=F)
    r = p.wait()
    assert r.n_ranges == w.n_ranges and r.pairs_device_ptr
    got = idx.gather_text(None, r.n_ranges)            # the ranges of the handle's last job are still on the device
    assert len(got) == w.bytes_out and got == idx.gather_text(w.ranges, w.n_ranges)


def test_two_lanes_overlap_and_keep_their_outputs(quirk):
    """queued jobs alternate between two lanes (own buffers, own streams): big enough to really run side by side; the device
    outputs of a job survive the job queued after it (the other lane's)"""
    if os.environ.get("GFFX_LANES") == "1":
        pytest.skip("one lane: a queued job's device outputs last until the next job")
    idx = quirk
    sets = [mixed_regions(idx, 40 + k, 300_000) for k in range(3)]
    dsets = [dev(s) for s in sets]
    cases = [(2, False), (0, False), (1, False), (0, True), (2, False), (1, True), (0, False)]
    want = [idx.intersect(sets[k % 3], m, i, False, None) for k, (m, i) in enumerate(cases)]
    for rep in range(2):
        pend = [idx.intersect(dsets[k % 3], m, i, False, None, flags=F) for k, (m, i) in enumerate(cases)]
        if rep == 0:   # a synchronous job in between runs on lane 0, behind its queued jobs
            mid = idx.intersect(sets[0], 2, False, False, None)
            assert mid.n_ranges == want[0].n_ranges and np.array_equal(mid.ranges, want[0].ranges)
        for k in (5, 6, 0, 1, 2, 3, 4):
            r = pend[k].wait()
            assert (r.n_pairs, r.n_roots, r.n_ranges, r.bytes_out) == (want[k].n_pairs, want[k].n_roots, want[k].n_ranges, want[k].bytes_out), (rep, k)
            if rep == 1 and k >= 5:   # the last job of either lane: its ranges are still on the device
                assert idx.gather_text(None, r.n_ranges) == idx.gather_text(want[k].ranges, want[k].n_ranges)


def test_error_surfaces_at_the_wait_and_ring_limit(gffx, quirk):
    idx = quirk
    bad = (torch.tensor([999], dtype=torch.int32).cuda(), torch.tensor([1], dtype=torch.int32).cuda(), torch.tensor([2], dtype=torch.int32).cuda())
    p = idx.intersect(bad, 2, False, False, None, flags=F)
    with pytest.raises(gffx.GffxError) as ei:
        p.wait()
    assert ei.value.code == -1
    regs = dev(mixed_regions(idx, 5, 500))
    pend = [idx.intersect(regs, 0, False, False, None, flags=F) for _ in range(32)]
    with pytest.raises(gffx.GffxError) as ei:          # 32 results may be pending per handle
        idx.intersect(regs, 0, False, False, None, flags=F)
    assert ei.value.code == -5
    for p in pend[:16]:
        p.close()                                      # abandoned without a wait: the slots free themselves
    idx.sync()
    more = [idx.intersect(regs, 0, False, False, None, flags=F) for _ in range(16)]
    n = {p.wait().n_pairs for p in pend[16:] + more}
    assert len(n) == 1


def test_epochs_wrap(gffx, synth_quirk_gff):
    """the rank-space arrays are tagged per job instead of cleared; GFFX_EPOCH_START puts both counters next to their wrap"""
    code = (
        "import os, sys, numpy as np\n"
        "sys.path.insert(0, os.getcwd())\n"
        "import gffx_b200 as g\n"
        "from tests._oracle import Oracle\n"
        "from tests.test_intersect_gpu import mixed_regions\n"
        f"gff = {synth_quirk_gff!r}\n"
        "idx = g.TreeIndexData.load_tree_index(gff, device=0)\n"
        "with Oracle(gff) as o:\n"
        "    for k in range(16):\n"
        "        regs = mixed_regions(idx, 50 + k, 1500)\n"
        "        mode, inv = [(2, False), (0, False), (2, False), (1, False), (2, False), (0, True), (2, False), (1, True)][k % 8]\n"
        "        r = idx.intersect(regs, mode, inv, False, None)\n"
        "        want, n_pairs, _ = o.intersect(regs, mode, inv, False, None)\n"
        "        assert r.n_pairs == n_pairs and np.array_equal(r.ranges, want), (k, mode, inv)\n"
        "print('ok')\n"
    )
    env = dict(os.environ, GFFX_EPOCH_START=str((1 << 23) - 1 - 4))   # wraps after 4 jobs; (2^23-5) % 256 = 251: the marks wrap too
    out = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=300,
                         cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    assert out.returncode == 0 and out.stdout.strip().endswith("ok"), out.stderr[-2000:]
