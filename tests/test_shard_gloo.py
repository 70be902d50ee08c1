"""Host-side logic of the multi-GPU path (gffx_b200/shard.py) on CPU: world_size-2 `gloo` processes run the accumulator
exchange protocol (all-to-all of slices, fold, all-gather) for every reduce op of the C ABI, the shard bounds, and the
"gather to rank 0" merge of depth rows checked against the oracle on the union of the shards' regions."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _parse_tsv(tsv: bytes) -> dict:
    rows = {}
    for line in tsv.split(b"\n")[1:]:
        if line:
            i, c, s, e, v = line.split(b"\t")
            rows[i] = (c, int(s), int(e), int(v))
    return rows


def _worker(rank, world, port, gff, q):
    try:
        sys.path.insert(0, ROOT)
        import torch.distributed as dist

        from gffx_b200 import shard
        from tests._oracle import Oracle

        dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
        # ---- the exchange protocol, every op, sizes that do not divide by the world size ----------------
        for op, (dt, fold) in shard.OPS.items():
            for n in (1, 7, 1000, 4097):
                per_rank = []
                for r in range(world):
                    rng = np.random.default_rng(1000 * op + 10 * n + r)
                    if np.dtype(dt) == np.int32:
                        a = rng.integers(-1000, 1000, size=n).astype(dt)
                    else:
                        a = rng.integers(0, np.iinfo(dt).max if op != 3 else 1 << 20, size=n, dtype=np.uint64).astype(dt)
                    per_rank.append(a)
                got = shard.combine_array(per_rank[rank], op)
                want = fold(np.stack(per_rank))
                assert got.dtype == np.dtype(dt) and np.array_equal(got, want), (op, n)
        # ---- depth rows: every rank counts its chunk with the oracle, rank 0 merges (depth.rs:501-508) --
        import gffx_b200 as g

        idx = g.TreeIndexData.load_tree_index(gff, device=-1)   # host-only: generators, no compute
        regs = idx.synth_regions(5, 30000, 150)
        with Oracle(gff) as o:
            lo, hi = shard.shard_bounds(30000, world, rank, align=4096)
            mine = _parse_tsv(o.depth(tuple(a[lo:hi] for a in regs)))
            gathered = [None] * world
            dist.all_gather_object(gathered, mine)
            if rank == 0:
                merged = shard.merge_count_rows(gathered, 0)
                assert merged == _parse_tsv(o.depth(regs))
        idx.close()
        dist.barrier()
        dist.destroy_process_group()
        q.put((rank, "ok"))
    except Exception as e:  # noqa: BLE001
        import traceback

        q.put((rank, traceback.format_exc() + repr(e)))


def test_shard_bounds_cover_everything():
    from gffx_b200.shard import shard_bounds

    for n in (0, 1, 4095, 4096, 4097, 10_000_000, 10_000_001):
        for w in (1, 2, 3, 4, 8):
            spans = [shard_bounds(n, w, r) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            for (a, b), (c, d) in zip(spans, spans[1:]):
                assert b == c and a <= b and c <= d
            assert all(lo % 4096 == 0 or lo == n for lo, _ in spans)


def test_merge_count_rows_rejects_coverage():
    from gffx_b200.shard import merge_count_rows

    with pytest.raises(ValueError):
        merge_count_rows([{}, {}], 1)


@pytest.mark.timeout(300)
def test_exchange_protocol_world2_gloo(synth_plain_gff):
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, synth_plain_gff, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(res) == [(0, "ok"), (1, "ok")], res
