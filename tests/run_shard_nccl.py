"""torchrun --nproc-per-node N tests/run_shard_nccl.py <gff with index files>: every rank joins its share of the regions,
the accumulators are exchanged over NCCL (gffx_b200.shard.Combiner), every rank finishes with the global state, and the
result is compared with the CPU oracle run on ALL the regions.  TEST INFRASTRUCTURE (imports the oracle)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist

    import gffx_b200 as g
    from gffx_b200.shard import Combiner, shard_bounds
    from tests._oracle import Oracle

    gff = sys.argv[1]
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    idx = g.TreeIndexData.load_tree_index(gff, device=local)
    comb = Combiner(idx)
    n = 40000
    a = idx.synth_regions(3, n, 20000)
    b = idx.synth_regions(4, n, 150)
    regs = tuple(np.concatenate([x, y]) for x, y in zip(a, b))
    lo, hi = shard_bounds(2 * n, world, rank)
    mine = tuple(x[lo:hi].copy() for x in regs)
    with Oracle(gff) as o:
        for mode, invert, entire, types in [(2, False, False, None), (0, False, False, "gene,exon"), (1, False, False, None),
                                            (0, True, False, None), (2, False, True, None)]:
            r = idx.shard_intersect(mine, mode, invert, entire, types, combine=comb, batches=2)
            want, n_pairs, _ = o.intersect(regs, mode, invert, entire, types)
            tot = torch.tensor([r.n_pairs], device="cuda", dtype=torch.int64)
            dist.all_reduce(tot)
            assert int(tot.item()) == n_pairs, (mode, invert, int(tot.item()), n_pairs)
            assert np.array_equal(r.ranges, want), (rank, mode, invert, entire, types)
        assert idx.depth(mine, combine=comb).to_tsv() == o.depth(regs)
        assert idx.coverage(mine, combine=comb).to_tsv() == o.coverage(regs)
    dist.barrier()
    if rank == 0:
        print(f"shard nccl ok: world {world}, {comb.bytes_exchanged} bytes exchanged per rank", flush=True)
    idx.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
