#!/usr/bin/env python3
"""Generates the committed fixtures under tests/golden/ from the reference checkout.

Run ONCE in the build container (needs /root/reference; the GPU box has no such path):

    python tests/golden/make_golden.py

What it writes
  ps/ps.gff.{sqs,rix,rit,gof}   byte copies of the reference's own index files for P. sativum
                                (benchmark/ps.gff.*, produced by the reference's `gffx index`).
  ps/input_001.bed.xz           columns 1-3 of benchmark/intersect/ps/input_001.txt (100,000
                                `bedtools random -l 20000` regions, benchmark/run_benchmark.sh:24).
  ps/golden.json                root-level join answers computed HERE by numpy brute force straight
                                from the bincode2 bytes of ps.gff.rit — independent of both the C++
                                oracle and the CUDA engine: for every mode x invert the number of
                                pairs, the number of distinct root fids and an order-independent
                                checksum of the pair multiset.
  kat.gff                       the hand-derived known-answer GFF of SURVEY.md App. C.

Predicates (relative to /root/reference): hit = iv.start < re && iv.end > rs (src/utils/tree.rs:110);
keep = Contained iv.start>=rs && iv.end<=re | ContainsRegion iv.start<=rs && iv.end>=re | Overlap true
(src/commands/intersect.rs:145-158); emitted iff hit && (invert ^ keep) (:161-163).
"""
import json
import lzma
import os
import shutil
import struct

import numpy as np

REF = "/root/reference/benchmark"
HERE = os.path.dirname(os.path.abspath(__file__))


def decode_rit(rit: bytes, offsets):
    """bincode2 grammar of SURVEY.md App. B -> list of (start,end,fid) arrays, one per tree."""
    trees = []
    for i, off in enumerate(offsets):
        end = offsets[i + 1] if i + 1 < len(offsets) else len(rit)
        pos, owed, ivs = off, 1, []
        while owed:
            tag = rit[pos]
            pos += 1
            owed -= 1
            if tag == 0:
                continue
            assert tag == 1
            _center, n = struct.unpack_from("<IQ", rit, pos)
            pos += 12
            ivs.append(np.frombuffer(rit, dtype="<u4", count=3 * n, offset=pos).reshape(n, 3))
            pos += 12 * n
            owed += 2
        assert pos == end, (i, pos, end)
        trees.append(np.concatenate(ivs) if ivs else np.zeros((0, 3), dtype=np.uint32))
    return trees


MIX = np.uint64(0x9E3779B97F4A7C15)


def pair_checksum(fid, s, e):
    """sum over pairs of a 64-bit mix of (fid,start,end), mod 2^64 — order independent"""
    with np.errstate(over="ignore"):
        x = fid.astype(np.uint64) * MIX + s.astype(np.uint64) * np.uint64(0xBF58476D1CE4E5B9) + e.astype(np.uint64) * np.uint64(0x94D049BB133111EB)
        x ^= x >> np.uint64(31)
        x *= MIX
        return int(x.sum(dtype=np.uint64))


def main():
    os.makedirs(os.path.join(HERE, "ps"), exist_ok=True)
    for ext in ("sqs", "rix", "rit", "gof"):
        shutil.copy(f"{REF}/ps.gff.{ext}", os.path.join(HERE, "ps", f"ps.gff.{ext}"))
        os.chmod(os.path.join(HERE, "ps", f"ps.gff.{ext}"), 0o644)
    names = open(f"{REF}/ps.gff.sqs").read().split("\n")
    if names and names[-1] == "":
        names.pop()
    num = {n: i for i, n in enumerate(names)}
    offsets = json.load(open(f"{REF}/ps.gff.rix"))
    trees = decode_rit(open(f"{REF}/ps.gff.rit", "rb").read(), offsets)
    bed_lines = []
    chr_, qs, qe = [], [], []
    for line in open(f"{REF}/intersect/ps/input_001.txt"):
        f = line.split()
        bed_lines.append("\t".join(f[:3]) + "\n")
        if f[0] in num:
            chr_.append(num[f[0]])
            qs.append(int(f[1]))
            qe.append(int(f[2]))
    with lzma.open(os.path.join(HERE, "ps", "input_001.bed.xz"), "wb", preset=9) as g:
        g.write("".join(bed_lines).encode())
    chr_, qs, qe = np.array(chr_), np.array(qs, dtype=np.int64), np.array(qe, dtype=np.int64)
    golden = {"n_regions": int(len(chr_)), "n_roots_total": int(sum(len(t) for t in trees)), "cases": {}}
    for mode, mname in ((0, "contained"), (1, "contains_region"), (2, "overlap")):
        for invert in (False, True):
            n_pairs, roots, csum = 0, set(), 0
            for c, t in enumerate(trees):
                sel = np.nonzero(chr_ == c)[0]
                if len(sel) == 0 or len(t) == 0:
                    continue
                ts, te, tf = t[:, 0].astype(np.int64), t[:, 1].astype(np.int64), t[:, 2]
                for b0 in range(0, len(sel), 2048):
                    q = sel[b0:b0 + 2048]
                    rs, re = qs[q][:, None], qe[q][:, None]
                    hit = (ts[None, :] < re) & (te[None, :] > rs)
                    if mode == 0:
                        keep = (ts[None, :] >= rs) & (te[None, :] <= re)
                    elif mode == 1:
                        keep = (ts[None, :] <= rs) & (te[None, :] >= re)
                    else:
                        keep = np.ones_like(hit)
                    m = hit & (keep != invert)
                    _, j = np.nonzero(m)
                    n_pairs += len(j)
                    roots.update(tf[j].tolist())
                    csum = (csum + pair_checksum(tf[j], t[j, 0], t[j, 1])) & 0xFFFFFFFFFFFFFFFF
            golden["cases"][f"{mname}{'+invert' if invert else ''}"] = {"pairs": n_pairs, "roots": len(roots), "checksum": csum}
            print(mname, invert, n_pairs, len(roots), csum)
    json.dump(golden, open(os.path.join(HERE, "ps", "golden.json"), "w"), indent=1, sort_keys=True)

    kat = [
        "##gff-version 3",
        "chr1\tsrc\tgene\t101\t200\t.\t+\t.\tID=g1",
        "chr1\tsrc\tmRNA\t101\t200\t.\t+\t.\tID=m1;Parent=g1",
        "chr1\tsrc\texon\t101\t150\t.\t+\t.\tID=e1;Parent=m1",
        "chr1\tsrc\texon\t181\t200\t.\t+\t.\tID=e2;Parent=m1",
        "chr1\tsrc\tgene\t301\t400\t.\t-\t.\tID=g2",
        "chr1\tsrc\texon\t301\t400\t.\t-\t.\tID=e3;Parent=g2",
        "chr2\tsrc\tgene\t1\t50\t.\t+\t.\tID=g3",
    ]
    open(os.path.join(HERE, "kat.gff"), "w").write("\n".join(kat) + "\n")


if __name__ == "__main__":
    main()
