"""An index with the shape of the reference's human index: tests/golden/hs/hs.gff.{sqs,rix} are byte copies of the
reference's own benchmark/hs.gff.sqs (583 sequence names) and benchmark/hs.gff.rix (583 offsets into a .rit that upstream
keeps out of the repository).  The .rit written here has EXACTLY the byte size the .rix offsets dictate for every tree
(bincode2 grammar of SURVEY.md App. B: a tree of k nodes and N intervals takes 1 + 14 k + 12 N bytes), filled with
synthetic intervals; the .gof gives every root a 24-byte record.  So the loaders read the reference's real header
files unchanged, and the join has 583 sequences to cope with.  TEST INFRASTRUCTURE ONLY."""
import json
import os
import shutil
import struct

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
HS = os.path.join(HERE, "golden", "hs")


def tree_shape(nbytes):
    """(k nodes, N intervals) with 1 + 14 k + 12 N == nbytes, smallest k >= 1"""
    for k in range(1, 7):
        rest = nbytes - 1 - 14 * k
        if rest >= 0 and rest % 12 == 0:
            return k, rest // 12
    raise ValueError(f"no tree of {nbytes} bytes")


def encode_chain(ivs, k):
    """k nodes chained through their LEFT child; the first k-1 hold one interval each, the last one the rest"""
    parts = [ivs[i:i + 1] for i in range(k - 1)] + [ivs[k - 1:]]
    out = bytearray()
    for p in parts:
        out += b"\x01" + struct.pack("<IQ", int(p[0, 0]) if len(p) else 0, len(p)) + p.astype("<u4").tobytes()
    out += b"\x00"                  # left child of the last node
    out += b"\x00" * k              # right children, innermost first
    return bytes(out)


def write_hs_like(dirpath, seed=20260583):
    """-> (gff path without text, per-sequence list of (start,end,fid) arrays)"""
    os.makedirs(dirpath, exist_ok=True)
    base = os.path.join(dirpath, "hs.gff")
    shutil.copy(os.path.join(HS, "hs.gff.sqs"), base + ".sqs")
    shutil.copy(os.path.join(HS, "hs.gff.rix"), base + ".rix")
    offs = json.load(open(base + ".rix"))
    sizes = [b - a for a, b in zip(offs, offs[1:])] + [1 + 14 + 12 * 40]   # the last tree's size is not recorded: 40 intervals
    rng = np.random.default_rng(seed)
    rit = bytearray()
    trees, fid = [], 0
    gof = bytearray()
    for nbytes in sizes:
        k, n = tree_shape(nbytes)
        k = min(k, max(n, 1)) if n else k
        if n < k:   # a node needs no interval, but keep the chain simple: k nodes, n intervals, n may be < k only when n == 0
            assert n == 0
        length = max(20000, n * 3000)
        s = np.sort(rng.integers(0, length, size=n)).astype(np.uint32)
        ln = np.minimum(rng.lognormal(9.5, 1.3, size=n), 2.0e6).astype(np.uint32) + 50
        iv = np.stack([s, s + ln, np.arange(fid, fid + n, dtype=np.uint32)], axis=1).astype(np.uint32) if n else np.zeros((0, 3), np.uint32)
        fid += n
        enc = encode_chain(iv, k) if n >= k else b"\x01" + struct.pack("<IQ", 0, 0) + b"\x00\x00" + b""
        if n >= k:
            assert len(enc) == nbytes, (len(enc), nbytes, k, n)
        rit += enc
        trees.append(iv)
        for f in iv[:, 2]:
            gof += struct.pack("<IIQQ", int(f), 0, 24 * int(f), 24 * int(f) + 24)
    open(base + ".rit", "wb").write(bytes(rit))
    open(base + ".gof", "wb").write(bytes(gof))
    return base, trees
