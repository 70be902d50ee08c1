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
  at/, dm/                      the reference's index files for A. thaliana and D. melanogaster (xz; `at` comes from an older
                                writer: .gof sequence numbers all 0, and 39,296 tree intervals over 39,289 distinct root
                                fids — DUPLICATE ROOT FIDS on real data), a seeded BED, and golden.json by the same numpy
                                brute force + the `-e` answer (blocks of the hit roots = LAST .gof record of each fid,
                                sorted, merged while s <= current end; src/utils/common.rs:199-229).
  quirk/                        a ~2.4 k-line synthetic GFF with every quirk of SURVEY.md App. D switched on (+ 12 scaffolds),
                                its index files, two BED files (intersect dialect / depth-coverage dialect) and what the
                                NAIVE evaluator tests/golden/naive_eval.py — pure Python, O(L x Q), written from the Rust
                                source independently of oracle/ — makes of them: per mode x invert x {plain,-e,-T,-e -T}
                                pairs / roots / ranges / emitted-text digests, and the sorted depth / coverage tables.

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


def brute_cases(trees, chr_, qs, qe):
    cases = {}
    for mode, mname in ((0, "contained"), (1, "contains_region"), (2, "overlap")):
        for invert in (False, True):
            n_pairs, roots, csum = 0, set(), 0
            for c, t in enumerate(trees):
                sel = np.nonzero(chr_ == c)[0]
                if len(sel) == 0 or len(t) == 0:
                    continue
                ts, te, tf = t[:, 0].astype(np.int64), t[:, 1].astype(np.int64), t[:, 2]
                for b0 in range(0, len(sel), 1024):
                    q = sel[b0:b0 + 1024]
                    rs, re = qs[q][:, None], qe[q][:, None]
                    hit = (ts[None, :] < re) & (te[None, :] > rs)
                    keep = (ts[None, :] >= rs) & (te[None, :] <= re) if mode == 0 else (ts[None, :] <= rs) & (te[None, :] >= re) if mode == 1 else np.ones_like(hit)
                    _, j = np.nonzero(hit & (keep != invert))
                    n_pairs += len(j)
                    roots.update(tf[j].tolist())
                    csum = (csum + pair_checksum(tf[j], t[j, 0], t[j, 1])) & 0xFFFFFFFFFFFFFFFF
            cases[f"{mname}{'+invert' if invert else ''}"] = {"pairs": n_pairs, "roots": len(roots), "checksum": csum, "root_fids": sorted(roots)}
    return cases


def real_index(name, n_regions, seed):
    """at / dm: index files as the reference wrote them, a seeded BED over their sequences, brute-force answers"""
    import hashlib

    d = os.path.join(HERE, name)
    os.makedirs(d, exist_ok=True)
    for ext in ("sqs", "rix", "rit", "gof"):
        with open(f"{REF}/{name}.gff.{ext}", "rb") as f, lzma.open(os.path.join(d, f"{name}.gff.{ext}.xz"), "wb", preset=9) as g:
            g.write(f.read())
    names = open(f"{REF}/{name}.gff.sqs").read().split("\n")
    if names and names[-1] == "":
        names.pop()
    offsets = json.load(open(f"{REF}/{name}.gff.rix"))
    trees = decode_rit(open(f"{REF}/{name}.gff.rit", "rb").read(), offsets)
    gof = np.frombuffer(open(f"{REF}/{name}.gff.gof", "rb").read(), dtype=np.dtype([("fid", "<u4"), ("seq", "<u4"), ("s", "<u8"), ("e", "<u8")]))
    rng = np.random.default_rng(seed)
    ext = np.array([int(t[:, 1].max()) if len(t) else 1 for t in trees], dtype=np.int64)
    w = ext / ext.sum()
    chr_ = rng.choice(len(trees), size=n_regions, p=w)
    ln = rng.choice(np.array([150, 2000, 20000, 20000, 20000, 500000]), size=n_regions)
    qs = (rng.random(n_regions) * ext[chr_]).astype(np.int64)
    qe = qs + ln
    flip = rng.random(n_regions) < 0.01
    qs, qe = np.where(flip, qe, qs), np.where(flip, qs, qe)
    with lzma.open(os.path.join(d, "regions.bed.xz"), "wb", preset=9) as g:
        g.write("".join(f"{names[c]}\t{s}\t{e}\n" for c, s, e in zip(chr_, qs, qe)).encode())
    cases = brute_cases(trees, chr_, qs, qe)
    all_fids = np.concatenate([t[:, 2] for t in trees])
    last = {}
    for i, f in enumerate(gof["fid"].tolist()):
        last[f] = i                                     # GofMap::index: the last record of a fid wins (gof.rs:32-37)
    for cname, c in cases.items():
        blocks = sorted((int(gof["s"][last[f]]), int(gof["e"][last[f]])) for f in c.pop("root_fids") if f in last)
        merged = []
        for s, e in blocks:                             # common.rs:199-229
            if merged and s <= merged[-1][1]:
                merged[-1][1] = max(merged[-1][1], e)
            else:
                merged.append([s, e])
        merged = [m for m in merged if m[0] < m[1]]
        h = hashlib.sha256()
        for s, e in merged:
            h.update(struct.pack("<QQ", s, e))
        c["entire_group"] = {"ranges": len(merged), "bytes": sum(e - s for s, e in merged), "sha256": h.hexdigest()}
    gold = {"n_regions": n_regions, "n_intervals": int(len(all_fids)), "n_distinct_root_fids": int(len(np.unique(all_fids))),
            "gof_records": int(len(gof)), "gff_len": int(gof["e"].max()), "cases": cases}
    json.dump(gold, open(os.path.join(d, "golden.json"), "w"), indent=1, sort_keys=True)
    print(name, {k: (v["pairs"], v["roots"]) for k, v in cases.items()}, gold["n_intervals"], gold["n_distinct_root_fids"])


def quirk_fixture():
    """tests/golden/quirk: the fixture and the naive evaluator's answers (see naive_eval.py)"""
    import sys

    sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
    import gffx_b200 as g
    from tests import _oracle
    from tests.golden import naive_eval as nv

    d = os.path.join(HERE, "quirk")
    os.makedirs(d, exist_ok=True)
    gff = os.path.join(d, "quirk.gff")
    g.synth_gff(gff, seed=20260099, n_genes=40, quirks=True, n_scaffolds=12)
    _oracle.build_index(gff)                            # the index writer is pinned by the byte-compatibility tests of tests/test_host.py
    for ext in ("fts", "prt", "a2f", "atn"):            # not read by the hot path
        os.remove(f"{gff}.{ext}")
    ix = nv.Index(gff)
    rng = np.random.default_rng(99)
    # regions placed around the fixture's own root intervals: every boundary relation at offsets -1 / 0 / +1
    lines_i, lines_c = ["track name=quirk description=\"intersect dialect\"\n", "# a comment\n", "\n"], ["# depth / coverage dialect\n"]
    ivs = []

    def walk(n, c):
        if n is None:
            return
        ivs.extend((c, iv[0], iv[1]) for iv in n.intervals)
        walk(n.left, c)
        walk(n.right, c)

    for c, t in ix.chr_entries.items():
        walk(t, c)
    ivs.sort()
    for k, (c, s, e) in enumerate(ivs):
        name = ix.num_to_seqid[c]
        cand = [(s, e), (s + 1, e), (s, e - 1), (max(s - 1, 0), e + 1), (e, e + 50), (e - 1, e + 50), (max(s - 50, 0), s), (max(s - 50, 0), s + 1),
                (s + (e - s) // 3, s + (e - s) // 2), (max(s - 20000, 0), s + 150)]
        for qs, qe in (cand if k % 3 == 0 else cand[k % 10:k % 10 + 2]):
            sep = "\t" if k % 2 else " "
            lines_i.append(f"{name}{sep}{qs}{sep}{qe}\tr{k}\t0\t+\n" if k % 5 == 0 else f"{name}{sep}{qs}{sep}{qe}\n")
            lines_c.append(f"{name}\t{qs}\t{qe}\n")
        if k % 7 == 0:                                   # inverted / empty regions: kept by intersect, dropped by depth / coverage
            lines_i.append(f"{name}\t{e}\t{s}\n")
            lines_i.append(f"{name}\t{s + 5}\t{s + 5}\n")
            lines_c.append(f"{name}\t{e}\t{s}\n")
            lines_c.append(f"{name}\t{s + 5}\t{s + 5}\n")
    for name in ix.num_to_seqid[:3] + ix.num_to_seqid[-3:]:
        lines_i.append(f"{name}\t0\t4294967295\n")      # the whole sequence and far beyond
        lines_c.append(f"{name}\t0\t4294967295\r\n")    # CRLF: trim_end() saves the number (depth.rs:486)
        lines_i.append(f"{name}\t4000000000\t4100000000\n")
        lines_c.append(f"{name} 4000000000   4100000000\n")
    lines_i += ["chrNope\tabc\t-5\n", "chr1\t10\n", "chr1 5000000 5001000 foreign\n"]
    lines_c += ["chrNope\t1\t2\n", "chr1\t10\n", "chr1\tx\t20\n", "chr1\t+100\t+2000000\n", "chr1\t5\t99999999999\n", "chr1 5000000 5001000\n"]
    for _ in range(60):                                  # background: 20 kb windows and reads anywhere
        c = int(rng.integers(0, 24))
        s = int(rng.integers(0, 50_000_000))
        ln = int(rng.choice([150, 20000, 3_000_000]))
        lines_i.append(f"{ix.num_to_seqid[c]}\t{s}\t{s + ln}\n")
        lines_c.append(f"{ix.num_to_seqid[c]}\t{s}\t{s + ln}\n")
    lines_i[-1] = lines_i[-1].rstrip("\n")               # last line without a newline
    bed_i, bed_c = "".join(lines_i).encode(), "".join(lines_c).encode()
    open(os.path.join(d, "intersect.bed"), "wb").write(bed_i)
    open(os.path.join(d, "counts.bed"), "wb").write(bed_c)
    regions = nv.parse_bed_file(bed_i, ix.seqid_to_num)
    gold = {"n_lines_gff": ix.gff.count(b"\n"), "n_regions_intersect": len(regions), "intersect": {}}
    for mode, mname in ((0, "contained"), (1, "contains_region"), (2, "overlap")):
        for invert in (False, True):
            for entire, types, tname in ((False, None, "plain"), (True, None, "e"), (False, "gene,mRNA,exon", "T"), (True, "exon, CDS ,,nosuchtype", "eT")):
                r = nv.intersect_run(ix, regions, mode, invert, entire, types)
                rd, td = nv.ranges_digest(ix, r["ranges"])
                gold["intersect"][f"{mname}{'+invert' if invert else ''}/{tname}"] = {
                    "pairs": len(r["pairs"]), "pair_checksum": nv.pair_checksum(r["pairs"]), "roots": len(r["roots"]),
                    "ranges": len(r["ranges"]), "bytes": sum(e - s for s, e in r["ranges"]), "ranges_sha256": rd, "text_sha256": td,
                    "warn": r["warn"], "first_ranges": [list(x) for x in r["ranges"][:4]]}
                print(mname, invert, tname, len(r["pairs"]), len(r["roots"]), len(r["ranges"]))
    cregs = nv.parse_bed_counts(bed_c, ix.seqid_to_num)
    gold["n_regions_counts"] = len(cregs)
    drows, damb = nv.depth_run(ix, cregs)
    crows, camb = nv.coverage_run(ix, cregs)
    gold["depth_rows"], gold["coverage_rows"] = len(drows), len(crows)
    gold["chrom_order_dependent_ids"] = sorted(x.decode() for x in damb | camb)
    with lzma.open(os.path.join(d, "depth.tsv.xz"), "wb", preset=9) as f:
        f.write(nv.depth_tsv(drows))
    with lzma.open(os.path.join(d, "coverage.tsv.xz"), "wb", preset=9) as f:
        f.write(nv.coverage_tsv(crows))
    json.dump(gold, open(os.path.join(d, "golden.json"), "w"), indent=1, sort_keys=True)
    with open(gff, "rb") as f, lzma.open(gff + ".xz", "wb", preset=9) as gz:
        gz.write(f.read())
    os.remove(gff)
    print("quirk:", gold["n_lines_gff"], "lines,", len(regions), "/", len(cregs), "regions,", len(drows), "depth rows,", len(crows), "coverage rows;",
          "order-dependent chr:", gold["chrom_order_dependent_ids"])


if __name__ == "__main__":
    import sys

    what = sys.argv[1:] or ["ps", "at", "dm", "quirk"]
    if "ps" in what:
        main()
    if "at" in what:
        real_index("at", 40000, 20260101)
    if "dm" in what:
        real_index("dm", 40000, 20260102)
    if "quirk" in what:
        quirk_fixture()
