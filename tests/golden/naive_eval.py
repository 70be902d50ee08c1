#!/usr/bin/env python3
"""A second, deliberately naive statement of GFFx's intersect / depth / coverage semantics — pure Python, O(L x Q),
written from the reference's Rust source one function at a time and sharing NO code with oracle/ (C++) or the engine
(CUDA).  It exists to pin the oracle: tests/golden/make_golden.py runs it once, in the build container, on the small
quirk fixture and commits what it produced; tests/test_golden_quirk.py then holds BOTH the oracle and the GPU engine to
those bytes.  (The Rust reference itself cannot be compiled in this image — no cargo/rustc — so "the reference's output"
is not obtainable; two independent readings that agree are the next best thing.  Parity stays "unpinned by the reference".)

Every function names the reference lines it restates (paths relative to the reference root).  Only the on-disk index
formats are read; nothing here knows about windows, rank spaces or virtual sequences.

TEST INFRASTRUCTURE ONLY: imported by tests/golden/make_golden.py and tests/test_golden_quirk.py's self-check.
"""
import hashlib
import json
import struct

MISSING = 0xFFFFFFFFFFFFFFFF
U32_MAX = 0xFFFFFFFF


# ---------------------------------------------------------------------------------------------------
# index files
# ---------------------------------------------------------------------------------------------------
class Node:
    __slots__ = ("center", "intervals", "left", "right")


def decode_tree(buf, pos):
    """bincode2 of Option<Box<Node<u32>>> (src/utils/tree.rs:13-23, src/utils/tree_io.rs:37-54): tag byte, then
    center u32, count u64, count x (start u32, end u32, root_fid u32), left, right"""
    tag = buf[pos]
    pos += 1
    if tag == 0:
        return None, pos
    assert tag == 1
    n = Node()
    n.center, cnt = struct.unpack_from("<IQ", buf, pos)
    pos += 12
    n.intervals = [struct.unpack_from("<III", buf, pos + 12 * i) for i in range(cnt)]
    pos += 12 * cnt
    n.left, pos = decode_tree(buf, pos)
    n.right, pos = decode_tree(buf, pos)
    return n, pos


class Index:
    """TreeIndexData (src/utils/tree_index.rs:12-82) + GofMap (src/index_loader/gof.rs:20-128) + the GFF bytes"""

    def __init__(self, gff_path):
        # load_sqs, src/index_loader/core.rs:19-34: BufRead::lines()
        raw = open(gff_path + ".sqs", "rb").read().decode("utf-8")
        names = raw.split("\n")
        if names and names[-1] == "":
            names.pop()
        self.num_to_seqid = [n[:-1] if n.endswith("\r") else n for n in names]
        self.seqid_to_num = {}
        for i, n in enumerate(self.num_to_seqid):
            self.seqid_to_num[n] = i          # a later duplicate overwrites (HashMap::insert)
        offsets = json.load(open(gff_path + ".rix"))
        rit = open(gff_path + ".rit", "rb").read()
        self.chr_entries = {}
        for i, off in enumerate(offsets):     # src/utils/tree_index.rs:64-79: tree i is sequence number i
            root, end = decode_tree(rit, off)
            assert end == (offsets[i + 1] if i + 1 < len(offsets) else len(rit))
            self.chr_entries[i] = root
        gof = open(gff_path + ".gof", "rb").read()
        assert len(gof) % 24 == 0
        self.gof_entries = [struct.unpack_from("<IIQQ", gof, 24 * i) for i in range(len(gof) // 24)]
        self.gof_index = {}
        for fid, _seq, s, e in self.gof_entries:   # GofMap::index, gof.rs:32-37: collect() keeps the LAST record of a fid
            self.gof_index[fid] = (s, e)
        try:
            self.gff = open(gff_path, "rb").read()
        except OSError:
            self.gff = None


def query_interval(node, start, end, out):
    """src/utils/tree.rs:98-121, literally — including the pruning on the node's center"""
    if node is None:
        return
    for iv in node.intervals:
        if iv[0] < end and iv[1] > start:
            out.append(iv)
    if start < node.center:
        query_interval(node.left, start, end, out)
    if end > node.center:
        query_interval(node.right, start, end, out)


# ---------------------------------------------------------------------------------------------------
# text helpers that mirror the Rust std calls
# ---------------------------------------------------------------------------------------------------
ASCII_WS = b" \t\n\r\x0c"                     # u8::is_ascii_whitespace
RUST_WS = {chr(c) for c in [0x09, 0x0A, 0x0B, 0x0C, 0x0D, 0x20, 0x85, 0xA0, 0x1680, 0x2028, 0x2029, 0x202F, 0x205F, 0x3000] + list(range(0x2000, 0x200B))}   # char::is_whitespace


def split_ascii_whitespace(b):
    out, cur = [], bytearray()
    for c in b:
        if c in ASCII_WS:
            if cur:
                out.append(bytes(cur))
                cur = bytearray()
        else:
            cur.append(c)
    if cur:
        out.append(bytes(cur))
    return out


def rust_trim(s):                              # str::trim (char::is_whitespace)
    a, b = 0, len(s)
    while a < b and s[a] in RUST_WS:
        a += 1
    while b > a and s[b - 1] in RUST_WS:
        b -= 1
    return s[a:b]


def rust_trim_end(s):
    b = len(s)
    while b > 0 and s[b - 1] in RUST_WS:
        b -= 1
    return s[:b]


def parse_u32_digits(b):
    """parse_u32_ascii (src/commands/intersect.rs:526-538) == parse_u32_fast (src/commands/depth.rs:86-95):
    non-empty, ASCII digits only, checked arithmetic"""
    if len(b) == 0:
        return None
    v = 0
    for c in b:
        if c < 0x30 or c > 0x39:
            return None
        v = v * 10 + (c - 0x30)
        if v > U32_MAX:
            return None
    return v


def parse_u32_std(s):
    """str::parse::<u32>: an optional '+', then at least one digit, no overflow"""
    if s.startswith("+"):
        s = s[1:]
    if not s or any(c < "0" or c > "9" for c in s):
        return None
    v = int(s)
    return v if v <= U32_MAX else None


class Abort(Exception):
    """the command ends with an error (anyhow::Error reaching main)"""


# ---------------------------------------------------------------------------------------------------
# intersect
# ---------------------------------------------------------------------------------------------------
def parse_bed_file(bed_bytes, seqid_map):
    """src/commands/intersect.rs:201-230"""
    regions = []
    for line in bed_bytes.split(b"\n"):
        if len(line) == 0 or line[0:1] == b"#":
            continue
        try:
            line_str = line.decode("utf-8")
        except UnicodeDecodeError:
            raise Abort("invalid utf-8 in BED")
        parts = split_ascii_whitespace(line_str.encode("utf-8"))
        if len(parts) < 3:
            continue
        seq = parts[0].decode("utf-8")
        if seq not in seqid_map:
            continue
        start = parse_u32_std(parts[1].decode("utf-8"))   # lexical_core::parse::<u32>: digits after at most one '+'
        end = parse_u32_std(parts[2].decode("utf-8"))
        if start is None or end is None:
            raise Abort("BED number")
        regions.append((seqid_map[seq], start, end))
    return regions


def query_features(index, regions, mode, invert):
    """src/commands/intersect.rs:105-169.  mode: 0 Contained, 1 ContainsRegion, 2 Overlap"""
    buckets = [[] for _ in range(len(index.seqid_to_num))]
    for chr_, s, e in regions:
        buckets[chr_].append((chr_, s, e))       # index out of range would panic in the reference
    results = []
    for seq_num, tree in index.chr_entries.items():
        if seq_num >= len(buckets):
            raise Abort("tree beyond seqid_to_num")
        chr_regs = buckets[seq_num]
        if not chr_regs:
            continue
        for _, rstart, rend in chr_regs:
            hits = []
            query_interval(tree, rstart, rend, hits)
            for iv in hits:
                if mode == 0:
                    keep = iv[0] >= rstart and iv[1] <= rend
                elif mode == 1:
                    keep = iv[0] <= rstart and iv[1] >= rend
                else:
                    keep = True
                if invert != keep:
                    results.append((iv[2], iv[0], iv[1]))
    return results


def gff_type_allowed(line, allow):
    """src/commands/intersect.rs:80-102"""
    off = 0
    for _ in range(2):
        i = line.find(b"\t", off)
        if i < 0:
            return False
        off = i + 1
    i2 = line.find(b"\t", off)
    if i2 < 0:
        return False
    try:
        return line[off:i2].decode("utf-8") in allow
    except UnicodeDecodeError:
        return False


def gff_line_overlaps_queries(line, ivmap, mode):
    """src/commands/intersect.rs:441-523"""
    cols, off = [], 0
    for _ in range(5):
        i = line.find(b"\t", off)
        if i < 0:
            return False
        cols.append(line[off:i])
        off = i + 1
    start = parse_u32_digits(cols[3])
    if start is None:
        return False
    end = parse_u32_digits(cols[4])
    if end is None:
        return False
    try:
        seq = cols[0].decode("utf-8")
    except UnicodeDecodeError:
        return False
    ivs = ivmap.get(seq)
    if ivs is None:
        return False
    for qs, qe in ivs:
        if mode == 0:
            keep = start >= qs and end <= qe
        elif mode == 1:
            keep = start <= qs and end >= qe
        else:
            keep = (qs <= start <= qe) or (qs <= end <= qe) or (start <= qs <= end) or (start <= qe <= end)
        if keep:
            return True
    return False


def write_gff_match_only_by_coords(index, blocks, ivmap, types_filter, mode, warn):
    """src/commands/intersect.rs:232-438 -> list of (abs_start, abs_end) in output order"""
    gff = index.gff
    file_len = len(gff)
    allow = None
    if types_filter is not None:
        allow = {rust_trim(t) for t in types_filter.split(",")}
        allow.discard("")
    parts = []
    for root, start, end in blocks:
        if start == MISSING:
            warn.append(root)
            continue
        s, e = start, min(end, file_len)
        if s >= e or e > file_len:
            continue
        src = gff[s:e]
        matched, pos = [], 0
        while pos < len(src):
            i = src.find(b"\n", pos)
            nl = i + 1 if i >= 0 else len(src)
            line = src[pos:nl]
            line_nocr = line[:-1] if line.endswith(b"\n") else line
            if len(line_nocr) != 0 and line_nocr[0:1] != b"#":
                ok = True
                if allow is not None and not gff_type_allowed(line_nocr, allow):
                    ok = False
                if ok and gff_line_overlaps_queries(line_nocr, ivmap, mode):
                    matched.append((start + pos, start + nl))
            pos = nl
        if matched:
            parts.append((start, matched))
    parts.sort(key=lambda p: p[0])               # sort_unstable_by_key on the block start: starts are distinct in a .gof
    return [r for _, m in parts for r in m]


def write_gff_output(index, blocks, warn):
    """src/utils/common.rs:188-245 -> merged (start, end) slices that are written"""
    file_len = len(index.gff)
    v = []
    for fid, s, e in blocks:
        if s == MISSING:
            warn.append(fid)
            continue
        v.append((s, e))
    v.sort(key=lambda x: x[0])
    merged = []
    if v:
        cs, ce = v[0]
        for s, e in v[1:]:
            if s <= ce:
                ce = max(ce, e)
            else:
                if cs < ce:
                    merged.append((cs, ce))
                cs, ce = s, e
        if cs < ce:
            merged.append((cs, ce))
    return [(so, eo) for so, eo in merged if so < eo and eo <= file_len]


def intersect_run(index, regions, mode, invert, entire_group, types):
    """the body of run() after BED parsing, src/commands/intersect.rs:586-654
    -> dict(pairs, roots, ranges, warn)"""
    feats = query_features(index, regions, mode, invert)
    roots = []
    seen = set()
    for root, _s, _e in feats:
        if root not in seen:
            seen.add(root)
            roots.append(root)
    blocks = []
    for r in roots:                               # GofMap::roots_to_offsets, src/index_loader/gof.rs:54-84
        if r in index.gof_index:
            s, e = index.gof_index[r]
            blocks.append((r, s, e))
        else:
            blocks.append((r, MISSING, MISSING))
    warn = []
    if not entire_group or types is not None:     # :619
        num_to_seq = {num: name for name, num in index.seqid_to_num.items()}
        ivmap = {}
        for chr_num, s, e in regions:
            name = num_to_seq.get(chr_num)
            if name is not None:
                ivmap.setdefault(name, []).append((s, e))
        ranges = write_gff_match_only_by_coords(index, blocks, ivmap, types, mode, warn)
    else:
        ranges = write_gff_output(index, blocks, warn)
    return {"pairs": feats, "roots": roots, "ranges": ranges, "warn": sorted(set(warn))}


# ---------------------------------------------------------------------------------------------------
# depth / coverage
# ---------------------------------------------------------------------------------------------------
def parse_bed_counts(bed_bytes, seqid_map):
    """the BED dialect of src/commands/depth.rs:449-495 == src/commands/coverage.rs:222-255"""
    offs = [0] + [i + 1 for i, b in enumerate(bed_bytes) if b == 0x0A]
    if offs[-1] != len(bed_bytes):
        offs.append(len(bed_bytes))
    regions = []
    for a, b in zip(offs, offs[1:]):
        if a >= b:
            continue
        line = bed_bytes[a:b]                      # keeps its '\n'
        if line[0:1] == b"#":
            continue
        fields, cur = [], bytearray()
        for c in line:
            if c == 0x09 or c == 0x20:
                if cur:
                    fields.append(bytes(cur))
                    cur = bytearray()
            else:
                cur.append(c)
        if cur:
            fields.append(bytes(cur))
        if len(fields) < 3:
            continue
        try:
            chrom = fields[0].decode("utf-8")
            s = parse_u32_std(fields[1].decode("utf-8"))
            e = parse_u32_std(rust_trim_end(fields[2].decode("utf-8")))
        except UnicodeDecodeError:
            continue
        if s is None or e is None or s >= e:
            continue
        if chrom not in seqid_map:
            continue
        regions.append((seqid_map[chrom], s, e))
    return regions


def fast_id(attrs):
    """src/commands/depth.rs:104-118 (bytes of a &str: the result is always valid UTF-8 when attrs is)"""
    b = attrs
    i = 0
    while i + 2 < len(b):
        if b[i] == 0x49 and b[i + 1] == 0x44 and b[i + 2] == 0x3D:
            j = i + 3
            while j < len(b) and b[j] not in (0x3B, 0x20, 0x09):
                j += 1
            return b[i + 3:j]
        i += 1
    return None


def block_features(gff_slice):
    """the line loop shared by compute_root_depth (src/commands/depth.rs:131-157) and compute_breadth_for_root
    (src/commands/coverage.rs:296-323) -> (id strings, id chroms, [(fstart0, fend0, id_idx)])"""
    ids, chroms, idx_of, feats = [], [], {}, []
    try:
        gff_slice.decode("utf-8")
    except UnicodeDecodeError:
        return ids, chroms, feats                   # `if let Ok(text) = str::from_utf8(..)`
    lines = gff_slice.split(b"\n")
    if lines and lines[-1] == b"":
        lines.pop()                                 # split_terminator
    for line in lines:
        if len(line) == 0 or line[0:1] == b"#":
            continue
        cols = line.split(b"\t", 8)                 # splitn(9, '\t')
        if len(cols) < 9:
            continue
        s1, e1 = parse_u32_digits(cols[3]), parse_u32_digits(cols[4])
        if s1 is None or e1 is None:
            continue
        if e1 == 0:
            continue
        if s1 > e1:
            s1, e1 = e1, s1
        fstart0 = max(s1 - 1, 0)
        fend0 = e1
        fid = fast_id(cols[8])
        if fid is not None:
            if fid not in idx_of:
                idx_of[fid] = len(ids)
                ids.append(fid)
                chroms.append(cols[0])
            feats.append((fstart0, fend0, idx_of[fid]))
    return ids, chroms, feats


def tree_hits_by_root(index, regions, need_valid_block):
    """routing shared by compute_hit_depth (src/commands/depth.rs:230-250, need_valid_block) and
    collect_by_root_from_bed (src/commands/coverage.rs:257-266) -> {root_fid: [(start,end), ...]} in region order"""
    by_root = {}
    for chr_, rstart, rend in regions:
        tree = index.chr_entries.get(chr_)
        if tree is None:
            continue
        hits = []
        query_interval(tree, rstart, rend, hits)
        seen = set()
        for h in hits:
            if h[2] in seen:
                continue
            seen.add(h[2])
            if need_valid_block:
                if h[2] not in index.gof_index:
                    continue
                s_off, e_off = index.gof_index[h[2]]
                if s_off == MISSING or e_off == MISSING or e_off <= s_off:
                    continue
            by_root.setdefault(h[2], []).append((rstart, rend))
    return by_root


def root_order(index, by_root):
    """The reference walks `by_root` in hash order and keeps the chromosome string of whichever root it meets first
    (depth.rs:266-271, coverage.rs:431-443): for an ID that sits under roots on different sequences the `chr` column is
    order-dependent.  Every deterministic restatement must pick an order; this one (and oracle/) takes the roots by
    ascending block start offset.  make_golden.py reports how many IDs that choice touches."""
    return sorted(by_root, key=lambda r: index.gof_index.get(r, (MISSING, MISSING))[0])


def merge_rows(out, partial, ambiguous):
    for id_, (chrom, s, e, v) in partial.items():
        if id_ in out:
            c0, s0, e0, v0 = out[id_]
            if c0 != chrom:
                ambiguous.add(id_)
            out[id_] = (c0, min(s0, s), max(e0, e), v0 + v)
        else:
            out[id_] = (chrom, s, e, v)


def compute_root_depth(gff_slice, regions):
    """src/commands/depth.rs:121-217; the bins (:163-191) only pre-select candidates and are left out: every feature line
    is tested against every region"""
    ids, chroms, feats = block_features(gff_slice)
    if not feats or not regions:
        return {}
    n = len(ids)
    min_s, max_e, depths = [U32_MAX] * n, [0] * n, [0] * n
    for rs, re in regions:
        hit_ids = set()
        for fs, fe, ii in feats:
            if max(fs, rs) < min(fe, re):            # overlaps(), :78-82
                hit_ids.add(ii)
                if fs < min_s[ii]:
                    min_s[ii] = fs
                if fe > max_e[ii]:
                    max_e[ii] = fe
        for ii in hit_ids:
            depths[ii] += 1
    out = {}
    for i, d in enumerate(depths):
        if d > 0:
            out[ids[i]] = (chroms[i], 0 if min_s[i] == U32_MAX else min_s[i], max_e[i], d)
    return out


def depth_run(index, regions, batch=100000):
    """process_bed's batching (src/commands/depth.rs:24,462-509) over already parsed regions + compute_hit_depth (:222-293)"""
    glob, ambiguous = {}, set()
    for b0 in range(0, len(regions), batch):
        by_root = tree_hits_by_root(index, regions[b0:b0 + batch], True)
        out = {}
        for root in root_order(index, by_root):
            s_off, e_off = index.gof_index[root]
            merge_rows(out, compute_root_depth(index.gff[s_off:e_off], by_root[root]), ambiguous)
        merge_rows(glob, out, ambiguous)
    return glob, ambiguous


def merge_intervals(ivs):
    """src/commands/coverage.rs:92-107"""
    if not ivs:
        return []
    ivs = sorted(ivs, key=lambda x: x[0])
    out = []
    cs, ce = ivs[0]
    for s, e in ivs[1:]:
        if s <= ce:
            if e > ce:
                ce = e
        else:
            out.append((cs, ce))
            cs, ce = s, e
    out.append((cs, ce))
    return out


def union_len(ivs):
    """src/commands/coverage.rs:110-121"""
    if not ivs:
        return 0
    ivs = sorted(ivs, key=lambda x: x[0])
    total = 0
    cs, ce = ivs[0]
    for s, e in ivs[1:]:
        if s <= ce:
            ce = max(ce, e)
        else:
            total += ce - cs
            cs, ce = s, e
    return total + (ce - cs)


def compute_breadth_for_root(gff_slice, cov):
    """src/commands/coverage.rs:277-377; the two-pointer sweep is replaced by testing every line against every merged
    coverage interval (same pieces: [max(fs,cs), min(fe,ce)) when non-empty)"""
    ids, chroms, lines = block_features(gff_slice)
    if not lines or not cov:
        return {}
    n = len(ids)
    min_s, max_e = [U32_MAX] * n, [0] * n
    pieces = [[] for _ in range(n)]
    for fs, fe, ii in lines:
        min_s[ii] = min(min_s[ii], fs)
        max_e[ii] = max(max_e[ii], fe)
        for cs, ce in cov:
            s, e = max(fs, cs), min(fe, ce)
            if e > s:
                pieces[ii].append((s, e))
    out = {}
    for i in range(n):
        length = max_e[i] - min_s[i] if max_e[i] > min_s[i] else 0
        breadth = union_len(pieces[i])
        if length > 0 or breadth > 0:
            out[ids[i]] = (chroms[i], min_s[i], max_e[i], breadth)
    return out


def coverage_run(index, regions):
    """collect_by_root_from_bed (src/commands/coverage.rs:208-274) + finalize_compute_breadth (:383-450)"""
    by_root = tree_hits_by_root(index, regions, False)
    glob, ambiguous = {}, set()
    for root in root_order(index, by_root):
        cov = merge_intervals(by_root[root])
        if root in index.gof_index:
            s_off, e_off = index.gof_index[root]
            if s_off != MISSING and e_off != MISSING and e_off > s_off:
                merge_rows(glob, compute_breadth_for_root(index.gff[s_off:e_off], cov), ambiguous)
    return glob, ambiguous


def fmt_fraction(breadth, length):
    """`{:.6}` of breadth as f64 / length as f64 (src/commands/coverage.rs:466-472): correctly rounded, ties to even —
    decimal arithmetic on the exact binary value of the quotient"""
    from decimal import ROUND_HALF_EVEN, Decimal

    q = (breadth / length) if length > 0 else 0.0     # IEEE double division, as in Rust
    return str(Decimal(q).quantize(Decimal("0.000001"), rounding=ROUND_HALF_EVEN))


def depth_tsv(rows):
    """write_depth_results (src/commands/depth.rs:515-545); rows sorted bytewise (the reference's order is hash order)"""
    body = sorted(id_ + b"\t" + c + b"\t%d\t%d\t%d\n" % (s, e, d) for id_, (c, s, e, d) in rows.items())
    return b"id\tchr\tstart\tend\tdepth\n" + b"".join(body)


def coverage_tsv(rows):
    """write_breadth_results (src/commands/coverage.rs:453-490); rows sorted bytewise"""
    body = []
    for id_, (c, s, e, b) in rows.items():
        length = e - s if e > s else 0                 # end.saturating_sub(start)
        body.append(id_ + b"\t" + c + b"\t%d\t%d\t%d\t" % (s, e, b) + fmt_fraction(b, length).encode() + b"\n")
    return b"id\tchr\tstart\tend\tbreadth\tfraction\n" + b"".join(sorted(body))


# ---------------------------------------------------------------------------------------------------
def pair_checksum(pairs):
    """order-independent 64-bit checksum of a pair multiset (same mix as make_golden.py's numpy version)"""
    M = (1 << 64) - 1
    tot = 0
    for fid, s, e in pairs:
        x = (fid * 0x9E3779B97F4A7C15 + s * 0xBF58476D1CE4E5B9 + e * 0x94D049BB133111EB) & M
        x ^= x >> 31
        x = (x * 0x9E3779B97F4A7C15) & M
        tot = (tot + x) & M
    return tot


def ranges_digest(index, ranges):
    h = hashlib.sha256()
    for s, e in ranges:
        h.update(struct.pack("<QQ", s, e))
    t = hashlib.sha256()
    for s, e in ranges:
        t.update(index.gff[s:e])
    return h.hexdigest(), t.hexdigest()
