"""Writers of small SAM / BAM files for the tests (SAMv1 specification: §1.4 text, §4.1 BGZF, §4.2 BAM), plus the expected
(tid, pos, reference_end) of every record as the test itself derives it — so the product's reader and the oracle's are each
compared with a third statement of the rules (src/commands/depth.rs:335-366 of the reference for the skips)."""
import random
import struct
import zlib

OPS = "MIDNSHP=XB"
REF_OPS = set("MDN=X")


def bgzf_block(data: bytes, level=6) -> bytes:
    assert len(data) <= 0xFF00
    c = zlib.compressobj(level, zlib.DEFLATED, -15)
    comp = c.compress(data) + c.flush()
    bsize = 12 + 6 + len(comp) + 8 - 1
    head = b"\x1f\x8b\x08\x04" + b"\0\0\0\0" + b"\0\xff" + struct.pack("<H", 6) + b"BC" + struct.pack("<HH", 2, bsize)
    return head + comp + struct.pack("<II", zlib.crc32(data) & 0xFFFFFFFF, len(data))


BGZF_EOF = bytes.fromhex("1f8b08040000000000ff0600424302001b0003000000000000000000")


def bgzf(data: bytes, block=0xFF00, levels=(6,), rng=None) -> bytes:
    out, o = [], 0
    while o < len(data):
        n = block if rng is None else rng.randint(1, block)
        out.append(bgzf_block(data[o:o + n], levels[len(out) % len(levels)]))
        o += n
    return b"".join(out) + BGZF_EOF


def cigar_ref_len(cigar):
    return sum(n for n, op in cigar if op in REF_OPS)


class Rec:
    def __init__(self, name, flag, tid, pos, cigar, l_seq=None, tags=b""):
        self.name, self.flag, self.tid, self.pos, self.cigar, self.tags = name, flag, tid, pos, cigar, tags
        q = sum(n for n, op in cigar if op in "MIS=X")
        self.l_seq = q if l_seq is None else l_seq

    def expected(self):
        """(tid, pos, end) or None, by the rules of depth.rs:335-366 on htslib's record"""
        if self.flag & 4 or self.tid < 0 or self.pos < 0:
            return None
        rlen = cigar_ref_len(self.cigar)
        return (self.tid, self.pos, self.pos + (rlen if rlen else 1))


def bam_record(r: Rec) -> bytes:
    name = r.name.encode() + b"\0"
    cig = r.cigar
    tags = r.tags
    if len(cig) > 65535:   # the real CIGAR goes to CG:B,I behind a placeholder (SAM spec §4.2.2)
        real = b"".join(struct.pack("<I", (n << 4) | OPS.index(op)) for n, op in cig)
        tags = tags + b"CGBI" + struct.pack("<I", len(cig)) + real
        cig = [(r.l_seq, "S"), (cigar_ref_len(r.cigar), "N")]
    cigb = b"".join(struct.pack("<I", (n << 4) | OPS.index(op)) for n, op in cig)
    seq = bytes((r.l_seq + 1) // 2)
    qual = b"\xff" * r.l_seq
    body = struct.pack("<iiBBHHHIiii", r.tid, r.pos, len(name), 30, 4680, len(cig), r.flag, r.l_seq, -1, -1, 0) + name + cigb + seq + qual + tags
    return struct.pack("<I", len(body)) + body


def bam_bytes(targets, recs, header_text=None) -> bytes:
    text = header_text if header_text is not None else "@HD\tVN:1.6\n" + "".join(f"@SQ\tSN:{n}\tLN:{l}\n" for n, l in targets)
    t = text.encode()
    out = b"BAM\1" + struct.pack("<I", len(t)) + t + struct.pack("<I", len(targets))
    for n, l in targets:
        nb = n.encode() + b"\0"
        out += struct.pack("<I", len(nb)) + nb + struct.pack("<I", l)
    return out + b"".join(bam_record(r) for r in recs)


def sam_text(targets, recs, crlf=False) -> bytes:
    nl = "\r\n" if crlf else "\n"
    lines = ["@HD\tVN:1.6\tSO:unsorted"] + [f"@SQ\tSN:{n}\tLN:{l}" for n, l in targets] + ["@PG\tID:test\tPN:test", "@CO\tSN:not_a_target"]
    for r in recs:
        rname = targets[r.tid][0] if 0 <= r.tid < len(targets) else "*"
        cig = "".join(f"{n}{op}" for n, op in r.cigar) or "*"
        seq = "A" * r.l_seq if r.l_seq else "*"
        lines.append("\t".join([r.name, str(r.flag), rname, str(r.pos + 1), "30", cig, "*", "0", "0", seq, "*" if not r.l_seq else "I" * r.l_seq]) +
                     ("\tNM:i:0" if r.pos % 2 else ""))
    return (nl.join(lines) + nl).encode()


def random_records(rng: random.Random, targets, n, max_pos):
    recs = []
    for i in range(n):
        k = rng.random()
        tid = rng.randrange(len(targets)) if k > 0.05 else -1
        flag = rng.choice([0, 16, 99, 147, 256, 2048, 1024])
        if k < 0.10:
            flag |= 4   # unmapped (may still carry a position: the mate's)
        ops = []
        for _ in range(rng.randint(0, 6)):
            ops.append((rng.randint(1, 200), rng.choice("MIDNSHP=X")))
        if rng.random() < 0.03:
            ops = []            # CIGAR '*': reference length 0 -> end = pos + 1
        if rng.random() < 0.03:
            ops = [(rng.randint(1, 50), "S"), (rng.randint(1, 50), "I")]   # nothing consumes the reference
        pos = rng.randint(0, max_pos) if rng.random() > 0.02 else -1      # SAM POS 0
        recs.append(Rec(f"r{i}", flag, tid, pos, ops))
    return recs
