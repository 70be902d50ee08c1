"""The reader's own DEFLATE decoder and CRC-32 (gffx_b200/csrc/inflate.h, through gffx_inflate_raw / gffx_crc32) against zlib:
every compression level and strategy over data of several kinds and sizes up to a BGZF block (stored, fixed and dynamic
blocks; literal runs, short and long matches, overlapping copies), exact output sizes as BGZF gives them, too-small outputs,
and damaged / truncated streams (an error or a wrong answer, never a crash).  CPU only: the library loads without a GPU."""
import ctypes as C
import random
import zlib

import numpy as np
import pytest

from gffx_b200 import _lib


def inflate(comp: bytes, cap: int):
    out = (C.c_uint8 * max(cap, 1))()
    src = (C.c_uint8 * max(len(comp), 1)).from_buffer_copy(comp or b"\0")
    r = _lib.lib.gffx_inflate_raw(src, len(comp), out, cap)
    return r, bytes(out[:max(r, 0)])


def deflate_raw(data: bytes, level: int, strategy=zlib.Z_DEFAULT_STRATEGY) -> bytes:
    c = zlib.compressobj(level, zlib.DEFLATED, -15, 9, strategy)
    return c.compress(data) + c.flush()


def samples(rng: random.Random):
    nrng = np.random.default_rng(rng.randrange(1 << 30))
    for n in (0, 1, 7, 8, 15, 16, 17, 255, 256, 319, 320, 321, 1000, 4096, 40000, 0xFF00):
        yield "random", nrng.integers(0, 256, n, dtype=np.uint8).tobytes()
        yield "few symbols", nrng.integers(0, 4, n, dtype=np.uint8).tobytes()
        yield "zeros", bytes(n)
        yield "period 3", (b"abc" * (n // 3 + 1))[:n]
        yield "period 9", (b"abcdefghi" * (n // 9 + 1))[:n]
        text = b"".join(b"read%07d\tchr%d\t%d\t60\t150M\n" % (i, i % 23, i * 37) for i in range(n // 20 + 1))[:n]
        yield "text", text
        q = nrng.integers(2, 41, n, dtype=np.uint8)
        yield "qualities", (np.repeat(q[::3], 3)[:n] + 33).astype(np.uint8).tobytes()


def test_every_level_and_strategy_round_trips():
    rng = random.Random(5)
    n_cases = 0
    for kind, data in samples(rng):
        for level in (0, 1, 4, 6, 9):
            for strategy in (zlib.Z_DEFAULT_STRATEGY, zlib.Z_FIXED, zlib.Z_HUFFMAN_ONLY, zlib.Z_RLE):
                comp = deflate_raw(data, level, strategy)
                r, out = inflate(comp, len(data))            # the exact size, as a BGZF block's ISIZE gives it
                assert r == len(data) and out == data, (kind, len(data), level, strategy)
                r, out = inflate(comp, len(data) + 1000)
                assert r == len(data) and out == data, (kind, len(data), level, strategy, "roomy")
                if len(data) > 1:
                    assert inflate(comp, len(data) - 1)[0] == -2
                assert _lib.lib.gffx_crc32((C.c_uint8 * max(len(data), 1)).from_buffer_copy(data or b"\0"), len(data)) == zlib.crc32(data) & 0xFFFFFFFF
                n_cases += 1
    assert n_cases > 2000


def test_multi_block_streams_and_flushes():
    rng = random.Random(6)
    for kind, data in samples(rng):
        if len(data) < 1000:
            continue
        c = zlib.compressobj(6, zlib.DEFLATED, -15)
        comp = b""
        for k in range(0, len(data), 777):
            comp += c.compress(data[k:k + 777])
            comp += c.flush(rng.choice((zlib.Z_SYNC_FLUSH, zlib.Z_FULL_FLUSH, zlib.Z_NO_FLUSH)))
        comp += c.flush()
        r, out = inflate(comp, len(data))
        assert r == len(data) and out == data, kind


def test_damaged_streams_do_not_crash():
    rng = random.Random(7)
    for kind, data in samples(rng):
        if not 256 <= len(data) <= 40000:
            continue
        comp = deflate_raw(data, 6)
        for _ in range(12):
            bad = bytearray(comp)
            for _ in range(rng.randint(1, 4)):
                bad[rng.randrange(len(bad))] ^= 1 << rng.randrange(8)
            r, out = inflate(bytes(bad), len(data))          # any answer but a crash; a right-sized answer may still be wrong bytes
            assert -2 <= r <= len(data)
            cut = comp[:rng.randrange(len(comp))]
            r, _ = inflate(cut, len(data))
            assert -2 <= r <= len(data)
