// inflate.h — a self-contained DEFLATE (RFC 1951) decoder and CRC-32 for BGZF blocks (the container of BAM files and of
// bgzip-ed SAM).  The reference reads alignments through rust-htslib (src/commands/depth.rs:309, src/commands/coverage.rs:132),
// which is absent from this image; this is the engine's own decoder, so the library links nothing beyond libstdc++ / cudart.
// The decoder is table driven: a 10-bit primary table for the literal/length code, 8 bits for distances, canonical
// count/symbol arrays for the (rare) longer codes.  One BGZF block is at most 64 KB of output, so blocks are inflated side by
// side on the host threads (reads.cpp).
#pragma once
#include <cstddef>
#include <cstdint>
#include <cstring>

namespace gffx {

struct InflateTables {
    uint16_t lit[1 << 10];   // sym << 4 | code length (1..10); 0: a longer code
    uint16_t dst[1 << 8];
    uint16_t lcnt[16], dcnt[16];
    uint16_t lsym[288], dsym[32];
};

namespace inflate_detail {

constexpr int LIT_BITS = 10, DST_BITS = 8;

// canonical Huffman code from code lengths: count[len], symbols ordered by (len, symbol); the primary table holds every
// code of at most `bits` bits at all the indices whose low bits are the (bit-reversed) code.  Returns false when the lengths
// over-subscribe the code space.
inline bool build(const uint8_t* len, int n, uint16_t* cnt, uint16_t* sym, uint16_t* table, int bits) {
    for (int i = 0; i < 16; i++) cnt[i] = 0;
    for (int i = 0; i < n; i++) cnt[len[i]]++;
    cnt[0] = 0;
    int left = 1;
    for (int l = 1; l < 16; l++) {
        left <<= 1;
        left -= cnt[l];
        if (left < 0) return false;
    }
    uint16_t offs[16];
    offs[1] = 0;
    for (int l = 1; l < 15; l++) offs[l + 1] = (uint16_t)(offs[l] + cnt[l]);
    for (int i = 0; i < n; i++)
        if (len[i]) sym[offs[len[i]]++] = (uint16_t)i;
    memset(table, 0, sizeof(uint16_t) << bits);
    uint32_t code = 0, idx = 0;
    for (int l = 1; l <= bits; l++) {
        for (uint32_t k = 0; k < cnt[l]; k++, code++, idx++) {
            uint32_t rev = 0;
            for (int b = 0; b < l; b++) rev |= ((code >> b) & 1u) << (l - 1 - b);
            const uint16_t e = (uint16_t)((sym[idx] << 4) | l);
            for (uint32_t f = rev; f < (1u << bits); f += 1u << l) table[f] = e;
        }
        code <<= 1;
    }
    return true;
}

struct Bits {
    const uint8_t* p;
    const uint8_t* end;
    uint64_t buf = 0;
    int n = 0;          // valid bits in buf; bits past the input read as zeros and are counted in `over`
    int over = 0;
    void refill() {
        while (n <= 56) {
            if (p < end) buf |= (uint64_t)*p++ << n;
            else over += 8;
            n += 8;
        }
    }
    uint32_t peek(int k) const { return (uint32_t)(buf & ((1ull << k) - 1)); }
    void drop(int k) { buf >>= k; n -= k; }
    uint32_t take(int k) { const uint32_t v = peek(k); drop(k); return v; }
    bool overrun() const { return over > n; }   // consumed bits that were never there
};

// one symbol of a code whose primary table has `bits` index bits
inline int decode(Bits& B, const uint16_t* table, int bits, const uint16_t* cnt, const uint16_t* sym) {
    const uint16_t e = table[B.peek(bits)];
    if (e & 15) { B.drop(e & 15); return e >> 4; }
    int code = 0, first = 0, index = 0;   // longer than the table's reach: canonical walk, a bit at a time
    for (int l = 1; l <= 15; l++) {
        code |= (int)((B.buf >> (l - 1)) & 1u);
        const int count = cnt[l];
        if (code - count < first) { B.drop(l); return sym[index + (code - first)]; }
        index += count; first += count;
        first <<= 1; code <<= 1;
    }
    return -1;
}

}  // namespace inflate_detail

// Inflates one raw DEFLATE stream.  Returns the number of bytes written, or -1 (malformed stream) / -2 (output too small).
inline long inflate_raw(const uint8_t* in, size_t in_len, uint8_t* out, size_t out_cap, InflateTables& T) {
    using namespace inflate_detail;
    static const uint16_t LBASE[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
    static const uint8_t LEXT[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
    static const uint16_t DBASE[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
    static const uint8_t DEXT[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};
    static const uint8_t ORDER[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
    Bits B{in, in + in_len};
    size_t o = 0;
    for (;;) {
        B.refill();
        const uint32_t last = B.take(1), type = B.take(2);
        if (type == 0) {   // stored: back to the byte stream
            B.drop(B.n & 7);
            const uint8_t* q = B.p - (B.n - B.over) / 8;   // unread whole input bytes sit in the buffer
            if (B.overrun() || B.end - q < 4) return -1;
            const uint32_t len = q[0] | (q[1] << 8), nlen = q[2] | (q[3] << 8);
            q += 4;
            if ((len ^ 0xFFFFu) != nlen || (size_t)(B.end - q) < len) return -1;
            if (o + len > out_cap) return -2;
            memcpy(out + o, q, len);
            o += len;
            B.p = q + len; B.buf = 0; B.n = 0; B.over = 0;
        } else if (type == 1 || type == 2) {
            uint8_t lens[320];
            int nl, nd;
            if (type == 1) {
                nl = 288; nd = 30;
                for (int i = 0; i < 144; i++) lens[i] = 8;
                for (int i = 144; i < 256; i++) lens[i] = 9;
                for (int i = 256; i < 280; i++) lens[i] = 7;
                for (int i = 280; i < 288; i++) lens[i] = 8;
                for (int i = 0; i < 30; i++) lens[288 + i] = 5;
            } else {
                nl = (int)B.take(5) + 257; nd = (int)B.take(5) + 1;
                const int nc = (int)B.take(4) + 4;
                if (nl > 286 || nd > 30) return -1;
                uint8_t cl[19] = {0};
                B.refill();
                for (int i = 0; i < nc; i++) {
                    cl[ORDER[i]] = (uint8_t)B.take(3);
                    if ((i & 7) == 7) B.refill();
                }
                uint16_t ccnt[16], csym[19], ctab[1 << 7];
                if (!build(cl, 19, ccnt, csym, ctab, 7)) return -1;
                int i = 0;
                while (i < nl + nd) {
                    B.refill();
                    const int s = decode(B, ctab, 7, ccnt, csym);
                    if (s < 0) return -1;
                    if (s < 16) { lens[i++] = (uint8_t)s; continue; }
                    int rep;
                    uint8_t v = 0;
                    if (s == 16) {
                        if (i == 0) return -1;
                        v = lens[i - 1];
                        rep = 3 + (int)B.take(2);
                    } else if (s == 17) rep = 3 + (int)B.take(3);
                    else rep = 11 + (int)B.take(7);
                    if (i + rep > nl + nd) return -1;
                    while (rep--) lens[i++] = v;
                }
                if (lens[256] == 0) return -1;   // no end-of-block code
                // the distance lengths follow the literal/length ones: move them to a fixed place
                memmove(lens + 288, lens + nl, (size_t)nd);
            }
            if (!build(lens, nl, T.lcnt, T.lsym, T.lit, LIT_BITS)) return -1;
            if (!build(lens + 288, nd, T.dcnt, T.dsym, T.dst, DST_BITS)) return -1;
            for (;;) {
                // fast path, away from both ends (16 input bytes, 320 output bytes): one 8-byte refill (>= 56 valid bits) pays for up to three literals or for one
                // whole length / distance pair (15 + 5 + 15 + 13 bits); matches are copied eight bytes at a time
                bool block_end = false;
                while (B.end - B.p >= 16 && o + 320 <= out_cap) {
                    uint64_t w;
                    memcpy(&w, B.p, 8);
                    B.buf |= w << B.n;
                    B.p += (63 - B.n) >> 3;
                    B.n |= 56;
                    int s = decode(B, T.lit, LIT_BITS, T.lcnt, T.lsym);
                    if (s < 256) {
                        if (s < 0) return -1;
                        out[o++] = (uint8_t)s;
                        s = decode(B, T.lit, LIT_BITS, T.lcnt, T.lsym);
                        if (s < 256) {
                            if (s < 0) return -1;
                            out[o++] = (uint8_t)s;
                            s = decode(B, T.lit, LIT_BITS, T.lcnt, T.lsym);
                            if (s < 256) {
                                if (s < 0) return -1;
                                out[o++] = (uint8_t)s;
                                continue;
                            }
                        }
                        if (B.n < 48) {   // a length / distance pair follows the literals: top the buffer up again (>= 9 input bytes are left)
                            memcpy(&w, B.p, 8);
                            B.buf |= w << B.n;
                            B.p += (63 - B.n) >> 3;
                            B.n |= 56;
                        }
                    }
                    if (s == 256) { block_end = true; break; }
                    s -= 257;
                    if (s >= 29) return -1;
                    const uint32_t len = LBASE[s] + B.take(LEXT[s]);
                    const int d = decode(B, T.dst, DST_BITS, T.dcnt, T.dsym);
                    if (d < 0 || d >= 30) return -1;
                    const uint32_t dist = DBASE[d] + B.take(DEXT[d]);
                    if (dist > o) return -1;
                    const uint8_t* src = out + o - dist;
                    uint8_t* dstp = out + o;
                    if (dist >= 8) {
                        for (uint32_t k = 0; k < len; k += 8) memcpy(dstp + k, src + k, 8);   // may run up to 7 bytes past len: inside the 320-byte margin
                    } else {
                        for (uint32_t k = 0; k < len; k++) dstp[k] = src[k];
                    }
                    o += len;
                }
                if (block_end) break;
                B.refill();
                if (B.overrun()) return -1;
                int s = decode(B, T.lit, LIT_BITS, T.lcnt, T.lsym);
                if (s < 0) return -1;
                if (s < 256) {
                    if (o >= out_cap) return -2;
                    out[o++] = (uint8_t)s;
                    continue;
                }
                if (s == 256) break;
                s -= 257;
                if (s >= 29) return -1;
                const uint32_t len = LBASE[s] + B.take(LEXT[s]);
                const int d = decode(B, T.dst, DST_BITS, T.dcnt, T.dsym);
                if (d < 0 || d >= 30) return -1;
                const uint32_t dist = DBASE[d] + B.take(DEXT[d]);
                if (dist > o) return -1;
                if (o + len > out_cap) return -2;
                const uint8_t* src = out + o - dist;
                uint8_t* dstp = out + o;
                if (dist >= len) memcpy(dstp, src, len);
                else for (uint32_t k = 0; k < len; k++) dstp[k] = src[k];   // overlapping: byte by byte, forwards
                o += len;
            }
            if (B.overrun()) return -1;
        } else {
            return -1;
        }
        if (last) break;
    }
    return (long)o;
}

// CRC-32 (IEEE 802.3, the gzip trailer's), slicing by 8
struct Crc32Table {
    uint32_t t[8][256];
    Crc32Table() {
        for (uint32_t i = 0; i < 256; i++) {
            uint32_t c = i;
            for (int k = 0; k < 8; k++) c = (c >> 1) ^ (0xEDB88320u & (0u - (c & 1u)));
            t[0][i] = c;
        }
        for (uint32_t i = 0; i < 256; i++)
            for (int k = 1; k < 8; k++) t[k][i] = (t[k - 1][i] >> 8) ^ t[0][t[k - 1][i] & 0xFF];
    }
};
inline uint32_t crc32_bytes(const uint8_t* p, size_t n) {
    static const Crc32Table tab;
    const uint32_t(*T)[256] = tab.t;
    uint32_t c = 0xFFFFFFFFu;
    while (n >= 8) {
        uint32_t a, b;
        memcpy(&a, p, 4);       // little-endian hosts (x86-64, aarch64 as built here)
        memcpy(&b, p + 4, 4);
        a ^= c;
        c = T[7][a & 0xFF] ^ T[6][(a >> 8) & 0xFF] ^ T[5][(a >> 16) & 0xFF] ^ T[4][a >> 24] ^ T[3][b & 0xFF] ^ T[2][(b >> 8) & 0xFF] ^ T[1][(b >> 16) & 0xFF] ^
            T[0][b >> 24];
        p += 8; n -= 8;
    }
    while (n--) c = (c >> 1 >> 7) ^ T[0][(c ^ *p++) & 0xFF];
    return ~c;
}

}  // namespace gffx
