/* inflate_oracle.c -- a plain RFC 1951 decoder.
 * TEST INFRASTRUCTURE ONLY (see oracle.h).  Its table builder and bit-serial symbol decoder (build / decode below)
 * follow the structure of Mark Adler's puff (the reference vendors it as lib/zlib/contrib/puff/puff.c:235-258,340-379):
 * a second, deliberately simple implementation of the format that shares no code with zlib's inflate.c.  The product's
 * own decoder (k_verify.cu) is a different design (sorted left-justified codes + binary search, built by a warp).  The reference decodes with zlib's inflate through
 * inflateGZIP (src/unzip.c:254-370), which is out of scope for the engine; this decoder exists so
 * the tests can (a) verify round trips without trusting a single implementation and (b) dissect a
 * stream (block types, token counts, where it breaks).  It is checked against the reference's
 * inflate (oracle/_ref/librefz.so) in tests/test_oracle.py.
 * Format facts: RFC 1951 3.2; tables as in lib/zlib/trees.c:62-72 and lib/zlib/inftrees.c:40-60.
 */
#include "oracle.h"
#include <string.h>

typedef struct { const uint8_t *in; size_t len; uint64_t bit; int err; } bits_t;

static uint32_t getbits(bits_t *b, int n)            /* LSB-first, RFC 1951 3.1.1 */
{
    uint32_t v = 0;
    for (int i = 0; i < n; i++, b->bit++) {
        size_t byte = (size_t)(b->bit >> 3);
        if (byte >= b->len) { b->err = 1; return 0; }
        v |= (uint32_t)((b->in[byte] >> (b->bit & 7)) & 1) << i;
    }
    return v;
}

typedef struct { uint16_t count[16], symbol[288]; } huff_t;

/* canonical code from lengths (RFC 1951 3.2.2); returns <0 over-subscribed, >0 incomplete, 0 complete */
static int build(huff_t *h, const uint8_t *len, int n)
{
    uint16_t offs[16];
    memset(h->count, 0, sizeof(h->count));
    for (int i = 0; i < n; i++) h->count[len[i]]++;
    if (h->count[0] == n) return 0;
    int left = 1;
    for (int l = 1; l <= 15; l++) { left <<= 1; left -= h->count[l]; if (left < 0) return left; }
    offs[1] = 0;
    for (int l = 1; l < 15; l++) offs[l + 1] = (uint16_t)(offs[l] + h->count[l]);
    for (int i = 0; i < n; i++) if (len[i]) h->symbol[offs[len[i]]++] = (uint16_t)i;
    return left;
}

static int decode(bits_t *b, const huff_t *h)        /* codes are packed MSB-first (RFC 1951 3.1.1) */
{
    int code = 0, first = 0, index = 0;
    for (int l = 1; l <= 15; l++) {
        code |= (int)getbits(b, 1);
        if (b->err) return -1;
        int count = h->count[l];
        if (code - count < first) return h->symbol[index + (code - first)];
        index += count; first += count; first <<= 1; code <<= 1;
    }
    return -1;
}

static const uint16_t LBASE[29] = {3,4,5,6,7,8,9,10,11,13,15,17,19,23,27,31,35,43,51,59,67,83,99,115,131,163,195,227,258};
static const uint8_t  LEXT[29]  = {0,0,0,0,0,0,0,0,1,1,1,1,2,2,2,2,3,3,3,3,4,4,4,4,5,5,5,5,0};
static const uint16_t DBASE[30] = {1,2,3,4,5,7,9,13,17,25,33,49,65,97,129,193,257,385,513,769,1025,1537,2049,3073,4097,6145,8193,12289,16385,24577};
static const uint8_t  DEXT[30]  = {0,0,0,0,1,1,2,2,3,3,4,4,5,5,6,6,7,7,8,8,9,9,10,10,11,11,12,12,13,13};
static const uint8_t  ORDER[19] = {16,17,18,0,8,7,9,6,10,5,11,4,12,3,13,2,14,1,15};

/* Returns the number of bytes produced, or a negative code:
 * -1 truncated input, -2 bad block type, -3 stored length check, -4 bad code lengths, -5 bad symbol,
 * -6 distance too far back, -7 output overflow. */
int64_t orc_inflate(const uint8_t *in, size_t in_len, uint8_t *out, size_t out_cap, orc_inflate_stats *st)
{
    bits_t b = {in, in_len, 0, 0};
    size_t o = 0;
    orc_inflate_stats s; memset(&s, 0, sizeof(s));
    int last;
    do {
        last = (int)getbits(&b, 1);
        int type = (int)getbits(&b, 2);
        if (b.err) return -1;
        if (type == 3) return -2;
        s.blocks[type]++;
        if (type == 0) {
            b.bit = (b.bit + 7) & ~7ull;
            uint32_t len = getbits(&b, 16), nlen = getbits(&b, 16);
            if (b.err) return -1;
            if ((len ^ 0xffffu) != nlen) return -3;
            size_t pos = (size_t)(b.bit >> 3);
            if (pos + len > in_len) return -1;
            if (o + len > out_cap) return -7;
            memcpy(out + o, in + pos, len); o += len; b.bit += (uint64_t)len * 8;
            continue;
        }
        huff_t hl, hd; uint8_t lens[320];
        if (type == 1) {
            for (int i = 0; i < 288; i++) lens[i] = (uint8_t)(i < 144 ? 8 : i < 256 ? 9 : i < 280 ? 7 : 8);
            build(&hl, lens, 288);
            for (int i = 0; i < 30; i++) lens[i] = 5;
            build(&hd, lens, 30);
        } else {
            int hlit = (int)getbits(&b, 5) + 257, hdist = (int)getbits(&b, 5) + 1, hclen = (int)getbits(&b, 4) + 4;
            if (b.err) return -1;
            if (hlit > 286 || hdist > 30) return -4;
            uint8_t cl[19]; memset(cl, 0, sizeof(cl));
            for (int i = 0; i < hclen; i++) cl[ORDER[i]] = (uint8_t)getbits(&b, 3);
            huff_t hc;
            if (build(&hc, cl, 19) != 0) return -4;
            int idx = 0;
            while (idx < hlit + hdist) {
                int sym = decode(&b, &hc);
                if (sym < 0) return b.err ? -1 : -5;
                if (sym < 16) lens[idx++] = (uint8_t)sym;
                else {
                    int rep, val = 0;
                    if (sym == 16) { if (!idx) return -4; val = lens[idx - 1]; rep = 3 + (int)getbits(&b, 2); }
                    else if (sym == 17) rep = 3 + (int)getbits(&b, 3);
                    else rep = 11 + (int)getbits(&b, 7);
                    if (idx + rep > hlit + hdist) return -4;
                    while (rep--) lens[idx++] = (uint8_t)val;
                }
            }
            if (lens[256] == 0) return -4;
            int e = build(&hl, lens, hlit);
            if (e < 0 || (e > 0 && hlit - hl.count[0] != 1)) return -4;
            e = build(&hd, lens + hlit, hdist);
            if (e < 0 || (e > 0 && hdist - hd.count[0] != 1)) return -4;
        }
        for (;;) {
            int sym = decode(&b, &hl);
            if (sym < 0) return b.err ? -1 : -5;
            if (sym < 256) { if (o >= out_cap) return -7; out[o++] = (uint8_t)sym; s.literals++; }
            else if (sym == 256) break;
            else {
                sym -= 257;
                if (sym >= 29) return -5;
                uint32_t len = LBASE[sym] + getbits(&b, LEXT[sym]);
                int ds = decode(&b, &hd);
                if (ds < 0 || ds >= 30) return b.err ? -1 : -5;
                uint32_t dist = DBASE[ds] + getbits(&b, DEXT[ds]);
                if (b.err) return -1;
                if (dist > o) return -6;
                if (o + len > out_cap) return -7;
                for (uint32_t k = 0; k < len; k++, o++) out[o] = out[o - dist];
                s.matches++;
            }
        }
    } while (!last);
    s.bfinal_seen = 1; s.out_bytes = o; s.in_bits = b.bit;
    if (st) *st = s;
    return (int64_t)o;
}
