/* crc32_oracle.c -- CPU restatement of the reference's CRC-32 and CRC combine.
 * TEST INFRASTRUCTURE ONLY (see oracle.h).  Pinned by tests/test_oracle.py against oracle/_ref
 * (the reference's own zlib 1.2.11) and Python's zlib.
 */
#include "oracle.h"
#include <pthread.h>

#define POLY 0xEDB88320u                 /* reflected CRC-32 polynomial, lib/zlib/crc32.c:60-89 */

static uint32_t T[4][256];               /* crc_table[0..3], lib/zlib/crc32.c:89-135 (make_crc_table) */
static pthread_once_t once = PTHREAD_ONCE_INIT;

static void make_tables(void)
{
    /* crc_table[0][n]: remainder of n * x^32; crc_table[k][n]: the same advanced by k more bytes
     * (lib/zlib/crc32.c:118-132) */
    for (uint32_t n = 0; n < 256; n++) {
        uint32_t c = n;
        for (int k = 0; k < 8; k++) c = (c & 1) ? (c >> 1) ^ POLY : c >> 1;
        T[0][n] = c;
    }
    for (uint32_t n = 0; n < 256; n++) {
        uint32_t c = T[0][n];
        for (int k = 1; k < 4; k++) { c = T[0][c & 0xff] ^ (c >> 8); T[k][n] = c; }
    }
}

/* crc32() -> crc32_z() -> crc32_little(): byte-wise until 4-aligned, then 4 bytes per step through
 * the four tables (DOLIT4), then the tail (lib/zlib/crc32.c:266-297) */
uint32_t orc_crc32(uint32_t crc, const uint8_t *buf, size_t len)
{
    pthread_once(&once, make_tables);
    if (!buf) return 0;                                     /* crc32.c:207 */
    uint32_t c = ~crc;
    while (len && ((uintptr_t)buf & 3)) { c = T[0][(c ^ *buf++) & 0xff] ^ (c >> 8); len--; }
    while (len >= 4) {
        uint32_t w = (uint32_t)buf[0] | (uint32_t)buf[1] << 8 | (uint32_t)buf[2] << 16 | (uint32_t)buf[3] << 24;
        c ^= w;
        c = T[3][c & 0xff] ^ T[2][(c >> 8) & 0xff] ^ T[1][(c >> 16) & 0xff] ^ T[0][c >> 24];
        buf += 4; len -= 4;
    }
    while (len--) c = T[0][(c ^ *buf++) & 0xff] ^ (c >> 8);
    return ~c;
}

/* updcrc(): the legacy byte-at-a-time table CRC with a global running value kept pre-inverted
 * (src/util.c:164-185).  Not on the compression path; kept as a second witness.
 * `running` is the raw register (0xffffffff to start); returns the new register; the CRC is ~register. */
uint32_t orc_updcrc(uint32_t running, const uint8_t *s, size_t n)
{
    pthread_once(&once, make_tables);
    uint32_t c = running;
    while (n--) c = T[0][(c ^ *s++) & 0xff] ^ (c >> 8);
    return c;
}

/* crc32_comb(): CRC(A||B) from CRC(A), CRC(B), len(B) by building the "append len2 zero bytes"
 * operator as a 32x32 GF(2) matrix through repeated squaring (src/parallel.c:393-474, a copy of
 * lib/zlib/crc32.c:344-425) */
static uint32_t mat_times(const uint32_t *mat, uint32_t vec)          /* gf2_matrix_times, parallel.c:393-407 */
{
    uint32_t sum = 0;
    for (; vec; vec >>= 1, mat++) if (vec & 1) sum ^= *mat;
    return sum;
}
static void mat_square(uint32_t *sq, const uint32_t *mat)             /* gf2_matrix_square, parallel.c:409-416 */
{
    for (int n = 0; n < 32; n++) sq[n] = mat_times(mat, mat[n]);
}
uint32_t orc_crc32_combine(uint32_t crc1, uint32_t crc2, uint64_t len2)
{
    uint32_t even[32], odd[32];
    if (len2 == 0) return crc1;                                        /* parallel.c:428-431 */
    odd[0] = POLY;                                                     /* operator for one zero bit */
    for (int n = 1; n < 32; n++) odd[n] = 1u << (n - 1);
    mat_square(even, odd);                                             /* two zero bits  */
    mat_square(odd, even);                                             /* four zero bits */
    for (;;) {                                                         /* parallel.c:447-468 */
        mat_square(even, odd);
        if (len2 & 1) crc1 = mat_times(even, crc1);
        len2 >>= 1;
        if (!len2) break;
        mat_square(odd, even);
        if (len2 & 1) crc1 = mat_times(odd, crc1);
        len2 >>= 1;
        if (!len2) break;
    }
    return crc1 ^ crc2;
}
