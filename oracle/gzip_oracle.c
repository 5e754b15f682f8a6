/* gzip_oracle.c -- CPU restatement of the reference's gzip member framing.
 * TEST INFRASTRUCTURE ONLY (see oracle.h). */
#include "oracle.h"
#include <string.h>

/* create_header() + the writes in write_thread(): 10 bytes, then ifname and NUL when the input is a
 * file (src/parallel.c:23-47, 496-507).  XFL: 2 if level >= 9, 4 if level == 1 (:43). */
size_t orc_gzip_header(int level, uint32_t mtime, const char *name, uint8_t *out)
{
    size_t n = 0;
    out[n++] = 31; out[n++] = 139; out[n++] = 8;                       /* magic, CM = deflate (:38-40) */
    out[n++] = name ? 8 : 0;                                           /* FLG.FNAME (:41) */
    for (int i = 0; i < 4; i++) out[n++] = (uint8_t)(mtime >> (8 * i));/* MTIME, little endian struct dump (:42,500) */
    out[n++] = (uint8_t)(level >= 9 ? 2 : level == 1 ? 4 : 0);         /* XFL (:43) */
    out[n++] = 3;                                                      /* OS = Unix (:44) */
    if (name) { size_t l = strlen(name) + 1; memcpy(out + n, name, l); n += l; }   /* :506 */
    return n;
}

/* zlib's own gzip wrapper, used by the single-stream path: FLG 0, MTIME 0,
 * XFL = 2 if level == 9, 4 if level < 2 (strategy default), OS = 3 (lib/zlib/deflate.c:849-866) */
size_t orc_gzip_header_single(int level, uint8_t *out)
{
    static const uint8_t h[8] = {31, 139, 8, 0, 0, 0, 0, 0};
    memcpy(out, h, 8);
    out[8] = (uint8_t)(level == 9 ? 2 : level < 2 ? 4 : 0);
    out[9] = 3;
    return 10;
}

/* create_trailer(): CRC-32 then length mod 2^32, both little endian (src/parallel.c:52-62, 555-556;
 * single-stream twin lib/zlib/deflate.c:1051-1060) */
size_t orc_gzip_trailer(uint32_t crc, uint64_t ulen, uint8_t *out)
{
    uint32_t l = (uint32_t)ulen;
    for (int i = 0; i < 4; i++) { out[i] = (uint8_t)(crc >> (8 * i)); out[4 + i] = (uint8_t)(l >> (8 * i)); }
    return 8;
}
