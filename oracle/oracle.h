/* oracle.h -- CPU restatement of the reference's algorithm for the compression hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference leg may load liboracle.so; the product (gzippier_b200/) never does.
 * Every function cites the reference file:line it restates.  Parity of this restatement is
 * PINNED against the reference itself (oracle/_ref: the reference's own sources compiled here)
 * and against vectors frozen from it in tests/golden/ -- see tests/test_oracle.py.
 */
#ifndef GZB_ORACLE_H
#define GZB_ORACLE_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* crc32_oracle.c */
uint32_t orc_crc32(uint32_t crc, const uint8_t *buf, size_t len);          /* lib/zlib/crc32.c:202-297 */
uint32_t orc_updcrc(uint32_t running, const uint8_t *s, size_t n);          /* src/util.c:164-185      */
uint32_t orc_crc32_combine(uint32_t crc1, uint32_t crc2, uint64_t len2);    /* src/parallel.c:393-474  */

/* gzip_oracle.c */
size_t   orc_gzip_header(int level, uint32_t mtime, const char *name, uint8_t *out);   /* src/parallel.c:23-47,496-507 */
size_t   orc_gzip_header_single(int level, uint8_t *out);                   /* lib/zlib/deflate.c:849-866 */
size_t   orc_gzip_trailer(uint32_t crc, uint64_t ulen, uint8_t *out);       /* src/parallel.c:52-62,555-556 */

/* inflate_oracle.c -- RFC 1951 decoder used to verify round trips and to dissect streams */
typedef struct orc_inflate_stats { uint64_t blocks[3], literals, matches, out_bytes, in_bits; int bfinal_seen; } orc_inflate_stats;
int64_t  orc_inflate(const uint8_t *in, size_t in_len, uint8_t *out, size_t out_cap, orc_inflate_stats *st);

/* The deflate arithmetic itself (lib/zlib deflate.c/trees.c) is exercised through oracle/_ref -- the
 * reference's own sources compiled by oracle/Makefile -- and through the sizes frozen from it in
 * tests/golden/; engine_model.[ch] is the CPU statement of the ENGINE's algorithm, not the reference's. */

#ifdef __cplusplus
}
#endif
#endif
