/* engine_model.h -- CPU model of the B200 engine's own algorithm (TEST INFRASTRUCTURE ONLY). */
#ifndef GZB_ENGINE_MODEL_H
#define GZB_ENGINE_MODEL_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define GZM_HBITS       14
#define GZM_H3BITS      11                        /* single-entry table of most recent 3-byte twins */
#define GZM_H4BITS      12                        /* single-entry table of most recent 4-byte twins */
#define GZM_SCAN_MIN    4096                      /* k_scan: shorter chunks always take the normal path       */
#define GZM_SCAN_SAMPLE 256                       /* one 16-byte word per 256 bytes goes into the histogram   */
#define GZM_SCAN_SLOTS  4096
#define GZM_SCAN_MAX_GRAMS 3072
#define GZM_WALK_STEPS  4                         /* chain steps per round of the walk (k_lz77 search_group)  */
#define GZM_TILE        2048                      /* positions per pipeline step of the LZ77 kernel */
#define GZM_MAXD        32768                     /* farthest match (the link ring is 32768+2*TILE) */
#define GZM_TOO_FAR     4096                      /* lib/zlib/deflate.c:110-112                      */
#define GZM_BLOCK_SPAN  65536                     /* input bytes per deflate block                   */

typedef struct gzm_params {
    int level, good, max_lazy, nice, ref_chain;   /* the reference's configuration_table row         */
    int depth;                                    /* candidates examined per position                */
    int lazy_parse;                               /* 0: greedy (levels 1-3), 1: lazy (4-9)           */
    int maxd, hbits, block_span, chunk, dict;
    int use_h3;                                   /* fall back to the most recent 3-byte twin when nothing longer exists */
    int use_h4;                                   /* ... and to the most recent 4-byte twin when the chain gives < 6  */
    int scan;                                     /* 1: incompressible chunks are stored without a search (k_scan)   */
    int good_cut;                                 /* quarter the remaining candidate budget once the best match reaches `good` */
} gzm_params;

typedef struct gzm_block_plan {
    uint8_t  type;                                /* 0 stored, 1 static, 2 dynamic                    */
    int      hlit, hdist, hclen;
    uint8_t  llen[288], dlen[30], blen[19];
    uint16_t lcode[288], dcode[30];
    uint8_t  header[640]; uint32_t header_bits;   /* HLIT..end of the code-length section             */
    uint64_t dyn_bits, static_bits;               /* block body + header, without the 3 type bits     */
} gzm_block_plan;

typedef struct gzm_chunk_stats { uint64_t tokens, matches, blocks[3]; } gzm_chunk_stats;

void     gzm_level_params(int level, gzm_params *p);
uint32_t gzm_hash6(uint32_t lo, uint32_t hi16, int hbits);
uint32_t gzm_hash4(uint32_t v);
uint32_t gzm_hash3(uint32_t v);
int      gzm_len_code(int len);
int      gzm_dist_code(int dist);
int      gzm_chunk_is_stored(const uint8_t *stream, int D, int n_chunk);
void     gzm_find_matches(const gzm_params *P, const uint8_t *stream, int D, int n, uint16_t *mlen, uint16_t *mdist);
int      gzm_parse(const gzm_params *P, const uint8_t *chunk, int n, const uint16_t *mlen, const uint16_t *mdist,
                   uint32_t *tokens, uint32_t *tok_pos);
int      gzm_huff_lengths(const uint32_t *freq, int nsym, int maxbits, uint8_t *lens);
void     gzm_huff_codes(const uint8_t *lens, int nsym, int maxbits, uint16_t *codes);
void     gzm_static_lengths(uint8_t *ll, uint8_t *dl);
void     gzm_plan_block(const uint32_t *lfreq, const uint32_t *dfreq, uint32_t in_bytes, gzm_block_plan *B);
size_t   gzm_deflate_chunk(const gzm_params *P, const uint8_t *stream, int D, int n_chunk, int final,
                           uint8_t *out, gzm_chunk_stats *st);
int64_t  gzm_deflate(const gzm_params *P, const uint8_t *in, uint64_t n, int halo, int final,
                     uint8_t *out, uint64_t cap, int nthreads, gzm_chunk_stats *stats);
#ifdef __cplusplus
}
#endif
#endif
