/* gzb_corpus.h -- synthetic corpora for the benchmark and the tests (libgzbcorpus.so, csrc/corpus.c).
 *
 * NOT part of the engine: a separate, host-only library, so that a process that only needs input bytes -- bench.py's
 * reference arm, the oracle tests -- maps nothing of libgzb200.so.  The reference's own benchmark inputs are downloads
 * that are not in its tree (benchmarks/Makefile.am:4-7); every number in this repository is on these synthetic bytes.
 */
#ifndef GZB_CORPUS_H
#define GZB_CORPUS_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif
#define GZB_CORPUS_TEXT    0
#define GZB_CORPUS_ZIPF    1
#define GZB_CORPUS_BINARY  2
#define GZB_CORPUS_RANDOM  3
#define GZB_CORPUS_MIXED   4
/* dst[0, n) = bytes [offset, offset + n) of the corpus (kind, seed); any range can be produced independently.
 * Returns 0, or -1 on a bad argument. */
int gzb_corpus_fill(int kind, uint64_t seed, uint64_t offset, uint8_t *dst, uint64_t n, int nthreads);
#ifdef __cplusplus
}
#endif
#endif /* GZB_CORPUS_H */
