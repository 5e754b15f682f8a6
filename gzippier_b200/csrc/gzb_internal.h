/* gzb_internal.h -- what the fd driver (host/gzb_zip.c) uses of the engine besides the public ABI (include/gzb200.h).
 * Not installed, not part of the drop-in boundary. */
#ifndef GZB_INTERNAL_H
#define GZB_INTERNAL_H
#include "gzb200.h"
#ifdef __cplusplus
extern "C" {
#endif

/* A sequence of batches that one context compresses through its own pinned staging buffers, with the same three-stream
 * overlap as gzb_deflate_host (H2D of batch k+1 || kernels of batch k || D2H of batch k-1) -- but the bytes come from and
 * go to callbacks, so a file can be read straight into the staging buffer by many threads, and the batches need not be
 * neighbours in the input (the fd driver deals them round-robin to the GPUs).
 *   fetch  fills dst[0, *len) with batch k (*len <= max_bytes); if the batch has a dictionary it also fills
 *          dst[-dict_sz, 0) and sets *have_dict; *final = 1 if the member ends with this batch.
 *   emit   gets the fragment of batch k (r->out_bytes bytes at src), in order of k, as soon as it has reached the host.
 *   emit_done (may be NULL) blocks until emit(k) is through with src: without it src is valid until emit returns, with it
 *          emit may hand the bytes to somebody else (a writer) and return; the engine asks before it reuses the buffer
 *          (two batches later) and once for every batch before the call returns.
 * All return GZB_OK or a GZB_E_* code, which ends the call.  With `blocking` the host waits sleep instead of spinning. */
typedef struct {
    void *user;
    uint64_t n_batches, max_bytes;
    int blocking;
    int (*fetch)(void *user, uint64_t k, uint8_t *dst, uint32_t dict_sz, uint64_t *len, int *have_dict, int *final);
    int (*emit)(void *user, uint64_t k, const uint8_t *src, const gzb_result *r);
    int (*emit_done)(void *user, uint64_t k);
} gzb_batch_io;
int gzb_internal_deflate_batches(gzb_ctx *ctx, int level, uint32_t flags, const gzb_batch_io *io);

#ifdef __cplusplus
}
#endif
#endif
