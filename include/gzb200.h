/* gzb200.h -- C ABI of the B200-native gzip compression engine.
 *
 * This is the drop-in boundary for the compression hot path of
 * wfehrnstrom/gzippier.  The reference has no plugin API; the seam is the
 * direct call in zip() (reference src/zip.c:149-153, the IBM_Z_DFLTCC
 * precedent) into deflateGZIP()/parallel_zip() (src/zip.c:40-125,
 * src/parallel.c:699-834), whose compress threads drive zlib
 * (src/parallel.c:561-696).  Every entry point below names the reference
 * interface it stands in for.  INTEGRATION.md shows the patch a maintainer of
 * the reference would apply.
 *
 * Plain C linkage, plain pointers and sizes; no C++/torch types.  All
 * functions return GZB_OK (0) or a negative GZB_E_* code -- unlike the
 * reference, which swallows errors on this path.
 * There is NO CPU fallback: without a CUDA device gzb_create() fails with
 * GZB_E_NO_DEVICE and nothing else can be called.
 */
#ifndef GZB200_H
#define GZB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GZB_VERSION "0.1.0"

/* ---- error codes (zlib-style: 0 ok, negative = error) ---- */
#define GZB_OK            0
#define GZB_E_ARG        (-1)   /* bad pointer / size / alignment                       */
#define GZB_E_LEVEL      (-2)   /* level outside 1..9 (reference: gzip.c:592-602)       */
#define GZB_E_NOMEM      (-3)   /* host or device allocation failed                     */
#define GZB_E_CUDA       (-4)   /* a CUDA call or kernel failed; see gzb_last_error()   */
#define GZB_E_NO_DEVICE  (-5)   /* no usable CUDA device: there is no CPU fallback      */
#define GZB_E_OUT_FULL   (-6)   /* out_cap < gzb_bound(n)                               */
#define GZB_E_IO         (-7)   /* read()/write() failed in the host driver             */
#define GZB_E_VERIFY     (-8)   /* GZB_F_VERIFY: the output did not decode back to the input */
#define GZB_E_TOO_LARGE  (-9)   /* gzb_crc32_device_noalloc: n > GZB_NOALLOC_MAX_BYTES  */

/* ---- the reference's chunk grid (src/parallel.c:13-15) ---- */
#define GZB_CHUNK        131072u   /* IN_BUF_SIZE                                       */
#define GZB_DICT          32768u   /* DICTIONARY_SIZE                                   */
#define GZB_RSYNC_CHUNK   16384u   /* --rsyncable: 16 KiB chunks, no dictionary (:333,:767) */

/* ---- flags ---- */
#define GZB_F_FINAL       1u    /* this call ends the deflate stream: BFINAL on the last block
                                   (the job with more==0, src/parallel.c:645-649)          */
#define GZB_F_RSYNCABLE   2u    /* 16 KiB independent chunks (src/parallel.c:333,767)      */
#define GZB_F_VERIFY      4u    /* decode every fragment again on the device and compare it with the input before
                                   returning (what `gzip -t` after the fact does through src/unzip.c:254-370, chunk
                                   parallel because our fragment boundaries are known): GZB_E_VERIFY on a mismatch */

typedef struct gzb_ctx gzb_ctx;

/* Result of one compress call.  crc32/isize are what create_trailer() takes
 * (src/parallel.c:52-62, 542-556). */
typedef struct gzb_result {
    uint64_t in_bytes;      /* bytes consumed (== n)                                     */
    uint64_t out_bytes;     /* raw-deflate bytes produced                                */
    uint32_t crc32;         /* CRC-32 of the n input bytes (zlib crc32 semantics)        */
    uint32_t n_chunks;      /* chunks on the reference grid                              */
    uint32_t n_blocks;      /* deflate blocks emitted                                    */
    int32_t  status;        /* GZB_OK or GZB_E_*                                         */
} gzb_result;

/* Per-stage device time of the last gzb_deflate_device() call, in milliseconds
 * (CUDA events on the context's stream).  Stands in for the reference's absent
 * profiler. */
typedef struct gzb_stage_ms {
    float crc, lz77, huffman, layout, pack, total;
} gzb_stage_ms;

/* ---- context: one per GPU (replaces deflateInit2/deflateEnd per compress
 *      thread, src/parallel.c:571,695, and the buffer pools :327-347) ---- */
int  gzb_create(int device, gzb_ctx **ctx);
void gzb_destroy(gzb_ctx *ctx);
const char *gzb_strerror(int code);
const char *gzb_last_error(const gzb_ctx *ctx);          /* detail text of the last failure  */
int  gzb_device_count(void);                             /* 0 when there is no usable GPU    */

/* Worst-case raw-deflate size for n input bytes (replaces deflateBound,
 * lib/zlib/deflate.c:652-709, and the grow-by-1.25 output pool,
 * src/parallel.c:220-259: overflow is impossible by construction). */
uint64_t gzb_bound(uint64_t n, uint32_t flags);

/* Compress n bytes that are ALREADY IN DEVICE MEMORY into raw-deflate
 * fragments on the reference's chunk grid: chunk i = in[i*C, (i+1)*C), primed
 * with the dictionary in[i*C-32768, i*C) (src/parallel.c:767-774); every
 * fragment ends on a byte boundary (:624-644); the last one carries BFINAL iff
 * GZB_F_FINAL (:645-649).  No gzip header or trailer is written -- the host
 * does that, as write_thread does (:496-507, :555-556).
 *
 * halo = number of valid bytes BEFORE d_in (0, or 32768 for a shard that does
 * not start the stream): d_in[-halo, n) must be readable device memory.
 * d_in and d_out must be 16-byte aligned.  Replaces the per-job body of
 * compress_thread (src/parallel.c:595-688: deflateReset/Params/SetDictionary/
 * deflate/crc32z).  Runs on `stream` (a cudaStream_t, or NULL for the
 * context's own stream) and returns after the result is known.  The context's
 * own stream is non-blocking: it is NOT ordered with the legacy default stream,
 * so a caller that fills d_in or reads d_out on another stream passes that
 * stream here or synchronises itself. */
int gzb_deflate_device(gzb_ctx *ctx, int level, uint32_t flags,
                       const uint8_t *d_in, uint64_t n, uint32_t halo,
                       uint8_t *d_out, uint64_t out_cap,
                       gzb_result *result, void *stream);

/* Same, but only enqueues the work on `stream`; the result is delivered to
 * ctx-owned pinned memory and fetched with gzb_deflate_wait().  Used to time
 * back-to-back launches with CUDA events on the caller's stream. */
int gzb_deflate_device_async(gzb_ctx *ctx, int level, uint32_t flags,
                             const uint8_t *d_in, uint64_t n, uint32_t halo,
                             uint8_t *d_out, uint64_t out_cap, void *stream);
int gzb_deflate_wait(gzb_ctx *ctx, gzb_result *result);
int gzb_last_stage_ms(gzb_ctx *ctx, gzb_stage_ms *ms);

/* Compress n bytes of HOST memory into host memory: pinned staging, H2D,
 * kernels, D2H, pipelined in batches of chunks.  `dict` is the 32 KiB that
 * precede `in` in the stream (NULL at the start of the stream).  This is the
 * call the reference-facing host driver makes once per read batch.
 * out_cap >= gzb_bound(n) always suffices; a smaller buffer is accepted and the
 * call fails with GZB_E_OUT_FULL (nothing written past out_cap) if the output
 * does not fit.  On any error nothing is left in flight: all copies into the
 * caller's buffers have completed when the call returns. */
int gzb_deflate_host(gzb_ctx *ctx, int level, uint32_t flags,
                     const uint8_t *in, uint64_t n, const uint8_t *dict,
                     uint8_t *out, uint64_t out_cap, gzb_result *result);

/* The same over SEVERAL GPUs in one call (one context per GPU, created by the caller): contiguous shards on chunk
 * boundaries, the 32 KiB before a shard as its dictionary, no collective -- what parallel_zip does with -j N compress
 * threads (src/parallel.c:699-834), with a GPU where a thread stood.  Shard i's body lands at out + shards[i].out_off
 * (out_cap >= gzb_bound(n) + 64 * n_ctx); compact != 0 moves the bodies together so that out[0, result->out_bytes) is the
 * fragment run of the whole input.  crc32 in *result is folded with gzb_crc32_combine.  shards may be NULL. */
typedef struct gzb_shard { uint64_t in_off, in_bytes, out_off, out_bytes; } gzb_shard;
int gzb_deflate_host_multi(gzb_ctx **ctx, int n_ctx, int level, uint32_t flags,
                           const uint8_t *in, uint64_t n, const uint8_t *dict,
                           uint8_t *out, uint64_t out_cap, gzb_result *result,
                           gzb_shard *shards, int compact);

/* CRC-32 on the device (replaces crc32z -> zlib crc32, src/parallel.c:476-486,
 * lib/zlib/crc32.c:202-297).  crc is the running value, 0 to start. */
int gzb_crc32_device(gzb_ctx *ctx, uint32_t crc, const uint8_t *d_in, uint64_t n,
                     uint32_t *crc_out, void *stream);
int gzb_crc32_host(gzb_ctx *ctx, uint32_t crc, const uint8_t *in, uint64_t n, uint32_t *crc_out);

/* The same CRC-32 with no allocation of any kind.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Needs no gzb_ctx.  The per-chunk values live in a file-scope __device__ array
 * (~260 KiB per GPU that loads the library, held until exit), which bounds the input: n <= GZB_NOALLOC_MAX_BYTES
 * (8 GiB), else GZB_E_TOO_LARGE and nothing is launched.  Works on the CURRENT device, which must own d_in
 * (16-byte aligned) and d_crc_out (a DEVICE word, written on `stream`; NULL = the default stream).  The call
 * enqueues one 4-byte memset and one kernel and returns without synchronising.  Every call on a GPU shares the
 * array: calls that overlap on the same GPU (two streams, two threads) race -- serialise them.  The first call
 * on a device uploads 4 KiB of tables synchronously: make it (n = 0 will do) before starting a capture. */
#define GZB_NOALLOC_MAX_BYTES 0x200000000ull
int gzb_crc32_device_noalloc(uint32_t crc, const uint8_t *d_in, uint64_t n,
                             uint32_t *d_crc_out, void *stream);

/* CRC(A||B) from CRC(A), CRC(B), len(B) (replaces crc32_comb,
 * src/parallel.c:393-474 == lib/zlib/crc32.c:344-425).  Host arithmetic. */
uint32_t gzb_crc32_combine(uint32_t crc1, uint32_t crc2, uint64_t len2);

/* gzip member framing (replaces create_header/create_trailer and the writes in
 * write_thread, src/parallel.c:23-62, 496-507, 555-556; single-stream twin:
 * lib/zlib/deflate.c:849-873, 1051-1060).  name may be NULL (stdin).  Returns
 * the number of bytes written or a negative error. */
int gzb_gzip_header(int level, uint32_t mtime, const char *name, uint8_t *out, size_t cap);
int gzb_gzip_trailer(uint32_t crc32, uint64_t isize, uint8_t *out, size_t cap);

/* ---- reference-facing host driver (host/gzb_zip.c) -----------------------
 * Mirror of `off_t parallel_zip(int pack_level)` / `off_t deflateGZIP(int)`
 * (src/gzip.h:286,338) with the reference's globals passed explicitly:
 * reads ifd to EOF, writes one gzip member to ofd.  n_gpus devices are used,
 * input sharded by contiguous chunk ranges, no collective.
 * Pipes and terminals go through one reader, workers and one in-order writer
 * (the reference's structure, src/parallel.c:106-143, 490-834).  Between two
 * regular files nothing is serial: every worker reads its own batches, and a
 * fragment's range of the output is reserved with fallocate() and filled
 * through a shared mapping (pwrite where that is not to be had); the input is
 * taken to end where the file ended when the call began. */
typedef struct gzb_zip_opts {
    int level;              /* `level`       src/gzip.c:211   */
    int rsyncable;          /* `rsync`       src/gzip.c:244   */
    int n_gpus;             /* stands where `threads` (-j N) stood, src/gzip.c:202 */
    const char *name;       /* `ifname` when input is a file, else NULL (src/parallel.c:496-507) */
    uint32_t mtime;         /* `time_stamp.tv_sec`           */
    uint64_t bytes_in, bytes_out;   /* filled on return (reference globals bytes_in/bytes_out) */
    int members;            /* 0: ONE member, fragments of all batches joined (the reference's layout, what its own
                             * `gzip -d` reads).  1: every batch a member of its own, no dictionary across batches
                             * ("per-GPU members concatenated"): legal RFC 1952 that gunzip, zlib's gzread and Python's
                             * gzip decode in full, but the reference's decoder stops after the first member
                             * (src/unzip.c:361, src/gzip.c:1024-1030) -- hence off by default. */
} gzb_zip_opts;
int gzb_zip_fd(int ifd, int ofd, gzb_zip_opts *opts);
/* gzb_zip_fd keeps its contexts and page-locked batch slots for the next call of the process (creating them costs more
 * than compressing a few GiB); this frees them.  Stands where the reference frees its pools at the end of parallel_zip
 * (src/parallel.c:820-833). */
void gzb_zip_release(void);

/* Page-locked host memory for the buffers handed to gzb_deflate_host (which then DMAs straight from / to them
 * instead of staging).  Stands where the reference's buffer pools stood (src/parallel.c:169-259: get_space /
 * drop_space / grow_space -- fixed sizes here, gzb_bound() makes growing unnecessary). */
int gzb_host_alloc(gzb_ctx *ctx, uint64_t bytes, void **p);
void gzb_host_free(void *p);

/* How gzb_deflate_host cuts n input bytes into batches: off[0] = 0 < off[1] < ... < off[k] = n, k returned (1 for n = 0;
 * GZB_E_ARG if cap < k + 1).  Host arithmetic only, no GPU needed.  The reader loop of the reference hands out fixed
 * 128 KiB jobs (src/parallel.c:699-834); here a batch is what one H2D copy / kernel sequence / D2H copy handles, every
 * boundary is a multiple of both chunk sizes, and a large input starts with batches of 64, 192 MiB ... and ends with one
 * of <= 64 MiB so that the copies nothing can hide behind stay small (GZB_HOST_BATCH_MB, GZB_HOST_EDGE_MB). */
int gzb_host_batch_plan(uint64_t n, uint64_t *off, int cap);

#ifdef __cplusplus
}
#endif
#endif /* GZB200_H */
