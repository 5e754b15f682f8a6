/* mock_engine.c -- TEST INFRASTRUCTURE ONLY: a stand-in for the engine entry points host/gzb_zip.c calls, so that the fd
 * driver's own logic (batching, dictionary hand-over, order, offsets of the file-to-file path, framing, CRC fold, error
 * paths) can be exercised on a machine without a GPU.  It "compresses" into stored deflate blocks on the CPU.  It is built
 * into tests/build/libzipmock.so together with gzb_zip.c by tests/test_zip_driver_cpu.py and nowhere else: the product
 * library never sees it, and the product has no CPU path.
 * The CRC-32 and its combine are the oracle's (oracle/crc32_oracle.c). */
#include "gzb200.h"
#include "../gzippier_b200/csrc/gzb_internal.h"
#include "../oracle/oracle.h"
#include <stdlib.h>
#include <string.h>

struct gzb_ctx { int device; long calls; };

static int env_int(const char *name, int dflt) { const char *e = getenv(name); return e ? atoi(e) : dflt; }

int gzb_device_count(void) { return env_int("MOCK_GPUS", 1); }
int gzb_create(int device, gzb_ctx **ctx)
{
    if (device < 0 || device >= gzb_device_count()) return GZB_E_NO_DEVICE;
    *ctx = (gzb_ctx *)calloc(1, sizeof(gzb_ctx));
    if (!*ctx) return GZB_E_NOMEM;
    (*ctx)->device = device;
    return GZB_OK;
}
void gzb_destroy(gzb_ctx *ctx) { free(ctx); }
const char *gzb_strerror(int code) { return code == GZB_OK ? "ok" : "mock engine error"; }
uint64_t gzb_bound(uint64_t n, uint32_t flags)
{
    const uint64_t chunk = (flags & GZB_F_RSYNCABLE) ? GZB_RSYNC_CHUNK : GZB_CHUNK;
    return n + 19 * ((n + chunk - 1) / chunk) + 64;
}
uint32_t gzb_crc32_combine(uint32_t crc1, uint32_t crc2, uint64_t len2) { return orc_crc32_combine(crc1, crc2, len2); }
int gzb_gzip_header(int level, uint32_t mtime, const char *name, uint8_t *out, size_t cap)
{
    const size_t nl = name ? strlen(name) + 1 : 0;
    if (!out || cap < 10 + nl) return GZB_E_ARG;
    out[0] = 0x1f; out[1] = 0x8b; out[2] = 8; out[3] = name ? 8 : 0;
    for (int i = 0; i < 4; i++) out[4 + i] = (uint8_t)(mtime >> (8 * i));
    out[8] = (uint8_t)(level >= 9 ? 2 : level == 1 ? 4 : 0); out[9] = 3;
    if (name) memcpy(out + 10, name, nl);
    return (int)(10 + nl);
}
int gzb_gzip_trailer(uint32_t crc32, uint64_t isize, uint8_t *out, size_t cap)
{
    if (!out || cap < 8) return GZB_E_ARG;
    for (int i = 0; i < 4; i++) { out[i] = (uint8_t)(crc32 >> (8 * i)); out[4 + i] = (uint8_t)((uint32_t)isize >> (8 * i)); }
    return 8;
}
int gzb_host_alloc(gzb_ctx *ctx, uint64_t bytes, void **p) { (void)ctx; *p = malloc(bytes ? bytes : 1); return *p ? GZB_OK : GZB_E_NOMEM; }
void gzb_host_free(void *p) { free(p); }

/* for the tests: xor of the CRC-32s of all dictionaries seen since the last reset, and how many there were */
static uint32_t g_dict_xor; static long g_dict_n;
uint32_t mock_dict_xor(void) { return g_dict_xor; }
long mock_dict_count(void) { return g_dict_n; }
void mock_reset(void) { g_dict_xor = 0; g_dict_n = 0; }

/* n bytes as stored blocks (a byte-aligned fragment, BFINAL on the last block if asked); returns the fragment's length */
static uint64_t store(const uint8_t *in, uint64_t n, int final, uint8_t *out)
{
    uint64_t o = 0, off = 0;
    if (n == 0) { if (final) { static const uint8_t e[5] = {1, 0, 0, 0xff, 0xff}; memcpy(out, e, 5); return 5; } return 0; }
    while (off < n) {
        const uint32_t len = n - off > 65535 ? 65535u : (uint32_t)(n - off);
        out[o++] = (uint8_t)((final && off + len == n) ? 1 : 0);
        out[o++] = (uint8_t)len; out[o++] = (uint8_t)(len >> 8); out[o++] = (uint8_t)~len; out[o++] = (uint8_t)(~len >> 8);
        memcpy(out + o, in + off, len); o += len; off += len;
    }
    return o;
}

int gzb_deflate_host(gzb_ctx *ctx, int level, uint32_t flags, const uint8_t *in, uint64_t n, const uint8_t *dict,
                     uint8_t *out, uint64_t out_cap, gzb_result *result)
{
    if (!ctx || !out || (n && !in)) return GZB_E_ARG;
    if (dict && !(flags & GZB_F_RSYNCABLE)) {                    /* what the driver handed over as the 32 KiB before the batch */
        __atomic_fetch_xor(&g_dict_xor, orc_crc32(0, dict, GZB_DICT), __ATOMIC_RELAXED);
        __atomic_fetch_add(&g_dict_n, 1, __ATOMIC_RELAXED);
    }
    if (level < 1 || level > 9) return GZB_E_LEVEL;
    if (out_cap < gzb_bound(n, flags)) return GZB_E_OUT_FULL;
    if (env_int("MOCK_FAIL_AFTER", -1) >= 0 && ctx->calls++ >= env_int("MOCK_FAIL_AFTER", -1)) return GZB_E_CUDA;
    memset(result, 0, sizeof(*result));
    result->in_bytes = n; result->out_bytes = store(in, n, (flags & GZB_F_FINAL) != 0, out);
    result->crc32 = orc_crc32(0, in, (size_t)n); result->n_chunks = (uint32_t)((n + GZB_CHUNK - 1) / GZB_CHUNK);
    return GZB_OK;
}

/* the batch interface, with the real one's habits: the fragment of batch k is passed on while batch k + 1 is being read,
 * and its buffer is only taken back (emit_done) two batches later and at the very end */
int gzb_internal_deflate_batches(gzb_ctx *ctx, int level, uint32_t flags, const gzb_batch_io *io)
{
    if (!ctx || !io || !io->fetch || !io->emit) return GZB_E_ARG;
    const uint32_t dsz = (flags & GZB_F_RSYNCABLE) ? 0u : GZB_DICT;
    uint8_t *in = (uint8_t *)malloc(dsz + io->max_bytes + 16), *out[2];
    out[0] = (uint8_t *)malloc(gzb_bound(io->max_bytes, flags)); out[1] = (uint8_t *)malloc(gzb_bound(io->max_bytes, flags));
    int rc = (in && out[0] && out[1]) ? GZB_OK : GZB_E_NOMEM, busy[2] = {0, 0};
    uint64_t held[2] = {0, 0};
    for (uint64_t k = 0; rc == GZB_OK && k < io->n_batches; k++) {
        const int sl = (int)(k & 1);
        uint64_t len = 0; int have_dict = 0, final = 0;
        memset(in, 0xEE, dsz);                                   /* a dictionary that is claimed must have been written */
        rc = io->fetch(io->user, k, in + dsz, dsz, &len, &have_dict, &final);
        if (rc != GZB_OK) break;
        if (len > io->max_bytes) { rc = GZB_E_ARG; break; }
        if (busy[sl] && io->emit_done) { rc = io->emit_done(io->user, held[sl]); busy[sl] = 0; if (rc != GZB_OK) break; }
        gzb_result r;
        rc = gzb_deflate_host(ctx, level, (flags & ~GZB_F_FINAL) | (final ? GZB_F_FINAL : 0u), in + dsz, len,
                              have_dict ? in : NULL, out[sl], gzb_bound(io->max_bytes, flags), &r);
        if (rc != GZB_OK) break;
        rc = io->emit(io->user, k, out[sl], &r);
        busy[sl] = io->emit_done != NULL; held[sl] = k;
    }
    for (int sl = 0; sl < 2; sl++) if (busy[sl] && io->emit_done) { const int r2 = io->emit_done(io->user, held[sl]); if (rc == GZB_OK) rc = r2; }
    free(in); free(out[0]); free(out[1]);
    return rc;
}
