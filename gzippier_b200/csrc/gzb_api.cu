// gzb_api.cu -- the C ABI of include/gzb200.h: context, scratch memory, kernel orchestration.
//
// Stands where compress_thread's per-job body stands in the reference (src/parallel.c:595-688) and
// where its buffer pools stand (:327-347).  Everything here is plumbing; the arithmetic is in k_*.cu.
#include "gzb_common.cuh"
#include "gzb200.h"
#include "gzb_internal.h"
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>
#include <thread>
#include <new>

using namespace gzb;

namespace {

constexpr int MAX_SUB = 64;                       // stage-timing slots (sub-batches timed per call)

struct Scratch {
    uint32_t*  tokens = nullptr;
    BlockRec*  blocks = nullptr;
    ChunkRec*  chunks = nullptr;
    uint32_t   cap_chunks = 0, chunk = 0;
};

}  // namespace

// Scheduling knobs, read from the environment ONCE, in gzb_create (DESIGN.md "Tuning knobs").  None changes a byte of output.
struct Tuning {
    int pw = 0;               // GZB_PW: parser group of 8 / 16 warps (0: by level)
    int help = -1;            // GZB_HELP: bit 0 inserters, bit 1 parsers take search groups when idle (-1: by level)
    int rolemap = -1;         // GZB_ROLEMAP: warp layout of k_lz77 (-1: by level)
    int max_run = 32;         // GZB_MAX_RUN
    int force_run = 0;        // GZB_FORCE_RUN (tests)
    int scan = 1;             // GZB_SCAN: 0 = no incompressibility test before the match finder (every chunk takes the normal path)
    uint64_t host_batch = 512ull << 20, host_edge = 64ull << 20;   // GZB_HOST_BATCH_MB, GZB_HOST_EDGE_MB
};

struct gzb_ctx {
    int device = 0, n_sm = 0;
    Tuning tune;
    ChunkRec* d_crc_recs = nullptr; ChunkRec* h_crc_recs = nullptr;   // gzb_crc32_device: one record per chunk, 65536 chunks per launch
#ifdef GZB_TESTING
    long long test_corrupt_at = -1;               // fault injection for the tests of k_verify (test build only)
#endif
    cudaStream_t stream = nullptr;
    uint32_t* d_tab = nullptr;                    // CRC tables (CRC_TAB_WORDS words)
    uint32_t  h_x2n[64];
    Scratch   scr;
    Cursor*   d_cursor = nullptr;
    Cursor*   h_cursor = nullptr;                 // pinned
    uint32_t* d_work = nullptr;                   // one work counter per sub-batch
    uint32_t  batch_chunks = 4096;                // sub-batch size (GZB_BATCH_CHUNKS)
    cudaEvent_t ev[MAX_SUB][6];
    int       n_sub_timed = 0;
    bool      pending = false;
    uint64_t  pending_n = 0;
    cudaStream_t pending_stream = nullptr;
    // host path: two slots of device buffers (+ pinned staging, only used for pageable caller memory),
    // copy streams and events for the H2D / kernels / D2H pipeline
    uint8_t *h_in[2] = {nullptr, nullptr}, *h_out[2] = {nullptr, nullptr}, *d_in[2] = {nullptr, nullptr}, *d_out[2] = {nullptr, nullptr};
    uint64_t dev_in_cap = 0, dev_out_cap = 0, pin_in_cap = 0, pin_out_cap = 0;
    cudaStream_t s_in = nullptr, s_out = nullptr;
    cudaEvent_t e_in[2] = {nullptr, nullptr}, e_cmp[2] = {nullptr, nullptr}, e_out[2] = {nullptr, nullptr};
    // gzb_internal_deflate_batches with `blocking` (the fd driver's workers): the host waits sleep on these instead of spinning
    cudaEvent_t e_blk_cmp = nullptr, e_blk_in = nullptr, e_blk_out[2] = {nullptr, nullptr};
    char      err[512];
    std::mutex mu;
};

namespace {

int fail(gzb_ctx* ctx, int code, const char* what, cudaError_t ce = cudaSuccess)
{
    if (ctx) {
        if (ce != cudaSuccess) snprintf(ctx->err, sizeof(ctx->err), "%s: %s", what, cudaGetErrorString(ce));
        else snprintf(ctx->err, sizeof(ctx->err), "%s", what);
    }
    return code;
}
#define CU(ctx, call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail((ctx), GZB_E_CUDA, #call, e_); } while (0)

void build_crc_tables(uint32_t* tab /* CRC_TAB_WORDS */)
{
    for (uint32_t n = 0; n < 256; n++) {
        uint32_t c = n;
        for (int k = 0; k < 8; k++) c = (c & 1u) ? CRC_POLY ^ (c >> 1) : c >> 1;
        tab[n] = c;
    }
    for (uint32_t n = 0; n < 256; n++) {
        uint32_t c = tab[n];
        for (int k = 1; k < 4; k++) { c = tab[c & 0xFFu] ^ (c >> 8); tab[k * 256 + n] = c; }
    }
    uint32_t p = 1u << 30;                        // x^1
    tab[1024] = p;
    for (int k = 1; k < 64; k++) { p = gf2_mul(p, p); tab[1024 + k] = p; }
}

LevelParams level_params(int level)
{
    // good/lazy/nice: the reference's configuration_table (lib/zlib/deflate.c:134-146).
    // depth: candidates examined per position (every position is searched, chains keyed on 6 bytes).
    static const int t[10][6] = {
        {0, 0, 0, 0, 0, 0},
        {1, 8, 4, 0, 1, 1},   {2, 16, 5, 0, 1, 1},   {4, 32, 6, 0, 1, 1},
        {4, 16, 4, 1, 1, 1},  {6, 32, 16, 1, 1, 1},  {8, 128, 16, 1, 1, 1},
        {16, 128, 32, 1, 1, 1}, {32, 258, 128, 1, 1, 1}, {64, 258, 258, 1, 1, 1}};
    LevelParams lp; memset(&lp, 0, sizeof(lp));
    lp.depth = t[level][0]; lp.nice = t[level][1]; lp.max_lazy = t[level][2]; lp.lazy_parse = t[level][3]; lp.use_h3 = t[level][4]; lp.use_h4 = t[level][5];
    return lp;
}

int ensure_scratch(gzb_ctx* ctx, uint32_t nchunks, uint32_t chunk)
{
    Scratch& s = ctx->scr;
    if (s.cap_chunks >= nchunks && s.chunk == chunk) return GZB_OK;
    cudaFree(s.tokens); cudaFree(s.blocks); cudaFree(s.chunks);
    s = Scratch();
    const uint32_t bpc = (chunk + BLOCK_SPAN - 1) / BLOCK_SPAN;
    cudaError_t e;
    if ((e = cudaMalloc(&s.tokens, (size_t)nchunks * chunk * sizeof(uint32_t))) != cudaSuccess ||
        (e = cudaMalloc(&s.blocks, (size_t)nchunks * bpc * sizeof(BlockRec))) != cudaSuccess ||
        (e = cudaMalloc(&s.chunks, (size_t)nchunks * sizeof(ChunkRec))) != cudaSuccess) {
        cudaFree(s.tokens); cudaFree(s.blocks); cudaFree(s.chunks); s = Scratch();
        (void)cudaGetLastError();
        return fail(ctx, GZB_E_NOMEM, "cudaMalloc(scratch)", e);
    }
    s.cap_chunks = nchunks; s.chunk = chunk;
    return GZB_OK;
}

}  // namespace

extern "C" {

const char* gzb_strerror(int code)
{
    switch (code) {
    case GZB_OK: return "ok";
    case GZB_E_ARG: return "bad argument";
    case GZB_E_LEVEL: return "compression level must be 1..9";
    case GZB_E_NOMEM: return "out of memory";
    case GZB_E_CUDA: return "CUDA error";
    case GZB_E_NO_DEVICE: return "no usable CUDA device (there is no CPU fallback)";
    case GZB_E_OUT_FULL: return "output buffer smaller than gzb_bound()";
    case GZB_E_IO: return "I/O error";
    case GZB_E_VERIFY: return "GZB_F_VERIFY: the output does not decode back to the input";
    case GZB_E_TOO_LARGE: return "input larger than GZB_NOALLOC_MAX_BYTES (the no-allocation entry point's static scratch)";
    default: return "unknown error";
    }
}

const char* gzb_last_error(const gzb_ctx* ctx) { return ctx ? ctx->err : "no context"; }

int gzb_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { (void)cudaGetLastError(); return 0; }
    return n;
}

uint64_t gzb_bound(uint64_t n, uint32_t flags)
{
    // worst case per block: stored, 5 bytes per <=65535-byte sub-block, + 5 bytes of alignment block per chunk
    const uint64_t chunk = (flags & GZB_F_RSYNCABLE) ? GZB_RSYNC_CHUNK : GZB_CHUNK;
    const uint64_t nchunks = n ? (n + chunk - 1) / chunk : 1;
    const uint64_t bpc = (chunk + BLOCK_SPAN - 1) / BLOCK_SPAN;
    return n + nchunks * (bpc * 2 * 5 + 5 + 4) + 64;
}

int gzb_create(int device, gzb_ctx** out)
{
    if (!out) return GZB_E_ARG;
    *out = nullptr;
    int ndev = gzb_device_count();
    if (ndev <= 0 || device < 0 || device >= ndev) return GZB_E_NO_DEVICE;
    gzb_ctx* ctx = new (std::nothrow) gzb_ctx();
    if (!ctx) return GZB_E_NOMEM;
    ctx->err[0] = 0; ctx->device = device;
    memset(ctx->ev, 0, sizeof(ctx->ev));
    cudaDeviceProp prop;
    if (cudaSetDevice(device) != cudaSuccess || cudaGetDeviceProperties(&prop, device) != cudaSuccess) { delete ctx; return GZB_E_NO_DEVICE; }
    if (prop.major < 10) { delete ctx; return GZB_E_NO_DEVICE; }     // sm_100a code only
    ctx->n_sm = prop.multiProcessorCount;
    auto env_int = [](const char* name, long long lo, long long hi, long long dflt) -> long long {
        const char* e = getenv(name);
        if (!e || !*e) return dflt;
        const long long v = atoll(e);
        return v < lo || v > hi ? dflt : v;                                  // out of range: ignored
    };
    ctx->batch_chunks = (uint32_t)env_int("GZB_BATCH_CHUNKS", 1, 1 << 20, 4096);
    ctx->tune.pw = (int)env_int("GZB_PW", 8, 16, 0); if (ctx->tune.pw != 8 && ctx->tune.pw != 16) ctx->tune.pw = 0;
    ctx->tune.help = (int)env_int("GZB_HELP", 0, 3, -1);
    ctx->tune.rolemap = (int)env_int("GZB_ROLEMAP", 0, 4, -1);
    ctx->tune.max_run = (int)env_int("GZB_MAX_RUN", 1, 64, 32);
    ctx->tune.force_run = (int)env_int("GZB_FORCE_RUN", 1, 4096, 0);
    ctx->tune.scan = (int)env_int("GZB_SCAN", 0, 1, 1);
    ctx->tune.host_batch = (uint64_t)env_int("GZB_HOST_BATCH_MB", 1, 1 << 16, 512) << 20;
    ctx->tune.host_edge = (uint64_t)env_int("GZB_HOST_EDGE_MB", 0, 1 << 16, 64) << 20;
    uint32_t tab[CRC_TAB_WORDS];
    build_crc_tables(tab);
    memcpy(ctx->h_x2n, tab + 1024, sizeof(ctx->h_x2n));
    bool ok = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) == cudaSuccess &&
              cudaStreamCreateWithFlags(&ctx->s_in, cudaStreamNonBlocking) == cudaSuccess &&
              cudaStreamCreateWithFlags(&ctx->s_out, cudaStreamNonBlocking) == cudaSuccess &&
              cudaMalloc(&ctx->d_tab, sizeof(tab)) == cudaSuccess &&
              cudaMemcpy(ctx->d_tab, tab, sizeof(tab), cudaMemcpyHostToDevice) == cudaSuccess &&
              cudaMalloc(&ctx->d_cursor, sizeof(Cursor)) == cudaSuccess &&
              cudaMallocHost(&ctx->h_cursor, sizeof(Cursor)) == cudaSuccess &&
              cudaMalloc(&ctx->d_work, sizeof(uint32_t) * 65536) == cudaSuccess &&
              lz77_configure() == cudaSuccess;
    for (int i = 0; ok && i < MAX_SUB; i++) for (int j = 0; j < 6; j++) ok = ok && cudaEventCreate(&ctx->ev[i][j]) == cudaSuccess;
    for (int i = 0; ok && i < 2; i++)
        ok = cudaEventCreateWithFlags(&ctx->e_in[i], cudaEventDisableTiming) == cudaSuccess &&
             cudaEventCreateWithFlags(&ctx->e_cmp[i], cudaEventDisableTiming) == cudaSuccess &&
             cudaEventCreateWithFlags(&ctx->e_out[i], cudaEventDisableTiming) == cudaSuccess &&
             cudaEventCreateWithFlags(&ctx->e_blk_out[i], cudaEventDisableTiming | cudaEventBlockingSync) == cudaSuccess;
    ok = ok && cudaEventCreateWithFlags(&ctx->e_blk_cmp, cudaEventDisableTiming | cudaEventBlockingSync) == cudaSuccess
            && cudaEventCreateWithFlags(&ctx->e_blk_in, cudaEventDisableTiming | cudaEventBlockingSync) == cudaSuccess;
    if (!ok) { (void)cudaGetLastError(); gzb_destroy(ctx); return GZB_E_CUDA; }
    *out = ctx;
    return GZB_OK;
}

void gzb_destroy(gzb_ctx* ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    cudaFree(ctx->scr.tokens); cudaFree(ctx->scr.blocks); cudaFree(ctx->scr.chunks);
    cudaFree(ctx->d_tab); cudaFree(ctx->d_cursor); cudaFree(ctx->d_work); cudaFree(ctx->d_crc_recs);
    if (ctx->h_crc_recs) cudaFreeHost(ctx->h_crc_recs);
    if (ctx->s_in) cudaStreamSynchronize(ctx->s_in);
    if (ctx->s_out) cudaStreamSynchronize(ctx->s_out);
    for (int i = 0; i < 2; i++) {
        cudaFree(ctx->d_in[i]); cudaFree(ctx->d_out[i]);
        if (ctx->h_in[i]) cudaFreeHost(ctx->h_in[i]);
        if (ctx->h_out[i]) cudaFreeHost(ctx->h_out[i]);
        if (ctx->e_in[i]) cudaEventDestroy(ctx->e_in[i]);
        if (ctx->e_cmp[i]) cudaEventDestroy(ctx->e_cmp[i]);
        if (ctx->e_out[i]) cudaEventDestroy(ctx->e_out[i]);
        if (ctx->e_blk_out[i]) cudaEventDestroy(ctx->e_blk_out[i]);
        if (i == 0 && ctx->e_blk_cmp) cudaEventDestroy(ctx->e_blk_cmp);
        if (i == 0 && ctx->e_blk_in) cudaEventDestroy(ctx->e_blk_in);
    }
    if (ctx->s_in) cudaStreamDestroy(ctx->s_in);
    if (ctx->s_out) cudaStreamDestroy(ctx->s_out);
    if (ctx->h_cursor) cudaFreeHost(ctx->h_cursor);
    for (int i = 0; i < MAX_SUB; i++) for (int j = 0; j < 6; j++) if (ctx->ev[i][j]) cudaEventDestroy(ctx->ev[i][j]);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    (void)cudaGetLastError();
    delete ctx;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// device-resident path
// ---------------------------------------------------------------------------------------------
namespace {

// Enqueue the whole pipeline for d_in[0, n) on `st`.  No host synchronisation.
int enqueue_deflate(gzb_ctx* ctx, int level, uint32_t flags, const uint8_t* d_in, uint64_t n, uint32_t halo,
                    uint8_t* d_out, uint64_t out_cap, cudaStream_t st)
{
    if (level < 1 || level > 9) return fail(ctx, GZB_E_LEVEL, "level must be 1..9");
    const bool rs = (flags & GZB_F_RSYNCABLE) != 0;
    const uint32_t chunk = rs ? GZB_RSYNC_CHUNK : GZB_CHUNK, dict = rs ? 0u : GZB_DICT;
    if (n && (!d_in || !d_out)) return fail(ctx, GZB_E_ARG, "null buffer");
    if (((uintptr_t)d_in & 15u) || ((uintptr_t)d_out & 15u)) return fail(ctx, GZB_E_ARG, "d_in and d_out must be 16-byte aligned");
    if (halo != 0u && halo != dict) return fail(ctx, GZB_E_ARG, "halo must be 0 or the dictionary size");
    if (out_cap < gzb_bound(n, flags)) return fail(ctx, GZB_E_OUT_FULL, "out_cap < gzb_bound(n)");
    CU(ctx, cudaSetDevice(ctx->device));
    const uint64_t nchunks64 = (n + chunk - 1) / chunk;
    const uint32_t bpc = (chunk + BLOCK_SPAN - 1) / BLOCK_SPAN;
    LevelParams lp = level_params(level);
    // scheduling by level, measured on B200 (tools/variant_sweep.py, profiles/)
    lp.parse_warps = ctx->tune.pw ? ctx->tune.pw : (level <= 2 ? 16 : 8);
    const int help = ctx->tune.help >= 0 ? ctx->tune.help : (level >= 7 ? 3 : 0);
    lp.help_ins = help & 1; lp.help_parse = (help >> 1) & 1;
    lp.max_run = ctx->tune.max_run; lp.force_run = ctx->tune.force_run;
    lz77_role_map(lp, ctx->tune.rolemap >= 0 ? ctx->tune.rolemap : 4);
    ctx->n_sub_timed = 0;

    CU(ctx, cudaMemsetAsync(ctx->d_cursor, 0, sizeof(Cursor), st));
    if (n == 0) {
        // empty input: the member body is one final static block holding only EOB -- "03 00", what the
        // reference's single-stream path writes (its -j path deadlocks here)
        static const uint8_t empty_final[2] = {0x03, 0x00};
        Cursor c; memset(&c, 0, sizeof(c));
        if (flags & GZB_F_FINAL) {
            if (!d_out) return fail(ctx, GZB_E_ARG, "null buffer");
            CU(ctx, cudaMemcpyAsync(d_out, empty_final, 2, cudaMemcpyHostToDevice, st));
            c.out_bytes = 2; c.n_blocks = 1;
        }
        CU(ctx, cudaStreamSynchronize(st));
        *ctx->h_cursor = c;
        return GZB_OK;
    }
    const uint32_t sub = (uint32_t)(nchunks64 < ctx->batch_chunks ? nchunks64 : ctx->batch_chunks);
    int rc = ensure_scratch(ctx, sub, chunk);
    if (rc != GZB_OK) return rc;
    const uint64_t nsub = (nchunks64 + sub - 1) / sub;
    if (nsub > 65536) return fail(ctx, GZB_E_ARG, "input too large for one call");
    CU(ctx, cudaMemsetAsync(ctx->d_work, 0, sizeof(uint32_t) * nsub, st));

    for (uint64_t sb = 0; sb < nsub; sb++) {
        const uint64_t c0 = sb * sub;
        const uint32_t nch = (uint32_t)((nchunks64 - c0) < sub ? (nchunks64 - c0) : sub);
        const uint64_t off = c0 * chunk;
        const uint64_t nb = (n - off) < (uint64_t)nch * chunk ? (n - off) : (uint64_t)nch * chunk;
        const uint32_t h = (sb == 0) ? halo : dict;
        const int fin = ((flags & GZB_F_FINAL) && sb == nsub - 1) ? 1 : 0;
        const bool timed = sb < (uint64_t)MAX_SUB;
        cudaEvent_t* ev = ctx->ev[timed ? sb : 0];
        if (timed) { CU(ctx, cudaEventRecord(ev[0], st)); }
        launch_scan(d_in + off, nb, h, chunk, dict, nch, bpc, ctx->tune.scan, ctx->scr.chunks, ctx->scr.blocks, ctx->d_tab, st);
        if (timed) { CU(ctx, cudaEventRecord(ev[1], st)); }
        launch_lz77(d_in + off, nb, h, chunk, dict, nch, ctx->scr.tokens, ctx->scr.blocks, ctx->scr.chunks, lp,
                    ctx->d_work + sb, ctx->n_sm, st);
        if (timed) { CU(ctx, cudaEventRecord(ev[2], st)); }
        launch_huffman(ctx->scr.blocks, ctx->scr.chunks, nch, bpc, st);
        if (timed) { CU(ctx, cudaEventRecord(ev[3], st)); }
        launch_layout(ctx->scr.blocks, ctx->scr.chunks, nch, bpc, fin, ctx->d_cursor, ctx->d_tab, st);
        if (timed) { CU(ctx, cudaEventRecord(ev[4], st)); }
        launch_pack(d_in + off, chunk, ctx->scr.tokens, ctx->scr.blocks, ctx->scr.chunks, nch, bpc, d_out, st);
        if (timed) { CU(ctx, cudaEventRecord(ev[5], st)); ctx->n_sub_timed = (int)sb + 1; }
#ifdef GZB_TESTING
        if ((flags & GZB_F_VERIFY) && sb == 0 && ctx->test_corrupt_at >= 0 && (uint64_t)ctx->test_corrupt_at < gzb_bound(n, flags))
            CU(ctx, cudaMemsetAsync(d_out + ctx->test_corrupt_at, 0xA5, 1, st));
#endif
        if (flags & GZB_F_VERIFY)
            launch_verify(d_in + off, chunk, dict, h, ctx->scr.chunks, nch, d_out, fin, c0, ctx->d_cursor, st);
    }
    CU(ctx, cudaGetLastError());
    CU(ctx, cudaMemcpyAsync(ctx->h_cursor, ctx->d_cursor, sizeof(Cursor), cudaMemcpyDeviceToHost, st));
    return GZB_OK;
}

int verify_status(gzb_ctx* ctx, const Cursor& c)
{
    if (c.verify_bad == 0u) return GZB_OK;
    char msg[160];
    snprintf(msg, sizeof(msg), "verify: %u chunk(s) do not decode back to their input; first: chunk %llu, code %u (k_verify.cu)",
             c.verify_bad, (unsigned long long)c.verify_first, c.verify_code);
    return fail(ctx, GZB_E_VERIFY, msg);
}

void cursor_to_result(const Cursor& c, gzb_result* r)
{
    r->in_bytes = c.in_bytes; r->out_bytes = c.out_bytes; r->crc32 = c.crc;
    r->n_chunks = c.n_chunks; r->n_blocks = c.n_blocks; r->status = c.status;
}

}  // namespace

extern "C" {

// the two halves of a device call; ctx->mu is held by the caller
static int submit_locked(gzb_ctx* ctx, int level, uint32_t flags, const uint8_t* d_in, uint64_t n, uint32_t halo,
                         uint8_t* d_out, uint64_t out_cap, void* stream)
{
    // one call in flight per context: its cursor, work counters and scratch are shared
    if (ctx->pending) return fail(ctx, GZB_E_ARG, "a call is already pending on this context: gzb_deflate_wait() first");
    cudaStream_t st = stream ? (cudaStream_t)stream : ctx->stream;
    int rc = enqueue_deflate(ctx, level, flags, d_in, n, halo, d_out, out_cap, st);
    if (rc == GZB_OK) { ctx->pending = true; ctx->pending_n = n; ctx->pending_stream = st; }
    else if (ctx->stream) { cudaStreamSynchronize(st); (void)cudaGetLastError(); }     // nothing of a failed submit stays in flight
    return rc;
}
static int wait_locked(gzb_ctx* ctx, gzb_result* result)
{
    if (!ctx->pending) return fail(ctx, GZB_E_ARG, "nothing pending");
    ctx->pending = false;
    CU(ctx, cudaSetDevice(ctx->device));
    CU(ctx, cudaStreamSynchronize(ctx->pending_stream));
    if (ctx->h_cursor->in_bytes != ctx->pending_n) return fail(ctx, GZB_E_CUDA, "pipeline consumed a different number of bytes than submitted");
    if (result) cursor_to_result(*ctx->h_cursor, result);
    if (ctx->h_cursor->status == GZB_OK && ctx->h_cursor->verify_bad) {
        if (result) result->status = GZB_E_VERIFY;
        return verify_status(ctx, *ctx->h_cursor);
    }
    return ctx->h_cursor->status;
}

int gzb_deflate_device_async(gzb_ctx* ctx, int level, uint32_t flags, const uint8_t* d_in, uint64_t n, uint32_t halo,
                             uint8_t* d_out, uint64_t out_cap, void* stream)
{
    if (!ctx) return GZB_E_ARG;
    std::lock_guard<std::mutex> g(ctx->mu);
    return submit_locked(ctx, level, flags, d_in, n, halo, d_out, out_cap, stream);
}

int gzb_deflate_wait(gzb_ctx* ctx, gzb_result* result)
{
    if (!ctx) return GZB_E_ARG;
    std::lock_guard<std::mutex> g(ctx->mu);
    return wait_locked(ctx, result);
}

int gzb_deflate_device(gzb_ctx* ctx, int level, uint32_t flags, const uint8_t* d_in, uint64_t n, uint32_t halo,
                       uint8_t* d_out, uint64_t out_cap, gzb_result* result, void* stream)
{
    if (!ctx) return GZB_E_ARG;
    std::lock_guard<std::mutex> g(ctx->mu);                                  // one critical section: nobody slips in between
    int rc = submit_locked(ctx, level, flags, d_in, n, halo, d_out, out_cap, stream);
    if (rc != GZB_OK) return rc;
    return wait_locked(ctx, result);
}

#ifdef GZB_TESTING
// test build only (csrc/build/libgzb200_testing.so): the next GZB_F_VERIFY calls overwrite one output byte before k_verify runs
int gzb_test_corrupt_at(gzb_ctx* ctx, long long offset) { if (!ctx) return GZB_E_ARG; ctx->test_corrupt_at = offset; return GZB_OK; }
#endif

// debugging aid (not in gzb200.h): cycle counters of k_lz77's pipeline phases, see k_lz77.cu
int gzb_debug_lzprof(gzb_ctx* ctx, unsigned long long* out16, int reset)
{
    if (!ctx || !out16) return GZB_E_ARG;
    CU(ctx, cudaSetDevice(ctx->device));
    CU(ctx, cudaDeviceSynchronize());
    CU(ctx, lz77_read_prof(out16, reset));
    return GZB_OK;
}

int gzb_last_stage_ms(gzb_ctx* ctx, gzb_stage_ms* ms)
{
    if (!ctx || !ms) return GZB_E_ARG;
    std::lock_guard<std::mutex> g(ctx->mu);
    memset(ms, 0, sizeof(*ms));
    for (int i = 0; i < ctx->n_sub_timed; i++) {
        float t[5];
        for (int j = 0; j < 5; j++) CU(ctx, cudaEventElapsedTime(&t[j], ctx->ev[i][j], ctx->ev[i][j + 1]));
        ms->crc += t[0]; ms->lz77 += t[1]; ms->huffman += t[2]; ms->layout += t[3]; ms->pack += t[4];
    }
    ms->total = ms->crc + ms->lz77 + ms->huffman + ms->layout + ms->pack;
    return GZB_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// host-buffer path, CRC entry points, gzip framing
// ---------------------------------------------------------------------------------------------
namespace {

int ensure_dev(gzb_ctx* ctx, uint64_t in_cap, uint64_t out_cap)
{
    cudaError_t e = cudaSuccess;
    if (ctx->dev_in_cap < in_cap) {
        for (int i = 0; i < 2; i++) { cudaFree(ctx->d_in[i]); ctx->d_in[i] = nullptr; }
        ctx->dev_in_cap = 0;
        for (int i = 0; i < 2 && e == cudaSuccess; i++) e = cudaMalloc(&ctx->d_in[i], in_cap);
        if (e != cudaSuccess) { (void)cudaGetLastError(); return fail(ctx, GZB_E_NOMEM, "device staging (in)", e); }
        ctx->dev_in_cap = in_cap;
    }
    if (ctx->dev_out_cap < out_cap) {
        for (int i = 0; i < 2; i++) { cudaFree(ctx->d_out[i]); ctx->d_out[i] = nullptr; }
        ctx->dev_out_cap = 0;
        for (int i = 0; i < 2 && e == cudaSuccess; i++) e = cudaMalloc(&ctx->d_out[i], out_cap);
        if (e != cudaSuccess) { (void)cudaGetLastError(); return fail(ctx, GZB_E_NOMEM, "device staging (out)", e); }
        ctx->dev_out_cap = out_cap;
    }
    return GZB_OK;
}

int ensure_pinned(gzb_ctx* ctx, uint64_t in_cap, uint64_t out_cap)
{
    cudaError_t e = cudaSuccess;
    if (ctx->pin_in_cap < in_cap) {
        for (int i = 0; i < 2; i++) { if (ctx->h_in[i]) cudaFreeHost(ctx->h_in[i]); ctx->h_in[i] = nullptr; }
        ctx->pin_in_cap = 0;
        for (int i = 0; i < 2 && e == cudaSuccess; i++) e = cudaMallocHost(&ctx->h_in[i], in_cap);
        if (e != cudaSuccess) { (void)cudaGetLastError(); return fail(ctx, GZB_E_NOMEM, "pinned staging (in)", e); }
        ctx->pin_in_cap = in_cap;
    }
    if (ctx->pin_out_cap < out_cap) {
        for (int i = 0; i < 2; i++) { if (ctx->h_out[i]) cudaFreeHost(ctx->h_out[i]); ctx->h_out[i] = nullptr; }
        ctx->pin_out_cap = 0;
        for (int i = 0; i < 2 && e == cudaSuccess; i++) e = cudaMallocHost(&ctx->h_out[i], out_cap);
        if (e != cudaSuccess) { (void)cudaGetLastError(); return fail(ctx, GZB_E_NOMEM, "pinned staging (out)", e); }
        ctx->pin_out_cap = out_cap;
    }
    return GZB_OK;
}

bool is_pinned(const void* p)
{
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { (void)cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeHost;
}

}  // namespace

// Batch boundaries of the host path.  The H2D copy of the first batch and the D2H copy of the last one have nothing to
// hide behind, so a large input starts with growing batches, ends with a small one (EDGE) and runs in HOST_BATCH pieces in
// between.
static std::vector<uint64_t> host_batch_plan(uint64_t n, uint64_t HOST_BATCH, uint64_t EDGE)
{
    std::vector<uint64_t> boff;
    boff.push_back(0);
    if (EDGE && EDGE < HOST_BATCH && n > 4 * EDGE) {
        // sizes EDGE, 3 EDGE, 9 EDGE ... up to HOST_BATCH (a batch's H2D copy hides behind the kernels of the one before
        // it as long as it is at most ~3 times as large: PCIe moves ~55 GB/s, the kernels take ~20 GB/s at -6), the
        // last batch <= EDGE
        const uint64_t rem = n - EDGE, tail = rem % EDGE ? rem % EDGE : EDGE, end_main = n - tail;
        uint64_t pos = 0, sz = EDGE;
        while (pos + sz < end_main) { pos += sz; boff.push_back(pos); sz = sz * 3 < HOST_BATCH ? sz * 3 : HOST_BATCH; }
        boff.push_back(end_main);
    } else for (uint64_t p = HOST_BATCH; p < n; p += HOST_BATCH) boff.push_back(p);
    boff.push_back(n);
    return boff;
}

// host arithmetic only (no context): the sizes come from the same environment variables gzb_create reads
extern "C" int gzb_host_batch_plan(uint64_t n, uint64_t* off, int cap)
{
    if (!off) return GZB_E_ARG;
    const char* eb = getenv("GZB_HOST_BATCH_MB"); const char* ee = getenv("GZB_HOST_EDGE_MB");
    const uint64_t mb = eb && atoll(eb) > 0 ? (uint64_t)atoll(eb) : 512ull, edge = ee && atoll(ee) >= 0 ? (uint64_t)atoll(ee) : 64ull;
    const std::vector<uint64_t> b = host_batch_plan(n, mb << 20, edge << 20);
    if ((size_t)cap < b.size()) return GZB_E_ARG;
    for (size_t i = 0; i < b.size(); i++) off[i] = b[i];
    return (int)b.size() - 1;
}

extern "C" {

// Host buffers in, host buffers out.  The input is cut into batches of HOST_BATCH bytes that flow through a
// three-stage pipeline on three streams -- H2D of batch i+1 || kernels of batch i || D2H of batch i-1 -- over two
// slots of device buffers.  Pinned caller memory (cudaHostAlloc / cudaHostRegister / torch pin_memory) is DMA'd
// directly; pageable memory goes through the context's pinned staging with one memcpy on each side.
int gzb_deflate_host(gzb_ctx* ctx, int level, uint32_t flags, const uint8_t* in, uint64_t n, const uint8_t* dict,
                     uint8_t* out, uint64_t out_cap, gzb_result* result)
{
    if (!ctx || (n && !in) || !out) return GZB_E_ARG;
    if (level < 1 || level > 9) return fail(ctx, GZB_E_LEVEL, "level must be 1..9");
    // out_cap >= gzb_bound(n) can never overflow; a smaller buffer is accepted and checked batch by batch (GZB_E_OUT_FULL
    // before anything is written past it): a caller that knows its ratio need not page-lock n bytes of output
    const bool rs = (flags & GZB_F_RSYNCABLE) != 0;
    const uint32_t dsz = rs ? 0u : GZB_DICT;
    std::lock_guard<std::mutex> g(ctx->mu);
    if (ctx->pending) return fail(ctx, GZB_E_ARG, "a device call is pending on this context: gzb_deflate_wait() first");
    CU(ctx, cudaSetDevice(ctx->device));
    const uint64_t HOST_BATCH = ctx->tune.host_batch;
    const std::vector<uint64_t> boff = host_batch_plan(n, HOST_BATCH, ctx->tune.host_edge);   // batch i = [boff[i], boff[i + 1])
    const uint64_t nbatch = boff.size() - 1;
    const uint64_t bmax = n < HOST_BATCH ? n : HOST_BATCH;
    const bool pin_in = n && is_pinned(in) && (!dict || is_pinned(dict)), pin_out = is_pinned(out);
    int rc = ensure_dev(ctx, dsz + bmax + 16, gzb_bound(bmax, flags));
    if (rc == GZB_OK) rc = ensure_pinned(ctx, pin_in ? 0 : dsz + bmax + 16, pin_out ? 0 : gzb_bound(bmax, flags));
    if (rc != GZB_OK) return rc;

    gzb_result total; memset(&total, 0, sizeof(total));
    uint64_t opos = 0;
    struct Pend { uint64_t bytes, opos; bool live; } pend[2] = {{0, 0, false}, {0, 0, false}};   // D2H in flight per slot

    auto batch_len = [&](uint64_t i) { return boff[i + 1] - boff[i]; };
    auto issue_h2d = [&](uint64_t i) -> int {
        const int sl = (int)(i & 1);
        const uint64_t off = boff[i], nb = batch_len(i);
        const uint8_t* halo_src = (off == 0) ? dict : in + off - dsz;       // the 32 KiB before the batch
        const uint32_t halo = (halo_src && dsz) ? dsz : 0u;
        CU(ctx, cudaStreamWaitEvent(ctx->s_in, ctx->e_cmp[sl], 0));        // kernels of batch i-2 have read this slot
        uint8_t* dst = ctx->d_in[sl] + (dsz - halo);                        // keeps d_in + dsz 16-byte aligned
        if (pin_in) {
            if (off == 0) {
                if (halo) CU(ctx, cudaMemcpyAsync(dst, halo_src, halo, cudaMemcpyHostToDevice, ctx->s_in));
                if (nb) CU(ctx, cudaMemcpyAsync(dst + halo, in, nb, cudaMemcpyHostToDevice, ctx->s_in));
            } else CU(ctx, cudaMemcpyAsync(dst, in + off - halo, halo + nb, cudaMemcpyHostToDevice, ctx->s_in));
        } else {
            CU(ctx, cudaEventSynchronize(ctx->e_in[sl]));                   // staging of batch i-2 has left the host
            uint8_t* hin = ctx->h_in[sl] + (dsz - halo);
            if (halo) memcpy(hin, halo_src, halo);
            if (nb) memcpy(hin + halo, in + off, nb);
            CU(ctx, cudaMemcpyAsync(dst, hin, halo + nb, cudaMemcpyHostToDevice, ctx->s_in));
        }
        CU(ctx, cudaEventRecord(ctx->e_in[sl], ctx->s_in));
        return GZB_OK;
    };
    auto drain = [&](int sl) -> int {                                      // complete the D2H of a slot
        if (!pend[sl].live) return GZB_OK;
        CU(ctx, cudaEventSynchronize(ctx->e_out[sl]));
        if (!pin_out) memcpy(out + pend[sl].opos, ctx->h_out[sl], pend[sl].bytes);
        pend[sl].live = false;
        return GZB_OK;
    };
    auto pipeline = [&]() -> int {
        int r;
        if ((r = issue_h2d(0)) != GZB_OK) return r;
        for (uint64_t i = 0; i < nbatch; i++) {
            const int sl = (int)(i & 1);
            const uint64_t off = boff[i], nb = batch_len(i);
            const uint32_t halo = ((off == 0 ? dict != nullptr : true) && dsz) ? dsz : 0u;
            const uint32_t f = (flags & ~GZB_F_FINAL) | ((i + 1 == nbatch && (flags & GZB_F_FINAL)) ? GZB_F_FINAL : 0u);
            if ((r = drain(sl)) != GZB_OK) return r;                        // d_out[sl] / h_out[sl] are free again
            CU(ctx, cudaStreamWaitEvent(ctx->stream, ctx->e_in[sl], 0));
            r = enqueue_deflate(ctx, level, f, ctx->d_in[sl] + dsz, nb, halo, ctx->d_out[sl], ctx->dev_out_cap, ctx->stream);
            if (r != GZB_OK) return r;
            CU(ctx, cudaEventRecord(ctx->e_cmp[sl], ctx->stream));
            if (i + 1 < nbatch && (r = issue_h2d(i + 1)) != GZB_OK) return r; // overlaps the kernels of batch i
            CU(ctx, cudaStreamSynchronize(ctx->stream));                    // out_bytes of batch i is known now
            const Cursor c = *ctx->h_cursor;
            if (c.status != GZB_OK) return fail(ctx, c.status, "device pipeline reported an error");
            if (c.verify_bad) return verify_status(ctx, c);
            if (c.in_bytes != nb) return fail(ctx, GZB_E_CUDA, "pipeline consumed a different number of bytes than submitted");
            if (opos + c.out_bytes > out_cap) return fail(ctx, GZB_E_OUT_FULL, "output overflow");
            if (c.out_bytes) CU(ctx, cudaMemcpyAsync(pin_out ? out + opos : ctx->h_out[sl], ctx->d_out[sl], c.out_bytes, cudaMemcpyDeviceToHost, ctx->s_out));
            CU(ctx, cudaEventRecord(ctx->e_out[sl], ctx->s_out));
            pend[sl].bytes = c.out_bytes; pend[sl].opos = opos; pend[sl].live = true;
            opos += c.out_bytes;
            total.crc32 = gzb_crc32_combine(total.crc32, c.crc, nb);
            total.in_bytes += nb; total.n_chunks += c.n_chunks; total.n_blocks += c.n_blocks;
        }
        for (int sl = 0; sl < 2; sl++) if ((r = drain(sl)) != GZB_OK) return r;
        return GZB_OK;
    };
    rc = pipeline();
    if (rc != GZB_OK) {
        // One exit for every failure: nothing stays in flight.  The H2D copy of the next batch and the D2H copies of the
        // earlier ones may still be running -- into the caller's own buffers when those are pinned -- so all three streams
        // are drained before the caller gets its memory back; the detail text of the first failure is kept.
        char keep[sizeof(ctx->err)]; memcpy(keep, ctx->err, sizeof(keep));
        cudaStreamSynchronize(ctx->s_in); cudaStreamSynchronize(ctx->stream); cudaStreamSynchronize(ctx->s_out);
        (void)cudaGetLastError();
        memcpy(ctx->err, keep, sizeof(keep));
        return rc;
    }
    total.out_bytes = opos; total.status = GZB_OK;
    if (result) *result = total;
    return GZB_OK;
}

// Host buffers over SEVERAL GPUs in one call: the input is cut into contiguous shards on chunk boundaries (one per
// context, sizes as even as the grid allows), every shard runs gzb_deflate_host on its own context and host thread with
// the 32 KiB before it as dictionary, and lands at out + shards[i].out_off (= gzb_bound of the bytes before it, so the
// shards cannot collide).  No collective: crc32_combine and concatenation are host arithmetic (src/parallel.c:538-546).
// With compact != 0 the bodies are moved together so that out[0, result->out_bytes) is the fragment run of the whole input.
int gzb_deflate_host_multi(gzb_ctx** ctxs, int n_ctx, int level, uint32_t flags, const uint8_t* in, uint64_t n, const uint8_t* dict,
                           uint8_t* out, uint64_t out_cap, gzb_result* result, gzb_shard* shards, int compact)
{
    if (!ctxs || n_ctx < 1 || n_ctx > 64 || (n && !in) || !out) return GZB_E_ARG;
    for (int i = 0; i < n_ctx; i++) if (!ctxs[i]) return GZB_E_ARG;
    const uint64_t chunk = (flags & GZB_F_RSYNCABLE) ? GZB_RSYNC_CHUNK : GZB_CHUNK;
    const uint64_t dsz = (flags & GZB_F_RSYNCABLE) ? 0u : GZB_DICT;
    const uint64_t nchunks = (n + chunk - 1) / chunk;
    std::vector<gzb_shard> sh((size_t)n_ctx);
    std::vector<gzb_result> res((size_t)n_ctx);
    std::vector<int> rcs((size_t)n_ctx, GZB_OK);
    uint64_t c0 = 0;
    for (int i = 0; i < n_ctx; i++) {
        const uint64_t per = nchunks / (uint64_t)n_ctx + ((uint64_t)i < nchunks % (uint64_t)n_ctx ? 1u : 0u);
        const uint64_t lo = c0 * chunk < n ? c0 * chunk : n, hi = (c0 + per) * chunk < n ? (c0 + per) * chunk : n;
        sh[(size_t)i].in_off = lo; sh[(size_t)i].in_bytes = hi - lo; sh[(size_t)i].out_off = i ? gzb_bound(lo, flags) : 0; sh[(size_t)i].out_bytes = 0;
        c0 += per;
    }
    if (out_cap < gzb_bound(n, flags) + (uint64_t)n_ctx * 64u) return fail(ctxs[0], GZB_E_OUT_FULL, "out_cap < gzb_bound(n) + 64 per shard");
    int last = 0;                                             // the last shard that holds bytes ends the stream
    for (int i = 0; i < n_ctx; i++) if (sh[(size_t)i].in_bytes || i == 0) last = i;
    auto work = [&](int i) {
        const gzb_shard& q = sh[(size_t)i];
        if (q.in_bytes == 0 && i != last) { memset(&res[(size_t)i], 0, sizeof(gzb_result)); return; }
        const uint8_t* d = q.in_off == 0 ? dict : (dsz ? in + q.in_off - dsz : nullptr);
        const uint32_t f = (flags & ~GZB_F_FINAL) | ((i == last && (flags & GZB_F_FINAL)) ? GZB_F_FINAL : 0u);
        rcs[(size_t)i] = gzb_deflate_host(ctxs[i], level, f, in + q.in_off, q.in_bytes, d, out + q.out_off, gzb_bound(q.in_bytes, flags), &res[(size_t)i]);
    };
    std::vector<std::thread> th;
    for (int i = 1; i < n_ctx; i++) th.emplace_back(work, i);
    work(0);
    for (auto& t : th) t.join();
    for (int i = 0; i < n_ctx; i++) if (rcs[(size_t)i] != GZB_OK) {
        if (i) snprintf(ctxs[0]->err, sizeof(ctxs[0]->err), "shard %d: %.480s", i, ctxs[i]->err);
        return rcs[(size_t)i];
    }
    gzb_result total; memset(&total, 0, sizeof(total));
    uint64_t opos = 0;
    for (int i = 0; i < n_ctx; i++) {
        const gzb_result& r = res[(size_t)i];
        sh[(size_t)i].out_bytes = r.out_bytes;
        if (compact) { if (sh[(size_t)i].out_off != opos && r.out_bytes) memmove(out + opos, out + sh[(size_t)i].out_off, r.out_bytes); sh[(size_t)i].out_off = opos; }
        opos += r.out_bytes;
        total.crc32 = gzb_crc32_combine(total.crc32, r.crc32, sh[(size_t)i].in_bytes);
        total.in_bytes += r.in_bytes; total.out_bytes += r.out_bytes; total.n_chunks += r.n_chunks; total.n_blocks += r.n_blocks;
    }
    if (shards) for (int i = 0; i < n_ctx; i++) shards[i] = sh[(size_t)i];
    if (result) *result = total;
    return GZB_OK;
}

// ctx->mu is held by the caller
static int crc32_device_locked(gzb_ctx* ctx, uint32_t crc, const uint8_t* d_in, uint64_t n, uint32_t* crc_out, cudaStream_t st)
{
    constexpr uint64_t SUB = 65536;                                         // chunks per launch
    const uint32_t chunk = GZB_CHUNK;
    if (n && !ctx->d_crc_recs) {                                            // scratch of the context, allocated on first use
        if (cudaMalloc(&ctx->d_crc_recs, sizeof(ChunkRec) * SUB) != cudaSuccess ||
            cudaMallocHost(&ctx->h_crc_recs, sizeof(ChunkRec) * SUB) != cudaSuccess) {
            cudaFree(ctx->d_crc_recs); ctx->d_crc_recs = nullptr; (void)cudaGetLastError();
            return fail(ctx, GZB_E_NOMEM, "crc32 scratch");
        }
    }
    uint32_t acc = crc;
    const uint64_t nchunks64 = (n + chunk - 1) / chunk;
    for (uint64_t c0 = 0; c0 < nchunks64; c0 += SUB) {
        const uint32_t nch = (uint32_t)((nchunks64 - c0) < SUB ? (nchunks64 - c0) : SUB);
        const uint64_t off = c0 * chunk, nb = (n - off) < (uint64_t)nch * chunk ? (n - off) : (uint64_t)nch * chunk;
        launch_scan(d_in + off, nb, 0u, chunk, 0u, nch, 1u, 0, ctx->d_crc_recs, nullptr, ctx->d_tab, st);
        CU(ctx, cudaMemcpyAsync(ctx->h_crc_recs, ctx->d_crc_recs, sizeof(ChunkRec) * nch, cudaMemcpyDeviceToHost, st));
        CU(ctx, cudaStreamSynchronize(st));
        for (uint32_t i = 0; i < nch; i++) {
            uint64_t len = (uint64_t)(i + 1) * chunk <= nb ? chunk : nb - (uint64_t)i * chunk;
            acc = crc_shift(ctx->h_x2n, acc, len) ^ ctx->h_crc_recs[i].crc;
        }
    }
    *crc_out = acc;
    return GZB_OK;
}

// internal (gzb_internal.h): batches through callbacks, for the fd driver.  One thread per context.
int gzb_internal_deflate_batches(gzb_ctx* ctx, int level, uint32_t flags, const gzb_batch_io* io)
{
    if (!ctx || !io || !io->fetch || !io->emit) return GZB_E_ARG;
    if (level < 1 || level > 9) return fail(ctx, GZB_E_LEVEL, "level must be 1..9");
    const uint32_t dsz = (flags & GZB_F_RSYNCABLE) ? 0u : GZB_DICT;
    std::lock_guard<std::mutex> g(ctx->mu);
    if (ctx->pending) return fail(ctx, GZB_E_ARG, "a device call is pending on this context: gzb_deflate_wait() first");
    CU(ctx, cudaSetDevice(ctx->device));
    const uint64_t bmax = io->max_bytes, ocap = gzb_bound(bmax, flags);
    int rc = ensure_dev(ctx, dsz + bmax + 16, ocap);
    if (rc == GZB_OK) rc = ensure_pinned(ctx, dsz + bmax + 16, ocap);
    if (rc != GZB_OK) return rc;
    const bool blk = io->blocking != 0;
    // live: the fragment is on its way to / in h_out and has not been emitted; busy: emitted, the caller may still read h_out
    struct Slot { uint64_t len = 0; int have_dict = 0, final = 0; gzb_result res; bool live = false, busy = false; uint64_t k = 0; } slot[2];

    auto wait_event = [&](cudaEvent_t spin, cudaEvent_t sleep, cudaStream_t st) -> cudaError_t {
        if (!blk) return cudaEventSynchronize(spin);
        cudaError_t e = cudaEventRecord(sleep, st);                         // behind `spin` on the same stream
        return e != cudaSuccess ? e : cudaEventSynchronize(sleep);
    };
    auto fetch_h2d = [&](uint64_t k) -> int {                               // batch k into slot k & 1
        const int sl = (int)(k & 1);
        Slot& S = slot[sl];
        CU(ctx, wait_event(ctx->e_in[sl], ctx->e_blk_in, ctx->s_in));      // the staging buffer's previous batch has left the host
        uint8_t* hin = ctx->h_in[sl] + dsz;
        S.len = 0; S.have_dict = 0; S.final = 0;
        int r = io->fetch(io->user, k, hin, dsz, &S.len, &S.have_dict, &S.final);
        if (r != GZB_OK) return fail(ctx, r, "fetch callback failed");
        if (S.len > bmax) return fail(ctx, GZB_E_ARG, "fetch returned more than max_bytes");
        const uint32_t halo = (S.have_dict && dsz) ? dsz : 0u;
        CU(ctx, cudaStreamWaitEvent(ctx->s_in, ctx->e_cmp[sl], 0));        // kernels of batch k-2 have read this device slot
        if (halo + S.len) CU(ctx, cudaMemcpyAsync(ctx->d_in[sl] + (dsz - halo), hin - halo, halo + S.len, cudaMemcpyHostToDevice, ctx->s_in));
        CU(ctx, cudaEventRecord(ctx->e_in[sl], ctx->s_in));
        return GZB_OK;
    };
    auto emit_slot = [&](int sl) -> int {                                   // hand the fragment of a slot to the caller
        Slot& S = slot[sl];
        if (!S.live) return GZB_OK;
        CU(ctx, wait_event(ctx->e_out[sl], ctx->e_blk_out[sl], ctx->s_out));
        S.live = false; S.busy = io->emit_done != nullptr;
        int r = io->emit(io->user, S.k, ctx->h_out[sl], &S.res);
        return r == GZB_OK ? GZB_OK : fail(ctx, r, "emit callback failed");
    };
    auto release_slot = [&](int sl) -> int {                                // the caller is through with h_out[sl]
        Slot& S = slot[sl];
        if (!S.busy) return GZB_OK;
        S.busy = false;
        int r = io->emit_done(io->user, S.k);
        return r == GZB_OK ? GZB_OK : fail(ctx, r, "emit_done callback failed");
    };
    auto pipeline = [&]() -> int {
        int r;
        if (io->n_batches == 0) return GZB_OK;
        if ((r = fetch_h2d(0)) != GZB_OK) return r;
        for (uint64_t k = 0; k < io->n_batches; k++) {
            const int sl = (int)(k & 1);
            Slot& S = slot[sl];
            const uint32_t halo = (S.have_dict && dsz) ? dsz : 0u;
            const uint32_t f = (flags & ~GZB_F_FINAL) | (S.final ? GZB_F_FINAL : 0u);
            // the kernels go first: the device output slot is free once the D2H of batch k-2 is through (a wait on the GPU) ...
            CU(ctx, cudaStreamWaitEvent(ctx->stream, ctx->e_in[sl], 0));
            CU(ctx, cudaStreamWaitEvent(ctx->stream, ctx->e_out[sl], 0));
            r = enqueue_deflate(ctx, level, f, ctx->d_in[sl] + dsz, S.len, halo, ctx->d_out[sl], ctx->dev_out_cap, ctx->stream);
            if (r != GZB_OK) return r;
            CU(ctx, cudaEventRecord(ctx->e_cmp[sl], ctx->stream));
            // ... and under them the host reads batch k+1 -- first, because read + H2D of a batch together take about as long as
            // the kernels of one, and the next kernels can only start when they are through -- and passes on batch k-1
            if (k + 1 < io->n_batches && (r = fetch_h2d(k + 1)) != GZB_OK) return r;
            if ((r = emit_slot(sl ^ 1)) != GZB_OK) return r;
            CU(ctx, wait_event(ctx->e_cmp[sl], ctx->e_blk_cmp, ctx->stream));   // out_bytes of batch k is known now
            if ((r = release_slot(sl)) != GZB_OK) return r;                  // batch k-2 has left the pinned slot
            const Cursor c = *ctx->h_cursor;
            if (c.status != GZB_OK) return fail(ctx, c.status, "device pipeline reported an error");
            if (c.verify_bad) return verify_status(ctx, c);
            if (c.in_bytes != S.len) return fail(ctx, GZB_E_CUDA, "pipeline consumed a different number of bytes than submitted");
            if (c.out_bytes > ocap) return fail(ctx, GZB_E_OUT_FULL, "output overflow");
            if (c.out_bytes) CU(ctx, cudaMemcpyAsync(ctx->h_out[sl], ctx->d_out[sl], c.out_bytes, cudaMemcpyDeviceToHost, ctx->s_out));
            CU(ctx, cudaEventRecord(ctx->e_out[sl], ctx->s_out));
            memset(&S.res, 0, sizeof(S.res));
            S.res.in_bytes = S.len; S.res.out_bytes = c.out_bytes; S.res.crc32 = c.crc; S.res.n_chunks = c.n_chunks; S.res.n_blocks = c.n_blocks;
            S.k = k; S.live = true;
        }
        const int last = (int)((io->n_batches - 1) & 1);
        if ((r = emit_slot(last)) != GZB_OK) return r;
        if ((r = release_slot(last ^ 1)) != GZB_OK) return r;
        return release_slot(last);
    };
    rc = pipeline();
    if (rc != GZB_OK) {                                                     // as in gzb_deflate_host: nothing stays in flight
        char keep[sizeof(ctx->err)]; memcpy(keep, ctx->err, sizeof(keep));
        for (int sl = 0; sl < 2; sl++) if (slot[sl].busy) { slot[sl].busy = false; (void)io->emit_done(io->user, slot[sl].k); }
        cudaStreamSynchronize(ctx->s_in); cudaStreamSynchronize(ctx->stream); cudaStreamSynchronize(ctx->s_out);
        (void)cudaGetLastError();
        memcpy(ctx->err, keep, sizeof(keep));
    }
    return rc;
}

int gzb_crc32_device(gzb_ctx* ctx, uint32_t crc, const uint8_t* d_in, uint64_t n, uint32_t* crc_out, void* stream)
{
    if (!ctx || !crc_out || (n && !d_in)) return GZB_E_ARG;
    if ((uintptr_t)d_in & 15u) return fail(ctx, GZB_E_ARG, "d_in must be 16-byte aligned");
    std::lock_guard<std::mutex> g(ctx->mu);
    CU(ctx, cudaSetDevice(ctx->device));
    return crc32_device_locked(ctx, crc, d_in, n, crc_out, stream ? (cudaStream_t)stream : ctx->stream);
}

int gzb_crc32_host(gzb_ctx* ctx, uint32_t crc, const uint8_t* in, uint64_t n, uint32_t* crc_out)
{
    if (!ctx || !crc_out || (n && !in)) return GZB_E_ARG;
    const uint64_t B = 128ull << 20;
    std::lock_guard<std::mutex> g(ctx->mu);                                 // held across staging, copy and kernel of every batch
    if (ctx->pending) return fail(ctx, GZB_E_ARG, "a device call is pending on this context: gzb_deflate_wait() first");
    CU(ctx, cudaSetDevice(ctx->device));
    uint32_t acc = crc;
    for (uint64_t off = 0; off < n; off += B) {
        const uint64_t nb = (n - off) < B ? (n - off) : B;
        int rc = ensure_dev(ctx, GZB_DICT + nb + 16, 0);
        if (rc == GZB_OK) rc = ensure_pinned(ctx, GZB_DICT + nb + 16, 0);
        if (rc != GZB_OK) return rc;
        memcpy(ctx->h_in[0], in + off, nb);
        CU(ctx, cudaMemcpyAsync(ctx->d_in[0], ctx->h_in[0], nb, cudaMemcpyHostToDevice, ctx->stream));
        rc = crc32_device_locked(ctx, acc, ctx->d_in[0], nb, &acc, ctx->stream);
        if (rc != GZB_OK) { cudaStreamSynchronize(ctx->stream); (void)cudaGetLastError(); return rc; }
    }
    *crc_out = acc;
    return GZB_OK;
}

int gzb_host_alloc(gzb_ctx* ctx, uint64_t bytes, void** p)
{
    if (!ctx || !p) return GZB_E_ARG;
    *p = nullptr;
    if (cudaSetDevice(ctx->device) != cudaSuccess) return fail(ctx, GZB_E_CUDA, "cudaSetDevice");
    if (cudaHostAlloc(p, bytes ? bytes : 1, cudaHostAllocPortable) != cudaSuccess) { cudaGetLastError(); *p = nullptr; return fail(ctx, GZB_E_NOMEM, "cudaHostAlloc"); }
    return GZB_OK;
}

void gzb_host_free(void* p)
{
    if (p) cudaFreeHost(p);
}

uint32_t gzb_crc32_combine(uint32_t crc1, uint32_t crc2, uint64_t len2)
{
    static uint32_t x2n[64];
    static std::once_flag once;
    std::call_once(once, [] { uint32_t t[CRC_TAB_WORDS]; build_crc_tables(t); memcpy(x2n, t + 1024, sizeof(x2n)); });
    return crc_shift(x2n, crc1, len2) ^ crc2;
}

int gzb_gzip_header(int level, uint32_t mtime, const char* name, uint8_t* out, size_t cap)
{
    const size_t nl = name ? strlen(name) + 1 : 0;
    if (!out || cap < 10 + nl) return GZB_E_ARG;
    out[0] = 0x1f; out[1] = 0x8b; out[2] = 8; out[3] = name ? 8 : 0;
    out[4] = (uint8_t)mtime; out[5] = (uint8_t)(mtime >> 8); out[6] = (uint8_t)(mtime >> 16); out[7] = (uint8_t)(mtime >> 24);
    out[8] = (uint8_t)(level >= 9 ? 2 : level == 1 ? 4 : 0);     // XFL, src/parallel.c:43
    out[9] = 3;                                                    // OS = Unix
    if (name) memcpy(out + 10, name, nl);
    return (int)(10 + nl);
}

int gzb_gzip_trailer(uint32_t crc32, uint64_t isize, uint8_t* out, size_t cap)
{
    if (!out || cap < 8) return GZB_E_ARG;
    const uint32_t sz = (uint32_t)isize;                           // mod 2^32, src/parallel.c:60
    for (int i = 0; i < 4; i++) { out[i] = (uint8_t)(crc32 >> (8 * i)); out[4 + i] = (uint8_t)(sz >> (8 * i)); }
    return 8;
}

}  // extern "C"
