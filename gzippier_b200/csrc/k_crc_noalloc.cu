// k_crc_noalloc.cu -- gzb_crc32_device_noalloc: CRC-32 of a device buffer without a single allocation.
//
// No counterpart in the reference: added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  It is a variant of the CRC pass of
// k_scan.cu (gzb_crc32_device): that one writes one record per 128 KiB chunk into scratch the context cudaMalloc'ed and
// folds the records on the host; this one keeps the per-chunk values in a file-scope __device__ array and folds them on
// the device, so the call is one memset node + one kernel on the caller's stream and the result lands in a device word
// the caller owns.
//
// The scratch array is shared by every call on a device and keeps the previous call's contents:
//   * the arrival counter is cleared on the stream at the start of every call;
//   * the per-chunk values are all rewritten before they are read (only the first nchunks are read);
//   * the tables are written once per device at set-up and never again.
// Two calls that overlap on the same GPU (different streams, or two host threads) race on it: serialise them.
#include <mutex>
#include "gzb_common.cuh"
#include "gzb200.h"

namespace gzb {
namespace {

constexpr int      NA_THREADS    = 256;
constexpr uint32_t NA_MAX_CHUNKS = (uint32_t)(GZB_NOALLOC_MAX_BYTES / GZB_CHUNK);            // 65536
constexpr uint32_t NA_TAB   = 0;                                                  // [0, 1088): slice-by-4 tables + x^(2^k)
constexpr uint32_t NA_COUNT = 1088;                                               // CTAs that have finished
constexpr uint32_t NA_CRCS  = 1152;                                               // one raw CRC register per chunk
constexpr uint32_t NA_WORDS = NA_CRCS + NA_MAX_CHUNKS;
static_assert(CRC_TAB_WORDS <= (int)NA_COUNT && GZB_NOALLOC_MAX_BYTES % GZB_CHUNK == 0, "scratch layout");

// ~260 KiB per GPU that loads the library, held until the process exits
static __device__ __align__(256) uint32_t g_noalloc_scratch[NA_WORDS];

struct NaSmem {
    uint32_t tab[4][256];
    uint32_t x2n[64];
    uint32_t crc[NA_THREADS];
    uint64_t len[NA_THREADS];
    uint32_t last;
};

// folds S.crc / S.len over the CTA: afterwards S.crc[0] is the register of the concatenation (a shift by the bytes to the
// right of each part, then xor), S.len[0] its length
__device__ __forceinline__ void fold_cta(NaSmem& S, uint32_t tid)
{
    for (uint32_t stride = 1; stride < (uint32_t)NA_THREADS; stride <<= 1) {
        if ((tid & (2u * stride - 1u)) == 0u) {
            const uint64_t lb = S.len[tid + stride];
            S.crc[tid] = crc_shift(S.x2n, S.crc[tid], lb) ^ S.crc[tid + stride];
            S.len[tid] += lb;
        }
        __syncthreads();
    }
}

// One CTA per chunk.  s is the device address of g_noalloc_scratch (device code could name the array itself; the address
// is passed so that what the host looked up is what the kernel uses).  The chunk's register (no preset, no final xor: the
// CRC is linear without them) goes to s[NA_CRCS + chunk]; the CTA that arrives last folds all of them, adds the shifted
// running value and writes *out.
__global__ void __launch_bounds__(NA_THREADS)
k_crc_noalloc(const uint8_t* __restrict__ in, uint64_t n_bytes, uint32_t crc_in, uint32_t* __restrict__ s, uint32_t* __restrict__ out)
{
    __shared__ __align__(16) NaSmem S;
    const uint32_t tid = threadIdx.x;
    for (uint32_t i = tid; i < 1024u; i += NA_THREADS) (&S.tab[0][0])[i] = s[NA_TAB + i];
    if (tid < 64u) S.x2n[tid] = s[NA_TAB + 1024u + tid];
    __syncthreads();

    const uint64_t lo = (uint64_t)blockIdx.x * GZB_CHUNK;
    const uint32_t nc = (uint32_t)min((uint64_t)GZB_CHUNK, n_bytes - lo);
    const uint32_t slice = GZB_CHUNK / NA_THREADS;              // 512 bytes per thread, 16-byte aligned
    const uint32_t a = min(tid * slice, nc), b = min(a + slice, nc);
    const uint8_t* p = in + lo;
    uint32_t crc = 0u, i = a;
    for (; i + 16u <= b; i += 16u) {
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(p + i));
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int k = 0; k < 4; k++) {
            crc ^= w[k];
            crc = S.tab[3][crc & 0xFFu] ^ S.tab[2][(crc >> 8) & 0xFFu] ^ S.tab[1][(crc >> 16) & 0xFFu] ^ S.tab[0][crc >> 24];
        }
    }
    for (; i < b; i++) crc = S.tab[0][(crc ^ p[i]) & 0xFFu] ^ (crc >> 8);
    S.crc[tid] = crc; S.len[tid] = b - a;
    __syncthreads();
    fold_cta(S, tid);

    if (tid == 0) {
        s[NA_CRCS + blockIdx.x] = S.crc[0];
        __threadfence();
        S.last = (atomicAdd(&s[NA_COUNT], 1u) == gridDim.x - 1u) ? 1u : 0u;
    }
    __syncthreads();
    if (!S.last) return;
    __threadfence();

    // the last CTA: thread t folds chunks [t * per, (t + 1) * per) left to right, then the tree over the threads
    const uint32_t nchunks = gridDim.x, per = (nchunks + NA_THREADS - 1u) / NA_THREADS;
    const uint32_t c0 = min(tid * per, nchunks), c1 = min(c0 + per, nchunks);
    uint32_t acc = 0u; uint64_t len = 0;
    for (uint32_t c = c0; c < c1; c++) {
        const uint64_t l = min((uint64_t)GZB_CHUNK, n_bytes - (uint64_t)c * GZB_CHUNK);
        acc = crc_shift(S.x2n, acc, l) ^ __ldcg(&s[NA_CRCS + c]);
        len += l;
    }
    S.crc[tid] = acc; S.len[tid] = len;
    __syncthreads();
    fold_cta(S, tid);
    // crc(running || data): the preset and the final xor cancel into  shift(crc_in ^ ~0, n) ^ raw(data) ^ ~0
    if (tid == 0) *out = crc_shift(S.x2n, crc_in ^ 0xFFFFFFFFu, n_bytes) ^ S.crc[0] ^ 0xFFFFFFFFu;
}

// ---- set-up, once per device -------------------------------------------------------------------------------------
constexpr int NA_MAX_DEVICES = 64;
std::mutex g_na_mu;
uint32_t*  g_na_ptr[NA_MAX_DEVICES];          // device address of g_noalloc_scratch on each device (0 = not set up yet)

void na_tables(uint32_t* tab)                 // the same tables gzb_create uploads (make_crc_table, lib/zlib/crc32.c:89-135)
{
    for (uint32_t n = 0; n < 256; n++) {
        uint32_t c = n;
        for (int k = 0; k < 8; k++) c = (c & 1u) ? CRC_POLY ^ (c >> 1) : c >> 1;
        tab[n] = c;
    }
    for (uint32_t n = 0; n < 256; n++)
        for (int k = 1; k < 4; k++) tab[k * 256 + n] = tab[tab[(k - 1) * 256 + n] & 0xFFu] ^ (tab[(k - 1) * 256 + n] >> 8);
    uint32_t p = 1u << 30;
    for (int k = 0; k < 64; k++) { tab[1024 + k] = p; p = gf2_mul(p, p); }
}

// The array's name is only a handle in host code: its device address comes from cudaGetSymbolAddress, asked for with the
// device current.  No allocation: the array is part of the module image.  The table upload is a synchronous copy, so the
// first call on a device must not be made under stream capture (call once with n = 0 beforehand).
int na_setup(int dev, uint32_t** ptr)
{
    std::lock_guard<std::mutex> g(g_na_mu);
    if (!g_na_ptr[dev]) {
        uint32_t* p = nullptr;
        if (cudaGetSymbolAddress((void**)&p, g_noalloc_scratch) != cudaSuccess || !p) { (void)cudaGetLastError(); return GZB_E_CUDA; }
        uint32_t tab[CRC_TAB_WORDS];
        na_tables(tab);
        if (cudaMemcpy(p + NA_TAB, tab, sizeof(tab), cudaMemcpyHostToDevice) != cudaSuccess) { (void)cudaGetLastError(); return GZB_E_CUDA; }
        g_na_ptr[dev] = p;
    }
    *ptr = g_na_ptr[dev];
    return GZB_OK;
}

}  // namespace
}  // namespace gzb

extern "C" int gzb_crc32_device_noalloc(uint32_t crc, const uint8_t* d_in, uint64_t n, uint32_t* d_crc_out, void* stream)
{
    using namespace gzb;
    if (!d_crc_out || (n && !d_in)) return GZB_E_ARG;
    if (n > GZB_NOALLOC_MAX_BYTES) return GZB_E_TOO_LARGE;                 // the static scratch holds 65536 chunk values
    if (((uintptr_t)d_in & 15u) || ((uintptr_t)d_crc_out & 3u)) return GZB_E_ARG;
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess) { (void)cudaGetLastError(); return GZB_E_NO_DEVICE; }
    if (dev < 0 || dev >= NA_MAX_DEVICES) return GZB_E_ARG;
    uint32_t* s = nullptr;
    const int rc = na_setup(dev, &s);
    if (rc != GZB_OK) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    const uint32_t nchunks = n ? (uint32_t)((n + GZB_CHUNK - 1) / GZB_CHUNK) : 1u;      // n = 0: one CTA that passes crc through
    if (cudaMemsetAsync(s + NA_COUNT, 0, sizeof(uint32_t), st) != cudaSuccess) { (void)cudaGetLastError(); return GZB_E_CUDA; }
    k_crc_noalloc<<<nchunks, NA_THREADS, 0, st>>>(d_in, n, crc, s, d_crc_out);
    if (cudaGetLastError() != cudaSuccess) return GZB_E_CUDA;
    return GZB_OK;
}
