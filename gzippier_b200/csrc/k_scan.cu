// k_scan.cu -- first pass over every chunk (one CTA per chunk): CRC-32, and the decision "this chunk is incompressible".
//
// CRC-32 replaces crc32z -> zlib crc32 (src/parallel.c:476-486; lib/zlib/crc32.c:202-297, slice-by-4
// tables in crc32.h) for the per-chunk checksum computed by compress_thread (parallel.c:677-688),
// and crc32_comb's 32x32 matrix squaring (parallel.c:393-474) for folding them: here a shift by
// len bytes is one multiplication by x^(8 len) mod P, taken from a table of x^(2^k).
// Each of the 256 threads runs slice-by-4 over its own contiguous slice of the chunk (tables in shared memory),
// then the slices are folded pairwise in a tree:
//   crc(A||B) = crc(A) * x^(8 len B)  xor  crc(B)       (mod P, reflected bit order)
// Thread 0 starts from the 0xFFFFFFFF preset so the root is the standard CRC before the final xor.
//
// The incompressibility test stands where zlib decides "stored" AFTER it has searched the block
// (_tr_flush_block, lib/zlib/trees.c:945-993: stored_len + 4 <= opt_lenb): the bytes pass through this kernel anyway, so
// the decision is taken BEFORE the match finder, which then does not run at all for such a chunk (k_lz77 skips it,
// k_huffman has nothing to plan, k_pack copies it as stored blocks).  A chunk is marked stored iff
//   (1) the order-0 entropy of a 1/16 sample of its bytes (one 16-byte word per 256 bytes) is at least 8 * 127/128
//       bits per byte, in Q8 fixed point (c_log2_q8): no Huffman code can save more than about 0.8 %, and
//   (2) its 8-byte grams do not repeat: of the grams at content-defined positions (about one in 64: those whose first
//       byte ends in the bits 101010) in the chunk, fewer than 1 in 64 equal a gram seen before in the chunk or its 32 KiB
//       dictionary -- otherwise the data is flat but repetitive (a compressed file stored twice) and LZ77 would win.
// Both are sums and set sizes: they do not depend on the order in which the threads get there, and
// oracle/engine_model.c (gzm_chunk_is_stored) computes the same integers sequentially.
#include "gzb_common.cuh"

namespace gzb {

// Tables live in global memory (built once per context by the host, gzb_api.cu):
//   tab[0..1023]    slice-by-4 tables (same construction as make_crc_table, lib/zlib/crc32.c:89-135)
//   tab[1024..1087] x^(2^k) mod P, k = 0..63

namespace {

constexpr int SCAN_THREADS = 256;
constexpr uint32_t GQ_CAP = 160;         // gram_scan's queue per warp: 31 left over + at most 4 * 32 new ones per step

__device__ __forceinline__ uint32_t lg_q8(uint32_t x)            // 256 * log2(x), x >= 1
{
    const uint32_t k = 31u - (uint32_t)__clz(x);
    const uint32_t m = k >= 8u ? (x >> (k - 8u)) & 0xFFu : (x << (8u - k)) & 0xFFu;
    return (k << 8) + c_log2_q8[m];
}

struct ScanSmem {
    uint32_t tab[4][256];
    uint32_t x2n[64];
    uint32_t s_crc[SCAN_THREADS];
    uint32_t s_len[SCAN_THREADS];
    uint32_t hist[SCAN_THREADS / 32][256];       // one byte histogram per warp
    uint32_t grams[SCAN_SLOTS];                  // fingerprints of the sampled 8-byte grams (0 = empty)
    uint32_t gq[SCAN_THREADS / 32][GQ_CAP];      // per warp: sampled positions waiting to be fingerprinted (gram_scan)
    uint32_t gqn[SCAN_THREADS / 32];
    unsigned long long bits_q8;                  // sum over byte values of count * (lg(n) - lg(count))
    uint32_t n_distinct, n_chunk_grams, n_chunk_new, overflow, verdict;
};

// insert one fingerprint (open addressing); returns 1 if it was not there yet.  Once more than SCAN_MAX_GRAMS distinct
// grams have turned up the chunk is not random-like: the verdict is "not stored" whatever else happens, and nothing more
// goes into the table (which therefore never fills up).
__device__ __forceinline__ uint32_t gram_insert(ScanSmem& S, uint32_t fp)
{
    if (*(volatile uint32_t*)&S.overflow) return 0u;
    uint32_t slot = (fp >> 8) & (SCAN_SLOTS - 1u);
    for (;;) {
        const uint32_t old = atomicCAS(&S.grams[slot], 0u, fp);
        if (old == fp) return 0u;
        if (old == 0u) {
            if (atomicAdd(&S.n_distinct, 1u) >= SCAN_MAX_GRAMS) S.overflow = 1u;
            return 1u;
        }
        slot = (slot + 1u) & (SCAN_SLOTS - 1u);
    }
}

// Grams of the stream positions [p0, p1) (multiples of 4).  The first bytes of four consecutive positions are the bytes
// of one aligned word, so one masked zero-byte test says whether any of the four is sampled (6 % of the words) -- but in
// 86 % of the warp steps SOME lane has such a word, so handling it on the spot would run the long path (two more loads,
// fingerprint, table insert) in nearly every step with one or two lanes active.  Instead the lanes only append the
// sampled positions to a queue of their warp, and whenever 32 are waiting the warp handles them with all lanes busy.
// This is a second pass over bytes that have partly left the caches: four independent loads are in flight per thread.
// (Measured on B200: 16-byte loads with four tests per load are slower.)  The result is a set and two counts: the order
// in which positions reach the table does not matter.

__device__ __forceinline__ void gram_take(ScanSmem& S, const uint8_t* sp, uint32_t p, uint32_t& ngram, uint32_t& nnew)
{
    const uint32_t* w = reinterpret_cast<const uint32_t*>(sp + (p & ~3u));
    const uint32_t sh = (p & 3u) * 8u;
    const uint32_t w0 = __ldg(w), w1 = __ldg(w + 1), w2 = __ldg(w + 2);
    const uint32_t a = sh ? __funnelshift_r(w0, w1, sh) : w0, b = sh ? __funnelshift_r(w1, w2, sh) : w1;
    ngram++; nnew += gram_insert(S, scan_fingerprint(a, b));
}

__device__ void gram_scan(ScanSmem& S, const uint8_t* sp, uint32_t p0, uint32_t p1, bool count, uint32_t tid)
{
    uint32_t ngram = 0, nnew = 0;
    const uint32_t lane = tid & 31u, warp = tid >> 5;
    const uint32_t* w = reinterpret_cast<const uint32_t*>(sp);
    const uint32_t qe = p1 / 4u;
    uint32_t* Q = S.gq[warp];
    uint32_t* qn = &S.gqn[warp];
    if (lane == 0) *qn = 0;
    __syncwarp();
    auto test = [&](uint32_t q, uint32_t w0, bool live) {        // whole warp; live: this lane's word exists
        const uint32_t m = (w0 & 0x3F3F3F3Fu) ^ 0x2A2A2A2Au;     // a zero byte <=> that position is sampled
        if (live && ((m - 0x01010101u) & ~m & 0x80808080u) != 0u) {
            uint32_t c = 0, pos[4];
#pragma unroll
            for (uint32_t j = 0; j < 4u; j++) if (((m >> (8u * j)) & 0xFFu) == 0u) pos[c++] = 4u * q + j;
            const uint32_t at = atomicAdd(qn, c);
#pragma unroll
            for (uint32_t j = 0; j < 4u; j++) if (j < c) Q[at + j] = pos[j];
        }
        __syncwarp();
        uint32_t n = *(volatile uint32_t*)qn;
        while (n >= 32u) {                                       // warp-uniform
            n -= 32u;
            gram_take(S, sp, Q[n + lane], ngram, nnew);
        }
        __syncwarp();
        if (lane == 0) *qn = n;
        __syncwarp();
    };
    uint32_t q = p0 / 4u + tid;
    const uint32_t q_warp = q - lane;                            // the loops' conditions are the same for a whole warp
    uint32_t base = q_warp;
    for (; base + 3u * SCAN_THREADS + 32u <= qe; base += 4u * SCAN_THREADS, q += 4u * SCAN_THREADS) {
        const uint32_t a0 = __ldg(w + q), a1 = __ldg(w + q + SCAN_THREADS), a2 = __ldg(w + q + 2u * SCAN_THREADS), a3 = __ldg(w + q + 3u * SCAN_THREADS);
        test(q, a0, true); test(q + SCAN_THREADS, a1, true); test(q + 2u * SCAN_THREADS, a2, true); test(q + 3u * SCAN_THREADS, a3, true);
    }
    for (; base < qe; base += SCAN_THREADS, q += SCAN_THREADS) {
        const bool live = q < qe;
        test(q, live ? __ldg(w + q) : 0u, live);
    }
    const uint32_t n = *(volatile uint32_t*)qn;                  // what is left: fewer than 32
    if (lane < n) gram_take(S, sp, Q[lane], ngram, nnew);
    if (count) { if (ngram) atomicAdd(&S.n_chunk_grams, ngram); if (nnew) atomicAdd(&S.n_chunk_new, nnew); }
}

template <bool SCAN>
__global__ void __launch_bounds__(SCAN_THREADS)
k_scan(const uint8_t* __restrict__ in, uint64_t n_bytes, uint32_t halo, uint32_t chunk, uint32_t dict, uint32_t bpc,
       ChunkRec* __restrict__ chunks, BlockRec* __restrict__ blocks, const uint32_t* __restrict__ g_tab)
{
    __shared__ __align__(16) ScanSmem S;
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    for (uint32_t i = tid; i < 1024u; i += SCAN_THREADS) (&S.tab[0][0])[i] = g_tab[i];
    if (tid < 64u) S.x2n[tid] = g_tab[1024u + tid];
    if (SCAN) {
        for (uint32_t i = tid; i < (SCAN_THREADS / 32u) * 256u; i += SCAN_THREADS) (&S.hist[0][0])[i] = 0;
        for (uint32_t i = tid; i < SCAN_SLOTS; i += SCAN_THREADS) S.grams[i] = 0;
        if (tid == 0) { S.bits_q8 = 0; S.n_distinct = 0; S.n_chunk_grams = 0; S.n_chunk_new = 0; S.overflow = 0; S.verdict = 0; }
    }
    __syncthreads();

    const uint64_t lo = (uint64_t)blockIdx.x * chunk;
    const uint32_t nc = (uint32_t)min((uint64_t)chunk, n_bytes - lo);
    const uint32_t slice = chunk / SCAN_THREADS;                // chunk is a multiple of 4096
    const uint32_t a = min(tid * slice, nc), b = min(a + slice, nc);
    const uint8_t* p = in + lo;

    uint32_t crc = (tid == 0) ? 0xFFFFFFFFu : 0u;
    uint32_t i = a;
    for (; i + 16u <= b; i += 16u) {                            // a and the buffer are 16-byte aligned
        uint4 v = __ldg(reinterpret_cast<const uint4*>(p + i));
        uint32_t w[4] = {v.x, v.y, v.z, v.w};
        if (SCAN && (i % (uint32_t)SCAN_SAMPLE) == 0u) {
#pragma unroll
            for (int k = 0; k < 4; k++) {
                atomicAdd(&S.hist[warp][w[k] & 0xFFu], 1u); atomicAdd(&S.hist[warp][(w[k] >> 8) & 0xFFu], 1u);
                atomicAdd(&S.hist[warp][(w[k] >> 16) & 0xFFu], 1u); atomicAdd(&S.hist[warp][w[k] >> 24], 1u);
            }
        }
#pragma unroll
        for (int k = 0; k < 4; k++) {
            crc ^= w[k];
            crc = S.tab[3][crc & 0xFFu] ^ S.tab[2][(crc >> 8) & 0xFFu] ^ S.tab[1][(crc >> 16) & 0xFFu] ^ S.tab[0][crc >> 24];
        }
    }
    for (; i < b; i++) crc = S.tab[0][(crc ^ p[i]) & 0xFFu] ^ (crc >> 8);
    S.s_crc[tid] = crc; S.s_len[tid] = b - a;
    __syncthreads();
    for (uint32_t stride = 1; stride < SCAN_THREADS; stride <<= 1) {
        if ((tid & (2u * stride - 1u)) == 0u) {
            uint32_t lenB = S.s_len[tid + stride];
            S.s_crc[tid] = crc_shift(S.x2n, S.s_crc[tid], lenB) ^ S.s_crc[tid + stride];
            S.s_len[tid] += lenB;
        }
        __syncthreads();
    }
    ChunkRec* cr = &chunks[blockIdx.x];
    if (tid == 0) { cr->crc = S.s_crc[0] ^ 0xFFFFFFFFu; cr->stored = 0; }
    if (!SCAN || nc < (uint32_t)SCAN_MIN) return;

    // (1) order-0 entropy of the sampled bytes, Q8: thread = byte value
    {
        uint32_t cnt = 0;
#pragma unroll
        for (int w = 0; w < SCAN_THREADS / 32; w++) cnt += S.hist[w][tid];
        const uint32_t ns = (nc / (uint32_t)SCAN_SAMPLE + ((nc % (uint32_t)SCAN_SAMPLE) >= 16u ? 1u : 0u)) * 16u;   // sampled bytes
        unsigned long long term = cnt ? (unsigned long long)cnt * (lg_q8(ns) - lg_q8(cnt)) : 0ull;
        for (int o = 16; o > 0; o >>= 1) term += __shfl_xor_sync(0xFFFFFFFFu, term, o);
        if (lane == 0 && term) atomicAdd(&S.bits_q8, term);
        __syncthreads();
        if (S.bits_q8 * SCAN_MIN_BITS_D < (unsigned long long)ns * 8ull * 256ull * SCAN_MIN_BITS_N) return;        // a code would pay: normal path
    }
    // (2) repeated 8-byte grams: dictionary first (insert only), then the chunk (count)
    {
        const uint32_t D = dict ? (uint32_t)min((uint64_t)dict, lo + halo) : 0u;
        const uint8_t* sp = p - D;                              // 4-byte aligned: D is 0 or 32768
        const uint32_t P = ((D + nc) / 4u - 2u) * 4u;           // positions whose three words lie inside the stream
        gram_scan(S, sp, 0u, D < P ? D : P, false, tid);
        __syncthreads();
        gram_scan(S, sp, D, P, true, tid);
        __syncthreads();
        if (tid == 0) {
            const uint32_t g = S.n_chunk_grams, dups = g - S.n_chunk_new;
            S.verdict = (!S.overflow && g != 0u && dups * 64u < g) ? 1u : 0u;
        }
        __syncthreads();
        if (!S.verdict) return;
    }
    // stored: the chunk's records, as k_lz77 would have left them -- blocks are plain 64 KiB spans, no tokens
    const uint32_t nblocks = (nc + BLOCK_SPAN - 1) / BLOCK_SPAN;
    if (tid == 0) { cr->stored = 1; cr->n_bytes = nc; cr->n_blocks = nblocks; cr->n_tokens = 0; }
    if (tid < nblocks) {
        BlockRec* B = &blocks[(size_t)blockIdx.x * bpc + tid];
        B->type = 0; B->header_bits = 0; B->body_bits = 0; B->tok_begin = 0; B->tok_count = 0;
        B->in_off = tid * (uint32_t)BLOCK_SPAN; B->in_bytes = min((uint32_t)BLOCK_SPAN, nc - tid * (uint32_t)BLOCK_SPAN);
    }
}

static_assert(sizeof(ScanSmem) <= 48 * 1024, "k_scan: static shared memory");

}  // namespace

void launch_scan(const uint8_t* in, uint64_t n_bytes, uint32_t halo, uint32_t chunk, uint32_t dict, uint32_t nchunks, uint32_t bpc, int scan,
                 ChunkRec* chunks, BlockRec* blocks, const uint32_t* d_tab, cudaStream_t st)
{
    if (nchunks == 0) return;
    if (scan) k_scan<true><<<nchunks, SCAN_THREADS, 0, st>>>(in, n_bytes, halo, chunk, dict, bpc, chunks, blocks, d_tab);
    else k_scan<false><<<nchunks, SCAN_THREADS, 0, st>>>(in, n_bytes, halo, chunk, dict, bpc, chunks, blocks, d_tab);
}

}  // namespace gzb
