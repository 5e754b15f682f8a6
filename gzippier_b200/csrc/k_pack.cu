// k_pack.cu -- bit packing: every chunk's fragment is written at its final position, in whole 16-byte lines.
//
// Replaces compress_block / send_bits / bi_flush / bi_windup / _tr_stored_block / copy_block
// (lib/zlib/trees.c:1064-1109, 211-223, 1173-1203, 863-881) and, because fragments land where the
// member needs them, the ordered concatenation done by write_thread (src/parallel.c:510-540).
// zlib appends codes one at a time to a 16-bit accumulator.  Here one CTA packs one chunk (its blocks
// are one contiguous bit run that starts and ends on a byte):
//   * the tokens of a block are taken in tiles of 2048, loaded with coalesced reads into shared memory (transposed, so
//     that each thread then picks up its own 8 consecutive tokens without bank conflicts); every thread adds up
//     the code lengths of its tokens, a block-wide prefix sum (warp shuffles + one shared-memory step) turns the
//     sums into bit offsets, and the thread then appends its codes to a 64-bit accumulator and stores whole
//     32-bit words into a shared-memory staging ring -- only its first and last word, which it shares with its
//     neighbours, are merged with a shared-memory atomic;
//   * complete 16-byte lines leave the ring with one 128-bit coalesced store each; the only byte-wise stores are
//     the ragged first and last line of the chunk (its neighbours' bytes live in the same lines);
//   * a stored block's payload is copied straight from the input, again in aligned 16-byte lines.
// Nothing is ORed into global memory, so the output needs no zeroing.
#include "gzb_common.cuh"

namespace gzb {

namespace {

constexpr int PACK_THREADS = 256;
constexpr int TPT = 8;                               // tokens per thread and tile
constexpr int TILE_TOK = PACK_THREADS * TPT;
constexpr int TSTRIDE = PACK_THREADS + 1;            // row stride of the transposed token tile
constexpr uint32_t RING_WORDS = 4096;                // staging ring: >= one tile at 48 bits per token + a dynamic header + a line
constexpr uint32_t FULL = 0xFFFFFFFFu;
static_assert(TILE_TOK * 48 / 32 + HDR_BYTES / 4 + 8 <= (int)RING_WORDS, "staging ring too small");

struct PackSmem {
    uint32_t ring[RING_WORDS];                       // word w of the output lives in ring[w % RING_WORDS] until its line is flushed
    uint32_t tile[2][TPT * TSTRIDE];                 // token i of a tile at [(i % TPT) * TSTRIDE + i / TPT]; double-buffered (cp.async)
    uint32_t first[512];                             // a token's first code, ready to append: bits | count << 24.  [0, 256): the literal's
                                                     // code; [256, 512): match length - 3 -> length code + its extra bits (<= 20 bits)
    uint32_t dst[32];                                // code | length << 16 of every distance symbol
    uint32_t eob;                                    // code | length << 16 of the end-of-block symbol
    uint32_t warp_sum[PACK_THREADS / 32];
};

// OR nbits (<= 32) of v into the ring at absolute bit position pos (shared-memory atomics: the words may be shared)
__device__ __forceinline__ void ring_or(PackSmem& S, uint64_t pos, uint32_t v, uint32_t nbits)
{
    if (nbits == 0u) return;
    const uint32_t w = (uint32_t)(pos >> 5), s = (uint32_t)pos & 31u;
    const uint64_t x = (uint64_t)v << s;
    if ((uint32_t)x) atomicOr(&S.ring[w % RING_WORDS], (uint32_t)x);
    if ((uint32_t)(x >> 32)) atomicOr(&S.ring[(w + 1u) % RING_WORDS], (uint32_t)(x >> 32));
}

// The ring holds output bytes [lo, ...).  Write bytes [lo, hi) to global memory and clear them in the ring: whole
// 16-byte lines with one 128-bit store, the ragged ends byte-wise.  All threads call; ends with a barrier.
__device__ void flush_ring(PackSmem& S, uint8_t* __restrict__ out, uint64_t lo, uint64_t hi, uint32_t tid)
{
    const uint64_t l0 = (lo + 15ull) >> 4, l1 = hi >> 4;          // whole lines [l0, l1)
    if (l1 > l0) {
        for (uint64_t L = l0 + tid; L < l1; L += PACK_THREADS) {
            uint4* r = reinterpret_cast<uint4*>(&S.ring[(uint32_t)((L * 4ull) % RING_WORDS)]);
            *reinterpret_cast<uint4*>(out + L * 16ull) = *r;
            *r = make_uint4(0, 0, 0, 0);
        }
    }
    // ragged head [lo, min(l0 * 16, hi)) and ragged tail [max(l1 * 16, lo), hi): at most 15 bytes each
    const uint64_t head_end = l0 * 16ull < hi ? l0 * 16ull : hi;
    const uint64_t tail_beg = l1 > l0 ? l1 * 16ull : head_end;    // no whole line: the head loop has covered [lo, head_end)
    const uint8_t* rb = reinterpret_cast<const uint8_t*>(S.ring);
    if (tid < 16u) { const uint64_t b = lo + tid; if (b < head_end) out[b] = rb[(uint32_t)(b % (RING_WORDS * 4u))]; }
    else if (tid < 32u) { const uint64_t b = tail_beg + (tid - 16u); if (b < hi) out[b] = rb[(uint32_t)(b % (RING_WORDS * 4u))]; }
    __syncthreads();
    // clear the words of the ragged ends (a word that straddles hi is cleared only when the caller says everything below hi
    // is final, which is always the case here: hi is either a line boundary or the end of a byte-aligned run)
    if (tid < 8u) {
        const uint64_t w0 = (tid < 4u ? lo : tail_beg) >> 2, w = w0 + (tid & 3u);
        const uint64_t wend = ((tid < 4u ? head_end : hi) + 3ull) >> 2;
        if (w < wend) S.ring[(uint32_t)(w % RING_WORDS)] = 0;
    }
    __syncthreads();
}

// payload of a stored block: out[dst, dst + len) = src[0, len), aligned 16-byte lines of the destination.  A line is
// assembled from aligned 32-bit words of the source (five when the two are not congruent mod 4), so the lines stop four
// bytes short of the payload's end: no word is read that does not hold a payload byte followed by payload.
__device__ void copy_stored(uint8_t* __restrict__ out, uint64_t dst, const uint8_t* __restrict__ src, uint32_t len, uint32_t tid)
{
    const uint64_t end = dst + len;
    const uint64_t l0 = (dst + 15ull) >> 4, l1 = len >= 4u ? (end - 4ull) >> 4 : 0ull;
    for (uint64_t L = l0 + tid; L < l1; L += PACK_THREADS) {
        const uint8_t* s = src + (L * 16ull - dst);
        const uint32_t sh = (uint32_t)((uintptr_t)s & 3u) * 8u;
        const uint32_t* sw = reinterpret_cast<const uint32_t*>((uintptr_t)s & ~(uintptr_t)3);
        const uint32_t w0 = __ldg(sw), w1 = __ldg(sw + 1), w2 = __ldg(sw + 2), w3 = __ldg(sw + 3);
        uint4 v;
        if (sh == 0u) v = make_uint4(w0, w1, w2, w3);
        else {
            const uint32_t w4 = __ldg(sw + 4);
            v = make_uint4(__funnelshift_r(w0, w1, sh), __funnelshift_r(w1, w2, sh), __funnelshift_r(w2, w3, sh), __funnelshift_r(w3, w4, sh));
        }
        *reinterpret_cast<uint4*>(out + L * 16ull) = v;
    }
    // ragged head [dst, head_end) (< 16 bytes) and tail [tail_beg, end) (< 36 bytes), byte-wise
    const uint64_t head_end = l0 * 16ull < end ? l0 * 16ull : end;
    const uint64_t tail_beg = l1 > l0 ? l1 * 16ull : head_end;
    if (tid < 16u) { const uint64_t b = dst + tid; if (b < head_end) out[b] = src[b - dst]; }
    else if (tid < 64u) { const uint64_t b = tail_beg + (tid - 16u); if (b < end) out[b] = src[b - dst]; }
}

__global__ void __launch_bounds__(PACK_THREADS, 4)
k_pack(const uint8_t* __restrict__ in, uint32_t chunk, const uint32_t* __restrict__ tokens,
       const BlockRec* __restrict__ blocks, const ChunkRec* __restrict__ chunks, uint32_t bpc, uint8_t* __restrict__ out)
{
    __shared__ __align__(16) PackSmem S;
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const uint32_t c = blockIdx.x;
    const ChunkRec* cr = &chunks[c];
    const uint32_t nblk = cr->n_blocks;
    for (uint32_t i = tid; i < RING_WORDS; i += PACK_THREADS) S.ring[i] = 0;
    uint64_t lo = cr->out_byte;                       // first byte the ring is responsible for
    uint64_t cur = lo * 8ull;                         // bit cursor (the same in every thread)
    __syncthreads();

    for (uint32_t b = 0; b < nblk; b++) {
        const BlockRec* B = &blocks[(size_t)c * bpc + b];
        const uint32_t type = B->type, bfinal = B->bfinal;
        cur = B->out_bit;                             // == cur by construction (k_layout)

        if (type == 0u) {
            // ---- stored: sub-blocks of <= 65535 bytes: BFINAL/00, pad, LEN, ~LEN, raw bytes ----
            const uint8_t* src = in + (size_t)c * chunk + B->in_off;
            const uint32_t n = B->in_bytes;
            const uint32_t nsub = n ? (n + 65534u) / 65535u : 1u;
            for (uint32_t sb = 0; sb < nsub; sb++) {
                const uint32_t off = sb * 65535u;
                const uint32_t len = min(65535u, n - off);
                const uint64_t hdr = (cur + 3ull + 7ull) & ~7ull;             // byte-aligned LEN field
                if (tid == 0) {
                    if (bfinal && sb == nsub - 1u) ring_or(S, cur, 1u, 1u);
                    ring_or(S, hdr, len | ((len ^ 0xFFFFu) << 16), 32u);
                }
                __syncthreads();
                const uint64_t dbyte = (hdr >> 3) + 4ull;                     // first payload byte
                flush_ring(S, out, lo, dbyte, tid);
                copy_stored(out, dbyte, src + off, len, tid);
                lo = dbyte + len; cur = lo * 8ull;
            }
            continue;
        }

        // ---- static / dynamic ----
        {
            static_assert(PACK_THREADS == 256, "one literal and one match length per thread");
            S.first[tid] = (uint32_t)B->lcode[tid] | ((uint32_t)B->llen[tid] << 24);
            const uint32_t len = tid + 3u, lc = len_code(len), ll = B->llen[257u + lc];
            S.first[256u + tid] = ((uint32_t)B->lcode[257u + lc] | ((len - len_base(lc)) << ll)) | ((ll + len_extra_bits(lc)) << 24);
            if (tid < 32u) S.dst[tid] = (uint32_t)B->dcode[tid] | ((uint32_t)B->dlen[tid] << 16);
            if (tid == 0) S.eob = (uint32_t)B->lcode[256] | ((uint32_t)B->llen[256] << 16);
        }
        if (tid == 0) ring_or(S, cur, bfinal | (type << 1), 3u);
        cur += 3ull;
        if (type == 2u) {
            const uint32_t hb = B->header_bits;
            for (uint32_t w = tid; w * 32u < hb; w += PACK_THREADS) {
                const uint32_t nb = min(32u, hb - w * 32u);
                const uint8_t* h = &B->header[w * 4u];
                uint32_t v = (uint32_t)h[0] | ((uint32_t)h[1] << 8) | ((uint32_t)h[2] << 16) | ((uint32_t)h[3] << 24);
                if (nb < 32u) v &= (1u << nb) - 1u;
                ring_or(S, cur + (uint64_t)w * 32ull, v, nb);
            }
            cur += hb;
        }
        __syncthreads();

        const uint32_t* tok = tokens + (size_t)c * chunk + B->tok_begin;
        const uint32_t nt = B->tok_count;
        // the next tile's tokens travel (cp.async, coalesced, straight into the transposed layout) while this one is packed
        auto prefetch = [&](uint32_t i0, uint32_t buf) {
#pragma unroll
            for (int k = 0; k < TPT; k++) {
                const uint32_t i = tid + (uint32_t)k * PACK_THREADS;
                if (i0 + i < nt) {
                    const uint32_t dsts = (uint32_t)__cvta_generic_to_shared(&S.tile[buf][(i % TPT) * TSTRIDE + i / TPT]);
                    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" :: "r"(dsts), "l"(tok + i0 + i) : "memory");
                }
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
        };
        if (nt) prefetch(0u, 0u);
        uint32_t buf = 0;
        for (uint32_t i0 = 0; i0 < nt; i0 += TILE_TOK, buf ^= 1u) {
            if (i0 + TILE_TOK < nt) { prefetch(i0 + TILE_TOK, buf ^ 1u); asm volatile("cp.async.wait_group 1;" ::: "memory"); }
            else asm volatile("cp.async.wait_group 0;" ::: "memory");
            __syncthreads();
            const uint32_t k0 = i0 + tid * TPT;
            const uint32_t cnt = k0 < nt ? min((uint32_t)TPT, nt - k0) : 0u;
            uint32_t t[TPT], e1[TPT];                                   // the token; its first code (bits | count << 24)
            uint32_t bits = 0;
#pragma unroll
            for (int j = 0; j < TPT; j++) {
                t[j] = S.tile[buf][j * TSTRIDE + tid];
                const bool m = (t[j] & 0x80000000u) != 0u;
                e1[j] = S.first[m ? 256u + ((t[j] >> 16) & 0xFFu) : (t[j] & 0xFFu)];
                if ((uint32_t)j < cnt) {
                    bits += e1[j] >> 24;
                    if (m) { const uint32_t dc = dist_code((t[j] & 0xFFFFu) + 1u); bits += (S.dst[dc] >> 16) + dist_extra_bits(dc); }
                }
            }
            // block-wide exclusive prefix sum of `bits`
            uint32_t incl = bits;
            for (int o = 1; o < 32; o <<= 1) { const uint32_t x = __shfl_up_sync(FULL, incl, o); if (lane >= (uint32_t)o) incl += x; }
            if (lane == 31) S.warp_sum[warp] = incl;
            __syncthreads();
            uint32_t woff = 0, total = 0;
#pragma unroll
            for (int w = 0; w < PACK_THREADS / 32; w++) { const uint32_t x = S.warp_sum[w]; if ((uint32_t)w < warp) woff += x; total += x; }
            // append the codes: whole words are stored, the first and the last word of the thread's run are merged
            if (cnt) {
                const uint64_t pos = cur + woff + (incl - bits);
                uint32_t w = (uint32_t)(pos >> 5);
                uint32_t nacc = (uint32_t)pos & 31u;                 // the low nacc bits of the first word belong to the thread before
                unsigned long long acc = 0;
                bool first = true;
                auto put = [&](uint32_t v, uint32_t nb) {            // nb <= 28: nacc + nb < 64
                    acc |= (unsigned long long)v << nacc; nacc += nb;
                    if (nacc >= 32u) {
                        if (first) { atomicOr(&S.ring[w % RING_WORDS], (uint32_t)acc); first = false; }
                        else S.ring[w % RING_WORDS] = (uint32_t)acc;
                        acc >>= 32; nacc -= 32u; w++;
                    }
                };
#pragma unroll
                for (int j = 0; j < TPT; j++) {
                    if ((uint32_t)j < cnt) {
                        put(e1[j] & 0xFFFFFFu, e1[j] >> 24);                                                          // <= 15 + 5 bits
                        if (t[j] & 0x80000000u) {
                            const uint32_t dist = (t[j] & 0xFFFFu) + 1u, dc = dist_code(dist), de = S.dst[dc], dl = de >> 16;
                            put((de & 0xFFFFu) | ((dist - dist_base(dc)) << dl), dl + dist_extra_bits(dc));           // <= 15 + 13 bits
                        }
                    }
                }
                if (nacc) atomicOr(&S.ring[w % RING_WORDS], (uint32_t)acc);
            }
            cur += total;
            __syncthreads();
            flush_ring(S, out, lo, (cur >> 7) << 4, tid);              // whole lines below the cursor
            if (((cur >> 7) << 4) > lo) lo = (cur >> 7) << 4;
        }
        if (tid == 0) ring_or(S, cur, S.eob & 0xFFFFu, S.eob >> 16);      // end of block
        cur += S.eob >> 16;
        __syncthreads();
    }
    // chunk-boundary alignment: an empty stored block (0 00, pad, 00 00 FF FF) -- cf. Z_SYNC_FLUSH, parallel.c:630-633
    if (cr->tail_kind) {
        const uint64_t p = (cr->tail_bit + 3ull + 7ull) & ~7ull;
        if (tid == 0) ring_or(S, p, 0xFFFF0000u, 32u);
        __syncthreads();
    }
    flush_ring(S, out, lo, cr->out_byte + cr->frag_bytes, tid);         // the rest, up to the fragment's last byte
}

}  // namespace

void launch_pack(const uint8_t* in, uint32_t chunk, const uint32_t* tokens, const BlockRec* blocks, const ChunkRec* chunks,
                 uint32_t nchunks, uint32_t bpc, uint8_t* out, cudaStream_t st)
{
    if (nchunks == 0) return;
    k_pack<<<nchunks, PACK_THREADS, 0, st>>>(in, chunk, tokens, blocks, chunks, bpc, out);
}

}  // namespace gzb
