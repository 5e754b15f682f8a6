// k_pack.cu -- bit packing: every deflate block is written at its final bit position.
//
// Replaces compress_block / send_bits / bi_flush / bi_windup / _tr_stored_block / copy_block
// (lib/zlib/trees.c:1064-1109, 211-223, 1173-1203, 863-881) and, because blocks land where the
// member needs them, the ordered concatenation done by write_thread (src/parallel.c:510-540).
// zlib appends codes one at a time to a 16-bit accumulator; here the bit length of every token is
// known from the code tables, a block-wide prefix sum turns lengths into bit offsets, and each thread
// ORs its token's <= 48 code bits into the (pre-zeroed) output words.
//
// One CTA per deflate block.
#include "gzb_common.cuh"

namespace gzb {

namespace {

constexpr int PACK_THREADS = 256;
constexpr uint32_t FULL = 0xFFFFFFFFu;

// OR `nbits` (<= 57) low bits of v into the bit stream at absolute bit position pos
__device__ __forceinline__ void or_bits(uint32_t* out32, uint64_t pos, uint64_t v, uint32_t nbits)
{
    if (nbits == 0u) return;
    const uint64_t w = pos >> 5;
    const uint32_t s = (uint32_t)pos & 31u;
    const uint64_t lo = v << s;
    const uint32_t hi = s ? (uint32_t)(v >> (64u - s)) : 0u;
    const uint32_t w0 = (uint32_t)lo, w1 = (uint32_t)(lo >> 32);
    if (w0) atomicOr(&out32[w], w0);
    if (w1) atomicOr(&out32[w + 1], w1);
    if (hi) atomicOr(&out32[w + 2], hi);
}

__global__ void __launch_bounds__(PACK_THREADS)
k_pack(const uint8_t* __restrict__ in, uint32_t chunk, const uint32_t* __restrict__ tokens,
       const BlockRec* __restrict__ blocks, const ChunkRec* __restrict__ chunks, uint32_t bpc, uint8_t* __restrict__ out)
{
    __shared__ uint16_t s_lcode[288];
    __shared__ uint16_t s_dcode[32];
    __shared__ uint8_t  s_llen[288];
    __shared__ uint8_t  s_dlen[32];
    __shared__ uint32_t s_warp[PACK_THREADS / 32];
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const uint32_t c = blockIdx.x / bpc, b = blockIdx.x % bpc;
    const ChunkRec* cr = &chunks[c];
    if (b >= cr->n_blocks) return;
    const BlockRec* B = &blocks[blockIdx.x];
    uint32_t* out32 = reinterpret_cast<uint32_t*>(out);
    const uint64_t bit0 = B->out_bit;
    const uint32_t type = B->type, bfinal = B->bfinal;

    // chunk-boundary alignment: an empty stored block (0 00, pad, 00 00 FF FF) -- cf. Z_SYNC_FLUSH, parallel.c:630-633
    if (tid == 0 && b == cr->n_blocks - 1u && cr->tail_kind) {
        uint64_t p = (cr->tail_bit + 3ull + 7ull) & ~7ull;
        or_bits(out32, p, 0xFFFF0000ull, 32);
    }

    if (type == 0u) {
        // ---- stored: sub-blocks of <= 65535 bytes: BFINAL/00, pad, LEN, ~LEN, raw bytes ----
        const uint8_t* src = in + (size_t)c * chunk + B->in_off;
        const uint32_t n = B->in_bytes;
        const uint32_t nsub = n ? (n + 65534u) / 65535u : 1u;
        uint64_t bit = bit0;
        for (uint32_t sb = 0; sb < nsub; sb++) {
            const uint32_t off = sb * 65535u;
            const uint32_t len = min(65535u, n - off);
            if (tid == 0 && bfinal && sb == nsub - 1u) or_bits(out32, bit, 1ull, 1);
            const uint64_t hdr = (bit + 3ull + 7ull) & ~7ull;         // byte-aligned LEN field
            if (tid == 0) or_bits(out32, hdr, (uint64_t)len | ((uint64_t)(len ^ 0xFFFFu) << 16), 32);
            const uint64_t dbyte = (hdr >> 3) + 4ull;                 // first data byte
            // every thread fills whole destination words; the ragged ends are ORed byte-wise
            const uint64_t dend = dbyte + len;
            const uint64_t wfirst = (dbyte + 3ull) >> 2, wlast = dend >> 2;   // full words [wfirst, wlast)
            for (uint64_t w = wfirst + tid; w < wlast; w += PACK_THREADS) {
                const uint8_t* s = src + off + (w * 4ull - dbyte);
                uint32_t v = (uint32_t)s[0] | ((uint32_t)s[1] << 8) | ((uint32_t)s[2] << 16) | ((uint32_t)s[3] << 24);
                out32[w] = v;
            }
            if (tid < 8u) {
                // head bytes [dbyte, wfirst*4) and tail bytes [max(wlast*4, dbyte), dend)
                const uint64_t wl4 = wlast * 4ull, wf4 = wfirst * 4ull;
                uint64_t pos = (tid < 4u) ? dbyte + tid : (wl4 > dbyte ? wl4 : dbyte) + (tid - 4u);
                bool ok = (tid < 4u) ? (pos < (wf4 < dend ? wf4 : dend)) : (pos < dend && wlast >= wfirst);
                if (ok) atomicOr(&out32[pos >> 2], (uint32_t)src[off + (pos - dbyte)] << ((pos & 3ull) * 8ull));
            }
            bit = (dbyte + len) * 8ull;
        }
        return;
    }

    // ---- static / dynamic ----
    for (uint32_t i = tid; i < 288u; i += PACK_THREADS) { s_lcode[i] = B->lcode[i]; s_llen[i] = B->llen[i]; }
    if (tid < 32u) { s_dcode[tid] = B->dcode[tid]; s_dlen[tid] = B->dlen[tid]; }
    if (tid == 0) or_bits(out32, bit0, (uint64_t)(bfinal | (type << 1)), 3);
    uint64_t base = bit0 + 3ull;
    if (type == 2u) {
        const uint32_t hb = B->header_bits;
        for (uint32_t w = tid; w * 32u < hb; w += PACK_THREADS) {
            uint32_t nb = min(32u, hb - w * 32u);
            const uint8_t* h = &B->header[w * 4u];
            uint32_t v = (uint32_t)h[0] | ((uint32_t)h[1] << 8) | ((uint32_t)h[2] << 16) | ((uint32_t)h[3] << 24);
            if (nb < 32u) v &= (1u << nb) - 1u;
            or_bits(out32, base + (uint64_t)w * 32ull, v, nb);
        }
        base += hb;
    }
    __syncthreads();

    const uint32_t* tok = tokens + (size_t)c * chunk + B->tok_begin;
    const uint32_t nt = B->tok_count;
    for (uint32_t i0 = 0; i0 < nt; i0 += PACK_THREADS) {
        const uint32_t i = i0 + tid;
        uint64_t v = 0; uint32_t nb = 0;
        if (i < nt) {
            const uint32_t t = tok[i];
            if (!(t & 0x80000000u)) { v = s_lcode[t & 0xFFu]; nb = s_llen[t & 0xFFu]; }
            else {
                const uint32_t len = ((t >> 16) & 0xFFu) + 3u, dist = (t & 0xFFFFu) + 1u;
                const uint32_t lc = len_code(len), dc = dist_code(dist);
                v = s_lcode[257u + lc]; nb = s_llen[257u + lc];
                v |= (uint64_t)(len - len_base(lc)) << nb; nb += len_extra_bits(lc);
                v |= (uint64_t)s_dcode[dc] << nb; nb += s_dlen[dc];
                v |= (uint64_t)(dist - dist_base(dc)) << nb; nb += dist_extra_bits(dc);
            }
        }
        // block-wide exclusive scan of nb
        uint32_t incl = nb;
        for (int o = 1; o < 32; o <<= 1) { uint32_t x = __shfl_up_sync(FULL, incl, o); if (lane >= (uint32_t)o) incl += x; }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        uint32_t woff = 0, total = 0;
#pragma unroll
        for (int w = 0; w < PACK_THREADS / 32; w++) { uint32_t x = s_warp[w]; if ((uint32_t)w < warp) woff += x; total += x; }
        __syncthreads();
        or_bits(out32, base + woff + (incl - nb), v, nb);
        base += total;
    }
    if (tid == 0) or_bits(out32, base, s_lcode[256], s_llen[256]);      // end of block
}

}  // namespace

void launch_pack(const uint8_t* in, uint32_t chunk, const uint32_t* tokens, const BlockRec* blocks, const ChunkRec* chunks,
                 uint32_t nchunks, uint32_t bpc, uint8_t* out, cudaStream_t st)
{
    if (nchunks == 0) return;
    k_pack<<<nchunks * bpc, PACK_THREADS, 0, st>>>(in, chunk, tokens, blocks, chunks, bpc, out);
}

}  // namespace gzb
