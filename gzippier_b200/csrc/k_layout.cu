// k_layout.cu -- where every deflate block goes, fragment sizes, chunk-boundary alignment, CRC fold.
//
// Replaces the ordering half of write_thread (src/parallel.c:510-551: emit jobs by seq, running
// crc32_comb and ulen), the byte-alignment dance of compress_thread (:624-644) and BFINAL on the
// last job (:645-649).  The reference serialises this on one thread; here block sizes are known
// exactly after k_huffman, so the output position of every block is a prefix sum and the per-chunk
// CRCs fold as  xor_i crc_i * x^(8 * bytes after chunk i)  (mod P), both computed in parallel.
//
// One CTA per sub-batch.  The running totals live in a device-side Cursor so consecutive
// sub-batches chain without a host round trip.
#include "gzb_common.cuh"

namespace gzb {

namespace {

constexpr int LAYOUT_THREADS = 1024;

__device__ __forceinline__ uint64_t align8(uint64_t b) { return (b + 7ull) & ~7ull; }

// size in bits of a stored block (possibly several sub-blocks of <= 65535 bytes) starting at bit `bit`
__device__ __forceinline__ uint64_t stored_end(uint64_t bit, uint32_t in_bytes)
{
    uint64_t nsub = in_bytes ? (in_bytes + 65534u) / 65535u : 1u;
    return align8(bit + 3) + 32ull * nsub + 8ull * (nsub - 1ull) + 8ull * in_bytes;
}

__global__ void __launch_bounds__(LAYOUT_THREADS)
k_layout(BlockRec* __restrict__ blocks, ChunkRec* __restrict__ chunks, uint32_t nchunks, uint32_t bpc, int final,
         Cursor* __restrict__ cur, const uint32_t* __restrict__ g_tab)
{
    __shared__ uint64_t s_warp[32];
    __shared__ uint64_t s_carry;
    __shared__ uint32_t s_x2n[64];
    __shared__ uint32_t s_crcw[32];
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    if (tid < 64u) s_x2n[tid] = g_tab[1024u + tid];
    if (tid == 0) s_carry = cur->out_bytes;
    __syncthreads();

    // total input bytes of the sub-batch (all chunks full except possibly the last)
    const uint64_t chunk_bytes = chunks[0].n_bytes;
    const uint64_t total_in = (uint64_t)(nchunks - 1u) * chunk_bytes + chunks[nchunks - 1u].n_bytes;
    uint32_t crc_acc = 0, nblk_acc = 0;

    for (uint32_t c0 = 0; c0 < nchunks; c0 += LAYOUT_THREADS) {
        const uint32_t c = c0 + tid;
        uint64_t frag = 0;
        if (c < nchunks) {
            ChunkRec* cr = &chunks[c];
            const bool last_chunk = final && (c == nchunks - 1u);
            uint64_t bit = 0;
            for (uint32_t b = 0; b < cr->n_blocks; b++) {
                BlockRec* B = &blocks[(size_t)c * bpc + b];
                B->out_bit = bit;                                   // relative for now
                B->bfinal = (last_chunk && b == cr->n_blocks - 1u) ? 1u : 0u;
                bit = (B->type == 0u) ? stored_end(bit, B->in_bytes) : bit + 3ull + B->body_bits;
            }
            uint32_t tail = 0;
            if (last_chunk) bit = align8(bit);
            else if (bit & 7ull) { tail = 1; cr->tail_bit = bit; bit = align8(bit + 3ull) + 32ull; }   // empty stored block
            cr->tail_kind = tail;
            frag = bit >> 3;
            cr->frag_bytes = frag;
            nblk_acc += cr->n_blocks;
            // CRC contribution: crc_c * x^(8 * bytes after this chunk inside the sub-batch)
            uint64_t after = total_in - ((uint64_t)c * chunk_bytes + cr->n_bytes);
            crc_acc ^= crc_shift(s_x2n, cr->crc, after);
        }
        // exclusive scan of frag over the 1024 chunks of this round
        uint64_t incl = frag;
        for (int o = 1; o < 32; o <<= 1) { uint64_t v = __shfl_up_sync(0xFFFFFFFFu, incl, o); if (lane >= (uint32_t)o) incl += v; }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            uint64_t w = s_warp[lane], wi = w;
            for (int o = 1; o < 32; o <<= 1) { uint64_t v = __shfl_up_sync(0xFFFFFFFFu, wi, o); if (lane >= (uint32_t)o) wi += v; }
            s_warp[lane] = wi - w;                                  // exclusive warp offsets
        }
        __syncthreads();
        const uint64_t base = s_carry;
        const uint64_t excl = base + s_warp[warp] + (incl - frag);
        if (c < nchunks) {
            ChunkRec* cr = &chunks[c];
            cr->out_byte = excl;
            for (uint32_t b = 0; b < cr->n_blocks; b++) blocks[(size_t)c * bpc + b].out_bit += excl * 8ull;
            if (cr->tail_kind) cr->tail_bit += excl * 8ull;
        }
        __syncthreads();
        if (tid == LAYOUT_THREADS - 1) s_carry = excl + frag;       // last thread holds the running total
        __syncthreads();
    }

    // fold the CRC contributions and update the cursor
    for (int o = 16; o > 0; o >>= 1) { crc_acc ^= __shfl_xor_sync(0xFFFFFFFFu, crc_acc, o); nblk_acc += __shfl_xor_sync(0xFFFFFFFFu, nblk_acc, o); }
    __shared__ uint32_t s_nb[32];
    if (lane == 0) { s_crcw[warp] = crc_acc; s_nb[warp] = nblk_acc; }
    __syncthreads();
    if (tid == 0) {
        uint32_t x = 0, nb = 0;
        for (int w = 0; w < LAYOUT_THREADS / 32; w++) { x ^= s_crcw[w]; nb += s_nb[w]; }
        cur->crc = crc_shift(s_x2n, cur->crc, total_in) ^ x;
        cur->in_bytes += total_in;
        cur->out_bytes = s_carry;
        cur->n_chunks += nchunks;
        cur->n_blocks += nb;
    }
}

}  // namespace

void launch_layout(BlockRec* blocks, ChunkRec* chunks, uint32_t nchunks, uint32_t bpc, int final, Cursor* cur, const uint32_t* d_tab, cudaStream_t st)
{
    if (nchunks == 0) return;
    k_layout<<<1, LAYOUT_THREADS, 0, st>>>(blocks, chunks, nchunks, bpc, final, cur, d_tab);
}

}  // namespace gzb
