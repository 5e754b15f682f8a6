// k_crc32.cu -- CRC-32 of every chunk (one CTA per chunk) + the GF(2) machinery to combine them.
//
// Replaces crc32z -> zlib crc32 (src/parallel.c:476-486; lib/zlib/crc32.c:202-297, slice-by-4
// tables in crc32.h) for the per-chunk checksum computed by compress_thread (parallel.c:677-688),
// and crc32_comb's 32x32 matrix squaring (parallel.c:393-474) for folding them: here a shift by
// len bytes is one multiplication by x^(8 len) mod P, taken from a table of x^(2^k).
//
// Each of the 256 threads of a CTA runs slice-by-4 over its own contiguous slice of the chunk
// (tables in shared memory), then the slices are folded pairwise in a tree:
//   crc(A||B) = crc(A) * x^(8 len B)  xor  crc(B)       (mod P, reflected bit order)
// Thread 0 starts from the 0xFFFFFFFF preset so the root is the standard CRC before the final xor.
#include "gzb_common.cuh"

namespace gzb {

// Tables live in global memory (built once per context by the host, gzb_api.cu):
//   tab[0..1023]    slice-by-4 tables (same construction as make_crc_table, lib/zlib/crc32.c:89-135)
//   tab[1024..1087] x^(2^k) mod P, k = 0..63

namespace {

constexpr int CRC_THREADS = 256;

__global__ void __launch_bounds__(CRC_THREADS)
k_crc32(const uint8_t* __restrict__ in, uint64_t n_bytes, uint32_t chunk, ChunkRec* __restrict__ chunks, const uint32_t* __restrict__ g_tab)
{
    __shared__ uint32_t tab[4][256];
    __shared__ uint32_t x2n[64];
    __shared__ uint32_t s_crc[CRC_THREADS];
    __shared__ uint32_t s_len[CRC_THREADS];
    const uint32_t tid = threadIdx.x;
    for (uint32_t i = tid; i < 1024u; i += CRC_THREADS) (&tab[0][0])[i] = g_tab[i];
    if (tid < 64u) x2n[tid] = g_tab[1024u + tid];
    __syncthreads();

    const uint64_t lo = (uint64_t)blockIdx.x * chunk;
    const uint32_t nc = (uint32_t)min((uint64_t)chunk, n_bytes - lo);
    const uint32_t slice = chunk / CRC_THREADS;                 // chunk is a multiple of 4096
    const uint32_t a = min(tid * slice, nc), b = min(a + slice, nc);
    const uint8_t* p = in + lo;

    uint32_t crc = (tid == 0) ? 0xFFFFFFFFu : 0u;
    uint32_t i = a;
    for (; i + 16u <= b; i += 16u) {                            // a and the buffer are 16-byte aligned
        uint4 v = __ldg(reinterpret_cast<const uint4*>(p + i));
        uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int k = 0; k < 4; k++) {
            crc ^= w[k];
            crc = tab[3][crc & 0xFFu] ^ tab[2][(crc >> 8) & 0xFFu] ^ tab[1][(crc >> 16) & 0xFFu] ^ tab[0][crc >> 24];
        }
    }
    for (; i < b; i++) crc = tab[0][(crc ^ p[i]) & 0xFFu] ^ (crc >> 8);
    s_crc[tid] = crc; s_len[tid] = b - a;
    __syncthreads();
    for (uint32_t stride = 1; stride < CRC_THREADS; stride <<= 1) {
        if ((tid & (2u * stride - 1u)) == 0u) {
            uint32_t lenB = s_len[tid + stride];
            s_crc[tid] = crc_shift(x2n, s_crc[tid], lenB) ^ s_crc[tid + stride];
            s_len[tid] += lenB;
        }
        __syncthreads();
    }
    if (tid == 0) chunks[blockIdx.x].crc = s_crc[0] ^ 0xFFFFFFFFu;
}

}  // namespace

void launch_crc32(const uint8_t* in, uint64_t n_bytes, uint32_t chunk, uint32_t nchunks, ChunkRec* chunks, const uint32_t* d_tab, cudaStream_t st)
{
    if (nchunks == 0) return;
    k_crc32<<<nchunks, CRC_THREADS, 0, st>>>(in, n_bytes, chunk, chunks, d_tab);
}

}  // namespace gzb
