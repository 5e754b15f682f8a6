// Microbenchmark + unit test of the pipelined hash-table insertion used by k_lz77 (round 2).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I include -I gzippier_b200/csrc -o /tmp/micro_insert \
//        tools/micro_insert.cu gzippier_b200/csrc/corpus.c && /tmp/micro_insert
// Per 2 KiB tile, one warp per table (6-byte chains, 3-byte and 4-byte single-entry tables) inserts the 2048 positions
// in stream order: 64 steps of 32 positions, lane = position.  The lanes of a step that share a bucket are found with
// MATCH.ANY; the step is then one predicated table load (lanes without a predecessor in the step) and one predicated
// table store (the last lane of every bucket) -- no read-back, so consecutive steps pipeline.
// Question: cycles per tile when the other warps of the CTA are idle / hammer shared memory, and with the three
// inserters on one scheduler or spread over three.  Results are compared with a sequential CPU insertion.
#include "gzb_common.cuh"
#include "gzb_corpus.h"
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

using namespace gzb;

constexpr int RING_MIRROR = 288;
struct Smem {
    uint32_t ring32[RING / 4 + RING_MIRROR / 4];
    uint16_t head[1 << HBITS], head3[1 << H3BITS], head4[1 << H4BITS];
    uint32_t junk[4096];
    uint32_t done;
};
constexpr uint32_t TABLE_FILL = 0x80008000u;

__device__ __forceinline__ uint32_t lds_u16(uint32_t a) { uint32_t v; asm volatile("ld.shared.u16 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
__device__ __forceinline__ uint32_t lds_u32(uint32_t a) { uint32_t v; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
__device__ __forceinline__ void sts_u16(uint32_t a, uint32_t v) { asm volatile("st.shared.u16 [%0], %1;" :: "r"(a), "h"((uint16_t)v) : "memory"); }

template <int KIND> __device__ __forceinline__ uint32_t to_delta(uint32_t p16, uint32_t raw)
{
    const uint32_t d = (p16 - raw) & 0xFFFFu;
    if (KIND == 0) return d - 1u < 32767u ? d : 0xFFFFu;
    return d - 1u < (KIND == 1 ? (uint32_t)TOO_FAR : 32767u) ? d : 0u;
}
__device__ __forceinline__ uint32_t sweep_pair(uint32_t w, uint32_t p16)
{
    const uint32_t mark = p16 ^ 0x8000u;
    uint32_t lo = w & 0xFFFFu, hi = w >> 16;
    if (((p16 - lo) & 0xFFFFu) > 32767u) lo = mark;
    if (((p16 - hi) & 0xFFFFu) > 32767u) hi = mark;
    return lo | (hi << 16);
}

// U = steps in flight
template <int KIND, int U>
__device__ void insert_tile(uint32_t sbase, uint32_t t, uint32_t lane, uint16_t* __restrict__ gout)
{
    const uint32_t ring_s = sbase + (uint32_t)offsetof(Smem, ring32);
    const uint32_t table_s = sbase + (uint32_t)(KIND == 0 ? offsetof(Smem, head) : KIND == 1 ? offsetof(Smem, head3) : offsetof(Smem, head4));
    const uint32_t base = t * TILE;
    // ageing sweep of one eighth of the table (see k_lz77.cu)
    {
        const uint32_t p16 = base & 0xFFFFu, slice = t & 7u;
        const uint32_t words = (KIND == 0 ? (1u << HBITS) : (1u << H3BITS)) / 2u / 8u;
        uint32_t* tab = reinterpret_cast<uint32_t*>(__cvta_shared_to_generic(table_s)) + slice * words;
        for (uint32_t i = lane * 4u; i < words; i += 128u) {
            uint4 v = *reinterpret_cast<uint4*>(tab + i);
            v.x = sweep_pair(v.x, p16); v.y = sweep_pair(v.y, p16); v.z = sweep_pair(v.z, p16); v.w = sweep_pair(v.w, p16);
            *reinterpret_cast<uint4*>(tab + i) = v;
        }
        __syncwarp();
    }
    const uint32_t lt_mask = (1u << lane) - 1u;
#pragma unroll 1
    for (uint32_t g0 = 0; g0 < (uint32_t)TILE / 32u; g0 += U) {
        uint32_t b[U], same[U];
#pragma unroll
        for (int u = 0; u < U; u++) {
            const uint32_t a = (base + (g0 + u) * 32u + lane) & 0xFFFFu;
            const uint32_t wa = ring_s + (a & 0xFFFCu), sh = (a & 3u) * 8u;
            const uint32_t w0 = lds_u32(wa), w1 = lds_u32(wa + 4u);
            const uint32_t v = __funnelshift_r(w0, w1, sh);
            if (KIND == 0) {
                const uint32_t w2 = lds_u32(wa + 8u);
                b[u] = hash6(v, __funnelshift_r(w1, w2, sh) & 0xFFFFu);
            } else b[u] = KIND == 1 ? hash3(v) : hash4(v);
            same[u] = __match_any_sync(0xFFFFFFFFu, b[u]);
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            const uint32_t p16 = (base + (g0 + u) * 32u + lane) & 0xFFFFu;
            const uint32_t ta = table_s + 2u * b[u];
            const uint32_t lower = same[u] & lt_mask;
            uint32_t raw = p16 - (lane - (uint32_t)(31 - __clz(lower)));        // the nearest lower lane of the bucket
            if (!lower) raw = lds_u16(ta);
            if ((same[u] >> lane) == 1u) sts_u16(ta, p16);                      // the last position of the step owns the bucket
            __syncwarp();
            gout[(size_t)base + (g0 + u) * 32u + lane] = (uint16_t)to_delta<KIND>(p16, raw);
        }
    }
}

template <int U>
__global__ void __launch_bounds__(1024, 1)
k(const uint8_t* __restrict__ in, uint32_t ntiles, uint16_t* o6, uint16_t* o3, uint16_t* o4, int busy, int spread, unsigned long long* cyc)
{
    extern __shared__ __align__(16) unsigned char raw[];
    Smem& S = *reinterpret_cast<Smem*>(raw);
    uint32_t sbase; asm volatile("{ .reg .u64 t; cvta.to.shared.u64 t, %1; cvt.u32.u64 %0, t; }" : "=r"(sbase) : "l"(raw));
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const size_t off = (size_t)blockIdx.x * ntiles * TILE;
    in += off; o6 += off; o3 += off; o4 += off;
    for (uint32_t i = tid; i < (1u << HBITS) / 2; i += 1024u) reinterpret_cast<uint32_t*>(S.head)[i] = TABLE_FILL;
    for (uint32_t i = tid; i < (1u << H3BITS) / 2; i += 1024u) { reinterpret_cast<uint32_t*>(S.head3)[i] = TABLE_FILL; reinterpret_cast<uint32_t*>(S.head4)[i] = TABLE_FILL; }
    for (uint32_t i = tid; i < 4096u; i += 1024u) S.junk[i] = i * 2654435761u;
    if (tid == 0) S.done = 0;
    int kind = -1;
    if (spread) { if (warp >= 1 && warp <= 3) kind = (int)warp - 1; }          // schedulers 1, 2, 3
    else if ((warp & 3u) == 3u && warp < 12u) kind = (int)(warp >> 2);         // all on scheduler 3
    unsigned long long acc = 0; uint32_t sink = 0;
    auto load = [&](uint32_t t) {
        if (t > ntiles) return;
        const uint32_t o = tid * 2u;                                           // 2 bytes per thread
        const uint32_t pos = t * TILE + o;
        uint16_t v = t < ntiles ? *reinterpret_cast<const uint16_t*>(in + pos) : (uint16_t)0;
        reinterpret_cast<uint16_t*>(S.ring32)[(pos & 0xFFFFu) >> 1] = v;
        if ((pos & 0xFFFFu) < (uint32_t)RING_MIRROR) reinterpret_cast<uint16_t*>(S.ring32)[(RING + (pos & 0xFFFFu)) >> 1] = v;
    };
    load(0);
    for (uint32_t t = 0; t < ntiles; t++) {
        load(t + 1);
        __syncthreads();
        if (kind >= 0) {
            const long long c0 = clock64();
            if (kind == 0) insert_tile<0, U>(sbase, t, lane, o6);
            else if (kind == 1) insert_tile<1, U>(sbase, t, lane, o3);
            else insert_tile<2, U>(sbase, t, lane, o4);
            acc += (unsigned long long)(clock64() - c0);
            __syncwarp();
            if (lane == 0) { __threadfence_block(); atomicAdd(&S.done, 1u); }
        } else if (busy) {
            // a stand-in for the search warps: dependent shared-memory loads + a little ALU
            uint32_t x = tid;
            while (*(volatile uint32_t*)&S.done < 3u * (t + 1u)) {
#pragma unroll
                for (int j = 0; j < 16; j++) x = S.junk[(x + (uint32_t)j) & 4095u] ^ (x >> 3);
            }
            sink ^= x;
        }
        __syncthreads();
    }
    if (kind >= 0 && lane == 0) atomicAdd(&cyc[kind], acc);
    if (sink == 0x12345u) cyc[3] = sink;
}

static void cpu_ref(const uint8_t* s, size_t n, std::vector<uint16_t>& o6, std::vector<uint16_t>& o3, std::vector<uint16_t>& o4)
{
    std::vector<int64_t> h6(1 << HBITS, -1), h3(1 << H3BITS, -1), h4(1 << H4BITS, -1);
    for (size_t p = 0; p < n; p++) {
        uint8_t tmp[8] = {0};
        for (int k = 0; k < 6 && p + k < n; k++) tmp[k] = s[p + k];
        uint32_t lo; memcpy(&lo, tmp, 4);
        const uint32_t hi = tmp[4] | (tmp[5] << 8);
        const uint32_t a = hash6(lo, hi), b = hash3(lo), c = hash4(lo);
        const int64_t d6 = h6[a] < 0 ? 0 : (int64_t)p - h6[a], d3 = h3[b] < 0 ? 0 : (int64_t)p - h3[b], d4 = h4[c] < 0 ? 0 : (int64_t)p - h4[c];
        o6[p] = (d6 >= 1 && d6 <= 32767) ? (uint16_t)d6 : 0xFFFF;
        o3[p] = (d3 >= 1 && d3 <= TOO_FAR) ? (uint16_t)d3 : 0;
        o4[p] = (d4 >= 1 && d4 <= 32767) ? (uint16_t)d4 : 0;
        h6[a] = h3[b] = h4[c] = (int64_t)p;
    }
}

template <int U>
static void run(const uint8_t* d_in, const std::vector<uint8_t>& host, uint32_t ntiles, int nblk, int busy, int spread)
{
    const size_t n = (size_t)nblk * ntiles * TILE;
    uint16_t *o6, *o3, *o4; unsigned long long* cyc;
    cudaMalloc(&o6, n * 2); cudaMalloc(&o3, n * 2); cudaMalloc(&o4, n * 2); cudaMalloc(&cyc, 32); cudaMemset(cyc, 0, 32);
    cudaFuncSetAttribute(k<U>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem));
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    cudaEventRecord(a);
    k<U><<<nblk, 1024, sizeof(Smem)>>>(d_in, ntiles, o6, o3, o4, busy, spread, cyc);
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    unsigned long long h[4]; cudaMemcpy(h, cyc, 32, cudaMemcpyDeviceToHost);
    // check block 0 and the last block
    size_t bad = 0;
    for (int blk : {0, nblk - 1}) {
        const size_t off = (size_t)blk * ntiles * TILE, len = (size_t)ntiles * TILE;
        std::vector<uint16_t> g6(len), g3(len), g4(len), r6(len), r3(len), r4(len);
        cudaMemcpy(g6.data(), o6 + off, len * 2, cudaMemcpyDeviceToHost); cudaMemcpy(g3.data(), o3 + off, len * 2, cudaMemcpyDeviceToHost);
        cudaMemcpy(g4.data(), o4 + off, len * 2, cudaMemcpyDeviceToHost);
        cpu_ref(host.data() + off, len, r6, r3, r4);
        for (size_t i = 0; i + 8 < len; i++) bad += (g6[i] != r6[i]) + (g3[i] != r3[i]) + (g4[i] != r4[i]);   // the last positions hash zeros on the CPU
    }
    const double per = 1.0 / ((double)nblk * ntiles);
    printf("U=%d busy=%d spread=%d: %.2f ms, cycles per tile: insert6 %.0f insert3 %.0f insert4 %.0f, mismatches %zu, err %d\n", U, busy, spread, ms,
           h[0] * per, h[1] * per, h[2] * per, bad, (int)cudaGetLastError());
    cudaFree(o6); cudaFree(o3); cudaFree(o4); cudaFree(cyc);
}

int main(int argc, char** argv)
{
    const int kind = argc > 1 ? atoi(argv[1]) : GZB_CORPUS_TEXT;
    const uint32_t ntiles = 256; const int nblk = 148;
    const size_t n = (size_t)nblk * ntiles * TILE;
    std::vector<uint8_t> host(n + 64);
    gzb_corpus_fill(kind, 1234, 0, host.data(), n, 16);
    uint8_t* d_in; cudaMalloc(&d_in, n + 64); cudaMemcpy(d_in, host.data(), n, cudaMemcpyHostToDevice);
    for (int busy = 0; busy < 2; busy++)
        for (int spread = 0; spread < 2; spread++) {
            run<1>(d_in, host, ntiles, nblk, busy, spread);
            run<2>(d_in, host, ntiles, nblk, busy, spread);
            run<4>(d_in, host, ntiles, nblk, busy, spread);
            run<8>(d_in, host, ntiles, nblk, busy, spread);
        }
    return 0;
}
