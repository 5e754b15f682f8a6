// Microbenchmark + unit test of the two-pass insertion for the single-entry 3-byte / 4-byte tables (k_lz77, round 2).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I include -I gzippier_b200/csrc -o tools/bin/micro_prepass \
//        tools/micro_prepass.cu gzippier_b200/csrc/corpus.c && tools/bin/micro_prepass
// Pass 1 (any warp, one 32-position group at a time, groups independent): MATCH.ANY on the bucket tells every lane its nearest
//   lower lane with the same bucket and whether it is the last of its bucket in the group -> one byte per position.
// Pass 2 (one warp per table, groups in order): lanes without a predecessor in the group read the table, the last lane of
//   every bucket writes it -- no read-back, no votes, so consecutive groups pipeline.
// Questions: cycles per tile of pass 2 under load; what the MATCH traffic of pass 1 costs the warps that do other work.
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
    uint16_t head3[1 << H3BITS], head4[1 << H4BITS];
    uint8_t side[2][TILE];
    uint32_t junk[4096];
    uint32_t done, pctr;
};
constexpr uint32_t TABLE_FILL = 0x80008000u;
__device__ __forceinline__ uint32_t lds_u16(uint32_t a) { uint32_t v; asm volatile("ld.shared.u16 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
__device__ __forceinline__ uint32_t lds_u8(uint32_t a) { uint32_t v; asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
__device__ __forceinline__ uint32_t lds_u32(uint32_t a) { uint32_t v; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
__device__ __forceinline__ void sts_u16(uint32_t a, uint32_t v) { asm volatile("st.shared.u16 [%0], %1;" :: "r"(a), "h"((uint16_t)v) : "memory"); }
__device__ __forceinline__ void sts_u8(uint32_t a, uint32_t v) { asm volatile("st.shared.u8 [%0], %1;" :: "r"(a), "r"(v) : "memory"); }
template <int KIND> __device__ __forceinline__ uint32_t to_delta(uint32_t p16, uint32_t raw)
{
    const uint32_t d = (p16 - raw) & 0xFFFFu;
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
__device__ __forceinline__ uint32_t word_at(uint32_t ring_s, uint32_t p)
{
    const uint32_t a = p & 0xFFFFu, wa = ring_s + (a & 0xFFFCu), sh = (a & 3u) * 8u;
    return __funnelshift_r(lds_u32(wa), lds_u32(wa + 4u), sh);
}
// pass 1, one group: side byte = distance to the nearest lower lane of the bucket (0: none) | 0x80 if last of the bucket
__device__ __forceinline__ void prep_group(uint32_t sbase, uint32_t t, uint32_t g, uint32_t lane)
{
    const uint32_t ring_s = sbase + (uint32_t)offsetof(Smem, ring32);
    const uint32_t v = word_at(ring_s, t * TILE + g * 32u + lane);
    const uint32_t b3 = hash3(v), b4 = hash4(v);
    const uint32_t s3 = __match_any_sync(0xFFFFFFFFu, b3), s4 = __match_any_sync(0xFFFFFFFFu, b4);
    const uint32_t lt = (1u << lane) - 1u;
    const uint32_t l3 = s3 & lt, l4 = s4 & lt;
    const uint32_t e3 = (l3 ? lane - (uint32_t)(31 - __clz(l3)) : 0u) | ((s3 >> lane) == 1u ? 0x80u : 0u);
    const uint32_t e4 = (l4 ? lane - (uint32_t)(31 - __clz(l4)) : 0u) | ((s4 >> lane) == 1u ? 0x80u : 0u);
    const uint32_t o = g * 32u + lane;
    sts_u8(sbase + (uint32_t)offsetof(Smem, side) + o, e3);
    sts_u8(sbase + (uint32_t)offsetof(Smem, side) + TILE + o, e4);
}
// pass 2
template <int KIND, int U>
__device__ void stream_tile(uint32_t sbase, uint32_t t, uint32_t lane, uint16_t* __restrict__ gout)
{
    const uint32_t ring_s = sbase + (uint32_t)offsetof(Smem, ring32);
    const uint32_t table_s = sbase + (uint32_t)(KIND == 1 ? offsetof(Smem, head3) : offsetof(Smem, head4));
    const uint32_t side_s = sbase + (uint32_t)offsetof(Smem, side) + (KIND == 1 ? 0u : (uint32_t)TILE);
    const uint32_t base = t * TILE;
    {
        const uint32_t p16 = base & 0xFFFFu, slice = t & 7u;
        uint4* tab = reinterpret_cast<uint4*>(__cvta_shared_to_generic(table_s)) + slice * 64u;
        for (uint32_t i = lane; i < 64u; i += 32u) {
            uint4 v = tab[i];
            v.x = sweep_pair(v.x, p16); v.y = sweep_pair(v.y, p16); v.z = sweep_pair(v.z, p16); v.w = sweep_pair(v.w, p16);
            tab[i] = v;
        }
        __syncwarp();
    }
#pragma unroll 1
    for (uint32_t g0 = 0; g0 < (uint32_t)TILE / 32u; g0 += U) {
        uint32_t ta[U], e[U];
#pragma unroll
        for (int u = 0; u < U; u++) {
            const uint32_t o = (g0 + u) * 32u + lane;
            const uint32_t v = word_at(ring_s, base + o);
            ta[u] = table_s + 2u * (KIND == 1 ? hash3(v) : hash4(v));
            e[u] = lds_u8(side_s + o);
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            const uint32_t p16 = (base + (g0 + u) * 32u + lane) & 0xFFFFu;
            const uint32_t dl = e[u] & 31u;
            uint32_t raw = (p16 - dl) & 0xFFFFu;
            if (dl == 0u) raw = lds_u16(ta[u]);
            if (e[u] & 0x80u) sts_u16(ta[u], p16);
            __syncwarp();
            gout[(size_t)base + (g0 + u) * 32u + lane] = (uint16_t)to_delta<KIND>(p16, raw);
        }
    }
}

// mode 0: the 30 other warps idle; 1: they run a fixed amount of LDS+ALU work per tile (timed) while pass 1 / pass 2 run on
// dedicated warps; 2: the same work with pass 1 spread over all of them (2 groups per warp) before they start
template <int U>
__global__ void __launch_bounds__(1024, 1)
k(const uint8_t* __restrict__ in, uint32_t ntiles, uint16_t* o3, uint16_t* o4, int mode, int work, unsigned long long* cyc)
{
    extern __shared__ __align__(16) unsigned char raw[];
    Smem& S = *reinterpret_cast<Smem*>(raw);
    uint32_t sbase; asm volatile("{ .reg .u64 t; cvta.to.shared.u64 t, %1; cvt.u32.u64 %0, t; }" : "=r"(sbase) : "l"(raw));
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const size_t off = (size_t)blockIdx.x * ntiles * TILE;
    in += off; o3 += off; o4 += off;
    for (uint32_t i = tid; i < (1u << H3BITS) / 2; i += 1024u) { reinterpret_cast<uint32_t*>(S.head3)[i] = TABLE_FILL; reinterpret_cast<uint32_t*>(S.head4)[i] = TABLE_FILL; }
    for (uint32_t i = tid; i < 4096u; i += 1024u) S.junk[i] = i * 2654435761u;
    unsigned long long a_prep = 0, a_ser = 0, a_work = 0; uint32_t sink = 0;
    auto load = [&](uint32_t t) {
        if (t > ntiles) return;
        const uint32_t pos = t * TILE + tid * 2u;
        uint16_t v = t < ntiles ? *reinterpret_cast<const uint16_t*>(in + pos) : (uint16_t)0;
        reinterpret_cast<uint16_t*>(S.ring32)[(pos & 0xFFFFu) >> 1] = v;
        if ((pos & 0xFFFFu) < (uint32_t)RING_MIRROR) reinterpret_cast<uint16_t*>(S.ring32)[(RING + (pos & 0xFFFFu)) >> 1] = v;
    };
    load(0);
    for (uint32_t t = 0; t < ntiles; t++) {
        load(t + 1);
        if (tid == 0) S.pctr = 0;
        __syncthreads();
        // pass 1
        long long c0 = clock64();
        if (mode == 2) { prep_group(sbase, t, warp * 2u, lane); prep_group(sbase, t, warp * 2u + 1u, lane); }
        else if (warp < 4u) for (uint32_t g = warp; g < 64u; g += 4u) prep_group(sbase, t, g, lane);   // one warp per scheduler
        if (warp == 0) a_prep += (unsigned long long)(clock64() - c0);
        if (mode != 1) __syncthreads();
        else asm volatile("bar.sync 1, 128;" ::: "memory");                    // mode 1: only the four pass-1 warps meet here
        // pass 2 on warps 1 and 2 (two schedulers), the work on warps 4..31
        c0 = clock64();
        if (warp == 1u) { stream_tile<1, U>(sbase, t, lane, o3); a_ser += (unsigned long long)(clock64() - c0); }
        else if (warp == 2u) { stream_tile<2, U>(sbase, t, lane, o4); a_ser += (unsigned long long)(clock64() - c0); }
        else if (warp >= 4u && mode != 0) {
            uint32_t x = tid;
            for (int it = 0; it < work; it++) {
#pragma unroll
                for (int j = 0; j < 16; j++) x = S.junk[(x + (uint32_t)j) & 4095u] ^ (x >> 3);
            }
            sink ^= x;
            if (warp == 4u) a_work += (unsigned long long)(clock64() - c0);
        }
        __syncthreads();
    }
    if (lane == 0 && warp == 0) atomicAdd(&cyc[0], a_prep);
    if (lane == 0 && (warp == 1 || warp == 2)) atomicAdd(&cyc[warp], a_ser);
    if (lane == 0 && warp == 4) atomicAdd(&cyc[3], a_work);
    if (sink == 0x12345u) cyc[4] = sink;
}

static void cpu_ref(const uint8_t* s, size_t n, std::vector<uint16_t>& o3, std::vector<uint16_t>& o4)
{
    std::vector<int64_t> h3(1 << H3BITS, -1), h4(1 << H4BITS, -1);
    for (size_t p = 0; p < n; p++) {
        uint8_t tmp[4] = {0};
        for (int k = 0; k < 4 && p + k < n; k++) tmp[k] = s[p + k];
        uint32_t lo; memcpy(&lo, tmp, 4);
        const uint32_t b = hash3(lo), c = hash4(lo);
        const int64_t d3 = h3[b] < 0 ? 0 : (int64_t)p - h3[b], d4 = h4[c] < 0 ? 0 : (int64_t)p - h4[c];
        o3[p] = (d3 >= 1 && d3 <= TOO_FAR) ? (uint16_t)d3 : 0;
        o4[p] = (d4 >= 1 && d4 <= 32767) ? (uint16_t)d4 : 0;
        h3[b] = h4[c] = (int64_t)p;
    }
}

template <int U>
static void run(const uint8_t* d_in, const std::vector<uint8_t>& host, uint32_t ntiles, int nblk, int mode, int work)
{
    const size_t n = (size_t)nblk * ntiles * TILE;
    uint16_t *o3, *o4; unsigned long long* cyc;
    cudaMalloc(&o3, n * 2); cudaMalloc(&o4, n * 2); cudaMalloc(&cyc, 64); cudaMemset(cyc, 0, 64);
    cudaFuncSetAttribute(k<U>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem));
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    cudaEventRecord(a);
    k<U><<<nblk, 1024, sizeof(Smem)>>>(d_in, ntiles, o3, o4, mode, work, cyc);
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    unsigned long long h[8]; cudaMemcpy(h, cyc, 64, cudaMemcpyDeviceToHost);
    size_t bad = 0;
    for (int blk : {0, nblk - 1}) {
        const size_t off = (size_t)blk * ntiles * TILE, len = (size_t)ntiles * TILE;
        std::vector<uint16_t> g3(len), g4(len), r3(len), r4(len);
        cudaMemcpy(g3.data(), o3 + off, len * 2, cudaMemcpyDeviceToHost); cudaMemcpy(g4.data(), o4 + off, len * 2, cudaMemcpyDeviceToHost);
        cpu_ref(host.data() + off, len, r3, r4);
        for (size_t i = 0; i + 8 < len; i++) bad += (g3[i] != r3[i]) + (g4[i] != r4[i]);
    }
    const double per = 1.0 / ((double)nblk * ntiles);
    printf("U=%d mode=%d work=%d: %.2f ms, cycles per tile: pass1 (warp 0) %.0f, pass2 h3 %.0f h4 %.0f, other work (warp 4) %.0f, mismatches %zu, err %d\n",
           U, mode, work, ms, h[0] * per, h[1] * per, h[2] * per, h[3] * per, bad, (int)cudaGetLastError());
    cudaFree(o3); cudaFree(o4); cudaFree(cyc);
}

int main(int argc, char** argv)
{
    const int kind = argc > 1 ? atoi(argv[1]) : GZB_CORPUS_TEXT;
    const uint32_t ntiles = 256; const int nblk = 148;
    const size_t n = (size_t)nblk * ntiles * TILE;
    std::vector<uint8_t> host(n + 64);
    gzb_corpus_fill(kind, 1234, 0, host.data(), n, 16);
    uint8_t* d_in; cudaMalloc(&d_in, n + 64); cudaMemcpy(d_in, host.data(), n, cudaMemcpyHostToDevice);
    run<4>(d_in, host, ntiles, nblk, 0, 0);
    run<8>(d_in, host, ntiles, nblk, 0, 0);
    for (int work : {100, 200}) {
        // reference: the work alone (pass 1 still runs on 4 warps but nobody waits for it; pass 2 runs beside the work)
        run<4>(d_in, host, ntiles, nblk, 1, work);
        run<4>(d_in, host, ntiles, nblk, 2, work);
        run<8>(d_in, host, ntiles, nblk, 2, work);
    }
    return 0;
}
