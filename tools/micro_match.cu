// Microbenchmark: how to find, for every lane of a warp, the nearest lower lane holding the same key.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/micro_match tools/micro_match.cu && /tmp/micro_match
// Variants: MATCH.ANY, a 31-step shuffle scan, and a broadcast-LDS scan; all 32 warps of a CTA busy, one CTA per SM.
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t mix(uint32_t x) { x ^= x >> 15; x *= 0x2c1b3c6du; x ^= x >> 12; x *= 0x297a2d39u; x ^= x >> 15; return x; }

template <int V>
__global__ void __launch_bounds__(1024, 1) k(uint32_t* out, int iters, uint32_t kmask)
{
    __shared__ uint16_t keys[1024];
    const uint32_t tid = threadIdx.x, lane = tid & 31u;
    uint32_t acc = 0, seed = mix(tid + blockIdx.x * 1024u + 1u);
    for (int it = 0; it < iters; it++) {
        seed = seed * 1664525u + 1013904223u;
        const uint32_t key = (seed >> 9) & kmask;
        uint32_t pred = 0;
        if (V == 0) {
            const uint32_t same = __match_any_sync(0xFFFFFFFFu, key);
            const uint32_t lower = same & ((1u << lane) - 1u);
            pred = lower ? lane - (31u - __clz(lower)) : 0u;
        } else if (V == 1) {
#pragma unroll
            for (int d = 31; d >= 1; d--) {
                const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, key, d);
                if (lane >= (uint32_t)d && o == key) pred = d;
            }
        } else {
            keys[tid] = (uint16_t)key;
            __syncwarp();
            const uint4* kp = reinterpret_cast<const uint4*>(&keys[tid & ~31u]);
            const uint32_t dup = key * 0x10001u;
            uint32_t eq = 0;
#pragma unroll
            for (int q = 0; q < 4; q++) {
                const uint4 v = kp[q];
                const uint32_t w[4] = {v.x ^ dup, v.y ^ dup, v.z ^ dup, v.w ^ dup};
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    if ((w[j] & 0xFFFFu) == 0u) eq |= 1u << (q * 8 + j * 2);
                    if (w[j] < 0x10000u) eq |= 1u << (q * 8 + j * 2 + 1);
                }
            }
            const uint32_t lower = eq & ((1u << lane) - 1u);
            pred = lower ? lane - (31u - __clz(lower)) : 0u;
            __syncwarp();
        }
        acc += pred;
    }
    out[blockIdx.x * 1024u + tid] = acc;
}

template <int V>
static void run(const char* name, uint32_t kmask)
{
    uint32_t* d; cudaMalloc(&d, 148 * 1024 * 4);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    const int iters = 20000;
    k<V><<<148, 1024>>>(d, 100, kmask);
    cudaEventRecord(a);
    k<V><<<148, 1024>>>(d, iters, kmask);
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    // per SM: 32 warps * iters group-ops in ms; cycles per group-op per SM at ~1.9 GHz
    printf("%-10s kmask=%#x: %.3f ms, %.1f ns per warp-op per SM (%.0f cycles @1.9GHz), err=%d\n", name, kmask, ms,
           ms * 1e6 / (32.0 * iters), ms * 1e6 / (32.0 * iters) * 1.9, (int)cudaGetLastError());
    cudaFree(d);
}

int main()
{
    for (uint32_t km : {0xFFFu, 0x3Fu}) {
        run<0>("match.any", km);
        run<1>("shfl-scan", km);
        run<2>("lds-scan", km);
    }
    return 0;
}
