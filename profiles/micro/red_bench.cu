// Microbenchmark: throughput of fire-and-forget global reductions (REDG) as a function of how many threads hit
// the same voxel at about the same time.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 red_bench.cu -o red_bench
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ unsigned mixu(unsigned h) { h ^= h >> 15; h *= 0x2C1B3C6Du; h ^= h >> 12; h *= 0x297A2D39u; h ^= h >> 15; return h; }

// n updates; update i targets voxel (i / dup) scrambled over `nvox` voxels; mode 0: float4 + float, 1: float2, 2: float
template <int MODE>
__global__ void red_kernel(float4* a4, float* a1, float2* a2, unsigned n, unsigned dup, unsigned nvox, int scramble) {
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        unsigned v = i / dup;
        if (scramble) v = mixu(v) % nvox; else v = v % nvox;
        if (MODE == 0) { atomicAdd(a4 + v, make_float4(1.f, 2.f, 3.f, 4.f)); atomicAdd(a1 + v, 1.f); }
        if (MODE == 1) atomicAdd(a2 + v, make_float2(1.f, 2.f));
        if (MODE == 2) atomicAdd(a1 + v, 1.f);
    }
}
// same but duplicates are far apart in thread index (stride n/dup): same address, different time/SM
template <int MODE>
__global__ void red_kernel_strided(float4* a4, float* a1, float2* a2, unsigned n, unsigned dup, unsigned nvox) {
    const unsigned per = n / dup;
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        unsigned v = mixu(i % per) % nvox;
        if (MODE == 0) { atomicAdd(a4 + v, make_float4(1.f, 2.f, 3.f, 4.f)); atomicAdd(a1 + v, 1.f); }
        if (MODE == 1) atomicAdd(a2 + v, make_float2(1.f, 2.f));
        if (MODE == 2) atomicAdd(a1 + v, 1.f);
    }
}

int main() {
    const unsigned nvox = 1u << 21;  // 2M voxels (32 MB of float4)
    float4* a4; float* a1; float2* a2;
    cudaMalloc(&a4, sizeof(float4) * nvox); cudaMalloc(&a1, sizeof(float) * nvox); cudaMalloc(&a2, sizeof(float2) * nvox);
    cudaMemset(a4, 0, sizeof(float4) * nvox); cudaMemset(a1, 0, sizeof(float) * nvox); cudaMemset(a2, 0, sizeof(float2) * nvox);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const unsigned n = 1u << 20;  // ~1M updates, like one frame
    const int grid = 148 * 8, block = 256;
    printf("n=%u updates, grid %d x %d\n", n, grid, block);
    for (int mode = 0; mode < 3; ++mode)
        for (int pattern = 0; pattern < 3; ++pattern)
            for (unsigned dup : {1u, 4u, 8u, 16u, 32u, 64u}) {
                float best = 1e9f;
                for (int rep = 0; rep < 5; ++rep) {
                    cudaEventRecord(e0);
                    if (pattern < 2) {
                        if (mode == 0) red_kernel<0><<<grid, block>>>(a4, a1, a2, n, dup, nvox, pattern);
                        if (mode == 1) red_kernel<1><<<grid, block>>>(a4, a1, a2, n, dup, nvox, pattern);
                        if (mode == 2) red_kernel<2><<<grid, block>>>(a4, a1, a2, n, dup, nvox, pattern);
                    } else {
                        if (mode == 0) red_kernel_strided<0><<<grid, block>>>(a4, a1, a2, n, dup, nvox);
                        if (mode == 1) red_kernel_strided<1><<<grid, block>>>(a4, a1, a2, n, dup, nvox);
                        if (mode == 2) red_kernel_strided<2><<<grid, block>>>(a4, a1, a2, n, dup, nvox);
                    }
                    cudaEventRecord(e1); cudaEventSynchronize(e1);
                    float ms; cudaEventElapsedTime(&ms, e0, e1);
                    if (ms < best) best = ms;
                }
                printf("mode %d (%s) pattern %d (%s) dup %2u : %8.2f us  %7.2f Gupd/s\n", mode,
                       mode == 0 ? "f4+f1" : mode == 1 ? "f2" : "f1", pattern,
                       pattern == 0 ? "adjacent lanes, contiguous voxels" : pattern == 1 ? "adjacent lanes, scattered voxels" : "duplicates far apart",
                       dup, best * 1e3f, n / (best * 1e-3f) / 1e9f);
            }
    return 0;
}
