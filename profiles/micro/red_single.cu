// Microbenchmark: cost of per-warp counter atomics that all hit ONE address (the fusion kernel's points/updates
// counters) — how long does a kernel take whose only work is `nwarps` reductions onto the same word?
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k_same(unsigned long long* c, int per_warp) {
    if ((threadIdx.x & 31) == 0) for (int i = 0; i < per_warp; ++i) atomicAdd(c + (i & 1), 1ull);
}
__global__ void k_spread(unsigned long long* c, int per_warp) {
    const unsigned w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if ((threadIdx.x & 31) == 0) for (int i = 0; i < per_warp; ++i) atomicAdd(c + 16 * (w & 1023) + (i & 1), 1ull);
}
__global__ void k_none(unsigned long long* c, int per_warp) { if (per_warp < 0) c[0] = 1; }
int main() {
    unsigned long long* c; cudaMalloc(&c, 8 * 16 * 1024 + 64); cudaMemset(c, 0, 8 * 16 * 1024 + 64);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int blocks : {148, 592, 1184, 2368}) for (int kind = 0; kind < 3; ++kind) {
        float best = 1e9f;
        for (int rep = 0; rep < 6; ++rep) {
            cudaEventRecord(e0);
            if (kind == 0) k_none<<<blocks, 256>>>(c, 2);
            if (kind == 1) k_same<<<blocks, 256>>>(c, 2);
            if (kind == 2) k_spread<<<blocks, 256>>>(c, 2);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
        }
        printf("blocks %4d (%5d warps) %-22s : %7.2f us\n", blocks, blocks * 8, kind == 0 ? "empty kernel" : kind == 1 ? "2 REDs/warp, 2 words" : "2 REDs/warp, spread", best * 1e3f);
    }
    return 0;
}
