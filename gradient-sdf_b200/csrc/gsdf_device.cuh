// Device-side building blocks of the Scan3D hot path: voxel-block hash, key arithmetic, voxel accumulators.
//
// HBM layout (one shard per GPU):
//   keys  : u64[cap]                    open-addressing table of packed 8^3-block keys, EMPTY = ~0
//   grad-sdf  acc4 : float4[cap*512]    per voxel {S = sum w*trunc(sdf), Gx, Gy, Gz = sum w*R*n}
//             accW : float [cap*512]    per voxel  W = sum w
//   base-sdf  acc2 : float2[cap*512]    per voxel {W, S}
// The hash slot IS the block's storage index (no second indirection, no allocator race): a block exists as
// soon as its key is CAS-ed into the table, and its voxels are zero from gsdf_create/gsdf_reset, so fusion
// is a single pass of fire-and-forget vector reductions (red.global.add.v4.f32 / .v2.f32 / .f32).
// dist = S / W is derived on read, which is algebraically the reference's running weighted mean
// (MapGradPixelSdfOmp.cpp:115-116) and order-independent up to fp32 rounding.
//
// Key-determining arithmetic (everything that feeds a branch or a round()) keeps the reference's fp32
// association order with every product and sum rounded separately.  The library is
// compiled with -fmad=false; the *_rn intrinsics below make that explicit where it matters.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace gsdf {

constexpr int kBlockLog2 = 3;                 // 8^3 voxels per block
constexpr int kBlockVoxels = 512;
constexpr unsigned long long kEmptyKey = ~0ull;
constexpr int kMaxProbe = 512;
constexpr int kKeyBias = 1 << 20;             // block coordinates in [-2^20, 2^20)

struct MapView {
    unsigned long long* keys;
    float4* acc4;
    float* accW;
    float2* acc2;
    uint32_t cap_mask;
    int32_t world, rank;
    // optional per-voxel frame visibility (the reference's vis_, MapGradPixelSdfOmp.cpp:118-120), off unless gsdf_vis_enable:
    // vis_stamp[voxel] = 1 + the last fused-frame counter that touched the voxel, vis_log = one {voxel address << 24 | frame}
    // word per (voxel, frame) pair, appended the first time a frame touches the voxel
    unsigned int* vis_stamp;
    unsigned long long* vis_log;
    unsigned long long vis_cap;
};

// Statistics and tickets.  The per-frame counters are spread over kCounterSlots slots (one per CTA modulo the
// slot count, after a CTA-level sum): measured on B200 (profiles/micro/red_single.cu), ~10 k warps each adding to
// the SAME two words cost 11 us of serialised L2 atomics — as much as the whole fusion kernel should take.
constexpr int kCounterSlots = 64;
struct Counters {
    unsigned int blocks;          // allocated blocks
    unsigned int overflow;        // find-or-insert gave up: GSDF_ERR_CAPACITY
    unsigned int ticket;          // last-block-done ticket of the reduction
    unsigned int ticket_snap;     // last-block-done ticket of the frame-result snapshot
    unsigned int barrier;         // arrival counter of track_kernel's grid barrier (0 between launches)
    unsigned int barrier_timeout; // a grid barrier gave up waiting: GSDF_ERR_CUDA
    unsigned int fused_frames;    // Sdf::counter_ (Sdf.h:65): frames integrated so far (increase_counter, MapGradPixelSdfOmp.cpp:128)
    unsigned int vis_overflow;    // the visibility log is full: entries were dropped
    unsigned long long vis_cursor;  // entries in the visibility log
    struct Slot {
        unsigned long long points;   // pixels that passed every validity test
        unsigned long long updates;  // committed (pixel,k) samples
    } slot[kCounterSlots];
};

__device__ __forceinline__ unsigned long long mix64(unsigned long long x) {
    x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
    return x;
}

__device__ __forceinline__ unsigned long long pack_block(int bx, int by, int bz) {
    return ((unsigned long long)(unsigned)(bx + kKeyBias) & 0x1FFFFFull) |
           (((unsigned long long)(unsigned)(by + kKeyBias) & 0x1FFFFFull) << 21) |
           (((unsigned long long)(unsigned)(bz + kKeyBias) & 0x1FFFFFull) << 42);
}
__device__ __forceinline__ void unpack_block(unsigned long long k, int& bx, int& by, int& bz) {
    bx = (int)(k & 0x1FFFFFull) - kKeyBias;
    by = (int)((k >> 21) & 0x1FFFFFull) - kKeyBias;
    bz = (int)((k >> 42) & 0x1FFFFFull) - kKeyBias;
}
__device__ __forceinline__ unsigned long long block_key_of(int vx, int vy, int vz) {
    return pack_block(vx >> kBlockLog2, vy >> kBlockLog2, vz >> kBlockLog2);  // arithmetic shift = floor
}
__device__ __forceinline__ int local_index(int vx, int vy, int vz) {
    return (vx & 7) | ((vy & 7) << 3) | ((vz & 7) << 6);
}
// Shard owner of a block (uses hash bits disjoint from the slot bits).
__device__ __forceinline__ int block_owner(unsigned long long key, int world) {
    return (int)((mix64(key) >> 40) % (unsigned long long)world);
}

// Which voxels a shard keeps (bit 0) and which of them it answers for (bit 1).  grad-sdf: the voxels of its own
// blocks.  base-sdf: a point's trilinear cell is evaluated on the shard owning the block of its FLOOR corner, so that
// shard also keeps copies of the low-face voxels (local coordinate 0) of the up-to-7 neighbour blocks at +1 (halo);
// every shard that keeps a voxel commits every sample of it, so the copies hold the same sums up to atomic order.
template <int VARIANT>
__device__ __forceinline__ int shard_keeps(int world, int rank, unsigned long long key, int vx, int vy, int vz) {
    if (world <= 1 || block_owner(key, world) == rank) return 3;
    if (VARIANT == 0) return 0;
    const int fx = (vx & 7) == 0, fy = (vy & 7) == 0, fz = (vz & 7) == 0;
    if (!(fx | fy | fz)) return 0;
    const int bx = vx >> kBlockLog2, by = vy >> kBlockLog2, bz = vz >> kBlockLog2;
#pragma unroll
    for (int o = 1; o < 8; ++o) {
        const int ox = o & 1, oy = (o >> 1) & 1, oz = o >> 2;
        if (ox <= fx && oy <= fy && oz <= fz && block_owner(pack_block(bx - ox, by - oy, bz - oz), world) == rank) return 1;
    }
    return 0;
}

// Table hash: 32-bit multiply/xorshift mixing (the 64-bit finaliser above costs ~4x the integer instructions and
// sits on the dependent path of every lookup; the hash function is not observable, only the key set is).
__device__ __forceinline__ uint32_t hash_key(unsigned long long key) {
    uint32_t h = (uint32_t)key * 0x9E3779B1u;
    h ^= (uint32_t)(key >> 32) * 0x85EBCA6Bu;
    h ^= h >> 15; h *= 0x2C1B3C6Du; h ^= h >> 12; h *= 0x297A2D39u; h ^= h >> 15;
    return h;
}

// Read-only probe.  Returns the slot or -1.
__device__ __forceinline__ int map_find(const MapView& m, unsigned long long key) {
    uint32_t slot = hash_key(key) & m.cap_mask;
#pragma unroll 1
    for (int probe = 0; probe < kMaxProbe; ++probe) {
        const unsigned long long k = __ldcg(m.keys + slot);
        if (k == key) return (int)slot;
        if (k == kEmptyKey) return -1;
        slot = (slot + 1) & m.cap_mask;
    }
    return -1;
}

// Probe that continues from a first slot whose key has already been loaded (lets a caller issue the first loads of
// several independent lookups together and resolve them afterwards).
__device__ __forceinline__ int map_find_resume(const MapView& m, unsigned long long key, uint32_t slot,
                                               unsigned long long first) {
    if (first == key) return (int)slot;
    if (first == kEmptyKey) return -1;
#pragma unroll 1
    for (int probe = 1; probe < kMaxProbe; ++probe) {
        slot = (slot + 1) & m.cap_mask;
        const unsigned long long k = __ldcg(m.keys + slot);
        if (k == key) return (int)slot;
        if (k == kEmptyKey) return -1;
    }
    return -1;
}

// Find-or-insert (the reference's tsdf_[vi], MapGradPixelSdfOmp.cpp:114, at block granularity).
// Returns the slot, or -1 if kMaxProbe slots were all taken by other keys (caller raises overflow).
__device__ __forceinline__ int map_find_or_insert(const MapView& m, unsigned long long key, Counters* ctr) {
    uint32_t slot = hash_key(key) & m.cap_mask;
#pragma unroll 1
    for (int probe = 0; probe < kMaxProbe; ++probe) {
        unsigned long long k = __ldcg(m.keys + slot);
        if (k == key) return (int)slot;
        if (k == kEmptyKey) {
            k = atomicCAS(m.keys + slot, kEmptyKey, key);
            if (k == kEmptyKey) { atomicAdd(&ctr->blocks, 1u); return (int)slot; }
            if (k == key) return (int)slot;
        }
        slot = (slot + 1) & m.cap_mask;
    }
    return -1;
}

// Fire-and-forget vector reductions into the voxel accumulators, spelled as PTX `red` so that the SASS is REDG (no return
// path) whatever the surrounding code looks like; atomicAdd(float4*) with an unused result came out as ATOMG ... PT, RZ
// inside the fusion loop.
__device__ __forceinline__ void red_add_f4(float4* p, float a, float b, float c, float d) {
    asm volatile("red.relaxed.gpu.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
__device__ __forceinline__ void red_add_f2(float2* p, float a, float b) {
    asm volatile("red.relaxed.gpu.global.add.v2.f32 [%0], {%1, %2};" ::"l"(p), "f"(a), "f"(b) : "memory");
}
__device__ __forceinline__ void red_add_f1(float* p, float a) {
    asm volatile("red.relaxed.gpu.global.add.f32 [%0], %1;" ::"l"(p), "f"(a) : "memory");
}

// Eigen fixed-size 3-term reduction: a0*b0 + (a1*b1 + a2*b2), each product and sum rounded, no FMA.
__device__ __forceinline__ float dot3_rn(float a0, float a1, float a2, float b0, float b1, float b2) {
    return __fadd_rn(__fmul_rn(a0, b0), __fadd_rn(__fmul_rn(a1, b1), __fmul_rn(a2, b2)));
}

// float2vox, MapGradPixelSdf.h:74-77: round-half-away-from-zero of vs_inv * p
__device__ __forceinline__ int float2vox1(float vs_inv, float p) { return (int)roundf(__fmul_rn(vs_inv, p)); }

// Sdf::weight, Sdf.h:76-85
__device__ __forceinline__ float sdf_weight(float sdf, float T, float inv_T) {
    float w = 0.f;
    if (sdf <= 0.f) w = 1.f;
    else if (sdf <= T) w = __fsub_rn(1.f, __fmul_rn(sdf, inv_T));
    return w;
}
// Sdf::truncate, Sdf.h:72-74
__device__ __forceinline__ float sdf_truncate(float sdf, float T) { return fmaxf(-T, fminf(T, sdf)); }

}  // namespace gsdf
