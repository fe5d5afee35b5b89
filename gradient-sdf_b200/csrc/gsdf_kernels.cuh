// Hand-written sm_100a kernels of the Scan3D hot path.  Included by gsdf_api.cu only.
// (normals and fusion live in gsdf_pipeline.cuh: normals_tile, fuse_work)
//   accumulate_kernel<V> RigidPointOptimizer per-pixel loop + sums   (RigidPointOptimizerOmp.cpp:69-91)
//   query_kernel<V>     Sdf::weights / Sdf::tsdf                     (MapGradPixelSdf.h:109-125,
//                                                                     MapPixelSdf.h:107-143, MapPixelSdfOmp.cpp:44-113)
//   count/download      map iteration of the writers                 (MapGradPixelSdfOmp.cpp:144-151,189,285)
#pragma once

#include "gsdf_device.cuh"

namespace gsdf {

struct FrameGeom {
    int w, h;
    float vs, vs_inv, T, inv_T, z_min, z_max;
    int factor;  // floor(T / vs), MapGradPixelSdfOmp.cpp:78
};

struct Pose {
    float R[9];
    float t[3];
};

// ---------------------------------------------------------------------------------------------------------
// Normal estimation (FALS), tile geometry shared by the kernels in gsdf_pipeline.cuh.
constexpr int NT_W = 32, NT_H = 16, N_MAXR = 7;

__device__ __forceinline__ int reflect101(int p, int len) {
    if (len == 1) return 0;
    while (p < 0 || p >= len) p = p < 0 ? -p : 2 * (len - 1) - p;
    return p;
}

// ---------------------------------------------------------------------------------------------------------
// Point queries shared by the tracker and gsdf_query.
template <int VARIANT>
struct Lookup;

template <>
struct Lookup<0> {  // MapGradPixelSdf.h:109-125
    // returns weight; if > 0 fills phi and grad
    // ck/cs: caller-held cache of the last probed block (key -> slot, slot < 0 = absent).  The map does not change
    // while it is being tracked against, so a thread that revisits the same pixel in the next Gauss-Newton pass
    // (the pose moves by far less than a block) skips the dependent key load.
    __device__ static __forceinline__ float eval(const MapView& m, const FrameGeom& g, float px, float py, float pz,
                                                 float& phi, float& g0, float& g1, float& g2,
                                                 unsigned long long& ck, int& cs) {
        const int vx = float2vox1(g.vs_inv, px), vy = float2vox1(g.vs_inv, py), vz = float2vox1(g.vs_inv, pz);
        const unsigned long long key = block_key_of(vx, vy, vz);
        if (m.world > 1 && block_owner(key, m.world) != m.rank) return 0.f;
        if (key != ck) { ck = key; cs = map_find(m, key); }
        const int slot = cs;
        if (slot < 0) return 0.f;
        const size_t addr = (size_t)slot * kBlockVoxels + local_index(vx, vy, vz);
        // both loads issued before either is looked at: one memory round trip instead of two dependent ones
        const float W = m.accW[addr];
        const float4 a = m.acc4[addr];
        if (!(W > 0.f)) return 0.f;
        // phi and grad are tolerance-level quantities (sums of them are compared at 1e-4), not key-determining:
        // approximate reciprocal / rsqrt (<= 2 ulp) instead of IEEE division on the tracker's critical path
        const float dist = __fdividef(a.x, W);
        const float sq = dot3_rn(a.y, a.z, a.w, a.y, a.z, a.w);
        float n0 = a.y, n1 = a.z, n2 = a.w;
        if (sq > 0.f) { const float il = rsqrtf(sq); n0 *= il; n1 *= il; n2 *= il; }
        const float dx = __fsub_rn(__fmul_rn(g.vs, (float)vx), px);
        const float dy = __fsub_rn(__fmul_rn(g.vs, (float)vy), py);
        const float dz = __fsub_rn(__fmul_rn(g.vs, (float)vz), pz);
        g0 = 1.2f * n0; g1 = 1.2f * n1; g2 = 1.2f * n2;
        // `1.2*v.grad.normalized().dot(..)` is a double product in the reference (MapGradPixelSdf.h:114)
        phi = (float)((double)dist + 1.2 * (double)dot3_rn(n0, n1, n2, dx, dy, dz));
        return W;
    }
};

template <>
struct Lookup<1> {  // MapPixelSdf.h:107-143 + interp3, MapPixelSdfOmp.cpp:44-113
    __device__ static __forceinline__ float eval(const MapView& m, const FrameGeom& g, float px, float py, float pz,
                                                 float& phi, float& g0, float& g1, float& g2,
                                                 unsigned long long& ck, int& cs) {
        const float fx = __fmul_rn(g.vs_inv, px), fy = __fmul_rn(g.vs_inv, py), fz = __fmul_rn(g.vs_inv, pz);
        const int im = (int)floorf(fx), jm = (int)floorf(fy), km = (int)floorf(fz);
        // sharded map: the cell is evaluated on the rank that owns its floor corner's block (which holds the halo)
        if (m.world > 1 && block_owner(block_key_of(im, jm, km), m.world) != m.rank) return 0.f;
        float d[8];
        float Wn = 0.f;
        const int rx = (int)roundf(fx) - im, ry = (int)roundf(fy) - jm, rz = (int)roundf(fz) - km;  // nearest corner
        // resolve the (at most 8, usually 1) blocks first, then issue all 8 voxel loads before looking at any of
        // them: one memory round trip instead of eight dependent ones
        const int cn = rx | (ry << 1) | (rz << 2);
        float2 v[8];
        const int lx = im & 7, ly = jm & 7, lz = km & 7;
        if (lx < 7 && ly < 7 && lz < 7) {  // (7/8)^3 = 67 % of the cells lie inside one block: one probe, fixed offsets
            const unsigned long long key = block_key_of(im, jm, km);
            if (key != ck) { ck = key; cs = map_find(m, key); }
            if (cs < 0) return 0.f;
            const float2* base = m.acc2 + (size_t)cs * kBlockVoxels + (lx | (ly << 3) | (lz << 6));
#pragma unroll
            for (int c = 0; c < 8; ++c) v[c] = base[(c & 1) + ((c >> 1) & 1) * 8 + (c >> 2) * 64];
        } else {
            // the cell straddles 2, 4 or 8 blocks: a corner (i0, j0, k0) lies in block base + (i0 & qx, j0 & qy, k0 & qz)
            // where q* tells whether that axis crosses; only those block combinations are probed (29 % of all cells
            // need 2 probes, 4 % need 4, 0.2 % need 8)
            const int qx = lx == 7, qy = ly == 7, qz = lz == 7;
            const int bx = im >> kBlockLog2, by = jm >> kBlockLog2, bz = km >> kBlockLog2;
            // first issue the key loads of ALL needed blocks, then resolve them: the probes are independent, and a
            // lane that needs 8 of them would otherwise pay 8 dependent L2 round trips (the slowest warp of a CTA sets
            // the length of every Gauss-Newton pass: measured 9 us against 3 us for grad-sdf before this change)
            int slots[8];
            uint32_t hs[8];
            unsigned long long want[8], got[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                const int ex = e & 1, ey = (e >> 1) & 1, ez = e >> 2;
                const bool need = (ex <= qx) && (ey <= qy) && (ez <= qz);
                want[e] = pack_block(bx + ex, by + ey, bz + ez);
                hs[e] = hash_key(want[e]) & m.cap_mask;
                got[e] = kEmptyKey;
                if (need && !(e == 0 && want[0] == ck)) got[e] = __ldcg(m.keys + hs[e]);
            }
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                const int ex = e & 1, ey = (e >> 1) & 1, ez = e >> 2;
                const bool need = (ex <= qx) && (ey <= qy) && (ez <= qz);
                slots[e] = 0;
                if (need) {
                    if (e == 0 && want[0] == ck) slots[0] = cs;
                    else slots[e] = map_find_resume(m, want[e], hs[e], got[e]);
                    if (e == 0) { ck = want[0]; cs = slots[0]; }
                    if (slots[e] < 0) return 0.f;  // weights(): all 8 corners must exist (MapPixelSdf.h:127-136)
                }
            }
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                const int i0 = c & 1, j0 = (c >> 1) & 1, k0 = c >> 2;
                const int e = (i0 & qx) | ((j0 & qy) << 1) | ((k0 & qz) << 2);
                v[c] = m.acc2[(size_t)slots[e] * kBlockVoxels + local_index(im + i0, jm + j0, km + k0)];
            }
        }
        bool all = true;
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            all = all && (v[c].x > 0.f);
            d[c] = __fdividef(v[c].y, v[c].x);  // tolerance-level quantity, see Lookup<0>
            Wn = (c == cn) ? v[c].x : Wn;
        }
        if (!all) return 0.f;
        const float dx = fx - (float)im, dy = fy - (float)jm, dz = fz - (float)km;
        const float d01 = (1.f - dx) * d[0] + dx * d[1], d23 = (1.f - dx) * d[2] + dx * d[3];
        const float d45 = (1.f - dx) * d[4] + dx * d[5], d67 = (1.f - dx) * d[6] + dx * d[7];
        const float d02 = (1.f - dy) * d[0] + dy * d[2], d13 = (1.f - dy) * d[1] + dy * d[3];
        const float d46 = (1.f - dy) * d[4] + dy * d[6], d57 = (1.f - dy) * d[5] + dy * d[7];
        g0 = g.vs_inv * ((((1.f - dz) * d13 + dz * d57) - (1.f - dz) * d02) - dz * d46);  // :97
        g1 = g.vs_inv * ((((1.f - dz) * d23 + dz * d67) - (1.f - dz) * d01) - dz * d45);  // :98
        g2 = g.vs_inv * ((((1.f - dy) * d45 + dy * d67) - (1.f - dy) * d01) - dy * d23);  // :99
        const float dy0 = (1.f - dy) * d01 + dy * d23, dy1 = (1.f - dy) * d45 + dy * d67;
        phi = (1.f - dz) * dy0 + dz * dy1;  // :107
        return Wn;
    }
};

template <int VARIANT>
__global__ void __launch_bounds__(256)
query_kernel(MapView m, FrameGeom g, const float* __restrict__ pts, long long n, float* __restrict__ phi,
             float* __restrict__ grad, float* __restrict__ w) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float f = 0.f, g0 = 0.f, g1 = 0.f, g2 = 0.f;
    unsigned long long ck = kEmptyKey;
    int cs = -1;
    const float w0 = Lookup<VARIANT>::eval(m, g, pts[3 * i], pts[3 * i + 1], pts[3 * i + 2], f, g0, g1, g2, ck, cs);
    if (!(w0 > 0.f)) { f = 0.f; g0 = g1 = g2 = 0.f; }
    w[i] = w0; phi[i] = f;
    grad[3 * i] = g0; grad[3 * i + 1] = g1; grad[3 * i + 2] = g2;
}

// ---------------------------------------------------------------------------------------------------------
// Tracker: per pixel residual + Jacobian, then the 29 sums (g[6], upper H[21], E, counter) by a
// warp-shuffle -> shared-memory -> grid (last-block-done, fixed order => bit-reproducible) tree reduction.
constexpr int kAccTerms = 29;
constexpr int kAccStride = 32;
constexpr int kAccThreads = 256;

struct TrackIntr {
    float fx_inv, fy_inv, cx, cy;  // RigidPointOptimizerOmp.cpp:48-52
    int sampling, ws, hs;          // sampled grid size
};

template <int VARIANT>
__global__ void __launch_bounds__(kAccThreads)
accumulate_kernel(MapView m, FrameGeom g, TrackIntr ti, const Pose* __restrict__ pose_ptr,
                  const float* __restrict__ depth, float* __restrict__ partials, Counters* ctr,
                  float* __restrict__ out29) {
    // pose_ptr: device-resident pose (kept on the device so a Gauss-Newton loop needs no host round trip)
    float acc[kAccTerms];
#pragma unroll
    for (int i = 0; i < kAccTerms; ++i) acc[i] = 0.f;
    __shared__ Pose pose;
    if (threadIdx.x < 12) reinterpret_cast<float*>(&pose)[threadIdx.x] = reinterpret_cast<const float*>(pose_ptr)[threadIdx.x];
    __syncthreads();
    const float* R = pose.R;
    const int total = ti.ws * ti.hs;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int ys = i / ti.ws, xs = i - ys * ti.ws;
        const int x = xs * ti.sampling, y = ys * ti.sampling;
        const float z = depth[(size_t)y * g.w + x];
        if (z <= g.z_min || z >= g.z_max) continue;  // :72-73
        const float x0 = ((float)x - ti.cx) * ti.fx_inv, y0 = ((float)y - ti.cy) * ti.fy_inv;  // :75-76
        const float cx_ = x0 * z, cy_ = y0 * z;
        const float px = dot3_rn(R[0], R[1], R[2], cx_, cy_, z) + pose.t[0];  // :78
        const float py = dot3_rn(R[3], R[4], R[5], cx_, cy_, z) + pose.t[1];
        const float pz = dot3_rn(R[6], R[7], R[8], cx_, cy_, z) + pose.t[2];
        float phi, g0, g1, g2;
        unsigned long long ck = kEmptyKey;
        int cs = -1;
        const float w0 = Lookup<VARIANT>::eval(m, g, px, py, pz, phi, g0, g1, g2, ck, cs);
        if (w0 > 0.f) {
            float J[6];
            J[0] = g0; J[1] = g1; J[2] = g2;
            J[3] = py * g2 - pz * g1;  // point.cross(grad), :86
            J[4] = pz * g0 - px * g2;
            J[5] = px * g1 - py * g0;
            int o = 6;
#pragma unroll
            for (int a = 0; a < 6; ++a) {
                acc[a] += phi * J[a];  // :87
#pragma unroll
                for (int b = a; b < 6; ++b) acc[o++] += J[a] * J[b];  // :88 (upper triangle)
            }
            acc[27] += phi * phi;  // :84
            acc[28] += 1.f;        // :89
        }
    }
    // warp tree
#pragma unroll
    for (int i = 0; i < kAccTerms; ++i) {
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) acc[i] += __shfl_xor_sync(0xffffffffu, acc[i], off);
    }
    __shared__ float sm[kAccThreads / 32][kAccStride];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) {
#pragma unroll
        for (int i = 0; i < kAccTerms; ++i) sm[warp][i] = acc[i];
    }
    __syncthreads();
    if (threadIdx.x < kAccTerms) {
        float s = 0.f;
#pragma unroll
        for (int wq = 0; wq < kAccThreads / 32; ++wq) s += sm[wq][threadIdx.x];
        partials[(size_t)blockIdx.x * kAccStride + threadIdx.x] = s;
    }
    // grid stage: the last block to arrive sums all partials in block order
    __shared__ bool is_last;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned int t = atomicInc(&ctr->ticket, gridDim.x - 1);  // wraps back to 0 for the next launch
        is_last = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    // warp wq reduces terms wq, wq+8, ...; lanes stride over blocks, then a fixed-shape shuffle tree
    for (int term = warp; term < kAccTerms; term += kAccThreads / 32) {
        float s = 0.f;
        for (int b = lane; b < (int)gridDim.x; b += 32) s += __ldcg(partials + (size_t)b * kAccStride + term);
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) s += __shfl_xor_sync(0xffffffffu, s, off);
        if (lane == 0) out29[term] = s;
    }
}

// ---------------------------------------------------------------------------------------------------------
// Normal-equation sums for callers that cannot allocate: same per-pixel maths and the same warp -> block ->
// last-block reduction as accumulate_kernel, but the per-block partial sums and the arrival ticket live in a
// caller-independent scratch array that the library owns statically (see gsdf_accumulate_noalloc in
// gsdf_api.cu), and the pose arrives by value.  One CTA per 256 sampled pixels, so the scratch needed is
// 32 floats per CTA; kStaticMaxTiles bounds it.
constexpr int kStaticMaxTiles = 4096;                               // 4096 * 256 = 1,048,576 sampled pixels
constexpr int kStaticTicketOffset = kStaticMaxTiles * kAccStride;   // float index of the ticket word
constexpr int kStaticScratchFloats = kStaticTicketOffset + 64;

template <int VARIANT>
__global__ void __launch_bounds__(kAccThreads)
accumulate_static_kernel(MapView m, FrameGeom g, TrackIntr ti, Pose pose, const float* __restrict__ depth,
                         float* __restrict__ scratch, float* __restrict__ out29) {
    float acc[kAccTerms];
#pragma unroll
    for (int i = 0; i < kAccTerms; ++i) acc[i] = 0.f;
    const float* R = pose.R;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < ti.ws * ti.hs) {
        const int ys = i / ti.ws, xs = i - ys * ti.ws;
        const int x = xs * ti.sampling, y = ys * ti.sampling;
        const float z = depth[(size_t)y * g.w + x];
        if (!(z <= g.z_min || z >= g.z_max)) {
            const float x0 = ((float)x - ti.cx) * ti.fx_inv, y0 = ((float)y - ti.cy) * ti.fy_inv;
            const float cx_ = x0 * z, cy_ = y0 * z;
            const float px = dot3_rn(R[0], R[1], R[2], cx_, cy_, z) + pose.t[0];
            const float py = dot3_rn(R[3], R[4], R[5], cx_, cy_, z) + pose.t[1];
            const float pz = dot3_rn(R[6], R[7], R[8], cx_, cy_, z) + pose.t[2];
            float phi, g0, g1, g2;
            unsigned long long ck = kEmptyKey;
            int cs = -1;
            const float w0 = Lookup<VARIANT>::eval(m, g, px, py, pz, phi, g0, g1, g2, ck, cs);
            if (w0 > 0.f) {
                float J[6];
                J[0] = g0; J[1] = g1; J[2] = g2;
                J[3] = py * g2 - pz * g1;
                J[4] = pz * g0 - px * g2;
                J[5] = px * g1 - py * g0;
                int o = 6;
#pragma unroll
                for (int a = 0; a < 6; ++a) {
                    acc[a] = phi * J[a];
#pragma unroll
                    for (int b = a; b < 6; ++b) acc[o++] = J[a] * J[b];
                }
                acc[27] = phi * phi;
                acc[28] = 1.f;
            }
        }
    }
#pragma unroll
    for (int t = 0; t < kAccTerms; ++t) {
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) acc[t] += __shfl_xor_sync(0xffffffffu, acc[t], off);
    }
    __shared__ float sm[kAccThreads / 32][kAccStride];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) {
#pragma unroll
        for (int t = 0; t < kAccTerms; ++t) sm[warp][t] = acc[t];
    }
    __syncthreads();
    if (threadIdx.x < kAccTerms) {
        float s = 0.f;
#pragma unroll
        for (int wq = 0; wq < kAccThreads / 32; ++wq) s += sm[wq][threadIdx.x];
        scratch[(size_t)blockIdx.x * kAccStride + threadIdx.x] = s;
    }
    __shared__ bool is_last;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned int* ticket = reinterpret_cast<unsigned int*>(scratch + kStaticTicketOffset);
        is_last = (atomicAdd(ticket, 1u) == gridDim.x - 1);  // the host clears the ticket before every launch
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    for (int term = warp; term < kAccTerms; term += kAccThreads / 32) {
        float s = 0.f;
        for (int b = lane; b < (int)gridDim.x; b += 32) s += __ldcg(scratch + (size_t)b * kAccStride + term);
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) s += __shfl_xor_sync(0xffffffffu, s, off);
        if (lane == 0) out29[term] = s;
    }
}

// ---------------------------------------------------------------------------------------------------------
// Map iteration: count / compact voxels with weight > 0.
template <int VARIANT>
__global__ void __launch_bounds__(256)
count_kernel(MapView m, unsigned long long* n_voxels, unsigned long long* n_blocks) {
    // one warp per slot
    const unsigned int warps_per_block = blockDim.x >> 5;
    const unsigned int lane = threadIdx.x & 31;
    unsigned long long cnt = 0, blk = 0;
    for (unsigned int slot = blockIdx.x * warps_per_block + (threadIdx.x >> 5); slot <= m.cap_mask;
         slot += gridDim.x * warps_per_block) {
        const unsigned long long key = m.keys[slot];
        if (key == kEmptyKey) continue;
        if (m.world > 1 && block_owner(key, m.world) != m.rank) continue;  // a halo block (sharded base-sdf): not reported
        if (lane == 0) ++blk;
        for (int v = lane; v < kBlockVoxels; v += 32) {
            const size_t addr = (size_t)slot * kBlockVoxels + v;
            const float W = VARIANT == 0 ? m.accW[addr] : m.acc2[addr].x;
            cnt += W > 0.f ? 1 : 0;
        }
    }
    for (int off = 16; off > 0; off >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, off);
    if (lane == 0) {
        if (cnt) atomicAdd(n_voxels, cnt);
        if (blk) atomicAdd(n_blocks, blk);
    }
}

template <int VARIANT>
__global__ void __launch_bounds__(256)
download_kernel(MapView m, unsigned long long* cursor, long long max_n, int* __restrict__ keys,
                float* __restrict__ dist, float* __restrict__ weight, float* __restrict__ grad) {
    const unsigned int warps_per_block = blockDim.x >> 5;
    const unsigned int lane = threadIdx.x & 31;
    for (unsigned int slot = blockIdx.x * warps_per_block + (threadIdx.x >> 5); slot <= m.cap_mask;
         slot += gridDim.x * warps_per_block) {
        const unsigned long long key = m.keys[slot];
        if (key == kEmptyKey) continue;
        if (m.world > 1 && block_owner(key, m.world) != m.rank) continue;  // a halo block (sharded base-sdf): not reported
        int bx, by, bz;
        unpack_block(key, bx, by, bz);
        for (int v = lane; v < kBlockVoxels; v += 32) {
            const size_t addr = (size_t)slot * kBlockVoxels + v;
            float W, S, gx = 0.f, gy = 0.f, gz = 0.f;
            if (VARIANT == 0) { W = m.accW[addr]; const float4 a = m.acc4[addr]; S = a.x; gx = a.y; gy = a.z; gz = a.w; }
            else { const float2 a = m.acc2[addr]; W = a.x; S = a.y; }
            const bool live = W > 0.f;
            const unsigned int mask = __ballot_sync(0xffffffffu, live);
            if (!mask) continue;
            unsigned long long base = 0;
            if (lane == 0) base = atomicAdd(cursor, (unsigned long long)__popc(mask));
            base = __shfl_sync(0xffffffffu, base, 0);
            if (live) {
                const long long o = (long long)base + __popc(mask & ((1u << lane) - 1));
                if (o < max_n) {
                    keys[3 * o] = (bx << 3) | (v & 7);
                    keys[3 * o + 1] = (by << 3) | ((v >> 3) & 7);
                    keys[3 * o + 2] = (bz << 3) | (v >> 6);
                    dist[o] = S / W;
                    weight[o] = W;
                    grad[3 * o] = gx; grad[3 * o + 1] = gy; grad[3 * o + 2] = gz;
                }
            }
        }
    }
}

// Visibility log -> (voxel key, frame) pairs: entry = voxel address << 24 | frame; the address is slot * 512 + local index.
__global__ void __launch_bounds__(256)
vis_decode_kernel(MapView m, long long n, int* __restrict__ keys, int* __restrict__ frames) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const unsigned long long e = m.vis_log[i];
    const unsigned long long vaddr = e >> 24;
    const unsigned int slot = (unsigned int)(vaddr >> 9), v = (unsigned int)(vaddr & 511);
    int bx, by, bz;
    unpack_block(m.keys[slot], bx, by, bz);
    keys[3 * i] = (bx << 3) | (v & 7);
    keys[3 * i + 1] = (by << 3) | ((v >> 3) & 7);
    keys[3 * i + 2] = (bz << 3) | (v >> 6);
    frames[i] = (int)(e & 0xFFFFFFull);
}

// getSdf(Vec3i) (MapGradPixelSdf.h:127-129, MapPixelSdf.h:123-125): the voxels at the listed keys; absent or empty -> zeros.
template <int VARIANT>
__global__ void __launch_bounds__(256)
get_voxels_kernel(MapView m, long long n, const int* __restrict__ keys, float* __restrict__ dist, float* __restrict__ weight,
                  float* __restrict__ grad) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int vx = keys[3 * i], vy = keys[3 * i + 1], vz = keys[3 * i + 2];
    float W = 0.f, S = 0.f, gx = 0.f, gy = 0.f, gz = 0.f;
    const int slot = map_find(m, block_key_of(vx, vy, vz));
    if (slot >= 0) {
        const size_t addr = (size_t)slot * kBlockVoxels + local_index(vx, vy, vz);
        if (VARIANT == 0) { W = m.accW[addr]; const float4 a = m.acc4[addr]; S = a.x; gx = a.y; gy = a.z; gz = a.w; }
        else { const float2 a = m.acc2[addr]; W = a.x; S = a.y; }
    }
    const bool live = W > 0.f;
    dist[i] = live ? S / W : 0.f;
    weight[i] = live ? W : 0.f;
    grad[3 * i] = live ? gx : 0.f; grad[3 * i + 1] = live ? gy : 0.f; grad[3 * i + 2] = live ? gz : 0.f;
}

// Growing the table (gsdf_grow): every block of the old table is re-inserted into the new one and its 512 voxels copied.
// One CTA per old slot; slot_map[old slot] = new slot (for the visibility log, whose entries hold voxel addresses).
template <int VARIANT>
__global__ void __launch_bounds__(128)
rehash_kernel(MapView old_m, MapView new_m, Counters* ctr, int* __restrict__ slot_map) {
    __shared__ int s_slot;
    for (unsigned int slot = blockIdx.x; slot <= old_m.cap_mask; slot += gridDim.x) {
        const unsigned long long key = old_m.keys[slot];
        if (key == kEmptyKey) continue;
        __syncthreads();
        if (threadIdx.x == 0) {
            // (ctr->blocks already counts this block: the insert below must not count it again)
            uint32_t ns = hash_key(key) & new_m.cap_mask;
            for (;;) {
                const unsigned long long k = atomicCAS(new_m.keys + ns, kEmptyKey, key);
                if (k == kEmptyKey || k == key) break;
                ns = (ns + 1) & new_m.cap_mask;
            }
            s_slot = (int)ns;
            if (slot_map) slot_map[slot] = (int)ns;
        }
        __syncthreads();
        const size_t src = (size_t)slot * kBlockVoxels, dst = (size_t)s_slot * kBlockVoxels;
        for (int v = threadIdx.x; v < kBlockVoxels; v += blockDim.x) {
            if (VARIANT == 0) { new_m.acc4[dst + v] = old_m.acc4[src + v]; new_m.accW[dst + v] = old_m.accW[src + v]; }
            else new_m.acc2[dst + v] = old_m.acc2[src + v];
            if (old_m.vis_stamp) new_m.vis_stamp[dst + v] = old_m.vis_stamp[src + v];
        }
    }
    (void)ctr;
}

__global__ void __launch_bounds__(256)
vis_remap_kernel(unsigned long long* __restrict__ log, long long n, const int* __restrict__ slot_map) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const unsigned long long e = log[i], vaddr = e >> 24;
    const unsigned long long nv = ((unsigned long long)slot_map[vaddr >> 9] << 9) | (vaddr & 511);
    log[i] = (nv << 24) | (e & 0xFFFFFFull);
}

// Restore voxels from a download (checkpoint / resume; no counterpart in the reference): one thread per listed voxel,
// the block is found or inserted, the voxel's accumulators are SET to {dist * W, grad, W}.  A sharded map keeps only
// what its ownership rule assigns to it (including base-sdf halo copies), so every rank may be handed the whole list.
template <int VARIANT>
__global__ void __launch_bounds__(256)
upload_kernel(MapView m, long long n, const int* __restrict__ keys, const float* __restrict__ dist,
              const float* __restrict__ weight, const float* __restrict__ grad, Counters* ctr) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float W = weight[i];
    if (!(W > 0.f)) return;
    const int vx = keys[3 * i], vy = keys[3 * i + 1], vz = keys[3 * i + 2];
    const unsigned long long key = block_key_of(vx, vy, vz);
    if (!shard_keeps<VARIANT>(m.world, m.rank, key, vx, vy, vz)) return;
    const int slot = map_find_or_insert(m, key, ctr);
    if (slot < 0) { atomicAdd(&ctr->overflow, 1u); return; }
    const size_t addr = (size_t)slot * kBlockVoxels + local_index(vx, vy, vz);
    const float S = __fmul_rn(dist[i], W);
    if (VARIANT == 0) {
        m.acc4[addr] = make_float4(S, grad[3 * i], grad[3 * i + 1], grad[3 * i + 2]);
        m.accW[addr] = W;
    } else {
        m.acc2[addr] = make_float2(W, S);
    }
}

}  // namespace gsdf
