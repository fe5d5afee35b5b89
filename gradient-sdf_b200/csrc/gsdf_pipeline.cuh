// Frame pipeline kernels: everything a frame needs runs on the streams without a host round trip.
//
//   normals_compact_kernel                    the frame-only work in one launch per frame (side stream): deterministic
//                                             single-pass compaction of the z-valid pixels into vlist = {(y<<16)|x, z}
//                                             in tile order (so that every lane of every warp downstream has work and a
//                                             warp's pixels are neighbours) + FALS normals of the tiles that hold any;
//   track_kernel<V, NT, STAMPS>               main stream, persistent cooperative launch over up to kMaxBatch frames
//                                             (NT threads per CTA; STAMPS = the twin with debug stamps compiled in);
//                                             per frame the WHOLE Gauss-Newton loop of
//                                             RigidPointOptimizer::optimize_sampled (RigidPointOptimizerOmp.cpp:44-106):
//                                             per-pixel lookup + Jacobian, warp -> CTA -> grid reduction of the 29 sums
//                                             (across GPUs too: allreduce over NVLink peer memory), and — redundantly in
//                                             every CTA, on identical data, so no broadcast is needed — the 6x6 LLT
//                                             solve, the stopping test and pose <- exp(-xi) * pose; then the fusion of
//                                             the frame iff converged (main_scan_3d.cpp:261), then the next frame;
//   compact_tiles_kernel, normals_tile_kernel the two halves of the frame-only work as separate launches (fine-grained
//                                             entry points, per-kernel profiling);
//   fuse_list_kernel<V>                       fusion over vlist as its own launch (gsdf_integrate*, profiling);
//   accumulate_list_kernel<V> + solve_kernel  the Gauss-Newton loop split into per-iteration launches, for sharded maps
//                                             without a peer connection: an NCCL all-reduce sits between the two.
#pragma once

#include "gsdf_kernels.cuh"
#include "gsdf_pose.cuh"

namespace gsdf {

__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
// debug frame log (gsdf_debug_frame_log): {first CTA entry, last CTA exit} of a kernel, into log[0], log[1]
__device__ __forceinline__ void flog_enter(unsigned long long* log) { if (log) atomicMin(log, global_ns()); }
__device__ __forceinline__ void flog_exit(unsigned long long* log) { if (log) atomicMax(log + 1, global_ns()); }


// Device-resident tracker / frame state (RigidOptimizer::pose_ and friends live here between calls).
struct FrameState {
    float pose7[7];             // tx ty tz qx qy qz qw, unit quaternion
    int converged;              // result of the last optimize
    int iters;                  // k at convergence, or max_iters
    int n_valid;                // compacted z-valid pixels of the current frame
    int done;                   // per-iteration path: loop finished (converged or exhausted)
    int k;                      // per-iteration path: iteration counter
    unsigned long long track_iters;   // cumulative Gauss-Newton iterations
    unsigned long long track_valid;   // cumulative contributing pixels (the reference's `counter`)
    float sums[32];             // last reduced sums
};

// What gsdf_scan_collect reports for a frame.  Written to device memory by the LAST CTA of the frame's last kernel and
// copied to pinned host memory on the result stream (see FrameSlot in gsdf_api.cu).
struct ResultSnapshot {
    FrameState st;
    unsigned long long points, updates;  // cumulative
    unsigned int blocks, overflow;
    unsigned int peer_error;             // sharded tracker: a peer did not show up in the exchange
    unsigned int barrier_timeout;        // the tracker's grid barrier gave up (its CTAs were not all resident)
};

// executed by one whole CTA (>= 96 threads) after every other CTA's writes are visible
__device__ __forceinline__ void write_snapshot(const FrameState* st, const Counters* ctr, ResultSnapshot* out,
                                               const int* peer_error = nullptr) {
    // (the host reads the snapshot only after the event that follows this launch: no system-scope fence is needed)
    static_assert(kCounterSlots == 64, "two slots per lane");
    const int t = threadIdx.x;
    constexpr int kWords = (int)(sizeof(FrameState) / sizeof(int));
    if (t >= 32 && t < 32 + kWords) reinterpret_cast<int*>(&out->st)[t - 32] = __ldcg(reinterpret_cast<const int*>(st) + t - 32);
    if (t < 32) {
        unsigned long long p = __ldcg(&ctr->slot[t].points) + __ldcg(&ctr->slot[t + 32].points);
        unsigned long long u = __ldcg(&ctr->slot[t].updates) + __ldcg(&ctr->slot[t + 32].updates);
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            p += __shfl_xor_sync(0xffffffffu, p, off);
            u += __shfl_xor_sync(0xffffffffu, u, off);
        }
        if (t == 0) {
            out->points = p; out->updates = u;
            out->blocks = __ldcg(&ctr->blocks); out->overflow = __ldcg(&ctr->overflow);
            out->peer_error = peer_error ? (unsigned int)__ldcg(peer_error) : 0u;
            out->barrier_timeout = __ldcg(&ctr->barrier_timeout);
        }
    }
}


constexpr int kTileW = 8, kTileH = 4;           // one warp <-> one 8x4 pixel tile
constexpr int kTilesPerWarp = 8;
constexpr int kCompactThreads = 256;
constexpr int kTilesPerBlock = (kCompactThreads / 32) * kTilesPerWarp;  // 64 tiles = 2048 pixels per CTA

__device__ __forceinline__ bool z_valid(float z, float z_min, float z_max) { return !(z <= z_min || z >= z_max); }

// Single-pass, deterministic compaction of the z-valid pixels into vlist = {(y << 16) | x, z}, in TILE order:
// 32 consecutive entries come from (at most a few) 8x4 pixel tiles, so the lanes of a warp downstream look at a
// compact 3-D neighbourhood — nearby voxels for the tracker's gathers, and many lanes hitting the SAME voxel in
// fusion, which the warp then merges into one reduction.
// The exclusive prefix over CTAs is a chained look-back: CTA b publishes {epoch, count} in flags[b] and sums the
// flags of all CTAs < b, spinning until they carry this launch's epoch (lower-numbered CTAs are dispatched first and
// wait only on even lower ones, so the chain always makes progress; flags need no clearing between launches).
// One segment = (NWARPS * kTilesPerWarp) tiles handled by one CTA of NWARPS warps.  The list order is the global tile
// order whatever the segment size, so the stand-alone kernel (8 warps) and the frame kernel (16 warps) agree.
constexpr int kFlagShift = 13;  // a segment holds at most 16 warps * 8 tiles * 32 = 4096 < 2^13 pixels
template <int NWARPS>
__device__ __forceinline__ void compact_segment(const float* __restrict__ depth, int w, int h, float z_min, float z_max,
                                                unsigned int* __restrict__ flags, unsigned int epoch,
                                                int2* __restrict__ vlist, int* __restrict__ n_valid_out, int seg,
                                                int nseg, int* s_warp /* [NWARPS] */, int* s_base) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int tiles_x = (w + kTileW - 1) / kTileW, tiles_y = (h + kTileH - 1) / kTileH, ntiles = tiles_x * tiles_y;
    const int tx = lane & 7, ty = lane >> 3;
    float z[kTilesPerWarp];
    int yx[kTilesPerWarp];
    unsigned int mask[kTilesPerWarp];
    int wtot = 0;
#pragma unroll
    for (int i = 0; i < kTilesPerWarp; ++i) {
        const int tile = (seg * NWARPS + warp) * kTilesPerWarp + i;
        const int tyi = tile / tiles_x, txi = tile - tyi * tiles_x;
        const int x = txi * kTileW + tx, y = tyi * kTileH + ty;
        const bool in = tile < ntiles && x < w && y < h;
        z[i] = in ? __ldg(depth + (size_t)y * w + x) : 0.f;
        yx[i] = (y << 16) | x;
        mask[i] = __ballot_sync(0xffffffffu, in && z_valid(z[i], z_min, z_max));
        wtot += __popc(mask[i]);
    }
    if (lane == 0) s_warp[warp] = wtot;
    __syncthreads();
    int before = 0, total = 0;
#pragma unroll
    for (int q = 0; q < NWARPS; ++q) { const int c = s_warp[q]; before += q < warp ? c : 0; total += c; }
    if (threadIdx.x == 0) {
        __threadfence();
        atomicExch(flags + seg, (epoch << kFlagShift) | (unsigned)total);
    }
    if (warp == 0) {
        int s = 0;
        for (int b = lane; b < seg; b += 32) {
            unsigned int f;
            do { f = *reinterpret_cast<volatile unsigned int*>(flags + b); } while ((f >> kFlagShift) != epoch);
            s += (int)(f & ((1u << kFlagShift) - 1));
        }
        s = __reduce_add_sync(0xffffffffu, s);
        if (lane == 0) *s_base = s;
    }
    __syncthreads();
    int off = *s_base + before;
#pragma unroll
    for (int i = 0; i < kTilesPerWarp; ++i) {
        if (mask[i] & (1u << lane)) vlist[off + __popc(mask[i] & ((1u << lane) - 1))] = make_int2(yx[i], __float_as_int(z[i]));
        off += __popc(mask[i]);
    }
    if (seg == nseg - 1 && threadIdx.x == 0) *n_valid_out = *s_base + total;
    __syncthreads();  // s_warp / s_base are reused by the caller's next segment
}

__global__ void __launch_bounds__(kCompactThreads)
compact_tiles_kernel(const float* __restrict__ depth, int w, int h, float z_min, float z_max,
                     unsigned int* __restrict__ flags, unsigned int epoch, int2* __restrict__ vlist,
                     int* __restrict__ n_valid_out) {
    __shared__ int s_warp[kCompactThreads / 32];
    __shared__ int s_base;
    compact_segment<kCompactThreads / 32>(depth, w, h, z_min, z_max, flags, epoch, vlist, n_valid_out, blockIdx.x,
                                          gridDim.x, s_warp, &s_base);
}

// ---------------------------------------------------------------------------------------------------------
// Per-pixel tracker term: residual, Jacobian, 29 accumulations (RigidPointOptimizerOmp.cpp:72-90).
template <int VARIANT>
__device__ __forceinline__ void track_pixel(const MapView& m, const FrameGeom& g, const TrackIntr& ti, const float* R,
                                            const float* t, int yx, float z, float* acc, unsigned long long& ck,
                                            int& cs) {
    const int y = yx >> 16, x = yx & 0xFFFF;
    const float x0 = ((float)x - ti.cx) * ti.fx_inv, y0 = ((float)y - ti.cy) * ti.fy_inv;
    const float cx_ = x0 * z, cy_ = y0 * z;
    const float px = dot3_rn(R[0], R[1], R[2], cx_, cy_, z) + t[0];
    const float py = dot3_rn(R[3], R[4], R[5], cx_, cy_, z) + t[1];
    const float pz = dot3_rn(R[6], R[7], R[8], cx_, cy_, z) + t[2];
    float phi, g0, g1, g2;
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
            acc[a] += phi * J[a];
#pragma unroll
            for (int b = a; b < 6; ++b) acc[o++] += J[a] * J[b];
        }
        acc[27] += phi * phi;
        acc[28] += 1.f;
    }
}

// Key of the block a pixel's lookup probes first (Lookup<V>::eval: the voxel's block / the block of the cell's floor corner),
// computed exactly as track_pixel_terms + eval compute it.
template <int VARIANT>
__device__ __forceinline__ unsigned long long first_block_key(const MapView& m, const FrameGeom& g, const TrackIntr& ti,
                                                              const float* R, const float* t, int yx, float z) {
    const int y = yx >> 16, x = yx & 0xFFFF;
    const float x0 = ((float)x - ti.cx) * ti.fx_inv, y0 = ((float)y - ti.cy) * ti.fy_inv;
    const float cx_ = x0 * z, cy_ = y0 * z;
    const float px = dot3_rn(R[0], R[1], R[2], cx_, cy_, z) + t[0];
    const float py = dot3_rn(R[3], R[4], R[5], cx_, cy_, z) + t[1];
    const float pz = dot3_rn(R[6], R[7], R[8], cx_, cy_, z) + t[2];
    if (VARIANT == 0) return block_key_of(float2vox1(g.vs_inv, px), float2vox1(g.vs_inv, py), float2vox1(g.vs_inv, pz));
    return block_key_of((int)floorf(__fmul_rn(g.vs_inv, px)), (int)floorf(__fmul_rn(g.vs_inv, py)),
                        (int)floorf(__fmul_rn(g.vs_inv, pz)));
}

// Same terms for the persistent tracker, which reduces every pixel's 29 products across the warp at once (the reduce-scatter
// butterfly below) instead of carrying 29 accumulators per thread from pixel to pixel: v[0..31] = this pixel's products
// (zeros if it does not contribute).
template <int VARIANT>
__device__ __forceinline__ bool track_pixel_terms(const MapView& m, const FrameGeom& g, const TrackIntr& ti, const float* R,
                                                  const float* t, bool has, int yx, float z, float* v,
                                                  unsigned long long& ck, int& cs) {
    const int y = yx >> 16, x = yx & 0xFFFF;
    const float x0 = ((float)x - ti.cx) * ti.fx_inv, y0 = ((float)y - ti.cy) * ti.fy_inv;
    const float cx_ = x0 * z, cy_ = y0 * z;
    const float px = dot3_rn(R[0], R[1], R[2], cx_, cy_, z) + t[0];
    const float py = dot3_rn(R[3], R[4], R[5], cx_, cy_, z) + t[1];
    const float pz = dot3_rn(R[6], R[7], R[8], cx_, cy_, z) + t[2];
    float phi = 0.f, g0 = 0.f, g1 = 0.f, g2 = 0.f;
    float w0 = 0.f;
    if (has) w0 = Lookup<VARIANT>::eval(m, g, px, py, pz, phi, g0, g1, g2, ck, cs);
    const bool on = w0 > 0.f;
    if (!on) { phi = 0.f; g0 = g1 = g2 = 0.f; }
    float J[6];
    J[0] = g0; J[1] = g1; J[2] = g2;
    J[3] = py * g2 - pz * g1;
    J[4] = pz * g0 - px * g2;
    J[5] = px * g1 - py * g0;
    int o = 6;
#pragma unroll
    for (int a = 0; a < 6; ++a) {
        v[a] = phi * J[a];
#pragma unroll
        for (int b = a; b < 6; ++b) v[o++] = J[a] * J[b];
    }
    v[27] = phi * phi;
    v[28] = on ? 1.f : 0.f;
    v[29] = v[30] = v[31] = 0.f;
    return on;
}

// The persistent tracker runs either two 384-thread CTAs or one 768-thread CTA per SM (same threads per SM; the second
// halves the rows of the grid-level exchange below).  64 registers per thread either way, so that one 512-thread,
// 32-register CTA of the frame-only kernel fits beside it (shadow mode).
constexpr int kTrackThreadsMax = 768;
constexpr int kMorePixels = 2;  // pixels per thread beyond the first whose probed block is remembered between passes

struct TrackParams {
    int max_iters;
    float thr_sq;
    float damping;
    unsigned int poll_ns;  // pause between two polls of the grid-level exchange (0 = spin)
};

// Warp stage of the tree as a reduce-scatter butterfly: instead of 29 independent 5-step shuffle reductions
// (145 SHFL per warp) the 32 (29 + 3 zero) values are halved at every step — lanes whose `off` bit is clear keep
// the lower half and receive the partner's lower half, the others the upper half — so 16+8+4+2+1 = 31 SHFL suffice
// and lane L ends up holding the warp total of term L.  Fixed pairing => bit-reproducible.
__device__ __forceinline__ float warp_reduce_scatter_32(float* v /* [32], destroyed */) {
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int half = 16; half >= 1; half >>= 1) {
        const bool up = (lane & half) != 0;
#pragma unroll
        for (int i = 0; i < half; ++i) {
            const float lo = v[i], hi = v[i + half];
            const float send = up ? lo : hi, keep = up ? hi : lo;
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, half);
        }
    }
    return v[0];
}

// block-level stage: acc[29] per thread -> 29 sums per block in dst[0..28] (dst may be global or shared)
template <int NWARPS>
__device__ __forceinline__ void block_reduce_29(const float* acc, float (*sm)[kAccStride], float* dst) {
    float v[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = i < kAccTerms ? acc[i] : 0.f;
    const float mine = warp_reduce_scatter_32(v);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    sm[warp][lane] = mine;
    __syncthreads();
    if (threadIdx.x < kAccTerms) {
        float s = 0.f;
#pragma unroll
        for (int wq = 0; wq < NWARPS; ++wq) s += sm[wq][threadIdx.x];
        dst[threadIdx.x] = s;
    }
}

// grid-level stage, executed identically by whoever calls it: tot[term] = sum over blocks, fixed grouping.
// Thread (grp, t) sums blocks grp, grp+NWARPS, ... of term t with all its loads in flight at once (coalesced
// 128 B rows), then the NWARPS group sums are added in order.
template <int NWARPS>
__device__ __forceinline__ void grid_reduce_29(const float* partials, int nblocks, float (*sm)[kAccStride], float* tot) {
    const int t = threadIdx.x & 31, grp = threadIdx.x >> 5;
    constexpr int kMaxJ = 12;  // up to 12 * NWARPS blocks with every load issued up front
    float s = 0.f;
    for (int b0 = grp; b0 < nblocks; b0 += NWARPS * kMaxJ) {
        float v[kMaxJ];
#pragma unroll
        for (int j = 0; j < kMaxJ; ++j) {
            const int b = b0 + j * NWARPS;
            v[j] = (b < nblocks) ? __ldcg(partials + (size_t)b * kAccStride + t) : 0.f;
        }
#pragma unroll
        for (int j = 0; j < kMaxJ; ++j) s += v[j];
    }
    __syncthreads();  // sm may still be read by the block stage of slower warps
    sm[grp][t] = s;
    __syncthreads();
    if (threadIdx.x < kAccTerms) {
        float r = 0.f;
#pragma unroll
        for (int wq = 0; wq < NWARPS; ++wq) r += sm[wq][threadIdx.x];
        tot[threadIdx.x] = r;
    }
}

// Grid-level stage of the persistent tracker: an all-to-all exchange of the CTA sums through tagged words, in place of
// "write row, grid barrier, read all rows".  CTA b publishes its 29 sums as 64-bit words {pass tag, float bits} in
// rows[tag & 1][b][0..28] (one store each, single-copy atomic: no fence, no separate flag, no arrival counter), and
// every CTA polls ALL rows until each word carries this pass's tag, then adds them in row order: identical totals in
// every CTA, bit-reproducible, ONE L2 round trip after the slowest CTA has published.  Two buffers suffice inside a
// launch (a CTA can publish pass k+2 only after it has seen everyone's words of pass k+1, which they publish after they
// have finished reading pass k); across launches the kernel boundary orders reuse, and tags never repeat (the host hands
// every launch a fresh range).  A row that does not show up within 2 s sets barrier_timeout instead of hanging.
__device__ __forceinline__ void st_word(unsigned long long* p, unsigned long long v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_word(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

// CTA stage: lane L of every warp holds the warp's sum of term L -> the CTA's sums, returned in threads 0..28
template <int NWARPS>
__device__ __forceinline__ float block_sum_29(float mine, float (*sm)[kAccStride]) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    sm[warp][lane] = mine;
    __syncthreads();
    float s = 0.f;
    if (threadIdx.x < kAccTerms) {
#pragma unroll
        for (int wq = 0; wq < NWARPS; ++wq) s += sm[wq][threadIdx.x];
    }
    return s;
}

struct NoWork { __device__ __forceinline__ void operator()() const {} };

// `between`: work that does not depend on the exchange, run after this CTA has published its sums and before it starts to
// poll (the words of the slowest CTA are at least one L2 round trip away).
template <int NWARPS, typename Between = NoWork>
__device__ __forceinline__ void exchange_29(unsigned long long* rows /* [2][gridDim.x][32] */, unsigned int tag, float mine,
                                            float (*sm)[kAccStride] /* not the CTA stage's array */, float* tot,
                                            Counters* ctr, unsigned int poll_ns, Between between = Between()) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nrows = gridDim.x;
    unsigned long long* buf = rows + (size_t)(tag & 1u) * nrows * kAccStride;
    const unsigned long long tg = (unsigned long long)tag << 32;
    if (threadIdx.x < kAccTerms) st_word(buf + (size_t)blockIdx.x * kAccStride + threadIdx.x, tg | __float_as_uint(mine));
    between();
    // warp w polls rows w, w + NWARPS, ...; lane = term.  All loads of a round are in flight together.
    constexpr int kJ = 8;
    float s = 0.f;
    unsigned long long t0 = 0;
    for (int r0 = warp; r0 < nrows; r0 += NWARPS * kJ) {
        unsigned long long w[kJ];
        unsigned int pend = 0;
#pragma unroll
        for (int j = 0; j < kJ; ++j)
            if (lane < kAccTerms && r0 + j * NWARPS < nrows) pend |= 1u << j;
        unsigned int spins = 0;
        while (pend) {
#pragma unroll
            for (int j = 0; j < kJ; ++j)
                if (pend & (1u << j)) w[j] = ld_word(buf + (size_t)(r0 + j * NWARPS) * kAccStride + lane);
#pragma unroll
            for (int j = 0; j < kJ; ++j)
                if ((pend & (1u << j)) && (w[j] >> 32 << 32) == tg) pend &= ~(1u << j);
            if (pend) {
                // optional pause between polls (GSDF_POLL_NS; 0..800 ns measured: no effect on the frame rate, so the
                // polling traffic is not what couples the tracker to the frame-only kernel beside it)
                if (poll_ns) __nanosleep(poll_ns);
                if ((++spins & 1023u) == 0) {
                    const unsigned long long now = global_ns();
                    if (t0 == 0) t0 = now;
                    else if (now - t0 > 2000000000ull) { ctr->barrier_timeout = 1u; break; }
                }
            }
        }
#pragma unroll
        for (int j = 0; j < kJ; ++j)
            if (lane < kAccTerms && r0 + j * NWARPS < nrows) s += __uint_as_float((unsigned int)w[j]);
    }
    sm[warp][lane] = s;
    __syncthreads();
    if (threadIdx.x < kAccTerms) {
        float r = 0.f;
#pragma unroll
        for (int wq = 0; wq < NWARPS; ++wq) r += sm[wq][threadIdx.x];
        tot[threadIdx.x] = r;
    }
}

// ---------------------------------------------------------------------------------------------------------
// Fusion over the compacted list: one lane per pixel walks its 2*factor+1 ray samples (MapGradPixelSdfOmp.cpp:103);
// at every sample the warp merges the lanes that hit the SAME voxel (list entries of a warp come from the same
// 8x4-pixel tiles, and a 1 cm voxel spans 3-5 pixels) and only the group leader issues the vector reductions.
// Measured on B200 (profiles/micro/red_bench.cu): global RED throughput is ~100 G ops/s regardless of how the
// duplicates are spread, so the number of REDs — not the bytes — bounds this kernel; merging cuts them 3-6x.
// pose_src: 0 = pose by value, 1 = pose from the device state, 2 = pose from the device state and only if the last
// optimize converged (main_scan_3d.cpp:261).
// The members of a group (lanes with the same voxel address) form a list in lane order; the list is summed by
// pointer jumping (v += v[next], next = next[next]): ceil(log2(group size)) rounds of NV + 1 shuffles, where a
// leader that collects its members one by one needs group size - 1 rounds — and at 1 cm voxels a voxel spans ~12
// pixels of an 8x4 tile.  After the last round the first member (the leader) holds the group's sum.
template <int NV>
__device__ __forceinline__ void warp_merge_by_key(unsigned long long key, bool narrow, bool has, float* v, bool& leader) {
    const int lane = threadIdx.x & 31;
    // lanes without a sample get a key no one else has; narrow: every key of this map fits in 32 bits
    const unsigned int peers = narrow ? __match_any_sync(0xffffffffu, has ? (unsigned int)key : (0xFFFFFFE0u | (unsigned)lane))
                                      : __match_any_sync(0xffffffffu, has ? key : (0xFFFFFFFF00000000ull | (unsigned)lane));
    leader = has && ((int)__ffs(peers) - 1 == lane);
    const unsigned int above = peers & (0xFFFFFFFEu << lane);
    int nxt = above ? __ffs(above) - 1 : -1;
    while (__any_sync(0xffffffffu, nxt >= 0)) {
        const int src = nxt >= 0 ? nxt : lane;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const float o = __shfl_sync(0xffffffffu, v[i], src);
            if (nxt >= 0) v[i] += o;
        }
        nxt = __shfl_sync(0xffffffffu, nxt, src);
    }
}

// Work decomposition: the unit of work is (chunk of 32 consecutive list entries, ONE ray sample); the 2*factor+1
// samples of a pixel are therefore spread over several warps instead of forming one long dependent instruction
// stream per lane (with ~2 k warps of 11 serial samples each the kernel was latency-bound at 25 us), and every warp
// gets the same number of units (+-1).  Called by fuse_list_kernel and by the tail of track_kernel (fused launch).
// COHERENT: the list and the normals were produced earlier in the SAME launch (frame kernel) -> L2 loads, not the
// non-coherent read-only path.
// STAMPS: debug stamps per warp (track_kernel<.., true>, launched only while gsdf_debug_track_timers is armed: the stamp
// tests alone are ~21 of the ~250 instructions of a unit).
// The calling warp's share [u0, u1) of the frame's units (unit = 32-pixel chunk of the list x one ray sample, chunk-major).
__device__ __forceinline__ void fuse_range(int n_valid, int S, int& u0, int& u1) {
    const int units = ((n_valid + 31) >> 5) * S;  // < 2^31: checked by gsdf_create
    const int warps_total = (gridDim.x * blockDim.x) >> 5;
    const int wid = (threadIdx.x >> 5) * gridDim.x + blockIdx.x;  // warp ranges are interleaved over the CTAs
    // warp w takes units [w*q + min(w, r), ...): q or q + 1 of them (one division, the same for every warp)
    const int q = units / warps_total, r = units - q * warps_total;
    u0 = wid * q + min(wid, r);
    u1 = u0 + q + (wid < r ? 1 : 0);
}

// The pose-independent pixel data of one list entry, as fusion needs it.  track_kernel fetches the data of each warp's
// FIRST unit while the sums of the frame's first two passes travel (the list entry in the first, the planes it indexes in the
// second: one L2 round trip each, less than the exchange takes anyway) and parks it in shared memory (kPfWords words per
// thread), so that the fusion phase does not open with two dependent L2 round trips that nothing overlaps.
constexpr int kPfWords = 7;  // z, x0, y0, normal (3), hom_inv
template <bool COHERENT>
__device__ __forceinline__ void fuse_pixel_planes(const FrameGeom& g, int2 e, const float* __restrict__ nx,
                                                  const float* __restrict__ ny, const float* __restrict__ nz,
                                                  const float* __restrict__ x0, const float* __restrict__ y0,
                                                  const float* __restrict__ hom_inv, float* o) {
    const int idx = (e.x >> 16) * g.w + (e.x & 0xFFFF);
    o[0] = __int_as_float(e.y);
    o[1] = x0[idx]; o[2] = y0[idx];
    o[3] = COHERENT ? __ldcg(nx + idx) : nx[idx];
    o[4] = COHERENT ? __ldcg(ny + idx) : ny[idx];
    o[5] = COHERENT ? __ldcg(nz + idx) : nz[idx];
    o[6] = hom_inv[idx];
}

template <int VARIANT, bool COHERENT, bool STAMPS = false>
__device__ __forceinline__ void fuse_work(const MapView& m, const FrameGeom& g, const Pose& pose,
                                          const int2* __restrict__ vlist, int n_valid, const float* __restrict__ nx,
                                          const float* __restrict__ ny, const float* __restrict__ nz,
                                          const float* __restrict__ x0, const float* __restrict__ y0,
                                          const float* __restrict__ hom_inv, Counters* ctr, unsigned int& n_pts,
                                          unsigned int& n_upd, unsigned long long* fdbg = nullptr /* debug: [warp][8] stamps */,
                                          const float* pf = nullptr /* shared: [kPfWords][blockDim.x], this thread's pixel of
                                                                       the warp's first unit (fuse_pixel_planes) */) {
    const float* R = pose.R;
    const int lane = threadIdx.x & 31;
    const int S = 2 * g.factor + 1;
    // Sdf::counter_ of this update(): constant while any CTA is still fusing (it is advanced after the grid barrier /
    // by the launch's last CTA)
    const unsigned int frame_no = m.vis_stamp != nullptr ? __ldcg(&ctr->fused_frames) : 0u;
    // Work unit = (32-pixel chunk of the list, one ray sample); every warp takes an equal, CONTIGUOUS range of the
    // chunk-major unit sequence (sizes differ by at most one unit; before, whole (chunk, sample-group) items were
    // dealt round-robin and a warp had 3 or 4 of them: the 4-item CTAs finished 6 us after the others), so a warp
    // loads the pixel data of only one or two chunks.  Warp ranges are interleaved over the CTAs (warp-major).
    const bool narrow = m.cap_mask < (1u << 23) - 32u;  // voxel addresses (slot * 512 + local) below 2^32 - 2^14
    const int wid = (threadIdx.x >> 5) * gridDim.x + blockIdx.x;
    int u0, u1;
    fuse_range(n_valid, S, u0, u1);
    // debug stamps of this warp (lane 0): 0 entry, 1 first pixel data there, 2 first slot known, 3 first merge done, 4 first
    // reductions issued, 5 all units done, 6 = units | probes << 32, 7 = reductions issued by the warp
    unsigned long long* dbg = (STAMPS && fdbg && lane == 0) ? fdbg + (size_t)wid * 8 : nullptr;
    unsigned int d_probes = 0, d_reds = 0;
    bool d_first = true;
#define GSDF_FSTAMP(i) do { if (STAMPS && dbg && d_first) dbg[i] = global_ns(); } while (0)
    if (dbg) { dbg[0] = global_ns(); dbg[6] = (unsigned long long)(u1 - u0); }
    for (int u = u0; u < u1;) {
        const int chunk = u / S, k_first = u - chunk * S;
        const int k_end = min(S, k_first + (u1 - u));  // exclusive
        u += k_end - k_first;
        const int j = chunk * 32 + lane;
        bool valid = j < n_valid;
        float z = 0.f, Rxy0 = 0.f, Rxy1 = 0.f, Rxy2 = 0.f, Rn0 = 0.f, Rn1 = 0.f, Rn2 = 0.f;
        if (valid) {
            float d7[kPfWords];
            if (pf != nullptr && u - (k_end - k_first) == u0) {  // the warp's first unit: fetched during the first passes
#pragma unroll
                for (int i = 0; i < kPfWords; ++i) d7[i] = pf[i * blockDim.x + threadIdx.x];
            } else {
                const int2 e = COHERENT ? __ldcg(vlist + j) : __ldg(vlist + j);
                fuse_pixel_planes<COHERENT>(g, e, nx, ny, nz, x0, y0, hom_inv, d7);
            }
            z = d7[0];
            const float hx = d7[1], hy = d7[2], a = d7[3], b = d7[4], c = d7[5], hi = d7[6];
            const float nsq = dot3_rn(a, b, c, a, b, c);
            const float d = dot3_rn(a, b, c, hx, hy, 1.f);
            if ((double)nsq < .1) valid = false;                                       // MapGradPixelSdfOmp.cpp:97
            else if ((double)__fmul_rn(__fmul_rn(d, d), hi) < .25) valid = false;      // :100
            Rxy0 = dot3_rn(R[0], R[1], R[2], hx, hy, 1.f);                             // :93
            Rxy1 = dot3_rn(R[3], R[4], R[5], hx, hy, 1.f);
            Rxy2 = dot3_rn(R[6], R[7], R[8], hx, hy, 1.f);
            if (VARIANT == 0) {
                Rn0 = dot3_rn(R[0], R[1], R[2], a, b, c);                              // :95
                Rn1 = dot3_rn(R[3], R[4], R[5], a, b, c);
                Rn2 = dot3_rn(R[6], R[7], R[8], a, b, c);
            }
        }
        n_pts += (valid && k_first == 0) ? 1u : 0u;
        GSDF_FSTAMP(1);
        // block of the previous committed sample of this ray and its slot.  Consecutive samples are one voxel apart and a
        // segment has at most 2*factor+1 of them, so the low 10 bits of each block coordinate identify the block here:
        // one 30-bit compare per sample instead of packing and comparing the 63-bit key (which is built only for a probe)
        unsigned int last_sig = 0xFFFFFFFFu;
        int slot = -1;
        for (int kk = k_first - g.factor; kk < k_end - g.factor; ++kk) {               // :103
            bool has = false;
            unsigned long long vaddr = 0;
            float v[VARIANT == 0 ? 5 : 2];
            float wq = 0.f, sq = 0.f;  // this lane's weight and sdf
            if (valid) {
                const float s = __fadd_rn(z, __fmul_rn((float)kk, g.vs));
                const float px = __fadd_rn(__fmul_rn(s, Rxy0), pose.t[0]);             // :105
                const float py = __fadd_rn(__fmul_rn(s, Rxy1), pose.t[1]);
                const float pz = __fadd_rn(__fmul_rn(s, Rxy2), pose.t[2]);
                const int vx = float2vox1(g.vs_inv, px), vy = float2vox1(g.vs_inv, py), vz = float2vox1(g.vs_inv, pz);
                const int bx = vx >> kBlockLog2, by = vy >> kBlockLog2, bz = vz >> kBlockLog2;  // arithmetic shift = floor
                const unsigned int sig = ((unsigned)bx & 1023u) | (((unsigned)by & 1023u) << 10) | (((unsigned)bz & 1023u) << 20);
                // sharded map: samples of other ranks' blocks (7 of 8 on 8 GPUs) are dropped before the sdf arithmetic
                const int keep = m.world > 1 ? shard_keeps<VARIANT>(m.world, m.rank, pack_block(bx, by, bz), vx, vy, vz) : 3;
                float sdf = 0.f, wgt = 0.f;
                if (keep) {
                    const float qx = __fsub_rn(__fmul_rn(g.vs, (float)vx), pose.t[0]);     // :107
                    const float qy = __fsub_rn(__fmul_rn(g.vs, (float)vy), pose.t[1]);
                    const float qz = __fsub_rn(__fmul_rn(g.vs, (float)vz), pose.t[2]);
                    sdf = __fsub_rn(dot3_rn(R[2], R[5], R[8], qx, qy, qz), z);             // :107-108
                    wgt = sdf_weight(sdf, g.T, g.inv_T);
                }
                if (wgt > 0.f) {                                                       // :110-111
                    if (sig != last_sig) {
                        last_sig = sig;
                        slot = map_find_or_insert(m, pack_block(bx, by, bz), ctr);
                        if (slot < 0) atomicAdd(&ctr->overflow, 1u);
                        if (STAMPS) ++d_probes;
                    }
                    if (slot >= 0) {
                        has = true;
                        vaddr = (unsigned long long)slot * kBlockVoxels + local_index(vx, vy, vz);
                        n_upd += keep >> 1;  // halo copies are not counted twice
                    }
                }
                wq = wgt; sq = sdf;
            }
            GSDF_FSTAMP(2);
            if (!__any_sync(0xffffffffu, has)) continue;  // nothing to commit in this unit (outside the band / not ours)
            // the summands, computed by every lane (lanes without a sample are in no one's group: theirs are never read)
            {
                const float ws = wq * sdf_truncate(sq, g.T);
                if (VARIANT == 0) { v[0] = ws; v[1] = wq * Rn0; v[2] = wq * Rn1; v[3] = wq * Rn2; v[4] = wq; }
                else { v[0] = wq; v[1] = ws; }
            }
            bool leader;
            warp_merge_by_key<(VARIANT == 0 ? 5 : 2)>(vaddr, narrow, has, v, leader);
            GSDF_FSTAMP(3);
            if (STAMPS && fdbg) d_reds += __popc(__ballot_sync(0xffffffffu, leader));
            if (leader) {
                if (VARIANT == 0) {
                    red_add_f4(m.acc4 + vaddr, v[0], v[1], v[2], v[3]);                // :115-117
                    red_add_f1(m.accW + vaddr, v[4]);
                } else {
                    red_add_f2(m.acc2 + vaddr, v[0], v[1]);                            // MapPixelSdfOmp.cpp:183-185
                }
            }
            if (m.vis_stamp != nullptr) {  // vis_[vi] gets `true` at this frame's counter (MapGradPixelSdfOmp.cpp:118-120)
                bool fresh = false;
                if (leader) fresh = atomicExch(m.vis_stamp + vaddr, frame_no + 1u) != frame_no + 1u;  // first touch in this frame
                const unsigned int fm = __ballot_sync(0xffffffffu, fresh);
                if (fm) {
                    unsigned long long base = 0;
                    if (lane == __ffs(fm) - 1) base = atomicAdd(&ctr->vis_cursor, (unsigned long long)__popc(fm));
                    base = __shfl_sync(0xffffffffu, base, __ffs(fm) - 1);
                    if (fresh) {
                        const unsigned long long o = base + __popc(fm & ((1u << lane) - 1));
                        if (o < m.vis_cap) m.vis_log[o] = (vaddr << 24) | (unsigned long long)(frame_no & 0xFFFFFFu);
                        else ctr->vis_overflow = 1u;
                    }
                }
            }
            if (STAMPS && dbg && d_first) { dbg[4] = global_ns(); d_first = false; }
        }
    }
    if (dbg) { dbg[5] = global_ns(); dbg[6] |= (unsigned long long)d_probes << 32; dbg[7] = d_reds; }
#undef GSDF_FSTAMP
}

// statistics: warp -> CTA (shared-memory atomics) -> one slot per CTA; never thousands of warps on one word
__device__ __forceinline__ void fuse_stats_flush(Counters* ctr, unsigned int n_pts, unsigned int n_upd) {
    __shared__ unsigned int s_pts, s_upd;
    const int lane = threadIdx.x & 31;
    if (threadIdx.x == 0) { s_pts = 0; s_upd = 0; }
    __syncthreads();
    n_pts = __reduce_add_sync(0xffffffffu, n_pts);
    n_upd = __reduce_add_sync(0xffffffffu, n_upd);
    if (lane == 0 && (n_pts | n_upd)) { atomicAdd(&s_pts, n_pts); atomicAdd(&s_upd, n_upd); }
    __syncthreads();
    if (threadIdx.x == 0 && (s_pts | s_upd)) {
        Counters::Slot* sl = ctr->slot + (blockIdx.x & (kCounterSlots - 1));
        atomicAdd(&sl->points, (unsigned long long)s_pts);
        atomicAdd(&sl->updates, (unsigned long long)s_upd);
    }
}

template <int VARIANT>
__global__ void __launch_bounds__(256)
fuse_list_kernel(MapView m, FrameGeom g, Pose pose_arg, int pose_src, const int2* __restrict__ vlist,
                 const int* __restrict__ n_valid_ptr, FrameState* st, const float* __restrict__ nx,
                 const float* __restrict__ ny, const float* __restrict__ nz, const float* __restrict__ x0,
                 const float* __restrict__ y0, const float* __restrict__ hom_inv, Counters* ctr,
                 ResultSnapshot* snap /* null, or where the last CTA leaves the frame result */) {
    if (pose_src == 2 && !st->converged) {  // nothing to integrate (main_scan_3d.cpp:261): only the result is published
        if (snap != nullptr && blockIdx.x == 0) write_snapshot(st, ctr, snap);
        return;
    }
    __shared__ Pose pose;
    if (threadIdx.x == 0) {
        if (pose_src == 0) pose = pose_arg;
        else {
            float q[4] = {st->pose7[3], st->pose7[4], st->pose7[5], st->pose7[6]};
            quat_to_R(q, pose.R);  // pose.rotationMatrix() of the stored SE3
            pose.t[0] = st->pose7[0]; pose.t[1] = st->pose7[1]; pose.t[2] = st->pose7[2];
        }
    }
    __syncthreads();
    const int n_valid = *n_valid_ptr;
    if (blockIdx.x == 0 && threadIdx.x == 0) st->n_valid = n_valid;  // reported with the frame result
    unsigned int n_upd = 0, n_pts = 0;
    fuse_work<VARIANT, false>(m, g, pose, vlist, n_valid, nx, ny, nz, x0, y0, hom_inv, ctr, n_pts, n_upd);
    fuse_stats_flush(ctr, n_pts, n_upd);
    {   // the last CTA advances Sdf::counter_ (increase_counter, MapGradPixelSdfOmp.cpp:128) and publishes the frame result
        __shared__ bool s_last;
        __threadfence();
        __syncthreads();
        if (threadIdx.x == 0) s_last = atomicInc(&ctr->ticket_snap, gridDim.x - 1) == gridDim.x - 1;
        __syncthreads();
        if (s_last) {
            __threadfence();
            if (threadIdx.x == 0) ctr->fused_frames += 1u;
            if (snap != nullptr) write_snapshot(st, ctr, snap);
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// FALS normals (cv::NormalEstimator<float>::compute, normals/NormalEstimator.h:179-204) per 32 x TH tile.  The masked
// reciprocal depth times the three cached planes is staged through shared memory with an r-pixel REFLECT_101 halo; the
// window sums follow the oracle's order exactly (double accumulation; rows left->right, then columns top->bottom, one
// rounding to float), so the normals — and therefore the validity tests that decide the voxel key set — are
// bit-identical to the reference build.
//   * tiles whose interior holds no z-valid pixel are skipped (fusion never reads normals of invalid pixels);
//   * the z_inv products are widened to double once, when staged, instead of once per window tap.
//   * RC > 0 fixes the window radius at compile time (RC = 5: the 11 x 11 window main_scan_3d.cpp:183 always uses): the tap
//     loops unroll, a thread's three planes run as three independent chains and all loads of a chain are in flight together
//     (with a runtime radius every tap cost a load, an add and three loop instructions: 2.4x the instructions);
//   * pixels outside the depth range get no normal (the pipeline's fusion reads normals of list entries only), and tiles
//     that do not touch the image border skip the reflection arithmetic.
template <int TH /* tile height: 32 x TH threads */, int RC /* window radius, 0 = runtime */>
__device__ __forceinline__ void normals_tile(const float* __restrict__ depth, const float* __restrict__ x0n,
                                             const float* __restrict__ y0n, const float* __restrict__ nsi,
                                             const float* __restrict__ Q, float* __restrict__ nx,
                                             float* __restrict__ ny, float* __restrict__ nz, int w, int h, int r_arg,
                                             float z_min, float z_max, int tile_x, int tile_y, int tix, int tiy,
                                             unsigned char* smem_raw) {
    const int r = RC ? RC : r_arg;
    const int tw = NT_W + 2 * r, th = TH + 2 * r;
    double* a = reinterpret_cast<double*>(smem_raw);   // [3][th][tw] products, widened
    double* rs = a + 3 * tw * th;                      // [3][th][NT_W] row sums
    const int x_base = tile_x * NT_W, y_base = tile_y * TH;
    const int x = x_base + tix, y = y_base + tiy;
    const size_t n = (size_t)w * h;
    const bool inside = x < w && y < h;
    const float zc = inside ? depth[(size_t)y * w + x] : 0.f;
    const bool wanted = inside && z_valid(zc, z_min, z_max);
    if (!__syncthreads_or(wanted)) return;

    const bool interior = x_base >= r && y_base >= r && x_base + NT_W + r <= w && y_base + TH + r <= h;
    for (int ly = tiy; ly < th; ly += TH) {
        const int gy = interior ? y_base + ly - r : reflect101(min(y_base + ly - r, h - 1 + r), h);
        for (int lx = tix; lx < tw; lx += NT_W) {
            const int gx = interior ? x_base + lx - r : reflect101(min(x_base + lx - r, w - 1 + r), w);
            const int gi = gy * w + gx;
            const float z = depth[gi];
            const float zi = z != 0.f ? __fdiv_rn(1.f, z) : 0.f;           // NormalEstimator.h:181-187
            const int i = ly * tw + lx;
            a[i] = (double)__fmul_rn(x0n[gi], zi);                          // :191
            a[tw * th + i] = (double)__fmul_rn(y0n[gi], zi);                // :192
            a[2 * tw * th + i] = (double)__fmul_rn(nsi[gi], zi);            // :193
        }
    }
    __syncthreads();
    for (int ly = tiy; ly < th; ly += TH) {
        const double* row = a + ly * tw + tix;
        double s1 = 0.0, s2 = 0.0, s3 = 0.0;
        if (RC) {
#pragma unroll
            for (int d = 0; d <= 2 * RC; ++d) {                             // left -> right
                s1 = __dadd_rn(s1, row[d]);
                s2 = __dadd_rn(s2, row[tw * th + d]);
                s3 = __dadd_rn(s3, row[2 * tw * th + d]);
            }
        } else {
            for (int d = 0; d <= 2 * r; ++d) {
                s1 = __dadd_rn(s1, row[d]);
                s2 = __dadd_rn(s2, row[tw * th + d]);
                s3 = __dadd_rn(s3, row[2 * tw * th + d]);
            }
        }
        rs[ly * NT_W + tix] = s1;
        rs[(th + ly) * NT_W + tix] = s2;
        rs[(2 * th + ly) * NT_W + tix] = s3;
    }
    __syncthreads();
    if (!wanted) return;
    double s1 = 0.0, s2 = 0.0, s3 = 0.0;
    if (RC) {
#pragma unroll
        for (int d = 0; d <= 2 * RC; ++d) {                                 // top -> bottom
            const int j = (tiy + d) * NT_W + tix;
            s1 = __dadd_rn(s1, rs[j]);
            s2 = __dadd_rn(s2, rs[th * NT_W + j]);
            s3 = __dadd_rn(s3, rs[2 * th * NT_W + j]);
        }
    } else {
        for (int d = 0; d <= 2 * r; ++d) {
            const int j = (tiy + d) * NT_W + tix;
            s1 = __dadd_rn(s1, rs[j]);
            s2 = __dadd_rn(s2, rs[th * NT_W + j]);
            s3 = __dadd_rn(s3, rs[2 * th * NT_W + j]);
        }
    }
    const float b1 = (float)s1, b2 = (float)s2, b3 = (float)s3;
    const size_t gi = (size_t)y * w + x;
    const float q11 = Q[gi], q12 = Q[n + gi], q13 = Q[2 * n + gi], q22 = Q[3 * n + gi], q23 = Q[4 * n + gi], q33 = Q[5 * n + gi];
    const float vx = __fadd_rn(__fadd_rn(__fmul_rn(b1, q11), __fmul_rn(b2, q12)), __fmul_rn(b3, q13));
    const float vy = __fadd_rn(__fadd_rn(__fmul_rn(b1, q12), __fmul_rn(b2, q22)), __fmul_rn(b3, q23));
    const float vz = __fadd_rn(__fadd_rn(__fmul_rn(b1, q13), __fmul_rn(b2, q23)), __fmul_rn(b3, q33));
    const float nn = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(vx, vx), __fmul_rn(vy, vy)), __fmul_rn(vz, vz)));
    nx[gi] = __fdiv_rn(vx, nn);
    ny[gi] = __fdiv_rn(vy, nn);
    nz[gi] = __fdiv_rn(vz, nn);
}

__global__ void __launch_bounds__(NT_W* NT_H)
normals_tile_kernel(const float* __restrict__ depth, const float* __restrict__ x0n, const float* __restrict__ y0n,
                    const float* __restrict__ nsi, const float* __restrict__ Q, float* __restrict__ nx,
                    float* __restrict__ ny, float* __restrict__ nz, int w, int h, int r, float z_min, float z_max) {
    extern __shared__ __align__(16) unsigned char smem_dyn[];
    if (r == 5) normals_tile<NT_H, 5>(depth, x0n, y0n, nsi, Q, nx, ny, nz, w, h, r, z_min, z_max, blockIdx.x, blockIdx.y, threadIdx.x,
                                      threadIdx.y, smem_dyn);
    else normals_tile<NT_H, 0>(depth, x0n, y0n, nsi, Q, nx, ny, nz, w, h, r, z_min, z_max, blockIdx.x, blockIdx.y, threadIdx.x,
                               threadIdx.y, smem_dyn);
}

// The frames one launch of track_kernel works through, in order (gsdf_scan_run hands it up to kMaxBatch frames whose
// lists and normals the side stream has prepared; every other caller passes one).  Entering, leaving and relaunching
// the kernel costs 7.6 us per launch at 640x480 (entry 1.2, result + exit 2.4, kernel-to-kernel gap 4); between two
// frames of a batch there is one grid barrier instead.
constexpr int kMaxBatch = 8;
struct FrameBatch {
    int n;                                  // frames in this launch
    int fuse_mode;                          // 0 = never, 1 = always (setup / ground-truth pose), 2 = iff the optimisation converged
    int2* vlist[kMaxBatch];                 // compacted valid pixels of frame f
    int* n_valid[kMaxBatch];
    float* normals[kMaxBatch];              // null: track only; else the 3 planes the fusion of frame f reads
    ResultSnapshot* snap[kMaxBatch];        // null, or where the result of frame f is left
    unsigned long long* flog[kMaxBatch];    // null, or frame f's entry of the debug frame log (words 2 .. 6)
    unsigned long long* dbg;                // null, or [gridDim.x][4 passes][6] + [gridDim.x][4] + [warps][8] %globaltimer stamps
};
// element f of a kernel-parameter array without a runtime-indexed access (which would copy the array to local memory)
template <typename T>
__device__ __forceinline__ T pick(T const (&a)[kMaxBatch], int f) {
    T r = a[0];
#pragma unroll
    for (int i = 1; i < kMaxBatch; ++i) r = (f == i) ? a[i] : r;
    return r;
}

#define GSDF_STAMP(i) do { if (STAMPS && fb.dbg && threadIdx.x == 0 && k < 4) fb.dbg[((size_t)blockIdx.x * 4 + k) * 6 + (i)] = global_ns(); } while (0)
// whole-launch stamps, after the per-pass ones: [gridDim.x][4] = entry, Gauss-Newton loop left, fusion done, exit
#define GSDF_STAMPX(i) do { if (STAMPS && fb.dbg && threadIdx.x == 0) fb.dbg[(size_t)gridDim.x * 24 + (size_t)blockIdx.x * 4 + (i)] = global_ns(); } while (0)

// Frame-only work of the pipeline in ONE launch: each CTA compacts the z-valid pixels of its 32 x 8 tile (8 warps <-> its
// eight 8x4 sub-tiles) and then computes the tile's normals.  List order: 32x8 tiles row-major, 8x4 sub-tiles row-major
// inside, pixels row-major inside.
// 256-thread CTAs, not 512: beside the resident tracker (768 threads x 64 registers) the register file has room for 512
// threads at 32 registers, and two independent CTAs hide each other's barriers and memory latency (the kernel is
// latency-bound: ~10 % issue utilisation), where one 512-thread CTA stalls as a whole at every __syncthreads.
// The exclusive prefix over CTAs is a decoupled look-back: a CTA publishes its count as {epoch, AGGREGATE, count}, looks
// at its 32 predecessors at once, adds their counts back to the nearest one that already carries a PREFIX, and publishes
// {epoch, PREFIX, inclusive sum} itself — one or two window reads instead of a poll of every lower-numbered CTA.
// Lower-numbered CTAs are dispatched first and wait only on even lower ones, so the chain always makes progress; the
// epoch makes clearing between launches unnecessary.
constexpr int NC_H = 8;                                // tile height of normals_compact_kernel
constexpr unsigned long long kLbPrefix = 1ull << 31;   // flag word = epoch << 32 | PREFIX bit | value (< 2^31)
__global__ void __launch_bounds__(NT_W* NC_H, 8)  // <= 32 registers: two CTAs fit beside the resident tracker CTA
normals_compact_kernel(const float* __restrict__ depth, const float* __restrict__ x0n, const float* __restrict__ y0n,
                       const float* __restrict__ nsi, const float* __restrict__ Q, float* __restrict__ nx,
                       float* __restrict__ ny, float* __restrict__ nz, int w, int h, int r, float z_min, float z_max,
                       unsigned long long* __restrict__ flags, unsigned int epoch, int2* __restrict__ vlist,
                       int* __restrict__ n_valid_out, unsigned long long* flog) {
    extern __shared__ __align__(16) unsigned char smem_dyn[];
    constexpr int NWARP = NT_W * NC_H / 32;
    __shared__ int s_warp[NWARP];
    __shared__ int s_base;
    const int tid = threadIdx.y * NT_W + threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) flog_enter(flog);
    const int cta = blockIdx.y * gridDim.x + blockIdx.x, ncta = gridDim.x * gridDim.y;
    const int px = blockIdx.x * NT_W + (warp & 3) * kTileW + (lane & 7);
    const int py = blockIdx.y * NC_H + (warp >> 2) * kTileH + (lane >> 3);
    const bool in = px < w && py < h;
    const float z = in ? __ldg(depth + (size_t)py * w + px) : 0.f;
    const unsigned int mask = __ballot_sync(0xffffffffu, in && z_valid(z, z_min, z_max));
    if (lane == 0) s_warp[warp] = __popc(mask);
    __syncthreads();
    int before = 0, total = 0;
#pragma unroll
    for (int q = 0; q < NWARP; ++q) { const int c = s_warp[q]; before += q < warp ? c : 0; total += c; }
    const unsigned long long ep = (unsigned long long)epoch << 32;
    if (warp == 0) {
        if (lane == 0) st_word(flags + cta, ep | (unsigned long long)total);
        int excl = 0;
        for (int look = cta - 1 - lane;; look -= 32) {     // window of 32 predecessors, nearest in lane 0
            unsigned long long f = ep | kLbPrefix;          // before the first CTA: prefix 0
            if (look >= 0) {
                do { f = ld_word(flags + look); } while ((f >> 32) != epoch);
            }
            const unsigned int pm = __ballot_sync(0xffffffffu, (f & kLbPrefix) != 0);
            const int nearest = pm ? __ffs(pm) - 1 : 31;    // lanes beyond the nearest prefix do not count
            excl += __reduce_add_sync(0xffffffffu, lane <= nearest ? (int)(f & (kLbPrefix - 1)) : 0);
            if (pm) break;
        }
        if (lane == 0) {
            st_word(flags + cta, ep | kLbPrefix | (unsigned long long)(excl + total));
            s_base = excl;
            if (cta == ncta - 1) *n_valid_out = excl + total;
        }
    }
    if (total == 0) {  // no valid pixel in the tile: fusion never reads its normals
        if (tid == 0) flog_exit(flog);
        return;
    }
    // warps 1.. start staging the tile while warp 0 is still looking back (a quarter of this kernel's stall samples sat in
    // that poll when every warp waited for it); the list entries are written afterwards — s_base is ordered before them by
    // the barriers inside normals_tile
    if (r == 5) normals_tile<NC_H, 5>(depth, x0n, y0n, nsi, Q, nx, ny, nz, w, h, r, z_min, z_max, blockIdx.x, blockIdx.y, threadIdx.x,
                                      threadIdx.y, smem_dyn);
    else normals_tile<NC_H, 0>(depth, x0n, y0n, nsi, Q, nx, ny, nz, w, h, r, z_min, z_max, blockIdx.x, blockIdx.y, threadIdx.x,
                               threadIdx.y, smem_dyn);
    if (mask & (1u << lane))
        vlist[s_base + before + __popc(mask & ((1u << lane) - 1))] = make_int2((py << 16) | px, __float_as_int(z));
    if (tid == 0) flog_exit(flog);
}

// Grid-wide barrier of the persistent tracker (cooperative launch: all CTAs are resident).  An arrival counter
// instead of cooperative_groups' grid.sync(): barrier number b is passed once the counter reaches (b + 1) * gridDim.x,
// the last CTA to leave the kernel resets it, and a wait that exceeds 2 s sets barrier_timeout (reported by
// gsdf_scan_collect as an error) instead of hanging the device.
// The barrier in two halves, so that work which does not depend on the other CTAs can be done between arriving and waiting.
__device__ __forceinline__ void grid_arrive(Counters* ctr) {
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(&ctr->barrier, 1u);
    }
}
__device__ __forceinline__ void grid_wait(Counters* ctr, unsigned int target) {
    if (threadIdx.x == 0) {
        unsigned int seen, spins = 0;
        unsigned long long t0 = 0;
        for (;;) {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(&ctr->barrier) : "memory");
            if (seen >= target) break;
            if ((++spins & 4095u) == 0) {
                const unsigned long long now = global_ns();
                if (t0 == 0) t0 = now;
                else if (now - t0 > 2000000000ull) { ctr->barrier_timeout = 1u; break; }
            }
        }
    }
    __syncthreads();
}

// Sharded maps: the all-reduce of the 29 sums INSIDE the tracker launch, over NVLink peer memory, with a
// low-latency (flag-in-data) protocol: every rank owns an exchange buffer [4][world][32] of 64-bit words;
// after its grid-level reduction CTA 0 stores {pass tag, float bits} for each of its 29 sums into slot `rank` of EVERY
// peer's buffer (one 8-byte store each, atomic, through the peer mapping) — no fence, no separate flag, so the cost
// is ONE NVLink traversal.  All CTAs then spin on the words in their OWN memory until each carries the tag of this
// pass and add the world partials in rank order: identical sums, hence identical poses and stopping decisions, on
// every rank, with no NCCL launch and no host involvement inside the Gauss-Newton loop.
constexpr int kMaxPeers = 8;
struct PeerExchange {
    unsigned long long* data[kMaxPeers];  // data[r]: rank r's exchange buffer (peer-mapped; data[rank] is local)
    int world, rank;                      // world <= 1: no exchange
    unsigned int epoch0;                  // pass k of this launch (counted over its frames) carries the tag epoch0 + k + 1
    unsigned int launch;                  // launch counter (buffer parity across launches)
    int* error;                           // set to 1 if a peer did not show up in time
};

__device__ __forceinline__ void peer_allreduce_29(const PeerExchange& px, int k, float* s_tot) {
    // four buffers: (launch parity, pass parity).  Inside a launch two suffice (see exchange_29); a rank that runs ahead
    // into launch L+1 writes the other launch parity, and it can reach launch L+2 only after every peer has finished L.
    const int par = ((int)(px.launch & 1u) << 1) | (k & 1);
    const unsigned long long ep = (unsigned long long)(px.epoch0 + (unsigned)k + 1u) << 32;
    __syncthreads();  // s_tot holds this rank's sums
    if (blockIdx.x == 0) {
        for (int i = threadIdx.x; i < px.world * kAccStride; i += blockDim.x) {
            const int peer = i / kAccStride, term = i - peer * kAccStride;
            if (term < kAccTerms)
                *reinterpret_cast<volatile unsigned long long*>(px.data[peer] + (par * px.world + px.rank) * kAccStride + term) =
                    ep | (unsigned long long)__float_as_uint(s_tot[term]);
        }
    }
    __syncthreads();
    if (threadIdx.x < kAccTerms) {
        const volatile unsigned long long* d = px.data[px.rank] + (size_t)par * px.world * kAccStride + threadIdx.x;
        const unsigned long long t0 = global_ns();
        float s = 0.f;
        for (int r = 0; r < px.world; ++r) {
            unsigned long long v;
            unsigned int spins = 0;
            while (((v = d[r * kAccStride]) >> 32 << 32) != ep) {
                if ((++spins & 1023u) == 0 && global_ns() - t0 > 500000000ull) { *px.error = 1; break; }  // 0.5 s
            }
            s += __uint_as_float((unsigned int)v);
        }
        s_tot[threadIdx.x] = s;
    }
    __syncthreads();
}

template <int VARIANT, int NT, bool STAMPS = false /* phase stamps into fb.dbg (gsdf_debug_track_timers armed) */>
__global__ void __launch_bounds__(NT == 384 ? 512 : 1024, NT == 384 ? 2 : 1)  // 64 registers per thread (see kTrackThreadsMax)
track_kernel(MapView m, FrameGeom g, TrackIntr ti, FrameState* st,
             unsigned long long* __restrict__ rows /* [2][gridDim.x][kAccStride] tagged words */, unsigned int tag0,
             TrackParams tp, const float* __restrict__ planes /* the 11 cached FALS planes */, Counters* ctr, FrameBatch fb,
             PeerExchange px) {
    constexpr int NW = NT / 32;
    unsigned int barriers = 0;  // grid barriers passed so far in this launch
    bool fused_prev = false;    // the frame just finished was integrated (every CTA takes the same decision)
    int pass_id = 0;            // Gauss-Newton passes run so far in this launch (tags of both exchanges)
    __shared__ float s_pose7[8];
    __shared__ Pose s_pose;
    __shared__ float s_tot[kAccStride];
    __shared__ float s_warp[NW][kAccStride];   // CTA stage of the reduction
    __shared__ float s_part[NW][kAccStride];   // grid stage
    __shared__ int s_conv;
    __shared__ float s_valid_sum;  // contributing pixels summed over the frame's passes (thread 0)
    __shared__ float s_pf[kPfWords * NT];  // pixel data of each warp's first fusion unit (fuse_pixel_planes)
    __shared__ unsigned long long s_ck[kMorePixels * NT];  // block probed for the thread's 2nd, 3rd pixel in the previous pass
    __shared__ int s_cs[kMorePixels * NT];                 // ... and its slot
    GSDF_STAMPX(0);
    // pixel j of the list -> warp-sized chunk j/32 -> CTA (chunk % grid), warp (chunk / grid): a frame with fewer valid
    // pixels than resident threads (the usual case: ~90 k of 114 k) is spread over ALL SMs instead of filling the
    // first CTAs completely and leaving the rest idle (the slowest CTA sets the length of every pass)
    const int nthreads = gridDim.x * blockDim.x;
    const int gtid = (((threadIdx.x >> 5) * gridDim.x + blockIdx.x) << 5) | (threadIdx.x & 31);
    // the pose the first frame starts from; later frames of the batch continue from the pose in shared memory
    if (threadIdx.x < 7) s_pose7[threadIdx.x] = __ldcg(&st->pose7[threadIdx.x]);
    for (int f = 0; f < fb.n; ++f) {
        int2* vlist = pick(fb.vlist, f);
        float* fuse_normals = pick(fb.normals, f);
        unsigned long long* flog = pick(fb.flog, f);
        // the loads every frame starts with (list length, the thread's list entry) do not depend on the map: they are
        // issued before the barrier that ends the previous frame; the entry is fetched before it is known to exist
        // (the list buffer holds one entry per pixel of the frame) and stays in registers over the passes (with fewer
        // valid pixels than resident threads it is the thread's only one)
        const int n_valid = __ldcg(pick(fb.n_valid, f));
        int2 e0 = make_int2(0, 0);
        if (gtid < g.w * g.h) e0 = __ldcg(vlist + gtid);
        unsigned long long ck = kEmptyKey;  // block probed for this thread's pixel in the previous pass
        int cs = -1;
        if (f > 0) {
            // every CTA's fusion of the previous frame (and its statistics) is complete before anyone reads the map again.
            // Between arriving and waiting: the block of this thread's pixel at the pose the frame starts from (the one the
            // previous frame ended with, still in shared memory) is looked up in the key table, so that the first pass
            // finds its slot cached like the later ones do.  Other CTAs may still be inserting: a block that is found stays
            // where it is, one that is not found is looked up again by the pass itself.
            grid_arrive(ctr);
            if (gtid < n_valid) {
                const unsigned long long key = first_block_key<VARIANT>(m, g, ti, s_pose.R, s_pose.t, e0.x, __int_as_float(e0.y));
                if (m.world <= 1 || block_owner(key, m.world) == m.rank) {
                    const int slot = map_find(m, key);
                    if (slot >= 0) { ck = key; cs = slot; }
                }
            }
            grid_wait(ctr, ++barriers * gridDim.x);
            // CTA 0 publishes the previous frame; it does so before it publishes its sums of this frame's first pass, which
            // every CTA waits for before it can reach this frame's fusion, so no statistics of this frame are in yet
            ResultSnapshot* prev = pick(fb.snap, f - 1);
            if (blockIdx.x == 0 && prev != nullptr) write_snapshot(st, ctr, prev, px.world > 1 ? px.error : nullptr);
            if (blockIdx.x == 0 && threadIdx.x == 0 && fused_prev) ctr->fused_frames += 1u;  // increase_counter()
        }
        if (threadIdx.x == 0 && flog) flog_enter(flog + 2);
        int converged = 0, k = 0;
        if (VARIANT == 0) {
#pragma unroll
            for (int i = 0; i < kMorePixels; ++i) s_ck[i * NT + threadIdx.x] = kEmptyKey;
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            quat_to_R(s_pose7 + 3, s_pose.R);  // pose_.rotationMatrix(), RigidPointOptimizerOmp.cpp:59
            s_pose.t[0] = s_pose7[0]; s_pose.t[1] = s_pose7[1]; s_pose.t[2] = s_pose7[2];
            s_valid_sum = 0.f;
        }
        __syncthreads();
        for (k = 0; k < tp.max_iters; ++k, ++pass_id) {
            GSDF_STAMP(0);
            // every pixel's 29 products are reduced across the warp at once: lane L keeps the warp's running sum of term L
            // (the loop bound is the chunk's first entry, so whole warps run it together)
            float wsum = 0.f;
            // A frame with more valid pixels than resident threads (room scale: 2.7 per thread) gives a thread several pixels:
            // (grad-sdf) the NEXT pixel's list entry is fetched while this one is looked up, and the probed block of the
            // thread's second and third pixel is remembered in shared memory like the first one's is in registers — otherwise
            // every further pixel costs three dependent L2 round trips (entry, key, voxel) in every pass.
            if (VARIANT == 1) {
                // base-sdf keeps the plain loop: a second inlined copy of its 8-corner lookup (or the state to share one copy)
                // costs the usual one-pixel-per-thread frame 3 % (measured), more than it gains on large frames
#pragma unroll 1
                for (int j = gtid; (j & ~31) < n_valid; j += nthreads) {
                    const bool has = j < n_valid;
                    const int2 e = j == gtid ? e0 : (has ? __ldcg(vlist + j) : make_int2(0, 0));
                    float v[32];
                    track_pixel_terms<VARIANT>(m, g, ti, s_pose.R, s_pose.t, has, e.x, __int_as_float(e.y), v, ck, cs);
                    wsum += warp_reduce_scatter_32(v);
                }
            } else {
                int2 e = e0;
                int it = 0;
#pragma unroll 1
                for (int j = gtid; (j & ~31) < n_valid; j += nthreads, ++it) {
                    const bool has = j < n_valid;
                    int2 en = make_int2(0, 0);
                    if (j + nthreads < n_valid) en = __ldcg(vlist + j + nthreads);
                    float v[32];
                    if (it == 0) {
                        track_pixel_terms<VARIANT>(m, g, ti, s_pose.R, s_pose.t, has, e.x, __int_as_float(e.y), v, ck, cs);
                    } else {
                        const bool slot = it <= kMorePixels;
                        unsigned long long ckx = slot ? s_ck[(it - 1) * NT + threadIdx.x] : kEmptyKey;
                        int csx = slot ? s_cs[(it - 1) * NT + threadIdx.x] : -1;
                        track_pixel_terms<VARIANT>(m, g, ti, s_pose.R, s_pose.t, has, e.x, __int_as_float(e.y), v, ckx, csx);
                        if (slot) { s_ck[(it - 1) * NT + threadIdx.x] = ckx; s_cs[(it - 1) * NT + threadIdx.x] = csx; }
                    }
                    wsum += warp_reduce_scatter_32(v);
                    e = en;
                }
            }
            GSDF_STAMP(1);
            const float mine = block_sum_29<NW>(wsum, s_warp);
            GSDF_STAMP(2);
            // while the sums travel: the pixel this thread's fusion starts with (depends on neither the pose nor the map)
            exchange_29<NW>(rows, tag0 + (unsigned)pass_id + 1u, mine, s_part, s_tot, ctr, tp.poll_ns, [&]() {
                if (k > 1 || fuse_normals == nullptr) return;
                if (k == 0) {  // its list entry
                    int u0, u1;
                    fuse_range(n_valid, 2 * g.factor + 1, u0, u1);
                    const int j = (u0 / (2 * g.factor + 1)) * 32 + (int)(threadIdx.x & 31);
                    int2 e = make_int2(-1, 0);
                    if (u0 < u1 && j < n_valid) e = __ldcg(vlist + j);
                    s_pf[threadIdx.x] = __int_as_float(e.x);
                    s_pf[NT + threadIdx.x] = __int_as_float(e.y);
                } else {       // the planes the entry indexes
                    const int2 e = make_int2(__float_as_int(s_pf[threadIdx.x]), __float_as_int(s_pf[NT + threadIdx.x]));
                    if (e.x >= 0) {
                        const size_t np = (size_t)g.w * g.h;
                        float d7[kPfWords];
                        fuse_pixel_planes<true>(g, e, fuse_normals, fuse_normals + np, fuse_normals + 2 * np, planes, planes + np,
                                                planes + 4 * np, d7);
#pragma unroll
                        for (int i = 0; i < kPfWords; ++i) s_pf[i * NT + threadIdx.x] = d7[i];
                    }
                }
            });
            GSDF_STAMP(3);
            if (px.world > 1) peer_allreduce_29(px, pass_id, s_tot);
            GSDF_STAMP(4);
            if (threadIdx.x < 32) {  // (s_tot was written by threads 0..28 of this warp)
                __syncwarp();
                if (threadIdx.x == 0) {
                    s_valid_sum += s_tot[28];
                    s_conv = gauss_newton_step(s_tot, tp.damping, tp.thr_sq, s_pose7);
                    quat_to_R(s_pose7 + 3, s_pose.R);
                    s_pose.t[0] = s_pose7[0]; s_pose.t[1] = s_pose7[1]; s_pose.t[2] = s_pose7[2];
                }
            }
            __syncthreads();
            GSDF_STAMP(5);
            if (k == 0 && threadIdx.x == 0 && flog) atomicMax(flog + 4, global_ns());  // first pass done (last CTA)
            if (s_conv) { converged = 1; ++pass_id; break; }
        }
        GSDF_STAMPX(1);
        if (threadIdx.x == 0 && flog) {  // Gauss-Newton loop left (last CTA), passes run
            atomicMax(flog + 5, global_ns());
            if (blockIdx.x == 0) flog[6] = (unsigned long long)(converged ? k + 1 : k);
        }
        if (blockIdx.x == 0) {
            if (threadIdx.x < 7) st->pose7[threadIdx.x] = s_pose7[threadIdx.x];
            if (threadIdx.x < kAccTerms) st->sums[threadIdx.x] = s_tot[threadIdx.x];
            if (threadIdx.x == 0) {
                st->n_valid = n_valid;
                st->converged = converged;
                st->iters = converged ? k : tp.max_iters;
                st->track_iters += (unsigned long long)(converged ? k + 1 : tp.max_iters);
                st->track_valid += (unsigned long long)s_valid_sum;
            }
            __threadfence();  // the snapshot of this frame is read back from L2 (by other threads of this CTA, or by the last CTA)
        }
        // main_scan_3d.cpp:261-264: integrate the frame at the pose just found, iff the optimisation converged — in the
        // same launch (every CTA holds the final pose in shared memory; the map is not read by anyone else meanwhile)
        fused_prev = fuse_normals != nullptr && (fb.fuse_mode == 1 || (fb.fuse_mode == 2 && converged));
        if (fused_prev) {
            const size_t np = (size_t)g.w * g.h;
            unsigned int n_upd = 0, n_pts = 0;
            fuse_work<VARIANT, true, STAMPS>(m, g, s_pose, vlist, n_valid, fuse_normals, fuse_normals + np, fuse_normals + 2 * np,
                                             planes, planes + np, planes + 4 * np, ctr, n_pts, n_upd,
                                             (STAMPS && fb.dbg) ? fb.dbg + (size_t)gridDim.x * 28 : nullptr,
                                             (converged ? k + 1 : k) >= 2 ? s_pf : nullptr);  // two passes have run
            fuse_stats_flush(ctr, n_pts, n_upd);
        }
        GSDF_STAMPX(2);
        if (threadIdx.x == 0 && flog) flog_exit(flog + 2);
    }
    {   // the last CTA to get here re-arms the grid barrier and publishes the result of the last frame
        __shared__ bool s_last;
        __threadfence();
        __syncthreads();
        if (threadIdx.x == 0) s_last = atomicInc(&ctr->ticket_snap, gridDim.x - 1) == gridDim.x - 1;
        __syncthreads();
        if (s_last) {
            __threadfence();
            if (threadIdx.x == 0) { ctr->barrier = 0u; if (fused_prev) ctr->fused_frames += 1u; }
            ResultSnapshot* snap = pick(fb.snap, fb.n - 1);
            if (snap != nullptr) write_snapshot(st, ctr, snap, px.world > 1 ? px.error : nullptr);
        }
    }
    GSDF_STAMPX(3);
}

// Per-iteration variant (sharded maps): sums of this rank's pixels -> out29 (last-block-done), the caller
// all-reduces out29 across ranks and then runs solve_kernel.  Exits at once when the loop is already done.
template <int VARIANT>
__global__ void __launch_bounds__(kAccThreads)
accumulate_list_kernel(MapView m, FrameGeom g, TrackIntr ti, const int2* __restrict__ vlist,
                       const int* __restrict__ n_valid_ptr, const FrameState* st, float* __restrict__ partials,
                       Counters* ctr, float* __restrict__ out29) {
    if (st->done) return;
    __shared__ Pose s_pose;
    __shared__ float sm[kAccThreads / 32][kAccStride];
    if (threadIdx.x == 0) {
        float q[4] = {st->pose7[3], st->pose7[4], st->pose7[5], st->pose7[6]};
        quat_to_R(q, s_pose.R);
        s_pose.t[0] = st->pose7[0]; s_pose.t[1] = st->pose7[1]; s_pose.t[2] = st->pose7[2];
    }
    __syncthreads();
    float acc[kAccTerms];
#pragma unroll
    for (int i = 0; i < kAccTerms; ++i) acc[i] = 0.f;
    const int n_valid = *n_valid_ptr;
    unsigned long long ck = kEmptyKey;
    int cs = -1;
    for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < n_valid; j += gridDim.x * blockDim.x) {
        const int2 e = __ldg(vlist + j);
        track_pixel<VARIANT>(m, g, ti, s_pose.R, s_pose.t, e.x, __int_as_float(e.y), acc, ck, cs);
    }
    block_reduce_29<kAccThreads / 32>(acc, sm, partials + (size_t)blockIdx.x * kAccStride);
    __shared__ bool is_last;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) is_last = (atomicInc(&ctr->ticket, gridDim.x - 1) == gridDim.x - 1);
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    grid_reduce_29<kAccThreads / 32>(partials, gridDim.x, sm, out29);
}

__global__ void solve_kernel(FrameState* st, const float* __restrict__ sums29, TrackParams tp) {
    if (threadIdx.x != 0 || st->done) return;
    float pose7[7], sums[kAccTerms];
    for (int i = 0; i < 7; ++i) pose7[i] = st->pose7[i];
    for (int i = 0; i < kAccTerms; ++i) { sums[i] = sums29[i]; st->sums[i] = sums[i]; }
    const int conv = gauss_newton_step(sums, tp.damping, tp.thr_sq, pose7);
    for (int i = 0; i < 7; ++i) st->pose7[i] = pose7[i];
    st->track_iters += 1;
    st->track_valid += (unsigned long long)sums[28];
    if (conv) { st->converged = 1; st->iters = st->k; st->done = 1; }
    else if (++st->k >= tp.max_iters) { st->converged = 0; st->iters = tp.max_iters; st->done = 1; }
}

__global__ void track_begin_kernel(FrameState* st) {
    if (threadIdx.x == 0) { st->done = 0; st->k = 0; st->converged = 0; st->iters = 0; }
}


}  // namespace gsdf
