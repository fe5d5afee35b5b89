// libgsdf_b200.so — C ABI (include/gsdf_b200.h) and host logic of the B200 Scan3D hot path.
// Host side: handle lifetime, the one-time FALS cache (cv::NormalEstimator::cache, double precision), stream and
// event plumbing of the frame pipeline, NCCL plumbing for sharded maps.
// Device side: gsdf_kernels.cuh (v1 kernels, still used by the fine-grained entry points) and gsdf_pipeline.cuh
// (compaction, persistent tracker, list-driven fusion, tile-skipping normals).
// There is no CPU implementation of the per-frame path in this library.
#include "../../include/gsdf_b200.h"

#include <cuda_runtime.h>
#include <dlfcn.h>

#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "gsdf_pipeline.cuh"
#include "gsdf_mesh.cuh"

using namespace gsdf;

// ----------------------------------------------------------------------------------------------------------
// NCCL through dlopen: single-GPU users need no NCCL at all, multi-GPU users get whichever libnccl.so.2 the
// process already has (torch's bundled one under torchrun) or the system one.
namespace {
typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
struct NcclApi {
    void* lib = nullptr;
    int (*GetUniqueId)(ncclUniqueId*) = nullptr;
    int (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    bool load(std::string& err) {
        if (lib) return true;
        const char* names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char* n : names) { lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL); if (lib) break; }
        if (!lib) { err = "cannot dlopen libnccl.so.2"; return false; }
        GetUniqueId = (int (*)(ncclUniqueId*))dlsym(lib, "ncclGetUniqueId");
        CommInitRank = (int (*)(ncclComm_t*, int, ncclUniqueId, int))dlsym(lib, "ncclCommInitRank");
        CommDestroy = (int (*)(ncclComm_t))dlsym(lib, "ncclCommDestroy");
        AllReduce = (int (*)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t))dlsym(lib, "ncclAllReduce");
        GetErrorString = (const char* (*)(int))dlsym(lib, "ncclGetErrorString");
        if (!GetUniqueId || !CommInitRank || !CommDestroy || !AllReduce) { err = "libnccl lacks expected symbols"; return false; }
        return true;
    }
};
NcclApi g_nccl;
constexpr int kNcclFloat = 7, kNcclSum = 0;  // ncclFloat32, ncclSum (nccl.h)
std::string g_create_error;
constexpr int kFlogFrames = 256;
constexpr int kRing = 24;       // frame slots: gsdf_scan_run keeps up to three batches of kMaxBatch frames in flight
constexpr int kMaxPending = 4;  // frames in flight through gsdf_scan_submit (the documented limit)
}  // namespace

// tracker state and statistics live side by side on the device so that a frame's result is ONE small copy
struct DeviceResult {
    FrameState st;
    Counters ctr;
};

struct FrameSlot {
    float* d_depth = nullptr;              // landing buffer for host uploads
    cudaEvent_t ev_up = nullptr, ev_norm = nullptr, ev_done = nullptr, ev_result = nullptr;
    // The frame's last kernel leaves the result in d_snap; the result stream copies it to h_snap (pinned) once ev_done
    // has fired.  Nothing of this sits on the main stream: a kernel that wrote the 256 bytes straight into mapped host
    // memory took 1.5 us longer to complete (the PCIe write has to drain before the next kernel may start), and a copy
    // operation on the main stream between two frames measured 8 us.
    // A frame submitted to an EMPTY pipeline (latency matters, nothing follows it yet) is written straight into h_snap,
    // which is mapped: 8 us sooner at the host than through the copy.
    ResultSnapshot* d_snap = nullptr;
    ResultSnapshot* h_snap = nullptr;
    bool direct = false;                   // this frame's result goes straight to h_snap
    cudaEvent_t wait_ev = nullptr;         // what gsdf_scan_collect waits for (the event of the frame's batch)
    int mode = 0;
};

struct gsdf_map {
    gsdf_config cfg;
    FrameGeom geom;
    MapView view;
    size_t cap = 0;
    cudaStream_t stream = nullptr;    // main: prepare, track, fuse, results
    cudaStream_t stream_n = nullptr;  // normals (overlaps tracking in the pipeline)
    cudaStream_t stream_c = nullptr;  // host -> device frame copies
    cudaStream_t stream_r = nullptr;  // device -> host frame results
    // device buffers
    float* d_planes = nullptr;        // 11 cached planes
    const float* d_depth = nullptr;   // current frame (ring slot or caller's device buffer)
    float* d_normals[kRing] = {};     // one set of frame-only buffers (normals, pixel list, count) per frame in flight
    int normals_buf = 0;              // buffer index (normals, vlist, nvalid) of the current frame
    bool normals_valid = false, prepared = false;
    int2* d_vlist[kRing] = {};              // compacted valid pixels, buffered like the normals
    int* d_nvalid = nullptr;                // [kRing] counts of the lists
    unsigned int* d_compact_flags = nullptr;  // [nseg] look-back flags {epoch, count} of compact_tiles_kernel
    unsigned long long* d_compact_flags2 = nullptr; // [tiles of normals_compact_kernel] look-back words {epoch, status, value}
    unsigned int compact_epoch = 0;
    int nseg = 0;                           // CTAs of the compaction (64 tiles of 8x4 pixels each)
    DeviceResult* d_res = nullptr;    // owns d_state and d_ctr
    FrameState* d_state = nullptr;
    Counters* d_ctr = nullptr;
    float* d_partials = nullptr;      // v1 accumulate / accumulate_list: [acc_blocks][32]
    int acc_blocks = 0;
    unsigned long long* d_track_rows = nullptr;  // track_kernel: [2][track_blocks][32] tagged words of the grid-level exchange
    unsigned int track_tag = 0;         // tags handed out so far (every launch gets a fresh range)
    int track_blocks = 0;
    int track_threads = kTrackThreadsMax;  // 768 (one CTA per SM) or 384 (two): GSDF_TRACK_THREADS
    Pose* d_pose = nullptr;
    float* d_out29 = nullptr;
    float* d_query = nullptr;         // staging of gsdf_query / gsdf_get_voxels (8 floats per point), grown on demand
    int64_t query_cap = 0;
    int carveout = 50;                // preferred shared-memory carveout (%) of the two pipeline kernels (GSDF_CARVEOUT)
    bool auto_grow = true;            // GSDF_AUTO_GROW=0: a full hash is an error instead of a reason to grow
    unsigned long long* d_count2 = nullptr;
    // pinned host staging
    float* h_out29 = nullptr;
    Counters* h_ctr = nullptr;
    Pose* h_pose = nullptr;
    FrameState* h_state = nullptr;
    unsigned long long* h_count2 = nullptr;
    std::vector<float> h_planes;
    // frame pipeline
    FrameSlot ring[kRing];
    cudaEvent_t ev_main = nullptr;    // orders the side streams after earlier synchronous-API work on `stream`
    int64_t seq_submitted = 0, seq_collected = 0;
    int sync_slot = 0;                // ring slot used by gsdf_upload_frame
    TrackParams scan_tp{25, 1e-6f, 1.f, 0};
    unsigned int poll_ns = 0;         // GSDF_POLL_NS: pause between polls of the tracker's grid-level exchange (measured: no effect)
    unsigned long long last_points = 0, last_updates = 0;  // cumulative counters at the last report
    int sm_count = 148;
    int64_t launches = 0;
    int64_t track_iterations = 0, track_valid = 0;
    bool profiling = false;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> prof_events[GSDF_PROF_KINDS];
    size_t prof_used[GSDF_PROF_KINDS] = {0, 0, 0, 0};
    // timeline tracing (off by default): labelled events on whichever stream the step runs on
    bool fuse_in_track = true;  // GSDF_FUSE_IN_TRACK=0: separate fusion launch instead of the tracker launch's tail
    bool merged_side = true;    // GSDF_MERGED_SIDE=0: compaction and normals as two launches
    bool serial = false;        // GSDF_SERIAL=1: frame-only kernels on the main stream instead of the side stream
    int batch_frames = kMaxBatch;  // GSDF_BATCH_FRAMES=1..8: frames per tracker launch in gsdf_scan_run
    unsigned long long* d_dbg = nullptr;  // tracker phase stamps (gsdf_debug_track_timers)
    int64_t dbg_only_seq = -1;            // >= 0: only the launch that holds this frame writes stamps (GSDF_DBG_SKIP)
    unsigned long long* d_flog = nullptr; // [kFlogFrames][8] kernel spans per pipeline frame (gsdf_debug_frame_log)
    // NVLink peer exchange of the sharded tracker (gsdf_p2p_export / gsdf_p2p_connect)
    void* d_p2p = nullptr;            // [4][world][32] 64-bit {tag, float} words, then one error word
    PeerExchange px{};
    bool p2p_connected = false;
    unsigned int p2p_launches = 0;
    void* p2p_peer_base[kMaxPeers] = {};
    bool tracing = false;
    std::vector<std::pair<std::string, cudaEvent_t>> trace;
    ncclComm_t comm = nullptr;
    std::string err;
};

namespace {

constexpr int kMaxDevices = 64;

#define CUDA_TRY(h, expr)                                                                      \
    do {                                                                                       \
        cudaError_t e_ = (expr);                                                               \
        if (e_ != cudaSuccess) {                                                               \
            (h)->err = std::string(#expr) + ": " + cudaGetErrorString(e_);                     \
            return GSDF_ERR_CUDA;                                                              \
        }                                                                                      \
    } while (0)

int reflect101_h(int p, int len) {
    if (len == 1) return 0;
    while (p < 0 || p >= len) p = p < 0 ? -p : 2 * (len - 1) - p;
    return p;
}

// Unnormalised window sums, REFLECT_101, double accumulation, rows then columns (cv::boxFilter semantics the
// reference relies on, normals/NormalEstimator.h:108-113).
void box_sum(const std::vector<double>& src, std::vector<double>& dst, int w, int h, int win) {
    const int r = win / 2;
    std::vector<double> rows((size_t)w * h);
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) {
            double s = 0.0;
            for (int d = -r; d <= r; ++d) s += src[(size_t)y * w + reflect101_h(x + d, w)];
            rows[(size_t)y * w + x] = s;
        }
    dst.resize((size_t)w * h);
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) {
            double s = 0.0;
            for (int d = -r; d <= r; ++d) s += rows[(size_t)reflect101_h(y + d, h) * w + x];
            dst[(size_t)y * w + x] = s;
        }
}

// cv::NormalEstimator<float>::cache (normals/NormalEstimator.h:81-154): one-time, double precision, host.
void build_normal_cache(const gsdf_config& c, std::vector<float>& P) {
    const int w = c.width, h = c.height;
    const size_t n = (size_t)w * h;
    const double fx_inv = 1. / (double)c.K[0], fy_inv = 1. / (double)c.K[4], cx = c.K[2], cy = c.K[5];
    std::vector<double> x0(n), y0(n), nsi(n), x0n(n), y0n(n), a11(n), a12(n), a22(n);
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) {
            const size_t i = (size_t)y * w + x;
            const double u = fx_inv * ((double)x - cx), v = fy_inv * ((double)y - cy);
            const double nsq = (1. + u * u) + v * v, inv = 1. / nsq;
            x0[i] = u; y0[i] = v; nsi[i] = inv; x0n[i] = u * inv; y0n[i] = v * inv;
            a11[i] = (u * u) * inv; a12[i] = (u * v) * inv; a22[i] = (v * v) * inv;
        }
    std::vector<double> M11, M12, M13, M22, M23, M33;
    box_sum(a11, M11, w, h, c.normal_window); box_sum(a12, M12, w, h, c.normal_window);
    box_sum(x0n, M13, w, h, c.normal_window); box_sum(a22, M22, w, h, c.normal_window);
    box_sum(y0n, M23, w, h, c.normal_window); box_sum(nsi, M33, w, h, c.normal_window);
    P.resize(n * 11);
    for (size_t i = 0; i < n; ++i) {
        const double det = (M11[i] * (M22[i] * M33[i]) + 2 * M12[i] * (M23[i] * M13[i])) -
                           ((M13[i] * (M13[i] * M22[i]) + M12[i] * (M12[i] * M33[i])) + M23[i] * (M23[i] * M11[i]));
        const double di = 1. / det;
        P[0 * n + i] = (float)x0[i]; P[1 * n + i] = (float)y0[i]; P[2 * n + i] = (float)x0n[i];
        P[3 * n + i] = (float)y0n[i]; P[4 * n + i] = (float)nsi[i];
        P[5 * n + i] = (float)(di * (M22[i] * M33[i] - M23[i] * M23[i]));
        P[6 * n + i] = (float)(di * (M13[i] * M23[i] - M12[i] * M33[i]));
        P[7 * n + i] = (float)(di * (M12[i] * M23[i] - M13[i] * M22[i]));
        P[8 * n + i] = (float)(di * (M11[i] * M33[i] - M13[i] * M13[i]));
        P[9 * n + i] = (float)(di * (M12[i] * M13[i] - M11[i] * M23[i]));
        P[10 * n + i] = (float)(di * (M11[i] * M22[i] - M12[i] * M12[i]));
    }
}

template <typename F>
gsdf_status by_variant(gsdf_map* h, F f) { return h->cfg.variant == GSDF_GRAD_SDF ? f(std::integral_constant<int, 0>()) : f(std::integral_constant<int, 1>()); }

gsdf_status check_launch(gsdf_map* h, const char* what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { h->err = std::string(what) + ": " + cudaGetErrorString(e); return GSDF_ERR_CUDA; }
    h->launches++;
    return GSDF_OK;
}

struct ProfScope {
    cudaStream_t st = nullptr;
    cudaEvent_t stop = nullptr;
    ProfScope(gsdf_map* h, int kind, cudaStream_t stream) : st(stream) {
        if (!h->profiling) return;
        auto& v = h->prof_events[kind];
        if (h->prof_used[kind] == v.size()) {
            cudaEvent_t a, b;
            cudaEventCreate(&a); cudaEventCreate(&b);
            v.push_back({a, b});
        }
        auto& pr = v[h->prof_used[kind]++];
        cudaEventRecord(pr.first, st);
        stop = pr.second;
    }
    ~ProfScope() { if (stop) cudaEventRecord(stop, st); }
};

void trace_mark(gsdf_map* h, const char* label, cudaStream_t stream) {
    if (!h->tracing) return;
    cudaEvent_t e;
    cudaEventCreate(&e);
    cudaEventRecord(e, stream);
    h->trace.push_back({label, e});
}

TrackIntr make_intr(const gsdf_map* h, int sampling) {
    TrackIntr ti;
    ti.fx_inv = 1.f / h->cfg.K[0]; ti.fy_inv = 1.f / h->cfg.K[4]; ti.cx = h->cfg.K[2]; ti.cy = h->cfg.K[5];
    ti.sampling = sampling;
    ti.ws = (h->cfg.width + sampling - 1) / sampling; ti.hs = (h->cfg.height + sampling - 1) / sampling;
    return ti;
}

size_t normals_smem(const gsdf_map* h, int tile_h = NT_H) {
    const int r = h->cfg.normal_window / 2, tw = NT_W + 2 * r, th = tile_h + 2 * r;
    return sizeof(double) * 3 * tw * th + sizeof(double) * 3 * th * NT_W;
}

// normals of `depth` into buffer `buf` on `stream`; skip_empty = false computes every tile (gsdf_compute_normals)
gsdf_status launch_normals(gsdf_map* h, const float* depth, int buf, cudaStream_t stream, bool skip_empty) {
    const int w = h->cfg.width, ht = h->cfg.height, r = h->cfg.normal_window / 2;
    const size_t n = (size_t)w * ht;
    dim3 grid((w + NT_W - 1) / NT_W, (ht + NT_H - 1) / NT_H), block(NT_W, NT_H);
    float* nrm = h->d_normals[buf];
    const float zmin = skip_empty ? h->geom.z_min : -INFINITY, zmax = skip_empty ? h->geom.z_max : INFINITY;
    ProfScope prof(h, GSDF_PROF_NORMALS, stream);
    normals_tile_kernel<<<grid, block, normals_smem(h), stream>>>(depth, h->d_planes + 2 * n, h->d_planes + 3 * n,
                                                                  h->d_planes + 4 * n, h->d_planes + 5 * n, nrm, nrm + n,
                                                                  nrm + 2 * n, w, ht, r, zmin, zmax);
    return check_launch(h, "normals_tile_kernel");
}

// compaction of a frame's z-valid pixels (deterministic, tile order) into d_vlist[buf], count into d_nvalid[buf]
gsdf_status launch_prepare(gsdf_map* h, const float* depth, int buf, cudaStream_t stream) {
    ProfScope prof(h, GSDF_PROF_PREPARE, stream);
    h->compact_epoch = (h->compact_epoch + 1) & 0x7FFFFu;  // 19 bits: flags hold {epoch, count < 2^13}
    compact_tiles_kernel<<<h->nseg, kCompactThreads, 0, stream>>>(depth, h->cfg.width, h->cfg.height, h->geom.z_min,
                                                                  h->geom.z_max, h->d_compact_flags, h->compact_epoch,
                                                                  h->d_vlist[buf], h->d_nvalid + buf);
    return check_launch(h, "compact_tiles_kernel");
}

// compaction + normals of a frame in one launch (the pipeline's frame-only work).  In the pipeline it runs in the shadow
// of the tracker: the tracker's 768 threads x 64 registers leave room for two of its 256-thread, 32-register CTAs per SM,
// which fill the issue slots the latency-bound tracker leaves idle; its CTAs live a few microseconds, so a cooperative
// tracker launch that arrives while more of them are resident (tracker absent) waits only that long.
gsdf_status launch_normals_compact(gsdf_map* h, const float* depth, int buf, cudaStream_t stream, int64_t seq = -1) {
    const int w = h->cfg.width, ht = h->cfg.height, r = h->cfg.normal_window / 2;
    const size_t n = (size_t)w * ht;
    dim3 grid((w + NT_W - 1) / NT_W, (ht + NC_H - 1) / NC_H), block(NT_W, NC_H);
    float* nrm = h->d_normals[buf];
    h->compact_epoch = (h->compact_epoch + 1) & 0x7FFFFu;
    normals_compact_kernel<<<grid, block, normals_smem(h, NC_H), stream>>>(
        depth, h->d_planes + 2 * n, h->d_planes + 3 * n, h->d_planes + 4 * n, h->d_planes + 5 * n, nrm, nrm + n,
        nrm + 2 * n, w, ht, r, h->geom.z_min, h->geom.z_max, h->d_compact_flags2, h->compact_epoch, h->d_vlist[buf],
        h->d_nvalid + buf, h->d_flog ? h->d_flog + 8 * ((seq < 0 ? h->seq_submitted : seq) % kFlogFrames) : nullptr);
    return check_launch(h, "normals_compact_kernel");
}

gsdf_status ensure_frame(gsdf_map* h) {
    if (!h->d_depth) { h->err = "no depth frame set"; return GSDF_ERR_NO_FRAME; }
    return GSDF_OK;
}

gsdf_status ensure_prepared(gsdf_map* h) {
    gsdf_status s = ensure_frame(h);
    if (s != GSDF_OK || h->prepared) return s;
    s = launch_prepare(h, h->d_depth, h->normals_buf, h->stream);
    if (s == GSDF_OK) h->prepared = true;
    return s;
}

gsdf_status ensure_normals(gsdf_map* h) {
    gsdf_status s = ensure_frame(h);
    if (s != GSDF_OK || h->normals_valid) return s;
    s = launch_normals(h, h->d_depth, h->normals_buf, h->stream, true);
    if (s == GSDF_OK) h->normals_valid = true;
    return s;
}

gsdf_status launch_fuse(gsdf_map* h, const Pose& pose, int pose_src, int nbuf, cudaStream_t stream,
                        ResultSnapshot* snap = nullptr) {
    const size_t np = (size_t)h->cfg.width * h->cfg.height;
    const int blocks = h->sm_count * 8;  // warps stride over (32-pixel chunk, sample group) work items
    const float* nrm = h->d_normals[nbuf];
    return by_variant(h, [&](auto V) {
        ProfScope prof(h, GSDF_PROF_FUSE, stream);
        fuse_list_kernel<decltype(V)::value><<<blocks, 256, 0, stream>>>(
            h->view, h->geom, pose, pose_src, h->d_vlist[nbuf], h->d_nvalid + nbuf, h->d_state, nrm, nrm + np,
            nrm + 2 * np, h->d_planes,
            h->d_planes + np, h->d_planes + 4 * np, h->d_ctr, snap);
        return check_launch(h, "fuse_list_kernel");
    });
}

// The Gauss-Newton loop from the pose in d_state for nb frames in ONE launch (frame i uses the frame-only buffers
// bufs[i] and leaves its result in snaps[i]); pose, converged flag and iteration count of the last frame stay in
// d_state.  fuse_mode 0/1/2 (never / always / iff converged) integrates each frame in the same launch.  Sharded maps
// without a peer connection take the per-iteration NCCL path below (nb = 1 only).
// The persistent tracker as a launchable function: grad-/base-sdf, 768 or 2 x 384 threads per SM, with or without the debug
// stamps (the stamped twin runs only while gsdf_debug_track_timers is armed: its stamp tests cost instructions in every unit).
const void* track_fn(int variant, int threads, bool stamps) {
    const bool two = threads == 384;
    if (variant == 0) {
        if (stamps) return two ? (const void*)track_kernel<0, 384, true> : (const void*)track_kernel<0, 768, true>;
        return two ? (const void*)track_kernel<0, 384, false> : (const void*)track_kernel<0, 768, false>;
    }
    if (stamps) return two ? (const void*)track_kernel<1, 384, true> : (const void*)track_kernel<1, 768, true>;
    return two ? (const void*)track_kernel<1, 384, false> : (const void*)track_kernel<1, 768, false>;
}

gsdf_status launch_track_batch(gsdf_map* h, const TrackParams& tp, const int* bufs, int nb, cudaStream_t stream, int fuse_mode,
                               ResultSnapshot* const* snaps, int64_t seq0);

gsdf_status launch_track(gsdf_map* h, const TrackParams& tp, int buf, cudaStream_t stream, int fuse_mode = 0,
                         ResultSnapshot* snap_out = nullptr) {
    return launch_track_batch(h, tp, &buf, 1, stream, fuse_mode, &snap_out, h->seq_submitted);
}

gsdf_status launch_track_batch(gsdf_map* h, const TrackParams& tp, const int* bufs, int nb, cudaStream_t stream, int fuse_mode,
                               ResultSnapshot* const* snaps, int64_t seq0) {
    const TrackIntr ti = make_intr(h, 1);
    const int buf = bufs[0];
    if (nb < 1 || nb > kMaxBatch) { h->err = "launch_track_batch: bad batch size"; return GSDF_ERR_INVALID; }
    if (h->cfg.world_size == 1 || h->p2p_connected) {
        return by_variant(h, [&](auto V) {
            ProfScope prof(h, GSDF_PROF_ACCUMULATE, stream);
            MapView view = h->view;
            FrameGeom geom = h->geom;
            TrackIntr ti_ = ti;
            FrameState* st = h->d_state;
            unsigned long long* rows = h->d_track_rows;
            unsigned int tag0 = h->track_tag;
            h->track_tag += (unsigned int)(nb * tp.max_iters);
            TrackParams tp_ = tp;
            tp_.poll_ns = h->poll_ns;
            const float* planes = h->d_planes;
            Counters* ctr = h->d_ctr;
            FrameBatch fb = {};
            fb.n = nb; fb.fuse_mode = fuse_mode;
            fb.dbg = (h->dbg_only_seq < 0 || (seq0 <= h->dbg_only_seq && h->dbg_only_seq < seq0 + nb)) ? h->d_dbg : nullptr;
            for (int i = 0; i < nb; ++i) {
                fb.vlist[i] = h->d_vlist[bufs[i]];
                fb.n_valid[i] = h->d_nvalid + bufs[i];
                fb.normals[i] = fuse_mode ? h->d_normals[bufs[i]] : nullptr;
                fb.snap[i] = snaps[i];
                fb.flog[i] = (h->d_flog && snaps[i]) ? h->d_flog + 8 * ((seq0 + i) % kFlogFrames) : nullptr;
            }
            PeerExchange px = h->px;  // world <= 1 unless gsdf_p2p_connect succeeded
            if (px.world > 1 && tp.max_iters > 63) { h->err = "sharded tracking over NVLink supports at most 63 iterations"; return GSDF_ERR_INVALID; }
            if (px.world > 1) { px.launch = ++h->p2p_launches; px.epoch0 = px.launch * 512u; }  // <= kMaxBatch * 63 passes per launch
            void* args[] = {&view, &geom, &ti_, &st, &rows, &tag0, &tp_, &planes, &ctr, &fb, &px};
            // cooperative launch: the driver guarantees that all CTAs are resident together (the exchange and grid_barrier
            // rely on it)
            const void* fn = track_fn(decltype(V)::value, h->track_threads, fb.dbg != nullptr);
            cudaError_t e = cudaLaunchCooperativeKernel(fn, dim3(h->track_blocks), dim3(h->track_threads), args, 0, stream);
            if (e != cudaSuccess) { h->err = std::string("launch of track_kernel: ") + cudaGetErrorString(e); return GSDF_ERR_CUDA; }
            return check_launch(h, "track_kernel");
        });
    }
    if (nb != 1) { h->err = "launch_track_batch: batches need one GPU or a peer-connected shard"; return GSDF_ERR_INVALID; }
    // sharded map: per-iteration launches with an all-reduce between the sums and the solve
    if (!h->comm) { h->err = "world_size > 1 but gsdf_comm_init was not called"; return GSDF_ERR_NCCL; }
    track_begin_kernel<<<1, 32, 0, stream>>>(h->d_state);
    gsdf_status s = check_launch(h, "track_begin_kernel");
    if (s != GSDF_OK) return s;
    const int batch = 4;
    for (int k0 = 0; k0 < tp.max_iters; k0 += batch) {
        for (int k = k0; k < k0 + batch && k < tp.max_iters; ++k) {
            s = by_variant(h, [&](auto V) {
                ProfScope prof(h, GSDF_PROF_ACCUMULATE, stream);
                accumulate_list_kernel<decltype(V)::value><<<h->acc_blocks, kAccThreads, 0, stream>>>(
                    h->view, h->geom, ti, h->d_vlist[buf], h->d_nvalid + buf, h->d_state, h->d_partials, h->d_ctr,
                    h->d_out29);
                return check_launch(h, "accumulate_list_kernel");
            });
            if (s != GSDF_OK) return s;
            const int rc = g_nccl.AllReduce(h->d_out29, h->d_out29, kAccTerms, kNcclFloat, kNcclSum, h->comm, stream);
            if (rc != 0) { h->err = std::string("ncclAllReduce: ") + (g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "?"); return GSDF_ERR_NCCL; }
            solve_kernel<<<1, 32, 0, stream>>>(h->d_state, h->d_out29, tp);
            s = check_launch(h, "solve_kernel");
            if (s != GSDF_OK) return s;
        }
        // every rank holds identical sums, hence identical `done` flags: all ranks leave the loop together
        CUDA_TRY(h, cudaMemcpyAsync(h->h_state, h->d_state, sizeof(FrameState), cudaMemcpyDeviceToHost, stream));
        CUDA_TRY(h, cudaStreamSynchronize(stream));
        if (h->h_state->done) break;
    }
    return GSDF_OK;
}

void sum_counters(const Counters& c, unsigned long long& points, unsigned long long& updates) {
    points = updates = 0;
    for (int i = 0; i < kCounterSlots; ++i) { points += c.slot[i].points; updates += c.slot[i].updates; }
}

gsdf_status clear_map(gsdf_map* h) {
    CUDA_TRY(h, cudaMemsetAsync(h->view.keys, 0xFF, sizeof(unsigned long long) * h->cap, h->stream));
    if (h->cfg.variant == GSDF_GRAD_SDF) {
        CUDA_TRY(h, cudaMemsetAsync(h->view.acc4, 0, sizeof(float4) * h->cap * kBlockVoxels, h->stream));
        CUDA_TRY(h, cudaMemsetAsync(h->view.accW, 0, sizeof(float) * h->cap * kBlockVoxels, h->stream));
    } else {
        CUDA_TRY(h, cudaMemsetAsync(h->view.acc2, 0, sizeof(float2) * h->cap * kBlockVoxels, h->stream));
    }
    CUDA_TRY(h, cudaMemsetAsync(h->d_ctr, 0, sizeof(Counters), h->stream));
    if (h->view.vis_stamp) CUDA_TRY(h, cudaMemsetAsync(h->view.vis_stamp, 0, sizeof(unsigned int) * h->cap * kBlockVoxels, h->stream));
    FrameState zero;
    std::memset(&zero, 0, sizeof(zero));
    zero.pose7[6] = 1.f;
    *h->h_state = zero;
    CUDA_TRY(h, cudaMemcpyAsync(h->d_state, h->h_state, sizeof(FrameState), cudaMemcpyHostToDevice, h->stream));
    CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    h->last_points = h->last_updates = 0;
    h->track_iterations = h->track_valid = 0;
    h->prepared = false;  // n_valid lives in the state that was just cleared
    return GSDF_OK;
}

void free_all(gsdf_map* h) {
    if (!h) return;
    cudaSetDevice(h->cfg.device);
    if (h->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(h->comm);
    for (int r = 0; r < kMaxPeers; ++r)
        if (h->p2p_peer_base[r] && r != h->cfg.rank) cudaIpcCloseMemHandle(h->p2p_peer_base[r]);
    cudaFree(h->d_p2p);
    cudaFree(h->view.keys); cudaFree(h->view.acc4); cudaFree(h->view.accW); cudaFree(h->view.acc2);
    cudaFree(h->view.vis_stamp); cudaFree(h->view.vis_log);
    cudaFree(h->d_planes); cudaFree(h->d_res);
    for (int i = 0; i < kRing; ++i) { cudaFree(h->d_normals[i]); cudaFree(h->d_vlist[i]); }
    cudaFree(h->d_nvalid); cudaFree(h->d_compact_flags); cudaFree(h->d_compact_flags2); cudaFree(h->d_track_rows);
    cudaFree(h->d_dbg); cudaFree(h->d_flog);
    cudaFree(h->d_partials); cudaFree(h->d_pose); cudaFree(h->d_out29); cudaFree(h->d_count2); cudaFree(h->d_query);
    cudaFreeHost(h->h_out29); cudaFreeHost(h->h_ctr); cudaFreeHost(h->h_pose); cudaFreeHost(h->h_count2);
    cudaFreeHost(h->h_state);
    for (auto& sl : h->ring) {
        cudaFree(sl.d_depth); cudaFree(sl.d_snap); cudaFreeHost(sl.h_snap);
        if (sl.ev_result) cudaEventDestroy(sl.ev_result);
        if (sl.ev_up) cudaEventDestroy(sl.ev_up);
        if (sl.ev_norm) cudaEventDestroy(sl.ev_norm);
        if (sl.ev_done) cudaEventDestroy(sl.ev_done);
    }
    for (auto& v : h->prof_events) for (auto& pr : v) { cudaEventDestroy(pr.first); cudaEventDestroy(pr.second); }
    if (h->ev_main) cudaEventDestroy(h->ev_main);
    if (h->stream) cudaStreamDestroy(h->stream);
    if (h->stream_n) cudaStreamDestroy(h->stream_n);
    if (h->stream_c) cudaStreamDestroy(h->stream_c);
    if (h->stream_r) cudaStreamDestroy(h->stream_r);
    delete h;
}

gsdf_status create_impl(gsdf_map* h) {
    const gsdf_config& c = h->cfg;
    CUDA_TRY(h, cudaSetDevice(c.device));
    cudaDeviceProp prop;
    CUDA_TRY(h, cudaGetDeviceProperties(&prop, c.device));
    h->sm_count = prop.multiProcessorCount;
    if (!prop.cooperativeLaunch) { h->err = "device lacks cooperative launch"; return GSDF_ERR_CUDA; }
    // the main stream carries the sequentially dependent work (track -> fuse -> next track): highest priority, so
    // that the frame-only work of the NEXT frame on the side stream fills the gaps instead of delaying it
    int prio_lo = 0, prio_hi = 0;
    CUDA_TRY(h, cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    CUDA_TRY(h, cudaStreamCreateWithPriority(&h->stream, cudaStreamNonBlocking, prio_hi));
    CUDA_TRY(h, cudaStreamCreateWithPriority(&h->stream_n, cudaStreamNonBlocking, prio_lo));
    CUDA_TRY(h, cudaStreamCreateWithPriority(&h->stream_c, cudaStreamNonBlocking, prio_lo));
    CUDA_TRY(h, cudaStreamCreateWithPriority(&h->stream_r, cudaStreamNonBlocking, prio_hi));
    const size_t n = (size_t)c.width * c.height;
    h->cap = (size_t)1 << c.log2_capacity;
    h->view.cap_mask = (uint32_t)(h->cap - 1);
    h->view.world = c.world_size; h->view.rank = c.rank;
    CUDA_TRY(h, cudaMalloc(&h->view.keys, sizeof(unsigned long long) * h->cap));
    if (c.variant == GSDF_GRAD_SDF) {
        CUDA_TRY(h, cudaMalloc(&h->view.acc4, sizeof(float4) * h->cap * kBlockVoxels));
        CUDA_TRY(h, cudaMalloc(&h->view.accW, sizeof(float) * h->cap * kBlockVoxels));
    } else {
        CUDA_TRY(h, cudaMalloc(&h->view.acc2, sizeof(float2) * h->cap * kBlockVoxels));
    }
    CUDA_TRY(h, cudaMalloc(&h->d_planes, sizeof(float) * n * 11));
    for (int i = 0; i < kRing; ++i) {
        CUDA_TRY(h, cudaMalloc(&h->d_normals[i], sizeof(float) * n * 3));
        CUDA_TRY(h, cudaMalloc(&h->d_vlist[i], sizeof(int2) * n));
    }
    CUDA_TRY(h, cudaMalloc(&h->d_nvalid, sizeof(int) * kRing));
    CUDA_TRY(h, cudaMemset(h->d_nvalid, 0, sizeof(int) * kRing));
    const int ntiles = ((c.width + kTileW - 1) / kTileW) * ((c.height + kTileH - 1) / kTileH);
    h->nseg = (ntiles + kTilesPerBlock - 1) / kTilesPerBlock;
    CUDA_TRY(h, cudaMalloc(&h->d_compact_flags, sizeof(unsigned int) * h->nseg));
    CUDA_TRY(h, cudaMemset(h->d_compact_flags, 0, sizeof(unsigned int) * h->nseg));
    const int n_ntiles = ((c.width + NT_W - 1) / NT_W) * ((c.height + NC_H - 1) / NC_H);
    CUDA_TRY(h, cudaMalloc(&h->d_compact_flags2, sizeof(unsigned long long) * n_ntiles));
    CUDA_TRY(h, cudaMemset(h->d_compact_flags2, 0, sizeof(unsigned long long) * n_ntiles));
    CUDA_TRY(h, cudaMalloc(&h->d_res, sizeof(DeviceResult)));
    h->d_state = &h->d_res->st;
    h->d_ctr = &h->d_res->ctr;
    h->acc_blocks = h->sm_count * 2;
    CUDA_TRY(h, cudaMalloc(&h->d_partials, sizeof(float) * kAccStride * h->sm_count * 4));
    // persistent tracker: 768 threads per SM (one CTA, or two of 384), all co-resident (cooperative launch)
    if (const char* e = std::getenv("GSDF_POLL_NS")) h->poll_ns = (unsigned int)std::max(0, std::atoi(e));
    if (const char* e = std::getenv("GSDF_TRACK_THREADS")) h->track_threads = std::atoi(e) == 384 ? 384 : 768;
    {
        const bool two = h->track_threads == 384;
        const int v = c.variant == GSDF_GRAD_SDF ? 0 : 1;
        const void* fn = track_fn(v, h->track_threads, false);
        // Shadow mode needs an SM whose L1/shared split already leaves room for the frame-only kernel's 116 KB when its CTA
        // arrives: the split is chosen when a kernel lands on an EMPTY SM, and the tracker alone would ask for a few KB, after
        // which no frame-only CTA fits beside it until the SM drains (measured: the two kernels then simply alternate).
        // 50 % of the maximum (a 132 KB shared / 124 KB L1 split): room for the tracker CTA and two frame-only CTAs, and an L1
        // that still holds the voxels a CTA re-reads pass after pass (measured 21.5 k frames/s against 21.1 k with the maximum
        // carveout; 40 % and 64 % fell back into the alternating state, profiles/r2t_carveout_sweep.txt)
        int carve = 50;
        if (const char* e = std::getenv("GSDF_CARVEOUT")) carve = std::min(100, std::max(0, std::atoi(e)));  // % of the maximum
        CUDA_TRY(h, cudaFuncSetAttribute(fn, cudaFuncAttributePreferredSharedMemoryCarveout, carve));
        CUDA_TRY(h, cudaFuncSetAttribute(track_fn(v, h->track_threads, true), cudaFuncAttributePreferredSharedMemoryCarveout, carve));
        h->carveout = carve;
        int per_sm = 0;
        CUDA_TRY(h, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, h->track_threads, 0));
        if (per_sm < 1) { h->err = "track_kernel does not fit on an SM"; return GSDF_ERR_CUDA; }
        h->track_blocks = h->sm_count * ((two && per_sm >= 2) ? 2 : 1);
    }
    CUDA_TRY(h, cudaMalloc(&h->d_track_rows, sizeof(unsigned long long) * 2 * kAccStride * h->track_blocks));
    CUDA_TRY(h, cudaMemset(h->d_track_rows, 0, sizeof(unsigned long long) * 2 * kAccStride * h->track_blocks));
    CUDA_TRY(h, cudaMalloc(&h->d_pose, sizeof(Pose)));
    CUDA_TRY(h, cudaMalloc(&h->d_out29, sizeof(float) * kAccStride));
    CUDA_TRY(h, cudaMalloc(&h->d_count2, sizeof(unsigned long long) * 4));
    CUDA_TRY(h, cudaHostAlloc(&h->h_out29, sizeof(float) * kAccStride, cudaHostAllocDefault));
    CUDA_TRY(h, cudaHostAlloc(&h->h_ctr, sizeof(Counters), cudaHostAllocDefault));
    CUDA_TRY(h, cudaHostAlloc(&h->h_pose, sizeof(Pose), cudaHostAllocDefault));
    CUDA_TRY(h, cudaHostAlloc(&h->h_state, sizeof(FrameState), cudaHostAllocDefault));
    CUDA_TRY(h, cudaHostAlloc(&h->h_count2, sizeof(unsigned long long) * 4, cudaHostAllocDefault));
    for (auto& sl : h->ring) {
        CUDA_TRY(h, cudaMalloc(&sl.d_depth, sizeof(float) * n));
        CUDA_TRY(h, cudaHostAlloc(&sl.h_snap, sizeof(ResultSnapshot), cudaHostAllocMapped));
        CUDA_TRY(h, cudaMalloc(&sl.d_snap, sizeof(ResultSnapshot)));
        CUDA_TRY(h, cudaEventCreateWithFlags(&sl.ev_result, cudaEventDisableTiming));
        CUDA_TRY(h, cudaEventCreateWithFlags(&sl.ev_up, cudaEventDisableTiming));
        CUDA_TRY(h, cudaEventCreateWithFlags(&sl.ev_norm, cudaEventDisableTiming));
        CUDA_TRY(h, cudaEventCreateWithFlags(&sl.ev_done, cudaEventDisableTiming));
    }
    CUDA_TRY(h, cudaEventCreateWithFlags(&h->ev_main, cudaEventDisableTiming));
    if (normals_smem(h) > 48 * 1024)
        CUDA_TRY(h, cudaFuncSetAttribute(normals_tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)normals_smem(h)));
    if (normals_smem(h, NC_H) > 48 * 1024)
        CUDA_TRY(h, cudaFuncSetAttribute(normals_compact_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)normals_smem(h, NC_H)));
    CUDA_TRY(h, cudaFuncSetAttribute(normals_compact_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, h->carveout));
    build_normal_cache(c, h->h_planes);
    CUDA_TRY(h, cudaMemcpyAsync(h->d_planes, h->h_planes.data(), sizeof(float) * n * 11, cudaMemcpyHostToDevice, h->stream));
    return clear_map(h);
}

// v1 reduction over the pixel grid (supports sampling): result lands in d_out29, pose taken from d_pose
gsdf_status accumulate_device(gsdf_map* h, int sampling, bool collective = true) {
    const TrackIntr ti = make_intr(h, sampling);
    const int total = ti.ws * ti.hs;
    int blocks = (total + kAccThreads - 1) / kAccThreads;
    if (blocks > h->sm_count * 4) blocks = h->sm_count * 4;
    gsdf_status s = by_variant(h, [&](auto V) {
        ProfScope prof(h, GSDF_PROF_ACCUMULATE, h->stream);
        accumulate_kernel<decltype(V)::value><<<blocks, kAccThreads, 0, h->stream>>>(
            h->view, h->geom, ti, h->d_pose, h->d_depth, h->d_partials, h->d_ctr, h->d_out29);
        return check_launch(h, "accumulate_kernel");
    });
    if (s != GSDF_OK) return s;
    if (h->cfg.world_size > 1 && collective) {
        if (!h->comm) { h->err = "world_size > 1 but gsdf_comm_init was not called"; return GSDF_ERR_NCCL; }
        const int rc = g_nccl.AllReduce(h->d_out29, h->d_out29, kAccTerms, kNcclFloat, kNcclSum, h->comm, h->stream);
        if (rc != 0) { h->err = std::string("ncclAllReduce: ") + (g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "?"); return GSDF_ERR_NCCL; }
    }
    return GSDF_OK;
}

// the overflow flag is sticky: once a sample was dropped the map no longer matches the reference key set
gsdf_status check_overflow(gsdf_map* h) {
    CUDA_TRY(h, cudaMemcpyAsync(h->h_ctr, h->d_ctr, sizeof(Counters), cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    if (h->h_ctr->overflow) {
        h->err = "voxel-block hash full: samples dropped; raise gsdf_config.log2_capacity";
        return GSDF_ERR_CAPACITY;
    }
    return GSDF_OK;
}

gsdf_status set_pose_device(gsdf_map* h, const float* R9, const float* t3) {
    std::memcpy(h->h_pose->R, R9, sizeof(float) * 9);
    std::memcpy(h->h_pose->t, t3, sizeof(float) * 3);
    CUDA_TRY(h, cudaMemcpyAsync(h->d_pose, h->h_pose, sizeof(Pose), cudaMemcpyHostToDevice, h->stream));
    return GSDF_OK;
}

// pose7 -> d_state.pose7 (stream-ordered; the rest of the state is left alone)
gsdf_status set_state_pose(gsdf_map* h, const float* pose7) {
    // pageable source would make the copy synchronous anyway; use the pinned pose buffer as staging
    std::memcpy(h->h_pose, pose7, sizeof(float) * 7);
    CUDA_TRY(h, cudaMemcpyAsync(h->d_state, h->h_pose, sizeof(float) * 7, cudaMemcpyHostToDevice, h->stream));
    return GSDF_OK;
}

// The reference's map never fills up; this one is rehashed into a table four times as large as soon as it is half full
// (called where the pipeline is idle).  A frame adds a few per cent of the blocks already there, so a table that starts with
// room for its first frame never drops a sample; GSDF_AUTO_GROW=0 keeps the capacity fixed (then a full table is an error).
gsdf_status maybe_grow(gsdf_map* h, int64_t blocks) {
    if (!h->auto_grow || blocks * 2 <= (int64_t)h->cap) return GSDF_OK;
    int lg = h->cfg.log2_capacity;
    while (lg < 26 && blocks * 4 > ((int64_t)1 << lg)) ++lg;
    if (lg == h->cfg.log2_capacity) return GSDF_OK;
    return gsdf_grow(h, lg);
}

gsdf_status pipeline_idle(gsdf_map* h) {
    if (h->seq_submitted != h->seq_collected) {
        h->err = "frames submitted with gsdf_scan_submit are still pending: collect them first";
        return GSDF_ERR_INVALID;
    }
    return GSDF_OK;
}

}  // namespace

extern "C" {

const char* gsdf_version(void) { return "gsdf_b200 0.2 (sm_100a)"; }

const char* gsdf_last_error(gsdf_handle h) { return h ? h->err.c_str() : g_create_error.c_str(); }

gsdf_status gsdf_create(const gsdf_config* cfg, gsdf_handle* out) {
    if (!cfg || !out) { g_create_error = "null argument"; return GSDF_ERR_INVALID; }
    *out = nullptr;
    gsdf_config c = *cfg;
    if (c.log2_capacity == 0) c.log2_capacity = 17;
    if (c.world_size <= 0) { c.world_size = 1; c.rank = 0; }
    if (c.normal_window == 0) c.normal_window = 11;
    if ((c.variant != GSDF_GRAD_SDF && c.variant != GSDF_BASE_SDF) || !(c.voxel_size > 0.f) || !(c.truncation > 0.f) ||
        c.width <= 0 || c.height <= 0 || c.normal_window < 1 || !(c.normal_window & 1) || c.normal_window / 2 > N_MAXR ||
        c.normal_window / 2 >= c.width || c.normal_window / 2 >= c.height || c.log2_capacity < 4 || c.log2_capacity > 26 ||
        c.rank < 0 || c.rank >= c.world_size || !(c.K[0] != 0.f) || !(c.K[4] != 0.f) || c.width > 65535 ||
        c.height > 32767) {
        g_create_error = "invalid gsdf_config";
        return GSDF_ERR_INVALID;
    }
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0 || c.device < 0 || c.device >= ndev || c.device >= kMaxDevices) {
        g_create_error = std::string("no usable CUDA device (there is no CPU fallback): ") + cudaGetErrorString(e);
        return GSDF_ERR_CUDA;
    }
    gsdf_map* h = new gsdf_map();
    h->cfg = c;
    if (const char* e = std::getenv("GSDF_BATCH_FRAMES")) h->batch_frames = std::min(kMaxBatch, std::max(1, std::atoi(e)));
    if (const char* e = std::getenv("GSDF_SERIAL")) h->serial = e[0] == '1';
    if (const char* e = std::getenv("GSDF_MERGED_SIDE")) h->merged_side = e[0] == '1';
    if (const char* e = std::getenv("GSDF_AUTO_GROW")) h->auto_grow = e[0] == '1';
    if (const char* e = std::getenv("GSDF_FUSE_IN_TRACK")) h->fuse_in_track = e[0] == '1';
    h->geom.w = c.width; h->geom.h = c.height;
    h->geom.vs = c.voxel_size;
    h->geom.vs_inv = (float)(1. / (double)c.voxel_size);  // MapGradPixelSdf.h:102
    h->geom.T = c.truncation;
    h->geom.inv_T = (float)(1. / (double)c.truncation);   // Sdf.h:105
    h->geom.z_min = c.z_min; h->geom.z_max = c.z_max;
    h->geom.factor = (int)std::floor(c.truncation / c.voxel_size);  // MapGradPixelSdfOmp.cpp:78
    // fusion counts its work units (32-pixel chunk x ray sample) in 32 bits
    if (((double)c.width * c.height / 32. + 1.) * (2. * h->geom.factor + 1.) >= 2147483648.) {
        g_create_error = "width * height * (2 * truncation / voxel_size + 1) is too large";
        free_all(h);
        return GSDF_ERR_INVALID;
    }
    gsdf_status s = create_impl(h);
    if (s != GSDF_OK) { g_create_error = h->err; free_all(h); return s; }
    *out = h;
    return GSDF_OK;
}

gsdf_status gsdf_destroy(gsdf_handle h) {
    if (!h) return GSDF_ERR_INVALID;
    cudaSetDevice(h->cfg.device);
    cudaStreamSynchronize(h->stream_c);
    cudaStreamSynchronize(h->stream_n);
    cudaStreamSynchronize(h->stream);
    cudaStreamSynchronize(h->stream_r);
    free_all(h);
    return GSDF_OK;
}

gsdf_status gsdf_reset(gsdf_handle h) {
    if (!h) return GSDF_ERR_INVALID;
    CUDA_TRY(h, cudaSetDevice(h->cfg.device));
    gsdf_status s = pipeline_idle(h);
    if (s != GSDF_OK) return s;
    CUDA_TRY(h, cudaStreamSynchronize(h->stream_n));
    return clear_map(h);
}

gsdf_status gsdf_get_normal_cache(gsdf_handle h, float* planes11) {
    if (!h || !planes11) return GSDF_ERR_INVALID;
    std::memcpy(planes11, h->h_planes.data(), sizeof(float) * h->h_planes.size());
    return GSDF_OK;
}

gsdf_status gsdf_upload_frame(gsdf_handle h, const float* depth_host) {
    if (!h || !depth_host) return GSDF_ERR_INVALID;
    CUDA_TRY(h, cudaSetDevice(h->cfg.device));
    gsdf_status s = pipeline_idle(h);
    if (s != GSDF_OK) return s;
    const size_t n = (size_t)h->cfg.width * h->cfg.height;
    float* dst = h->ring[h->sync_slot].d_depth;
    CUDA_TRY(h, cudaMemcpyAsync(dst, depth_host, sizeof(float) * n, cudaMemcpyHostToDevice, h->stream));
    h->d_depth = dst;
    h->normals_valid = false;
    h->prepared = false;
    return GSDF_OK;
}

gsdf_status gsdf_set_frame_device(gsdf_handle h, const float* depth_dev) {
    if (!h || !depth_dev) return GSDF_ERR_INVALID;
    if ((uintptr_t)depth_dev & 15) { h->err = "device frame must be 16-byte aligned"; return GSDF_ERR_INVALID; }
    gsdf_status s = pipeline_idle(h);
    if (s != GSDF_OK) return s;
    h->d_depth = depth_dev;
    h->normals_valid = false;
    h->prepared = false;
    return GSDF_OK;
}

gsdf_status gsdf_compute_normals(gsdf_handle h, float* nx, float* ny, float* nz) {
    if (!h || !nx || !ny || !nz) return GSDF_ERR_INVALID;
    CUDA_TRY(h, cudaSetDevice(h->cfg.device));
    gsdf_status s = ensure_frame(h);
    if (s != GSDF_OK) return s;
    // every tile, including those without a valid pixel (the pipeline skips them): compute() of the whole image
    s = launch_normals(h, h->d_depth, h->normals_buf, h->stream, false);
    if (s != GSDF_OK) return s;
    h->normals_valid = true;
    const size_t n = (size_t)h->cfg.width * h->cfg.height;
    const float* nrm = h->d_normals[h->normals_buf];
    CUDA_TRY(h, cudaMemcpyAsync(nx, nrm, sizeof(float) * n, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(h, cudaMemcpyAsync(ny, nrm + n, sizeof(float) * n, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(h, cudaMemcpyAsync(nz, nrm + 2 * n, sizeof(float) * n, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    return GSDF_OK;
}

gsdf_status gsdf_integrate(gsdf_handle h, const float* R9, const float* t3, gsdf_integrate_stats* stats) {
    if (!h || !R9 || !t3) return GSDF_ERR_INVALID;
    CUDA_TRY(h, cudaSetDevice(h->cfg.device));
    gsdf_status s = pipeline_idle(h);
    if (s != GSDF_OK) return s;
    s = ensure_prepared(h);
    if (s != GSDF_OK) return s;
    s = ensure_normals(h);  // MapGradPixelSdfOmp.cpp:59
    if (s != GSDF_OK) return s;
    Pose pose;
    std::memcpy(pose.R, R9, sizeof(float) * 9);
    std::memcpy(pose.t, t3, sizeof(float) * 3);
    s = launch_fuse(h, pose, 0, h->normals_buf, h->stream);
    if (s != GSDF_OK) return s;
    if (stats) {
        CUDA_TRY(h, cudaMemcpyAsync(h->h_ctr, h->d_ctr, sizeof(Counters), cudaMemcpyDeviceToHost, h->stream));
        CUDA_TRY(h, cudaStreamSynchronize(h->stream));
        unsigned long long pts, upd;
        sum_counters(*h->h_ctr, pts, upd);
        stats->points = (int64_t)(pts - h->last_points);
        stats->updates = (int64_t)(upd - h->last_updates);
        stats->blocks = (int64_t)h->h_ctr->blocks;
        h->last_points = pts;
        h->last_updates = upd;
        if (h->h_ctr->overflow) {
            h->err = "voxel-block hash full: samples dropped; raise gsdf_config.log2_capacity";
            return GSDF_ERR_CAPACITY;
        }
        return maybe_grow(h, stats->blocks);
    }
    return GSDF_OK;
}

gsdf_status gsdf_integrate_pose(gsdf_handle h, const float* pose7, gsdf_integrate_stats* stats) {
    if (!h || !pose7) return GSDF_ERR_INVALID;
    float R[9];
    quat_to_R(pose7 + 3, R);  // pose.rotationMatrix() of a stored SE3: no renormalisation
    return gsdf_integrate(h, R, pose7, stats);
}

gsdf_status gsdf_accumulate(gsdf_handle h, const float* R9, const float* t3, int32_t sampling, float* out29) {
    if (!h || !R9 || !t3 || !out29 || sampling < 1) return GSDF_ERR_INVALID;
    if (!h->d_depth) { h->err = "no depth frame set"; return GSDF_ERR_NO_FRAME; }
    CUDA_TRY(h, cudaSetDevice(h->cfg.device));
    gsdf_status s = pipeline_idle(h);
    if (s != GSDF_OK) return s;
    s = set_pose_device(h, R9, t3);
    if (s != GSDF_OK) return s;
    s = accumulate_device(h, sampling);
    if (s != GSDF_OK) return s;
    CUDA_TRY(h, cudaMemcpyAsync(h->h_out29, h->d_out29, sizeof(float) * kAccTerms, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    std::memcpy(out29, h->h_out29, sizeof(float) * kAccTerms);
    return GSDF_OK;
}

gsdf_status gsdf_accumulate_local(gsdf_handle h, const float* R9, const float* t3, int32_t sampling, float* out29) {
    if (!h || !R9 || !t3 || !out29 || sampling < 1) return GSDF_ERR_INVALID;
    if (!h->d_depth) { h->err = "no depth frame set"; return GSDF_ERR_NO_FRAME; }
    CUDA_TRY(h, cudaSetDevice(h->cfg.device));
    gsdf_status s = pipeline_idle(h);
    if (s != GSDF_OK) return s;
    s = set_pose_device(h, R9, t3);
    if (s != GSDF_OK) return s;
    s = accumulate_device(h, sampling, false);
    if (s != GSDF_OK) return s;
    CUDA_TRY(h, cudaMemcpyAsync(h->h_out29, h->d_out29, sizeof(float) * kAccTerms, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    std::memcpy(out29, h->h_out29, sizeof(float) * kAccTerms);
    return GSDF_OK;
}

gsdf_status gsdf_track(gsdf_handle h, float* pose7, int32_t max_iters, float conv_threshold, float damping,
                       int32_t* converged, int32_t* iters_used) {
    if (!h || !pose7 || !converged || max_iters < 0) return GSDF_ERR_INVALID;
    CUDA_TRY(h, cudaSetDevice(h->cfg.device));
    gsdf_status s = pipeline_idle(h);
    if (s != GSDF_OK) return s;
    *converged = 0;
    if (iters_used) *iters_used = 0;
    if (max_iters == 0) return ensure_frame(h);
    s = ensure_prepared(h);
    if (s != GSDF_OK) return s;
    s = set_state_pose(h, pose7);
    if (s != GSDF_OK) return s;
    TrackParams tp;
    tp.max_iters = max_iters;
    tp.thr_sq = conv_threshold * conv_threshold;  // RigidOptimizer.h:72
    tp.damping = damping;
    s = launch_track(h, tp, h->normals_buf, h->stream);
    if (s != GSDF_OK) return s;
    CUDA_TRY(h, cudaMemcpyAsync(h->h_state, h->d_state, sizeof(FrameState), cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    if (h->p2p_connected) {
        int perr = 0;
        CUDA_TRY(h, cudaMemcpy(&perr, h->px.error, sizeof(int), cudaMemcpyDeviceToHost));
        if (perr) { h->err = "sharded tracker: a peer rank did not reach the NVLink exchange within 0.5 s"; return GSDF_ERR_NCCL; }
    }
    std::memcpy(pose7, h->h_state->pose7, sizeof(float) * 7);
    *converged = h->h_state->converged;
    if (iters_used) *iters_used = h->h_state->iters;
    h->track_iterations = (int64_t)h->h_state->track_iters;
    h->track_valid = (int64_t)h->h_state->track_valid;
    return GSDF_OK;
}

namespace {
// staging buffer of 8 floats per point, kept between calls (Sdf::tsdf / weights / getSdf are called point by point by the
// reference's callers: no cudaMalloc per call)
gsdf_status query_staging(gsdf_map* h, int64_t n) {
    if (n <= h->query_cap) return GSDF_OK;
    CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    cudaFree(h->d_query);
    h->d_query = nullptr; h->query_cap = 0;
    const int64_t cap = std::max<int64_t>(n, 4096);
    CUDA_TRY(h, cudaMalloc(&h->d_query, sizeof(float) * 8 * cap));
    h->query_cap = cap;
    return GSDF_OK;
}
}  // namespace

gsdf_status gsdf_query(gsdf_handle h, const float* pts, int64_t n, float* phi, float* grad, float* w) {
    if (!h || n < 0 || (n > 0 && (!pts || !phi || !grad || !w))) return GSDF_ERR_INVALID;
    if (n == 0) return GSDF_OK;
    CUDA_TRY(h, cudaSetDevice(h->cfg.device));
    gsdf_status s = query_staging(h, n);
    if (s != GSDF_OK) return s;
    float *d_pts = h->d_query, *d_out = h->d_query + 3 * n;
    CUDA_TRY(h, cudaMemcpyAsync(d_pts, pts, sizeof(float) * 3 * n, cudaMemcpyHostToDevice, h->stream));
    const int blocks = (int)((n + 255) / 256);
    s = by_variant(h, [&](auto V) {
        query_kernel<decltype(V)::value><<<blocks, 256, 0, h->stream>>>(h->view, h->geom, d_pts, (long long)n, d_out, d_out + n, d_out + 4 * n);
        return check_launch(h, "query_kernel");
    });
    if (s != GSDF_OK) return s;
    CUDA_TRY(h, cudaMemcpyAsync(phi, d_out, sizeof(float) * n, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(h, cudaMemcpyAsync(grad, d_out + n, sizeof(float) * 3 * n, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(h, cudaMemcpyAsync(w, d_out + 4 * n, sizeof(float) * n, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    return GSDF_OK;
}

// getSdf(Vec3i) (MapGradPixelSdf.h:127-129, MapPixelSdf.h:123-125), batched: the voxels stored at the given keys
gsdf_status gsdf_get_voxels(gsdf_handle h, int64_t n, const int32_t* keys, float* dist, float* weight, float* grad) {
    if (!h || n < 0 || (n > 0 && (!keys || !dist || !weight || !grad))) return GSDF_ERR_INVALID;
    if (n == 0) return GSDF_OK;
    CUDA_TRY(h, cudaSetDevice(h->cfg.device));
    gsdf_status s = pipeline_idle(h);
    if (s != GSDF_OK) return s;
    s = query_staging(h, n);
    if (s != GSDF_OK) return s;
    int* d_keys = reinterpret_cast<int*>(h->d_query);
    float* d_out = h->d_query + 3 * n;
    CUDA_TRY(h, cudaMemcpyAsync(d_keys, keys, sizeof(int) * 3 * n, cudaMemcpyHostToDevice, h->stream));
    s = by_variant(h, [&](auto V) {
        get_voxels_kernel<decltype(V)::value><<<(unsigned)((n + 255) / 256), 256, 0, h->stream>>>(h->view, (long long)n, d_keys, d_out,
                                                                                              d_out + n, d_out + 2 * n);
        return check_launch(h, "get_voxels_kernel");
    });
    if (s != GSDF_OK) return s;
    CUDA_TRY(h, cudaMemcpyAsync(dist, d_out, sizeof(float) * n, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(h, cudaMemcpyAsync(weight, d_out + n, sizeof(float) * n, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(h, cudaMemcpyAsync(grad, d_out + 2 * n, sizeof(float) * 3 * n, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    return GSDF_OK;
}

// The reference's map grows without bound (tsdf_[vi], MapGradPixelSdfOmp.cpp:114); this one grows by rehashing into a larger
// table: new arrays, every block re-inserted and copied on the device, the old arrays freed.
gsdf_status gsdf_grow(gsdf_handle h, int32_t new_log2_capacity) {
    if (!h) return GSDF_ERR_INVALID;
    if (new_log2_capacity <= h->cfg.log2_capacity || new_log2_capacity > 26) { h->err = "gsdf_grow: new_log2_capacity must be larger than the current one and at most 26"; return GSDF_ERR_INVALID; }
    CUDA_TRY(h, cudaSetDevice(h->cfg.device));
    gsdf_status s = pipeline_idle(h);
    if (s != GSDF_OK) return s;
    CUDA_TRY(h, cudaStreamSynchronize(h->stream_n));
    CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    const size_t ncap = (size_t)1 << new_log2_capacity;
    MapView nv = h->view;
    nv.keys = nullptr; nv.acc4 = nullptr; nv.accW = nullptr; nv.acc2 = nullptr; nv.vis_stamp = nullptr;
    nv.cap_mask = (uint32_t)(ncap - 1);
    int* d_slot_map = nullptr;
    auto fail = [&](const char* what) {
        cudaFree(nv.keys); cudaFree(nv.acc4); cudaFree(nv.accW); cudaFree(nv.acc2); cudaFree(nv.vis_stamp); cudaFree(d_slot_map);
        h->err = std::string("gsdf_grow: ") + what;
        return GSDF_ERR_CUDA;
    };
    const bool grad = h->cfg.variant == GSDF_GRAD_SDF;
    if (cudaMalloc(&nv.keys, sizeof(unsigned long long) * ncap) != cudaSuccess) return fail("out of device memory (keys)");
    if (grad) {
        if (cudaMalloc(&nv.acc4, sizeof(float4) * ncap * kBlockVoxels) != cudaSuccess || cudaMalloc(&nv.accW, sizeof(float) * ncap * kBlockVoxels) != cudaSuccess)
            return fail("out of device memory (voxels)");
        cudaMemsetAsync(nv.acc4, 0, sizeof(float4) * ncap * kBlockVoxels, h->stream);
        cudaMemsetAsync(nv.accW, 0, sizeof(float) * ncap * kBlockVoxels, h->stream);
    } else {
        if (cudaMalloc(&nv.acc2, sizeof(float2) * ncap * kBlockVoxels) != cudaSuccess) return fail("out of device memory (voxels)");
        cudaMemsetAsync(nv.acc2, 0, sizeof(float2) * ncap * kBlockVoxels, h->stream);
    }
    cudaMemsetAsync(nv.keys, 0xFF, sizeof(unsigned long long) * ncap, h->stream);
    if (h->view.vis_stamp) {
        if (cudaMalloc(&nv.vis_stamp, sizeof(unsigned int) * ncap * kBlockVoxels) != cudaSuccess || cudaMalloc(&d_slot_map, sizeof(int) * h->cap) != cudaSuccess)
            return fail("out of device memory (visibility)");
        cudaMemsetAsync(nv.vis_stamp, 0, sizeof(unsigned int) * ncap * kBlockVoxels, h->stream);
    }
    const int grid = (int)std::min<size_t>(h->cap, (size_t)h->sm_count * 32);
    s = by_variant(h, [&](auto V) {
        rehash_kernel<decltype(V)::value><<<grid, 128, 0, h->stream>>>(h->view, nv, h->d_ctr, d_slot_map);
        return check_launch(h, "rehash_kernel");
    });
    if (s == GSDF_OK && h->view.vis_stamp) {
        if (cudaMemcpyAsync(h->h_ctr, h->d_ctr, sizeof(Counters), cudaMemcpyDeviceToHost, h->stream) != cudaSuccess ||
            cudaStreamSynchronize(h->stream) != cudaSuccess) return fail("reading the counters failed");
        const long long n = (long long)std::min<unsigned long long>(h->h_ctr->vis_cursor, h->view.vis_cap);
        if (n > 0) {
            vis_remap_kernel<<<(unsigned)((n + 255) / 256), 256, 0, h->stream>>>(h->view.vis_log, n, d_slot_map);
            s = check_launch(h, "vis_remap_kernel");
        }
    }
    if (s != GSDF_OK || cudaStreamSynchronize(h->stream) != cudaSuccess) return fail("rehash failed");
    cudaFree(h->view.keys); cudaFree(h->view.acc4); cudaFree(h->view.accW); cudaFree(h->view.acc2); cudaFree(h->view.vis_stamp);
    cudaFree(d_slot_map);
    h->view = nv;
    h->cap = ncap;
    h->cfg.log2_capacity = new_log2_capacity;
    return GSDF_OK;
}

gsdf_status gsdf_voxel_count(gsdf_handle h, int64_t* n_voxels, int64_t* n_blocks) {
    if (!h) return GSDF_ERR_INVALID;
    CUDA_TRY(h, cudaSetDevice(h->cfg.device));
    CUDA_TRY(h, cudaMemsetAsync(h->d_count2, 0, sizeof(unsigned long long) * 4, h->stream));
    gsdf_status s = by_variant(h, [&](auto V) {
        count_kernel<decltype(V)::value><<<h->sm_count * 8, 256, 0, h->stream>>>(h->view, h->d_count2, h->d_count2 + 1);
        return check_launch(h, "count_kernel");
    });
    if (s != GSDF_OK) return s;
    CUDA_TRY(h, cudaMemcpyAsync(h->h_count2, h->d_count2, sizeof(unsigned long long) * 2, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    if (n_voxels) *n_voxels = (int64_t)h->h_count2[0];
    if (n_blocks) *n_blocks = (int64_t)h->h_count2[1];
    return check_overflow(h);
}

gsdf_status gsdf_download(gsdf_handle h, int64_t max_n, int32_t* keys, float* dist, float* weight, float* grad,
                          int64_t* n_out) {
    if (!h || max_n < 0 || !n_out) return GSDF_ERR_INVALID;
    *n_out = 0;
    if (max_n == 0) return GSDF_OK;
    if (!keys || !dist || !weight || !grad) return GSDF_ERR_INVALID;
    CUDA_TRY(h, cudaSetDevice(h->cfg.device));
    int* d_keys = nullptr;
    float* d_f = nullptr;
    CUDA_TRY(h, cudaMalloc(&d_keys, sizeof(int) * 3 * max_n));
    if (cudaMalloc(&d_f, sizeof(float) * 5 * max_n) != cudaSuccess) { cudaFree(d_keys); h->err = "cudaMalloc(download)"; return GSDF_ERR_CUDA; }
    gsdf_status s = GSDF_OK;
    do {
        cudaMemsetAsync(h->d_count2 + 2, 0, sizeof(unsigned long long), h->stream);
        s = by_variant(h, [&](auto V) {
            download_kernel<decltype(V)::value><<<h->sm_count * 8, 256, 0, h->stream>>>(
                h->view, h->d_count2 + 2, (long long)max_n, d_keys, d_f, d_f + max_n, d_f + 2 * max_n);
            return check_launch(h, "download_kernel");
        });
        if (s != GSDF_OK) break;
        cudaMemcpyAsync(h->h_count2 + 2, h->d_count2 + 2, sizeof(unsigned long long), cudaMemcpyDeviceToHost, h->stream);
        if (cudaStreamSynchronize(h->stream) != cudaSuccess) { h->err = "download: stream sync failed"; s = GSDF_ERR_CUDA; break; }
        const int64_t total = (int64_t)h->h_count2[2];
        if (total > max_n) { h->err = "gsdf_download: max_n too small"; *n_out = total; s = GSDF_ERR_INVALID; break; }
        cudaMemcpyAsync(keys, d_keys, sizeof(int) * 3 * total, cudaMemcpyDeviceToHost, h->stream);
        cudaMemcpyAsync(dist, d_f, sizeof(float) * total, cudaMemcpyDeviceToHost, h->stream);
        cudaMemcpyAsync(weight, d_f + max_n, sizeof(float) * total, cudaMemcpyDeviceToHost, h->stream);
        cudaMemcpyAsync(grad, d_f + 2 * max_n, sizeof(float) * 3 * total, cudaMemcpyDeviceToHost, h->stream);
        if (cudaStreamSynchronize(h->stream) != cudaSuccess) { h->err = "download: copy failed"; s = GSDF_ERR_CUDA; break; }
        *n_out = total;
    } while (0);
    cudaFree(d_keys); cudaFree(d_f);
    return s;
}

gsdf_status gsdf_host_alloc(size_t bytes, void** out) {
    if (!out || bytes == 0) return GSDF_ERR_INVALID;
    *out = nullptr;
    if (cudaHostAlloc(out, bytes, cudaHostAllocDefault) != cudaSuccess) { g_create_error = "cudaHostAlloc failed"; *out = nullptr; return GSDF_ERR_CUDA; }
    return GSDF_OK;
}
gsdf_status gsdf_host_free(void* p) {
    if (!p) return GSDF_OK;
    return cudaFreeHost(p) == cudaSuccess ? GSDF_OK : GSDF_ERR_CUDA;
}

gsdf_status gsdf_upload(gsdf_handle h, int64_t n, const int32_t* keys, const float* dist, const float* weight, const float* grad) {
    if (!h || n < 0) return GSDF_ERR_INVALID;
    if (n == 0) return GSDF_OK;
    if (!keys || !dist || !weight || (h->cfg.variant == GSDF_GRAD_SDF && !grad)) return GSDF_ERR_INVALID;
    if (h->seq_submitted != h->seq_collected) { h->err = "gsdf_upload: frames are pending, collect them first"; return GSDF_ERR_INVALID; }
    CUDA_TRY(h, cudaSetDevice(h->cfg.device));
    int* d_keys = nullptr;
    float* d_f = nullptr;
    CUDA_TRY(h, cudaMalloc(&d_keys, sizeof(int) * 3 * n));
    if (cudaMalloc(&d_f, sizeof(float) * 5 * n) != cudaSuccess) { cudaFree(d_keys); h->err = "cudaMalloc(upload)"; return GSDF_ERR_CUDA; }
    gsdf_status s = GSDF_OK;
    do {
        cudaMemcpyAsync(d_keys, keys, sizeof(int) * 3 * n, cudaMemcpyHostToDevice, h->stream);
        cudaMemcpyAsync(d_f, dist, sizeof(float) * n, cudaMemcpyHostToDevice, h->stream);
        cudaMemcpyAsync(d_f + n, weight, sizeof(float) * n, cudaMemcpyHostToDevice, h->stream);
        if (grad) cudaMemcpyAsync(d_f + 2 * n, grad, sizeof(float) * 3 * n, cudaMemcpyHostToDevice, h->stream);
        s = by_variant(h, [&](auto V) {
            upload_kernel<decltype(V)::value><<<(unsigned)((n + 255) / 256), 256, 0, h->stream>>>(
                h->view, (long long)n, d_keys, d_f, d_f + n, d_f + 2 * n, h->d_ctr);
            return check_launch(h, "upload_kernel");
        });
        if (s != GSDF_OK) break;
        if (cudaStreamSynchronize(h->stream) != cudaSuccess) { h->err = "upload: stream sync failed"; s = GSDF_ERR_CUDA; break; }
        s = check_overflow(h);
    } while (0);
    cudaFree(d_keys); cudaFree(d_f);
    return s;
}

// Sdf::extract_mesh (MapGradPixelSdfOmp.cpp:132-183, MapPixelSdfOmp.cpp:197-248): marching cubes over the block hash on the
// device (gsdf_mesh.cuh), triangles returned in the reference's sweep order.
gsdf_status gsdf_extract_mesh(gsdf_handle h, int64_t max_triangles, float* vertices9, int64_t* n_triangles) {
    if (!h || !n_triangles || max_triangles < 0) return GSDF_ERR_INVALID;
    *n_triangles = 0;
    if (h->cfg.world_size > 1) { h->err = "gsdf_extract_mesh: single-GPU maps only (merge the shards with gsdf_download / gsdf_upload first)"; return GSDF_ERR_INVALID; }
    CUDA_TRY(h, cudaSetDevice(h->cfg.device));
    gsdf_status s = pipeline_idle(h);
    if (s != GSDF_OK) return s;
    int* d_box = nullptr;
    unsigned long long* d_keys = nullptr;
    float* d_verts = nullptr;
    CUDA_TRY(h, cudaMalloc(&d_box, sizeof(int) * 6 + sizeof(unsigned long long) * 2));
    unsigned long long* d_total = reinterpret_cast<unsigned long long*>(d_box + 6);
    const int grid = (int)std::min<size_t>(h->cap, (size_t)h->sm_count * 16);
    do {
        const int init[6] = {INT_MAX, INT_MAX, INT_MAX, INT_MIN, INT_MIN, INT_MIN};
        int box6[6];
        if (cudaMemcpyAsync(d_box, init, sizeof(init), cudaMemcpyHostToDevice, h->stream) != cudaSuccess ||
            cudaMemsetAsync(d_total, 0, sizeof(unsigned long long) * 2, h->stream) != cudaSuccess) { h->err = "extract_mesh: init failed"; s = GSDF_ERR_CUDA; break; }
        s = by_variant(h, [&](auto V) {
            mesh_bbox_kernel<decltype(V)::value><<<h->sm_count * 8, 256, 0, h->stream>>>(h->view, d_box);
            return check_launch(h, "mesh_bbox_kernel");
        });
        if (s != GSDF_OK) break;
        if (cudaMemcpyAsync(box6, d_box, sizeof(box6), cudaMemcpyDeviceToHost, h->stream) != cudaSuccess ||
            cudaStreamSynchronize(h->stream) != cudaSuccess) { h->err = "extract_mesh: bounding box failed"; s = GSDF_ERR_CUDA; break; }
        if (box6[0] == INT_MAX) break;  // empty map: no triangles
        MeshBox box;
        for (int c = 0; c < 3; ++c) { box.min[c] = box6[c]; box.dim[c] = box6[3 + c] - box6[c] + 1; }
        s = by_variant(h, [&](auto V) {
            mesh_kernel<decltype(V)::value, false><<<grid, 256, 0, h->stream>>>(h->view, box, h->geom.vs, d_total, 0, nullptr, nullptr);
            return check_launch(h, "mesh_kernel(count)");
        });
        if (s != GSDF_OK) break;
        unsigned long long total = 0;
        if (cudaMemcpyAsync(&total, d_total, sizeof(total), cudaMemcpyDeviceToHost, h->stream) != cudaSuccess ||
            cudaStreamSynchronize(h->stream) != cudaSuccess) { h->err = "extract_mesh: count failed"; s = GSDF_ERR_CUDA; break; }
        *n_triangles = (int64_t)total;
        if (!vertices9 || total == 0) break;  // count only
        if ((int64_t)total > max_triangles) { h->err = "gsdf_extract_mesh: max_triangles too small"; s = GSDF_ERR_INVALID; break; }
        if (cudaMalloc(&d_keys, sizeof(unsigned long long) * total) != cudaSuccess ||
            cudaMalloc(&d_verts, sizeof(float) * 9 * total) != cudaSuccess) { h->err = "extract_mesh: cudaMalloc failed"; s = GSDF_ERR_CUDA; break; }
        s = by_variant(h, [&](auto V) {
            mesh_kernel<decltype(V)::value, true><<<grid, 256, 0, h->stream>>>(h->view, box, h->geom.vs, d_total + 1, (long long)total,
                                                                               d_keys, d_verts);
            return check_launch(h, "mesh_kernel(emit)");
        });
        if (s != GSDF_OK) break;
        std::vector<unsigned long long> keys(total);
        std::vector<float> verts(9 * total);
        if (cudaMemcpyAsync(keys.data(), d_keys, sizeof(unsigned long long) * total, cudaMemcpyDeviceToHost, h->stream) != cudaSuccess ||
            cudaMemcpyAsync(verts.data(), d_verts, sizeof(float) * 9 * total, cudaMemcpyDeviceToHost, h->stream) != cudaSuccess ||
            cudaStreamSynchronize(h->stream) != cudaSuccess) { h->err = "extract_mesh: copy failed"; s = GSDF_ERR_CUDA; break; }
        // the reference's order: cells in its sweep order (z, y, x), triangles of a cell in case-table order
        std::vector<uint32_t> order(total);
        for (size_t i = 0; i < total; ++i) order[i] = (uint32_t)i;
        std::sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return keys[a] < keys[b]; });
        for (size_t i = 0; i < total; ++i) std::memcpy(vertices9 + 9 * i, verts.data() + 9 * (size_t)order[i], sizeof(float) * 9);
    } while (0);
    cudaFree(d_box); cudaFree(d_keys); cudaFree(d_verts);
    return s;
}

// vis_ (MapGradPixelSdf.h:70, MapGradPixelSdfOmp.cpp:118-120; read by PhotoBA, main_photo_ba.cpp:300-301): which fused
// frames touched which voxel.  Off by default (it costs 4 bytes per voxel slot plus 8 bytes per (voxel, frame) pair).
gsdf_status gsdf_vis_enable(gsdf_handle h, int64_t max_entries) {
    if (!h || max_entries <= 0) return GSDF_ERR_INVALID;
    if (h->view.vis_stamp) { h->err = "gsdf_vis_enable: already enabled"; return GSDF_ERR_INVALID; }
    CUDA_TRY(h, cudaSetDevice(h->cfg.device));
    gsdf_status s = pipeline_idle(h);
    if (s != GSDF_OK) return s;
    CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    unsigned int* stamp = nullptr;
    unsigned long long* log = nullptr;
    CUDA_TRY(h, cudaMalloc(&stamp, sizeof(unsigned int) * h->cap * kBlockVoxels));
    if (cudaMalloc(&log, sizeof(unsigned long long) * (size_t)max_entries) != cudaSuccess) { cudaFree(stamp); h->err = "gsdf_vis_enable: cudaMalloc failed"; return GSDF_ERR_CUDA; }
    CUDA_TRY(h, cudaMemset(stamp, 0, sizeof(unsigned int) * h->cap * kBlockVoxels));
    h->view.vis_stamp = stamp; h->view.vis_log = log; h->view.vis_cap = (unsigned long long)max_entries;
    return GSDF_OK;
}

gsdf_status gsdf_vis_download(gsdf_handle h, int64_t max_n, int32_t* keys, int32_t* frames, int64_t* n_out, int64_t* frames_fused) {
    if (!h || !n_out || max_n < 0) return GSDF_ERR_INVALID;
    *n_out = 0;
    if (!h->view.vis_stamp) { h->err = "gsdf_vis_download: call gsdf_vis_enable first"; return GSDF_ERR_INVALID; }
    CUDA_TRY(h, cudaSetDevice(h->cfg.device));
    gsdf_status s = pipeline_idle(h);
    if (s != GSDF_OK) return s;
    CUDA_TRY(h, cudaMemcpyAsync(h->h_ctr, h->d_ctr, sizeof(Counters), cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    if (frames_fused) *frames_fused = (int64_t)h->h_ctr->fused_frames;
    if (h->h_ctr->vis_overflow) { h->err = "visibility log full: entries dropped; raise max_entries of gsdf_vis_enable"; return GSDF_ERR_CAPACITY; }
    const int64_t n = (int64_t)h->h_ctr->vis_cursor;
    *n_out = n;
    if (n == 0 || !keys || !frames) return GSDF_OK;  // count only
    if (n > max_n) { h->err = "gsdf_vis_download: max_n too small"; return GSDF_ERR_INVALID; }
    int* d_k = nullptr;
    CUDA_TRY(h, cudaMalloc(&d_k, sizeof(int) * 4 * n));
    vis_decode_kernel<<<(unsigned)((n + 255) / 256), 256, 0, h->stream>>>(h->view, (long long)n, d_k, d_k + 3 * n);
    s = check_launch(h, "vis_decode_kernel");
    if (s == GSDF_OK) {
        if (cudaMemcpyAsync(keys, d_k, sizeof(int) * 3 * n, cudaMemcpyDeviceToHost, h->stream) != cudaSuccess ||
            cudaMemcpyAsync(frames, d_k + 3 * n, sizeof(int) * n, cudaMemcpyDeviceToHost, h->stream) != cudaSuccess ||
            cudaStreamSynchronize(h->stream) != cudaSuccess) { h->err = "gsdf_vis_download: copy failed"; s = GSDF_ERR_CUDA; }
    }
    cudaFree(d_k);
    return s;
}

// ---- the asynchronous frame pipeline ------------------------------------------------------------------------
gsdf_status gsdf_scan_set_pose(gsdf_handle h, const float* pose7) {
    if (!h || !pose7) return GSDF_ERR_INVALID;
    CUDA_TRY(h, cudaSetDevice(h->cfg.device));
    float R[9], pn[7];
    gsdf_pose_matrix(pose7, R, pn);  // SE3(quaternion, t): normalised once, like RigidOptimizer::set_pose(SE3)
    // stream-ordered after the frames already submitted; staging must outlive the copy -> synchronise the copy
    CUDA_TRY(h, cudaMemcpyAsync(h->d_state, pn, sizeof(float) * 7, cudaMemcpyHostToDevice, h->stream));
    CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    return GSDF_OK;
}

gsdf_status gsdf_scan_set_tracker(gsdf_handle h, int32_t max_iters, float conv_threshold, float damping) {
    if (!h || max_iters < 1) return GSDF_ERR_INVALID;
    h->scan_tp.max_iters = max_iters;
    h->scan_tp.thr_sq = conv_threshold * conv_threshold;
    h->scan_tp.damping = damping;
    return GSDF_OK;
}

namespace {

// can this handle run several frames per tracker launch?  (one GPU or a peer-connected shard, fused track+fuse launch,
// merged frame-only kernel, no per-kernel profiling / tracing / serial debugging mode)
bool batch_capable(const gsdf_map* h) {
    return (h->cfg.world_size == 1 || h->p2p_connected) && !h->profiling && !h->serial && !h->tracing && h->fuse_in_track &&
           h->merged_side;
}

// Enqueue nb frames: their uploads (copy stream), their frame-only kernels (side stream) and ONE tracker launch that
// works through them in order (main stream).  nb > 1 requires batch_capable() and GSDF_SCAN_TRACK_FUSE.  The caller has
// checked that nb free slots exist.
gsdf_status submit_frames(gsdf_map* h, const float* const* host, const float* const* dev, int nb, int32_t mode) {
    CUDA_TRY(h, cudaSetDevice(h->cfg.device));
    const size_t n = (size_t)h->cfg.width * h->cfg.height;
    const bool idle = h->seq_submitted == h->seq_collected;
    if (idle) {
        // pipeline was idle: whatever the synchronous entry points left running on the main stream may still use
        // the buffers the side streams are about to overwrite
        CUDA_TRY(h, cudaEventRecord(h->ev_main, h->stream));
        CUDA_TRY(h, cudaStreamWaitEvent(h->stream_n, h->ev_main, 0));
        CUDA_TRY(h, cudaStreamWaitEvent(h->stream_c, h->ev_main, 0));
    }
    // single cooperative launch for track + fuse: one GPU, or a sharded map whose ranks are connected over NVLink
    const bool single = (h->cfg.world_size == 1 || h->p2p_connected) && !h->profiling;
    // per-kernel profiling runs everything on one stream so that the event brackets do not overlap
    const bool serial = h->serial || h->profiling;
    cudaStream_t side = serial ? h->stream : h->stream_n;
    const int64_t seq0 = h->seq_submitted;
    int bufs[kMaxBatch];
    ResultSnapshot* snaps[kMaxBatch];
    const float* frame = nullptr;
    gsdf_status s = GSDF_OK;
    for (int i = 0; i < nb; ++i) {
        const int64_t seq = seq0 + i;
        FrameSlot& sl = h->ring[seq % kRing];
        frame = dev ? dev[i] : nullptr;
        if (host) {
            CUDA_TRY(h, cudaMemcpyAsync(sl.d_depth, host[i], sizeof(float) * n, cudaMemcpyHostToDevice, h->stream_c));
            CUDA_TRY(h, cudaEventRecord(sl.ev_up, h->stream_c));
            CUDA_TRY(h, cudaStreamWaitEvent(h->stream, sl.ev_up, 0));
            CUDA_TRY(h, cudaStreamWaitEvent(h->stream_n, sl.ev_up, 0));
            frame = sl.d_depth;
        }
        const int nbuf = (int)(seq % kRing);
        sl.direct = idle && nb == 1;
        bufs[i] = nbuf;
        snaps[i] = sl.direct ? sl.h_snap : sl.d_snap;
        // Pixel compaction and normals depend only on the frame, not on the map or the pose: they run on the side
        // stream, frames ahead of the main stream (which has the higher priority: the side kernels fill what the
        // tracker leaves idle instead of competing with it for the moment a frame ends).  Their buffers (index
        // seq % kRing) were last read by the tracker/fusion of frame seq - kRing, collected long ago.
        trace_mark(h, "side:begin", side);
        if (h->profiling || !h->merged_side) {  // separate launches (also keeps the per-kernel brackets meaningful)
            s = launch_prepare(h, frame, nbuf, side);
            if (s != GSDF_OK) return s;
            s = launch_normals(h, frame, nbuf, side, true);
        } else {
            s = launch_normals_compact(h, frame, nbuf, side, seq);
        }
        if (s != GSDF_OK) return s;
        trace_mark(h, "side:normals", side);
        if (!serial) {
            CUDA_TRY(h, cudaEventRecord(sl.ev_norm, side));
            CUDA_TRY(h, cudaStreamWaitEvent(h->stream, sl.ev_norm, 0));
        }
        sl.mode = mode;
    }
    trace_mark(h, "main:begin", h->stream);
    if (single && mode == GSDF_SCAN_TRACK_FUSE && h->fuse_in_track) {
        // track, fuse iff converged and leave the result, for every frame of the batch, in ONE cooperative launch
        s = launch_track_batch(h, h->scan_tp, bufs, nb, h->stream, 2, snaps, seq0);
        if (s != GSDF_OK) return s;
    } else {
        if (nb != 1) { h->err = "submit_frames: this configuration takes one frame per launch"; return GSDF_ERR_INVALID; }
        if (mode == GSDF_SCAN_TRACK_FUSE) {
            s = launch_track(h, h->scan_tp, bufs[0], h->stream);
            if (s != GSDF_OK) return s;
            trace_mark(h, "main:tracked", h->stream);
        }
        // fusion (predicated on the device-side converged flag); its last CTA publishes the frame result
        Pose unused;
        std::memset(&unused, 0, sizeof(unused));
        s = launch_fuse(h, unused, mode == GSDF_SCAN_TRACK_FUSE ? 2 : 1, bufs[0], h->stream, snaps[0]);
        if (s != GSDF_OK) return s;
    }
    trace_mark(h, "main:tracked+fused", h->stream);
    FrameSlot& last = h->ring[(seq0 + nb - 1) % kRing];
    CUDA_TRY(h, cudaEventRecord(last.ev_done, h->stream));
    if (!last.direct) {
        CUDA_TRY(h, cudaStreamWaitEvent(h->stream_r, last.ev_done, 0));
        for (int i = 0; i < nb; ++i) {
            FrameSlot& sl = h->ring[(seq0 + i) % kRing];
            CUDA_TRY(h, cudaMemcpyAsync(sl.h_snap, sl.d_snap, sizeof(ResultSnapshot), cudaMemcpyDeviceToHost, h->stream_r));
        }
        CUDA_TRY(h, cudaEventRecord(last.ev_result, h->stream_r));
    }
    for (int i = 0; i < nb; ++i) h->ring[(seq0 + i) % kRing].wait_ev = last.direct ? last.ev_done : last.ev_result;
    trace_mark(h, "main:result", h->stream);
    h->seq_submitted += nb;
    // the synchronous entry points see the last frame as current once the pipeline has drained
    h->d_depth = frame;
    h->normals_buf = bufs[nb - 1];
    h->normals_valid = true;
    h->prepared = true;
    return GSDF_OK;
}

}  // namespace

gsdf_status gsdf_scan_submit(gsdf_handle h, const float* depth_host, const float* depth_dev, int32_t mode) {
    if (!h || (!depth_host && !depth_dev) || (depth_host && depth_dev) || (mode != GSDF_SCAN_FUSE && mode != GSDF_SCAN_TRACK_FUSE))
        return GSDF_ERR_INVALID;
    if (depth_dev && ((uintptr_t)depth_dev & 15)) { h->err = "device frame must be 16-byte aligned"; return GSDF_ERR_INVALID; }
    if (h->seq_submitted - h->seq_collected >= kMaxPending) {
        h->err = "gsdf_scan_submit: 4 frames already in flight, call gsdf_scan_collect";
        return GSDF_ERR_INVALID;
    }
    return submit_frames(h, depth_host ? &depth_host : nullptr, depth_dev ? &depth_dev : nullptr, 1, mode);
}

gsdf_status gsdf_scan_collect(gsdf_handle h, gsdf_frame_result* out) {
    if (!h || !out) return GSDF_ERR_INVALID;
    if (h->seq_collected == h->seq_submitted) { h->err = "gsdf_scan_collect: nothing pending"; return GSDF_ERR_INVALID; }
    CUDA_TRY(h, cudaSetDevice(h->cfg.device));
    FrameSlot& sl = h->ring[h->seq_collected % kRing];
    CUDA_TRY(h, cudaEventSynchronize(sl.wait_ev));
    h->seq_collected++;
    const FrameState& st = sl.h_snap->st;
    const ResultSnapshot& c = *sl.h_snap;
    std::memcpy(out->pose7, st.pose7, sizeof(float) * 7);
    const bool tracked = sl.mode == GSDF_SCAN_TRACK_FUSE;
    out->converged = tracked ? st.converged : 1;
    out->iters = tracked ? st.iters : 0;
    out->fused = out->converged;
    out->n_valid = st.n_valid;
    const unsigned long long pts = c.points, upd = c.updates;
    out->points = (int64_t)(pts - h->last_points);
    out->updates = (int64_t)(upd - h->last_updates);
    out->blocks = (int64_t)c.blocks;
    out->track_iterations = (int64_t)st.track_iters - h->track_iterations;
    out->track_valid = (int64_t)st.track_valid - h->track_valid;
    h->last_points = pts;
    h->last_updates = upd;
    h->track_iterations = (int64_t)st.track_iters;
    h->track_valid = (int64_t)st.track_valid;
    if (c.overflow) {
        h->err = "voxel-block hash full: samples dropped; raise gsdf_config.log2_capacity";
        return GSDF_ERR_CAPACITY;
    }
    if (c.barrier_timeout) {
        h->err = "tracker grid barrier timed out: its CTAs were not all resident (another persistent kernel on this device?)";
        return GSDF_ERR_CUDA;
    }
    if (c.peer_error) {
        h->err = "sharded tracker: a peer rank did not reach the NVLink exchange within 0.5 s";
        return GSDF_ERR_NCCL;
    }
    if (h->seq_collected == h->seq_submitted) return maybe_grow(h, out->blocks);  // pipeline empty: the table may grow now
    return GSDF_OK;
}

int32_t gsdf_scan_pending(gsdf_handle h) { return h ? (int32_t)(h->seq_submitted - h->seq_collected) : 0; }

gsdf_status gsdf_scan_run(gsdf_handle h, const float* const* frames, int64_t n, int32_t on_device, int32_t mode,
                          gsdf_frame_result* results, int64_t* n_done) {
    if (!h || !frames || !results || n < 0 || (mode != GSDF_SCAN_FUSE && mode != GSDF_SCAN_TRACK_FUSE)) return GSDF_ERR_INVALID;
    if (n_done) *n_done = 0;
    if (h->seq_submitted != h->seq_collected) { h->err = "gsdf_scan_run: frames are pending, collect them first"; return GSDF_ERR_INVALID; }
    for (int64_t i = 0; i < n; ++i) {
        if (!frames[i]) { h->err = "gsdf_scan_run: null frame"; return GSDF_ERR_INVALID; }
        if (on_device && ((uintptr_t)frames[i] & 15)) { h->err = "device frame must be 16-byte aligned"; return GSDF_ERR_INVALID; }
    }
    // Knowing the frames ahead, the loop hands the tracker up to kMaxBatch frames per launch (one grid barrier between two
    // frames instead of leaving and relaunching the kernel) and keeps up to three such batches in flight: the batch being
    // tracked, and two whose uploads and frame-only kernels run ahead of it.
    const bool batches = batch_capable(h) && mode == GSDF_SCAN_TRACK_FUSE && h->batch_frames > 1;
    int64_t submitted = 0, collected = 0;
    int launches = 0;
    gsdf_status s = GSDF_OK;
    while (collected < n) {
        // half full: submit nothing new, so that the pipeline drains and gsdf_scan_collect grows the table
        const bool hold = h->auto_grow && collected > 0 && results[collected - 1].blocks * 2 > (int64_t)h->cap && submitted > collected;
        while (submitted < n && !hold) {
            // batch sizes ramp up 1, 2, 4, ...: the first tracker launch waits for ONE frame-only kernel, not for four
            const int ramp = std::min(h->batch_frames, 1 << std::min(launches, 3));
            const int nb = batches ? (int)std::min<int64_t>(ramp, n - submitted) : 1;
            if (submitted - collected + nb > (batches ? kRing : kMaxPending)) break;
            s = submit_frames(h, on_device ? nullptr : frames + submitted, on_device ? frames + submitted : nullptr, nb, mode);
            if (s != GSDF_OK) break;
            submitted += nb;
            ++launches;
        }
        if (s != GSDF_OK) break;
        s = gsdf_scan_collect(h, results + collected);
        if (s != GSDF_OK) break;
        ++collected;
        if (n_done) *n_done = collected;
    }
    if (s != GSDF_OK) {  // drain what is still in flight so that the handle stays usable; keep the first error
        const std::string first = h->err;
        gsdf_frame_result scratch;
        while (h->seq_collected < h->seq_submitted) gsdf_scan_collect(h, &scratch);
        h->err = first;
    }
    return s;
}

// ---- pose algebra ---------------------------------------------------------------------------------------------
void gsdf_pose_matrix(const float* pose7, float* R9, float* pose7_normalised) {
    float q[4] = {pose7[3], pose7[4], pose7[5], pose7[6]};
    quat_normalize(q);  // SE3(quaternion, t) constructor normalises
    quat_to_R(q, R9);
    if (pose7_normalised) {
        pose7_normalised[0] = pose7[0]; pose7_normalised[1] = pose7[1]; pose7_normalised[2] = pose7[2];
        for (int i = 0; i < 4; ++i) pose7_normalised[3 + i] = q[i];
    }
}

void gsdf_pose_apply_increment(const float* pose7_in, const float* xi6, float* pose7_out) {
    pose_apply_increment(pose7_in, xi6, pose7_out);
}

void gsdf_llt_solve6(const float* H36, const float* g6, float* xi6) { llt_solve6(H36, g6, xi6); }

gsdf_status gsdf_synchronize(gsdf_handle h) {
    if (!h) return GSDF_ERR_INVALID;
    CUDA_TRY(h, cudaSetDevice(h->cfg.device));
    CUDA_TRY(h, cudaStreamSynchronize(h->stream_c));
    CUDA_TRY(h, cudaStreamSynchronize(h->stream_n));
    return check_overflow(h);
}

}  // extern "C" (reopened below)

// ----------------------------------------------------------------------------------------------------------
// gsdf_accumulate_noalloc — the one entry point that never allocates.
//
// Who needs it: callers that must not (or cannot) reach the CUDA allocator when they call into the tracker —
// a process whose device heap has been handed over to another allocator (a framework's caching pool that
// already holds all free memory, so a cudaMalloc inside a library call would simply fail), a real-time scan
// loop that cannot afford cudaMalloc's implicit device-wide synchronisation, or a host that pre-budgets every
// byte of HBM per tenant.  gsdf_create reserves this handle's map and frame buffers up front, but the grid-level
// reduction of the normal-equation sums also needs per-CTA scratch; here that scratch is NOT taken from the
// handle: it is the file-scope array below, part of the library image, resident on every GPU that loads the
// library until the process exits (kStaticScratchFloats * 4 B = 512.25 KiB).  The caller passes the inputs
// (handle = map + current frame, pose, sampling) and the output buffer (29 floats in DEVICE memory).
//
// Limit: one CTA per 256 sampled pixels and at most kStaticMaxTiles (4096) CTAs, i.e. at most 1,048,576 sampled
// pixels per call (1280x720 at sampling 1 fits); larger frames are rejected with GSDF_ERR_INVALID.
// Set-up: the array's name is only a host-side stand-in for the device symbol, so the first call on each device
// resolves the real device address once with cudaGetSymbolAddress (that device current) and caches it.  That first
// call — and, under CUDA's lazy module loading, the first launch of this kernel for each map variant — may load
// the library's device code, which is driver work that can allocate and synchronise: make one warm-up call per
// device and variant (grad-sdf / base-sdf) at start-up, before any stream capture and before the loop that must
// not allocate.  After that a call issues one 4-byte memset and one kernel launch, nothing else.  All calls on a device share the one array and it keeps
// the previous call's contents: the arrival ticket is cleared on the stream before every launch (the partial
// sums are fully overwritten by each launch and need no clearing), and two calls whose kernels overlap on the
// GPU (different handles/streams on the same device) race on it — serialise them.
static __device__ __align__(256) float g_noalloc_scratch[kStaticScratchFloats];

namespace {
float* g_noalloc_ptr[kMaxDevices] = {};
std::mutex g_noalloc_mutex;
}  // namespace

extern "C" gsdf_status gsdf_accumulate_noalloc(gsdf_handle h, const float* R9, const float* t3, int32_t sampling,
                                               float* out29_dev) {
    if (!h || !R9 || !t3 || !out29_dev || sampling < 1) return GSDF_ERR_INVALID;
    if (!h->d_depth) { h->err = "no depth frame set"; return GSDF_ERR_NO_FRAME; }
    if (h->cfg.world_size > 1) { h->err = "gsdf_accumulate_noalloc: single-GPU maps only"; return GSDF_ERR_INVALID; }
    // like gsdf_accumulate: refused while pipeline frames are pending (they own the current frame and the map)
    if (gsdf_status idle = pipeline_idle(h); idle != GSDF_OK) return idle;
    const int dev = h->cfg.device;
    if (dev < 0 || dev >= kMaxDevices) { h->err = "gsdf_accumulate_noalloc: device ordinal out of range"; return GSDF_ERR_INVALID; }
    TrackIntr ti;
    ti.fx_inv = 1.f / h->cfg.K[0]; ti.fy_inv = 1.f / h->cfg.K[4]; ti.cx = h->cfg.K[2]; ti.cy = h->cfg.K[5];
    ti.sampling = sampling;
    ti.ws = (h->cfg.width + sampling - 1) / sampling; ti.hs = (h->cfg.height + sampling - 1) / sampling;
    const long long total = (long long)ti.ws * ti.hs;
    const long long tiles = (total + kAccThreads - 1) / kAccThreads;
    if (tiles > kStaticMaxTiles) {
        h->err = "gsdf_accumulate_noalloc: " + std::to_string(total) + " sampled pixels exceed the static scratch limit of " +
                 std::to_string((long long)kStaticMaxTiles * kAccThreads) + " (use gsdf_accumulate or a larger sampling)";
        return GSDF_ERR_INVALID;
    }
    CUDA_TRY(h, cudaSetDevice(dev));
    float* scratch = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_noalloc_mutex);
        if (!g_noalloc_ptr[dev]) {  // once per device, with that device current
            void* p = nullptr;
            CUDA_TRY(h, cudaGetSymbolAddress(&p, g_noalloc_scratch));
            g_noalloc_ptr[dev] = static_cast<float*>(p);
        }
        scratch = g_noalloc_ptr[dev];
    }
    Pose pose;
    std::memcpy(pose.R, R9, sizeof(float) * 9);
    std::memcpy(pose.t, t3, sizeof(float) * 3);
    CUDA_TRY(h, cudaMemsetAsync(scratch + kStaticTicketOffset, 0, sizeof(unsigned int), h->stream));
    return by_variant(h, [&](auto V) {
        accumulate_static_kernel<decltype(V)::value><<<(unsigned)tiles, kAccThreads, 0, h->stream>>>(
            h->view, h->geom, ti, pose, h->d_depth, scratch, out29_dev);
        return check_launch(h, "accumulate_static_kernel");
    });
}

extern "C" {

// Debug: %globaltimer stamps of the next gsdf_track's phases per CTA and pass: out[n_ctas][4][6] nanoseconds
// (0 lookups begin, 1 lookups end, 2 CTA reduce end, 3 grid barrier passed, 4 grid reduce end, 5 solve end).
// Then out[n_ctas][4] whole-launch stamps (entry, Gauss-Newton loop left, fusion done, exit) and out[n_ctas * 24][8]
// per-warp fusion stamps (see fuse_work; warp id = warp_in_cta * n_ctas + cta).
// Call once with out = NULL to arm (allocates), run gsdf_track, then call with a buffer of *n_ctas * 220 values.
gsdf_status gsdf_debug_track_timers(gsdf_handle h, unsigned long long* out, int32_t* n_ctas) {
    if (!h || !n_ctas) return GSDF_ERR_INVALID;
    CUDA_TRY(h, cudaSetDevice(h->cfg.device));
    *n_ctas = h->track_blocks;
    const size_t bytes = sizeof(unsigned long long) * h->track_blocks * (28 + 24 * 8);
    if (!h->d_dbg) {
        CUDA_TRY(h, cudaMalloc(&h->d_dbg, bytes));
        CUDA_TRY(h, cudaMemset(h->d_dbg, 0, bytes));
        // GSDF_DBG_SKIP=n: stamps only from the launch that holds the n-th frame submitted from now on (a launch in the
        // middle of a long gsdf_scan_run, with the frame-only kernels of later frames running beside it)
        const char* e = std::getenv("GSDF_DBG_SKIP");
        h->dbg_only_seq = e ? h->seq_submitted + std::max(0, std::atoi(e)) : -1;
    }
    if (out) {
        CUDA_TRY(h, cudaStreamSynchronize(h->stream));
        CUDA_TRY(h, cudaMemcpy(out, h->d_dbg, bytes, cudaMemcpyDeviceToHost));
        CUDA_TRY(h, cudaFree(h->d_dbg));
        h->d_dbg = nullptr;
    }
    return GSDF_OK;
}

gsdf_status gsdf_debug_frame_log(gsdf_handle h, unsigned long long* out /* null: arm; else kFlogFrames * 8 words */) {
    if (!h) return GSDF_ERR_INVALID;
    CUDA_TRY(h, cudaSetDevice(h->cfg.device));
    const size_t bytes = sizeof(unsigned long long) * 8 * kFlogFrames;
    if (!out) {
        if (!h->d_flog) CUDA_TRY(h, cudaMalloc(&h->d_flog, bytes));
        std::vector<unsigned long long> init(8 * kFlogFrames, 0ull);
        for (int i = 0; i < kFlogFrames; ++i) init[8 * i] = init[8 * i + 2] = ~0ull;
        CUDA_TRY(h, cudaDeviceSynchronize());
        CUDA_TRY(h, cudaMemcpy(h->d_flog, init.data(), bytes, cudaMemcpyHostToDevice));
        return GSDF_OK;
    }
    if (!h->d_flog) { h->err = "gsdf_debug_frame_log: not armed"; return GSDF_ERR_INVALID; }
    CUDA_TRY(h, cudaDeviceSynchronize());
    CUDA_TRY(h, cudaMemcpy(out, h->d_flog, bytes, cudaMemcpyDeviceToHost));
    CUDA_TRY(h, cudaFree(h->d_flog));
    h->d_flog = nullptr;
    return GSDF_OK;
}

gsdf_status gsdf_trace_enable(gsdf_handle h, int32_t on) {
    if (!h) return GSDF_ERR_INVALID;
    for (auto& t : h->trace) cudaEventDestroy(t.second);
    h->trace.clear();
    h->tracing = on != 0;
    return GSDF_OK;
}

// "label microseconds-since-first-mark" per line; returns the number of bytes needed (including the final NUL)
int64_t gsdf_trace_dump(gsdf_handle h, char* buf, int64_t size) {
    if (!h) return 0;
    cudaSetDevice(h->cfg.device);
    cudaDeviceSynchronize();
    std::string out;
    for (size_t i = 0; i < h->trace.size(); ++i) {
        float ms = 0.f;
        cudaEventElapsedTime(&ms, h->trace[0].second, h->trace[i].second);
        char line[96];
        std::snprintf(line, sizeof(line), "%s %.2f\n", h->trace[i].first.c_str(), ms * 1e3);
        out += line;
    }
    if (buf && size > 0) {
        const size_t n = std::min((size_t)size - 1, out.size());
        std::memcpy(buf, out.data(), n);
        buf[n] = 0;
    }
    return (int64_t)out.size() + 1;
}

gsdf_status gsdf_track_stats(gsdf_handle h, int64_t* iterations, int64_t* valid_pixels) {
    if (!h) return GSDF_ERR_INVALID;
    if (iterations) *iterations = h->track_iterations;
    if (valid_pixels) *valid_pixels = h->track_valid;
    return GSDF_OK;
}

gsdf_status gsdf_profile_enable(gsdf_handle h, int32_t on) {
    if (!h) return GSDF_ERR_INVALID;
    h->profiling = on != 0;
    for (int k = 0; k < GSDF_PROF_KINDS; ++k) h->prof_used[k] = 0;
    return GSDF_OK;
}

gsdf_status gsdf_profile_read(gsdf_handle h, int32_t kind, int64_t* launches, double* total_ms) {
    if (!h || kind < 0 || kind >= GSDF_PROF_KINDS || !launches || !total_ms) return GSDF_ERR_INVALID;
    CUDA_TRY(h, cudaSetDevice(h->cfg.device));
    CUDA_TRY(h, cudaStreamSynchronize(h->stream_n));
    CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    double tot = 0.0;
    for (size_t i = 0; i < h->prof_used[kind]; ++i) {
        float ms = 0.f;
        CUDA_TRY(h, cudaEventElapsedTime(&ms, h->prof_events[kind][i].first, h->prof_events[kind][i].second));
        tot += ms;
    }
    *launches = (int64_t)h->prof_used[kind];
    *total_ms = tot;
    return GSDF_OK;
}

void* gsdf_stream(gsdf_handle h) { return h ? (void*)h->stream : nullptr; }

int64_t gsdf_launch_count(gsdf_handle h) { return h ? h->launches : 0; }

// ---- NVLink peer exchange: the sharded tracker's all-reduce inside the cooperative launch --------------------------
static size_t p2p_bytes(int world) { return sizeof(unsigned long long) * 4 * world * kAccStride + 64; }

gsdf_status gsdf_p2p_export(gsdf_handle h, uint8_t handle64[64]) {
    if (!h || !handle64) return GSDF_ERR_INVALID;
    if (h->cfg.world_size < 2 || h->cfg.world_size > kMaxPeers) { h->err = "gsdf_p2p_export: world_size must be 2..8"; return GSDF_ERR_INVALID; }
    CUDA_TRY(h, cudaSetDevice(h->cfg.device));
    if (!h->d_p2p) {
        CUDA_TRY(h, cudaMalloc(&h->d_p2p, p2p_bytes(h->cfg.world_size)));
        CUDA_TRY(h, cudaMemset(h->d_p2p, 0, p2p_bytes(h->cfg.world_size)));
        CUDA_TRY(h, cudaDeviceSynchronize());
    }
    cudaIpcMemHandle_t mh;
    CUDA_TRY(h, cudaIpcGetMemHandle(&mh, h->d_p2p));
    static_assert(sizeof(mh) == 64, "cudaIpcMemHandle_t is 64 bytes");
    std::memcpy(handle64, &mh, 64);
    return GSDF_OK;
}

gsdf_status gsdf_p2p_connect(gsdf_handle h, const uint8_t* handles /* world_size * 64 bytes, in rank order */) {
    if (!h || !handles) return GSDF_ERR_INVALID;
    const int world = h->cfg.world_size, rank = h->cfg.rank;
    if (world < 2 || world > kMaxPeers || !h->d_p2p) { h->err = "gsdf_p2p_connect: call gsdf_p2p_export on every rank first"; return GSDF_ERR_INVALID; }
    CUDA_TRY(h, cudaSetDevice(h->cfg.device));
    for (int r = 0; r < world; ++r) {
        void* base = h->d_p2p;
        if (r != rank) {
            cudaIpcMemHandle_t mh;
            std::memcpy(&mh, handles + 64 * r, 64);
            CUDA_TRY(h, cudaIpcOpenMemHandle(&base, mh, cudaIpcMemLazyEnablePeerAccess));
        }
        h->p2p_peer_base[r] = base;
        h->px.data[r] = reinterpret_cast<unsigned long long*>(base);
    }
    h->px.world = world;
    h->px.rank = rank;
    h->px.error = reinterpret_cast<int*>(reinterpret_cast<char*>(h->d_p2p) + sizeof(unsigned long long) * 4 * world * kAccStride);
    h->p2p_connected = true;
    return GSDF_OK;
}

gsdf_status gsdf_comm_unique_id(uint8_t id128[128]) {
    std::string err;
    if (!id128) return GSDF_ERR_INVALID;
    if (!g_nccl.load(err)) { g_create_error = err; return GSDF_ERR_NCCL; }
    ncclUniqueId id;
    if (g_nccl.GetUniqueId(&id) != 0) { g_create_error = "ncclGetUniqueId failed"; return GSDF_ERR_NCCL; }
    std::memcpy(id128, id.internal, 128);
    return GSDF_OK;
}

gsdf_status gsdf_comm_init(gsdf_handle h, const uint8_t id128[128]) {
    if (!h || !id128) return GSDF_ERR_INVALID;
    if (h->cfg.world_size <= 1) return GSDF_OK;
    if (!g_nccl.load(h->err)) return GSDF_ERR_NCCL;
    CUDA_TRY(h, cudaSetDevice(h->cfg.device));
    ncclUniqueId id;
    std::memcpy(id.internal, id128, 128);
    const int rc = g_nccl.CommInitRank(&h->comm, h->cfg.world_size, id, h->cfg.rank);
    if (rc != 0) { h->err = std::string("ncclCommInitRank: ") + (g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "?"); return GSDF_ERR_NCCL; }
    return GSDF_OK;
}

}  // extern "C"
