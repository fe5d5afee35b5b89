/* gsdf_b200.h — C ABI of the B200-native Scan3D track+fuse hot path (libgsdf_b200.so).
 *
 * Drop-in boundary for c-sommer/gradient-sdf's depth-scanning path.  The reference has no FFI: its boundary is
 * the C++ class surface consumed by cpp/depth_scanning/src/main_scan_3d.cpp:208-266.  Each entry point below
 * names the reference interface it replaces (paths relative to /root/reference/cpp/include).  The host-side
 * forwarding shells with the reference's class names live in gradient-sdf_b200/host/; INTEGRATION.md shows the
 * binding a maintainer would add to the reference tree.
 *
 * Conventions: plain pointers and sizes only; every call returns a gsdf_status (0 = ok) and never throws;
 * host buffers are borrowed for the duration of the call; one handle <-> one host thread <-> one device and one
 * CUDA stream; a handle owns all its device memory.  Rotations are row-major 3x3 float, poses are 7 floats
 * "tx ty tz qx qy qz qw" (the reference's pose-file order, img_loader/ImageLoader.h:231-259), camera-to-world.
 * There is no CPU fallback: without a CUDA device gsdf_create fails with GSDF_ERR_CUDA.
 */
#ifndef GSDF_B200_H_
#define GSDF_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct gsdf_map* gsdf_handle;

typedef enum gsdf_status {
    GSDF_OK = 0,
    GSDF_ERR_INVALID = 1,   /* bad argument / null pointer / call order */
    GSDF_ERR_CUDA = 2,      /* CUDA runtime error, see gsdf_last_error */
    GSDF_ERR_CAPACITY = 3,  /* block hash full: samples would have been dropped (never silent) */
    GSDF_ERR_NCCL = 4,
    GSDF_ERR_NO_FRAME = 5   /* integrate/accumulate/track before a depth frame was set */
} gsdf_status;

enum { GSDF_GRAD_SDF = 0, GSDF_BASE_SDF = 1 };

/* Replaces the constructor arguments of MapGradPixelSdf(float voxel_size, float T) / MapPixelSdf(float, float)
 * (sdf_tracker/MapGradPixelSdf.h:99-103, MapPixelSdf.h:97-101), cv::NormalEstimator<float>(w, h, K, Size)
 * (normals/NormalEstimator.h:158-165) and the constants of Sdf (sdf_tracker/Sdf.h:63-68). */
typedef struct gsdf_config {
    int32_t variant;        /* GSDF_GRAD_SDF | GSDF_BASE_SDF (main_scan_3d.cpp --scan-type) */
    float voxel_size;       /* metres */
    float truncation;       /* T in metres = trunc_factor * voxel_size computed in float (main_scan_3d.cpp:191) */
    float z_min, z_max;     /* 0.5, 3.5 (Sdf.h:67-68) */
    int32_t width, height;  /* depth image size (main_scan_3d.cpp:183 hard-codes 640x480) */
    float K[9];             /* row-major intrinsics */
    int32_t normal_window;  /* 11 (main_scan_3d.cpp:183) */
    int32_t log2_capacity;  /* voxel-block hash slots = 2^log2_capacity blocks of 8^3 voxels; 0 -> default 17 */
    int32_t device;         /* CUDA device ordinal */
    int32_t world_size;     /* >1: map sharded by block key, this handle owns blocks with owner(key) == rank;
                             * owner(block) = (mix64(packed 21-bit block coords) >> 40) % world_size, see
                             * shard_owner() in gradient-sdf_b200/__init__.py for the host-side restatement */
    int32_t rank;
} gsdf_config;

typedef struct gsdf_integrate_stats {
    int64_t points;   /* the reference's "points integrated" counter (MapGradPixelSdfOmp.cpp:125-129) */
    int64_t updates;  /* (pixel,k) samples with w>0 committed by this rank = voxel updates */
    int64_t blocks;   /* voxel blocks allocated so far on this rank */
} gsdf_integrate_stats;

/* -- lifetime ------------------------------------------------------------------------------------------- */
gsdf_status gsdf_create(const gsdf_config* cfg, gsdf_handle* out);
gsdf_status gsdf_destroy(gsdf_handle h);
const char* gsdf_last_error(gsdf_handle h); /* h may be NULL: error of the last failed gsdf_create */
/* Empties the map (all voxels back to weight 0) without reallocating. */
gsdf_status gsdf_reset(gsdf_handle h);

/* -- cv::NormalEstimator<float> (normals/NormalEstimator.h) ------------------------------------------------ */
/* cache(): computed at gsdf_create in double on the host exactly as NormalEstimator.h:81-154 and uploaded.
 * gsdf_get_normal_cache returns the 11 float planes (x0, y0, x0/n^2, y0/n^2, 1/n^2, Q11, Q12, Q13, Q22, Q23,
 * Q33), i.e. what x0_ptr()/y0_ptr()/n_sq_inv_ptr() expose (NormalEstimator.h:206-210) plus the private ones. */
gsdf_status gsdf_get_normal_cache(gsdf_handle h, float* planes11);
/* compute(depth, nx, ny, nz) (NormalEstimator.h:179-204) on the current frame; planes of width*height floats. */
gsdf_status gsdf_compute_normals(gsdf_handle h, float* nx, float* ny, float* nz);

/* -- the depth frame (cv::Mat depth, CV_32FC1 metres, 0 = invalid, img_loader/ImageLoader.h:159-175) ------- */
/* Host buffer -> device (async on the handle's stream, complete before the next call that uses it). */
gsdf_status gsdf_upload_frame(gsdf_handle h, const float* depth_host);
/* Frame already resident in HBM (device pointer, 16-byte aligned, width*height floats, must stay valid while it
 * is current). */
gsdf_status gsdf_set_frame_device(gsdf_handle h, const float* depth_dev);

/* -- Sdf::update / Sdf::setup (sdf_tracker/Sdf.h:117-121; MapGradPixelSdfOmp.cpp:44-130,
 *    MapPixelSdfOmp.cpp:116-195): fuse the current frame at pose (R, t) = (pose.rotationMatrix(),
 *    pose.translation()).  stats may be NULL (then no device->host read happens). */
gsdf_status gsdf_integrate(gsdf_handle h, const float* R9, const float* t3, gsdf_integrate_stats* stats);
/* Same with the pose as a stored Sophus::SE3f (7 floats, UNIT quaternion, e.g. the output of gsdf_track or of
 * gsdf_pose_matrix): R = pose.rotationMatrix() without renormalising, as main_scan_3d.cpp:263 passes pOpt->pose(). */
gsdf_status gsdf_integrate_pose(gsdf_handle h, const float* pose7, gsdf_integrate_stats* stats);

/* -- RigidPointOptimizer (sdf_tracker/RigidOptimizer.h:90-110, RigidPointOptimizerOmp.cpp:44-106) ---------- */
/* One pass of the per-pixel loop (RigidPointOptimizerOmp.cpp:69-91) over the current frame at (R, t):
 * out29 = g[6], upper triangle of H row-major [21], E, counter.  With world_size > 1 the partials of all
 * ranks are summed (NCCL) so every rank receives the same totals. */
gsdf_status gsdf_accumulate(gsdf_handle h, const float* R9, const float* t3, int32_t sampling, float* out29);
/* This rank's partial sums only, no collective: for callers that reduce across shards themselves (any transport;
 * tests/test_sharding.py sums them with torch.distributed gloo) and for inspecting a shard.  No counterpart in the
 * reference (it has one map). */
gsdf_status gsdf_accumulate_local(gsdf_handle h, const float* R9, const float* t3, int32_t sampling, float* out29);
/* optimize(depth, K): Gauss-Newton from pose7 (updated in place) for at most max_iters iterations with the
 * reference's stopping rule |xi|^2 < conv_threshold^2 (last xi not applied) and NaN guard.
 * *converged = the bool optimize() returns; *iters_used = k at convergence (or max_iters). */
gsdf_status gsdf_track(gsdf_handle h, float* pose7, int32_t max_iters, float conv_threshold, float damping,
                       int32_t* converged, int32_t* iters_used);

/* No counterpart in the reference (it has no device and no allocator to avoid): gsdf_accumulate for callers
 * that cannot allocate device memory when they call — the CUDA heap is owned by another allocator, or the
 * call sits in a real-time loop where cudaMalloc's device-wide synchronisation is not acceptable.  It performs no
 * cudaMalloc/cudaFree and uses no handle-owned scratch: the reduction's per-CTA partial sums live in a static
 * 512 KiB device array inside the library.  Same sums as gsdf_accumulate, written to out29_dev (29 floats in
 * DEVICE memory, ready when the handle's stream reaches that point; gsdf_synchronize waits for it).
 * Limit: at most 1,048,576 sampled pixels (ceil(width/sampling)*ceil(height/sampling), the sampled grid of
 * RigidPointOptimizerOmp.cpp:70-71), else GSDF_ERR_INVALID; single-GPU maps only; refused (GSDF_ERR_INVALID) while
 * frames submitted with gsdf_scan_submit are pending, like gsdf_accumulate.
 * The first call on a device resolves the array's device address, and the first launch for each map variant may
 * load the library's device code (CUDA loads modules lazily; that driver work can allocate and synchronise): make
 * one warm-up call per device and variant at start-up, before any stream capture and before the loop that must
 * not allocate.  All calls on a device share the array: calls that overlap on the GPU race — keep them on one
 * stream. */
gsdf_status gsdf_accumulate_noalloc(gsdf_handle h, const float* R9, const float* t3, int32_t sampling,
                                    float* out29_dev);

/* -- Sdf::weights / Sdf::tsdf (sdf_tracker/Sdf.h:113-115; MapGradPixelSdf.h:109-125; MapPixelSdf.h:107-143,
 *    MapPixelSdfOmp.cpp:44-113), batched: per point w = weights(p); phi, grad = tsdf(p, &grad) if w > 0 else 0.
 *    pts n*3, phi n, grad n*3, w n (host pointers). */
gsdf_status gsdf_query(gsdf_handle h, const float* pts, int64_t n, float* phi, float* grad, float* w);

/* -- map access (getSdf/get_tsdf, MapGradPixelSdf.h:127-138; the writers iterate the map,
 *    MapGradPixelSdfOmp.cpp:144-151,189,285) ---------------------------------------------------------------- */
gsdf_status gsdf_voxel_count(gsdf_handle h, int64_t* n_voxels, int64_t* n_blocks);
/* Voxels with weight > 0 (this rank's shard): keys max_n*3 int32, dist, weight max_n, grad max_n*3 (zeros for
 * base-sdf).  *n_out = number written; GSDF_ERR_INVALID if max_n is too small.  Order unspecified. */
gsdf_status gsdf_download(gsdf_handle h, int64_t max_n, int32_t* keys, float* dist, float* weight, float* grad,
                          int64_t* n_out);

/* Sdf::extract_mesh (sdf_tracker/MapGradPixelSdfOmp.cpp:132-183, MapPixelSdfOmp.cpp:197-248 ->
 * mesh/LayeredMarchingCubesNoColor.cpp:354-719): marching cubes at iso value 0 over the voxels with weight > 0, with the
 * reference's rules (a cell is meshed only if all 8 corners have weight != 0; its corner / edge numbering, case table and
 * interpolate(); degenerate triangles dropped; vertices not shared between triangles).  vertices9 receives 9 floats per
 * triangle (three xyz points, world coordinates) in the reference's order — cells swept z, then y, then x over the map's
 * bounding box, triangles of a cell in case-table order — so that vertex i of the reference's PLY is point i here and face
 * f is (3f, 3f+1, 3f+2) (savePly, LayeredMarchingCubesNoColor.cpp:719-757).  vertices9 == NULL (or max_triangles == 0):
 * only counts.  *n_triangles = triangles in the map; GSDF_ERR_INVALID if max_triangles is smaller.  Single-GPU maps. */
gsdf_status gsdf_extract_mesh(gsdf_handle h, int64_t max_triangles, float* vertices9, int64_t* n_triangles);

/* vis_ — per-voxel frame visibility (sdf_tracker/MapGradPixelSdf.h:70,140-142 get_vis(); written by update(),
 * MapGradPixelSdfOmp.cpp:118-120: vis_[vi] gets `true` at index counter_, the number of update() calls before this one; read
 * only by PhotoBA, photometric_opt/src/main_photo_ba.cpp:300-301).  Off by default: gsdf_vis_enable reserves 4 bytes per voxel
 * slot and 8 bytes per (voxel, frame) entry, up to max_entries; call it before the frames whose visibility is wanted
 * (gsdf_reset clears the log and the frame counter).  gsdf_vis_download returns one entry per (voxel, frame) pair — keys
 * max_n*3, frames max_n: voxel key and the counter_ value of an update() that touched it — in no particular order, and
 * *frames_fused = counter_; the reference's std::vector<bool> of a voxel has `true` exactly at its entries' frames and size
 * 1 + the largest of them.  keys == NULL: only counts.  GSDF_ERR_CAPACITY if entries were dropped. */
gsdf_status gsdf_vis_enable(gsdf_handle h, int64_t max_entries);
gsdf_status gsdf_vis_download(gsdf_handle h, int64_t max_n, int32_t* keys, int32_t* frames, int64_t* n_out,
                              int64_t* frames_fused);

/* getSdf(Vec3i) (MapGradPixelSdf.h:127-129, MapPixelSdf.h:123-125), batched: dist, weight and grad (n*3, zeros for
 * base-sdf) of the voxels at keys (n*3); zeros where the voxel does not exist. */
gsdf_status gsdf_get_voxels(gsdf_handle h, int64_t n, const int32_t* keys, float* dist, float* weight, float* grad);

/* The reference's hash map grows without bound (tsdf_[vi], MapGradPixelSdfOmp.cpp:114).  Here the table is rehashed into a
 * larger one: gsdf_grow moves every block into a table of 2^new_log2_capacity slots on the device (new arrays are allocated,
 * the old ones freed; both exist for the duration of the call).  It happens by itself — to four times the blocks in use, as
 * soon as the table is half full, at the points where no frame is in flight (gsdf_integrate with stats, gsdf_scan_collect of
 * the last pending frame, inside gsdf_scan_run, which stops submitting and drains first) — unless GSDF_AUTO_GROW=0 is set
 * in the environment at gsdf_create; then, and when growth fails for lack of memory, or when one frame alone overflows the
 * table, a full table is GSDF_ERR_CAPACITY as before (never a silent drop). */
gsdf_status gsdf_grow(gsdf_handle h, int32_t new_log2_capacity);

/* Checkpoint / resume, no counterpart in the reference (which can only dump the final map as text, save_sdf): sets the
 * listed voxels to the given state — the arrays are what gsdf_download returns (grad may be NULL for base-sdf); voxels
 * with weight <= 0 are skipped.  Into an empty map (gsdf_reset) this restores a downloaded one: same key set, same
 * weight and grad, dist within one ulp (the map stores dist * weight).  A sharded map keeps only what its ownership
 * rule assigns to this rank, so every rank may be given the whole list.  GSDF_ERR_CAPACITY if the block hash fills up. */
gsdf_status gsdf_upload(gsdf_handle h, int64_t n, const int32_t* keys, const float* dist, const float* weight,
                        const float* grad);

/* Pinned (page-locked) host memory for frames handed to gsdf_scan_submit / gsdf_scan_run, so that their uploads are
 * asynchronous copies that overlap the GPU work; no counterpart in the reference (its frames live in cv::Mat).  Plain
 * wrappers of the CUDA runtime for hosts that link only this library. */
gsdf_status gsdf_host_alloc(size_t bytes, void** out);
gsdf_status gsdf_host_free(void* p);

/* -- the frame pipeline: main_scan_3d.cpp:208-266 without the host in the loop --------------------------------
 * The reference's loop body is "optimize(depth); if (converged) update(depth, pose())" with the pose living in the
 * optimizer object.  Here the pose, the converged flag and the frame's compacted pixel list live on the device, so a
 * whole frame (upload, pixel compaction, the complete Gauss-Newton loop, normals, conditional fusion, result
 * read-back) is enqueued by ONE call and up to 4 frames may be in flight: the host only feeds frames and collects
 * results, it never gates the GPU.  Frame i+1 is tracked from the pose frame i ended with (as pose_ persists in
 * RigidOptimizer, RigidOptimizer.h:66).  Do not mix with the synchronous calls while frames are pending. */
enum { GSDF_SCAN_FUSE = 0,        /* fuse at the current pose without tracking: setup() / ground-truth poses */
       GSDF_SCAN_TRACK_FUSE = 1   /* optimize, then update iff converged (main_scan_3d.cpp:258-264) */ };

typedef struct gsdf_frame_result {
    float pose7[7];            /* pose after this frame (RigidOptimizer::pose()) */
    int32_t converged;         /* optimize()'s return value (1 for GSDF_SCAN_FUSE frames) */
    int32_t iters;             /* k at convergence, or max_iters */
    int32_t fused;             /* whether the frame was integrated */
    int32_t n_valid;           /* pixels inside the depth range */
    int64_t points, updates;   /* this frame's "points integrated" and voxel updates */
    int64_t blocks;            /* blocks allocated so far */
    int64_t track_iterations;  /* Gauss-Newton passes run for this frame (iters + 1 when converged) */
    int64_t track_valid;       /* contributing pixels summed over those passes */
} gsdf_frame_result;

/* RigidOptimizer::set_pose (RigidOptimizer.h:104): pose7 goes through the SE3(quaternion, t) constructor. */
gsdf_status gsdf_scan_set_pose(gsdf_handle h, const float* pose7);
/* set_num_iterations / set_conv_threshold / set_damping (RigidOptimizer.h:90-102); defaults 25, 1e-3, 1. */
gsdf_status gsdf_scan_set_tracker(gsdf_handle h, int32_t max_iters, float conv_threshold, float damping);
/* Enqueue one frame: exactly one of depth_host (pinned memory recommended; copied on a separate copy stream, so
 * the upload of frame i+1 overlaps the work on frame i) or depth_dev (already in HBM, must stay valid until the
 * frame is collected).  Returns at once; GSDF_ERR_INVALID if 4 frames are already pending. */
gsdf_status gsdf_scan_submit(gsdf_handle h, const float* depth_host, const float* depth_dev, int32_t mode);
/* Wait for the oldest pending frame and return its result. */
gsdf_status gsdf_scan_collect(gsdf_handle h, gsdf_frame_result* out);
int32_t gsdf_scan_pending(gsdf_handle h);
/* The Scan3D frame loop (main_scan_3d.cpp:208-266) for n frames in one call; results[i] is filled in frame order
 * (the pipeline must be empty on entry and is empty on return).  Knowing the frames ahead, the loop gives the tracker
 * up to 8 frames per launch and keeps up to 24 frames in flight (uploads and frame-only kernels run ahead); the result
 * of every frame is the same as through gsdf_scan_submit / gsdf_scan_collect.  frames[i] are host pointers
 * (on_device = 0, pinned memory recommended) or 16-byte aligned device pointers (on_device = 1), all valid until the
 * call returns; mode as gsdf_scan_submit.  Stops at the first error (results of the frames collected so far are
 * valid, *n_done says how many). */
gsdf_status gsdf_scan_run(gsdf_handle h, const float* const* frames, int64_t n, int32_t on_device, int32_t mode,
                          gsdf_frame_result* results, int64_t* n_done);

/* -- pose algebra used by the tracker (Sophus::SE3f semantics: so3.hpp:297-303,325-339,356-366,583-619,
 *    se3.hpp:761-782; Eigen toRotationMatrix), exported so callers without Sophus get identical poses -------- */
void gsdf_pose_matrix(const float* pose7, float* R9, float* pose7_normalised /* may be NULL */);
void gsdf_pose_apply_increment(const float* pose7_in, const float* xi6, float* pose7_out); /* exp(-xi) * pose */
/* xi = LLT(H)^-1 g, H row-major 6x6 (RigidPointOptimizerOmp.cpp:93) */
void gsdf_llt_solve6(const float* H36, const float* g6, float* xi6);

/* -- plumbing ------------------------------------------------------------------------------------------- */
gsdf_status gsdf_synchronize(gsdf_handle h);
/* The handle's cudaStream_t (as void*), for CUDA-event timing and for ordering caller-owned device buffers. */
void* gsdf_stream(gsdf_handle h);
/* Kernel launches issued by this handle since creation (bench.py reports it as gpu_launches). */
int64_t gsdf_launch_count(gsdf_handle h);
/* Cumulative Gauss-Newton iterations run by gsdf_track on this handle and the pixels that contributed to them
 * (the reference's `counter`, RigidPointOptimizerOmp.cpp:66,89), for bytes-per-launch accounting. */
gsdf_status gsdf_track_stats(gsdf_handle h, int64_t* iterations, int64_t* valid_pixels);
/* Per-kernel device timing (the reference's only metrics are Timer tic/toc prints, Timer.h:59-78): when
 * enabled, a CUDA-event pair brackets every launch of that kind on the handle's stream; gsdf_profile_read
 * synchronises and returns the launches and their summed duration since gsdf_profile_enable(h, 1). */
enum { GSDF_PROF_NORMALS = 0, GSDF_PROF_FUSE = 1, GSDF_PROF_ACCUMULATE = 2 /* tracker */, GSDF_PROF_PREPARE = 3,
       GSDF_PROF_KINDS = 4 };
gsdf_status gsdf_profile_enable(gsdf_handle h, int32_t on);
gsdf_status gsdf_profile_read(gsdf_handle h, int32_t kind, int64_t* launches, double* total_ms);
/* Timeline tracing of the frame pipeline (the reference prints Timer tic/toc lines, Timer.h:59-78): when enabled,
 * gsdf_scan_submit drops labelled CUDA events on its streams; gsdf_trace_dump synchronises the device and writes
 * "label microseconds" lines (time since the first mark) into buf, returning the size needed. */
gsdf_status gsdf_trace_enable(gsdf_handle h, int32_t on);
int64_t gsdf_trace_dump(gsdf_handle h, char* buf, int64_t size);
/* Debug aid, no reference counterpart: per-CTA %globaltimer stamps of the tracker's phases and per-warp stamps of its fusion
 * (layout in gsdf_api.cu).  out == NULL arms it; while armed the tracker runs its stamped twin (track_kernel<V, NT, true>:
 * the stamp tests cost ~2 % of the frame rate, which is why the regular kernel does not carry them). */
gsdf_status gsdf_debug_track_timers(gsdf_handle h, unsigned long long* out, int32_t* n_ctas);
/* Debug aid, no reference counterpart: device-clock spans of the pipeline's two kernels per submitted frame.
 * out == NULL arms the log; otherwise out receives 256 x 8 words per frame (frame seq modulo 256): {frame-only kernel first
 * entry, last exit, track+fuse first entry, last exit, first Gauss-Newton pass done, Gauss-Newton loop left (fusion starts),
 * passes run, spare} in ns of %globaltimer (entry = ~0 / the others 0 where nothing ran), and the log is released. */
gsdf_status gsdf_debug_frame_log(gsdf_handle h, unsigned long long* out);
/* Multi-GPU: rank 0 calls gsdf_comm_unique_id, the 128 bytes are distributed by the caller (any transport),
 * then every rank calls gsdf_comm_init on its handle (cfg.world_size/rank).  Needed only for world_size > 1. */
gsdf_status gsdf_comm_unique_id(uint8_t id128[128]);
/* Optional, one node: the sharded tracker's all-reduce inside its cooperative launch, over NVLink peer memory (CUDA
 * IPC) instead of an NCCL launch per Gauss-Newton pass.  Every rank calls gsdf_p2p_export, the 64-byte handles are
 * gathered by the caller (any transport) and passed, in rank order, to gsdf_p2p_connect on every rank.  Afterwards
 * gsdf_track / gsdf_scan_submit run track + fuse as one launch per frame on every rank; all ranks must submit the
 * same frames in the same order (a rank that does not show up makes its peers fail with GSDF_ERR_NCCL after 0.5 s per
 * pass instead of hanging).  At most 63 iterations per optimize. */
gsdf_status gsdf_p2p_export(gsdf_handle h, uint8_t handle64[64]);
gsdf_status gsdf_p2p_connect(gsdf_handle h, const uint8_t* handles);
gsdf_status gsdf_comm_init(gsdf_handle h, const uint8_t id128[128]);
const char* gsdf_version(void);

#ifdef __cplusplus
}
#endif
#endif /* GSDF_B200_H_ */
