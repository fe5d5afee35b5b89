/* TEST INFRASTRUCTURE — CPU oracle for the Scan3D track+fuse hot path.  NOT part of the product:
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it.
 *
 * Plain-C, single-threaded restatement of the reference's algorithm (c-sommer/gradient-sdf @ 8e777231),
 * written from the reference's behaviour; every function cites the file:line it follows
 * (paths relative to /root/reference/cpp/include unless stated).
 *
 * Parity pin: tests/test_oracle_vs_ref.py checks this restatement BIT-FOR-BIT (keys, dist, weight, grad,
 * normals, cached planes, rotation matrices) and to 1e-6 (solve / pose algebra) against the reference's own
 * OMP build (oracle/_ref/libgsdf_ref.so, OMP_THREAD_LIMIT=1), and tests/golden/ holds outputs of that
 * reference build for the GPU box where /root/reference is absent.  The only unpinned link is OpenCV
 * itself (not installed): oracle/shim defines boxFilter/divide semantics for both.
 * Build: gcc -O2 -std=c11 -ffp-contract=off (no FMA contraction: the reference build has none).
 */
#ifndef GSDF_ORACLE_H_
#define GSDF_ORACLE_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct orc_ctx orc_ctx;

/* variant 0 = grad-sdf (MapGradPixelSdf), 1 = base-sdf (MapPixelSdf).  K row-major 3x3.
 * T = trunc_factor * voxel_size in float (depth_scanning/src/main_scan_3d.cpp:191). */
orc_ctx* orc_create(int variant, float voxel_size, float trunc_factor, int w, int h, const float* K9, int normal_win);
void orc_destroy(orc_ctx* c);
void orc_set_zrange(orc_ctx* c, float z_min, float z_max);
/* Emulation of the multi-GPU sharding protocol (no counterpart in the reference): this context commits only samples
 * whose 8^3 block it owns and its tracker counts only pixels whose nearest voxel it owns.  base-sdf: a pixel is
 * counted on the shard owning the block of its trilinear cell's floor corner, and that shard also keeps halo copies of
 * the low-face voxels of the +1 neighbour blocks (not counted, not downloaded). */
void orc_set_shard(orc_ctx* c, int world, int rank);
int orc_voxel_owner(const int32_t* v3, int world);

/* NormalEstimator::cache (normals/NormalEstimator.h:81-154): 11 planes of w*h floats, order
 * x0, y0, x0/n^2, y0/n^2, 1/n^2, Q11, Q12, Q13, Q22, Q23, Q33 */
void orc_normal_cache(const orc_ctx* c, float* out11);
/* NormalEstimator::compute (normals/NormalEstimator.h:179-204) */
void orc_normals(const orc_ctx* c, const float* depth, float* nx, float* ny, float* nz);

/* Map*PixelSdf::update (MapGradPixelSdfOmp.cpp:44-130, MapPixelSdfOmp.cpp:116-195), pixels row-major,
 * k ascending (= the OMP build with one thread).  R row-major, t translation (pose.rotationMatrix(),
 * pose.translation()).  Returns the number of (pixel,k) samples with w>0 ("voxel updates");
 * *n_points (optional) = the reference's "points integrated" counter. */
long long orc_update(orc_ctx* c, const float* depth, const float* R9, const float* t3, long long* n_points);

long long orc_voxel_count(const orc_ctx* c);
/* keys N*3, dist N, weight N, grad N*3 (zeros for base-sdf); table order (unspecified) */
void orc_download(const orc_ctx* c, int32_t* keys, float* dist, float* weight, float* grad);

/* per point: w = weights(p); phi/grad = tsdf(p,&grad) if w>0 else 0
 * (MapGradPixelSdf.h:109-125, MapPixelSdf.h:107-143 + MapPixelSdfOmp.cpp:44-113) */
void orc_query(const orc_ctx* c, const float* pts, long long n, float* phi, float* grad, float* w);

/* One pass of the tracker's per-pixel loop (RigidPointOptimizerOmp.cpp:70-91), sequential order.
 * out29 = g[6], H upper triangle row-major [21], E, counter. */
void orc_accumulate(const orc_ctx* c, const float* depth, const float* R9, const float* t3, int sampling, float* out29);

/* RigidPointOptimizer::optimize_sampled (RigidPointOptimizerOmp.cpp:44-106).  pose7 = tx ty tz qx qy qz qw,
 * updated in place.  Returns converged flag; *iters_used = k at convergence (or iters). */
int orc_optimize(const orc_ctx* c, const float* depth, float* pose7, int iters, float conv_threshold, float damping,
                 int sampling, int* iters_used);

/* host pose algebra (Sophus so3.hpp:297-303,325-339,356-366,583-619; se3.hpp:761-782; Eigen Quaternion
 * toRotationMatrix) */
void orc_pose_matrix(const float* pose7, float* R9, float* pose7_normalised);
void orc_apply_increment(const float* pose7_in, const float* xi6, float* pose7_out);
/* xi = LLT(H)^-1 g (RigidPointOptimizerOmp.cpp:93), H row-major 6x6; returns 0 if H is not positive definite
 * (xi then holds NaN, as Eigen's LLT propagates) */
int orc_llt_solve(const float* H36, const float* g6, float* xi6);

#ifdef __cplusplus
}
#endif
#endif
