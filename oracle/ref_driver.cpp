// TEST INFRASTRUCTURE — not part of the product.
//
// C entry points around the UNMODIFIED reference classes (compiled from /root/reference, see
// oracle/Makefile) so that tests, __graft_entry__.smoke() and bench.py's CPU-baseline legs can run
// the reference's own OMP implementation of the Scan3D hot path:
//   cv::NormalEstimator<float>            (cpp/include/normals/NormalEstimator.h:49-212)
//   MapGradPixelSdf / MapPixelSdf         (cpp/include/sdf_tracker/MapGradPixelSdfOmp.cpp:44-130,
//                                          MapPixelSdfOmp.cpp:44-195, *.h tsdf()/weights())
//   RigidPointOptimizer::optimize         (cpp/include/sdf_tracker/RigidPointOptimizerOmp.cpp:44-106)
// The per-frame ordering of main_scan_3d.cpp:208-266 (track, then fuse if converged) is driven by the
// caller (Python) through these calls.  Nothing here is linked into libgsdf_b200.so.
#include <algorithm>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <iostream>
#include <sstream>
#include <streambuf>
#include <string>
#include <vector>

#include <opencv2/core/core.hpp>
#include <opencv2/highgui/highgui.hpp>
#include <opencv2/imgproc/imgproc.hpp>
#include "mat.h"
#include "hash_map.h"
#include "sdf_voxel/SdfVoxel.h"

#include "normals/NormalEstimator.h"
#include "sdf_tracker/Sdf.h"
#include "sdf_tracker/MapGradPixelSdf.h"
#include "sdf_tracker/MapPixelSdf.h"
#include "sdf_tracker/RigidPointOptimizer.h"

// The cached FALS planes (Q11.., x0/n^2 ...) and MapPixelSdf's voxel map are private members with no
// getter, and the driver has to dump them.  Explicit template instantiation may name private members
// ([temp.spec]/6), which gives a standard-conforming way to obtain member pointers without touching
// the reference headers.
typedef phmap::parallel_node_hash_map<Vec3i, SdfVoxel, phmap::priv::hash_default_hash<Vec3i>,
                                      phmap::priv::hash_default_eq<Vec3i>,
                                      Eigen::aligned_allocator<std::pair<const Vec3i, SdfVoxel>>> VoxelMap;
typedef cv::NormalEstimator<float> NEstF;

template <typename Tag, typename Tag::type M>
struct MemberLeak {
    friend typename Tag::type leak(Tag) { return M; }
};
#define GSDF_LEAK(TAG, CLASS, TYPE, MEMBER)                 \
    struct TAG { typedef TYPE CLASS::*type; friend type leak(TAG); }; \
    template struct MemberLeak<TAG, &CLASS::MEMBER>;
GSDF_LEAK(GradMapTag, MapGradPixelSdf, VoxelMap, tsdf_)
GSDF_LEAK(BaseMapTag, MapPixelSdf, VoxelMap, tsdf_)
GSDF_LEAK(X0Tag, NEstF, cv::Mat_<float>, x0_)
GSDF_LEAK(Y0Tag, NEstF, cv::Mat_<float>, y0_)
GSDF_LEAK(X0nTag, NEstF, cv::Mat_<float>, x0_n_sq_inv_)
GSDF_LEAK(Y0nTag, NEstF, cv::Mat_<float>, y0_n_sq_inv_)
GSDF_LEAK(NTag, NEstF, cv::Mat_<float>, n_sq_inv_)
GSDF_LEAK(Q11Tag, NEstF, cv::Mat_<float>, Q11_)
GSDF_LEAK(Q12Tag, NEstF, cv::Mat_<float>, Q12_)
GSDF_LEAK(Q13Tag, NEstF, cv::Mat_<float>, Q13_)
GSDF_LEAK(Q22Tag, NEstF, cv::Mat_<float>, Q22_)
GSDF_LEAK(Q23Tag, NEstF, cv::Mat_<float>, Q23_)
GSDF_LEAK(Q33Tag, NEstF, cv::Mat_<float>, Q33_)

namespace {

struct NullBuf : std::streambuf {
    int overflow(int c) override { return c; }
};

// the reference prints per frame / per iteration; silence it while a call runs
struct Quiet {
    NullBuf nb;
    std::streambuf* old_out;
    explicit Quiet(bool on) : old_out(nullptr) { if (on) old_out = std::cout.rdbuf(&nb); }
    ~Quiet() { if (old_out) std::cout.rdbuf(old_out); }
};

struct RefScan {
    int variant;  // 0 = grad-sdf, 1 = base-sdf
    int w, h;
    Mat3f K;
    float voxel_size;
    cv::NormalEstimator<float>* nest;
    Sdf* sdf;
    RigidPointOptimizer* opt;
    bool quiet;
};

const VoxelMap& voxel_map(const RefScan* s) {
    if (s->variant == 0) return ((MapGradPixelSdf*)s->sdf)->*leak(GradMapTag());
    return ((MapPixelSdf*)s->sdf)->*leak(BaseMapTag());
}

SE3 pose_from7(const float* p) {
    // TUM order: tx ty tz qx qy qz qw  (ImageLoader.h:231-259)
    Eigen::Quaternionf q(p[6], p[3], p[4], p[5]);
    return SE3(q, Vec3f(p[0], p[1], p[2]));
}

void pose_to7(const SE3& T, float* p) {
    const Vec3f t = T.translation();
    const Eigen::Quaternionf& q = T.unit_quaternion();
    p[0] = t[0]; p[1] = t[1]; p[2] = t[2];
    p[3] = q.x(); p[4] = q.y(); p[5] = q.z(); p[6] = q.w();
}

void rot_to9(const SE3& T, float* R9) {
    const Mat3f R = T.rotationMatrix();
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) R9[3 * i + j] = R(i, j);
}

}  // namespace

extern "C" {

// K row-major 3x3.  trunc_factor as on the CLI (--trunc): T = trunc_factor * voxel_size in float
// (main_scan_3d.cpp:191).  normal_win = 11 in the reference (main_scan_3d.cpp:183).
void* gsdf_ref_create(int variant, float voxel_size, float trunc_factor, int w, int h, const float* K9,
                      int normal_win, int quiet) {
    RefScan* s = new RefScan();
    s->variant = variant; s->w = w; s->h = h; s->voxel_size = voxel_size; s->quiet = quiet != 0;
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) s->K(i, j) = K9[3 * i + j];
    s->nest = new cv::NormalEstimator<float>(w, h, s->K, cv::Size(normal_win, normal_win));
    const float truncation = trunc_factor * voxel_size;
    if (variant == 0) s->sdf = new MapGradPixelSdf(voxel_size, truncation);
    else s->sdf = new MapPixelSdf(voxel_size, truncation);
    s->opt = new RigidPointOptimizer(s->sdf);
    return s;
}

void gsdf_ref_destroy(void* h) {
    RefScan* s = (RefScan*)h;
    if (!s) return;
    delete s->opt; delete s->sdf; delete s->nest; delete s;
}

// 11 planes of w*h floats: x0, y0, x0/n^2, y0/n^2, 1/n^2, Q11, Q12, Q13, Q22, Q23, Q33
void gsdf_ref_normal_cache(void* h, float* out) {
    RefScan* s = (RefScan*)h;
    const size_t n = (size_t)s->w * s->h;
    NEstF& e = *s->nest;
    const cv::Mat* planes[11] = {&(e.*leak(X0Tag())), &(e.*leak(Y0Tag())), &(e.*leak(X0nTag())), &(e.*leak(Y0nTag())),
                                 &(e.*leak(NTag())), &(e.*leak(Q11Tag())), &(e.*leak(Q12Tag())), &(e.*leak(Q13Tag())),
                                 &(e.*leak(Q22Tag())), &(e.*leak(Q23Tag())), &(e.*leak(Q33Tag()))};
    for (int i = 0; i < 11; ++i) std::memcpy(out + i * n, planes[i]->data, n * sizeof(float));
}

void gsdf_ref_normals(void* h, const float* depth, float* nx, float* ny, float* nz) {
    RefScan* s = (RefScan*)h;
    cv::Mat d(s->h, s->w, CV_32FC1, (void*)depth), mx, my, mz;
    s->nest->compute(d, mx, my, mz);
    const size_t n = (size_t)s->w * s->h;
    std::memcpy(nx, mx.data, n * sizeof(float));
    std::memcpy(ny, my.data, n * sizeof(float));
    std::memcpy(nz, mz.data, n * sizeof(float));
}

// Sdf::update at the given pose (7 floats, TUM order).  R9_out (optional) receives the rotation
// matrix the reference used (pose.rotationMatrix(), row-major), t3_out the translation.
void gsdf_ref_update(void* h, const float* depth, const float* pose7, float* R9_out, float* t3_out) {
    RefScan* s = (RefScan*)h;
    Quiet q(s->quiet);
    cv::Mat d(s->h, s->w, CV_32FC1, (void*)depth), color;
    const SE3 T = pose_from7(pose7);
    if (R9_out) rot_to9(T, R9_out);
    if (t3_out) { const Vec3f t = T.translation(); t3_out[0] = t[0]; t3_out[1] = t[1]; t3_out[2] = t[2]; }
    s->sdf->update(color, d, s->K, T, s->nest);
}

void gsdf_ref_set_pose(void* h, const float* pose7) { ((RefScan*)h)->opt->set_pose(pose_from7(pose7)); }

void gsdf_ref_get_pose(void* h, float* pose7, float* R9_out) {
    RefScan* s = (RefScan*)h;
    const SE3 T = s->opt->pose();
    pose_to7(T, pose7);
    if (R9_out) rot_to9(T, R9_out);
}

void gsdf_ref_set_opt(void* h, int iters, float thr, float damping) {
    RefScan* s = (RefScan*)h;
    s->opt->set_num_iterations(iters);
    s->opt->set_conv_threshold(thr);
    s->opt->set_damping(damping);
}

// RigidPointOptimizer::optimize from the optimizer's current pose; returns the converged flag and,
// parsed from the reference's own "Convergence after k iterations" print, the iteration count (-1 if none).
int gsdf_ref_optimize(void* h, const float* depth, int* iters_out) {
    RefScan* s = (RefScan*)h;
    cv::Mat d(s->h, s->w, CV_32FC1, (void*)depth);
    std::ostringstream cap;
    std::streambuf* old = std::cout.rdbuf(cap.rdbuf());
    const bool conv = s->opt->optimize(d, s->K);
    std::cout.rdbuf(old);
    if (!s->quiet) std::cout << cap.str();
    if (iters_out) {
        *iters_out = -1;
        const std::string txt = cap.str();
        const size_t p = txt.find("after ");
        if (p != std::string::npos) *iters_out = std::atoi(txt.c_str() + p + 6);
    }
    return conv ? 1 : 0;
}

long long gsdf_ref_voxel_count(void* h) {
    RefScan* s = (RefScan*)h;
    return (long long)voxel_map(s).size();
}

// keys N*3 int32, dist N, weight N, grad N*3 (zeros for base-sdf), map iteration order
void gsdf_ref_download(void* h, int32_t* keys, float* dist, float* weight, float* grad) {
    RefScan* s = (RefScan*)h;
    size_t i = 0;
    auto emit = [&](const Vec3i& k, const SdfVoxel& v) {
        keys[3 * i] = k[0]; keys[3 * i + 1] = k[1]; keys[3 * i + 2] = k[2];
        dist[i] = v.dist; weight[i] = v.weight;
        grad[3 * i] = v.grad[0]; grad[3 * i + 1] = v.grad[1]; grad[3 * i + 2] = v.grad[2];
        ++i;
    };
    for (const auto& el : voxel_map(s)) emit(el.first, el.second);
}

// Per point: w = weights(p); if w > 0 then phi = tsdf(p, &grad) else phi = 0, grad = 0
// (the guard the tracker uses, RigidPointOptimizerOmp.cpp:80-83).
void gsdf_ref_query(void* h, const float* pts, long long n, float* phi, float* grad, float* w) {
    RefScan* s = (RefScan*)h;
    for (long long i = 0; i < n; ++i) {
        const Vec3f p(pts[3 * i], pts[3 * i + 1], pts[3 * i + 2]);
        const float w0 = s->sdf->weights(p);
        Vec3f g = Vec3f::Zero();
        float f = 0.f;
        if (w0 > 0) f = s->sdf->tsdf(p, &g);
        w[i] = w0; phi[i] = f;
        grad[3 * i] = g[0]; grad[3 * i + 1] = g[1]; grad[3 * i + 2] = g[2];
    }
}

// pose7 -> the rotation matrix Sophus/Eigen derive from it (row-major) and the normalised pose
void gsdf_ref_pose_matrix(const float* pose7, float* R9_out, float* pose7_norm_out) {
    const SE3 T = pose_from7(pose7);
    rot_to9(T, R9_out);
    if (pose7_norm_out) pose_to7(T, pose7_norm_out);
}

// pose_out = SE3::exp(-xi) * pose_in  (RigidPointOptimizerOmp.cpp:102), for checking the host algebra
void gsdf_ref_apply_increment(const float* pose7_in, const float* xi6, float* pose7_out) {
    Vec6f xi;
    for (int i = 0; i < 6; ++i) xi[i] = xi6[i];
    const SE3 T = SE3::exp(-xi) * pose_from7(pose7_in);
    pose_to7(T, pose7_out);
}

// xi = H.llt().solve(g) with Eigen (RigidPointOptimizerOmp.cpp:93), H row-major 6x6
void gsdf_ref_llt_solve(const float* H36, const float* g6, float* xi6) {
    Mat6f H; Vec6f g;
    for (int i = 0; i < 6; ++i) { g[i] = g6[i]; for (int j = 0; j < 6; ++j) H(i, j) = H36[6 * i + j]; }
    const Vec6f xi = H.llt().solve(g);
    for (int i = 0; i < 6; ++i) xi6[i] = xi[i];
}

// The reference's own output writers (MapGradPixelSdfOmp.cpp:132-303, MapPixelSdfOmp.cpp:197-351; marching cubes:
// mesh/LayeredMarchingCubesNoColor.cpp:354-757) on the map built so far: which = 0 extract_mesh, 1 extract_pc, 2 save_sdf.
int gsdf_ref_write(void* h, int which, const char* filename) {
    RefScan* s = (RefScan*)h;
    Quiet q(s->quiet);
    const std::string f(filename);
    if (which == 0) return s->sdf->extract_mesh(f) ? 1 : 0;
    if (which == 1) return s->sdf->extract_pc(f) ? 1 : 0;
    return s->sdf->save_sdf(f) ? 1 : 0;
}

// Upload a map (what gsdf_ref_download returns) into the reference's hash, so that its writers can be run on ANY map
// — e.g. the one the GPU built — and file-level differences are then differences of the writers alone.
void gsdf_ref_upload(void* h, long long n, const int32_t* keys, const float* dist, const float* weight, const float* grad) {
    RefScan* s = (RefScan*)h;
    VoxelMap& m = const_cast<VoxelMap&>(voxel_map(s));
    m.clear();
    for (long long i = 0; i < n; ++i) {
        SdfVoxel v;
        v.dist = dist[i]; v.weight = weight[i];
        v.grad = Vec3f(grad[3 * i], grad[3 * i + 1], grad[3 * i + 2]);
        m[Vec3i(keys[3 * i], keys[3 * i + 1], keys[3 * i + 2])] = v;
    }
}

// vis_ (MapGradPixelSdf.h:70, get_vis() :140-142): one (voxel key, frame index) pair per `true` bit; returns the number of
// pairs (call with keys == NULL to count); *max_size_out = the largest vector size (= 1 + last frame that touched a voxel).
long long gsdf_ref_vis(void* h, int32_t* keys, int32_t* frames, long long* max_size_out) {
    RefScan* s = (RefScan*)h;
    if (s->variant != 0) return 0;
    const auto vis = ((MapGradPixelSdf*)s->sdf)->get_vis();
    long long n = 0, mx = 0;
    for (const auto& el : vis) {
        mx = std::max<long long>(mx, (long long)el.second.size());
        for (size_t f = 0; f < el.second.size(); ++f)
            if (el.second[f]) {
                if (keys) { keys[3 * n] = el.first[0]; keys[3 * n + 1] = el.first[1]; keys[3 * n + 2] = el.first[2]; frames[n] = (int32_t)f; }
                ++n;
            }
    }
    if (max_size_out) *max_size_out = mx;
    return n;
}

}  // extern "C"
