// libgsdf_b200.so — C ABI (include/gsdf_b200.h) and host logic of the B200 Scan3D hot path.
// Host side: handle lifetime, the one-time FALS cache (cv::NormalEstimator::cache, double precision),
// pose algebra with Sophus::SE3f semantics, the Gauss-Newton driver loop, NCCL plumbing for sharded maps.
// Device side: gsdf_kernels.cuh.  There is no CPU implementation of the per-frame path in this library.
#include "../../include/gsdf_b200.h"

#include <cuda_runtime.h>
#include <dlfcn.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "gsdf_kernels.cuh"

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
}  // namespace

struct gsdf_map {
    gsdf_config cfg;
    FrameGeom geom;
    MapView view;
    size_t cap = 0;
    cudaStream_t stream = nullptr;
    // device buffers
    float* d_planes = nullptr;   // 11 cached planes
    float* d_depth_own = nullptr;
    const float* d_depth = nullptr;  // current frame (own buffer or caller's)
    float* d_normals = nullptr;  // 3 planes
    bool normals_valid = false;
    Counters* d_ctr = nullptr;
    float* d_partials = nullptr;
    int acc_blocks = 0;
    Pose* d_pose = nullptr;
    float* d_out29 = nullptr;
    // pinned host staging
    float* h_out29 = nullptr;
    Counters* h_ctr = nullptr;
    Pose* h_pose = nullptr;
    unsigned long long* d_count2 = nullptr;
    unsigned long long* h_count2 = nullptr;
    std::vector<float> h_planes;
    int sm_count = 148;
    int64_t launches = 0;
    int64_t track_iterations = 0, track_valid = 0;  // cumulative, for bytes-per-launch accounting
    // per-kernel CUDA-event timing (off by default): pairs recorded around each launch, resolved on read
    bool profiling = false;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> prof_events[GSDF_PROF_KINDS];
    size_t prof_used[GSDF_PROF_KINDS] = {0, 0, 0};
    ncclComm_t comm = nullptr;
    std::string err;
};

namespace {

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

// ---- pose algebra, Sophus::SE3f semantics (quaternion x,y,z,w + translation) ------------------------------
float dot3h(const float* a, const float* b) { const float p0 = a[0] * b[0], s = a[1] * b[1] + a[2] * b[2]; return p0 + s; }
void matvec3h(const float* R, const float* v, float* o) { o[0] = dot3h(R, v); o[1] = dot3h(R + 3, v); o[2] = dot3h(R + 6, v); }
void quat_normalize(float* q) {  // so3.hpp:297-303; Eigen packet sum pairs (x^2+z^2)+(y^2+w^2)
    const float len = std::sqrt((q[0] * q[0] + q[2] * q[2]) + (q[1] * q[1] + q[3] * q[3]));
    for (int i = 0; i < 4; ++i) q[i] /= len;
}
void quat_to_R(const float* q, float* R) {  // Eigen QuaternionBase::toRotationMatrix
    const float x = q[0], y = q[1], z = q[2], w = q[3];
    const float tx = 2.f * x, ty = 2.f * y, tz = 2.f * z;
    const float twx = tx * w, twy = ty * w, twz = tz * w;
    const float txx = tx * x, txy = ty * x, txz = tz * x, tyy = ty * y, tyz = tz * y, tzz = tz * z;
    R[0] = 1.f - (tyy + tzz); R[1] = txy - twz; R[2] = txz + twy;
    R[3] = txy + twz; R[4] = 1.f - (txx + tzz); R[5] = tyz - twx;
    R[6] = txz - twy; R[7] = tyz + twx; R[8] = 1.f - (txx + tyy);
}
void cross3(const float* a, const float* b, float* o) {
    o[0] = a[1] * b[2] - a[2] * b[1]; o[1] = a[2] * b[0] - a[0] * b[2]; o[2] = a[0] * b[1] - a[1] * b[0];
}
void quat_rotate(const float* q, const float* p, float* o) {  // so3.hpp:356-366
    float uv[3], c2[3];
    cross3(q, p, uv);
    for (int i = 0; i < 3; ++i) uv[i] += uv[i];
    cross3(q, uv, c2);
    for (int i = 0; i < 3; ++i) o[i] = (p[i] + q[3] * uv[i]) + c2[i];
}
void quat_mul(const float* a, const float* b, float* o) {  // so3.hpp:325-339, then SO3(quat) normalises
    const float ax = a[0], ay = a[1], az = a[2], aw = a[3], bx = b[0], by = b[1], bz = b[2], bw = b[3];
    o[3] = aw * bw - ax * bx - ay * by - az * bz;
    o[0] = aw * bx + ax * bw + ay * bz - az * by;
    o[1] = aw * by + ay * bw + az * bx - ax * bz;
    o[2] = aw * bz + az * bw + ax * by - ay * bx;
    quat_normalize(o);
}
void se3_exp(const float* a, float* q, float* tr) {  // se3.hpp:761-782, so3.hpp:583-619
    const float* om = a + 3;
    const float eps = 1e-5f;
    const float theta_sq = dot3h(om, om);
    float theta, imag, real;
    if (theta_sq < eps * eps) {
        theta = 0.f;
        const float p4 = theta_sq * theta_sq;
        imag = 0.5f - (float)(1.0 / 48.0) * theta_sq + (float)(1.0 / 3840.0) * p4;
        real = 1.f - (float)(1.0 / 8.0) * theta_sq + (float)(1.0 / 384.0) * p4;
    } else {
        theta = std::sqrt(theta_sq);
        const float half = 0.5f * theta;
        imag = std::sin(half) / theta;
        real = std::cos(half);
    }
    q[0] = imag * om[0]; q[1] = imag * om[1]; q[2] = imag * om[2]; q[3] = real;
    const float O[9] = {0.f, -om[2], om[1], om[2], 0.f, -om[0], -om[1], om[0], 0.f};
    float O2[9], V[9];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) { const float col[3] = {O[j], O[3 + j], O[6 + j]}; O2[3 * i + j] = dot3h(O + 3 * i, col); }
    if (theta < eps) quat_to_R(q, V);
    else {
        const float tsq = theta * theta;
        const float c1 = (1.f - std::cos(theta)) / tsq, c2 = (theta - std::sin(theta)) / (tsq * theta);
        for (int i = 0; i < 9; ++i) V[i] = ((i % 4 == 0 ? 1.f : 0.f) + c1 * O[i]) + c2 * O2[i];
    }
    matvec3h(V, a, tr);
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
    gsdf_map* h; int kind; cudaEvent_t stop = nullptr;
    ProfScope(gsdf_map* h_, int kind_) : h(h_), kind(kind_) {
        if (!h->profiling) return;
        auto& v = h->prof_events[kind];
        if (h->prof_used[kind] == v.size()) {
            cudaEvent_t a, b;
            cudaEventCreate(&a); cudaEventCreate(&b);
            v.push_back({a, b});
        }
        auto& pr = v[h->prof_used[kind]++];
        cudaEventRecord(pr.first, h->stream);
        stop = pr.second;
    }
    ~ProfScope() { if (stop) cudaEventRecord(stop, h->stream); }
};

gsdf_status ensure_normals(gsdf_map* h) {
    if (!h->d_depth) { h->err = "no depth frame set"; return GSDF_ERR_NO_FRAME; }
    if (h->normals_valid) return GSDF_OK;
    const int w = h->cfg.width, ht = h->cfg.height, r = h->cfg.normal_window / 2;
    const size_t n = (size_t)w * ht;
    const int tw = NT_W + 2 * r, th = NT_H + 2 * r;
    const size_t smem = ((sizeof(float) * 3 * tw * th + 7) & ~size_t(7)) + sizeof(double) * 3 * th * NT_W;
    dim3 grid((w + NT_W - 1) / NT_W, (ht + NT_H - 1) / NT_H), block(NT_W, NT_H);
    ProfScope prof(h, GSDF_PROF_NORMALS);
    normals_kernel<<<grid, block, smem, h->stream>>>(h->d_depth, h->d_planes + 2 * n, h->d_planes + 3 * n,
                                                     h->d_planes + 4 * n, h->d_planes + 5 * n, h->d_normals,
                                                     h->d_normals + n, h->d_normals + 2 * n, w, ht, r);
    gsdf_status s = check_launch(h, "normals_kernel");
    if (s == GSDF_OK) h->normals_valid = true;
    return s;
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
    return GSDF_OK;
}

void free_all(gsdf_map* h) {
    if (!h) return;
    cudaSetDevice(h->cfg.device);
    if (h->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(h->comm);
    cudaFree(h->view.keys); cudaFree(h->view.acc4); cudaFree(h->view.accW); cudaFree(h->view.acc2);
    cudaFree(h->d_planes); cudaFree(h->d_depth_own); cudaFree(h->d_normals); cudaFree(h->d_ctr);
    cudaFree(h->d_partials); cudaFree(h->d_pose); cudaFree(h->d_out29); cudaFree(h->d_count2);
    cudaFreeHost(h->h_out29); cudaFreeHost(h->h_ctr); cudaFreeHost(h->h_pose); cudaFreeHost(h->h_count2);
    for (auto& v : h->prof_events) for (auto& pr : v) { cudaEventDestroy(pr.first); cudaEventDestroy(pr.second); }
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
}

gsdf_status create_impl(gsdf_map* h) {
    const gsdf_config& c = h->cfg;
    CUDA_TRY(h, cudaSetDevice(c.device));
    cudaDeviceProp prop;
    CUDA_TRY(h, cudaGetDeviceProperties(&prop, c.device));
    h->sm_count = prop.multiProcessorCount;
    CUDA_TRY(h, cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
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
    CUDA_TRY(h, cudaMalloc(&h->d_depth_own, sizeof(float) * n));
    CUDA_TRY(h, cudaMalloc(&h->d_normals, sizeof(float) * n * 3));
    CUDA_TRY(h, cudaMalloc(&h->d_ctr, sizeof(Counters)));
    h->acc_blocks = h->sm_count * 4;
    CUDA_TRY(h, cudaMalloc(&h->d_partials, sizeof(float) * kAccStride * h->acc_blocks));
    CUDA_TRY(h, cudaMalloc(&h->d_pose, sizeof(Pose)));
    CUDA_TRY(h, cudaMalloc(&h->d_out29, sizeof(float) * kAccStride));
    CUDA_TRY(h, cudaMalloc(&h->d_count2, sizeof(unsigned long long) * 4));
    CUDA_TRY(h, cudaHostAlloc(&h->h_out29, sizeof(float) * kAccStride, cudaHostAllocDefault));
    CUDA_TRY(h, cudaHostAlloc(&h->h_ctr, sizeof(Counters), cudaHostAllocDefault));
    CUDA_TRY(h, cudaHostAlloc(&h->h_pose, sizeof(Pose), cudaHostAllocDefault));
    CUDA_TRY(h, cudaHostAlloc(&h->h_count2, sizeof(unsigned long long) * 4, cudaHostAllocDefault));
    build_normal_cache(c, h->h_planes);
    CUDA_TRY(h, cudaMemcpyAsync(h->d_planes, h->h_planes.data(), sizeof(float) * n * 11, cudaMemcpyHostToDevice, h->stream));
    gsdf_status s = clear_map(h);
    if (s != GSDF_OK) return s;
    CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    return GSDF_OK;
}

gsdf_status accumulate_device(gsdf_map* h, int sampling) {
    // launches the reduction; result lands in d_out29 (all 29 terms), pose taken from d_pose
    TrackIntr ti;
    ti.fx_inv = 1.f / h->cfg.K[0]; ti.fy_inv = 1.f / h->cfg.K[4]; ti.cx = h->cfg.K[2]; ti.cy = h->cfg.K[5];
    ti.sampling = sampling;
    ti.ws = (h->cfg.width + sampling - 1) / sampling; ti.hs = (h->cfg.height + sampling - 1) / sampling;
    const int total = ti.ws * ti.hs;
    int blocks = (total + kAccThreads - 1) / kAccThreads;
    if (blocks > h->acc_blocks) blocks = h->acc_blocks;
    gsdf_status s = by_variant(h, [&](auto V) {
        ProfScope prof(h, GSDF_PROF_ACCUMULATE);
        accumulate_kernel<decltype(V)::value><<<blocks, kAccThreads, 0, h->stream>>>(
            h->view, h->geom, ti, h->d_pose, h->d_depth, h->d_partials, h->d_ctr, h->d_out29);
        return check_launch(h, "accumulate_kernel");
    });
    if (s != GSDF_OK) return s;
    if (h->cfg.world_size > 1) {
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

}  // namespace

extern "C" {

const char* gsdf_version(void) { return "gsdf_b200 0.1 (sm_100a)"; }

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
        c.rank < 0 || c.rank >= c.world_size || !(c.K[0] != 0.f) || !(c.K[4] != 0.f)) {
        g_create_error = "invalid gsdf_config";
        return GSDF_ERR_INVALID;
    }
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0 || c.device < 0 || c.device >= ndev) {
        g_create_error = std::string("no usable CUDA device (there is no CPU fallback): ") + cudaGetErrorString(e);
        return GSDF_ERR_CUDA;
    }
    gsdf_map* h = new gsdf_map();
    h->cfg = c;
    h->geom.w = c.width; h->geom.h = c.height;
    h->geom.vs = c.voxel_size;
    h->geom.vs_inv = (float)(1. / (double)c.voxel_size);  // MapGradPixelSdf.h:102
    h->geom.T = c.truncation;
    h->geom.inv_T = (float)(1. / (double)c.truncation);   // Sdf.h:105
    h->geom.z_min = c.z_min; h->geom.z_max = c.z_max;
    h->geom.factor = (int)std::floor(c.truncation / c.voxel_size);  // MapGradPixelSdfOmp.cpp:78
    gsdf_status s = create_impl(h);
    if (s != GSDF_OK) { g_create_error = h->err; free_all(h); return s; }
    *out = h;
    return GSDF_OK;
}

gsdf_status gsdf_destroy(gsdf_handle h) {
    if (!h) return GSDF_ERR_INVALID;
    cudaSetDevice(h->cfg.device);
    cudaStreamSynchronize(h->stream);
    free_all(h);
    return GSDF_OK;
}

gsdf_status gsdf_reset(gsdf_handle h) {
    if (!h) return GSDF_ERR_INVALID;
    CUDA_TRY(h, cudaSetDevice(h->cfg.device));
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
    const size_t n = (size_t)h->cfg.width * h->cfg.height;
    CUDA_TRY(h, cudaMemcpyAsync(h->d_depth_own, depth_host, sizeof(float) * n, cudaMemcpyHostToDevice, h->stream));
    h->d_depth = h->d_depth_own;
    h->normals_valid = false;
    return GSDF_OK;
}

gsdf_status gsdf_set_frame_device(gsdf_handle h, const float* depth_dev) {
    if (!h || !depth_dev) return GSDF_ERR_INVALID;
    h->d_depth = depth_dev;
    h->normals_valid = false;
    return GSDF_OK;
}

gsdf_status gsdf_compute_normals(gsdf_handle h, float* nx, float* ny, float* nz) {
    if (!h || !nx || !ny || !nz) return GSDF_ERR_INVALID;
    CUDA_TRY(h, cudaSetDevice(h->cfg.device));
    gsdf_status s = ensure_normals(h);
    if (s != GSDF_OK) return s;
    const size_t n = (size_t)h->cfg.width * h->cfg.height;
    CUDA_TRY(h, cudaMemcpyAsync(nx, h->d_normals, sizeof(float) * n, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(h, cudaMemcpyAsync(ny, h->d_normals + n, sizeof(float) * n, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(h, cudaMemcpyAsync(nz, h->d_normals + 2 * n, sizeof(float) * n, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    return GSDF_OK;
}

gsdf_status gsdf_integrate(gsdf_handle h, const float* R9, const float* t3, gsdf_integrate_stats* stats) {
    if (!h || !R9 || !t3) return GSDF_ERR_INVALID;
    CUDA_TRY(h, cudaSetDevice(h->cfg.device));
    gsdf_status s = ensure_normals(h);  // MapGradPixelSdfOmp.cpp:59
    if (s != GSDF_OK) return s;
    Pose pose;
    std::memcpy(pose.R, R9, sizeof(float) * 9);
    std::memcpy(pose.t, t3, sizeof(float) * 3);
    const int n = h->cfg.width * h->cfg.height;
    const size_t np = (size_t)n;
    const int blocks = (n + 255) / 256;
    s = by_variant(h, [&](auto V) {
        ProfScope prof(h, GSDF_PROF_FUSE);
        fuse_kernel<decltype(V)::value><<<blocks, 256, 0, h->stream>>>(
            h->view, h->geom, pose, h->d_depth, h->d_normals, h->d_normals + np, h->d_normals + 2 * np, h->d_planes,
            h->d_planes + np, h->d_planes + 4 * np, h->d_ctr);
        return check_launch(h, "fuse_kernel");
    });
    if (s != GSDF_OK) return s;
    if (stats) {
        CUDA_TRY(h, cudaMemcpyAsync(h->h_ctr, h->d_ctr, sizeof(Counters), cudaMemcpyDeviceToHost, h->stream));
        CUDA_TRY(h, cudaMemsetAsync(h->d_ctr, 0, 2 * sizeof(unsigned long long), h->stream));  // points, updates
        CUDA_TRY(h, cudaStreamSynchronize(h->stream));
        stats->points = (int64_t)h->h_ctr->points;
        stats->updates = (int64_t)h->h_ctr->updates;
        stats->blocks = (int64_t)h->h_ctr->blocks;
        if (h->h_ctr->overflow) {
            h->err = "voxel-block hash full: samples dropped; raise gsdf_config.log2_capacity";
            return GSDF_ERR_CAPACITY;
        }
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
    gsdf_status s = set_pose_device(h, R9, t3);
    if (s != GSDF_OK) return s;
    s = accumulate_device(h, sampling);
    if (s != GSDF_OK) return s;
    CUDA_TRY(h, cudaMemcpyAsync(h->h_out29, h->d_out29, sizeof(float) * kAccTerms, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    std::memcpy(out29, h->h_out29, sizeof(float) * kAccTerms);
    return GSDF_OK;
}

gsdf_status gsdf_track(gsdf_handle h, float* pose7, int32_t max_iters, float conv_threshold, float damping,
                       int32_t* converged, int32_t* iters_used) {
    if (!h || !pose7 || !converged) return GSDF_ERR_INVALID;
    const float thr_sq = conv_threshold * conv_threshold;  // RigidOptimizer.h:72
    *converged = 0;
    for (int k = 0; k < max_iters; ++k) {
        float R[9], out[kAccTerms], H[36], xi[6];
        quat_to_R(pose7 + 3, R);  // pose_.rotationMatrix(), RigidPointOptimizerOmp.cpp:59
        gsdf_status s = gsdf_accumulate(h, R, pose7, 1, out);
        if (s != GSDF_OK) return s;
        h->track_iterations++;
        h->track_valid += (int64_t)out[28];
        int o = 6;
        for (int a = 0; a < 6; ++a)
            for (int b = a; b < 6; ++b) { H[6 * a + b] = out[o]; H[6 * b + a] = out[o]; ++o; }
        gsdf_llt_solve6(H, out, xi);
        bool has_nan = false;
        float nrm = 0.f;
        for (int i = 0; i < 6; ++i) { xi[i] = damping * xi[i]; has_nan |= std::isnan(xi[i]); nrm += xi[i] * xi[i]; }
        if (nrm < thr_sq) { *converged = 1; if (iters_used) *iters_used = k; return GSDF_OK; }  // :95-98
        if (!has_nan) { float np[7]; gsdf_pose_apply_increment(pose7, xi, np); std::memcpy(pose7, np, sizeof(np)); }  // :101-102
    }
    if (iters_used) *iters_used = max_iters;
    return GSDF_OK;
}

gsdf_status gsdf_query(gsdf_handle h, const float* pts, int64_t n, float* phi, float* grad, float* w) {
    if (!h || n < 0 || (n > 0 && (!pts || !phi || !grad || !w))) return GSDF_ERR_INVALID;
    if (n == 0) return GSDF_OK;
    CUDA_TRY(h, cudaSetDevice(h->cfg.device));
    float *d_pts = nullptr, *d_out = nullptr;
    CUDA_TRY(h, cudaMalloc(&d_pts, sizeof(float) * 3 * n));
    cudaError_t e = cudaMalloc(&d_out, sizeof(float) * 5 * n);
    if (e != cudaSuccess) { cudaFree(d_pts); h->err = "cudaMalloc(query)"; return GSDF_ERR_CUDA; }
    gsdf_status s = GSDF_OK;
    do {
        if (cudaMemcpyAsync(d_pts, pts, sizeof(float) * 3 * n, cudaMemcpyHostToDevice, h->stream) != cudaSuccess) { s = GSDF_ERR_CUDA; break; }
        const int blocks = (int)((n + 255) / 256);
        s = by_variant(h, [&](auto V) {
            query_kernel<decltype(V)::value><<<blocks, 256, 0, h->stream>>>(h->view, h->geom, d_pts, (long long)n, d_out,
                                                                            d_out + n, d_out + 4 * n);
            return check_launch(h, "query_kernel");
        });
        if (s != GSDF_OK) break;
        cudaMemcpyAsync(phi, d_out, sizeof(float) * n, cudaMemcpyDeviceToHost, h->stream);
        cudaMemcpyAsync(grad, d_out + n, sizeof(float) * 3 * n, cudaMemcpyDeviceToHost, h->stream);
        cudaMemcpyAsync(w, d_out + 4 * n, sizeof(float) * n, cudaMemcpyDeviceToHost, h->stream);
        if (cudaStreamSynchronize(h->stream) != cudaSuccess) { h->err = "query: stream sync failed"; s = GSDF_ERR_CUDA; }
    } while (0);
    cudaFree(d_pts); cudaFree(d_out);
    return s;
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
    float a[6], qa[4], ta[3], qo[4], rt[3];
    const float qb[4] = {pose7_in[3], pose7_in[4], pose7_in[5], pose7_in[6]};
    for (int i = 0; i < 6; ++i) a[i] = -xi6[i];
    se3_exp(a, qa, ta);
    quat_mul(qa, qb, qo);
    quat_rotate(qa, pose7_in, rt);
    for (int i = 0; i < 3; ++i) pose7_out[i] = ta[i] + rt[i];
    for (int i = 0; i < 4; ++i) pose7_out[3 + i] = qo[i];
}

void gsdf_llt_solve6(const float* H36, const float* g6, float* xi6) {
    // unblocked lower Cholesky in place, stopping at the first non-positive pivot as Eigen's LLT does, then
    // forward/back substitution
    float A[36], y[6];
    std::memcpy(A, H36, sizeof(A));
    for (int k = 0; k < 6; ++k) {
        float x = A[6 * k + k];
        for (int j = 0; j < k; ++j) x -= A[6 * k + j] * A[6 * k + j];
        if (x <= 0.f) break;
        x = std::sqrt(x);
        A[6 * k + k] = x;
        for (int i = k + 1; i < 6; ++i) {
            float s = A[6 * i + k];
            for (int j = 0; j < k; ++j) s -= A[6 * i + j] * A[6 * k + j];
            A[6 * i + k] = s / x;
        }
    }
    for (int i = 0; i < 6; ++i) {
        float s = g6[i];
        for (int j = 0; j < i; ++j) s -= A[6 * i + j] * y[j];
        y[i] = s / A[6 * i + i];
    }
    for (int i = 5; i >= 0; --i) {
        float s = y[i];
        for (int j = i + 1; j < 6; ++j) s -= A[6 * j + i] * xi6[j];
        xi6[i] = s / A[6 * i + i];
    }
}

gsdf_status gsdf_synchronize(gsdf_handle h) {
    if (!h) return GSDF_ERR_INVALID;
    CUDA_TRY(h, cudaSetDevice(h->cfg.device));
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
// resolves the real device address once with cudaGetSymbolAddress (that device current) and caches it; make
// that first call before starting any stream capture.  All calls on a device share the one array and it keeps
// the previous call's contents: the arrival ticket is cleared on the stream before every launch (the partial
// sums are fully overwritten by each launch and need no clearing), and two calls whose kernels overlap on the
// GPU (different handles/streams on the same device) race on it — serialise them.
static __device__ __align__(256) float g_noalloc_scratch[kStaticScratchFloats];

namespace {
constexpr int kMaxDevices = 64;
float* g_noalloc_ptr[kMaxDevices] = {};
std::mutex g_noalloc_mutex;
}  // namespace

extern "C" gsdf_status gsdf_accumulate_noalloc(gsdf_handle h, const float* R9, const float* t3, int32_t sampling,
                                               float* out29_dev) {
    if (!h || !R9 || !t3 || !out29_dev || sampling < 1) return GSDF_ERR_INVALID;
    if (!h->d_depth) { h->err = "no depth frame set"; return GSDF_ERR_NO_FRAME; }
    if (h->cfg.world_size > 1) { h->err = "gsdf_accumulate_noalloc: single-GPU maps only"; return GSDF_ERR_INVALID; }
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
