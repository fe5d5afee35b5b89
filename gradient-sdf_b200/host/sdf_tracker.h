// Forwarding shells with the reference's class names and signatures for the Scan3D hot path; every method body is a
// call into the C ABI (include/gsdf_b200.h).  Mirrors, in /root/reference/cpp/include:
//   normals/NormalEstimator.h:49-212        cv::NormalEstimator<float>
//   sdf_tracker/Sdf.h:50-149                Sdf
//   sdf_tracker/MapGradPixelSdf.h:51-152    MapGradPixelSdf   (+ writers MapGradPixelSdfOmp.cpp:186-303)
//   sdf_tracker/MapPixelSdf.h:52-155        MapPixelSdf       (+ writers MapPixelSdfOmp.cpp:250-351)
//   sdf_tracker/RigidOptimizer.h:49-112     RigidOptimizer
//   sdf_tracker/RigidPointOptimizer.h:48-73 RigidPointOptimizer
// Same argument meaning and error behaviour: update() reports on std::cerr and returns, optimize() returns the
// converged flag, the writers return bool.  The volume is created on the GPU at the first update(), because only
// then are the image size and the intrinsics known (the reference, too, builds its map lazily).
#ifndef GSDF_HOST_SDF_TRACKER_H_
#define GSDF_HOST_SDF_TRACKER_H_

#include <cmath>
#include <cstdint>
#include <fstream>
#include <iostream>
#include <array>
#include <limits>
#include <map>
#include <string>
#include <vector>

#include "gsdf_b200.h"
#include "mat.h"

namespace cv {

template <typename T>
class NormalEstimator {
public:
    // NormalEstimator(width, height, K, window_size), NormalEstimator.h:158-165
    NormalEstimator(int width, int height, const Mat3f& K, Size window_size)
        : width_(width), height_(height), K_(K), window_(window_size.width) {}
    ~NormalEstimator() { if (own_) gsdf_destroy(own_); }

    // compute(depth, nx, ny, nz), NormalEstimator.h:179-204; the three outputs are width*height floats each
    bool compute(const Mat& depth, std::vector<float>& nx, std::vector<float>& ny, std::vector<float>& nz) {
        if (!own_ && !make_private_handle()) return false;
        const size_t n = (size_t)width_ * height_;
        nx.resize(n); ny.resize(n); nz.resize(n);
        if (gsdf_upload_frame(own_, (const float*)depth.data) != GSDF_OK) return false;
        return gsdf_compute_normals(own_, nx.data(), ny.data(), nz.data()) == GSDF_OK;
    }

    int width() const { return width_; }
    int height() const { return height_; }
    int window() const { return window_; }
    const Mat3f& K() const { return K_; }

private:
    bool make_private_handle() {
        gsdf_config c{};
        c.variant = GSDF_GRAD_SDF; c.voxel_size = 0.01f; c.truncation = 0.05f; c.z_min = 0.5f; c.z_max = 3.5f;
        c.width = width_; c.height = height_;
        for (int i = 0; i < 9; ++i) c.K[i] = K_.m[i];
        c.normal_window = window_; c.log2_capacity = 4; c.world_size = 1;
        if (gsdf_create(&c, &own_) != GSDF_OK) { std::cerr << gsdf_last_error(nullptr) << std::endl; return false; }
        return true;
    }
    int width_, height_;
    Mat3f K_;
    int window_;
    gsdf_handle own_ = nullptr;  // only for stand-alone compute(); fusion uses the volume's own cache
};

}  // namespace cv

struct SdfVoxel {  // sdf_voxel/SdfVoxel.h:45-57
    float dist = 0.f;
    Vec3f grad;
    float weight = 0.f;
};

class Sdf {
protected:
    friend class RigidPointOptimizer;
    float T_, inv_T_;
    size_t counter_ = 0;
    float z_min_ = 0.5f, z_max_ = 3.5f;  // Sdf.h:67-68
    const float voxel_size_;
    const int variant_;
    int log2_capacity_ = 17;
    gsdf_handle gpu_ = nullptr;

    void increase_counter() { ++counter_; }

    bool ensure_gpu(int width, int height, const Mat3f& K, int window) {
        if (gpu_) return true;
        gsdf_config c{};
        c.variant = variant_; c.voxel_size = voxel_size_; c.truncation = T_; c.z_min = z_min_; c.z_max = z_max_;
        c.width = width; c.height = height;
        for (int i = 0; i < 9; ++i) c.K[i] = K.m[i];
        c.normal_window = window; c.log2_capacity = log2_capacity_; c.device = 0; c.world_size = 1; c.rank = 0;
        if (gsdf_create(&c, &gpu_) != GSDF_OK) {
            std::cerr << "gsdf_create: " << gsdf_last_error(nullptr) << std::endl;
            return false;
        }
        return true;
    }

    struct HostMap {
        std::vector<int32_t> keys;
        std::vector<float> dist, weight, grad;
        int64_t n = 0;
    };
    bool download(HostMap& m) const {
        if (!gpu_) return false;
        int64_t nv = 0, nb = 0;
        if (gsdf_voxel_count(gpu_, &nv, &nb) != GSDF_OK) return false;
        m.keys.resize(3 * nv); m.dist.resize(nv); m.weight.resize(nv); m.grad.resize(3 * nv);
        m.n = 0;
        if (nv == 0) return true;
        return gsdf_download(gpu_, nv, m.keys.data(), m.dist.data(), m.weight.data(), m.grad.data(), &m.n) == GSDF_OK;
    }

public:
    Sdf(int variant, float voxel_size, float T)
        : T_(T), inv_T_((float)(1. / (double)T)), voxel_size_(voxel_size), variant_(variant) {}
    virtual ~Sdf() { if (gpu_) gsdf_destroy(gpu_); }

    gsdf_handle gpu() const { return gpu_; }
    void set_capacity_log2(int log2_blocks) { log2_capacity_ = log2_blocks; }  // before the first update()

    // float tsdf(Vec3f point, Vec3f* grad_ptr) / float weights(Vec3f point), Sdf.h:113-115
    virtual float tsdf(Vec3f point, Vec3f* grad_ptr = nullptr) const {
        float phi = 0.f, g[3] = {0, 0, 0}, w = 0.f;
        if (gpu_) gsdf_query(gpu_, point.data(), 1, &phi, g, &w);
        if (grad_ptr) *grad_ptr = Vec3f(g[0], g[1], g[2]);
        return phi;
    }
    virtual float weights(Vec3f point) const {
        float phi = 0.f, g[3], w = 0.f;
        if (gpu_) gsdf_query(gpu_, point.data(), 1, &phi, g, &w);
        return w;
    }

    // update(color, depth, K, pose, NEst), Sdf.h:117; color is accepted and ignored, as in the reference
    virtual void update(const cv::Mat& color, const cv::Mat& depth, const Mat3f K, const SE3& pose,
                        cv::NormalEstimator<float>* NEst = nullptr) {
        (void)color;
        if (!NEst) {  // MapGradPixelSdfOmp.cpp:54-57
            std::cerr << "No normal estimation possible - cannot update SDF volume!" << std::endl;
            return;
        }
        if (!ensure_gpu(depth.cols, depth.rows, K, NEst->window())) return;
        gsdf_integrate_stats st;
        if (gsdf_upload_frame(gpu_, (const float*)depth.data) != GSDF_OK ||
            gsdf_integrate_pose(gpu_, pose.data(), &st) != GSDF_OK) {
            std::cerr << "gsdf_integrate: " << gsdf_last_error(gpu_) << std::endl;
            return;
        }
        increase_counter();
        std::cout << st.points << " points integrated into tSDF volume!" << std::endl;  // MapGradPixelSdfOmp.cpp:129
    }
    // fuse the frame the tracker has just uploaded (saves the second host->device copy of the same depth image)
    void update_current(const SE3& pose) {
        gsdf_integrate_stats st;
        if (!gpu_ || gsdf_integrate_pose(gpu_, pose.data(), &st) != GSDF_OK) {
            std::cerr << "gsdf_integrate: " << (gpu_ ? gsdf_last_error(gpu_) : "no volume") << std::endl;
            return;
        }
        increase_counter();
        std::cout << st.points << " points integrated into tSDF volume!" << std::endl;
    }
    virtual void setup(const cv::Mat& color, const cv::Mat& depth, const Mat3f K, cv::NormalEstimator<float>* NEst = nullptr) {
        update(color, depth, K, SE3(), NEst);  // Sdf.h:119-121
    }
    void set_zmin(float z_min) { z_min_ = z_min; }
    void set_zmax(float z_max) { z_max_ = z_max; }

    SdfVoxel getSdf(Vec3i idx) const {  // MapGradPixelSdf.h:127-129: the voxel stored at idx (zeros if there is none)
        SdfVoxel v;
        float g[3] = {0.f, 0.f, 0.f};
        const int32_t key[3] = {idx[0], idx[1], idx[2]};
        if (gpu_ && gsdf_get_voxels(gpu_, 1, key, &v.dist, &v.weight, g) == GSDF_OK) v.grad = Vec3f(g[0], g[1], g[2]);
        return v;
    }

    // extract_mesh, MapGradPixelSdfOmp.cpp:132-183 / MapPixelSdfOmp.cpp:197-248: marching cubes on the device
    // (gsdf_extract_mesh), ascii PLY exactly as LayeredMarchingCubesNoColor::savePly (mesh/LayeredMarchingCubesNoColor.cpp:
    // 719-757): three unshared vertices per face, false if there is nothing to write
    virtual bool extract_mesh(std::string filename = "mesh.ply") {
        if (!gpu_) return false;
        int64_t n = 0;
        if (gsdf_extract_mesh(gpu_, 0, nullptr, &n) != GSDF_OK || n == 0) return false;
        std::vector<float> v((size_t)9 * n);
        if (gsdf_extract_mesh(gpu_, n, v.data(), &n) != GSDF_OK) { std::cerr << gsdf_last_error(gpu_) << std::endl; return false; }
        std::ofstream ply(filename.c_str());
        if (!ply.is_open()) return false;
        ply << "ply" << std::endl << "format ascii 1.0" << std::endl << "element vertex " << 3 * n << std::endl
            << "property float x" << std::endl << "property float y" << std::endl << "property float z" << std::endl
            << "element face " << (int)n << std::endl << "property list uchar int vertex_indices" << std::endl
            << "end_header" << std::endl;
        for (int64_t i = 0; i < 3 * n; ++i) ply << v[3 * i] << " " << v[3 * i + 1] << " " << v[3 * i + 2] << "\n";
        for (int64_t f = 0; f < n; ++f) ply << "3 " << (int)(3 * f) << " " << (int)(3 * f + 1) << " " << (int)(3 * f + 2) << "\n";
        return true;
    }

    // extract_pc, MapGradPixelSdfOmp.cpp:186-233: voxels with weight >= 5 whose surface point lies inside the voxel
    virtual bool extract_pc(std::string filename = "cloud.ply") {
        HostMap m;
        if (!download(m)) return false;
        std::vector<float> pn;
        for (int64_t i = 0; i < m.n; ++i) {
            if (m.weight[i] < 5) continue;
            float g[3] = {m.grad[3 * i], m.grad[3 * i + 1], m.grad[3 * i + 2]};
            const float sq = g[0] * g[0] + (g[1] * g[1] + g[2] * g[2]);
            if (sq > 0.f) { const float l = std::sqrt(sq); for (float& c : g) c /= l; }
            float d[3], half = .5f * voxel_size_;
            for (int c = 0; c < 3; ++c) { g[c] = 1.2f * g[c]; d[c] = m.dist[i] * g[c]; }
            if (std::fabs(d[0]) < half && std::fabs(d[1]) < half && std::fabs(d[2]) < half) {
                for (int c = 0; c < 3; ++c) pn.push_back(voxel_size_ * (float)m.keys[3 * i + c] - d[c]);
                for (int c = 0; c < 3; ++c) pn.push_back(-g[c]);
            }
        }
        std::ofstream ply(filename.c_str());
        if (!ply.is_open()) return false;
        ply << "ply\nformat ascii 1.0\nelement vertex " << pn.size() / 6
            << "\nproperty float x\nproperty float y\nproperty float z\nproperty float nx\nproperty float ny\nproperty float nz\nend_header\n";
        for (size_t i = 0; i < pn.size(); i += 6)
            ply << pn[i] << " " << pn[i + 1] << " " << pn[i + 2] << " " << pn[i + 3] << " " << pn[i + 4] << " " << pn[i + 5] << "\n";
        return true;
    }

    // save_sdf, MapGradPixelSdfOmp.cpp:235-303: <prefix>_grid_info.txt, _sdf_d.txt, _sdf_weight.txt, _sdf_n{0,1,2}.txt
    virtual bool save_sdf(std::string filename = "sdf.txt") { return save_sdf_files(filename, true); }

protected:
    bool save_sdf_files(const std::string& filename, bool with_normals) {
        HostMap m;
        if (!download(m)) return false;
        int lo[3] = {std::numeric_limits<int>::max(), std::numeric_limits<int>::max(), std::numeric_limits<int>::max()};
        int hi[3] = {std::numeric_limits<int>::min(), std::numeric_limits<int>::min(), std::numeric_limits<int>::min()};
        for (int64_t i = 0; i < m.n; ++i)
            for (int c = 0; c < 3; ++c) { lo[c] = std::min(lo[c], m.keys[3 * i + c]); hi[c] = std::max(hi[c], m.keys[3 * i + c]); }
        const int dim[3] = {hi[0] - lo[0] + 1, hi[1] - lo[1] + 1, hi[2] - lo[2] + 1};
        std::ofstream grid((filename + "_grid_info.txt").c_str());
        if (!grid.is_open()) { std::cerr << "couldn't save grid_info file!" << std::endl; return false; }
        grid << "voxel size: " << voxel_size_ << std::endl
             << "voxel dim: " << dim[0] << " " << dim[1] << " " << dim[2] << std::endl
             << "voxel min: " << lo[0] << " " << lo[1] << " " << lo[2] << std::endl
             << "voxel max: " << hi[0] << " " << hi[1] << " " << hi[2] << std::endl;
        std::ofstream fd((filename + "_sdf_d.txt").c_str()), fw((filename + "_sdf_weight.txt").c_str());
        std::ofstream n0, n1, n2;
        if (with_normals) {
            n0.open((filename + "_sdf_n0.txt").c_str()); n1.open((filename + "_sdf_n1.txt").c_str()); n2.open((filename + "_sdf_n2.txt").c_str());
        }
        if (!fd.is_open() || !fw.is_open() || (with_normals && (!n0.is_open() || !n1.is_open() || !n2.is_open()))) {
            std::cerr << "couldn't save sdf or sdf weight file!" << std::endl;
            return false;
        }
        for (int64_t i = 0; i < m.n; ++i) {
            const int lin = dim[0] * dim[1] * (m.keys[3 * i + 2] - lo[2]) + dim[0] * (m.keys[3 * i + 1] - lo[1]) + m.keys[3 * i] - lo[0];
            fd << lin << " " << m.dist[i] << "\n";
            fw << lin << " " << m.weight[i] << "\n";
            if (with_normals) {
                n0 << lin << " " << m.grad[3 * i] << "\n";
                n1 << lin << " " << m.grad[3 * i + 1] << "\n";
                n2 << lin << " " << m.grad[3 * i + 2] << "\n";
            }
        }
        return true;
    }
};

class MapGradPixelSdf : public Sdf {
public:
    MapGradPixelSdf() : Sdf(GSDF_GRAD_SDF, 0.02f, 0.05f) {}
    explicit MapGradPixelSdf(float voxel_size) : Sdf(GSDF_GRAD_SDF, voxel_size, 0.05f) {}
    MapGradPixelSdf(float voxel_size, float T) : Sdf(GSDF_GRAD_SDF, voxel_size, T) {}

    // vis_ / get_vis(), MapGradPixelSdf.h:70,140-142: per voxel, which update() calls touched it.  The reference always
    // records it (only PhotoBA reads it); here it costs device memory, so it is recorded from record_vis() on: call it after
    // the first update() (the volume exists from then on), max_entries = (voxel, frame) pairs to make room for.
    bool record_vis(int64_t max_entries) { return gpu_ && gsdf_vis_enable(gpu_, max_entries) == GSDF_OK; }
    std::map<std::array<int, 3>, std::vector<bool>> get_vis() const {
        std::map<std::array<int, 3>, std::vector<bool>> out;
        int64_t n = 0, nf = 0;
        if (!gpu_ || gsdf_vis_download(gpu_, 0, nullptr, nullptr, &n, &nf) != GSDF_OK || n == 0) return out;
        std::vector<int32_t> keys(3 * n), frames(n);
        if (gsdf_vis_download(gpu_, n, keys.data(), frames.data(), &n, &nf) != GSDF_OK) return out;
        for (int64_t i = 0; i < n; ++i) {
            std::vector<bool>& v = out[{keys[3 * i], keys[3 * i + 1], keys[3 * i + 2]}];
            if ((int64_t)v.size() <= frames[i]) v.resize(frames[i] + 1, false);  // vis.resize(counter_); push_back(true)
            v[frames[i]] = true;
        }
        return out;
    }
};

class MapPixelSdf : public Sdf {
public:
    MapPixelSdf() : Sdf(GSDF_BASE_SDF, 0.02f, 0.05f) {}
    explicit MapPixelSdf(float voxel_size) : Sdf(GSDF_BASE_SDF, voxel_size, 0.05f) {}
    MapPixelSdf(float voxel_size, float T) : Sdf(GSDF_BASE_SDF, voxel_size, T) {}

    // extract_pc, MapPixelSdfOmp.cpp:250-288: voxel centres (xyz only, 7-line header) of the voxels with weight >= 5 and
    // |dist| < sqrt(3) * voxel_size (the comparison is in double there: std::sqrt(3) * voxel_size_)
    bool extract_pc(std::string filename = "cloud.ply") override {
        HostMap m;
        if (!download(m)) return false;
        std::vector<int64_t> keep;
        for (int64_t i = 0; i < m.n; ++i)
            if (!(m.weight[i] < 5) && (double)std::fabs(m.dist[i]) < std::sqrt(3.0) * (double)voxel_size_) keep.push_back(i);
        std::ofstream ply(filename.c_str());
        if (!ply.is_open()) return false;
        ply << "ply\nformat ascii 1.0\nelement vertex " << keep.size()
            << "\nproperty float x\nproperty float y\nproperty float z\nend_header\n";
        for (int64_t i : keep)  // vox2float, MapPixelSdf.h:74-76
            ply << voxel_size_ * (float)m.keys[3 * i] << " " << voxel_size_ * (float)m.keys[3 * i + 1] << " "
                << voxel_size_ * (float)m.keys[3 * i + 2] << "\n";
        return true;
    }

    // save_sdf, MapPixelSdfOmp.cpp:290-351: <prefix>_grid_info.txt, _sdf_d.txt, _sdf_weight.txt — and no normal files
    bool save_sdf(std::string filename = "sdf.txt") override { return save_sdf_files(filename, false); }
};

class RigidOptimizer {
protected:
    int num_iterations_;
    float conv_threshold_, conv_threshold_sq_, damping_;
    Sdf* tSDF_;
    SE3 pose_;

public:
    explicit RigidOptimizer(Sdf* tSDF)  // RigidOptimizer.h:70-76
        : num_iterations_(25), conv_threshold_(1e-3f), conv_threshold_sq_(1e-6f), damping_(1.f), tSDF_(tSDF) {}
    RigidOptimizer(int num_iterations, float conv_threshold, float damping, Sdf* tSDF)
        : num_iterations_(num_iterations), conv_threshold_(conv_threshold),
          conv_threshold_sq_(conv_threshold * conv_threshold), damping_(damping), tSDF_(tSDF) {}
    virtual ~RigidOptimizer() {}
    void set_num_iterations(int n) { num_iterations_ = n; }
    void set_conv_threshold(float t) { conv_threshold_ = t; conv_threshold_sq_ = t * t; }
    void set_damping(float d) { damping_ = d; }
    void set_pose(SE3 pose) { pose_ = pose; }
    SE3 pose() { return pose_; }
    virtual bool optimize(const cv::Mat& depth, const Mat3f K) = 0;
};

class RigidPointOptimizer : public RigidOptimizer {
public:
    explicit RigidPointOptimizer(Sdf* tSDF) : RigidOptimizer(tSDF) {}
    RigidPointOptimizer(int num_iterations, float conv_threshold, float damping, Sdf* tSDF)
        : RigidOptimizer(num_iterations, conv_threshold, damping, tSDF) {}
    int last_iterations() const { return last_iterations_; }

    // optimize(depth, K), RigidPointOptimizer.h:69-72 -> RigidPointOptimizerOmp.cpp:44-106
    bool optimize(const cv::Mat& depth, const Mat3f /*K*/) override {
        gsdf_handle gpu = tSDF_->gpu_;
        if (!gpu) { std::cerr << "RigidPointOptimizer: the volume holds no data yet" << std::endl; return false; }
        if (gsdf_upload_frame(gpu, (const float*)depth.data) != GSDF_OK) { std::cerr << gsdf_last_error(gpu) << std::endl; return false; }
        int32_t conv = 0, iters = 0;
        if (gsdf_track(gpu, pose_.data(), num_iterations_, conv_threshold_, damping_, &conv, &iters) != GSDF_OK) {
            std::cerr << "gsdf_track: " << gsdf_last_error(gpu) << std::endl;
            return false;
        }
        last_iterations_ = iters;
        if (conv) std::cout << "... Convergence after " << iters << " iterations!" << std::endl;  // :96
        return conv != 0;
    }

private:
    int last_iterations_ = 0;
};

#endif  // GSDF_HOST_SDF_TRACKER_H_
