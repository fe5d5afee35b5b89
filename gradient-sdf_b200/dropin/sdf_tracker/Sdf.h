// Drop-in for cpp/include/sdf_tracker/Sdf.h (:50-149): the reference's abstract volume with its Eigen / Sophus signatures,
// every body a call into the C ABI (include/gsdf_b200.h).  With this directory first on the include path the reference's own
// cpp/depth_scanning/src/main_scan_3d.cpp compiles UNMODIFIED against libgsdf_b200.so (Makefile target host/scan3d_ref_main).
// The writers are shared with the stand-alone shells (host/sdf_tracker.h), which are compiled in under a namespace.
#ifndef GSDF_DROPIN_SDF_H_
#define GSDF_DROPIN_SDF_H_

#include <iostream>
#include <string>

#include <opencv2/core/core.hpp>

#include "mat.h"                       // the reference's own typedefs (Mat3f, Vec3f, SE3 = Sophus::SE3<float>)
#include "normals/NormalEstimator.h"

// The Eigen-free shells (volume handling, tracker call, output writers) are compiled in under a namespace; every header they
// include is included here first, so that inside the namespace only their own declarations remain.
#include <array>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <fstream>
#include <limits>
#include <map>
#include <vector>

#include "gsdf_b200.h"
namespace gsdf_shell {
#include "../../host/mat.h"
#include "../../host/sdf_tracker.h"
}  // namespace gsdf_shell

class RigidPointOptimizer;

class Sdf {
protected:
    friend class RigidPointOptimizer;
    gsdf_shell::Sdf* impl_;
    float voxel_size_;

    static gsdf_shell::SE3 to_shell(const SE3& pose) {
        const Eigen::Quaternionf& q = pose.unit_quaternion();
        const Vec3f t = pose.translation();
        const float p[7] = {t[0], t[1], t[2], q.x(), q.y(), q.z(), q.w()};
        return gsdf_shell::SE3::fromStored(p);  // Sophus keeps the quaternion normalised
    }
    static gsdf_shell::Mat3f to_shell(const Mat3f& K) {
        gsdf_shell::Mat3f m;
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) m(i, j) = K(i, j);
        return m;
    }

public:
    explicit Sdf(gsdf_shell::Sdf* impl, float voxel_size) : impl_(impl), voxel_size_(voxel_size) {}
    virtual ~Sdf() { delete impl_; }

    virtual float tsdf(Vec3f point, Vec3f* grad_ptr = nullptr) const {  // Sdf.h:113
        gsdf_shell::Vec3f g;
        const float phi = impl_->tsdf(gsdf_shell::Vec3f(point[0], point[1], point[2]), &g);
        if (grad_ptr) *grad_ptr = Vec3f(g[0], g[1], g[2]);
        return phi;
    }
    virtual float weights(Vec3f point) const { return impl_->weights(gsdf_shell::Vec3f(point[0], point[1], point[2])); }  // :115

    virtual void update(const cv::Mat& color, const cv::Mat& depth, const Mat3f K, const SE3& pose,
                        cv::NormalEstimator<float>* NEst = nullptr) {  // Sdf.h:117
        (void)color;
        if (!NEst) { std::cerr << "No normal estimation possible - cannot update SDF volume!" << std::endl; return; }
        gsdf_shell::cv::NormalEstimator<float> ne(depth.cols, depth.rows, to_shell(K), gsdf_shell::cv::Size(NEst->window(), NEst->window()));
        const gsdf_shell::cv::Mat d(depth.rows, depth.cols, CV_32FC1, depth.data);
        impl_->update(gsdf_shell::cv::Mat(), d, to_shell(K), to_shell(pose), &ne);
    }
    virtual void setup(const cv::Mat& color, const cv::Mat& depth, const Mat3f K, cv::NormalEstimator<float>* NEst = nullptr) {
        update(color, depth, K, SE3(), NEst);  // Sdf.h:119-121
    }
    void set_zmin(float z) { impl_->set_zmin(z); }
    void set_zmax(float z) { impl_->set_zmax(z); }

    virtual bool extract_mesh(std::string filename = "mesh.ply") { return impl_->extract_mesh(filename); }
    virtual bool extract_pc(std::string filename = "cloud.ply") { return impl_->extract_pc(filename); }
    virtual bool save_sdf(std::string filename = "sdf.txt") { return impl_->save_sdf(filename); }
};

#endif
