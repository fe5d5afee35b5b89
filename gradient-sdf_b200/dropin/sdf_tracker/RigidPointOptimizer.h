// Drop-in for cpp/include/sdf_tracker/RigidPointOptimizer.h (:48-73): optimize() runs the whole Gauss-Newton loop of
// RigidPointOptimizerOmp.cpp:44-106 on the GPU (gsdf_track) and returns its converged flag; the pose lives here as a Sophus SE3.
#ifndef GSDF_DROPIN_RIGID_POINT_OPTIMIZER_H_
#define GSDF_DROPIN_RIGID_POINT_OPTIMIZER_H_
#include "RigidOptimizer.h"
class RigidPointOptimizer : public RigidOptimizer {
public:
    explicit RigidPointOptimizer(Sdf* tSDF) : RigidOptimizer(tSDF) {}
    RigidPointOptimizer(int num_iterations, float conv_threshold, float damping, Sdf* tSDF)
        : RigidOptimizer(num_iterations, conv_threshold, damping, tSDF) {}

    bool optimize(const cv::Mat& depth, const Mat3f K) override {
        gsdf_shell::RigidPointOptimizer opt(num_iterations_, conv_threshold_, damping_, tSDF_->impl_);
        opt.set_pose(Sdf::to_shell(pose_));
        const gsdf_shell::cv::Mat d(depth.rows, depth.cols, CV_32FC1, depth.data);
        const bool conv = opt.optimize(d, Sdf::to_shell(K));
        const gsdf_shell::SE3 found = opt.pose();
        const float* p = found.data();  // tx ty tz qx qy qz qw, unit quaternion
        pose_ = SE3(Eigen::Quaternionf(p[6], p[3], p[4], p[5]), Vec3f(p[0], p[1], p[2]));
        return conv;
    }
    bool optimize_sampled(const cv::Mat& depth, const Mat3f K, int /*sampling*/) { return optimize(depth, K); }
};
#endif
