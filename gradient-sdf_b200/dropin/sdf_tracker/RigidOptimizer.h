// Drop-in for cpp/include/sdf_tracker/RigidOptimizer.h (:49-112): tracker state with the reference's setters.
#ifndef GSDF_DROPIN_RIGID_OPTIMIZER_H_
#define GSDF_DROPIN_RIGID_OPTIMIZER_H_
#include "Sdf.h"
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
        : num_iterations_(num_iterations), conv_threshold_(conv_threshold), conv_threshold_sq_(conv_threshold * conv_threshold),
          damping_(damping), tSDF_(tSDF) {}
    virtual ~RigidOptimizer() {}
    void set_num_iterations(int n) { num_iterations_ = n; }
    void set_conv_threshold(float t) { conv_threshold_ = t; conv_threshold_sq_ = t * t; }
    void set_damping(float d) { damping_ = d; }
    void set_pose(SE3 pose) { pose_ = pose; }
    SE3 pose() { return pose_; }
    virtual bool optimize(const cv::Mat& depth, const Mat3f K) = 0;
};
#endif
