// Drop-in for cpp/include/normals/NormalEstimator.h (cv::NormalEstimator<T>, :49-212) when the reference's own sources are
// built against the B200 library: same constructor, but the cache() and compute() work happens on the GPU inside the volume
// (gsdf_create builds the cached planes; the frame pipeline computes the normals), so this object only carries the
// constructor arguments to Sdf::update().
#ifndef GSDF_DROPIN_NORMAL_ESTIMATOR_H_
#define GSDF_DROPIN_NORMAL_ESTIMATOR_H_

#include <Eigen/Dense>
#include <opencv2/core/core.hpp>

namespace cv {

template <typename T>
class NormalEstimator {
public:
    NormalEstimator(int width, int height, Eigen::Matrix<T, 3, 3> K, Size window_size)  // NormalEstimator.h:158-165
        : width_(width), height_(height), K_(K), window_(window_size.width) {}
    int width() const { return width_; }
    int height() const { return height_; }
    int window() const { return window_; }

private:
    int width_, height_;
    Eigen::Matrix<T, 3, 3> K_;
    int window_;
};

}  // namespace cv

#endif
