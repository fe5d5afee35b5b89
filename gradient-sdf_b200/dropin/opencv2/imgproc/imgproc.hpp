// cv::cvtColor as img_loader/ImageLoader.h:208 calls it (GRAY2BGR on a colour image that came back single-channel): a no-op
// here, the Scan3D path never reads colour.
#ifndef GSDF_DROPIN_OPENCV_IMGPROC_HPP_
#define GSDF_DROPIN_OPENCV_IMGPROC_HPP_
#include "../core/core.hpp"
namespace cv {
enum { COLOR_GRAY2BGR = 8 };
inline void cvtColor(const Mat& src, Mat& dst, int /*code*/) { dst = src; }
}  // namespace cv
#endif
