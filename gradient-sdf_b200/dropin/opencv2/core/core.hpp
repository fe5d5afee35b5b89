// Minimal cv:: surface for building the reference's OWN Scan3D driver and dataset loaders
// (cpp/depth_scanning/src/main_scan_3d.cpp, cpp/include/img_loader/*.h) against the B200 library without OpenCV:
// cv::Mat as a reference-counted single-plane image (8/16-bit integer or 32-bit float), convertTo, Size.  Only what those
// translation units call; the per-frame image processing of the reference (box filters, median) is not here — it runs on the GPU
// behind the class shells in ../sdf_tracker and ../normals.
#ifndef GSDF_DROPIN_OPENCV_CORE_HPP_
#define GSDF_DROPIN_OPENCV_CORE_HPP_

#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

#define CV_8U 0
#define CV_16U 2
#define CV_32F 5
#define CV_8UC1 CV_8U
#define CV_16UC1 CV_16U
#define CV_32FC1 CV_32F
#define CV_8UC3 16

namespace cv {

struct Size {
    int width, height;
    Size() : width(0), height(0) {}
    Size(int w, int h) : width(w), height(h) {}
};

class Mat {
public:
    int rows = 0, cols = 0;
    unsigned char* data = nullptr;

    Mat() {}
    Mat(int r, int c, int type) { create(r, c, type); }
    Mat(int r, int c, int type, void* ext) : rows(r), cols(c), data((unsigned char*)ext), type_(type) {}  // caller-owned memory
    void create(int r, int c, int type) {
        rows = r; cols = c; type_ = type;
        buf_ = std::shared_ptr<std::vector<unsigned char>>(new std::vector<unsigned char>((size_t)r * c * elem_size(type)));
        data = buf_->data();
    }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    int type() const { return type_; }
    int channels() const { return type_ == CV_8UC3 ? 3 : 1; }
    Size size() const { return Size(cols, rows); }
    size_t total() const { return (size_t)rows * cols; }
    // convertTo(dst, CV_32FC1, alpha): float(v) * float(alpha) per element (img_loader/ImageLoader.h:172,186,210).  A colour
    // image converts to a float image of the same size whose content is unspecified: Sdf::update() never reads colour
    // (MapGradPixelSdfOmp.cpp:44-130), and no colour decoder is linked.
    void convertTo(Mat& dst, int rtype, double alpha = 1.0) const {
        if (empty()) { dst = Mat(); return; }
        Mat out(rows, cols, rtype == CV_32F ? CV_32F : rtype);
        float* o = (float*)out.data;
        const float a = (float)alpha;
        const size_t n = total();
        if (type_ == CV_16U) for (size_t i = 0; i < n; ++i) o[i] = (float)((const uint16_t*)data)[i] * a;
        else if (type_ == CV_8U) for (size_t i = 0; i < n; ++i) o[i] = (float)data[i] * a;
        else if (type_ == CV_32F) for (size_t i = 0; i < n; ++i) o[i] = ((const float*)data)[i] * a;
        else std::memset(o, 0, n * sizeof(float));
        dst = out;
    }

private:
    static size_t elem_size(int type) { return type == CV_32F ? 4 : type == CV_16U ? 2 : type == CV_8UC3 ? 3 : 1; }
    int type_ = CV_8U;
    std::shared_ptr<std::vector<unsigned char>> buf_;
};

}  // namespace cv

#endif
