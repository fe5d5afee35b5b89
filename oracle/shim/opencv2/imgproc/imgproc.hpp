// TEST INFRASTRUCTURE — not part of the product.  See opencv2/core/core.hpp in this shim.
// boxFilter / medianBlur stand-ins for the reference call sites NormalEstimator.h:105-110,191-193
// and MapGradPixelSdfOmp.cpp:52 / MapPixelSdfOmp.cpp:122.
#ifndef GSDF_ORACLE_SHIM_OPENCV_IMGPROC_HPP_
#define GSDF_ORACLE_SHIM_OPENCV_IMGPROC_HPP_

#include "../core/core.hpp"

namespace cv {

// Unnormalised (or normalised) box filter, BORDER_REFLECT_101, anchor at the window centre.
// Accumulation is in double for every input depth (OpenCV picks sumType=CV_64F for 32F/64F input):
//   pass 1: rowsum(y,x) = sum_{dx=-r..r} src(y, reflect(x+dx))   added left→right
//   pass 2: dst(y,x)    = sum_{dy=-r..r} rowsum(reflect(y+dy), x) added top→bottom
// and a single rounding to the destination depth.  The CUDA normal kernel follows the same order.
inline void boxFilter(const Mat& src, Mat& dst, int ddepth, Size ksize, Point anchor = Point(-1, -1),
                      bool normalize = true) {
    (void)anchor;
    const int rx = ksize.width / 2, ry = ksize.height / 2;
    const int w = src.cols, h = src.rows;
    const int dtype = ddepth < 0 ? src.type() : ddepth;
    std::vector<double> rowsum((size_t)w * h);
    std::vector<int> xi((size_t)w * (2 * rx + 1)), yi((size_t)h * (2 * ry + 1));
    for (int x = 0; x < w; ++x)
        for (int d = -rx; d <= rx; ++d) xi[(size_t)x * (2 * rx + 1) + d + rx] = shim_detail::reflect101(x + d, w);
    for (int y = 0; y < h; ++y)
        for (int d = -ry; d <= ry; ++d) yi[(size_t)y * (2 * ry + 1) + d + ry] = shim_detail::reflect101(y + d, h);
    for (int y = 0; y < h; ++y) {
        for (int x = 0; x < w; ++x) {
            double s = 0.0;
            const int* xx = &xi[(size_t)x * (2 * rx + 1)];
            if (src.type() == CV_32F) {
                const float* p = (const float*)src.data + (size_t)y * w;
                for (int d = 0; d <= 2 * rx; ++d) s += (double)p[xx[d]];
            } else if (src.type() == CV_64F) {
                const double* p = (const double*)src.data + (size_t)y * w;
                for (int d = 0; d <= 2 * rx; ++d) s += p[xx[d]];
            } else {
                throw std::runtime_error("shim: boxFilter depth");
            }
            rowsum[(size_t)y * w + x] = s;
        }
    }
    Mat r(h, w, dtype);
    const double scale = normalize ? 1.0 / ((2 * rx + 1) * (2 * ry + 1)) : 1.0;
    for (int y = 0; y < h; ++y) {
        const int* yy = &yi[(size_t)y * (2 * ry + 1)];
        for (int x = 0; x < w; ++x) {
            double s = 0.0;
            for (int d = 0; d <= 2 * ry; ++d) s += rowsum[(size_t)yy[d] * w + x];
            if (normalize) s *= scale;
            if (dtype == CV_32F) ((float*)r.data)[(size_t)y * w + x] = (float)s;
            else if (dtype == CV_64F) ((double*)r.data)[(size_t)y * w + x] = s;
            else throw std::runtime_error("shim: boxFilter ddepth");
        }
    }
    dst.assign(r);
}

// 5x5 (or 3x3) median, BORDER_REPLICATE, CV_32F only (the only use on the path).
inline void medianBlur(const Mat& src, Mat& dst, int ksize) {
    if (src.type() != CV_32F) throw std::runtime_error("shim: medianBlur depth");
    const int r = ksize / 2, w = src.cols, h = src.rows, n = ksize * ksize;
    Mat out(h, w, CV_32F);
    const float* s = (const float*)src.data;
    float* o = (float*)out.data;
    std::vector<float> win(n);
    for (int y = 0; y < h; ++y) {
        for (int x = 0; x < w; ++x) {
            int k = 0;
            for (int dy = -r; dy <= r; ++dy) {
                const int yy = std::min(h - 1, std::max(0, y + dy));
                for (int dx = -r; dx <= r; ++dx) {
                    const int xx = std::min(w - 1, std::max(0, x + dx));
                    win[k++] = s[(size_t)yy * w + xx];
                }
            }
            std::nth_element(win.begin(), win.begin() + n / 2, win.end());
            o[(size_t)y * w + x] = win[n / 2];
        }
    }
    dst.assign(out);
}

}  // namespace cv

#endif
