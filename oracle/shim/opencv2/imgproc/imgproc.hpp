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
// Every output keeps exactly that order of additions (so the results are bit-identical to a naive double loop); the
// loops are arranged tap-outermost over a reflect-padded row so that the compiler vectorises ACROSS pixels, and rows
// run in parallel — a stand-in for real OpenCV must not inflate the timed CPU baseline (real cv::boxFilter takes well
// under a millisecond per 640x480 plane).
inline void boxFilter(const Mat& src, Mat& dst, int ddepth, Size ksize, Point anchor = Point(-1, -1),
                      bool normalize = true) {
    (void)anchor;
    const int rx = ksize.width / 2, ry = ksize.height / 2;
    const int w = src.cols, h = src.rows;
    const int dtype = ddepth < 0 ? src.type() : ddepth;
    if (src.type() != CV_32F && src.type() != CV_64F) throw std::runtime_error("shim: boxFilter depth");
    if (dtype != CV_32F && dtype != CV_64F) throw std::runtime_error("shim: boxFilter ddepth");
    std::vector<double> rowsum((size_t)w * h);
    std::vector<int> xr((size_t)w + 2 * rx);
    for (int x = -rx; x < w + rx; ++x) xr[x + rx] = shim_detail::reflect101(x, w);
    const bool f32 = src.type() == CV_32F;
#pragma omp parallel
    {
        std::vector<double> pad((size_t)w + 2 * rx);
#pragma omp for schedule(static)
        for (int y = 0; y < h; ++y) {
            if (f32) {
                const float* p = (const float*)src.data + (size_t)y * w;
                for (int i = 0; i < w + 2 * rx; ++i) pad[i] = (double)p[xr[i]];
            } else {
                const double* p = (const double*)src.data + (size_t)y * w;
                for (int i = 0; i < w + 2 * rx; ++i) pad[i] = p[xr[i]];
            }
            double* rs = &rowsum[(size_t)y * w];
            const double* pd = pad.data();
            for (int x = 0; x < w; ++x) rs[x] = 0.0;
            for (int d = 0; d <= 2 * rx; ++d)
                for (int x = 0; x < w; ++x) rs[x] += pd[x + d];
        }
    }
    Mat r(h, w, dtype);
    const double scale = normalize ? 1.0 / ((2 * rx + 1) * (2 * ry + 1)) : 1.0;
#pragma omp parallel
    {
        std::vector<double> acc((size_t)w);
#pragma omp for schedule(static)
        for (int y = 0; y < h; ++y) {
            double* s = acc.data();
            for (int x = 0; x < w; ++x) s[x] = 0.0;
            for (int d = -ry; d <= ry; ++d) {
                const double* rs = &rowsum[(size_t)shim_detail::reflect101(y + d, h) * w];
                for (int x = 0; x < w; ++x) s[x] += rs[x];
            }
            if (normalize)
                for (int x = 0; x < w; ++x) s[x] *= scale;
            if (dtype == CV_32F) {
                float* o = (float*)r.data + (size_t)y * w;
                for (int x = 0; x < w; ++x) o[x] = (float)s[x];
            } else {
                double* o = (double*)r.data + (size_t)y * w;
                for (int x = 0; x < w; ++x) o[x] = s[x];
            }
        }
    }
    dst.assign(r);
}

namespace shim_detail {
typedef float v4f __attribute__((vector_size(16)));
static inline void cswap(v4f& a, v4f& b) {
    const v4f lo = a < b ? a : b, hi = a < b ? b : a;
    a = lo; b = hi;
}
// 99 compare-exchanges that leave the median of 25 values in slot 12 (checked against std::nth_element on 2 million
// random windows, ties included); the median is a value of the input, so any exact selection gives identical bits.
static const unsigned char kMed25[99][2] = {
    {0,1},{3,4},{2,4},{2,3},{6,7},{5,7},{5,6},{9,10},{8,10},{8,9},{12,13},{11,13},{11,12},{15,16},{14,16},{14,15},{18,19},
    {17,19},{17,18},{21,22},{20,22},{20,21},{23,24},{2,5},{3,6},{0,6},{0,3},{4,7},{1,7},{1,4},{11,14},{8,14},{8,11},{12,15},
    {9,15},{9,12},{13,16},{10,16},{10,13},{20,23},{17,23},{17,20},{21,24},{18,24},{18,21},{19,22},{8,17},{9,18},{0,18},{0,9},
    {10,19},{1,19},{1,10},{11,20},{2,20},{2,11},{12,21},{3,21},{3,12},{13,22},{4,22},{4,13},{14,23},{5,23},{5,14},{15,24},
    {6,24},{6,15},{7,16},{7,19},{13,21},{15,23},{7,13},{7,15},{1,9},{3,11},{5,17},{11,17},{9,17},{4,10},{6,12},{7,14},{4,6},
    {4,7},{12,14},{10,14},{6,7},{10,12},{6,10},{6,17},{12,17},{7,17},{7,10},{12,18},{7,12},{10,18},{12,20},{10,20},{10,12}};
}  // namespace shim_detail

// 5x5 (or 3x3) median, BORDER_REPLICATE, CV_32F only (the only use on the path).  5x5: a selection network over four
// pixels at a time on a replicate-padded copy, rows in parallel (real cv::medianBlur takes ~1 ms per 640x480 frame; the
// per-pixel std::nth_element this replaces took 43 ms and was charged to the reference's timed fusion).
inline void medianBlur(const Mat& src, Mat& dst, int ksize) {
    if (src.type() != CV_32F) throw std::runtime_error("shim: medianBlur depth");
    const int r = ksize / 2, w = src.cols, h = src.rows, n = ksize * ksize;
    Mat out(h, w, CV_32F);
    const float* s = (const float*)src.data;
    float* o = (float*)out.data;
    if (ksize == 5) {
        const int pw = w + 4 + 4, ph = h + 4;  // 4 spare columns so that 4-wide loads at the right edge stay inside
        std::vector<float> pad((size_t)pw * ph);
        for (int y = 0; y < ph; ++y) {
            const float* row = s + (size_t)std::min(h - 1, std::max(0, y - 2)) * w;
            float* pr = &pad[(size_t)y * pw];
            for (int x = 0; x < pw; ++x) pr[x] = row[std::min(w - 1, std::max(0, x - 2))];
        }
#pragma omp parallel for schedule(static)
        for (int y = 0; y < h; ++y) {
            for (int x = 0; x < w; x += 4) {
                shim_detail::v4f p[25];
                for (int dy = 0; dy < 5; ++dy)
                    for (int dx = 0; dx < 5; ++dx)
                        std::memcpy(&p[dy * 5 + dx], &pad[(size_t)(y + dy) * pw + x + dx], 16);
                for (const auto& e : shim_detail::kMed25) shim_detail::cswap(p[e[0]], p[e[1]]);
                float m[4];
                std::memcpy(m, &p[12], 16);
                for (int i = 0; i < 4 && x + i < w; ++i) o[(size_t)y * w + x + i] = m[i];
            }
        }
        dst.assign(out);
        return;
    }
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
