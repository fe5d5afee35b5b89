// TEST INFRASTRUCTURE — not part of the product.
//
// Minimal single-channel stand-in for the slice of OpenCV 4 that the reference's
// hot-path translation units touch (cv::Mat, cv::Mat_<T>, element-wise arithmetic,
// divide, sqrt, repeat, boxFilter, medianBlur).  OpenCV's C++ library is not in this
// image, so this header DEFINES the oracle's semantics at the OpenCV boundary
// ("parity unpinned" there):
//   * element-wise float ops are done in float, double ops in double (as OpenCV does);
//   * boxFilter(normalize=false) = BORDER_REFLECT_101 window sum, accumulated in double
//     (OpenCV: sumType CV_64F for CV_32F input), separable: row sums left→right, then
//     column sums top→bottom, rounded once to the destination depth;
//   * medianBlur(5) = BORDER_REPLICATE 5x5 median.
// Call sites it serves: normals/NormalEstimator.h:65-204, MapGradPixelSdfOmp.cpp:51-52,
// MapPixelSdfOmp.cpp:121-122 (paths relative to /root/reference/cpp/include).
#ifndef GSDF_ORACLE_SHIM_OPENCV_CORE_HPP_
#define GSDF_ORACLE_SHIM_OPENCV_CORE_HPP_

#include <algorithm>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <type_traits>
#include <vector>

#define CV_8U 0
#define CV_16U 2
#define CV_32F 5
#define CV_64F 6
#define CV_8UC1 CV_8U
#define CV_16UC1 CV_16U
#define CV_32FC1 CV_32F
#define CV_64FC1 CV_64F

namespace cv {

typedef unsigned char uchar;

struct Size {
    int width, height;
    Size() : width(0), height(0) {}
    Size(int w, int h) : width(w), height(h) {}
    bool operator==(const Size& o) const { return width == o.width && height == o.height; }
};

struct Point {
    int x, y;
    Point() : x(0), y(0) {}
    Point(int x_, int y_) : x(x_), y(y_) {}
};

template <typename T> struct DepthOf;
template <> struct DepthOf<uchar> { enum { value = CV_8U }; };
template <> struct DepthOf<unsigned short> { enum { value = CV_16U }; };
template <> struct DepthOf<float> { enum { value = CV_32F }; };
template <> struct DepthOf<double> { enum { value = CV_64F }; };

inline size_t elem_size(int type) {
    switch (type) {
        case CV_8U: return 1;
        case CV_16U: return 2;
        case CV_32F: return 4;
        case CV_64F: return 8;
    }
    throw std::runtime_error("shim: unsupported depth");
}

class Mat {
public:
    int rows, cols;
    uchar* data;

    Mat() : rows(0), cols(0), data(nullptr), type_(CV_8U) {}
    Mat(int r, int c, int type) : rows(0), cols(0), data(nullptr), type_(type) { create(r, c, type); }
    Mat(Size s, int type) : rows(0), cols(0), data(nullptr), type_(type) { create(s.height, s.width, type); }
    // borrowed buffer (no ownership), as in cv::Mat(rows, cols, type, void*)
    Mat(int r, int c, int type, void* ext) : rows(r), cols(c), data((uchar*)ext), type_(type) {}
    template <typename T>
    explicit Mat(const std::vector<T>& v) : rows(0), cols(0), data(nullptr), type_(DepthOf<T>::value) {
        create((int)v.size(), 1, DepthOf<T>::value);
        if (!v.empty()) std::memcpy(data, v.data(), v.size() * sizeof(T));
    }
    Mat(const Mat& m) : rows(m.rows), cols(m.cols), data(m.data), type_(m.type_), buf_(m.buf_) {}
    virtual ~Mat() {}

    Mat& operator=(const Mat& m) {
        assign(m);
        return *this;
    }

    int type() const { return type_; }
    int depth() const { return type_; }
    int channels() const { return 1; }
    Size size() const { return Size(cols, rows); }
    size_t total() const { return (size_t)rows * (size_t)cols; }
    bool empty() const { return data == nullptr || total() == 0; }
    bool isContinuous() const { return true; }

    void create(int r, int c, int type) {
        const int ft = fixed_type();
        if (ft >= 0) type = ft;
        if (data && r == rows && c == cols && type == type_ && buf_) return;
        rows = r; cols = c; type_ = type;
        const size_t bytes = (size_t)r * c * elem_size(type);
        buf_.reset(new uchar[bytes ? bytes : 1], std::default_delete<uchar[]>());
        data = buf_.get();
    }

    static Mat zeros(Size s, int type) { return zeros(s.height, s.width, type); }
    static Mat zeros(int r, int c, int type) {
        Mat m(r, c, type);
        std::memset(m.data, 0, m.total() * elem_size(type));
        return m;
    }

    template <typename T> T* ptr(int r = 0) { return (T*)data + (size_t)r * cols; }
    template <typename T> const T* ptr(int r = 0) const { return (const T*)data + (size_t)r * cols; }
    template <typename T> T& at(int r, int c) { return ((T*)data)[(size_t)r * cols + c]; }
    template <typename T> const T& at(int r, int c) const { return ((const T*)data)[(size_t)r * cols + c]; }

    Mat clone() const {
        Mat m(rows, cols, type_);
        std::memcpy(m.data, data, total() * elem_size(type_));
        return m;
    }

    // single channel only: cn must be 1; shares the buffer
    Mat reshape(int cn, int new_rows) const {
        if (cn != 1 || new_rows <= 0 || total() % (size_t)new_rows) throw std::runtime_error("shim: reshape");
        Mat m(*this);
        m.rows = new_rows;
        m.cols = (int)(total() / (size_t)new_rows);
        return m;
    }

    Mat t() const {
        Mat m(cols, rows, type_);
        const size_t es = elem_size(type_);
        for (int r = 0; r < rows; ++r)
            for (int c = 0; c < cols; ++c)
                std::memcpy(m.data + ((size_t)c * rows + r) * es, data + ((size_t)r * cols + c) * es, es);
        return m;
    }

    Mat mul(const Mat& o) const;

    // dst = saturate_cast<rtype>(src * alpha + beta); for 32F destinations from integer depths the
    // product is formed in float (float(v) * float(alpha)), which is what ImageLoader.h:172 relies on.
    void convertTo(Mat& dst, int rtype, double alpha = 1.0, double beta = 0.0) const;

    void copyTo(Mat& dst) const { dst.assign(clone()); }
    void copyTo(Mat& dst, const Mat& mask) const {
        if (dst.rows != rows || dst.cols != cols || dst.type() != type_) dst.assign(zeros(rows, cols, type_));
        const size_t es = elem_size(type_);
        const size_t n = total();
        for (size_t i = 0; i < n; ++i)
            if (mask.data[i]) std::memcpy(dst.data + i * es, data + i * es, es);
    }

    // "assign keeping my depth": plain Mat shares the buffer; Mat_<T> converts when depths differ.
    void assign(const Mat& m) {
        const int ft = fixed_type();
        if (ft < 0 || ft == m.type_ || m.data == nullptr) {
            rows = m.rows; cols = m.cols; data = m.data; buf_ = m.buf_;
            type_ = (ft >= 0) ? ft : m.type_;
        } else {
            Mat tmp;
            m.convertTo(tmp, ft);
            rows = tmp.rows; cols = tmp.cols; data = tmp.data; buf_ = tmp.buf_; type_ = ft;
        }
    }

protected:
    virtual int fixed_type() const { return -1; }
    int type_;
    std::shared_ptr<uchar> buf_;
};

template <typename T>
class Mat_ : public Mat {
public:
    Mat_() : Mat() { type_ = DepthOf<T>::value; }
    Mat_(int r, int c) : Mat() { type_ = DepthOf<T>::value; create(r, c, DepthOf<T>::value); }
    Mat_(const Mat& m) : Mat() { type_ = DepthOf<T>::value; assign(m); }
    Mat_(const Mat_& m) : Mat() { type_ = DepthOf<T>::value; assign(m); }
    template <typename U>
    Mat_(const Mat_<U>& m) : Mat() { type_ = DepthOf<T>::value; assign(m); }
    Mat_& operator=(const Mat& m) { assign(m); return *this; }
    Mat_& operator=(const Mat_& m) { assign(m); return *this; }
    template <typename U>
    Mat_& operator=(const Mat_<U>& m) { assign(m); return *this; }
    T& operator()(int r, int c) { return ((T*)data)[(size_t)r * cols + c]; }
    const T& operator()(int r, int c) const { return ((const T*)data)[(size_t)r * cols + c]; }

protected:
    int fixed_type() const override { return DepthOf<T>::value; }
};

namespace shim_detail {

inline void check_same(const Mat& a, const Mat& b) {
    if (a.rows != b.rows || a.cols != b.cols || a.type() != b.type())
        throw std::runtime_error("shim: size/type mismatch");
}

template <typename F32, typename F64>
inline Mat binary(const Mat& a, const Mat& b, F32 f32, F64 f64) {
    check_same(a, b);
    Mat r(a.rows, a.cols, a.type());
    const size_t n = a.total();
    if (a.type() == CV_32F) {
        const float* x = (const float*)a.data; const float* y = (const float*)b.data; float* z = (float*)r.data;
        for (size_t i = 0; i < n; ++i) z[i] = f32(x[i], y[i]);
    } else if (a.type() == CV_64F) {
        const double* x = (const double*)a.data; const double* y = (const double*)b.data; double* z = (double*)r.data;
        for (size_t i = 0; i < n; ++i) z[i] = f64(x[i], y[i]);
    } else {
        throw std::runtime_error("shim: arithmetic on non-float depth");
    }
    return r;
}

template <typename F32, typename F64>
inline Mat unary(const Mat& a, F32 f32, F64 f64) {
    Mat r(a.rows, a.cols, a.type());
    const size_t n = a.total();
    if (a.type() == CV_32F) {
        const float* x = (const float*)a.data; float* z = (float*)r.data;
        for (size_t i = 0; i < n; ++i) z[i] = f32(x[i]);
    } else if (a.type() == CV_64F) {
        const double* x = (const double*)a.data; double* z = (double*)r.data;
        for (size_t i = 0; i < n; ++i) z[i] = f64(x[i]);
    } else {
        throw std::runtime_error("shim: arithmetic on non-float depth");
    }
    return r;
}

template <typename S>
inline double load_as_double(const uchar* p, int type, size_t i) {
    (void)sizeof(S);
    switch (type) {
        case CV_8U: return (double)p[i];
        case CV_16U: return (double)((const unsigned short*)p)[i];
        case CV_32F: return (double)((const float*)p)[i];
        case CV_64F: return ((const double*)p)[i];
    }
    throw std::runtime_error("shim: depth");
}

inline int reflect101(int p, int len) {
    if (len == 1) return 0;
    while (p < 0 || p >= len) {
        if (p < 0) p = -p;
        else p = 2 * (len - 1) - p;
    }
    return p;
}

}  // namespace shim_detail

inline Mat Mat::mul(const Mat& o) const {
    return shim_detail::binary(*this, o, [](float a, float b) { return a * b; }, [](double a, double b) { return a * b; });
}

inline void Mat::convertTo(Mat& dst, int rtype, double alpha, double beta) const {
    if (rtype < 0) rtype = type_;
    Mat r(rows, cols, rtype);
    const size_t n = total();
    const bool scaled = !(alpha == 1.0 && beta == 0.0);
    for (size_t i = 0; i < n; ++i) {
        if (rtype == CV_32F) {
            float v;
            if (type_ == CV_64F) {
                const double s = ((const double*)data)[i];
                v = (float)(scaled ? s * alpha + beta : s);
            } else {
                const float s = (type_ == CV_32F) ? ((const float*)data)[i]
                              : (type_ == CV_16U) ? (float)((const unsigned short*)data)[i]
                                                  : (float)data[i];
                v = scaled ? s * (float)alpha + (float)beta : s;
            }
            ((float*)r.data)[i] = v;
        } else if (rtype == CV_64F) {
            const double s = shim_detail::load_as_double<void>(data, type_, i);
            ((double*)r.data)[i] = scaled ? s * alpha + beta : s;
        } else {
            const double s = shim_detail::load_as_double<void>(data, type_, i) * alpha + beta;
            const double rr = std::nearbyint(s);
            if (rtype == CV_8U) r.data[i] = (uchar)std::min(255.0, std::max(0.0, rr));
            else ((unsigned short*)r.data)[i] = (unsigned short)std::min(65535.0, std::max(0.0, rr));
        }
    }
    dst.assign(r);
}

inline Mat operator+(const Mat& a, const Mat& b) {
    return shim_detail::binary(a, b, [](float x, float y) { return x + y; }, [](double x, double y) { return x + y; });
}
inline Mat operator-(const Mat& a, const Mat& b) {
    return shim_detail::binary(a, b, [](float x, float y) { return x - y; }, [](double x, double y) { return x - y; });
}
inline Mat operator*(double s, const Mat& a) {
    const float sf = (float)s;
    return shim_detail::unary(a, [sf](float x) { return x * sf; }, [s](double x) { return x * s; });
}
inline Mat operator*(const Mat& a, double s) { return s * a; }
inline Mat operator+(double s, const Mat& a) {
    const float sf = (float)s;
    return shim_detail::unary(a, [sf](float x) { return x + sf; }, [s](double x) { return x + s; });
}
inline Mat operator+(const Mat& a, double s) { return s + a; }

// comparison → CV_8U mask (0 / 255)
inline Mat operator!=(const Mat& a, double s) {
    Mat r(a.rows, a.cols, CV_8U);
    const size_t n = a.total();
    for (size_t i = 0; i < n; ++i) r.data[i] = (shim_detail::load_as_double<void>(a.data, a.type(), i) != s) ? 255 : 0;
    return r;
}

// dst = scale / src, plain IEEE division (OpenCV 4: no divide-by-zero special case for floats)
inline void divide(double scale, const Mat& src, Mat& dst) {
    const float sf = (float)scale;
    dst.assign(shim_detail::unary(src, [sf](float x) { return sf / x; }, [scale](double x) { return scale / x; }));
}
inline void divide(const Mat& a, const Mat& b, Mat& dst) {
    dst.assign(shim_detail::binary(a, b, [](float x, float y) { return x / y; }, [](double x, double y) { return x / y; }));
}
inline void sqrt(const Mat& src, Mat& dst) {
    dst.assign(shim_detail::unary(src, [](float x) { return std::sqrt(x); }, [](double x) { return std::sqrt(x); }));
}

inline void repeat(const Mat& src, int ny, int nx, Mat& dst) {
    Mat r(src.rows * ny, src.cols * nx, src.type());
    const size_t es = elem_size(src.type());
    for (int y = 0; y < r.rows; ++y)
        for (int x = 0; x < r.cols; ++x)
            std::memcpy(r.data + ((size_t)y * r.cols + x) * es,
                        src.data + ((size_t)(y % src.rows) * src.cols + (x % src.cols)) * es, es);
    dst.assign(r);
}

}  // namespace cv

#endif  // GSDF_ORACLE_SHIM_OPENCV_CORE_HPP_
