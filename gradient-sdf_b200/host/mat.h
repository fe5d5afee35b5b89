// Minimal stand-ins for the types the reference's class surface uses (cpp/include/mat.h:47-66: Eigen Vec3f / Mat3f /
// Vec6f typedefs and Sophus::SE3f) and for the slice of cv::Mat its hot path touches, so that the forwarding
// shells in this directory build on a machine without Eigen, Sophus and OpenCV (the GPU box has none of them).
// Inside the reference tree use the real types instead: INTEGRATION.md shows the two-line conversions.
#ifndef GSDF_HOST_MAT_H_
#define GSDF_HOST_MAT_H_

#include <cstddef>
#include <cmath>
#include <cstring>

#include "gsdf_b200.h"

struct Vec3f {
    float v[3];
    Vec3f() : v{0.f, 0.f, 0.f} {}
    Vec3f(float x, float y, float z) : v{x, y, z} {}
    float& operator[](int i) { return v[i]; }
    const float& operator[](int i) const { return v[i]; }
    const float* data() const { return v; }
    static Vec3f Zero() { return Vec3f(); }
};

struct Vec3i {
    int v[3];
    Vec3i() : v{0, 0, 0} {}
    Vec3i(int x, int y, int z) : v{x, y, z} {}
    int& operator[](int i) { return v[i]; }
    const int& operator[](int i) const { return v[i]; }
};

struct Vec6f {
    float v[6];
    Vec6f() : v{0, 0, 0, 0, 0, 0} {}
    float& operator[](int i) { return v[i]; }
    const float& operator[](int i) const { return v[i]; }
};

struct Mat3f {  // row-major storage, (i, j) access like Eigen
    float m[9];
    Mat3f() : m{1, 0, 0, 0, 1, 0, 0, 0, 1} {}
    float& operator()(int i, int j) { return m[3 * i + j]; }
    const float& operator()(int i, int j) const { return m[3 * i + j]; }
    static Mat3f Identity() { return Mat3f(); }
};

// Sophus::SE3f as the reference uses it: unit quaternion + translation, camera-to-world.
class SE3 {
public:
    SE3() : p_{0, 0, 0, 0, 0, 0, 1} {}
    // SE3(quaternion (w, x, y, z), translation): normalises the quaternion like Sophus' constructor
    SE3(float qw, float qx, float qy, float qz, const Vec3f& t) {
        const float raw[7] = {t[0], t[1], t[2], qx, qy, qz, qw};
        float R[9];
        gsdf_pose_matrix(raw, R, p_);
    }
    // a stored pose "tx ty tz qx qy qz qw" with a unit quaternion (no renormalisation)
    static SE3 fromStored(const float* pose7) { SE3 s; std::memcpy(s.p_, pose7, sizeof(s.p_)); return s; }
    // rotationMatrix() / matrix().topLeftCorner(3,3): Eigen's toRotationMatrix of the STORED unit quaternion (Sophus does
    // not renormalise here)
    Mat3f rotationMatrix() const {
        Mat3f R;
        const float x = p_[3], y = p_[4], z = p_[5], w = p_[6];
        const float tx = 2.f * x, ty = 2.f * y, tz = 2.f * z;
        const float twx = tx * w, twy = ty * w, twz = tz * w;
        const float txx = tx * x, txy = ty * x, txz = tz * x, tyy = ty * y, tyz = tz * y, tzz = tz * z;
        R.m[0] = 1.f - (tyy + tzz); R.m[1] = txy - twz; R.m[2] = txz + twy;
        R.m[3] = txy + twz; R.m[4] = 1.f - (txx + tzz); R.m[5] = tyz - twx;
        R.m[6] = txz - twy; R.m[7] = tyz + twx; R.m[8] = 1.f - (txx + tyy);
        return R;
    }
    // Eigen::Quaternion<float> q(R) (Eigen/src/Geometry/Quaternion.h, quaternionbase_assign_impl<Other,3,3>): what
    // main_scan_3d.cpp:275-276 writes into <results>_poses.txt; out = (x, y, z, w)
    static void quaternionFromMatrix(const Mat3f& m, float* out) {
        float t = m(0, 0) + (m(1, 1) + m(2, 2));
        if (t > 0.f) {
            t = std::sqrt(t + 1.f);
            out[3] = 0.5f * t;
            t = 0.5f / t;
            out[0] = (m(2, 1) - m(1, 2)) * t;
            out[1] = (m(0, 2) - m(2, 0)) * t;
            out[2] = (m(1, 0) - m(0, 1)) * t;
        } else {
            int i = 0;
            if (m(1, 1) > m(0, 0)) i = 1;
            if (m(2, 2) > m(i, i)) i = 2;
            const int j = (i + 1) % 3, k = (j + 1) % 3;
            t = std::sqrt(m(i, i) - m(j, j) - m(k, k) + 1.f);
            out[i] = 0.5f * t;
            t = 0.5f / t;
            out[3] = (m(k, j) - m(j, k)) * t;
            out[j] = (m(j, i) + m(i, j)) * t;
            out[k] = (m(k, i) + m(i, k)) * t;
        }
    }
    Vec3f translation() const { return Vec3f(p_[0], p_[1], p_[2]); }
    static SE3 exp(const Vec6f& xi) {  // se3.hpp:761-782
        Vec6f neg;
        for (int i = 0; i < 6; ++i) neg[i] = -xi[i];
        SE3 id, out;
        gsdf_pose_apply_increment(id.p_, neg.v, out.p_);  // exp(-(-xi)) * identity
        return out;
    }
    const float* data() const { return p_; }  // tx ty tz qx qy qz qw
    float* data() { return p_; }

private:
    float p_[7];
};

#ifndef CV_32FC1
#define CV_32FC1 5
#define CV_32F 5
#endif

namespace cv {
struct Size {
    int width, height;
    Size() : width(0), height(0) {}
    Size(int w, int h) : width(w), height(h) {}
};
// single-channel float image over caller-owned memory (the depth frames of ImageLoader.h:159-175)
struct Mat {
    int rows = 0, cols = 0;
    unsigned char* data = nullptr;
    Mat() {}
    Mat(int r, int c, int /*type*/, void* ext) : rows(r), cols(c), data((unsigned char*)ext) {}
    bool empty() const { return data == nullptr; }
};
}  // namespace cv

#endif  // GSDF_HOST_MAT_H_
