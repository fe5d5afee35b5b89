// Pose algebra with Sophus::SE3f semantics and the 6x6 normal-equation solve, compiled for BOTH host and device:
// the synchronous C-ABI helpers (gsdf_pose_matrix, gsdf_pose_apply_increment, gsdf_llt_solve6) and the on-device
// Gauss-Newton step of track_kernel run the same code.  Poses are stored as Sophus stores them: unit quaternion
// (x, y, z, w) + translation; arrays are "tx ty tz qx qy qz qw".
//   Sophus so3.hpp:297-303 (normalize), :325-339 (product), :356-366 (rotate a point), :583-619 (expAndTheta);
//   se3.hpp:761-782 (exp); Eigen QuaternionBase::toRotationMatrix; Eigen LLT (RigidPointOptimizerOmp.cpp:93).
// The library is compiled with -fmad=false, so host and device round every product and sum the same way; only
// sinf/cosf differ by an ulp or two between glibc and CUDA.
#pragma once

#include <cuda_runtime.h>
#include <math.h>

namespace gsdf {

#define GSDF_HD __host__ __device__ __forceinline__

// On the device the solve and the pose update sit on the critical path of every Gauss-Newton pass of the persistent
// tracker (one thread per CTA, everyone else waits): fused multiply-adds and reciprocal-based divisions there.  These are
// tolerance-level quantities (poses are compared at 1e-4 .. 2e-3), not key-determining ones; the host versions keep
// Eigen's / Sophus' separate roundings (the library is compiled with -fmad=false, explicit fmaf still fuses).
#ifdef __CUDA_ARCH__
#define GSDF_FMA(a, b, c) fmaf((a), (b), (c))
#define GSDF_DIV(a, b) __fdividef((a), (b))
#else
#define GSDF_FMA(a, b, c) ((a) * (b) + (c))
#define GSDF_DIV(a, b) ((a) / (b))
#endif

GSDF_HD float dot3p(const float* a, const float* b) {
    const float p0 = a[0] * b[0], s = a[1] * b[1] + a[2] * b[2];  // Eigen: a0*b0 + (a1*b1 + a2*b2)
    return p0 + s;
}
GSDF_HD void matvec3p(const float* R, const float* v, float* o) {
    o[0] = dot3p(R, v); o[1] = dot3p(R + 3, v); o[2] = dot3p(R + 6, v);
}
GSDF_HD void quat_normalize(float* q) {  // Eigen sums the 4-float packet as (x^2 + z^2) + (y^2 + w^2)
    const float n2 = (q[0] * q[0] + q[2] * q[2]) + (q[1] * q[1] + q[3] * q[3]);
#ifdef __CUDA_ARCH__
    const float il = rsqrtf(n2);
#pragma unroll
    for (int i = 0; i < 4; ++i) q[i] *= il;
#else
    const float len = sqrtf(n2);
#pragma unroll
    for (int i = 0; i < 4; ++i) q[i] /= len;
#endif
}
GSDF_HD void quat_to_R(const float* q, float* R) {
    const float x = q[0], y = q[1], z = q[2], w = q[3];
    const float tx = 2.f * x, ty = 2.f * y, tz = 2.f * z;
    const float twx = tx * w, twy = ty * w, twz = tz * w;
    const float txx = tx * x, txy = ty * x, txz = tz * x, tyy = ty * y, tyz = tz * y, tzz = tz * z;
    R[0] = 1.f - (tyy + tzz); R[1] = txy - twz; R[2] = txz + twy;
    R[3] = txy + twz; R[4] = 1.f - (txx + tzz); R[5] = tyz - twx;
    R[6] = txz - twy; R[7] = tyz + twx; R[8] = 1.f - (txx + tyy);
}
GSDF_HD void cross3p(const float* a, const float* b, float* o) {
    o[0] = a[1] * b[2] - a[2] * b[1]; o[1] = a[2] * b[0] - a[0] * b[2]; o[2] = a[0] * b[1] - a[1] * b[0];
}
GSDF_HD void quat_rotate(const float* q, const float* p, float* o) {
    float uv[3], c2[3];
    cross3p(q, p, uv);
#pragma unroll
    for (int i = 0; i < 3; ++i) uv[i] += uv[i];
    cross3p(q, uv, c2);
#pragma unroll
    for (int i = 0; i < 3; ++i) o[i] = (p[i] + q[3] * uv[i]) + c2[i];
}
GSDF_HD void quat_mul(const float* a, const float* b, float* o) {  // product, then SO3(quaternion) normalises
    const float ax = a[0], ay = a[1], az = a[2], aw = a[3], bx = b[0], by = b[1], bz = b[2], bw = b[3];
    o[3] = aw * bw - ax * bx - ay * by - az * bz;
    o[0] = aw * bx + ax * bw + ay * bz - az * by;
    o[1] = aw * by + ay * bw + az * bx - ax * bz;
    o[2] = aw * bz + az * bw + ax * by - ay * bx;
    quat_normalize(o);
}
GSDF_HD void se3_exp(const float* a, float* q, float* tr) {
    const float* om = a + 3;
    const float eps = 1e-5f;  // Sophus::Constants<float>::epsilon()
    const float theta_sq = dot3p(om, om);
#ifdef __CUDA_ARCH__
    // On the device, for |omega| < 0.5 rad (a Gauss-Newton increment is a fraction of a degree): the four coefficients
    // sin(t/2)/t, cos(t/2), (1 - cos t)/t^2, (t - sin t)/t^3 as series in t^2 — Sophus' own small-angle branch carried further.
    // Truncation error < 1e-10, i.e. more accurate in float than the closed forms (the last two cancel), and no square
    // root, sine, cosine or division on the one thread every CTA waits for.
    if (theta_sq < 0.25f) {
        const float x = theta_sq;
        const float imag = fmaf(x, fmaf(x, fmaf(x, -1.f / 645120.f, 1.f / 3840.f), -1.f / 48.f), 0.5f);
        const float real = fmaf(x, fmaf(x, fmaf(x, fmaf(x, 1.f / 10321920.f, -1.f / 46080.f), 1.f / 384.f), -0.125f), 1.f);
        const float c1 = fmaf(x, fmaf(x, fmaf(x, fmaf(x, 1.f / 3628800.f, -1.f / 40320.f), 1.f / 720.f), -1.f / 24.f), 0.5f);
        const float c2 = fmaf(x, fmaf(x, fmaf(x, fmaf(x, 1.f / 39916800.f, -1.f / 362880.f), 1.f / 5040.f), -1.f / 120.f), 1.f / 6.f);
        q[0] = imag * om[0]; q[1] = imag * om[1]; q[2] = imag * om[2]; q[3] = real;
        // V = I + c1 * hat(om) + c2 * hat(om)^2, hat(om)^2 = om om^T - |om|^2 I
        const float V[9] = {1.f + c2 * (om[0] * om[0] - x), c2 * om[0] * om[1] - c1 * om[2], c2 * om[0] * om[2] + c1 * om[1],
                            c2 * om[0] * om[1] + c1 * om[2], 1.f + c2 * (om[1] * om[1] - x), c2 * om[1] * om[2] - c1 * om[0],
                            c2 * om[0] * om[2] - c1 * om[1], c2 * om[1] * om[2] + c1 * om[0], 1.f + c2 * (om[2] * om[2] - x)};
        matvec3p(V, a, tr);
        return;
    }
#endif
    float theta, imag, real;
    if (theta_sq < eps * eps) {
        theta = 0.f;
        const float p4 = theta_sq * theta_sq;
        imag = 0.5f - (float)(1.0 / 48.0) * theta_sq + (float)(1.0 / 3840.0) * p4;
        real = 1.f - (float)(1.0 / 8.0) * theta_sq + (float)(1.0 / 384.0) * p4;
    } else {
        theta = sqrtf(theta_sq);
        const float half = 0.5f * theta;
        imag = GSDF_DIV(sinf(half), theta);
        real = cosf(half);
    }
    q[0] = imag * om[0]; q[1] = imag * om[1]; q[2] = imag * om[2]; q[3] = real;
    const float O[9] = {0.f, -om[2], om[1], om[2], 0.f, -om[0], -om[1], om[0], 0.f};  // hat(omega)
    float O2[9], V[9];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j)
            O2[3 * i + j] = O[3 * i] * O[j] + (O[3 * i + 1] * O[3 + j] + O[3 * i + 2] * O[6 + j]);
    if (theta < eps) {
        quat_to_R(q, V);
    } else {
        const float tsq = theta * theta;
        const float c1 = GSDF_DIV(1.f - cosf(theta), tsq), c2 = GSDF_DIV(theta - sinf(theta), tsq * theta);
#pragma unroll
        for (int i = 0; i < 9; ++i) V[i] = ((i % 4 == 0 ? 1.f : 0.f) + c1 * O[i]) + c2 * O2[i];
    }
    matvec3p(V, a, tr);
}
// pose_out = SE3::exp(-xi) * pose_in   (RigidPointOptimizerOmp.cpp:102)
GSDF_HD void pose_apply_increment(const float* pose7_in, const float* xi6, float* pose7_out) {
    float a[6], qa[4], ta[3], qo[4], rt[3];
    const float qb[4] = {pose7_in[3], pose7_in[4], pose7_in[5], pose7_in[6]};
#pragma unroll
    for (int i = 0; i < 6; ++i) a[i] = -xi6[i];
    se3_exp(a, qa, ta);
    quat_mul(qa, qb, qo);
    quat_rotate(qa, pose7_in, rt);
#pragma unroll
    for (int i = 0; i < 3; ++i) pose7_out[i] = ta[i] + rt[i];
#pragma unroll
    for (int i = 0; i < 4; ++i) pose7_out[3 + i] = qo[i];
}
// xi = LLT(H)^-1 g: unblocked lower Cholesky in place, stopping at the first non-positive pivot as Eigen's LLT
// does (the rest of the matrix is then left as is and the substitution yields inf/NaN for H = 0), then
// forward/back substitution.  H row-major 6x6.
GSDF_HD void llt_solve6(const float* H36, const float* g6, float* xi6) {
    // every loop has constant bounds and is fully unrolled, so that on the device the matrix lives in registers
    // (the solve runs on one thread per CTA between two barriers of the persistent tracker: it is on the
    // critical path of every Gauss-Newton pass)
    // (loops run over the full fixed range with a predicate instead of triangular bounds: nvcc keeps A in
    // registers only then)
    float A[36], y[6], x6[6], inv[6];
#pragma unroll
    for (int i = 0; i < 36; ++i) A[i] = H36[i];
    bool ok = true;  // false once a pivot was non-positive: Eigen stops there and leaves the remaining columns as is
    // Divisions by a pivot are done as one reciprocal per column and multiplications (<= 1 ulp away from Eigen's
    // divisions, far inside the solve's tolerance): the 12 dependent IEEE divisions of the two substitutions were
    // a third of the single-thread critical path that every Gauss-Newton pass of the persistent tracker waits for.
#pragma unroll
    for (int k = 0; k < 6; ++k) {
        float x = A[6 * k + k];
#pragma unroll
        for (int j = 0; j < 6; ++j)
            if (j < k) x = GSDF_FMA(-A[6 * k + j], A[6 * k + j], x);
        const bool good = ok && !(x <= 0.f);
#ifdef __CUDA_ARCH__
        // on the device one MUFU.RSQ (2 ulp) replaces the IEEE square root and reciprocal (two ~60-cycle sequences on
        // the dependent chain of each of the 6 columns); the solve is a tolerance-level quantity (tests: 1e-4 relative)
        const float ir = rsqrtf(x);
        const float r = x * ir;
#else
        const float r = sqrtf(x);
        const float ir = 1.f / r;
#endif
        A[6 * k + k] = good ? r : A[6 * k + k];
        inv[k] = good ? ir : GSDF_DIV(1.f, A[6 * k + k]);
#pragma unroll
        for (int i = 0; i < 6; ++i)
            if (i > k) {
                float s = A[6 * i + k];
#pragma unroll
                for (int j = 0; j < 6; ++j)
                    if (j < k) s = GSDF_FMA(-A[6 * i + j], A[6 * k + j], s);
                A[6 * i + k] = good ? s * ir : A[6 * i + k];
            }
        ok = good;
    }
#pragma unroll
    for (int i = 0; i < 6; ++i) {
        float s = g6[i];
#pragma unroll
        for (int j = 0; j < 6; ++j)
            if (j < i) s = GSDF_FMA(-A[6 * i + j], y[j], s);
        y[i] = s * inv[i];
    }
#pragma unroll
    for (int ii = 0; ii < 6; ++ii) {
        const int i = 5 - ii;
        float s = y[i];
#pragma unroll
        for (int j = 0; j < 6; ++j)
            if (j > i) s = GSDF_FMA(-A[6 * j + i], x6[j], s);
        x6[i] = s * inv[i];
    }
#pragma unroll
    for (int i = 0; i < 6; ++i) xi6[i] = x6[i];
}

// One Gauss-Newton step of RigidPointOptimizer::optimize_sampled (RigidPointOptimizerOmp.cpp:93-102) from the
// 29 reduced sums (g[6], upper H[21], E, counter).  Returns 1 if |xi|^2 < thr_sq (converged: pose NOT updated),
// else 0 after pose7 <- exp(-xi) * pose7 (skipped when xi has a NaN).
GSDF_HD int gauss_newton_step(const float* sums29, float damping, float thr_sq, float* pose7) {
    float H[36], xi[6];
    int o = 6;
#pragma unroll
    for (int a = 0; a < 6; ++a)
#pragma unroll
        for (int b = a; b < 6; ++b) { H[6 * a + b] = sums29[o]; H[6 * b + a] = sums29[o]; ++o; }
    llt_solve6(H, sums29, xi);
    bool has_nan = false;
    float nrm = 0.f;
#pragma unroll
    for (int i = 0; i < 6; ++i) { xi[i] = damping * xi[i]; has_nan = has_nan || (xi[i] != xi[i]); nrm += xi[i] * xi[i]; }
    if (nrm < thr_sq) return 1;
    if (!has_nan) {
        float np[7];
        pose_apply_increment(pose7, xi, np);
        for (int i = 0; i < 7; ++i) pose7[i] = np[i];
    }
    return 0;
}

}  // namespace gsdf
