// Keypoint undistortion: Frame::UndistortKeyPoints / ComputeImageBounds (src/datastructures/frame.cpp:758-791,
// 793-821), i.e. cv::undistortPoints(pts, pts, mK, mDistCoef, Mat(), mK) for the radtan model with OpenCV's default
// termination (5 fixed-point iterations), in double like cvUndistortPointsInternal.  Every product, sum and quotient
// goes through a round-to-nearest intrinsic so that nothing is contracted into an FMA: the result is the same
// double, hence the same float, as OpenCV's.  A few hundred FP64 operations per keypoint; the kernels that call
// this are bound by their keypoint loads/stores.
#pragma once
#include "viorb_common.cuh"

struct CamModel {
    double fx, fy, cx, cy, ifx, ify;   // mK (CV_32F in the reference, widened) and 1/fx, 1/fy
    double k[12];                      // OpenCV coefficient order k1 k2 p1 p2 k3 k4 k5 k6 s1 s2 s3 s4
    int mode;                          // 0 = none, 1 = copy (k1 == 0, frame.cpp:762), 2 = undistort
};

#define UD_MUL(a, b) __dmul_rn((a), (b))
#define UD_ADD(a, b) __dadd_rn((a), (b))
#define UD_SUB(a, b) __dsub_rn((a), (b))

__device__ __forceinline__ float2 undistort_point_dev(float px, float py, const CamModel& c) {
    const double u = (double)px, v = (double)py;
    double x = UD_MUL(UD_SUB(u, c.cx), c.ifx), y = UD_MUL(UD_SUB(v, c.cy), c.ify);
    const double x0 = x, y0 = y;
    const double* k = c.k;
#pragma unroll 1
    for (int j = 0; j < 5; ++j) {
        const double r2 = UD_ADD(UD_MUL(x, x), UD_MUL(y, y));
        const double num = UD_ADD(1.0, UD_MUL(UD_ADD(UD_MUL(UD_ADD(UD_MUL(k[7], r2), k[6]), r2), k[5]), r2));
        const double den = UD_ADD(1.0, UD_MUL(UD_ADD(UD_MUL(UD_ADD(UD_MUL(k[4], r2), k[1]), r2), k[0]), r2));
        const double icdist = __ddiv_rn(num, den);
        if (icdist < 0) {
            x = x0; y = y0;     // == (u - cx) * ifx, (v - cy) * ify
            break;
        }
        const double xx2 = UD_MUL(UD_MUL(2.0, x), x), yy2 = UD_MUL(UD_MUL(2.0, y), y);
        // deltaX = 2*k2*x*y + k3*(r2 + 2*x*x) + k8*r2 + k9*r2*r2, left to right
        double dX = UD_MUL(UD_MUL(UD_MUL(2.0, k[2]), x), y);
        dX = UD_ADD(dX, UD_MUL(k[3], UD_ADD(r2, xx2)));
        dX = UD_ADD(dX, UD_MUL(k[8], r2));
        dX = UD_ADD(dX, UD_MUL(UD_MUL(k[9], r2), r2));
        // deltaY = k2*(r2 + 2*y*y) + 2*k3*x*y + k10*r2 + k11*r2*r2
        double dY = UD_MUL(k[2], UD_ADD(r2, yy2));
        dY = UD_ADD(dY, UD_MUL(UD_MUL(UD_MUL(2.0, k[3]), x), y));
        dY = UD_ADD(dY, UD_MUL(k[10], r2));
        dY = UD_ADD(dY, UD_MUL(UD_MUL(k[11], r2), r2));
        x = UD_MUL(UD_SUB(x0, dX), icdist);
        y = UD_MUL(UD_SUB(y0, dY), icdist);
    }
    // P = K, R = I:  xx = fx*x + 0*y + cx, yy = 0*x + fy*y + cy, ww = 1/(0*x + 0*y + 1)
    const double xx = UD_ADD(UD_ADD(UD_MUL(c.fx, x), UD_MUL(0.0, y)), c.cx);
    const double yy = UD_ADD(UD_ADD(UD_MUL(0.0, x), UD_MUL(c.fy, y)), c.cy);
    const double ww = __ddiv_rn(1.0, UD_ADD(UD_ADD(UD_MUL(0.0, x), UD_MUL(0.0, y)), 1.0));
    return make_float2(__double2float_rn(UD_MUL(xx, ww)), __double2float_rn(UD_MUL(yy, ww)));
}

// Raw points: xy [n] -> out [n].
__global__ void __launch_bounds__(256) k_undistort_points(const float2* __restrict__ xy, int n, CamModel cam, float2* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float2 p = xy[i];
    out[i] = cam.mode == 2 ? undistort_point_dev(p.x, p.y, cam) : p;
}

// Keypoint arrays: nsets sets of `cap` slots, counts[s] valid ones; every other field is copied (frame.cpp:784-790).
__global__ void __launch_bounds__(256) k_undistort_keypoints(const viorb_keypoint* __restrict__ kps, const int* __restrict__ counts,
                                                             int cap, CamModel cam, viorb_keypoint* __restrict__ out) {
    const int s = blockIdx.y;
    const int n = counts ? min(counts[s], cap) : cap;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    viorb_keypoint kp = kps[(size_t)s * cap + i];
    if (cam.mode == 2) {
        const float2 p = undistort_point_dev(kp.x, kp.y, cam);
        kp.x = p.x; kp.y = p.y;
    }
    out[(size_t)s * cap + i] = kp;
}
