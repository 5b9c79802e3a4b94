// Fisheye stereo epilogue: KannalaBrandt8::TriangulateMatches (src/geometry/cameramodels/kannalabrandt8.cpp:324-393)
// with unproject (:82-109), project (:11-27) and Triangulate (:485-498), and the loop of
// Frame::ComputeStereoFishEyeMatches that calls it (src/datastructures/frame.cpp:1172-1188).
// One thread per match.  The camera model is evaluated in float with the same operation order as the reference
// (CUDA's tanf / atan2f / sinf / cosf are within 2 ulp of glibc's); the 4x4 null vector that the reference takes from
// cv::SVD (float Jacobi) is computed with a one-sided Jacobi iteration in double, i.e. at least as accurately.
// Parity is therefore a tolerance: 1e-3 relative on depth and point (observed 6e-5), identical accept / reject decisions
// except where the tested quantity sits within that tolerance of its threshold (tests/test_kb8.py).
#pragma once
#include "viorb_common.cuh"

struct Kb8Rig {
    float prm1[8], prm2[8];   // fx fy cx cy k0 k1 k2 k3 (Camera::mvParameters)
    float R12[9], t12[3];     // mRlr (row-major), mtlr
    float precision;          // KannalaBrandt8::precision (1e-6)
};

__device__ __forceinline__ void kb8_unproject_dev(const float* prm, float px, float py, float precision, float r[3]) {
    const float pwx = __fdiv_rn(__fsub_rn(px, prm[2]), prm[0]), pwy = __fdiv_rn(__fsub_rn(py, prm[3]), prm[1]);
    float scale = 1.f;
    float theta_d = sqrtf(__fadd_rn(__fmul_rn(pwx, pwx), __fmul_rn(pwy, pwy)));
    theta_d = fminf(fmaxf(-1.57079632679489661923f, theta_d), 1.57079632679489661923f);
    if (theta_d > 1e-8f) {
        float theta = theta_d;
        for (int j = 0; j < 10; ++j) {
            const float t2 = __fmul_rn(theta, theta), t4 = __fmul_rn(t2, t2), t6 = __fmul_rn(t4, t2), t8 = __fmul_rn(t4, t4);
            const float a = __fmul_rn(prm[4], t2), b = __fmul_rn(prm[5], t4), c = __fmul_rn(prm[6], t6), d = __fmul_rn(prm[7], t8);
            const float num = __fsub_rn(__fmul_rn(theta, __fadd_rn(__fadd_rn(__fadd_rn(__fadd_rn(1.f, a), b), c), d)), theta_d);
            const float den = __fadd_rn(__fadd_rn(__fadd_rn(__fadd_rn(1.f, __fmul_rn(3.f, a)), __fmul_rn(5.f, b)), __fmul_rn(7.f, c)),
                                        __fmul_rn(9.f, d));
            const float fix = __fdiv_rn(num, den);
            theta = __fsub_rn(theta, fix);
            if (fabsf(fix) < precision) break;
        }
        scale = __fdiv_rn(tanf(theta), theta_d);
    }
    r[0] = __fmul_rn(pwx, scale); r[1] = __fmul_rn(pwy, scale); r[2] = 1.f;
}

__device__ __forceinline__ void kb8_project_dev(const float* prm, const float p[3], float uv[2]) {
    const float x2y2 = __fadd_rn(__fmul_rn(p[0], p[0]), __fmul_rn(p[1], p[1]));
    const float theta = atan2f(sqrtf(x2y2), p[2]);
    const float psi = atan2f(p[1], p[0]);
    const float t2 = __fmul_rn(theta, theta), t3 = __fmul_rn(theta, t2), t5 = __fmul_rn(t3, t2), t7 = __fmul_rn(t5, t2),
                t9 = __fmul_rn(t7, t2);
    const float r = __fadd_rn(__fadd_rn(__fadd_rn(__fadd_rn(theta, __fmul_rn(prm[4], t3)), __fmul_rn(prm[5], t5)), __fmul_rn(prm[6], t7)),
                              __fmul_rn(prm[7], t9));
    uv[0] = __fadd_rn(__fmul_rn(__fmul_rn(prm[0], r), cosf(psi)), prm[2]);
    uv[1] = __fadd_rn(__fmul_rn(__fmul_rn(prm[1], r), sinf(psi)), prm[3]);
}

// Right singular vector of the smallest singular value of the 4x4 matrix A (columns W[.][j]): rotate column pairs until
// orthogonal, accumulate the rotations in V, take the column of smallest norm.  Fully unrolled: registers only.
__device__ __forceinline__ void kb8_null_vector4(double W[4][4], double x[4]) {
    double V[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) V[i][j] = i == j ? 1.0 : 0.0;
#pragma unroll 1
    for (int sweep = 0; sweep < 12; ++sweep) {
        bool changed = false;
#pragma unroll
        for (int i = 0; i < 3; ++i) {
#pragma unroll
            for (int j = i + 1; j < 4; ++j) {
                double a = 0, b = 0, p = 0;
#pragma unroll
                for (int k = 0; k < 4; ++k) { a += W[k][i] * W[k][i]; b += W[k][j] * W[k][j]; p += W[k][i] * W[k][j]; }
                if (fabs(p) > 1e-15 * sqrt(a * b)) {
                    p *= 2;
                    const double beta = a - b, gamma = sqrt(p * p + beta * beta);
                    double c, s;
                    if (beta < 0) { const double delta = (gamma - beta) * 0.5; s = sqrt(delta / gamma); c = p / (gamma * s * 2); }
                    else { c = sqrt((gamma + beta) / (gamma * 2)); s = p / (gamma * c * 2); }
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const double w0 = c * W[k][i] + s * W[k][j], w1 = -s * W[k][i] + c * W[k][j];
                        W[k][i] = w0; W[k][j] = w1;
                        const double v0 = c * V[k][i] + s * V[k][j], v1 = -s * V[k][i] + c * V[k][j];
                        V[k][i] = v0; V[k][j] = v1;
                    }
                    changed = true;
                }
            }
        }
        if (!changed) break;
    }
    double bn = 1e300;
#pragma unroll
    for (int k = 0; k < 4; ++k) x[k] = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        double nrm = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) nrm += W[k][j] * W[k][j];
        if (nrm < bn) {
            bn = nrm;
#pragma unroll
            for (int k = 0; k < 4; ++k) x[k] = V[k][j];
        }
    }
}

// TriangulateMatches for one pair; returns its return value (-1 = rejected) and the point.
__device__ __forceinline__ float kb8_triangulate_match_dev(const Kb8Rig& g, float k1x, float k1y, float k2x, float k2y, float sigmaLevel,
                                                           float unc, float p3D[3]) {
    float r1[3], r2[3];
    kb8_unproject_dev(g.prm1, k1x, k1y, g.precision, r1);
    kb8_unproject_dev(g.prm2, k2x, k2y, g.precision, r2);
    double r21[3], dot = 0, n1 = 0, n2 = 0;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        r21[i] = (double)(float)((double)g.R12[3 * i] * r2[0] + (double)g.R12[3 * i + 1] * r2[1] + (double)g.R12[3 * i + 2] * r2[2]);
        dot += (double)r1[i] * r21[i]; n1 += (double)r1[i] * r1[i]; n2 += r21[i] * r21[i];
    }
    const float cosPar = (float)(dot / (sqrt(n1) * sqrt(n2)));
    p3D[0] = p3D[1] = p3D[2] = 0.f;
    if ((double)cosPar > 0.9998) return -1.f;
    float R21[9], t21[3];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) R21[3 * i + j] = g.R12[3 * j + i];
#pragma unroll
    for (int i = 0; i < 3; ++i)
        t21[i] = -(float)((double)R21[3 * i] * g.t12[0] + (double)R21[3 * i + 1] * g.t12[1] + (double)R21[3 * i + 2] * g.t12[2]);
    // A = [p1.x*T1.row(2) - T1.row(0); p1.y*T1.row(2) - T1.row(1); p2.x*T2.row(2) - T2.row(0); p2.y*T2.row(2) - T2.row(1)] (float)
    double W[4][4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const float t1r0 = j == 0 ? 1.f : 0.f, t1r1 = j == 1 ? 1.f : 0.f, t1r2 = j == 2 ? 1.f : 0.f;
        const float t2r0 = j < 3 ? R21[j] : t21[0], t2r1 = j < 3 ? R21[3 + j] : t21[1], t2r2 = j < 3 ? R21[6 + j] : t21[2];
        W[0][j] = (double)__fsub_rn(__fmul_rn(r1[0], t1r2), t1r0);
        W[1][j] = (double)__fsub_rn(__fmul_rn(r1[1], t1r2), t1r1);
        W[2][j] = (double)__fsub_rn(__fmul_rn(r2[0], t2r2), t2r0);
        W[3][j] = (double)__fsub_rn(__fmul_rn(r2[1], t2r2), t2r1);
    }
    double xh[4];
    kb8_null_vector4(W, xh);
    const float x3D[3] = {(float)(xh[0] / xh[3]), (float)(xh[1] / xh[3]), (float)(xh[2] / xh[3])};
    const float z1 = x3D[2];
    if (z1 <= 0.f) return -1.f;
    const float z2 = (float)(((double)R21[6] * x3D[0] + (double)R21[7] * x3D[1] + (double)R21[8] * x3D[2]) + (double)t21[2]);
    if (z2 <= 0.f) return -1.f;
    float uv[2];
    kb8_project_dev(g.prm1, x3D, uv);
    const float e1x = __fsub_rn(uv[0], k1x), e1y = __fsub_rn(uv[1], k1y);
    if ((double)__fadd_rn(__fmul_rn(e1x, e1x), __fmul_rn(e1y, e1y)) > 5.991 * (double)sigmaLevel) return -1.f;
    float x2[3];
#pragma unroll
    for (int i = 0; i < 3; ++i)
        x2[i] = __fadd_rn((float)((double)R21[3 * i] * x3D[0] + (double)R21[3 * i + 1] * x3D[1] + (double)R21[3 * i + 2] * x3D[2]), t21[i]);
    kb8_project_dev(g.prm2, x2, uv);
    const float e2x = __fsub_rn(uv[0], k2x), e2y = __fsub_rn(uv[1], k2y);
    if ((double)__fadd_rn(__fmul_rn(e2x, e2x), __fmul_rn(e2y, e2y)) > 5.991 * (double)unc) return -1.f;
    p3D[0] = x3D[0]; p3D[1] = x3D[1]; p3D[2] = x3D[2];
    return z1;
}

// Bare TriangulateMatches over n pairs.  pt1 / pt2 [n] (kp.pt), sigma1 / sigma2 [n] -> depth [n], p3d [n][3].
__global__ void __launch_bounds__(128) k_kb8_triangulate(Kb8Rig g, const float2* __restrict__ pt1, const float2* __restrict__ pt2,
                                                         const float* __restrict__ sigma1, const float* __restrict__ sigma2, int n,
                                                         float* __restrict__ depth, float* __restrict__ p3d) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float p[3];
    const float2 a = pt1[i], b = pt2[i];
    depth[i] = kb8_triangulate_match_dev(g, a.x, a.y, b.x, b.y, sigma1[i], sigma2[i], p);
    p3d[3 * i] = p[0]; p3d[3 * i + 1] = p[1]; p3d[3 * i + 2] = p[2];
}

// The match loop of Frame::ComputeStereoFishEyeMatches (frame.cpp:1172-1188) after knnMatch: query row q is left keypoint
// q + monoLeft, train row t is right keypoint t + monoRight.  knnIdx / knnDist [nq][2] from the kNN-2 kernel (-1 = no
// such neighbour).  Outputs are pre-set by the caller: l2r / r2l = -1, depth = -1, p3d = 0.  mvRightToLeftMatch keeps the
// LAST accepted left index in the reference's sequential loop = the largest one: atomicMax.
__global__ void __launch_bounds__(128) k_fisheye_epilogue(Kb8Rig g, const viorb_keypoint* __restrict__ kpsL, int monoLeft,
                                                          const viorb_keypoint* __restrict__ kpsR, int monoRight, int nq,
                                                          const int32_t* __restrict__ knnIdx, const int32_t* __restrict__ knnDist,
                                                          const float* __restrict__ levelSigma2, int32_t* __restrict__ l2r,
                                                          int32_t* __restrict__ r2l, float* __restrict__ depth, float* __restrict__ p3d,
                                                          int* __restrict__ counters) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= nq) return;
    const int t0 = knnIdx[2 * q], t1 = knnIdx[2 * q + 1];
    if (t0 < 0 || t1 < 0) return;                                                  // (*it).size() >= 2
    if (!((double)(float)knnDist[2 * q] < (double)(float)knnDist[2 * q + 1] * 0.7)) return;
    atomicAdd(&counters[0], 1);                                                    // descMatches
    const viorb_keypoint a = kpsL[q + monoLeft], b = kpsR[t0 + monoRight];
    float p[3];
    const float z = kb8_triangulate_match_dev(g, a.x, a.y, b.x, b.y, levelSigma2[a.octave], levelSigma2[b.octave], p);
    if (z > 0.0001f) {
        const int il = q + monoLeft, ir = t0 + monoRight;
        l2r[il] = ir;
        atomicMax(&r2l[ir], il);
        depth[il] = z;
        p3d[3 * il] = p[0]; p3d[3 * il + 1] = p[1]; p3d[3 * il + 2] = p[2];
        atomicAdd(&counters[1], 1);                                                // nMatches
    }
}
