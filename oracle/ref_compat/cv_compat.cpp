// cv_compat.cpp — the OpenCV primitives behind cv_compat.hpp.
// TEST INFRASTRUCTURE ONLY.  The arithmetic is NOT restated here: every primitive forwards to the
// cv2-pinned restatement in oracle/orb_oracle.cpp (orc_fast, orc_resize_linear, orc_gaussian7,
// orc_fast_atan2; pinned by tests/test_oracle_golden.py against cv2 4.13 outputs).  copyMakeBorder is the
// only function written out (it is pure indexing; OpenCV modules/core/src/copy.cpp behaviour).
#include "cv_compat.hpp"

#include <algorithm>
#include <cstdint>
#include <stdexcept>

extern "C" {
void orc_resize_linear(const uint8_t* s, int sw, int sh, int sstep, uint8_t* d, int dw, int dh, int dstep);
void orc_gaussian7(const uint8_t* s, int w, int h, int sstep, uint8_t* d, int dstep);
int orc_fast(const uint8_t* img, int w, int h, int step, int th, int32_t* xys, int cap);
float orc_fast_atan2(float y, float x);
}

namespace cv_compat {
static thread_local std::vector<FastCall> g_trace;
static thread_local bool g_trace_on = false;
std::vector<FastCall>& fast_trace() { return g_trace; }
void fast_trace_enable(bool on) { g_trace_on = on; }
}

namespace cv {

// cv::FAST(image, keypoints, threshold, nonmaxSuppression=true), TYPE_9_16.  KeyPoint(x, y, 7.f, -1, score).
void FAST(InputArray image, std::vector<KeyPoint>& keypoints, int threshold, bool nonmaxSuppression) {
    if (!nonmaxSuppression) throw std::runtime_error("cv_compat: FAST without NMS is not on the reference path");
    Mat img = image.getMat();
    keypoints.clear();
    if (!img.empty()) {
        std::vector<int32_t> xys((size_t)img.rows * img.cols * 3 + 3);
        int n = orc_fast(img.data, img.cols, img.rows, (int)(size_t)img.step, threshold, xys.data(), img.rows * img.cols);
        keypoints.reserve(n);
        for (int i = 0; i < n; ++i)
            keypoints.push_back(KeyPoint((float)xys[3 * i], (float)xys[3 * i + 1], 7.f, -1.f, (float)xys[3 * i + 2]));
    }
    if (cv_compat::g_trace_on)
        cv_compat::g_trace.push_back({img.data, (size_t)img.step, img.cols, img.rows, threshold, keypoints});
}

void resize(InputArray src_, OutputArray dst_, Size dsize, double fx, double fy, int interpolation) {
    if (interpolation != INTER_LINEAR || fx != 0 || fy != 0)
        throw std::runtime_error("cv_compat: only resize(dsize, INTER_LINEAR) is on the reference path");
    Mat src = src_.getMat();
    dst_.create(dsize, CV_8UC1);
    Mat dst = dst_.getMat();
    orc_resize_linear(src.data, src.cols, src.rows, (int)(size_t)src.step, dst.data, dst.cols, dst.rows, (int)(size_t)dst.step);
}

// OpenCV borderInterpolate (modules/core/src/copy.cpp) for the reflecting / replicating types
int borderInterpolate(int p, int len, int borderType) {
    if ((unsigned)p < (unsigned)len) return p;
    if (borderType == BORDER_REPLICATE) return p < 0 ? 0 : len - 1;
    if (borderType == BORDER_REFLECT || borderType == BORDER_REFLECT_101) {
        int delta = borderType == BORDER_REFLECT_101;
        if (len == 1) return 0;
        do {
            if (p < 0) p = -p - 1 + delta;
            else p = len - 1 - (p - len) - delta;
        } while ((unsigned)p >= (unsigned)len);
        return p;
    }
    if (borderType == BORDER_WRAP) {
        if (p < 0) p -= ((p - len + 1) / len) * len;
        if (p >= len) p %= len;
        return p;
    }
    throw std::runtime_error("cv_compat: border type not supported");
}

void copyMakeBorder(InputArray src_, OutputArray dst_, int top, int bottom, int left, int right, int borderType) {
    Mat src = src_.getMat();
    if (src.isSubmatrix() && (borderType & BORDER_ISOLATED) == 0) {     // take real pixels around the ROI first
        Size ws; Point ofs;
        src.locateROI(ws, ofs);
        int dtop = std::min(ofs.y, top), dbottom = std::min(ws.height - src.rows - ofs.y, bottom);
        int dleft = std::min(ofs.x, left), dright = std::min(ws.width - src.cols - ofs.x, right);
        src.adjustROI(dtop, dbottom, dleft, dright);
        top -= dtop; left -= dleft; bottom -= dbottom; right -= dright;
    }
    dst_.create(src.rows + top + bottom, src.cols + left + right, CV_8UC1);
    Mat dst = dst_.getMat();
    if (top == 0 && left == 0 && bottom == 0 && right == 0) {
        if (src.data != dst.data || (size_t)src.step != (size_t)dst.step) src.copyTo(dst);
        return;
    }
    borderType &= ~BORDER_ISOLATED;
    std::vector<int> tab(left + right);
    for (int i = 0; i < left; ++i) tab[i] = borderInterpolate(i - left, src.cols, borderType);
    for (int i = 0; i < right; ++i) tab[left + i] = borderInterpolate(src.cols + i, src.cols, borderType);
    for (int y = 0; y < src.rows; ++y) {
        const uchar* s = src.ptr(y);
        uchar* d = dst.ptr(y + top);
        if (d + left != s) std::memmove(d + left, s, (size_t)src.cols);
        for (int i = 0; i < left; ++i) d[i] = s[tab[i]];
        for (int i = 0; i < right; ++i) d[left + src.cols + i] = s[tab[left + i]];
    }
    for (int i = 0; i < top; ++i)
        std::memcpy(dst.ptr(i), dst.ptr(top + borderInterpolate(i - top, src.rows, borderType)), (size_t)dst.cols);
    for (int i = 0; i < bottom; ++i)
        std::memcpy(dst.ptr(top + src.rows + i), dst.ptr(top + borderInterpolate(src.rows + i, src.rows, borderType)), (size_t)dst.cols);
}

void GaussianBlur(InputArray src_, OutputArray dst_, Size ksize, double sigmaX, double sigmaY, int borderType) {
    if (ksize.width != 7 || ksize.height != 7 || sigmaX != 2 || sigmaY != 2 || (borderType & ~BORDER_ISOLATED) != BORDER_REFLECT_101)
        throw std::runtime_error("cv_compat: only GaussianBlur(7x7, 2, 2, BORDER_REFLECT_101) is on the reference path");
    Mat src = src_.getMat().clone();          // the reference blurs in place
    dst_.create(src.rows, src.cols, CV_8UC1);
    Mat dst = dst_.getMat();
    orc_gaussian7(src.data, src.cols, src.rows, (int)(size_t)src.step, dst.data, (int)(size_t)dst.step);
}

float fastAtan2(float y, float x) { return orc_fast_atan2(y, x); }

void KeyPointsFilter::retainBest(std::vector<KeyPoint>& keypoints, int n_points) {
    if (n_points < 0 || (int)keypoints.size() <= n_points) return;
    if (n_points == 0) { keypoints.clear(); return; }
    std::nth_element(keypoints.begin(), keypoints.begin() + n_points - 1, keypoints.end(),
                     [](const KeyPoint& a, const KeyPoint& b) { return a.response > b.response; });
    float ambiguous = keypoints[n_points - 1].response;
    auto new_end = std::partition(keypoints.begin() + n_points, keypoints.end(),
                                  [ambiguous](const KeyPoint& k) { return k.response >= ambiguous; });
    keypoints.resize(new_end - keypoints.begin());
}

}  // namespace cv
