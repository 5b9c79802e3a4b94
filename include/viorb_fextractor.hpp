// viorb_fextractor.hpp — header-only C++ shim with the reference's extractor interface
// (reference include/vi_slam/geometry/fextractor.h:26-64) forwarding to the C ABI in viorb.h.
//
// In the reference tree: replace the body of src/geometry/fextractor.cpp by
//     #include "viorb_fextractor.hpp"
// with VIORB_WITH_OPENCV defined, link libviorb.so, and Frame / Tracking compile unchanged
// (see INTEGRATION.md).  Without OpenCV (this repository's own builds) the same class is
// available over the tiny POD stand-ins below, which is what tests/test_cpp_shim.py compiles.
#pragma once
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "viorb.h"

#ifdef VIORB_WITH_OPENCV
#include <opencv2/core.hpp>
namespace viorb_cv = cv;
#else
namespace viorb_cv {
struct Point2f { float x, y; };
struct KeyPoint {                 // cv::KeyPoint layout (28 bytes)
    Point2f pt; float size, angle, response; int octave, class_id;
};
struct Mat {                      // minimal CV_8UC1 view / owner
    int rows = 0, cols = 0; size_t step = 0; uint8_t* data = nullptr; std::vector<uint8_t> own;
    Mat() {}
    Mat(int r, int c, uint8_t* d, size_t s) : rows(r), cols(c), step(s), data(d) {}
    void create(int r, int c) { rows = r; cols = c; step = (size_t)c; own.assign((size_t)r * c, 0); data = own.data(); }
    void release() { rows = cols = 0; step = 0; own.clear(); data = nullptr; }
    bool empty() const { return rows == 0 || cols == 0 || data == nullptr; }
};
}  // namespace viorb_cv
#endif

namespace vi_slam {
namespace geometry {

static_assert(sizeof(viorb_cv::KeyPoint) == sizeof(viorb_keypoint), "cv::KeyPoint must be the 28-byte POD");

class FExtractor {
public:
    enum { HARRIS_SCORE = 0, FAST_SCORE = 1 };

    // fextractor.cpp:401-461.  max_width / max_height size the device arenas (the reference allocates per call).
    FExtractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST,
               int max_width = 1280, int max_height = 720, int device = 0)
        : nlevels_(nlevels), scaleFactor_(scaleFactor) {
        viorb_params p{nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, max_width, max_height, 1, device};
        if (viorb_extractor_create(&p, &h_) != VIORB_OK) throw std::runtime_error(viorb_last_error());
        mvScaleFactor.resize(nlevels); mvInvScaleFactor.resize(nlevels);
        mvLevelSigma2.resize(nlevels); mvInvLevelSigma2.resize(nlevels);
        viorb_extractor_tables(h_, nullptr, mvScaleFactor.data(), mvInvScaleFactor.data(), mvLevelSigma2.data(),
                               mvInvLevelSigma2.data(), nullptr);
        mvImagePyramid.resize(nlevels);
        cap_ = viorb_extractor_capacity(h_);
    }
    ~FExtractor() { viorb_extractor_destroy(h_); }
    FExtractor(const FExtractor&) = delete;
    FExtractor& operator=(const FExtractor&) = delete;

    // fextractor.cpp:1034-1133.  `mask` is ignored like in the reference.  Returns monoIndex, -1 on an empty image.
#ifdef VIORB_WITH_OPENCV
    int compute(cv::InputArray _image, cv::InputArray /*mask*/, std::vector<cv::KeyPoint>& keypoints,
                cv::OutputArray _descriptors, std::vector<int>& vLappingArea) {
        if (_image.empty()) return -1;
        cv::Mat image = _image.getMat();
        CV_Assert(image.type() == CV_8UC1);                       // fextractor.cpp:1041
        int n = 0;
        const int mono = run(image.data, image.cols, image.rows, (int)image.step, vLappingArea, keypoints, n);
        if (n == 0) _descriptors.release();
        else {
            _descriptors.create(n, 32, CV_8U);
            cv::Mat d = _descriptors.getMat();
            for (int i = 0; i < n; ++i) std::memcpy(d.ptr(i), &desc_[(size_t)i * 32], 32);
        }
        fill_pyramid();
        return mono;
    }
#else
    int compute(const viorb_cv::Mat& image, const viorb_cv::Mat& /*mask*/, std::vector<viorb_cv::KeyPoint>& keypoints,
                viorb_cv::Mat& descriptors, std::vector<int>& vLappingArea) {
        if (image.empty()) return -1;
        int n = 0;
        const int mono = run(image.data, image.cols, image.rows, (int)image.step, vLappingArea, keypoints, n);
        if (n == 0) descriptors.release();
        else {
            descriptors.create(n, 32);
            std::memcpy(descriptors.data, desc_.data(), (size_t)n * 32);
        }
        fill_pyramid();
        return mono;
    }
#endif

    int GetLevels() { return nlevels_; }
    float GetScaleFactor() { return scaleFactor_; }
    std::vector<float> GetScaleFactors() { return mvScaleFactor; }
    std::vector<float> GetInverseScaleFactors() { return mvInvScaleFactor; }
    std::vector<float> GetScaleSigmaSquares() { return mvLevelSigma2; }
    std::vector<float> GetInverseScaleSigmaSquares() { return mvInvLevelSigma2; }
    viorb_extractor* handle() { return h_; }
    // When ComputeStereoMatches is rerouted to viorb_stereo_match nothing on the host reads mvImagePyramid:
    // switch the per-frame device->host copy of the 8 levels off.
    void SetPyramidReadback(bool on) { readback_ = on; }
    // Frame::UndistortKeyPoints (frame.cpp:758-791) fused into compute(): give the extractor the Frame's mK
    // (fx fy cx cy) and mDistCoef (4 or 5 floats) once; after every compute() UndistortedKeyPoints() returns
    // ukeypoints_ of that frame.  dist == nullptr switches it off.
    void SetCamera(const float* cam, const float* dist, int ndist) {
        if (viorb_extractor_set_camera(h_, cam, dist, ndist) != VIORB_OK) throw std::runtime_error(viorb_last_error());
    }
    void UndistortedKeyPoints(std::vector<viorb_cv::KeyPoint>& ukeypoints) {
        int n = 0;
        kp_.resize(cap_);
        if (viorb_fetch_undistorted(h_, 0, kp_.data(), cap_, &n) != VIORB_OK) throw std::runtime_error(viorb_last_error());
        ukeypoints.resize(n);
        if (n) std::memcpy((void*)ukeypoints.data(), kp_.data(), sizeof(viorb_keypoint) * (size_t)n);
    }

    std::vector<viorb_cv::Mat> mvImagePyramid;

protected:
    int run(const uint8_t* data, int cols, int rows, int step, const std::vector<int>& lap,
            std::vector<viorb_cv::KeyPoint>& keypoints, int& n) {
        kp_.resize(cap_); desc_.resize((size_t)cap_ * 32);
        int mono = -1;
        const int rc = viorb_extract(h_, data, cols, rows, step, lap.at(0), lap.at(1), kp_.data(), desc_.data(), cap_, &n, &mono);
        if (rc == VIORB_E_EMPTY) { n = 0; keypoints.clear(); return -1; }
        if (rc != VIORB_OK) throw std::runtime_error(viorb_last_error());
        keypoints.resize(n);
        if (n) std::memcpy((void*)keypoints.data(), kp_.data(), sizeof(viorb_keypoint) * (size_t)n);
        return mono;
    }
    // mvImagePyramid stays valid until the next compute (read by Frame::ComputeStereoMatches, frame.cpp:830,920,937)
    void fill_pyramid() {
        if (!readback_) return;
        for (int l = 0; l < nlevels_; ++l) {
            int w = 0, hgt = 0;
            viorb_pyramid_level(h_, 0, l, nullptr, 0, &w, &hgt);
#ifdef VIORB_WITH_OPENCV
            mvImagePyramid[l].create(hgt, w, CV_8U);
#else
            mvImagePyramid[l].create(hgt, w);
#endif
            viorb_pyramid_level(h_, 0, l, mvImagePyramid[l].data, (int)mvImagePyramid[l].step, &w, &hgt);
        }
    }
    viorb_extractor* h_ = nullptr;
    int nlevels_, cap_ = 0;
    bool readback_ = true;
    float scaleFactor_;
    std::vector<float> mvScaleFactor, mvInvScaleFactor, mvLevelSigma2, mvInvLevelSigma2;
    std::vector<viorb_keypoint> kp_;
    std::vector<uint8_t> desc_;
};

// FMatcher::DescriptorDistance keeps its scalar host form in the reference's windowed matchers (it is called
// per candidate pair inside host loops, fmatcher.cpp:2859); bulk distance / kNN work goes through viorb_knn2,
// viorb_match_ratio and viorb_stereo_match.

}  // namespace geometry
}  // namespace vi_slam
