// dropin_driver.cpp — ONE source, compiled twice by tests/test_cpp_shim.py:
//   (a) against the reference's own headers and object code:  -I/root/reference/include -Ioracle/ref_compat, linked with
//       oracle/_ref/liborb_ref.so (vi_slam::geometry::FExtractor as compiled from the reference's fextractor.cpp);
//   (b) with -DUSE_VIORB -DVIORB_WITH_OPENCV against include/viorb_fextractor.hpp + viorb_fmatcher.hpp, linked with
//       libviorb (the emulation build on CPU, the CUDA build on the GPU box).
// It uses the class exactly as the reference's callers do: `new FExtractor(...)` as Tracking::ParseORBParamFile
// (tracking.cpp:1087-1093), compute(im, cv::Mat(), keypoints, descriptors, vLapping) as Frame::ExtractORB
// (frame.cpp:416-426), the six getters as the Frame ctor (frame.cpp:94-100), mvImagePyramid as ComputeStereoMatches
// (frame.cpp:830).  Both builds must write byte-identical dumps.
// usage: dropin_driver <in.raw> <w> <h> <nfeatures> <lap0> <lap1> <out.bin>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#ifdef USE_VIORB
#include "viorb_fmatcher.hpp"
#else
#include "vi_slam/geometry/fextractor.h"
// DistributeOctTree breaks size ties by node ADDRESS (fextractor.cpp:675): serve the reference's allocations in
// allocation order (oracle/ref_compat/ref_driver.cpp), the documented tie rule of the oracle and of the CUDA path
extern "C" void orb_ref_monotone(int on);
#endif

using vi_slam::geometry::FExtractor;

int main(int argc, char** argv) {
    if (argc != 8) return 2;
    const int w = std::atoi(argv[2]), h = std::atoi(argv[3]), nf = std::atoi(argv[4]);
    std::vector<unsigned char> buf((size_t)w * h);
    FILE* f = std::fopen(argv[1], "rb");
    if (!f || std::fread(buf.data(), 1, buf.size(), f) != buf.size()) return 3;
    std::fclose(f);
#ifndef USE_VIORB
    orb_ref_monotone(1);
#endif
#ifdef USE_VIORB
    FExtractor* mpORBextractorLeft = new FExtractor(nf, 1.2f, 8, 20, 7, w, h);
#else
    FExtractor* mpORBextractorLeft = new FExtractor(nf, 1.2f, 8, 20, 7);
#endif
    const cv::Mat im(h, w, CV_8UC1, buf.data(), (size_t)w);
    std::vector<cv::KeyPoint> keypoints_;
    cv::Mat descriptors_;
    std::vector<int> vLapping = {std::atoi(argv[5]), std::atoi(argv[6])};
    const int monoLeft = mpORBextractorLeft->compute(im, cv::Mat(), keypoints_, descriptors_, vLapping);
    const int mnScaleLevels = mpORBextractorLeft->GetLevels();
    const float mfScaleFactor = mpORBextractorLeft->GetScaleFactor();
    std::vector<float> mvScaleFactors = mpORBextractorLeft->GetScaleFactors(), mvInvScaleFactors = mpORBextractorLeft->GetInverseScaleFactors(),
                       mvLevelSigma2 = mpORBextractorLeft->GetScaleSigmaSquares(), mvInvLevelSigma2 = mpORBextractorLeft->GetInverseScaleSigmaSquares();
    FILE* o = std::fopen(argv[7], "wb");
    const int n = (int)keypoints_.size();
    std::fwrite(&monoLeft, 4, 1, o); std::fwrite(&n, 4, 1, o); std::fwrite(&mnScaleLevels, 4, 1, o); std::fwrite(&mfScaleFactor, 4, 1, o);
    for (auto* v : {&mvScaleFactors, &mvInvScaleFactors, &mvLevelSigma2, &mvInvLevelSigma2}) std::fwrite(v->data(), 4, v->size(), o);
    if (n) {
        std::fwrite(keypoints_.data(), sizeof(cv::KeyPoint), n, o);
        for (int i = 0; i < n; ++i) std::fwrite(descriptors_.ptr(i), 1, 32, o);
    }
    for (int l = 0; l < mnScaleLevels; ++l) {
        const cv::Mat& m = mpORBextractorLeft->mvImagePyramid[l];
        std::fwrite(&m.cols, 4, 1, o); std::fwrite(&m.rows, 4, 1, o);
        for (int y = 0; y < m.rows; ++y) std::fwrite(m.ptr(y), 1, (size_t)m.cols, o);
    }
#ifdef USE_VIORB
    // the matcher shims with the reference's signatures (fmatcher.h:34-42, :77; frame.cpp:1167)
    if (n >= 2) {
        using namespace vi_slam::geometry;
        if (FMatcher::DescriptorDistance(descriptors_.row(0), descriptors_.row(0)) != 0) return 5;
        std::vector<std::vector<cv::DMatch>> knn;
        knnMatchHamming2(descriptors_, descriptors_, knn);
        if ((int)knn.size() != n || knn[0].size() != 2 || knn[0][0].distance != 0.f) return 6;
        if ((int)knn[1][0].distance != FMatcher::DescriptorDistance(descriptors_.row(1), descriptors_.row(knn[1][0].trainIdx))) return 7;
        std::vector<cv::DMatch> matches;
        matchFeatures(cv::Mat1b(descriptors_), cv::Mat1b(descriptors_), matches, 2);
        for (size_t i = 1; i < matches.size(); ++i) if (matches[i].trainIdx <= matches[i - 1].trainIdx) return 8;     // sorted, unique
        bool threw = false;
        try { matchFeatures(cv::Mat1b(descriptors_), cv::Mat1b(descriptors_), matches, 7); } catch (const std::runtime_error&) { threw = true; }
        if (!threw) return 9;                                                                                          // fmatcher.cpp:231
        const int nm = (int)matches.size();
        std::fwrite(&nm, 4, 1, o);
    }
#endif
    std::fclose(o);
    return 0;
}
