// viorb_fmatcher.hpp — header-only C++ shims with the reference's MATCHER signatures, forwarding to the C ABI (viorb.h):
//
//   static int FMatcher::DescriptorDistance(const cv::Mat&, const cv::Mat&)      include/vi_slam/geometry/fmatcher.h:77,
//                                                                                src/geometry/fmatcher.cpp:2859-2875
//   void matchFeatures(const cv::Mat1b&, const cv::Mat1b&, vector<cv::DMatch>&, int method_index, bool is_print_res, ...)
//                                                                                fmatcher.h:34-42, fmatcher.cpp:136-245 (method 2)
//   cv::BFMatcher(NORM_HAMMING).knnMatch(q, t, matches, 2)                       src/datastructures/frame.cpp:1167, fmatcher.cpp:214
//
// In the reference tree define VIORB_WITH_OPENCV (the real cv:: types are used); without it (this repository's own
// builds) the same code runs over the small POD stand-ins of viorb_fextractor.hpp plus DMatch below.
// See INTEGRATION.md sections 4 and 6.
#pragma once
#include <algorithm>
#include <cstdio>
#include <stdexcept>
#include <vector>

#include "viorb_fextractor.hpp"

#ifndef VIORB_WITH_OPENCV
namespace viorb_cv {
struct DMatch {                       // cv::DMatch
    int queryIdx = -1, trainIdx = -1, imgIdx = -1;
    float distance = 3.402823466e+38f;
    DMatch() {}
    DMatch(int q, int t, float d) : queryIdx(q), trainIdx(t), imgIdx(-1), distance(d) {}
};
typedef Mat Mat1b;
}  // namespace viorb_cv
#endif

namespace vi_slam {
namespace geometry {

// lowe_method_dist_ratio of matchFeatures method 2.  The reference reads it once with Config::get<int>
// (fmatcher.cpp:144), which truncates the shipped 0.8 to... whatever the yaml holds as an int; the host sets the value
// it actually runs with here before the first call.
inline double& viorb_lowe_method_dist_ratio() { static double r = 0.8; return r; }
inline int& viorb_matcher_device() { static int d = 0; return d; }

namespace viorb_detail {
inline const uint8_t* rows_u8(const viorb_cv::Mat& m, std::vector<uint8_t>& pack) {
    // descriptors are N x 32 CV_8U; a continuous matrix is passed as it is, a strided one is packed
    if (m.rows == 0) return nullptr;
    if ((size_t)m.step == (size_t)m.cols) return m.data;
    pack.resize((size_t)m.rows * m.cols);
    for (int y = 0; y < m.rows; ++y) std::memcpy(&pack[(size_t)y * m.cols], m.data + (size_t)y * m.step, (size_t)m.cols);
    return pack.data();
}
}  // namespace viorb_detail

class FMatcher {
public:
    // One pair of 256-bit rows: a scalar host loop, like the reference (it is called per candidate inside host loops).
    // Bulk work goes through DescriptorDistances / viorb_knn2 / viorb_match_ratio / the windowed matchers.
    static int DescriptorDistance(const viorb_cv::Mat& a, const viorb_cv::Mat& b) {
        uint32_t wa[8], wb[8];
        std::memcpy(wa, a.data, 32);
        std::memcpy(wb, b.data, 32);
        int dist = 0;
        for (int i = 0; i < 8; ++i) dist += __builtin_popcount(wa[i] ^ wb[i]);
        return dist;
    }
    // n row pairs at once on the device (k_desc_dist): dist[i] = DescriptorDistance(A.row(i), B.row(i)).
    static void DescriptorDistances(const viorb_cv::Mat& A, const viorb_cv::Mat& B, std::vector<int>& dist) {
        if (A.rows != B.rows || A.cols != 32 || B.cols != 32) throw std::runtime_error("DescriptorDistances: N x 32 matrices of equal height");
        std::vector<uint8_t> pa, pb;
        dist.assign((size_t)A.rows, 0);
        if (A.rows == 0) return;
        if (viorb_descriptor_distance(viorb_detail::rows_u8(A, pa), viorb_detail::rows_u8(B, pb), A.rows, dist.data(), viorb_matcher_device()) != VIORB_OK)
            throw std::runtime_error(viorb_last_error());
    }
};

// cv::BFMatcher(cv::NORM_HAMMING).knnMatch(queryDescriptors, trainDescriptors, matches, 2) (frame.cpp:1167): per query the
// two nearest train rows, ascending distance, ties -> lower trainIdx; fewer entries where the train set is smaller than k.
inline void knnMatchHamming2(const viorb_cv::Mat& query, const viorb_cv::Mat& train, std::vector<std::vector<viorb_cv::DMatch>>& matches) {
    std::vector<uint8_t> pq, pt;
    const int nq = query.rows, nt = train.rows;
    matches.assign((size_t)nq, std::vector<viorb_cv::DMatch>());
    if (nq == 0) return;
    std::vector<int32_t> idx((size_t)nq * 2), dist((size_t)nq * 2);
    if (viorb_knn2(viorb_detail::rows_u8(query, pq), nq, viorb_detail::rows_u8(train, pt), nt, idx.data(), dist.data(), viorb_matcher_device()) != VIORB_OK)
        throw std::runtime_error(viorb_last_error());
    for (int i = 0; i < nq; ++i)
        for (int k = 0; k < 2; ++k)
            if (idx[2 * i + k] >= 0) matches[i].push_back(viorb_cv::DMatch(i, idx[2 * i + k], (float)dist[2 * i + k]));
}

// matchFeatures (fmatcher.h:34-42).  Method 2 (Lowe's ratio on brute-force kNN-2, then sort by trainIdx and
// removeDuplicatedMatches, fmatcher.cpp:204-229, 247-266) runs on the device.  Methods 1 and 3 stay with the host
// implementation (FLANN-LSH / radius search: approximate and randomised, used by disabled tests only); any other index
// throws like the reference does (fmatcher.cpp:231).
inline void matchFeatures(const viorb_cv::Mat1b& descriptors_1, const viorb_cv::Mat1b& descriptors_2, std::vector<viorb_cv::DMatch>& matches,
                          int method_index = 1, bool is_print_res = false,
                          const std::vector<viorb_cv::KeyPoint>& = std::vector<viorb_cv::KeyPoint>(),
                          const std::vector<viorb_cv::KeyPoint>& = std::vector<viorb_cv::KeyPoint>(), float = 0.0f) {
    if (method_index != 2) {
        if (method_index == 1 || method_index == 3)
            throw std::runtime_error("viorb matchFeatures: methods 1 and 3 are not rerouted (keep the host implementation); use method 2");
        throw std::runtime_error("feature_match.cpp::matchFeatures: wrong method index.");
    }
    std::vector<uint8_t> p1, p2;
    const int n1 = descriptors_1.rows, n2 = descriptors_2.rows;
    matches.clear();
    if (n1 == 0) return;
    std::vector<viorb_match> out((size_t)n1);
    int n = 0;
    if (viorb_match_ratio(viorb_detail::rows_u8(descriptors_1, p1), n1, viorb_detail::rows_u8(descriptors_2, p2), n2,
                          viorb_lowe_method_dist_ratio(), /* rule */ 1, /* dedup */ 1, out.data(), n1, &n, viorb_matcher_device()) != VIORB_OK)
        throw std::runtime_error(viorb_last_error());
    matches.reserve((size_t)n);
    for (int i = 0; i < n; ++i) matches.push_back(viorb_cv::DMatch(out[i].query_idx, out[i].train_idx, (float)out[i].distance));
    if (is_print_res) std::printf("Matching features:\nUsing method %d\nNumber of matches: %d\n", method_index, n);
}

}  // namespace geometry
}  // namespace vi_slam
