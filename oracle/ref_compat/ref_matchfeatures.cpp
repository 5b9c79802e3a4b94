// ref_matchfeatures.cpp — the reference's OWN geometry::matchFeatures (method 2: kNN-2 + Lowe ratio) and
// removeDuplicatedMatches (row a-9), compiled from where they lie.  TEST INFRASTRUCTURE ONLY (part of
// oracle/_ref/liborb_ref.so, built by oracle/ref.mk; see ref_slices.cpp for the method).
//     geometry/fmatcher.cpp:132-245   matchFeatures (whole function; methods 1 / 3 compile against throwing stand-ins)
//     geometry/fmatcher.cpp:247-266   removeDuplicatedMatches
// Stand-ins: cv::DescriptorMatcher::create("BruteForce-Hamming")->knnMatch = the oracle's kNN-2 (pinned against cv2),
// basics::Config::get<int> = values set by the test (the reference reads its ratios with get<int>: OpenCV's FileNode -> int
// conversion rounds, so a configured 0.8 arrives as 1).
#include "cv_compat.hpp"

#include <cstdint>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

extern "C" void orc_knn2(const uint8_t* q, int nq, const uint8_t* t, int nt, int32_t* idx, int32_t* dist, int nthreads);

namespace refmf {
namespace cv {
using namespace ::cv;
template <typename T> struct Ptr {
    std::shared_ptr<T> p;
    Ptr() {}
    Ptr(T* raw) : p(raw) {}
    T* operator->() const { return p.get(); }
};
class DescriptorMatcher {
public:
    static Ptr<DescriptorMatcher> create(const std::string& name) {
        if (name != "BruteForce-Hamming") throw std::runtime_error("stand-in: only BruteForce-Hamming");
        return Ptr<DescriptorMatcher>(new DescriptorMatcher());
    }
    void add(const std::vector<Mat>&) {}
    void train() {}
    void knnMatch(const Mat& q, const Mat& t, std::vector<std::vector<DMatch>>& matches, int k) const {
        if (k != 2) throw std::runtime_error("stand-in: k = 2 only");
        matches.assign(q.rows, std::vector<DMatch>());
        if (q.rows == 0 || t.rows == 0) return;
        Mat qc = q.isContinuous() ? q : q.clone(), tc = t.isContinuous() ? t : t.clone();
        std::vector<int32_t> idx(2 * (size_t)q.rows), dist(2 * (size_t)q.rows);
        orc_knn2(qc.ptr(0), q.rows, tc.ptr(0), t.rows, idx.data(), dist.data(), 1);
        for (int i = 0; i < q.rows; ++i)
            for (int j = 0; j < 2; ++j)
                if (idx[2 * i + j] >= 0) matches[i].push_back(DMatch(i, idx[2 * i + j], (float)dist[2 * i + j]));
    }
};
namespace flann { struct LshIndexParams { LshIndexParams(int, int, int) {} }; }
class FlannBasedMatcher {
public:
    explicit FlannBasedMatcher(flann::LshIndexParams* p) { delete p; }
    void match(const Mat&, const Mat&, std::vector<DMatch>&) { throw std::runtime_error("stand-in: FLANN-LSH is out of scope (SURVEY section 2)"); }
};
}  // namespace cv

namespace vi_slam {
namespace basics {
struct Config {
    static std::map<std::string, double>& values() { static std::map<std::string, double> v; return v; }
    template <typename T> static T get(const std::string& key) { return (T)values()[key]; }
};
}  // namespace basics
namespace geometry {
using namespace std;
vector<cv::DMatch> matchByRadiusAndBruteForce(const vector<cv::KeyPoint>&, const vector<cv::KeyPoint>&, const cv::Mat1b&, const cv::Mat1b&, float) {
    throw std::runtime_error("stand-in: method 3 is out of scope");
}
void removeDuplicatedMatches(vector<cv::DMatch>& matches);
#include "gen/fmatcher_132_245.inc"
#include "gen/fmatcher_247_266.inc"
}  // namespace geometry
}  // namespace vi_slam
}  // namespace refmf

extern "C" {
// matchFeatures(descriptors_1, descriptors_2, matches, 2, false) with lowe_method_dist_ratio as the reference would read it
// (an int).  The reference caches the value in a function-local static on the first call: one ratio per process.
int orb_refm_match_features(const uint8_t* d1, int n1, const uint8_t* d2, int n2, int loweRatioAsInt, int32_t* out /* [n1][3] */) {
    using namespace refmf;
    vi_slam::basics::Config::values()["lowe_method_dist_ratio"] = loweRatioAsInt;
    vi_slam::basics::Config::values()["xiang_gao_method_match_ratio"] = 2;
    vi_slam::basics::Config::values()["method_3_feature_dist_threshold"] = 50;
    ::cv::Mat1b a(::cv::Mat(n1, 32, CV_8UC1, (void*)d1, 32)), b(::cv::Mat(n2, 32, CV_8UC1, (void*)d2, 32));
    std::vector<::cv::DMatch> matches;
    vi_slam::geometry::matchFeatures(a, b, matches, 2, false, std::vector<::cv::KeyPoint>(), std::vector<::cv::KeyPoint>(), 0.f);
    for (size_t i = 0; i < matches.size(); ++i) {
        out[3 * i] = matches[i].queryIdx; out[3 * i + 1] = matches[i].trainIdx; out[3 * i + 2] = (int32_t)matches[i].distance;
    }
    return (int)matches.size();
}
}  // extern "C"
