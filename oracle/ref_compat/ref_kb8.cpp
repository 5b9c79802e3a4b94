// ref_kb8.cpp — the reference's OWN fisheye stereo path (row f-4), compiled from where it lies.
// TEST INFRASTRUCTURE ONLY (part of oracle/_ref/liborb_ref.so, built by oracle/ref.mk; see ref_slices.cpp for the method).
// Slices, unmodified:
//     cameramodels/kannalabrandt8.cpp:11-27    KannalaBrandt8::project(Point3f)        :33-37   project(Mat)
//     cameramodels/kannalabrandt8.cpp:70-74    unprojectMat                            :82-109  unproject (Newton iteration)
//     cameramodels/kannalabrandt8.cpp:324-393  TriangulateMatches                      :485-498 Triangulate
//     datastructures/frame.cpp:1149-1189       Frame::ComputeStereoFishEyeMatches
// Stand-ins: the class declarations below; the small-matrix algebra of oracle/ref_compat/cv_compat.hpp; cv::SVD::compute =
// the oracle's one-sided Jacobi null vector (orb_oracle.cpp; the real cv::SVD is pinned through tests/golden/ref_kb8.py);
// cv::BFMatcher(NORM_HAMMING).knnMatch = the oracle's kNN-2 (pinned against cv2 in tests/test_oracle_golden.py).
#include "cv_compat.hpp"

#include <cmath>
#include <cstdint>
#include <vector>

extern "C" void orc_null_vector4(const float* A16, float* x4);
extern "C" void orc_knn2(const uint8_t* q, int nq, const uint8_t* t, int nt, int32_t* idx, int32_t* dist, int nthreads);

namespace cv {
struct SVD {
    enum Flags { MODIFY_A = 1, NO_UV = 2, FULL_UV = 4 };
    // only vt.row(3) - the right singular vector of the smallest singular value - is read by the caller
    static void compute(const Mat& A, Mat& w, Mat& u, Mat& vt, int = 0) {
        float a[16], x[4];
        for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) a[4 * i + j] = A.at<float>(i, j);
        orc_null_vector4(a, x);
        w = Mat::zeros(4, 1, CV_32F);
        u = Mat::zeros(4, 4, CV_32F);
        vt = Mat::zeros(4, 4, CV_32F);
        for (int j = 0; j < 4; ++j) vt.at<float>(3, j) = x[j];
    }
};
class BFMatcher {
public:
    explicit BFMatcher(int normType = NORM_HAMMING) { (void)normType; }
    void knnMatch(const Mat& q, const Mat& t, std::vector<std::vector<DMatch>>& matches, int k) const {
        assert(k == 2); (void)k;
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
}  // namespace cv

// The stand-in Camera / Frame declarations differ from those of ref_slices.cpp (same library): an outer namespace keeps their
// implicit members and vtables apart; inside it the slices still see vi_slam::geometry / vi_slam::datastructures.
namespace refkb8 {
namespace vi_slam {
namespace geometry {
using namespace std;
// reference include/vi_slam/geometry/cameramodels/{camera,kannalabrandt8}.h: what the slices use
class Camera {
public:
    virtual ~Camera() {}
    virtual cv::Point2f project(const cv::Point3f& p3D) = 0;
    virtual cv::Point2f project(const cv::Mat& m3D) = 0;
    virtual cv::Point3f unproject(const cv::Point2f& p2D) = 0;
    virtual cv::Mat unprojectMat(const cv::Point2f& p2D) = 0;
    std::vector<float> mvParameters;
};
class KannalaBrandt8 : public Camera {
public:
    explicit KannalaBrandt8(float _precision) : precision(_precision) {}
    cv::Point2f project(const cv::Point3f& p3D);
    cv::Point2f project(const cv::Mat& m3D);
    cv::Point3f unproject(const cv::Point2f& p2D);
    cv::Mat unprojectMat(const cv::Point2f& p2D);
    float TriangulateMatches(Camera* pCamera2, const cv::KeyPoint& kp1, const cv::KeyPoint& kp2, const cv::Mat& R12, const cv::Mat& t12,
                             const float sigmaLevel, const float unc, cv::Mat& p3D);
    const float precision;

private:
    void Triangulate(const cv::Point2f& p1, const cv::Point2f& p2, const cv::Mat& Tcw1, const cv::Mat& Tcw2, cv::Mat& x3D);
};
#include "gen/kannalabrandt8_11_27.inc"
#include "gen/kannalabrandt8_33_37.inc"
#include "gen/kannalabrandt8_70_74.inc"
#include "gen/kannalabrandt8_82_109.inc"
#include "gen/kannalabrandt8_324_393.inc"
#include "gen/kannalabrandt8_485_498.inc"
}  // namespace geometry

namespace datastructures {
using namespace std;
// the members of datastructures::Frame (frame.h) ComputeStereoFishEyeMatches reads or writes
class Frame {
public:
    int Nleft = 0, Nright = 0, monoLeft = 0, monoRight = 0, mnCloseMPs = 0;
    std::vector<cv::KeyPoint> keypoints_, keypointsRight_;
    cv::Mat descriptors_, descriptorsRight_;
    std::vector<int> mvLeftToRightMatch, mvRightToLeftMatch;
    std::vector<float> mvDepth, mvuRight, mvLevelSigma2;
    std::vector<cv::Mat> mvStereo3Dpoints;
    geometry::Camera *mpCamera = nullptr, *mpCamera2 = nullptr;
    cv::Mat mRlr, mtlr;
    static cv::BFMatcher BFmatcher;
    void ComputeStereoFishEyeMatches();
};
cv::BFMatcher Frame::BFmatcher = cv::BFMatcher(cv::NORM_HAMMING);      // frame.cpp:366
#include "gen/frame_1149_1189.inc"
}  // namespace datastructures
}  // namespace vi_slam
}  // namespace refkb8

extern "C" {

// KannalaBrandt8::TriangulateMatches per match.  rig = prm1[8] prm2[8] R12[9] (row-major) t12[3] precision.
void orb_refk_triangulate(const float* rig, const float* pt1, const float* pt2, const float* sigma1, const float* sigma2, int n, float* depth,
                          float* p3d) {
    using namespace refkb8::vi_slam::geometry;
    KannalaBrandt8 c1(rig[28]), c2(rig[28]);
    c1.mvParameters.assign(rig, rig + 8);
    c2.mvParameters.assign(rig + 8, rig + 16);
    cv::Mat R12 = cv::Mat(3, 3, CV_32F, (void*)(rig + 16), 12).clone(), t12 = cv::Mat(3, 1, CV_32F, (void*)(rig + 25), 4).clone();
    for (int i = 0; i < n; ++i) {
        cv::KeyPoint k1(pt1[2 * i], pt1[2 * i + 1], 1.f), k2(pt2[2 * i], pt2[2 * i + 1], 1.f);
        cv::Mat p;
        depth[i] = c1.TriangulateMatches(&c2, k1, k2, R12, t12, sigma1[i], sigma2[i], p);
        for (int j = 0; j < 3; ++j) p3d[3 * i + j] = (depth[i] > 0 && !p.empty()) ? p.at<float>(j) : 0.f;
    }
}

// Frame::ComputeStereoFishEyeMatches -> mvLeftToRightMatch, mvRightToLeftMatch, mvDepth, mvStereo3Dpoints
void orb_refk_fisheye_matches(const float* rig, const cv::KeyPoint* kl, const uint8_t* dl, int nl, int monoLeft, const cv::KeyPoint* kr,
                              const uint8_t* dr, int nr, int monoRight, const float* levelSigma2, int nLevels, int32_t* l2r, int32_t* r2l,
                              float* depth, float* p3d) {
    using namespace refkb8::vi_slam::geometry;
    KannalaBrandt8 c1(rig[28]), c2(rig[28]);
    c1.mvParameters.assign(rig, rig + 8);
    c2.mvParameters.assign(rig + 8, rig + 16);
    refkb8::vi_slam::datastructures::Frame F;
    F.Nleft = nl; F.Nright = nr; F.monoLeft = monoLeft; F.monoRight = monoRight;
    F.keypoints_.assign(kl, kl + nl);
    F.keypointsRight_.assign(kr, kr + nr);
    F.descriptors_ = cv::Mat(nl, 32, CV_8UC1, (void*)dl, 32);
    F.descriptorsRight_ = cv::Mat(nr, 32, CV_8UC1, (void*)dr, 32);
    F.mvLevelSigma2.assign(levelSigma2, levelSigma2 + nLevels);
    F.mpCamera = &c1;
    F.mpCamera2 = &c2;
    F.mRlr = cv::Mat(3, 3, CV_32F, (void*)(rig + 16), 12).clone();
    F.mtlr = cv::Mat(3, 1, CV_32F, (void*)(rig + 25), 4).clone();
    F.ComputeStereoFishEyeMatches();
    for (int i = 0; i < nl; ++i) {
        l2r[i] = F.mvLeftToRightMatch[i];
        depth[i] = F.mvDepth[i];
        for (int j = 0; j < 3; ++j) p3d[3 * i + j] = F.mvStereo3Dpoints[i].empty() ? 0.f : F.mvStereo3Dpoints[i].at<float>(j);
    }
    for (int i = 0; i < nr; ++i) r2l[i] = F.mvRightToLeftMatch[i];
}

}  // extern "C"
