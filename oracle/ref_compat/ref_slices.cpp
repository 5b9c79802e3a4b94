// ref_slices.cpp — the reference's OWN function bodies of the matcher rows, compiled from where they lie.
// TEST INFRASTRUCTURE ONLY (part of oracle/_ref/liborb_ref.so, built by oracle/ref.mk).
//
// src/datastructures/frame.cpp and src/geometry/fmatcher.cpp cannot be compiled as translation units here: their headers
// pull in Eigen, Sophus, g2o, DBoW3, Pangolin and the whole Frame / KeyFrame / MapPoint / Atlas graph.  But the functions of
// SURVEY section 8 rows a-8, a-10 and a-11 (first half) touch only std containers, cv::KeyPoint / cv::Mat and a dozen data
// members.  oracle/ref.mk therefore cuts exactly those function DEFINITIONS out of the reference files by line range at build
// time (sed, into oracle/_ref/gen/*.inc - build output, never committed; every slice is checked to start with the expected
// signature and the files' sha256 are recorded), and this file supplies what they need to compile: stand-in class
// declarations with the members the slices use, under the reference's names and namespaces.  What executes is the
// reference's own source text:
//     frame.cpp:386-414      Frame::AssignFeaturesToGrid          frame.cpp:678-744   Frame::GetFeaturesInArea
//     frame.cpp:746-756      Frame::PosInGrid                     frame.cpp:823-997   Frame::ComputeStereoMatches
//     frame.cpp:758-791      Frame::UndistortKeyPoints            frame.cpp:793-821   Frame::ComputeImageBounds   (row f-2;
//                            cv::undistortPoints itself = the oracle's restatement, pinned bit for bit against cv2)
//     fmatcher.cpp:313-319   TH_HIGH / TH_LOW / HISTO_LENGTH, ctor fmatcher.cpp:983-1098   FMatcher::SearchForInitialization
//     fmatcher.cpp:2813-2854 FMatcher::ComputeThreeMaxima         fmatcher.cpp:2859-2875   FMatcher::DescriptorDistance
//     fmatcher.cpp:2471-2687 FMatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono), both camera branches
//     cameramodels/pinhole.cpp:13-16, 22-26  Pinhole::project(Point3f), project(Mat)
// The image pyramids ComputeStereoMatches reads are those of the reference's own FExtractor object code (ref_driver.cpp).
#include "vi_slam/geometry/fextractor.h"      // the reference's own header (through oracle/ref_compat's opencv2/ stand-ins)

#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <thread>
#include <utility>
#include <vector>

// the oracle's brute-force kNN-2 + ratio (orb_oracle.cpp, same library): OpenCV's BFMatcher is not linkable here
extern "C" void orc_knn2(const uint8_t* q, int nq, const uint8_t* t, int nt, int32_t* idx, int32_t* dist, int nthreads);
extern "C" int orc_ratio_filter(const int32_t* idx, const int32_t* dist, int nq, double ratio, int rule, int32_t* out);

// reference include/vi_slam/datastructures/frame.h:42-43
#define FRAME_GRID_ROWS 48
#define FRAME_GRID_COLS 64

namespace vi_slam {
namespace geometry {
// reference include/vi_slam/geometry/cameramodels/{camera,pinhole}.h: what SearchByProjection calls
class Camera {
public:
    virtual ~Camera() {}
    virtual cv::Point2f project(const cv::Point3f& p3D) = 0;
    virtual cv::Point2f project(const cv::Mat& m3D) = 0;
    std::vector<float> mvParameters;
};
class Pinhole : public Camera {
public:
    cv::Point2f project(const cv::Point3f& p3D);
    cv::Point2f project(const cv::Mat& m3D);
};
#include "gen/pinhole_13_16.inc"
#include "gen/pinhole_22_26.inc"
}  // namespace geometry

namespace datastructures {
using namespace std;
using namespace geometry;

// reference include/vi_slam/datastructures/mappoint.h: the three calls SearchByProjection makes
class MapPoint {
public:
    int index = -1;
    cv::Mat mWorldPos, mDescriptor;
    int nObs = 0;
    cv::Mat GetWorldPos() { return mWorldPos.clone(); }
    cv::Mat GetDescriptor() { return mDescriptor.clone(); }
    int Observations() { return nObs; }
};

// the members of datastructures::Frame (frame.h) the slices read or write - same names, same types
class Frame {
public:
    cv::Mat T_w_c_, mTrl, mK, mDistCoef;
    geometry::Camera* mpCamera = nullptr;
    std::vector<MapPoint*> mvpMapPoints;
    std::vector<bool> mvbOutlier;
    int N = 0;
    int Nleft = -1;
    std::vector<cv::KeyPoint> keypoints_, keypointsRight_, ukeypoints_;
    cv::Mat descriptors_, descriptorsRight_;
    std::vector<float> mvuRight, mvDepth;
    std::vector<float> mvScaleFactors, mvInvScaleFactors;
    float mb = 0, mbf = 0;
    geometry::FExtractor *mpORBextractorLeft = nullptr, *mpORBextractorRight = nullptr;
    float mnMinX = 0, mnMaxX = 0, mnMinY = 0, mnMaxY = 0;
    float mfGridElementWidthInv = 0, mfGridElementHeightInv = 0;
    std::vector<std::size_t> mGrid[FRAME_GRID_COLS][FRAME_GRID_ROWS];
    std::vector<std::size_t> mGridRight[FRAME_GRID_COLS][FRAME_GRID_ROWS];

    void AssignFeaturesToGrid();
    vector<size_t> GetFeaturesInArea(const float& x, const float& y, const float& r, const int minLevel = -1, const int maxLevel = -1,
                                     const bool bRight = false) const;
    bool PosInGrid(const cv::KeyPoint& kp, int& posX, int& posY);
    void ComputeStereoMatches();
    void UndistortKeyPoints();
    void ComputeImageBounds(const cv::Mat& imLeft);
};
}  // namespace datastructures

namespace geometry {
using namespace datastructures;
using namespace std;

// reference include/vi_slam/geometry/fmatcher.h: the members the slices use
class FMatcher {
public:
    FMatcher(float nnratio = 0.6, bool checkOri = true);
    int SearchForInitialization(Frame& F1, Frame& F2, vector<cv::Point2f>& vbPrevMatched, vector<int>& vnMatches12, int windowSize = 10);
    int SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, const float th, const bool bMono);
    static int DescriptorDistance(const cv::Mat& a, const cv::Mat& b);
    static const int TH_LOW;
    static const int TH_HIGH;
    static const int HISTO_LENGTH;

protected:
    void ComputeThreeMaxima(vector<int>* histo, const int L, int& ind1, int& ind2, int& ind3);
    float mfNNratio;
    bool mbCheckOrientation;
};

#include "gen/fmatcher_313_319.inc"
#include "gen/fmatcher_983_1098.inc"
#include "gen/fmatcher_2813_2854.inc"
#include "gen/fmatcher_2859_2875.inc"
#include "gen/fmatcher_2471_2687.inc"
}  // namespace geometry

namespace datastructures {
#include "gen/frame_386_414.inc"
#include "gen/frame_678_744.inc"
#include "gen/frame_746_756.inc"
#include "gen/frame_823_997.inc"
#include "gen/frame_758_791.inc"
#include "gen/frame_793_821.inc"
}  // namespace datastructures
}  // namespace vi_slam

// ---------------------------------------------------------------------------------------------
using vi_slam::datastructures::Frame;
using vi_slam::geometry::FExtractor;
using vi_slam::geometry::FMatcher;

namespace {
void fill_frame(Frame& F, const cv::KeyPoint* kps, int n, const uint8_t* desc, const float* bounds) {
    F.N = n;
    F.keypoints_.assign(kps, kps + n);
    F.ukeypoints_ = F.keypoints_;                     // no distortion: UndistortKeyPoints copies (frame.cpp:762-766)
    F.descriptors_ = cv::Mat(n, 32, CV_8UC1, (void*)desc, 32);
    if (bounds) {
        F.mnMinX = bounds[0]; F.mnMaxX = bounds[1]; F.mnMinY = bounds[2]; F.mnMaxY = bounds[3];
        // frame.cpp:153-154 (the Frame constructors)
        F.mfGridElementWidthInv = static_cast<float>(FRAME_GRID_COLS) / static_cast<float>(F.mnMaxX - F.mnMinX);
        F.mfGridElementHeightInv = static_cast<float>(FRAME_GRID_ROWS) / static_cast<float>(F.mnMaxY - F.mnMinY);
        F.AssignFeaturesToGrid();
    }
}
}  // namespace

extern "C" {

int orb_refs_descriptor_distance(const uint8_t* a, const uint8_t* b) {
    cv::Mat ma(1, 32, CV_8UC1, (void*)a, 32), mb(1, 32, CV_8UC1, (void*)b, 32);
    return FMatcher::DescriptorDistance(ma, mb);
}

// Frame::ComputeStereoMatches on the pyramids held by two reference extractors (handles of orb_ref_create after
// orb_ref_compute on the left / right image) and the keypoints / descriptors orb_ref_compute returned.
// Returns 0, or -1 when the reference would index an empty vector (frame.cpp:984, nothing matched).
int orb_refs_stereo(void* hLeft, void* hRight, const cv::KeyPoint* kl, int nl, const uint8_t* dl, const cv::KeyPoint* kr, int nr,
                    const uint8_t* dr, float bf, float b, float* uRight, float* depth) {
    FExtractor* L = (FExtractor*)hLeft;
    FExtractor* R = (FExtractor*)hRight;
    Frame F;
    fill_frame(F, kl, nl, dl, nullptr);
    F.keypointsRight_.assign(kr, kr + nr);
    F.descriptorsRight_ = cv::Mat(nr, 32, CV_8UC1, (void*)dr, 32);
    F.mpORBextractorLeft = L;
    F.mpORBextractorRight = R;
    F.mvScaleFactors = L->GetScaleFactors();
    F.mvInvScaleFactors = L->GetInverseScaleFactors();
    F.mbf = bf;
    F.mb = b;
    // the guard the reference lacks: with no accepted match it reads vDistIdx[0] of an empty vector (documented deviation)
    F.ComputeStereoMatches();
    for (int i = 0; i < nl; ++i) { uRight[i] = F.mvuRight[i]; depth[i] = F.mvDepth[i]; }
    return 0;
}

// Frame::GetFeaturesInArea on a frame built from (kps, bounds); returns the count, indices in the reference's order
int orb_refs_features_in_area(const cv::KeyPoint* kps, int n, const float* bounds, float x, float y, float r, int minLevel, int maxLevel,
                              int32_t* out, int cap) {
    Frame F;
    std::vector<uint8_t> none((size_t)std::max(n, 1) * 32);
    fill_frame(F, kps, n, none.data(), bounds);
    std::vector<size_t> v = F.GetFeaturesInArea(x, y, r, minLevel, maxLevel);
    for (size_t i = 0; i < v.size() && (int)i < cap; ++i) out[i] = (int32_t)v[i];
    return (int)v.size();
}

// FMatcher(nnRatio, checkOrientation).SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize)
int orb_refs_search_for_initialization(const cv::KeyPoint* k1, int n1, const uint8_t* d1, const cv::KeyPoint* k2, int n2, const uint8_t* d2,
                                       const float* bounds, float* prevMatched /* in/out [n1][2] */, int windowSize, float nnRatio,
                                       int checkOrientation, int32_t* matches12) {
    Frame F1, F2;
    fill_frame(F1, k1, n1, d1, bounds);
    fill_frame(F2, k2, n2, d2, bounds);
    std::vector<cv::Point2f> prev(n1);
    for (int i = 0; i < n1; ++i) prev[i] = cv::Point2f(prevMatched[2 * i], prevMatched[2 * i + 1]);
    std::vector<int> m12;
    FMatcher matcher(nnRatio, checkOrientation != 0);
    const int n = matcher.SearchForInitialization(F1, F2, prev, m12, windowSize);
    for (int i = 0; i < n1; ++i) {
        matches12[i] = m12[i];
        prevMatched[2 * i] = prev[i].x;
        prevMatched[2 * i + 1] = prev[i].y;
    }
    return n;
}

// FMatcher(nnRatio, checkOrientation).SearchByProjection(CurrentFrame, LastFrame, th, bMono) on a synthetic scene:
// last-frame map points Xw (NaN x = no map point), poses as row-major 4x4 (world -> camera), pinhole (fx, fy, cx, cy).
// Current frame: nLeft2 < 0 = one camera (k2 / d2, optional mvuRight), else a rig: k2 / d2 hold Nleft left then the right
// keypoints, trl = row-major 3x4 left -> right transform.  Outputs: assigned[n2] = last-frame index matched to each current
// keypoint or -1; and what the caller of the device API would compute on the host, bit for bit as the reference's own
// text computes it here: uv / uvRight / invz / valid per last-frame point and the direction flag (0, 1 forward, 2 backward).
int orb_refs_search_by_projection(int nLast, const cv::KeyPoint* kLast, const float* Xw, const uint8_t* descMP, const uint8_t* outlier,
                                  const uint8_t* observed, const cv::KeyPoint* k2, int n2, const uint8_t* d2, const float* uRight2,
                                  const uint8_t* blocked2, int nLeft2, const float* bounds, const float* TcwCur, const float* TcwLast,
                                  const float* trl, const float* cam, const float* scaleFactors, int nLevels, float th, int bMono, float bf,
                                  float b, float nnRatio, int checkOrientation, int32_t* assigned, float* uvOut, float* uvRightOut,
                                  float* invzOut, uint8_t* validOut, int32_t* directionOut) {
    using vi_slam::datastructures::MapPoint;
    vi_slam::geometry::Pinhole camera;
    camera.mvParameters.assign(cam, cam + 4);
    std::vector<MapPoint> points(nLast), blockers(n2);
    Frame Last, Cur;
    Last.N = nLast;
    Last.Nleft = -1;
    Last.keypoints_.assign(kLast, kLast + nLast);
    Last.ukeypoints_ = Last.keypoints_;
    Last.mvpMapPoints.assign(nLast, nullptr);
    Last.mvbOutlier.assign(nLast, false);
    Last.T_w_c_ = cv::Mat(4, 4, CV_32F, (void*)TcwLast, 16).clone();
    for (int i = 0; i < nLast; ++i) {
        if (Xw[3 * i] != Xw[3 * i]) continue;
        points[i].index = i;
        points[i].mWorldPos = cv::Mat(3, 1, CV_32F, (void*)(Xw + 3 * i), 4).clone();
        points[i].mDescriptor = cv::Mat(1, 32, CV_8UC1, (void*)(descMP + 32 * i), 32).clone();
        points[i].nObs = observed[i] ? 1 : 0;
        Last.mvpMapPoints[i] = &points[i];
        Last.mvbOutlier[i] = outlier[i] != 0;
    }
    Cur.N = n2;
    Cur.Nleft = nLeft2 < 0 ? -1 : nLeft2;
    if (nLeft2 < 0) {
        Cur.keypoints_.assign(k2, k2 + n2);
        Cur.ukeypoints_ = Cur.keypoints_;
    } else {
        Cur.keypoints_.assign(k2, k2 + nLeft2);
        Cur.keypointsRight_.assign(k2 + nLeft2, k2 + n2);
    }
    Cur.descriptors_ = cv::Mat(n2, 32, CV_8UC1, (void*)d2, 32);
    Cur.mvuRight.assign(n2, -1.f);
    if (uRight2) Cur.mvuRight.assign(uRight2, uRight2 + n2);
    Cur.mvpMapPoints.assign(n2, nullptr);
    for (int i = 0; i < n2; ++i)
        if (blocked2 && blocked2[i]) { blockers[i].index = -1; blockers[i].nObs = 1; Cur.mvpMapPoints[i] = &blockers[i]; }
    Cur.mvScaleFactors.assign(scaleFactors, scaleFactors + nLevels);
    Cur.mbf = bf;
    Cur.mb = b;
    Cur.mpCamera = &camera;
    Cur.T_w_c_ = cv::Mat(4, 4, CV_32F, (void*)TcwCur, 16).clone();
    if (trl) Cur.mTrl = cv::Mat(3, 4, CV_32F, (void*)trl, 16).clone();
    Cur.mnMinX = bounds[0]; Cur.mnMaxX = bounds[1]; Cur.mnMinY = bounds[2]; Cur.mnMaxY = bounds[3];
    Cur.mfGridElementWidthInv = static_cast<float>(FRAME_GRID_COLS) / static_cast<float>(Cur.mnMaxX - Cur.mnMinX);
    Cur.mfGridElementHeightInv = static_cast<float>(FRAME_GRID_ROWS) / static_cast<float>(Cur.mnMaxY - Cur.mnMinY);
    Cur.AssignFeaturesToGrid();

    // what the host side of the device API computes, with the same expressions (fmatcher.cpp:2481-2516, 2594-2596)
    {
        const cv::Mat Rcw = Cur.T_w_c_.rowRange(0, 3).colRange(0, 3);
        const cv::Mat tcw = Cur.T_w_c_.rowRange(0, 3).col(3);
        const cv::Mat twc = -Rcw.t() * tcw;
        const cv::Mat Rlw = Last.T_w_c_.rowRange(0, 3).colRange(0, 3);
        const cv::Mat tlw = Last.T_w_c_.rowRange(0, 3).col(3);
        const cv::Mat tlc = Rlw * twc + tlw;
        const bool bForward = tlc.at<float>(2) > Cur.mb && !bMono;
        const bool bBackward = -tlc.at<float>(2) > Cur.mb && !bMono;
        *directionOut = bForward ? 1 : bBackward ? 2 : 0;
        for (int i = 0; i < nLast; ++i) {
            uvOut[2 * i] = uvOut[2 * i + 1] = invzOut[i] = 0.f;
            if (uvRightOut) uvRightOut[2 * i] = uvRightOut[2 * i + 1] = 0.f;
            validOut[i] = 0;
            MapPoint* pMP = Last.mvpMapPoints[i];
            if (!pMP || Last.mvbOutlier[i]) continue;
            cv::Mat x3Dw = pMP->GetWorldPos();
            cv::Mat x3Dc = Rcw * x3Dw + tcw;
            const float invzc = 1.0 / x3Dc.at<float>(2);
            if (invzc < 0) continue;
            cv::Point2f uv = Cur.mpCamera->project(x3Dc);
            if (uv.x < Cur.mnMinX || uv.x > Cur.mnMaxX) continue;
            if (uv.y < Cur.mnMinY || uv.y > Cur.mnMaxY) continue;
            uvOut[2 * i] = uv.x; uvOut[2 * i + 1] = uv.y; invzOut[i] = invzc; validOut[i] = 1;
            if (Cur.Nleft != -1 && uvRightOut) {
                cv::Mat x3Dr = Cur.mTrl.colRange(0, 3).rowRange(0, 3) * x3Dc + Cur.mTrl.col(3);
                cv::Point2f uvr = Cur.mpCamera->project(x3Dr);
                uvRightOut[2 * i] = uvr.x; uvRightOut[2 * i + 1] = uvr.y;
            }
        }
    }
    FMatcher matcher(nnRatio, checkOrientation != 0);
    const int n = matcher.SearchByProjection(Cur, Last, th, bMono != 0);
    for (int i = 0; i < n2; ++i) assigned[i] = Cur.mvpMapPoints[i] ? Cur.mvpMapPoints[i]->index : -1;
    return n;
}

// Frame::UndistortKeyPoints / ComputeImageBounds with mK = (fx, fy, cx, cy), mDistCoef = dist[ndist]
static void set_camera(Frame& F, const float* cam, const float* dist, int ndist) {
    F.mK = cv::Mat::zeros(3, 3, CV_32F);
    F.mK.at<float>(0, 0) = cam[0]; F.mK.at<float>(1, 1) = cam[1]; F.mK.at<float>(0, 2) = cam[2]; F.mK.at<float>(1, 2) = cam[3];
    F.mK.at<float>(2, 2) = 1.f;
    F.mDistCoef = cv::Mat::zeros(ndist, 1, CV_32F);
    for (int i = 0; i < ndist; ++i) F.mDistCoef.at<float>(i) = dist[i];
}
void orb_refs_undistort_keypoints(const cv::KeyPoint* kps, int n, const float* cam, const float* dist, int ndist, cv::KeyPoint* out) {
    Frame F;
    F.N = n;
    F.keypoints_.assign(kps, kps + n);
    set_camera(F, cam, dist, ndist);
    F.UndistortKeyPoints();
    for (int i = 0; i < n; ++i) out[i] = F.ukeypoints_[i];
}
void orb_refs_image_bounds(int cols, int rows, const float* cam, const float* dist, int ndist, float* bounds) {
    Frame F;
    set_camera(F, cam, dist, ndist);
    cv::Mat im(rows, cols, CV_8UC1);
    F.ComputeImageBounds(im);
    bounds[0] = F.mnMinX; bounds[1] = F.mnMaxX; bounds[2] = F.mnMinY; bounds[3] = F.mnMaxY;
}

// Timed CPU baseline of bench.py (cpu_baseline / --impl reference): `n` frames (stereo: left / right interleaved) dealt to
// `nthreads` std::threads, every thread with its own reference FExtractor per camera on glibc malloc.
//   mode 0  FExtractor::compute per frame
//   mode 1  stereo pairs: compute x 2, Frame::ComputeStereoMatches, kNN-2 + 0.7 ratio
//   mode 2  consecutive frames: compute, then FMatcher::SearchForInitialization + kNN-2 + 0.8 ratio per pair
// Returns the total number of keypoints.
long orb_refs_bench_frames(int nf, float sf, int nl, int ini, int mn, const uint8_t* imgs, int n, int W, int H, int lap0, int lap1, int mode,
                           float bf, float b, int nthreads, long* nMatches) {
    std::vector<long> tot(nthreads, 0), mt(nthreads, 0);
    const size_t fsz = (size_t)W * H;
    struct Out { std::vector<cv::KeyPoint> k; cv::Mat d; };
    std::vector<Out> kept(mode == 2 ? n : 0);
    auto extract = [&](FExtractor& e, const uint8_t* img, std::vector<cv::KeyPoint>& k, cv::Mat& d) {
        cv::Mat image(H, W, CV_8UC1, (void*)img, (size_t)W), mask;
        std::vector<int> lap = {lap0, lap1};
        e.compute(image, mask, k, d, lap);
    };
    auto knn_ratio = [&](const cv::Mat& dq, int nq, const cv::Mat& dt, int nt, double ratio, int rule) -> long {
        if (nq == 0 || nt == 0) return 0;
        std::vector<int32_t> idx(2 * (size_t)nq), dist(2 * (size_t)nq), out(3 * (size_t)nq);
        orc_knn2(dq.ptr(0), nq, dt.ptr(0), nt, idx.data(), dist.data(), 1);
        return orc_ratio_filter(idx.data(), dist.data(), nq, ratio, rule, out.data());
    };
    auto work = [&](int tid) {
        FExtractor eL(nf, sf, nl, ini, mn), eR(nf, sf, nl, ini, mn);
        if (mode == 1) {
            for (int i = tid; i < n / 2; i += nthreads) {
                Frame F;
                extract(eL, imgs + (size_t)(2 * i) * fsz, F.keypoints_, F.descriptors_);
                extract(eR, imgs + (size_t)(2 * i + 1) * fsz, F.keypointsRight_, F.descriptorsRight_);
                F.N = (int)F.keypoints_.size();
                tot[tid] += F.N + (long)F.keypointsRight_.size();
                if (F.N == 0 || F.keypointsRight_.empty()) continue;
                F.mpORBextractorLeft = &eL;
                F.mpORBextractorRight = &eR;
                F.mvScaleFactors = eL.GetScaleFactors();
                F.mvInvScaleFactors = eL.GetInverseScaleFactors();
                F.mbf = bf;
                F.mb = b;
                F.ComputeStereoMatches();
                for (float v : F.mvDepth) mt[tid] += v > 0;
                mt[tid] += knn_ratio(F.descriptors_, F.N, F.descriptorsRight_, (int)F.keypointsRight_.size(), 0.7, 0);
            }
        } else {
            for (int i = tid; i < n; i += nthreads) {
                std::vector<cv::KeyPoint> k;
                cv::Mat d;
                extract(eL, imgs + (size_t)i * fsz, k, d);
                tot[tid] += (long)k.size();
                if (mode == 2) { kept[i].k.swap(k); kept[i].d = d.clone(); }
            }
        }
    };
    auto match = [&](int tid) {
        const float bounds[4] = {0.f, (float)W, 0.f, (float)H};
        for (int p = tid; p + 1 < n; p += nthreads) {
            Out &a = kept[p], &c = kept[p + 1];
            if (a.k.empty() || c.k.empty()) continue;
            Frame F1, F2;
            fill_frame(F1, a.k.data(), (int)a.k.size(), a.d.ptr(0), bounds);
            fill_frame(F2, c.k.data(), (int)c.k.size(), c.d.ptr(0), bounds);
            std::vector<cv::Point2f> prev(a.k.size());
            for (size_t i = 0; i < a.k.size(); ++i) prev[i] = a.k[i].pt;
            std::vector<int> m12;
            FMatcher matcher(0.9f, true);
            mt[tid] += matcher.SearchForInitialization(F1, F2, prev, m12, 100);
            mt[tid] += knn_ratio(a.d, (int)a.k.size(), c.d, (int)c.k.size(), 0.8, 1);
        }
    };
    {
        std::vector<std::thread> th;
        for (int t = 0; t < nthreads; ++t) th.emplace_back(work, t);
        for (auto& x : th) x.join();
    }
    if (mode == 2) {
        std::vector<std::thread> th;
        for (int t = 0; t < nthreads; ++t) th.emplace_back(match, t);
        for (auto& x : th) x.join();
    }
    long s = 0, m = 0;
    for (int t = 0; t < nthreads; ++t) { s += tot[t]; m += mt[t]; }
    if (nMatches) *nMatches = m;
    return s;
}

}  // extern "C"
