// ref_driver.cpp — C entry points around the reference's OWN FExtractor object code.
// TEST INFRASTRUCTURE ONLY.  Built by oracle/ref.mk into oracle/_ref/liborb_ref.so together with
//   /root/reference/src/geometry/fextractor.cpp      (compiled where it lies, byte for byte)
//   oracle/ref_compat/cv_compat.cpp                  (OpenCV stand-in; primitives = cv2-pinned restatements)
//   oracle/orb_oracle.cpp                            (only its orc_fast / orc_resize_linear / orc_gaussian7 /
//                                                     orc_fast_atan2 primitives are reached from here)
// so tests can pin oracle/orb_oracle.cpp's restatement of the reference's control flow — and the CUDA
// path — against the reference itself (reference fextractor.cpp:401-461, 472-844, 1034-1160).
//
// The one place the reference is undefined is DistributeOctTree's sort of (size, node POINTER)
// (fextractor.cpp:675): equal-size nodes expand in heap-address order.  Two allocator modes make that
// measurable: `monotone = 0` runs on glibc malloc as shipped; `monotone = 1` serves every operator new of
// this library from a bump arena that never reuses an address, so address order == allocation order,
// which is the rule the oracle and the CUDA kernel implement ("later-created node first").
#include "vi_slam/geometry/fextractor.h"

#include <sys/mman.h>

#include <atomic>
#include <cstdint>
#include <mutex>
#include <new>

// ---------------------------------------------------------------------------------------------
// allocator: replaces operator new/delete for this shared object only (linked -Bsymbolic)
// ---------------------------------------------------------------------------------------------
namespace {
const size_t kArenaBytes = (size_t)24 << 30;          // virtual, MAP_NORESERVE; pages are dropped after each call
uint8_t* g_arena = nullptr;
std::atomic<size_t> g_top{0};
std::atomic<long> g_fallbacks{0};
thread_local bool t_monotone = false;
std::mutex g_monotone_mutex;                           // one monotone call at a time

void arena_init() {
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = mmap(nullptr, kArenaBytes, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
        g_arena = p == MAP_FAILED ? nullptr : (uint8_t*)p;
    });
}
inline bool in_arena(const void* p) { return g_arena && (const uint8_t*)p >= g_arena && (const uint8_t*)p < g_arena + kArenaBytes; }
void* ref_alloc(size_t n) {
    if (t_monotone && g_arena) {
        size_t need = (n + 15) & ~(size_t)15;
        size_t at = g_top.fetch_add(need);
        if (at + need <= kArenaBytes) return g_arena + at;
        ++g_fallbacks;
    }
    void* p = std::malloc(n ? n : 1);
    if (!p) throw std::bad_alloc();
    return p;
}
void ref_free(void* p) { if (p && !in_arena(p)) std::free(p); }
void arena_reset() {
    size_t used = g_top.exchange(0);
    if (g_arena && used) madvise(g_arena, (used + 4095) & ~(size_t)4095, MADV_DONTNEED);
}
}  // namespace

void* operator new(size_t n) { return ref_alloc(n); }
void* operator new[](size_t n) { return ref_alloc(n); }
void operator delete(void* p) noexcept { ref_free(p); }
void operator delete[](void* p) noexcept { ref_free(p); }
void operator delete(void* p, size_t) noexcept { ref_free(p); }
void operator delete[](void* p, size_t) noexcept { ref_free(p); }

// ---------------------------------------------------------------------------------------------
namespace {
using vi_slam::geometry::FExtractor;

struct LevelList { std::vector<int32_t> xys; std::vector<float> angle; };

// FExtractor's stages are protected; a derived class may call them.
struct Probe : FExtractor {
    Probe(int nf, float sf, int nl, int ini, int mn) : FExtractor(nf, sf, nl, ini, mn) {}
    std::vector<LevelList> cand, kept;         // of the last call
    bool consistent = true;

    const std::vector<int>& featuresPerLevel() const { return mnFeaturesPerLevel; }
    const std::vector<int>& uMax() const { return umax; }

    // candidates (vToDistributeKeys) per level from the FAST calls the reference made
    void collectCandidates() {
        auto& tr = cv_compat::fast_trace();
        cand.assign(nlevels, LevelList());
        for (size_t c = 0; c < tr.size(); ++c) {
            const cv_compat::FastCall& fc = tr[c];
            // a cell whose iniThFAST call came back empty is called again with minThFAST on the same ROI
            if (c + 1 < tr.size() && tr[c + 1].origin == fc.origin && fc.out.empty()) continue;
            int level = -1, ox = 0, oy = 0;
            for (int l = 0; l < nlevels; ++l) {
                const cv::Mat& m = mvImagePyramid[l];
                if (fc.origin < m.data || fc.origin >= m.data + (size_t)m.rows * m.step) continue;
                size_t d = (size_t)(fc.origin - m.data);
                level = l; oy = (int)(d / m.step); ox = (int)(d % m.step);
                break;
            }
            if (level < 0) { consistent = false; continue; }
            for (const cv::KeyPoint& k : fc.out) {
                cand[level].xys.push_back(ox - 16 + (int)k.pt.x);
                cand[level].xys.push_back(oy - 16 + (int)k.pt.y);
                cand[level].xys.push_back((int)k.response);
            }
        }
    }

    // the per-level retained lists (order, angle) by running the two protected stages directly
    void collectKept(const cv::Mat& image) {
        ComputePyramid(image);
        std::vector<std::vector<cv::KeyPoint>> all;
        ComputeKeyPointsOctTree(all);
        kept.assign(nlevels, LevelList());
        for (int l = 0; l < nlevels; ++l)
            for (const cv::KeyPoint& k : all[l]) {
                kept[l].xys.push_back((int)k.pt.x);
                kept[l].xys.push_back((int)k.pt.y);
                kept[l].xys.push_back((int)k.response);
                kept[l].angle.push_back(k.angle);
            }
    }
};

struct MonotoneScope {
    bool on;
    explicit MonotoneScope(bool m) : on(m) {
        if (on) { arena_init(); g_monotone_mutex.lock(); t_monotone = true; }
    }
    ~MonotoneScope() {
        if (on) { t_monotone = false; arena_reset(); g_monotone_mutex.unlock(); }
    }
};
}  // namespace

extern "C" {

void* orb_ref_create(int nf, float sf, int nl, int ini, int mn) { return new Probe(nf, sf, nl, ini, mn); }
void orb_ref_destroy(void* h) { delete (Probe*)h; }

int orb_ref_tables(void* h, float* scale, float* inv, float* sig2, float* invsig2, int* feat, int* umax) {
    Probe* e = (Probe*)h;
    int n = e->GetLevels();
    std::vector<float> a = e->GetScaleFactors(), b = e->GetInverseScaleFactors(), c = e->GetScaleSigmaSquares(),
                       d = e->GetInverseScaleSigmaSquares();
    for (int i = 0; i < n; ++i) { scale[i] = a[i]; inv[i] = b[i]; sig2[i] = c[i]; invsig2[i] = d[i]; feat[i] = e->featuresPerLevel()[i]; }
    for (int i = 0; i < 16; ++i) umax[i] = e->uMax()[i];
    return n;
}

// FExtractor::compute on a caller image.  Returns the keypoint count (-1 on an empty image); *mono = compute's
// return value.  Afterwards the level images, the candidates and the per-level retained lists of this image
// can be read with the functions below.
int orb_ref_compute(void* h, const uint8_t* img, int W, int H, int step, int lap0, int lap1, cv::KeyPoint* kps,
                    uint8_t* desc, int cap, int* mono, int monotone) {
    Probe* e = (Probe*)h;
    int n = -1;
    {
        // everything the reference allocates during the call must die before the arena is recycled, so the
        // persistent pyramid is dropped and rebuilt outside the scope below
        MonotoneScope scope(monotone != 0);
        {
            cv::Mat image = (img && W > 0 && H > 0) ? cv::Mat(H, W, CV_8UC1, (void*)img, (size_t)step) : cv::Mat();
            std::vector<cv::KeyPoint> k;
            cv::Mat d, mask;
            std::vector<int> lap = {lap0, lap1};
            cv_compat::fast_trace().clear();
            cv_compat::fast_trace_enable(true);
            int m = e->compute(image, mask, k, d, lap);
            cv_compat::fast_trace_enable(false);
            if (mono) *mono = m;
            if (m >= 0) {
                n = (int)k.size();
                for (int i = 0; i < std::min(n, cap); ++i) {
                    kps[i] = k[i];
                    std::memcpy(desc + (size_t)32 * i, d.ptr(i), 32);
                }
                e->consistent = true;
                e->collectCandidates();
                cv_compat::fast_trace().clear();
                e->collectKept(image);
                // the second pass must have kept the same keypoints as compute() did
                size_t total = 0;
                for (auto& l : e->kept) total += l.angle.size();
                if ((int)total != n) e->consistent = false;
            }
            std::vector<cv_compat::FastCall>().swap(cv_compat::fast_trace());
        }
        if (monotone) {
            // what outlives the call must not live in the arena: deep-copy it onto malloc; the arena-backed
            // originals are dropped (operator delete ignores arena addresses)
            t_monotone = false;
            std::vector<LevelList> c2 = e->cand, k2 = e->kept;
            for (size_t l = 0; l < e->mvImagePyramid.size(); ++l) e->mvImagePyramid[l] = e->mvImagePyramid[l].clone();
            e->cand.swap(c2);
            e->kept.swap(k2);
            t_monotone = true;
        }
    }
    return n;
}

// FExtractor::compute and nothing else (no FAST trace, no second probing pass, glibc malloc): the timed CPU baseline of
// bench.py.  Thread-safe across handles.  Returns the keypoint count, -1 on an empty image.
int orb_ref_compute_plain(void* h, const uint8_t* img, int W, int H, int step, int lap0, int lap1, cv::KeyPoint* kps, uint8_t* desc,
                          int cap, int* mono) {
    Probe* e = (Probe*)h;
    cv::Mat image = (img && W > 0 && H > 0) ? cv::Mat(H, W, CV_8UC1, (void*)img, (size_t)step) : cv::Mat();
    std::vector<cv::KeyPoint> k;
    cv::Mat d, mask;
    std::vector<int> lap = {lap0, lap1};
    cv_compat::fast_trace_enable(false);
    const int m = e->compute(image, mask, k, d, lap);
    if (mono) *mono = m;
    if (m < 0) return -1;
    const int n = (int)k.size();
    for (int i = 0; i < std::min(n, cap); ++i) {
        kps[i] = k[i];
        std::memcpy(desc + (size_t)32 * i, d.ptr(i), 32);
    }
    return n;
}

int orb_ref_consistent(void* h) { return ((Probe*)h)->consistent ? 1 : 0; }

int orb_ref_level_size(void* h, int l, int* w, int* hh) {
    Probe* e = (Probe*)h; *w = e->mvImagePyramid[l].cols; *hh = e->mvImagePyramid[l].rows; return 0;
}
int orb_ref_level_image(void* h, int l, uint8_t* dst) {
    const cv::Mat& m = ((Probe*)h)->mvImagePyramid[l];
    for (int y = 0; y < m.rows; ++y) std::memcpy(dst + (size_t)y * m.cols, m.ptr(y), (size_t)m.cols);
    return 0;
}
// vToDistributeKeys of level l in the reference's order: x, y relative to the 16-px FAST border, score
int orb_ref_level_candidates(void* h, int l, int32_t* xys, int cap) {
    const LevelList& c = ((Probe*)h)->cand[l];
    int n = (int)c.xys.size() / 3;
    std::memcpy(xys, c.xys.data(), sizeof(int32_t) * 3 * std::min(n, cap));
    return n;
}
// retained keypoints of level l in DistributeOctTree's list order: x, y in level pixels, score; angle in degrees
int orb_ref_level_keypoints(void* h, int l, int32_t* xys, float* angle, int cap) {
    const LevelList& c = ((Probe*)h)->kept[l];
    int n = (int)c.angle.size();
    std::memcpy(xys, c.xys.data(), sizeof(int32_t) * 3 * std::min(n, cap));
    std::memcpy(angle, c.angle.data(), sizeof(float) * std::min(n, cap));
    return n;
}
long orb_ref_arena_fallbacks() { return g_fallbacks.load(); }

// For programs that use vi_slam::geometry::FExtractor from this library DIRECTLY (tests/cpp/dropin_driver.cpp): switch
// the calling thread to the allocation-ordered arena for the rest of the process (nothing is recycled).
void orb_ref_monotone(int on) {
    if (on) arena_init();
    t_monotone = on != 0;
}

}  // extern "C"
