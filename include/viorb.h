/* viorb.h — C ABI of the B200-native ORB front-end (libviorb.so).
 *
 * Drop-in boundary for the feature path of KMS-TEAM/vi_slam.  Every entry point
 * names the reference interface it replaces (paths relative to the reference
 * repository).  Plain pointers and sizes only; no C++/torch types; no exceptions
 * cross this boundary.  All functions return VIORB_OK (0) or a negative
 * viorb_status; viorb_last_error() gives a thread-local message.
 *
 * Threading contract (reference src/datastructures/frame.cpp:107-111): one
 * extractor handle per camera; different handles may be driven concurrently
 * from different host threads (each owns its CUDA stream and arenas); a single
 * handle must not be entered concurrently.
 */
#ifndef VIORB_H_
#define VIORB_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum viorb_status {
    VIORB_OK = 0,
    VIORB_E_EMPTY = -1,        /* empty image: FExtractor::compute returns -1 (fextractor.cpp:1037) */
    VIORB_E_ARG = -2,
    VIORB_E_CUDA = -3,
    VIORB_E_CAPACITY = -4,     /* caller buffer / handle arena too small; sizes are reported */
    VIORB_E_NO_DEVICE = -5
} viorb_status;

/* cv::KeyPoint, 28-byte POD (what FExtractor::compute fills, fextractor.cpp:1108-1129):
 * pt in level-0 pixels, size = int(31*scale[octave]), angle in degrees [0,360),
 * response = FAST score, class_id = -1. */
typedef struct viorb_keypoint {
    float x, y, size, angle, response;
    int32_t octave, class_id;
} viorb_keypoint;

/* cv::DMatch-shaped result row (queryIdx, trainIdx, distance as integer Hamming). */
typedef struct viorb_match {
    int32_t query_idx, train_idx, distance;
} viorb_match;

/* FExtractor ctor arguments (include/vi_slam/geometry/fextractor.h:30-31, values from
 * Tracking::ParseORBParamFile, src/core/tracking.cpp:1020-1100) plus arena sizing. */
typedef struct viorb_params {
    int32_t nfeatures;      /* ORBextractor.nFeatures (2000; 5x for the mono initialiser) */
    float scale_factor;     /* ORBextractor.scaleFactor (1.2) */
    int32_t nlevels;        /* ORBextractor.nLevels (8), <= 16 */
    int32_t ini_th_fast;    /* ORBextractor.iniThFAST (20) */
    int32_t min_th_fast;    /* ORBextractor.minThFAST (7) */
    int32_t max_width;      /* largest image this handle will see */
    int32_t max_height;
    int32_t max_batch;      /* frames per extract_batch call (>= 1) */
    int32_t device;         /* CUDA device ordinal */
} viorb_params;

typedef struct viorb_extractor viorb_extractor;

const char* viorb_last_error(void);
const char* viorb_version(void);

/* ---- extractor: replaces class FExtractor (fextractor.h:26-64) ------------------------- */

/* FExtractor::FExtractor (fextractor.cpp:401-461). */
int viorb_extractor_create(const viorb_params* params, viorb_extractor** out);
void viorb_extractor_destroy(viorb_extractor* ex);

/* GetLevels/GetScaleFactor(s)/GetInverseScaleFactors/GetScaleSigmaSquares/
 * GetInverseScaleSigmaSquares (fextractor.h:40-62) and mnFeaturesPerLevel. Any pointer may be NULL. */
int viorb_extractor_tables(const viorb_extractor* ex, int32_t* nlevels, float* scale, float* inv_scale,
                           float* sigma2, float* inv_sigma2, int32_t* features_per_level);

/* Upper bound of keypoints one frame can produce (sum over levels of N_l + slack). */
int viorb_extractor_capacity(const viorb_extractor* ex);

/* Page-locked host memory.  Images and result buffers that live in memory from viorb_host_alloc (or any
 * cudaHostRegister'ed block) are DMA'd directly; pageable buffers go through the handle's staging area. */
void* viorb_host_alloc(size_t bytes);
void viorb_host_free(void* p);

/* FExtractor::compute (fextractor.cpp:1034-1133) for one host image (CV_8UC1, row step `step`).
 * lap0/lap1 = vLappingArea.  Writes up to `cap` keypoints/descriptors (row i of desc belongs to
 * keypoint i), *n_out = keypoint count, *mono_index_out = compute()'s return value.
 * Returns VIORB_E_EMPTY for an empty image (the reference returns -1). */
int viorb_extract(viorb_extractor* ex, const uint8_t* img, int width, int height, int step,
                  int lap0, int lap1, viorb_keypoint* kps, uint8_t* desc, int cap,
                  int* n_out, int* mono_index_out);

/* The same for n host images of one size (frame-batch path of BASELINE.json config 4).
 * imgs[i] points to image i; outputs are [n][cap] / [n][cap][32] / [n]. */
int viorb_extract_batch(viorb_extractor* ex, const uint8_t* const* imgs, int n, int width, int height, int step,
                        int lap0, int lap1, viorb_keypoint* kps, uint8_t* desc, int cap,
                        int* n_out, int* mono_index_out);

/* Device-resident variant: n images already in HBM at d_imgs + i*frame_stride, row step `step`
 * (step % 16 == 0 and 16-byte aligned base).  Results stay on the device inside the handle until the
 * next extract call; fetch them with viorb_fetch / use them with the *_resident matchers.
 * The image buffer must stay valid while the results are used (it is pyramid level 0). */
int viorb_extract_device(viorb_extractor* ex, const uint8_t* d_imgs, int n, int width, int height, int step,
                         size_t frame_stride, int lap0, int lap1);
/* Copies the resident results of frame `frame` to host buffers. */
int viorb_fetch(viorb_extractor* ex, int frame, viorb_keypoint* kps, uint8_t* desc, int cap,
                int* n_out, int* mono_index_out);
/* Waits for all work queued on the handle's stream. */
int viorb_extractor_sync(viorb_extractor* ex);
/* The handle's CUDA stream (a cudaStream_t as void*), and a way to run the handle on a caller-owned
 * stream instead (so that a host framework can time or order the work with its own events). */
void* viorb_extractor_stream(viorb_extractor* ex);
int viorb_extractor_set_stream(viorb_extractor* ex, void* stream);
/* Stage timing with CUDA events on the handle's stream (the reference's REGISTER_TIMES spans,
 * frame.cpp:103-116,124-132).  stage_ms: ms[7] = resize, fast, octree, orient, emit, blur, describe of
 * the last extract; viorb_stereo_stage_ms: ms[5] = coarse, refine, cut, knn2, ratio of the last
 * viorb_stereo_batch (left handle). */
int viorb_extractor_set_timing(viorb_extractor* ex, int enable);
int viorb_extractor_stage_ms(viorb_extractor* ex, float* ms);
int viorb_stereo_stage_ms(viorb_extractor* left, float* ms);

/* mvImagePyramid[level] of the last compute (fextractor.h:64; read by Frame::ComputeStereoMatches,
 * frame.cpp:830,920,937).  Copies the level image (without the 19-px border, which this path never
 * reads) into dst (row step dst_step); reports its size. dst may be NULL to query the size. */
int viorb_pyramid_level(viorb_extractor* ex, int frame, int level, uint8_t* dst, int dst_step,
                        int* width, int* height);

/* Stage inspection for parity tests (not used by the reference-facing path):
 * FAST candidates of one level = vToDistributeKeys (fextractor.cpp:756-820), as (x, y, score)
 * triples relative to the 16-px border, sorted in the reference's cell-major raster order;
 * retained keypoints after DistributeOctTree + IC_Angle = (x, y, score) in level pixels + angle. */
int viorb_debug_candidates(viorb_extractor* ex, int frame, int level, int32_t* xys, int cap, int* n_out);
int viorb_debug_kept(viorb_extractor* ex, int frame, int level, int32_t* xys, float* angle, int cap, int* n_out);
int viorb_debug_blurred(viorb_extractor* ex, int frame, int level, uint8_t* dst, int dst_step);
/* Test hook of the kNN-2 launcher: force the number of train-range splits per query tile (0 = the cost model decides,
 * the default).  Process-wide; only the parity tests set it, to reach the split / merge path on small inputs. */
void viorb_debug_knn_split(int n_split);

/* ---- matching ---------------------------------------------------------------------------- */

/* FMatcher::DescriptorDistance (src/geometry/fmatcher.cpp:2859-2875) for n row pairs (host buffers). */
int viorb_descriptor_distance(const uint8_t* a, const uint8_t* b, int n, int32_t* dist, int device);

/* cv::BFMatcher(NORM_HAMMING).knnMatch(q, t, matches, 2) (frame.cpp:1167, fmatcher.cpp:214):
 * idx[nq][2], dist[nq][2]; ascending distance, ties -> lower train index; -1 where nt < 2. Host buffers. */
int viorb_knn2(const uint8_t* q, int nq, const uint8_t* t, int nt, int32_t* idx, int32_t* dist, int device);
/* Same with device buffers, asynchronous on `stream` (a cudaStream_t cast to void*; NULL = default). */
int viorb_knn2_device(const uint8_t* d_q, int nq, const uint8_t* d_t, int nt, int32_t* d_idx, int32_t* d_dist,
                      void* stream);

/* viorb_knn2_device without ANY allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same inputs, outputs and results as viorb_knn2_device.  The scratch rows of the kernel's train-range split live in a
 * file-scope device array of 4 MiB that every GPU loading the library holds until the process exits; the call makes no
 * cudaMalloc / cudaFree / pool allocation (the first call on a device does one-time set-up: make it before a capture).
 * Limit: nq <= VIORB_KNN2_STATIC_MAX_QUERIES rows per call, larger calls fail with VIORB_E_CAPACITY and a message (nt is
 * bounded by 2^22 - 1 as everywhere).  All calls share the array: calls that overlap on the GPU (different streams, or
 * several host threads) race - serialise them on one stream.  Nothing in the array needs clearing between calls. */
#define VIORB_KNN2_STATIC_MAX_QUERIES 16384
int viorb_knn2_device_static(const uint8_t* d_q, int nq, const uint8_t* d_t, int nt, int32_t* d_idx, int32_t* d_dist,
                             void* stream);
/* Same, one 16-byte row (idx0, idx1, dist0, dist1) per query at d_rows[nq][4]: the row format of the sharded matcher. */
int viorb_knn2_packed_device(const uint8_t* d_q, int nq, const uint8_t* d_t, int nt, int32_t* d_rows, void* stream);

/* ---- row-sharded brute-force matcher over the GPUs of one box (BASELINE.json configs[4]) -------------
 * The reference has no multi-GPU path; this is the partitioning north_star asks for: query rows in contiguous blocks of
 * rows_per_rank = ceil(nq_total / world), train set replicated, one all-gather of 16-byte result rows.
 * A viorb_comm is one rank's end of the exchange:
 *   - viorb_comm_create_nccl: rank 0 calls viorb_comm_unique_id, the host distributes the 128 bytes by its own means,
 *     every rank then creates the communicator (ncclCommInitRank; collective).  libnccl.so.2 is resolved at run time.
 *   - viorb_comm_adopt_nccl: wrap an ncclComm_t the host application already owns (not destroyed with the viorb_comm).
 *   - viorb_comm_create_callback: the host supplies the in-place all-gather itself (d_buf holds world blocks of
 *     bytes_per_rank; block `rank` is filled; must be stream-ordered or complete on return; 0 = success). */
typedef struct viorb_comm viorb_comm;
typedef int (*viorb_allgather_fn)(void* user, void* d_buf, size_t bytes_per_rank, int rank, int world, void* stream);
int viorb_comm_unique_id(uint8_t id[128]);
int viorb_comm_create_nccl(const uint8_t id[128], int rank, int world, int device, viorb_comm** out);
int viorb_comm_adopt_nccl(void* nccl_comm, int rank, int world, int device, viorb_comm** out);
int viorb_comm_create_callback(viorb_allgather_fn fn, void* user, int rank, int world, int device, viorb_comm** out);
void viorb_comm_destroy(viorb_comm* comm);
int viorb_comm_rank(const viorb_comm* comm);
int viorb_comm_world(const viorb_comm* comm);
/* The partition every rank must use: rank owns query rows [row0, row0 + rows); rows_per_rank = block size of the result. */
void viorb_shard_rows(int nq_total, int world, int rank, int* row0, int* rows, int* rows_per_rank);
/* cv::BFMatcher(NORM_HAMMING).knnMatch(q, t, 2) (frame.cpp:1167) with q sharded by rows.  d_q_local: THIS rank's
 * `rows` query rows; d_t: all nt train rows; d_out: int32[world * rows_per_rank][4] on the device, complete on every
 * rank when the stream reaches the end of the call; row r of the full query set is d_out[r] (rows past nq_total: -1).
 * Collective: every rank of the communicator must call it with the same nq_total / nt. */
int viorb_knn2_sharded(viorb_comm* comm, const uint8_t* d_q_local, int nq_total, const uint8_t* d_t, int nt,
                       int32_t* d_out, void* stream);

/* kNN-2 followed by the ratio test, accepted matches in query order.
 * rule 0: frame.cpp:1170   d0 < d1 * ratio (float * double; ratio 0.7 there);
 * rule 1: fmatcher.cpp:220-223 (matchFeatures method 2) (double)d0 < ratio * d1, then with
 *         dedup != 0 removeDuplicatedMatches (fmatcher.cpp:247-266; stable, first of each trainIdx). */
int viorb_match_ratio(const uint8_t* q, int nq, const uint8_t* t, int nt, double ratio, int rule, int dedup,
                      viorb_match* out, int cap, int* n_out, int device);

/* Frame::ComputeStereoMatches (src/datastructures/frame.cpp:823-997) for frame `frame` of two
 * extractors holding the left/right results of the same extract call index.  bf = mbf, b = mb.
 * u_right/depth: [n_left] (= mvuRight / mvDepth, -1 where unmatched). best_idx/best_dist may be NULL
 * (coarse row-band Hamming winner per left keypoint, -1 if none < TH_HIGH). Host output buffers. */
int viorb_stereo_match(viorb_extractor* left, viorb_extractor* right, int frame, float bf, float b,
                       float* u_right, float* depth, int32_t* best_idx, int32_t* best_dist, int cap, int* n_left);

/* Batched, device-resident stereo front-end for frames [0, n) of the last extract calls:
 * ComputeStereoMatches + ComputeStereoFishEyeMatches' kNN-2/ratio (frame.cpp:1167-1170) left->right.
 * Results stay on the device; *_fetch copies one frame's rows out. */
int viorb_stereo_batch(viorb_extractor* left, viorb_extractor* right, int n, float bf, float b, double ratio);
int viorb_stereo_fetch(viorb_extractor* left, int frame, float* u_right, float* depth, int32_t* lr_match,
                       int cap, int* n_left, int* n_stereo, int* n_ratio);
/* Whole-batch read-back: arrays are [n][cap] / [n], cap >= viorb_extractor_capacity(). */
int viorb_stereo_fetch_all(viorb_extractor* left, int n, float* u_right, float* depth, int32_t* lr_match,
                           int cap, int* n_left, int* n_stereo, int* n_ratio);

/* Frame-to-frame brute-force kNN-2 + ratio between the resident descriptors of two extractors'
 * frames (cur.frame_q -> prev.frame_t); result rows in query order. */
int viorb_match_resident(viorb_extractor* cur, int frame_q, viorb_extractor* prev, int frame_t, double ratio,
                         int rule, viorb_match* out, int cap, int* n_out);

/* ---- windowed (grid) matchers: SURVEY.md section 8 row a-11 ------------------------------------ */

/* Frame::mnMinX, mnMaxX, mnMinY, mnMaxY (src/datastructures/frame.cpp:133-145).  The cell-size inverses
 * mfGridElementWidthInv / mfGridElementHeightInv are derived as 64 / (max_x - min_x), 48 / (max_y - min_y)
 * (frame.cpp:235-236, FRAME_GRID_COLS / FRAME_GRID_ROWS of frame.h:42-43). */
typedef struct viorb_grid_bounds { float min_x, max_x, min_y, max_y; } viorb_grid_bounds;

/* Frame::AssignFeaturesToGrid + Frame::GetFeaturesInArea (frame.cpp:386-414, 678-757) for one query:
 * indices of the keypoints inside the square of half size r around (x, y) with min_level <= octave <=
 * max_level (the reference's "-1 = no bound" convention), in the reference's order.  Host buffers. */
int viorb_features_in_area(const viorb_keypoint* kps, int n, const viorb_grid_bounds* bounds, float x, float y, float r,
                           int min_level, int max_level, int32_t* out, int cap, int* n_out, int device);

/* FMatcher::SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize) (src/geometry/fmatcher.cpp:983-1098)
 * for npairs frame pairs.  Pair p: F1 = rows [p*cap1, p*cap1 + n1[p]) of kps1 (F1.ukeypoints_) / desc1, F2 likewise
 * with cap2 / n2 (< 65536 keypoints).  prev_matched [npairs][cap1][2] is vbPrevMatched (in/out), matches12
 * [npairs][cap1] is vnMatches12, nmatches[p] the return value.  nn_ratio = mfNNratio, check_orientation =
 * mbCheckOrientation (FMatcher ctor).  bounds: F2's grid.  Host buffers; npairs = 1, cap = n for one pair. */
int viorb_search_for_initialization(int npairs, int cap1, int cap2, const viorb_keypoint* kps1, const uint8_t* desc1,
                                    const int32_t* n1, const viorb_keypoint* kps2, const uint8_t* desc2, const int32_t* n2,
                                    const viorb_grid_bounds* bounds, float* prev_matched, int window_size, float nn_ratio,
                                    int check_orientation, int32_t* matches12, int32_t* nmatches, int device);

/* FMatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono) (fmatcher.cpp:2471-2688), both branches.
 * Pose, map points and the camera model stay with the caller, which passes per LastFrame keypoint i (i runs over
 * LastFrame.N, i.e. keypoints_ then keypointsRight_ for a two-camera LastFrame):
 * uv[i][2] = mpCamera->project(Rcw*x3Dw+tcw), invz[i], last_octave[i], last_angle[i] (kpLF.angle), desc_mp[i][32] =
 * pMP->GetDescriptor(), valid[i] = map point present && !mvbOutlier && invz >= 0 && uv inside [mnMinX,mnMaxX]x[mnMinY,mnMaxY]
 * (fmatcher.cpp:2496-2517), observed[i] = pMP->Observations() > 0.  Current frame: kps2 / desc2 / u_right2 (mvuRight, NULL
 * for monocular) / blocked2 (mvpMapPoints[i2] && Observations() > 0; NULL = none).  direction 0: octave window
 * nLastOctave-1..+1, 1: bForward, 2: bBackward.
 * Two-camera rigs (CurrentFrame.Nleft != -1, fmatcher.cpp:2593-2660): uv_right[i][2] = mpCamera->project(mTrl * x3Dc) and
 * nleft2[p] = CurrentFrame.Nleft; kps2 / desc2 / blocked2 / assigned then hold keypoints_ followed by keypointsRight_
 * (n2 = N) and u_right2 must be NULL.  uv_right == NULL: the single-camera branch.
 * assigned [npairs][cap2]: -1 untouched, -2 set to NULL by the rotation check, i >= 0: CurrentFrame.mvpMapPoints[i2] =
 * LastFrame.mvpMapPoints[i].  nmatches[p] = return value. */
int viorb_search_by_projection(int npairs, int cap1, int cap2, const float* uv, const float* invz, const int32_t* last_octave,
                               const float* last_angle, const uint8_t* desc_mp, const uint8_t* valid, const uint8_t* observed,
                               const int32_t* n1, const viorb_keypoint* kps2, const uint8_t* desc2, const float* u_right2,
                               const uint8_t* blocked2, const int32_t* n2, const viorb_grid_bounds* bounds,
                               const float* scale_factors, int nlevels, float th, float bf, int direction,
                               int check_orientation, const float* uv_right, const int32_t* nleft2, int32_t* assigned,
                               int32_t* nmatches, int device);

/* Frame-to-frame matching between CONSECUTIVE resident frames of the extractor's last batch (the monocular path,
 * BASELINE.json configs[0]): for p in [0, n - 1), F1 = frame p, F2 = frame p + 1:
 *   FMatcher::SearchForInitialization(F1, F2, vbPrevMatched = F1 keypoint positions (src/core/tracking.cpp:2286-2288),
 *   vnMatches12, window_size) (fmatcher.cpp:983-1098), and brute-force kNN-2 + ratio test F1 -> F2 (rule as in
 *   viorb_match_ratio; matchFeatures method 2, fmatcher.cpp:204-229).  bounds: the frames' grid (NULL = (0, W, 0, H),
 *   ComputeImageBounds without distortion).  Asynchronous on the handle's stream; results stay on the device. */
int viorb_f2f_batch(viorb_extractor* ex, int n, const viorb_grid_bounds* bounds, int window_size, float nn_ratio,
                    int check_orientation, double ratio, int rule);
/* Read-back: matches12 [npairs][cap] = vnMatches12, nmatches[p] = SearchForInitialization's return value, knn_match
 * [npairs][cap] = accepted train index per F1 keypoint or -1, n_ratio[p] = accepted count.  cap >= viorb_extractor_capacity().
 * VIORB_E_CAPACITY if a pair's candidate lists exceeded the handle's budget (that pair reports no window matches). */
int viorb_f2f_fetch_all(viorb_extractor* ex, int npairs, int32_t* matches12, int32_t* nmatches, int32_t* knn_match,
                        int32_t* n_ratio, int cap);

/* ---- keypoint undistortion: SURVEY.md section 8 row f-2 ----------------------------------------- */

/* Camera of Frame::UndistortKeyPoints / ComputeImageBounds (src/datastructures/frame.cpp:758-791, 793-821):
 * cam = fx fy cx cy of mK, dist = mDistCoef = k1 k2 p1 p2 [k3] (ndist 4 or 5), both CV_32F in the reference
 * (src/core/tracking.cpp:490-593).  The computation is cv::undistortPoints(pts, pts, mK, mDistCoef, Mat(), mK) with
 * OpenCV's default 5 iterations, in double; k1 == 0 copies the keypoints (frame.cpp:762-766).
 *
 * viorb_extractor_set_camera: from the next extract on, the extractor's epilogue also writes ukeypoints_ (same order
 * as the keypoints) and keeps them resident; read them with viorb_fetch_undistorted.  dist == NULL switches it off. */
int viorb_extractor_set_camera(viorb_extractor* ex, const float* cam, const float* dist, int ndist);
int viorb_fetch_undistorted(viorb_extractor* ex, int frame, viorb_keypoint* kps_un, int cap, int* n_out);
/* Host arrays: keypoints_ -> ukeypoints_ (every field but pt copied), and bare points xy [n][2] -> out [n][2]. */
int viorb_undistort_keypoints(const viorb_keypoint* kps, int n, const float* cam, const float* dist, int ndist,
                              viorb_keypoint* out, int device);
int viorb_undistort_points(const float* xy, int n, const float* cam, const float* dist, int ndist, float* out, int device);
/* Frame::ComputeImageBounds(imLeft): the undistorted image corners -> mnMinX, mnMaxX, mnMinY, mnMaxY
 * ((0, cols, 0, rows) when k1 == 0). */
int viorb_image_bounds(int cols, int rows, const float* cam, const float* dist, int ndist, viorb_grid_bounds* bounds,
                       int device);

/* ---- bag of words: SURVEY.md section 8 row f-3 --------------------------------------------------- */

/* The vocabulary tree of DBoW3::Vocabulary (thirdparty/DBoW3/DBoW3/src/Vocabulary.h:323-355: m_nodes, m_L, m_weighting,
 * m_scoring) as flat arrays; node 0 is the root.  children[child_start[i] .. child_start[i+1]) = m_nodes[i].children in
 * their stored order (the order decides ties), descriptors[i][32] = m_nodes[i].descriptor (CV_8U, 32 bytes), weight[i] =
 * m_nodes[i].weight, word_id[i] = m_nodes[i].word_id for leaves (-1 otherwise).  weighting: 0 TF_IDF, 1 TF, 2 IDF,
 * 3 BINARY; scoring: 0 L1_NORM, 1 L2_NORM, 2 CHI_SQUARE, 3 KL, 4 BHATTACHARYYA, 5 DOT_PRODUCT (BowVector.h:38-55).
 * The arrays are copied to the device; the caller may free them after the call. */
typedef struct viorb_vocab_desc {
    int32_t nnodes, levels, weighting, scoring;
    const int32_t* child_start;
    const int32_t* children;
    const uint8_t* descriptors;
    const double* weight;
    const int32_t* word_id;
} viorb_vocab_desc;
typedef struct viorb_vocabulary viorb_vocabulary;
int viorb_vocabulary_create(const viorb_vocab_desc* desc, int device, viorb_vocabulary** out);
void viorb_vocabulary_destroy(viorb_vocabulary* voc);

/* Frame::ComputeBoW (src/datastructures/frame.cpp:455-462) = Vocabulary::transform(features, mBowVec, mFeatVec, levelsup)
 * (Vocabulary.cpp:754-820) for nframes descriptor sets: set f = rows [f*cap, f*cap + n[f]) of desc (host, 32 bytes per
 * row, cap <= 16384).  Per set, with row stride cap (cap + 1 for fv_start):
 *   mBowVec  = { bow_word[j] -> bow_value[j] : j < nbow[f] } in std::map (ascending word id) order, doubles bit-identical;
 *   mFeatVec = { fv_node[j] -> fv_feature[fv_start[j] .. fv_start[j+1]) : j < nfv[f] } in ascending node id order, feature
 *              indices ascending inside a node.
 * A leaf above level levels - levelsup leaves the node id unset in the reference; 0 is reported here. */
int viorb_bow_transform(const viorb_vocabulary* voc, int nframes, int cap, const uint8_t* desc, const int32_t* n, int levelsup,
                        uint32_t* bow_word, double* bow_value, int32_t* nbow, uint32_t* fv_node, int32_t* fv_start,
                        uint32_t* fv_feature, int32_t* nfv);
/* The same on the descriptors the extractor still holds on the device for its last batch (queued on the handle's stream,
 * no host pass); viorb_bow_fetch copies one frame's result (buffers of viorb_extractor_capacity entries, + 1 for fv_start). */
int viorb_bow_resident(viorb_extractor* ex, const viorb_vocabulary* voc, int levelsup);
int viorb_bow_fetch(viorb_extractor* ex, int frame, uint32_t* bow_word, double* bow_value, int32_t* nbow, uint32_t* fv_node,
                    int32_t* fv_start, uint32_t* fv_feature, int32_t* nfv);

/* ---- fisheye stereo epilogue: SURVEY.md section 8 row f-4 ---------------------------------------- */

/* The two KannalaBrandt8 cameras of a fisheye stereo rig: cam1 / cam2 = Camera::mvParameters (fx fy cx cy k0 k1 k2 k3) of
 * mpCamera / mpCamera2, R12 (row-major) / t12 = Frame::mRlr / mtlr, precision = KannalaBrandt8::precision (1e-6,
 * include/vi_slam/geometry/cameramodels/kannalabrandt8.h:29). */
typedef struct viorb_kb8_rig { float cam1[8], cam2[8], R12[9], t12[3], precision; } viorb_kb8_rig;

/* KannalaBrandt8::TriangulateMatches (src/geometry/cameramodels/kannalabrandt8.cpp:324-393) for n keypoint pairs:
 * pt1 / pt2 [n][2] = kp1.pt / kp2.pt, sigma1 / sigma2 [n] = sigmaLevel / unc -> depth[n] = the return value (-1 when the
 * parallax, depth or reprojection checks reject the pair), p3d [n][3] = p3D (0 when rejected).  Float: 1e-3 relative. */
int viorb_kb8_triangulate_matches(const viorb_kb8_rig* rig, const float* pt1, const float* pt2, const float* sigma1,
                                  const float* sigma2, int n, float* depth, float* p3d, int device);

/* Frame::ComputeStereoFishEyeMatches (src/datastructures/frame.cpp:1148-1189): brute-force kNN-2 between the lapping-area
 * descriptors (rows mono_left.. of the left frame, mono_right.. of the right), the 0.7 ratio rule, TriangulateMatches with
 * level_sigma2[octave] (mvLevelSigma2), depth > 0.0001.  Outputs: left_to_right [n_left] = mvLeftToRightMatch,
 * right_to_left [n_right] = mvRightToLeftMatch, depth [n_left] = mvDepth (-1), p3d [n_left][3] = mvStereo3Dpoints (valid
 * where depth > 0), n_desc_matches = descMatches, n_matches = nMatches.  Host buffers. */
int viorb_fisheye_stereo_matches(const viorb_kb8_rig* rig, const viorb_keypoint* kps_left, const uint8_t* desc_left, int n_left,
                                 int mono_left, const viorb_keypoint* kps_right, const uint8_t* desc_right, int n_right,
                                 int mono_right, const float* level_sigma2, int nlevels, int32_t* left_to_right,
                                 int32_t* right_to_left, float* depth, float* p3d, int32_t* n_desc_matches, int32_t* n_matches,
                                 int device);

#ifdef __cplusplus
}
#endif
#endif /* VIORB_H_ */
