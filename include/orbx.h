/*
 * orbx.h — C ABI of liborbx: B200-native (sm_100a) ORB extraction + brute-force Hamming matching.
 *
 * Drop-in boundary for the `feature` module of MostlyKIGuess/slam-rs.  The reference has no FFI of
 * its own on this path: the seam is the Rust API of src/feature/{detector,matcher}.rs, which today
 * forwards to OpenCV's cv::ORB / cv::BFMatcher through the `opencv` crate.  Every entry point below
 * names the reference call it replaces (paths are relative to the reference root).
 *
 * Conventions
 *  - every `int` return: ORBX_OK (0) or a negative ORBX_E_* code; orbx_last_error() gives the text
 *    (thread-local).  Empty inputs are not errors (matcher.rs:29-31, detector.rs:58-67).
 *  - handles are not re-entrant (mirrors `&mut self` on every reference method); distinct handles are
 *    independent and own their CUDA stream and device scratch.
 *  - image / descriptor / result pointers may be HOST or DEVICE memory (detected with
 *    cudaPointerGetAttributes).  Host pointers: the call copies in/out and returns when results are
 *    ready.  Device pointers: work is enqueued on the handle's stream and the call returns without
 *    synchronising (use orbx_*_sync).  Count outputs (`int*`) follow the same rule.
 *  - descriptors are u8 rows of exactly 32 bytes, row pitch 32 (cv::Mat K x 32 CV_8U, contiguous).
 *  - there is no CPU fallback: without a CUDA device every create call fails with ORBX_E_CUDA.
 */
#ifndef ORBX_H
#define ORBX_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORBX_OK 0
#define ORBX_E_BADARG (-1)
#define ORBX_E_CAPACITY (-2) /* result larger than the caller's buffer; count outputs hold the needed size */
#define ORBX_E_CUDA (-3)
#define ORBX_E_NOMEM (-4)
#define ORBX_E_DATA (-5) /* valid call, unusable data: the reference's Err(String) cases (pose.rs:52-59,75-77,104-106) */

#define ORBX_HARRIS_SCORE 0

/* == cv::KeyPoint (28 B): consumed at src/odometry/pose.rs:36-41 (`kp.get(i).pt()`) */
typedef struct {
  float x, y;      /* level coordinates x level scale (fp32) */
  float size;      /* 31 * level scale */
  float angle;     /* degrees, [0,360) */
  float response;  /* Harris response */
  int32_t octave;  /* pyramid level */
  int32_t class_id; /* -1 */
} orbx_keypoint;

/* == cv::DMatch (16 B): consumed at src/odometry/pose.rs:36-41, src/mapping/map.rs:111-115 */
typedef struct {
  int32_t query_idx, train_idx, img_idx; /* img_idx = 0 */
  float distance;                        /* (float) Hamming distance */
} orbx_dmatch;

/* cv::ORB::create parameters.  orbx_default_params() = ORB::create_def() (detector.rs:16):
 * 500, 1.2f, 8, 31, 0, 2, HARRIS_SCORE, 31, 20.  Only nfeatures and fast_threshold may differ
 * from the defaults in this version (others -> ORBX_E_BADARG). */
typedef struct {
  int32_t nfeatures;
  float scale_factor;
  int32_t nlevels, edge_threshold, first_level, wta_k, score_type, patch_size, fast_threshold;
} orbx_params;

typedef struct orbx_detector orbx_detector;
typedef struct orbx_matcher orbx_matcher;

void orbx_default_params(orbx_params* p);

/* ------------------------------------------------------------------ OrbDetector
 * replaces OrbDetector::new (detector.rs:15-18).  Scratch is sized for images up to max_w x max_h
 * and max_batch frames per call.  `params` may be NULL (defaults). */
int orbx_detector_create(const orbx_params* params, int device, int max_w, int max_h, int max_batch,
                         orbx_detector** out);
void orbx_detector_destroy(orbx_detector* d);
/* replaces OrbDetector::max_features (detector.rs:42-44) */
int orbx_detector_max_features(const orbx_detector* d);

/* replaces OrbDetector::detect (detector.rs:20-24): S1-S7, 8-bit single channel image, no mask.
 * kps: cap entries.  *n = number of keypoints (if > cap: ORBX_E_CAPACITY, first cap are written). */
int orbx_detect(orbx_detector* d, const uint8_t* img, int w, int h, size_t pitch, orbx_keypoint* kps,
                int cap, int* n);
/* replaces OrbDetector::detect_and_compute (detector.rs:26-40): S1-S9.  desc: cap x 32 bytes. */
int orbx_detect_and_compute(orbx_detector* d, const uint8_t* img, int w, int h, size_t pitch,
                            orbx_keypoint* kps, uint8_t* desc, int cap, int* n);
/* the same over `batch` equally sized frames (frame i at imgs + i*frame_stride): the data-parallel
 * form of the per-frame loop at examples/visual_odometry.rs:84-218.  kps: batch x cap, desc: batch x
 * cap x 32 (may be NULL = detect only), n_per_frame: batch.  Keypoint order inside a frame: level
 * major, then response descending, then (y, x) — OpenCV's intra-level order is a std::nth_element
 * artefact, so parity is on sets keyed by (octave, x, y). */
int orbx_detect_and_compute_batch(orbx_detector* d, const uint8_t* imgs, int w, int h, size_t pitch,
                                  size_t frame_stride, int batch, orbx_keypoint* kps, uint8_t* desc,
                                  int cap_per_frame, int* n_per_frame);
/* the same for interleaved 8UC3 BGR frames (pitch in bytes >= 3*w): the frame ingest in front of S1.  Replaces
 * imgproc::cvt_color(COLOR_BGR2GRAY) at examples/visual_odometry.rs:93-99 / src/main.rs:43-49 and the implicit
 * conversion cv::ORB applies to non-8UC1 input (detector.rs:22,32): gray = (B*3735 + G*19235 + R*9798 + 16384) >> 15,
 * computed on the device straight into the pitched level-0 plane. */
int orbx_detect_and_compute_batch_bgr(orbx_detector* d, const uint8_t* imgs, int w, int h, size_t pitch,
                                      size_t frame_stride, int batch, orbx_keypoint* kps, uint8_t* desc,
                                      int cap_per_frame, int* n_per_frame);
/* waits for the handle's stream; returns ORBX_E_CAPACITY (once) if an asynchronous call since the last sync had to
 * truncate a frame to cap rows (n is clamped to cap on that path) or exhausted the tie arena */
int orbx_detector_sync(orbx_detector* d);
/* run on a caller-owned CUDA stream (cudaStream_t as void*); NULL restores the handle's own stream */
int orbx_detector_set_stream(orbx_detector* d, void* stream);
/* enable=1: single-frame calls with host buffers (orbx_detect / orbx_detect_and_compute - the shape of every reference
 * call site, examples/visual_odometry.rs:102) replay a CUDA graph captured once per (w, h, cap): one graph launch instead
 * of ~13 dependent kernel launches, three copies and a memset.  Results are byte-identical to the stream path. */
int orbx_detector_set_graph(orbx_detector* d, int enable);

/* Stage access for parity tests and profiling (device buffers owned by the handle, valid until
 * the next call): level geometry and pointers of the pyramid (blurred=0) or of the blurred
 * pyramid (blurred=1) of the last batch. */
int orbx_detector_level_info(const orbx_detector* d, int level, int blurred, int* w, int* h, size_t* pitch,
                             size_t* frame_stride, const uint8_t** dev_ptr);
/* copies level `level` of frame `frame` of the last batch to host memory (rows of `w` bytes at out_pitch) */
int orbx_detector_read_level(orbx_detector* d, int level, int blurred, int frame, uint8_t* host_out,
                             size_t out_pitch);
/* FAST+NMS candidates of the last batch: copies up to cap packed (score<<24|y<<12|x) entries of
 * (frame, level) to host memory in unspecified order; *n = count. */
int orbx_detector_candidates(orbx_detector* d, int frame, int level, uint32_t* out, int cap, int* n);

/* Per-stage device timing (CUDA events on the handle's stream).  enable=1 makes every later call
 * record events around each stage; orbx_detector_timing() synchronises and returns accumulated
 * milliseconds and launch counts per stage since the last reset. */
#define ORBX_STAGE_PYRAMID 0
#define ORBX_STAGE_FAST 1
#define ORBX_STAGE_HARRIS 2
#define ORBX_STAGE_SELECT 3
#define ORBX_STAGE_ORIENT 4
#define ORBX_STAGE_BLUR 5
#define ORBX_STAGE_BRIEF 6
#define ORBX_STAGE_COUNT 7
int orbx_detector_set_timing(orbx_detector* d, int enable);
int orbx_detector_timing(orbx_detector* d, float* ms_per_stage, int* launches_per_stage, int reset);

/* ------------------------------------------------------------------ FeatureMatcher
 * replaces FeatureMatcher::new (matcher.rs:15-21) = BFMatcher::create(NORM_HAMMING, crossCheck).
 * The reference always passes crossCheck=false. */
int orbx_matcher_create(int device, int cross_check, orbx_matcher** out);
void orbx_matcher_destroy(orbx_matcher* m);
int orbx_matcher_sync(orbx_matcher* m);
int orbx_matcher_set_stream(orbx_matcher* m, void* stream);

/* replaces FeatureMatcher::match_descriptors (matcher.rs:24-37) = DescriptorMatcher::match, k=1:
 * for each query row the train row of minimum Hamming distance, ties -> lowest train index.
 * out: nq entries in query order; *n = nq (0 when either side is empty).  With cross_check=1 only
 * mutual first-arg-min pairs survive, in ascending query_idx (cv::BFMatcher crossCheck). */
int orbx_match(orbx_matcher* m, const uint8_t* q, int nq, const uint8_t* t, int nt, orbx_dmatch* out,
               int* n);
/* BFMatcher::knnMatch(k=2) (north_star stage 5): out: nq x 2, n_per_q[i] = min(2, nt).  With
 * cross_check=1, n_per_q[i] is zeroed for queries failing the mutual-nearest test (OpenCV asserts
 * on this combination). */
int orbx_knn_match2(orbx_matcher* m, const uint8_t* q, int nq, const uint8_t* t, int nt,
                    orbx_dmatch* out, int* n_per_q);
/* Batched k=2 over `npairs` independent problems, device pointers only: problem p matches
 * q + p*q_stride (nq_dev[p] rows) against t + p*t_stride (nt_dev[p] rows); strides in bytes; counts
 * are device int arrays (clamped to max_rows).  out: npairs x max_rows x 2, n_per_q: npairs x max_rows.
 * This is the frame-to-frame loop of examples/visual_odometry.rs:106 over a whole sequence. */
int orbx_knn_match2_batch(orbx_matcher* m, const uint8_t* q, size_t q_stride, const int* nq_dev,
                          const uint8_t* t, size_t t_stride, const int* nt_dev, int npairs, int max_rows,
                          orbx_dmatch* out, int* n_per_q);
/* Lowe ratio test on a k=2 result: keeps knn2[2i] iff n_per_q[i] == 2 and d1 < ratio*d2 (fp32,
 * strict).  Host or device arrays; order preserved. */
int orbx_ratio_test(orbx_matcher* m, const orbx_dmatch* knn2, const int* n_per_q, int nq, float ratio,
                    orbx_dmatch* out, int* n);
/* batched device form of orbx_ratio_test: problem p has nq_dev[p] queries at knn2 + p*max_rows*2 */
int orbx_ratio_test_batch(orbx_matcher* m, const orbx_dmatch* knn2, const int* n_per_q, const int* nq_dev,
                          int npairs, int max_rows, float ratio, orbx_dmatch* out, int* n_out_dev);
/* replaces FeatureMatcher::filter_good_matches (matcher.rs:40-66): thr = max(ratio*min_dist, 30);
 * keep distance < thr (strict), order preserved.  Host or device arrays. */
int orbx_filter_good_matches(orbx_matcher* m, const orbx_dmatch* in, int n, float ratio, orbx_dmatch* out,
                             int* n_out);
/* batched device form: problem p has n_dev[p] matches at in + p*max_rows; out likewise. */
int orbx_filter_good_matches_batch(orbx_matcher* m, const orbx_dmatch* in, const int* n_dev, int npairs,
                                   int max_rows, float ratio, orbx_dmatch* out, int* n_out_dev);
/* k=1 batched (device pointers), same addressing as orbx_knn_match2_batch; out: npairs x max_rows */
int orbx_match_batch(orbx_matcher* m, const uint8_t* q, size_t q_stride, const int* nq_dev,
                     const uint8_t* t, size_t t_stride, const int* nt_dev, int npairs, int max_rows,
                     orbx_dmatch* out, int* n_out_dev);

/* BFMatcher::knnMatch(k=2) like orbx_knn_match2, without any allocation.  This entry point has NO counterpart in the
 * reference: it was added for callers that cannot call cudaMalloc (code under
 * CUDA-graph capture, tight serving loops, harnesses that forbid it).  No handle: the partial-key scratch of the
 * train-split search lives in a file-scope device array inside liborbx (8 MiB, held by every GPU that loads the library
 * until exit), so the call makes no cudaMalloc / cudaFree and creates no stream or event.  q (nq x 32), t (nt x 32), out
 * (nq x 2) and n_per_q (nq) must be DEVICE memory of `device`; the work is enqueued on `stream` (a cudaStream_t; NULL =
 * the default stream) and the call returns without synchronising.  Limits: nq <= 8192 and nt <= 64 x 65536 rows per call
 * (ORBX_E_CAPACITY beyond, with a clear message); no cross-check.  Results are bit-identical to orbx_knn_match2.  All
 * calls on one GPU share the one scratch array: calls whose kernels may overlap on the GPU (different streams) race -
 * serialise them.  The array keeps the previous call's contents; nothing in it has to start at zero (every slot a call
 * reads is written by that call first). */
int orbx_knn_match2_static(int device, const uint8_t* q, int nq, const uint8_t* t, int nt, orbx_dmatch* out, int* n_per_q,
                           void* stream);

/* replaces PoseEstimator::extract_matched_points (src/odometry/pose.rs:27-44; SURVEY.md §8 f2, first step):
 * pts1[i] = kp1[matches[i].query_idx].pt, pts2[i] = kp2[matches[i].train_idx].pt as (x, y) fp32 pairs, match
 * order preserved.  An index outside its keypoint array is the reference's `Vector::get(..)?` error:
 * ORBX_E_BADARG (outputs unspecified).  Host or device arrays. */
int orbx_extract_matched_points(orbx_matcher* m, const orbx_keypoint* kp1, int n1, const orbx_keypoint* kp2, int n2,
                                const orbx_dmatch* matches, int n, float* pts1, float* pts2);
/* batched device form (sequence front end): problem p reads kp1 + p*kp_stride (n1_dev[p] keypoints), kp2 +
 * p*kp_stride (n2_dev[p]) and matches + p*max_rows (n_dev[p]); writes pts1 / pts2 + p*max_rows*2 floats;
 * *bad_dev becomes non-zero if any index is out of range (the caller clears it).  Asynchronous. */
int orbx_extract_matched_points_batch(orbx_matcher* m, const orbx_keypoint* kp1, const int* n1_dev,
                                      const orbx_keypoint* kp2, const int* n2_dev, size_t kp_stride,
                                      const orbx_dmatch* matches, const int* n_dev, int npairs, int max_rows,
                                      float* pts1, float* pts2, int* bad_dev);

/* Which kernel runs the Hamming search: 0 = automatic (default: the tcgen05 int8 tensor-core kernel, csrc/match_mma.cu,
 * for problems of >= 2^27 descriptor pairs per call, the XOR+POPC kernel below that), 1 = POPC kernel only, 2 =
 * tensor-core kernel always.  Results are bit-identical; the environment variable ORBX_MATCHER=popc|mma sets the
 * default of new handles. */
int orbx_matcher_set_mode(orbx_matcher* m, int mode);
/* matcher timing: accumulated ms / launches of the Hamming search (expansion + kernel + merge) */
int orbx_matcher_set_timing(orbx_matcher* m, int enable);
int orbx_matcher_timing(orbx_matcher* m, float* ms, int* launches, int reset);

/* ------------------------------------------------------------------ Map
 * Device-resident replacement of the hot part of mapping::Map (src/mapping/map.rs): positions, descriptors and
 * observation counts live in HBM; point id == slot index, handed out sequentially (map.rs:31-37).
 * R is the 3x3 rotation in ROW-major order (nalgebra stores column-major), t the translation. */
typedef struct orbx_map orbx_map;
/* Map::new(intrinsics) (map.rs:21-28) with a fixed capacity (total points ever added) */
int orbx_map_create(int device, int capacity, double fx, double fy, double cx, double cy, orbx_map** out);
void orbx_map_destroy(orbx_map* m);
/* Map::size (map.rs:45-47) / Map::clear (:147-150) */
int orbx_map_size(const orbx_map* m);
int orbx_map_clear(orbx_map* m);
/* Map::add_points (map.rs:31-37): xyz n x 3 doubles; desc n x 32 or NULL (MapPoint::new has no descriptor);
 * has_desc (host, n bytes) or NULL = every point has one; observations (n ints) or NULL = 1 each.
 * xyz/desc/observations may be host or device memory.  *first_id = id of the first new point. */
int orbx_map_add_points(orbx_map* m, const double* xyz, const uint8_t* desc, const uint8_t* has_desc,
                        const int* observations, int n, int64_t* first_id);
/* Map::project_point (map.rs:50-69) + the bounds test of find_matches (:90) for every slot (host outputs):
 * xy: used x 2 (0,0 when behind the camera); vis bit0 = in front and 0<=x<4000, 0<=y<3000; bit1 = alive;
 * bit2 = has a descriptor.  Parity hook for the tests. */
int orbx_map_project(orbx_map* m, const double* R, const double* t, double* xy, uint8_t* vis);
/* Map::find_matches (map.rs:72-118): cull by projection, gather the visible descriptors, k=1 Hamming search
 * with query = map rows, train = frame rows (map.rs:107), filter_good_matches(2.0) (:108).
 * pairs: cap x 2 ints (map_id, kp_idx) in ascending map_id (host or device); frame_desc host or device. */
int orbx_map_find_matches(orbx_map* m, orbx_matcher* matcher, const double* R, const double* t,
                          const uint8_t* frame_desc, int n_frame, int* pairs, int cap, int* n_pairs, int* n_visible);
/* Map::update_observations (map.rs:121-127): pairs n x 2 as returned above (host or device) */
int orbx_map_update_observations(orbx_map* m, const int* pairs, int n);
/* Map::prune_outliers (map.rs:130-137): drops points with observations < 2, *removed = how many */
int orbx_map_prune_outliers(orbx_map* m, int* removed);
/* number of Map::stable_points (map.rs:140-145) */
int orbx_map_stable_count(orbx_map* m, int* count);
/* test hook: observation counts and alive flags of the first n slots (host outputs, either may be NULL) */
int orbx_map_read_observations(orbx_map* m, int* obs, uint8_t* alive, int n);

/* ------------------------------------------------------------------ PoseEstimator
 * Device form of odometry::PoseEstimator::compute_essential_matrix / recover_pose (src/odometry/pose.rs:47-114):
 * cv::findEssentialMat(points1, points2, K, RANSAC, 0.999, 1.0, 1000, mask) and cv::recoverPose(E, points1, points2,
 * K, R, t, mask), reproduced down to the RANSAC index draws (cv::RNG seeded as OpenCV seeds it), so the winning
 * hypothesis is OpenCV's; see csrc/pose.cu.  Points are (x, y) fp32 pairs (orbx_extract_matched_points). */
typedef struct orbx_pose_estimator orbx_pose_estimator;
enum { ORBX_POSE_OK = 0, ORBX_POSE_INSUFFICIENT_POINTS = 1, ORBX_POSE_NO_MODEL = 2, ORBX_POSE_TOO_FEW_INLIERS = 3 };
/* PoseEstimator::new(intrinsics) (pose.rs:17-23; min_matches = 8).  max_points bounds n of one problem,
 * max_problems the batch of orbx_estimate_pose_batch. */
int orbx_pose_estimator_create(int device, int max_points, int max_problems, double fx, double fy, double cx, double cy,
                               orbx_pose_estimator** out);
void orbx_pose_estimator_destroy(orbx_pose_estimator* pe);
int orbx_pose_estimator_set_stream(orbx_pose_estimator* pe, void* cuda_stream);
int orbx_pose_estimator_sync(orbx_pose_estimator* pe);
/* compute_essential_matrix (pose.rs:47-80).  E: 9 doubles, row major, unit Frobenius norm (as OpenCV returns it);
 * mask (n bytes, optional): RANSAC inliers; *iterations (optional): RANSAC iterations run.  Fewer than 8 points is
 * the reference's "Insufficient points" error, no model its "Failed to compute essential matrix": ORBX_E_DATA.
 * pts / mask may be host or device memory. */
int orbx_compute_essential_matrix(orbx_pose_estimator* pe, const float* pts1, int n1, const float* pts2, int n2,
                                  double* E, uint8_t* mask, int* n_inliers, int* iterations);
/* recover_pose (pose.rs:83-114): R 9 doubles row major, t 3 doubles (unit), mask (optional) = points in front of both
 * cameras and nearer than 50 for the chosen (R, t); *n_good < 8 is the reference's "Too few inliers" (ORBX_E_DATA,
 * R and t are still written). */
int orbx_recover_pose(orbx_pose_estimator* pe, const double* E, const float* pts1, int n1, const float* pts2, int n2,
                      double* R, double* t, uint8_t* mask, int* n_good);
/* both steps for a batch of frame pairs (visual_odometry.rs:111-117 over a sequence), device pointers, asynchronous:
 * problem p reads pts1 / pts2 + p*max_rows*2 floats (n_dev[p] points) and writes E + 9p, R + 9p, t + 3p,
 * n_inliers[p], n_good[p], status[p] (ORBX_POSE_*), masks + p*max_rows (optional). */
int orbx_estimate_pose_batch(orbx_pose_estimator* pe, const float* pts1, const float* pts2, const int* n_dev, int npairs,
                             int max_rows, double* E, double* R, double* t, int* n_inliers, int* n_good, int* status,
                             uint8_t* ransac_mask, uint8_t* pose_mask);

/* ------------------------------------------------------------------ Triangulator
 * Device form of mapping::Triangulator (src/mapping/triangulation.rs:52-198).  Poses are world-to-camera
 * (R 3x3 ROW-major, t 3), points are (x, y) fp32 pairs as produced by orbx_extract_matched_points. */
typedef struct orbx_triangulator orbx_triangulator;
/* Triangulator::new(intrinsics) (triangulation.rs:62-68); max_points bounds n of one call */
int orbx_triangulator_create(int device, int max_points, double fx, double fy, double cx, double cy,
                             orbx_triangulator** out);
void orbx_triangulator_destroy(orbx_triangulator* t);
/* with_min_parallax / with_max_reproj_error (triangulation.rs:71-80): stored and returned, not applied — the
 * reference does not apply them either (:138 "skipping this check") */
int orbx_triangulator_set_thresholds(orbx_triangulator* t, double min_parallax_deg, double max_reproj_error);
int orbx_triangulator_get_thresholds(const orbx_triangulator* t, double* min_parallax_deg, double* max_reproj_error);
/* Triangulator::triangulate (triangulation.rs:92-158): P = K [R | t] (:161-183), cv::triangulatePoints (4x4 DLT
 * system per correspondence, right singular vector of the smallest singular value, rounded to fp32 like the
 * CV_32F result the reference reads), skip |w| < 1e-10 (:127), X = (x, y, z) / w in f64 (:131), keep points in
 * front of both cameras (:134-139).  Survivors keep their input order; ids[k] = input index i (:140);
 * has_desc[k] = 1 and out_desc row k = desc row i when i < desc_rows (:141-150), else 0 / zeros.
 * n1 != n2 is the reference's "Point arrays must have the same length" error (ORBX_E_BADARG); n == 0 returns 0
 * points.  points4d (n x 4 fp32, optional) receives the homogeneous points of ALL n inputs (sign is arbitrary,
 * as in any SVD) — parity hook.  Every array may be host or device memory; device outputs need cap >= n.
 * *n_out > cap -> ORBX_E_CAPACITY. */
int orbx_triangulate(orbx_triangulator* t, const double* R1, const double* t1, const double* R2, const double* t2,
                     const float* pts1, int n1, const float* pts2, int n2, const uint8_t* desc, int desc_rows,
                     int32_t* ids, double* xyz, uint8_t* out_desc, uint8_t* has_desc, float* points4d, int cap,
                     int* n_out);
/* batched device form (sequence front end: the pairs of a whole sequence, poses straight from
 * orbx_estimate_pose_batch): problem p triangulates n_dev[p] correspondences at pts1 / pts2 + p*max_rows*2 floats
 * between the poses (R1 + 9p, t1 + 3p) — NULL = identity / zero for every problem — and (R2 + 9p, t2 + 3p), and writes
 * the survivors' ids + p*max_rows, xyz + p*max_rows*3 and n_out[p].  Device pointers, asynchronous on the handle's
 * stream; npairs * max_rows <= max_points. */
int orbx_triangulate_batch(orbx_triangulator* t, const double* R1, const double* t1, const double* R2, const double* t2,
                           const float* pts1, const float* pts2, const int* n_dev, int npairs, int max_rows, int32_t* ids,
                           double* xyz, int* n_out);
int orbx_triangulator_set_stream(orbx_triangulator* t, void* cuda_stream);
/* transform_points_to_world (examples/point_cloud.rs:522-544), the step between triangulate and Map::add_points:
 * xyz[i] = R^T * xyz[i] + (-R^T * t), in place (host or device), nalgebra's accumulation order, unfused f64 */
int orbx_transform_points_to_world(orbx_triangulator* t, const double* R, const double* tv, double* xyz, int n);

/* ------------------------------------------------------------------ sequence front end
 * The per-frame loop of examples/visual_odometry.rs:84-218 (detect_and_compute :102, match_descriptors :106, match
 * filter :107) over a whole sequence, scheduled natively (csrc/sequence.cu): frames -> T x (keypoints, descriptors) ->
 * T-1 frame-to-frame problems, BFMatcher k=2 (+ cross-check) + Lowe ratio (north_star stage 5) or, after
 * orbx_sequence_set_reference_filter(), the reference's own k=1 + filter_good_matches.  Row capacity per frame is
 * orbx_sequence_capacity() (>= nfeatures: OpenCV keeps ties; a frame with more keypoints than that - a tie flood -
 * is truncated to the capacity and orbx_sequence_sync() returns ORBX_E_CAPACITY).  Result row t of `good` / `n_good`
 * describes the pair (t-1, t); row 0 is unused.  All run calls are asynchronous: orbx_sequence_sync() before reading
 * results. */
typedef struct orbx_sequence orbx_sequence;
/* chunk: frames per launch set of orbx_sequence_run_device; host_chunk: pipeline granularity of orbx_sequence_run_host */
int orbx_sequence_create(int device, int width, int height, int nfeatures, int max_frames, int chunk, int host_chunk,
                         float ratio, int cross_check, orbx_sequence** out);
void orbx_sequence_destroy(orbx_sequence* s);
int orbx_sequence_capacity(const orbx_sequence* s);
/* enable=1: the pairs are matched exactly as the reference's loop does - match_descriptors (k=1, no cross-check,
 * visual_odometry.rs:106) then filter_good_matches(matches, ratio) (:107, ratio 2.0 there) - instead of the
 * north_star's k=2 + cross-check + Lowe composition; `good` / `n_good` then hold that filter's survivors. */
int orbx_sequence_set_reference_filter(orbx_sequence* s, int enable, float ratio);
/* device-resident frames (8UC1, `pitch` bytes per row, frame t at frames_dev + t*frame_stride); results stay in HBM */
int orbx_sequence_run_device(orbx_sequence* s, const uint8_t* frames_dev, size_t pitch, size_t frame_stride, int T);
/* HOST frames, tightly packed [T][height][width] (pinned memory for copy / compute overlap), results to host arrays
 * (pinned likewise; any may be NULL): kps T x cap, desc T x cap x 32, n T, good T x cap, n_good T.  Upload, kernels
 * and download of consecutive chunks overlap on four streams.  Consecutive calls are pipelined into each other (the
 * device staging and result arrays exist twice and the calls alternate between them, so the next sequence's uploads and
 * kernels run under the tail of the last one; chunk-level events keep every device buffer safe): a host that streams many sequences (64 independent sequences, BASELINE configs[4]) alternates
 * two sets of host result arrays and calls orbx_sequence_wait_pass(s, 1) to consume sequence k while k+1 runs.  The
 * host arrays of a call belong to the library until that call's pass has finished. */
int orbx_sequence_run_host(orbx_sequence* s, const uint8_t* frames, int T, orbx_keypoint* kps, uint8_t* desc, int* n,
                           orbx_dmatch* good, int* n_good);
/* Blocks until the run call issued `back` calls before the most recent one has finished, downloads included (back = 0:
 * the most recent call; 0 <= back <= 2).  Unlike orbx_sequence_sync() it does not wait for later calls and does not
 * report truncation (ORBX_E_CAPACITY comes from orbx_sequence_sync()). */
int orbx_sequence_wait_pass(orbx_sequence* s, int back);
/* device pointers of the last pass's results (layout as above; orbx_sequence_run_host alternates between two sets of
 * device arrays, so ask again after every run call) */
int orbx_sequence_results(orbx_sequence* s, orbx_keypoint** kps, uint8_t** desc, int** n, orbx_dmatch** good,
                          int** n_good);
int orbx_sequence_sync(orbx_sequence* s);
/* Device-side timing of whole passes: with enable=1, CUDA events bracket everything the run calls queue on the handle's
 * four streams.  orbx_sequence_elapsed() waits for the last pass and returns the milliseconds from the start of the
 * first pass since the last reset to the end of the last one (first upload / first kernel -> last download; consecutive
 * run_host passes overlap at their seams as described above), and how many passes that covers. */
int orbx_sequence_set_timing(orbx_sequence* s, int enable);
int orbx_sequence_elapsed(orbx_sequence* s, float* ms, int* passes, int reset);

/* ------------------------------------------------------------------ multi-GPU: query-sharded map-to-frame matching
 * (BASELINE configs[3]; the call it serves is match_descriptors at src/mapping/map.rs:107 over a map
 * too large for one GPU's time budget).  One process drives `ndev` GPUs of one box: the train set is resident on every
 * GPU (orbx_comm_set_train: one ncclBroadcast over NVLink per map update), the query rows are cut into contiguous
 * slices, and every GPU's kernel stores its DMatch records straight into device `devices[0]`'s result buffer through
 * peer-mapped pointers - there is no gather collective.  Results are identical to orbx_match / orbx_knn_match2 on one
 * GPU (query_idx is the global row).  Listing the same device more than once makes "virtual ranks" on one GPU (no NCCL;
 * for tests). */
typedef struct orbx_comm orbx_comm;
/* devices: ndev CUDA device indices (NULL = 0..ndev-1); needs peer access from every device to devices[0] */
int orbx_comm_create(const int* devices, int ndev, orbx_comm** out);
void orbx_comm_destroy(orbx_comm* c);
int orbx_comm_size(const orbx_comm* c);
/* train: nt x 32 bytes, host memory or device memory of devices[0]; copied, so the caller may reuse it after
 * orbx_comm_sync().  Asynchronous. */
int orbx_comm_set_train(orbx_comm* c, const uint8_t* train, int nt);
/* BFMatcher.knnMatch(k=2) of nq query rows (host memory or device memory of devices[0]) against the resident train
 * set: out nq x 2, n_per_q nq (host, or device memory of devices[0]).  Host outputs: returns when they are ready;
 * device outputs: asynchronous on the handle's streams (orbx_comm_sync). */
int orbx_comm_knn_match2(orbx_comm* c, const uint8_t* q, int nq, orbx_dmatch* out, int* n_per_q);
/* k = 1 (match_descriptors, matcher.rs:24-37): out nq, *n = nq (0 when either side is empty) */
int orbx_comm_match(orbx_comm* c, const uint8_t* q, int nq, orbx_dmatch* out, int* n);
int orbx_comm_sync(orbx_comm* c);
/* device time of the calls (CUDA events on devices[0]'s stream, from "queries staged" to "every rank's records have
 * landed"); with timing on, every call synchronises.  kernel_ms_rank0: accumulated Hamming-kernel time of rank 0. */
int orbx_comm_set_timing(orbx_comm* c, int enable);
int orbx_comm_timing(orbx_comm* c, float* ms, int* calls, float* kernel_ms_rank0, int reset);

/* ------------------------------------------------------------------ misc */
const char* orbx_last_error(void);
const char* orbx_version(void);
/* number of CUDA kernels this library has launched in this process */
uint64_t orbx_launch_count(void);
/* POPC-pipe microbenchmark (roofline denominator for S10): returns measured 32-bit POPC/s of
 * `device` for mode 0 (POPC only) or 1 (POPC interleaved with the XOR/add mix of the match loop) */
int orbx_popc_microbench(int device, int mode, double* popc_per_s);

#ifdef __cplusplus
}
#endif
#endif /* ORBX_H */
