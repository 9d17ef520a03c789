// Header-only C++ mirror of slamkit-rs `feature::{OrbDetector, FeatureMatcher}`, `odometry::PoseEstimator` and
// `mapping::Triangulator` over the C ABI (orbx.h).  Same names and argument meaning as
// /root/reference/src/feature/detector.rs:15-44, matcher.rs:15-66, src/odometry/pose.rs:17-114 and
// src/mapping/triangulation.rs:52-198; errors become std::runtime_error (the reference returns Err(Box<dyn Error>)).
#pragma once
#include <stdexcept>
#include <string>
#include <utility>
#include <array>
#include <vector>

#include "orbx.h"

namespace orbx {

inline void check(int rc) {
  if (rc != ORBX_OK) throw std::runtime_error(std::string("liborbx: ") + orbx_last_error());
}

struct Image {  // borrowed 8-bit single-channel image (== const cv::Mat& CV_8UC1)
  const uint8_t* data;
  int width, height;
  size_t pitch;
};

class OrbDetector {
 public:
  explicit OrbDetector(int max_features, int device = 0, int max_w = 4095, int max_h = 4095) : max_features_(max_features) {
    orbx_params p;
    orbx_default_params(&p);
    p.nfeatures = max_features;
    check(orbx_detector_create(&p, device, max_w, max_h, 1, &h_));
  }
  ~OrbDetector() { orbx_detector_destroy(h_); }
  OrbDetector(const OrbDetector&) = delete;
  OrbDetector& operator=(const OrbDetector&) = delete;

  std::vector<orbx_keypoint> detect(const Image& im) { return run(im, nullptr); }
  std::pair<std::vector<orbx_keypoint>, std::vector<uint8_t>> detect_and_compute(const Image& im) {
    std::vector<uint8_t> desc;
    auto k = run(im, &desc);
    return {std::move(k), std::move(desc)};
  }
  int max_features() const { return max_features_; }

 private:
  std::vector<orbx_keypoint> run(const Image& im, std::vector<uint8_t>* desc) {
    int cap = max_features_ + max_features_ / 4 + 256;
    for (;;) {
      std::vector<orbx_keypoint> k(cap);
      if (desc) desc->assign((size_t)cap * 32, 0);
      int n = 0;
      int rc = desc ? orbx_detect_and_compute(h_, im.data, im.width, im.height, im.pitch, k.data(), desc->data(), cap, &n)
                    : orbx_detect(h_, im.data, im.width, im.height, im.pitch, k.data(), cap, &n);
      if (rc == ORBX_E_CAPACITY && n > cap) {
        cap = n;
        continue;
      }
      check(rc);
      k.resize(n);
      if (desc) desc->resize((size_t)n * 32);
      return k;
    }
  }
  orbx_detector* h_ = nullptr;
  int max_features_;
};

class FeatureMatcher {
 public:
  explicit FeatureMatcher(int device = 0, bool cross_check = false) { check(orbx_matcher_create(device, cross_check, &h_)); }
  ~FeatureMatcher() { orbx_matcher_destroy(h_); }
  FeatureMatcher(const FeatureMatcher&) = delete;
  FeatureMatcher& operator=(const FeatureMatcher&) = delete;

  // desc: N x 32 bytes row-major.  Either side empty -> empty result (matcher.rs:29-31).
  std::vector<orbx_dmatch> match_descriptors(const std::vector<uint8_t>& d1, const std::vector<uint8_t>& d2) {
    const int nq = (int)(d1.size() / 32), nt = (int)(d2.size() / 32);
    std::vector<orbx_dmatch> out(nq);
    int n = 0;
    check(orbx_match(h_, d1.data(), nq, d2.data(), nt, out.data(), &n));
    out.resize(n);
    return out;
  }
  std::vector<orbx_dmatch> filter_good_matches(const std::vector<orbx_dmatch>& m, float ratio) const {
    std::vector<orbx_dmatch> out(m.size());
    int n = 0;
    check(orbx_filter_good_matches(h_, m.data(), (int)m.size(), ratio, out.data(), &n));
    out.resize(n);
    return out;
  }
  // PoseEstimator::extract_matched_points (src/odometry/pose.rs:27-44): (x, y) pairs, match order; throws when a
  // match index is outside its keypoint vector (the reference returns Err)
  std::pair<std::vector<float>, std::vector<float>> extract_matched_points(const std::vector<orbx_keypoint>& kp1,
                                                                          const std::vector<orbx_keypoint>& kp2,
                                                                          const std::vector<orbx_dmatch>& m) {
    std::vector<float> p1(2 * m.size()), p2(2 * m.size());
    check(orbx_extract_matched_points(h_, kp1.data(), (int)kp1.size(), kp2.data(), (int)kp2.size(), m.data(),
                                      (int)m.size(), p1.data(), p2.data()));
    return {p1, p2};
  }

 private:
  orbx_matcher* h_ = nullptr;
};

struct CameraIntrinsics {  // odometry::CameraIntrinsics (camera.rs:5-55)
  double fx, fy, cx, cy;
  static CameraIntrinsics kitti() { return {718.856, 718.856, 607.1928, 185.2157}; }
  static CameraIntrinsics webcam_vga() { return {500.0, 500.0, 320.0, 240.0}; }                   // camera.rs:29-36
  std::array<double, 9> to_matrix() const { return {fx, 0.0, cx, 0.0, fy, cy, 0.0, 0.0, 1.0}; }  // camera.rs:39-54, row-major
};

struct Pose {  // (R, t), world to camera; R row-major
  double R[9];
  double t[3];
};

// odometry::PoseEstimator (pose.rs:10-114); points are (x, y) pairs as returned by extract_matched_points
class PoseEstimator {
 public:
  explicit PoseEstimator(const CameraIntrinsics& k, int device = 0, int max_points = 1 << 14) {
    check(orbx_pose_estimator_create(device, max_points, 1, k.fx, k.fy, k.cx, k.cy, &h_));
  }
  ~PoseEstimator() { orbx_pose_estimator_destroy(h_); }
  PoseEstimator(const PoseEstimator&) = delete;
  PoseEstimator& operator=(const PoseEstimator&) = delete;

  // pose.rs:47-80: E row-major (9 doubles); throws "Insufficient points: n (need 8)" / "Failed to compute essential matrix"
  std::vector<double> compute_essential_matrix(const std::vector<float>& p1, const std::vector<float>& p2) {
    std::vector<double> E(9);
    check(orbx_compute_essential_matrix(h_, p1.data(), (int)(p1.size() / 2), p2.data(), (int)(p2.size() / 2), E.data(),
                                        nullptr, nullptr, nullptr));
    return E;
  }
  // pose.rs:83-114: throws "Too few inliers: k"
  Pose recover_pose(const std::vector<double>& E, const std::vector<float>& p1, const std::vector<float>& p2) {
    Pose out{};
    check(orbx_recover_pose(h_, E.data(), p1.data(), (int)(p1.size() / 2), p2.data(), (int)(p2.size() / 2), out.R, out.t,
                            nullptr, nullptr));
    return out;
  }

 private:
  orbx_pose_estimator* h_ = nullptr;
};

struct MapPoint {  // mapping::MapPoint (triangulation.rs:11-22)
  double position[3];
  std::vector<uint8_t> descriptor;  // empty == None
  size_t observations = 1;
  size_t id = 0;
};

// mapping::Triangulator (triangulation.rs:52-198)
class Triangulator {
 public:
  explicit Triangulator(const CameraIntrinsics& k, int device = 0, int max_points = 1 << 16) {
    check(orbx_triangulator_create(device, max_points, k.fx, k.fy, k.cx, k.cy, &h_));
  }
  ~Triangulator() { orbx_triangulator_destroy(h_); }
  Triangulator(const Triangulator&) = delete;
  Triangulator& operator=(const Triangulator&) = delete;

  Triangulator& with_min_parallax(double deg) {
    check(orbx_triangulator_set_thresholds(h_, deg, max_reproj_error()));
    return *this;
  }
  Triangulator& with_max_reproj_error(double e) {
    check(orbx_triangulator_set_thresholds(h_, min_parallax_deg(), e));
    return *this;
  }
  double min_parallax_deg() const {
    double a = 0, b = 0;
    check(orbx_triangulator_get_thresholds(h_, &a, &b));
    return a;
  }
  double max_reproj_error() const {
    double a = 0, b = 0;
    check(orbx_triangulator_get_thresholds(h_, &a, &b));
    return b;
  }
  // triangulation.rs:92-158; descriptors: rows x 32 bytes or empty (None); throws on mismatched point arrays
  std::vector<MapPoint> triangulate(const Pose& pose1, const Pose& pose2, const std::vector<float>& p1,
                                    const std::vector<float>& p2, const std::vector<uint8_t>& descriptors = {}) {
    const int n1 = (int)(p1.size() / 2), n2 = (int)(p2.size() / 2), cap = n1 > 0 ? n1 : 1;
    std::vector<int32_t> ids(cap);
    std::vector<double> xyz(3 * (size_t)cap);
    std::vector<uint8_t> desc(32 * (size_t)cap), has(cap);
    int m = 0;
    check(orbx_triangulate(h_, pose1.R, pose1.t, pose2.R, pose2.t, p1.data(), n1, p2.data(), n2,
                           descriptors.empty() ? nullptr : descriptors.data(), (int)(descriptors.size() / 32), ids.data(),
                           xyz.data(), desc.data(), has.data(), nullptr, cap, &m));
    std::vector<MapPoint> out(m);
    for (int k = 0; k < m; ++k) {
      for (int j = 0; j < 3; ++j) out[k].position[j] = xyz[3 * k + j];
      out[k].id = (size_t)ids[k];
      if (has[k]) out[k].descriptor.assign(desc.begin() + 32 * k, desc.begin() + 32 * (k + 1));
    }
    return out;
  }

 private:
  orbx_triangulator* h_ = nullptr;
};

// The per-frame loop of examples/visual_odometry.rs:84-218 over a whole sequence (csrc/sequence.cu): host frames in,
// keypoints / descriptors of every frame and the good matches of every consecutive pair out.
class Sequence {
 public:
  struct Result {  // row t of good / n_good describes the pair (t-1, t); rows hold capacity() entries, n[t] / n_good[t] valid
    std::vector<orbx_keypoint> kps;
    std::vector<uint8_t> desc;
    std::vector<int> n;
    std::vector<orbx_dmatch> good;
    std::vector<int> n_good;
  };
  Sequence(int width, int height, int nfeatures, int max_frames, int device = 0, int chunk = 1000, int host_chunk = 250,
           float ratio = 0.75f, bool cross_check = true)
      : max_frames_(max_frames) {
    check(orbx_sequence_create(device, width, height, nfeatures, max_frames, chunk, host_chunk, ratio, cross_check ? 1 : 0,
                               &h_));
  }
  ~Sequence() { orbx_sequence_destroy(h_); }
  Sequence(const Sequence&) = delete;
  Sequence& operator=(const Sequence&) = delete;

  int capacity() const { return orbx_sequence_capacity(h_); }
  // frames: T tightly packed 8-bit images (pinned memory lets the upload overlap the kernels); blocks until done
  Result run(const uint8_t* frames, int T) {
    const size_t cap = (size_t)capacity(), n = (size_t)T;
    Result r;
    r.kps.resize(n * cap), r.desc.resize(n * cap * 32), r.n.resize(n), r.good.resize(n * cap), r.n_good.resize(n);
    check(orbx_sequence_run_host(h_, frames, T, r.kps.data(), r.desc.data(), r.n.data(), r.good.data(), r.n_good.data()));
    check(orbx_sequence_sync(h_));
    return r;
  }
  // Streaming form: queue a sequence into caller-owned (pinned) result arrays and return at once; consecutive calls are
  // pipelined into each other.  wait_pass(1) blocks until the call before the most recent one has finished.
  void run_async(const uint8_t* frames, int T, orbx_keypoint* kps, uint8_t* desc, int* n, orbx_dmatch* good, int* n_good) {
    check(orbx_sequence_run_host(h_, frames, T, kps, desc, n, good, n_good));
  }
  void wait_pass(int back = 0) { check(orbx_sequence_wait_pass(h_, back)); }
  void sync() { check(orbx_sequence_sync(h_)); }

 private:
  orbx_sequence* h_ = nullptr;
  int max_frames_;
};

}  // namespace orbx
