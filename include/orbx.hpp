// Header-only C++ mirror of slamkit-rs `feature::{OrbDetector, FeatureMatcher}` over the C ABI (orbx.h).
// Same names and argument meaning as /root/reference/src/feature/detector.rs:15-44 and matcher.rs:15-66;
// errors become std::runtime_error (the reference returns Err(Box<dyn Error>)).
#pragma once
#include <stdexcept>
#include <string>
#include <utility>
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

}  // namespace orbx
