// The reference's hot-path unit tests (detector.rs:51-80, matcher.rs:73-86, camera.rs:58-85, triangulation.rs:243-354, pose.rs:52-59)
// against the C++ mirror.
// Built and run by tests/test_gpu_hostmirror.py on the GPU box.
#include <cassert>
#include <cstdio>
#include <string>
#include <vector>

#include "../include/orbx.hpp"

int main() {
  {  // test_orb_detector_creation
    orbx::OrbDetector d(500, 0, 640, 480);
    assert(d.max_features() == 500);
  }
  std::vector<uint8_t> blank(480 * 640, 0);
  orbx::Image im{blank.data(), 640, 480, 640};
  {  // test_detect_on_blank_image
    orbx::OrbDetector d(100, 0, 640, 480);
    assert(d.detect(im).size() == 0);
  }
  {  // test_detect_and_compute
    orbx::OrbDetector d(500, 0, 640, 480);
    auto r = d.detect_and_compute(im);
    assert(r.first.size() == r.second.size() / 32);
  }
  {  // test_matcher_creation + test_empty_match
    orbx::FeatureMatcher m;
    std::vector<uint8_t> e;
    assert(m.match_descriptors(e, e).size() == 0);
    assert(m.filter_good_matches({}, 2.0f).size() == 0);
  }
  {  // camera.rs:58-85 test_camera_creation / test_kitti_preset / test_to_matrix
    orbx::CameraIntrinsics cam{500.0, 500.0, 320.0, 240.0};
    assert(cam.fx == 500.0 && cam.fy == 500.0 && cam.cx == 320.0 && cam.cy == 240.0);
    assert(orbx::CameraIntrinsics::kitti().fx > 700.0 && orbx::CameraIntrinsics::kitti().fy > 700.0);
    auto k = orbx::CameraIntrinsics{100.0, 200.0, 50.0, 75.0}.to_matrix();
    assert(k.size() == 9 && k[0] == 100.0 && k[2] == 50.0 && k[4] == 200.0 && k[5] == 75.0 && k[8] == 1.0 && k[1] == 0.0);
  }
  {  // triangulation.rs:243-258 test_triangulator_creation / _builder
    orbx::Triangulator t(orbx::CameraIntrinsics::kitti(), 0, 64);
    assert(t.min_parallax_deg() == 1.0 && t.max_reproj_error() == 4.0);
    t.with_min_parallax(2.0).with_max_reproj_error(5.0);
    assert(t.min_parallax_deg() == 2.0 && t.max_reproj_error() == 5.0);
  }
  {  // triangulation.rs:261-354: empty, mismatched, two synthetic points at depth 10
    orbx::Triangulator t(orbx::CameraIntrinsics{500.0, 500.0, 320.0, 240.0}, 0, 64);
    orbx::Pose p1{{1, 0, 0, 0, 1, 0, 0, 0, 1}, {0, 0, 0}}, p2{{1, 0, 0, 0, 1, 0, 0, 0, 1}, {1, 0, 0}};
    assert(t.triangulate(p1, p2, {}, {}).size() == 0);
    bool threw = false;
    try {
      t.triangulate(p1, p2, {100.0f, 100.0f}, {});
    } catch (const std::runtime_error&) {
      threw = true;
    }
    assert(threw);
    const double X[2][3] = {{0, 0, 10}, {2, 1, 10}};
    std::vector<float> a, b;
    for (auto& x : X) {
      a.push_back((float)(500.0 * x[0] / x[2] + 320.0));
      a.push_back((float)(500.0 * x[1] / x[2] + 240.0));
      b.push_back((float)(500.0 * (x[0] + 1.0) / x[2] + 320.0));
      b.push_back((float)(500.0 * x[1] / x[2] + 240.0));
    }
    auto pts = t.triangulate(p1, p2, a, b);
    assert(pts.size() == 2);
    for (auto& p : pts) assert(p.position[2] > 5.0 && p.position[2] < 15.0 && p.observations == 1 && p.descriptor.empty());
  }
  {  // pose.rs:52-59: fewer than 8 points is an error
    orbx::PoseEstimator pe(orbx::CameraIntrinsics::kitti(), 0, 256);
    bool threw = false;
    try {
      pe.compute_essential_matrix(std::vector<float>(14, 0.0f), std::vector<float>(14, 0.0f));
    } catch (const std::runtime_error& e) {
      threw = std::string(e.what()).find("Insufficient points: 7") != std::string::npos;
    }
    assert(threw);
  }
  std::puts("host mirror ok");
  return 0;
}
