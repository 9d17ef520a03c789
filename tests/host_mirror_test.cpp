// The reference's five hot-path unit tests (detector.rs:51-80, matcher.rs:73-86) against the C++ mirror.
// Built and run by tests/test_gpu_hostmirror.py on the GPU box.
#include <cassert>
#include <cstdio>
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
  std::puts("host mirror ok");
  return 0;
}
