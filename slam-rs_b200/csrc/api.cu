// liborbx — process-wide pieces of the C ABI: error string, version, launch counter.
#include <stdarg.h>

#include "common.cuh"

namespace orbx {

static thread_local char g_err[512] = "";
std::atomic<uint64_t> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

}  // namespace orbx

extern "C" {

const char* orbx_last_error(void) { return orbx::g_err; }
const char* orbx_version(void) { return "orbx 0.1.0 (sm_100a)"; }
uint64_t orbx_launch_count(void) { return orbx::g_launches.load(); }

void orbx_default_params(orbx_params* p) {
  if (!p) return;
  // == cv::ORB::create() defaults reached through ORB::create_def() (detector.rs:16)
  p->nfeatures = 500;
  p->scale_factor = 1.2f;
  p->nlevels = 8;
  p->edge_threshold = 31;
  p->first_level = 0;
  p->wta_k = 2;
  p->score_type = ORBX_HARRIS_SCORE;
  p->patch_size = 31;
  p->fast_threshold = 20;
}

}  // extern "C"
