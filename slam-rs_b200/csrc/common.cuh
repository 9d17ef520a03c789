// Shared helpers for liborbx (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <atomic>
#include <string>
#include <vector>

#include "../../include/orbx.h"

namespace orbx {

void set_error(const char* fmt, ...);
extern std::atomic<uint64_t> g_launches;

#define ORBX_CUDA(call)                                                                       \
  do {                                                                                        \
    cudaError_t e_ = (call);                                                                  \
    if (e_ != cudaSuccess) {                                                                  \
      orbx::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
      return ORBX_E_CUDA;                                                                     \
    }                                                                                         \
  } while (0)

// ORBX_DEBUG_SYNC=1 in the environment: synchronise the device after every launch so that a faulting kernel is
// reported at its own launch site
inline bool debug_sync() {
  static const bool on = [] {
    const char* v = getenv("ORBX_DEBUG_SYNC");
    return v && *v && *v != '0';
  }();
  return on;
}

#define ORBX_LAUNCH_CHECK()                                                                  \
  do {                                                                                       \
    orbx::g_launches.fetch_add(1, std::memory_order_relaxed);                                \
    cudaError_t e_ = cudaGetLastError();                                                     \
    if (e_ == cudaSuccess && orbx::debug_sync()) e_ = cudaDeviceSynchronize();               \
    if (e_ != cudaSuccess) {                                                                 \
      orbx::set_error("%s:%d: kernel launch -> %s", __FILE__, __LINE__, cudaGetErrorString(e_)); \
      return ORBX_E_CUDA;                                                                    \
    }                                                                                        \
  } while (0)

// true if p is device (or managed) memory reachable by kernels
inline bool is_device_ptr(const void* p) {
  if (!p) return false;
  cudaPointerAttributes a;
  cudaError_t e = cudaPointerGetAttributes(&a, p);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

inline int div_up(int a, int b) { return (a + b - 1) / b; }
inline size_t align_up(size_t a, size_t b) { return (a + b - 1) / b * b; }

constexpr int kMaxLevels = 8;

// internal entry points shared between translation units
int match_slice(orbx_matcher* m, int K, const uint8_t* q, int nq, const uint8_t* t, int nt, orbx_dmatch* out, int* counts,
                int index_base);
cudaStream_t matcher_stream(orbx_matcher* m);
int matcher_cache_train(orbx_matcher* m, const uint8_t* t, int nt);

// CUDA-event stage timer: events are recorded on the launching stream around each stage and only
// read back (with a synchronise) when the caller asks for the totals.
struct EventTimer {
  static constexpr int kStages = 8;
  bool enabled = false;
  struct Pending {
    cudaEvent_t a, b;
    int stage;
  };
  std::vector<Pending> pending;
  std::vector<cudaEvent_t> free_ev;
  double ms[kStages] = {0};
  int launches[kStages] = {0};
  cudaEvent_t get() {
    if (!free_ev.empty()) {
      cudaEvent_t e = free_ev.back();
      free_ev.pop_back();
      return e;
    }
    cudaEvent_t e;
    cudaEventCreate(&e);
    return e;
  }
  void begin(int stage, cudaStream_t s) {
    if (!enabled) return;
    Pending p{get(), get(), stage};
    cudaEventRecord(p.a, s);
    pending.push_back(p);
  }
  void end(cudaStream_t s, int nlaunches = 1) {
    if (!enabled) return;
    cudaEventRecord(pending.back().b, s);
    launches[pending.back().stage] += nlaunches;
  }
  void collect() {
    for (auto& p : pending) {
      cudaEventSynchronize(p.b);
      float t = 0;
      cudaEventElapsedTime(&t, p.a, p.b);
      ms[p.stage] += t;
      free_ev.push_back(p.a);
      free_ev.push_back(p.b);
    }
    pending.clear();
  }
  void reset() {
    for (int i = 0; i < kStages; ++i) ms[i] = 0, launches[i] = 0;
  }
  void destroy() {
    collect();
    for (auto e : free_ev) cudaEventDestroy(e);
    free_ev.clear();
  }
};

}  // namespace orbx
