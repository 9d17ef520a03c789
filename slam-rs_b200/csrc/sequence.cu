// liborbx — the sequence front end as a native entry of the C ABI.
//
// /root/reference/examples/visual_odometry.rs:84-218 runs, per frame, detect_and_compute (:102), match_descriptors
// (:106) and a match filter (:107).  Over a whole sequence the extractions are independent and so are the T-1
// frame-to-frame matching problems; this file runs them as batched launches of the detector / matcher handles with the
// pipeline that slam-rs_b200/frontend.py::SequenceFrontEnd measured best on B200: frames go up in chunks on a copy
// stream, two detector + matcher pairs on two streams alternate over the chunks (the tail wave of every kernel of one
// chunk overlaps the next chunk's kernels), results go back on a fourth stream while the next chunk computes.
// No kernels of its own: host-side scheduling only (streams, events, the chunk schedule).
#include <vector>

#include "common.cuh"

struct orbx_sequence {
  int device = 0, w = 0, h = 0, nfeatures = 0, max_frames = 0, chunk = 0, host_chunk = 0, cap = 0;
  float ratio = 0.75f;
  int reference_filter = 0;  // 1: match_descriptors (k=1) + filter_good_matches(filter_ratio), the reference's own loop
  float filter_ratio = 2.0f;
  int* n_k1 = nullptr;       // [T] match counts of the k=1 pass (reference_filter)
  orbx_detector* det[2] = {nullptr, nullptr};
  orbx_matcher* mat[2] = {nullptr, nullptr};
  cudaStream_t st[2] = {nullptr, nullptr}, copy = nullptr, d2h = nullptr;
  // device-resident results of one pass (row t of knn / good / n_good describes the pair (t-1, t))
  uint8_t* frames = nullptr;  // [T][h][w] staging of run_host, tightly packed
  orbx_keypoint* kps = nullptr;
  uint8_t* desc = nullptr;
  int* n = nullptr;
  orbx_dmatch* knn = nullptr;
  int* n_per_q = nullptr;
  orbx_dmatch* good = nullptr;
  int* n_good = nullptr;
  // event pools: a pass takes its events from pool `pool`; the events of the two passes before stay intact in the others
  std::vector<cudaEvent_t> events[3];
  int pool = 0;
  size_t next_event = 0;
  // run_host pipelines consecutive passes: the frame staging and every result array exist twice and the passes
  // alternate between the two sets (`spare` = the set the current pass does not use), so pass k+1 shares no buffer with
  // pass k - its uploads and kernels run under pass k's tail - and only waits, chunk by chunk, for what pass k-1
  // (hist[1], same set) still does with its buffers
  struct Set {
    uint8_t* frames = nullptr;
    orbx_keypoint* kps = nullptr;
    uint8_t* desc = nullptr;
    int* n = nullptr;
    orbx_dmatch* knn = nullptr;
    int* n_per_q = nullptr;
    orbx_dmatch* good = nullptr;
    int* n_good = nullptr;
    int* n_k1 = nullptr;
  } spare;
  bool have_spare = false;
  struct ChunkEvents {
    std::vector<std::pair<int, int>> bounds;            // frames [first, second) of every chunk
    std::vector<cudaEvent_t> extracted, matched, done;  // per chunk: extraction / matching finished, results downloaded
  } hist[2];                                            // [0] the pass before, [1] two passes back
  bool pipelining = false;                              // the last call was run_host and nothing joined since
  unsigned chunk_seq = 0;                               // chunks alternate between the two compute streams across passes
  cudaEvent_t pass_done[4] = {nullptr, nullptr, nullptr, nullptr};  // ring, recorded at the end of every pass
  unsigned long long pass_seq = 0;
  // device-side timing of the passes (orbx_sequence_set_timing): t_begin is recorded when the first pass since the last
  // reset starts, t_end when everything a pass queued has finished (re-recorded by every pass)
  bool timing = false, have_begin = false;
  cudaEvent_t t_begin = nullptr, t_end = nullptr;
  int passes = 0;
};

namespace {

using orbx::set_error;

int take_event(orbx_sequence* s, cudaEvent_t* e) {
  std::vector<cudaEvent_t>& pool = s->events[s->pool];
  if (s->next_event == pool.size()) {
    cudaEvent_t ev;
    ORBX_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    pool.push_back(ev);
  }
  *e = pool[s->next_event++];
  return ORBX_OK;
}

// every pass starts here: fresh event pool (the previous pass keeps its own)
void begin_pass(orbx_sequence* s) {
  s->pool = (s->pool + 1) % 3;
  s->next_event = 0;
}

void swap_sets(orbx_sequence* s) {
  std::swap(s->frames, s->spare.frames);
  std::swap(s->kps, s->spare.kps);
  std::swap(s->desc, s->spare.desc);
  std::swap(s->n, s->spare.n);
  std::swap(s->knn, s->spare.knn);
  std::swap(s->n_per_q, s->spare.n_per_q);
  std::swap(s->good, s->spare.good);
  std::swap(s->n_good, s->spare.n_good);
  std::swap(s->n_k1, s->spare.n_k1);
}

int mark_begin(orbx_sequence* s) {
  if (s->timing && !s->have_begin) {
    ORBX_CUDA(cudaEventRecord(s->t_begin, s->copy));  // the copy stream carries the first upload of the pass
    s->have_begin = true;
  }
  return ORBX_OK;
}

// `waiter` continues after everything queued on `on` so far
int wait_stream(orbx_sequence* s, cudaStream_t waiter, cudaStream_t on) {
  cudaEvent_t e;
  if (int rc = take_event(s, &e)) return rc;
  ORBX_CUDA(cudaEventRecord(e, on));
  ORBX_CUDA(cudaStreamWaitEvent(waiter, e, 0));
  return ORBX_OK;
}

// a pass starts after everything the previous pass queued on any of the four streams
int join_all(orbx_sequence* s) {
  s->pipelining = false;
  s->hist[0] = s->hist[1] = orbx_sequence::ChunkEvents();
  cudaStream_t all[4] = {s->st[0], s->st[1], s->copy, s->d2h};
  cudaEvent_t e[4];
  for (int i = 0; i < 4; ++i) {
    if (int rc = take_event(s, &e[i])) return rc;
    ORBX_CUDA(cudaEventRecord(e[i], all[i]));
  }
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j)
      if (i != j) ORBX_CUDA(cudaStreamWaitEvent(all[i], e[j], 0));
  return mark_begin(s);  // the copy stream has just joined everything queued before
}

// end of a pass: the d2h stream waits for the other three and carries the end-of-pass timestamp
int finish_pass(orbx_sequence* s) {
  cudaStream_t others[3] = {s->st[0], s->st[1], s->copy};
  for (cudaStream_t o : others)
    if (int rc = wait_stream(s, s->d2h, o)) return rc;
  cudaEvent_t& done = s->pass_done[s->pass_seq & 3];
  if (!done) ORBX_CUDA(cudaEventCreateWithFlags(&done, cudaEventDisableTiming));
  ORBX_CUDA(cudaEventRecord(done, s->d2h));
  ++s->pass_seq;
  if (!s->timing) return ORBX_OK;
  ORBX_CUDA(cudaEventRecord(s->t_end, s->d2h));
  ++s->passes;
  return ORBX_OK;
}

// frames a..b-1 through detector `k` in launch sets of at most `step` frames
int extract(orbx_sequence* s, int k, const uint8_t* frames, size_t pitch, size_t stride, int a, int b, int step) {
  for (int f = a; f < b; f += step) {
    const int nb = std::min(step, b - f);
    const int rc = orbx_detect_and_compute_batch(s->det[k], frames + (size_t)f * stride, s->w, s->h, pitch, stride, nb,
                                                 s->kps + (size_t)f * s->cap, s->desc + (size_t)f * s->cap * 32, s->cap,
                                                 s->n + f);
    if (rc) return rc;
  }
  return ORBX_OK;
}

// pairs (t-1, t) for t in [max(a, 1), b), batched: BFMatcher k=2 (+ cross-check) and Lowe's ratio test (north_star
// stage 5), or - reference_filter - exactly what the reference's loop runs: k=1 + filter_good_matches
int match(orbx_sequence* s, int k, int a, int b) {
  a = std::max(a, 1);
  if (b <= a) return ORBX_OK;
  const int npairs = b - a;
  const size_t cap = (size_t)s->cap, stride = cap * 32;
  if (s->reference_filter) {
    // visual_odometry.rs:106-107: match_descriptors(&prev_desc, &desc) then filter_good_matches(&matches, 2.0)
    int rc = orbx_match_batch(s->mat[k], s->desc + (size_t)(a - 1) * stride, stride, s->n + (a - 1),
                              s->desc + (size_t)a * stride, stride, s->n + a, npairs, s->cap, s->knn + (size_t)a * cap,
                              s->n_k1 + a);
    if (rc) return rc;
    return orbx_filter_good_matches_batch(s->mat[k], s->knn + (size_t)a * cap, s->n_k1 + a, npairs, s->cap,
                                          s->filter_ratio, s->good + (size_t)a * cap, s->n_good + a);
  }
  int rc = orbx_knn_match2_batch(s->mat[k], s->desc + (size_t)(a - 1) * stride, stride, s->n + (a - 1),
                                 s->desc + (size_t)a * stride, stride, s->n + a, npairs, s->cap, s->knn + (size_t)a * cap * 2,
                                 s->n_per_q + (size_t)a * cap);
  if (rc) return rc;
  return orbx_ratio_test_batch(s->mat[k], s->knn + (size_t)a * cap * 2, s->n_per_q + (size_t)a * cap, s->n + (a - 1), npairs,
                               s->cap, s->ratio, s->good + (size_t)a * cap, s->n_good + a);
}

template <typename T>
int dev_alloc(T** p, size_t count) {
  cudaError_t e = cudaMalloc((void**)p, count * sizeof(T));
  if (e != cudaSuccess) {
    set_error("cudaMalloc(%zu) failed: %s", count * sizeof(T), cudaGetErrorString(e));
    return ORBX_E_NOMEM;
  }
  e = cudaMemset(*p, 0, count * sizeof(T));
  if (e != cudaSuccess) {
    set_error("cudaMemset failed: %s", cudaGetErrorString(e));
    return ORBX_E_CUDA;
  }
  return ORBX_OK;
}

// The chunk schedule of run_host (measured in frontend.py): three short, doubling first chunks so that compute starts
// early, full chunks in the middle, a short last chunk so that little is left to copy back after the last kernel.
std::vector<std::pair<int, int>> chunk_schedule(int T, int c) {
  std::vector<std::pair<int, int>> bounds;
  int a = 0;
  const int first[3] = {std::max(c / 8, 1), std::max(c / 4, 1), std::max(c / 2, 1)};
  for (int size : first)
    if (T - a > c) {
      bounds.push_back({a, a + size});
      a += size;
    }
  while (T - a > c + c / 2) {
    bounds.push_back({a, a + c});
    a += c;
  }
  if (T - a > c) {
    bounds.push_back({a, T - c / 2});
    a = T - c / 2;
  }
  bounds.push_back({a, T});
  return bounds;
}

// Steady state of pipelined passes: chunks of equal size (at most c frames each)
std::vector<std::pair<int, int>> even_schedule(int T, int c) {
  const int nc = (T + c - 1) / c;
  std::vector<std::pair<int, int>> bounds;
  for (int i = 0; i < nc; ++i) bounds.push_back({(int)((long)T * i / nc), (int)((long)T * (i + 1) / nc)});
  return bounds;
}

}  // namespace

extern "C" {

void orbx_sequence_destroy(orbx_sequence* s) {
  if (!s) return;
  cudaSetDevice(s->device);
  cudaDeviceSynchronize();
  for (int k = 0; k < 2; ++k) {
    if (s->det[k]) orbx_detector_destroy(s->det[k]);
    if (s->mat[k]) orbx_matcher_destroy(s->mat[k]);
    if (s->st[k]) cudaStreamDestroy(s->st[k]);
  }
  if (s->copy) cudaStreamDestroy(s->copy);
  if (s->d2h) cudaStreamDestroy(s->d2h);
  for (auto& pool : s->events)
    for (cudaEvent_t e : pool) cudaEventDestroy(e);
  for (cudaEvent_t e : s->pass_done)
    if (e) cudaEventDestroy(e);
  if (s->t_begin) cudaEventDestroy(s->t_begin);
  if (s->t_end) cudaEventDestroy(s->t_end);
  cudaFree(s->frames);
  cudaFree(s->kps);
  cudaFree(s->desc);
  cudaFree(s->n);
  cudaFree(s->knn);
  cudaFree(s->n_per_q);
  cudaFree(s->good);
  cudaFree(s->n_good);
  cudaFree(s->n_k1);
  cudaFree(s->spare.frames);
  cudaFree(s->spare.kps);
  cudaFree(s->spare.desc);
  cudaFree(s->spare.n);
  cudaFree(s->spare.knn);
  cudaFree(s->spare.n_per_q);
  cudaFree(s->spare.good);
  cudaFree(s->spare.n_good);
  cudaFree(s->spare.n_k1);
  delete s;
}

int orbx_sequence_create(int device, int width, int height, int nfeatures, int max_frames, int chunk, int host_chunk,
                         float ratio, int cross_check, orbx_sequence** out) {
  if (!out || width < 16 || height < 16 || nfeatures < 1 || max_frames < 1 || chunk < 1 || host_chunk < 1) {
    set_error("orbx_sequence_create: bad argument");
    return ORBX_E_BADARG;
  }
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) {
    cudaGetLastError();
    set_error("orbx_sequence_create: no CUDA device %d (there is no CPU fallback)", device);
    return ORBX_E_CUDA;
  }
  ORBX_CUDA(cudaSetDevice(device));
  auto* s = new orbx_sequence();
  s->device = device, s->w = width, s->h = height, s->nfeatures = nfeatures, s->max_frames = max_frames;
  s->chunk = std::min(chunk, max_frames);
  s->host_chunk = std::min(host_chunk, s->chunk);
  s->ratio = ratio;
  // OpenCV keeps ties at the selection threshold, so a frame can return a little more than nfeatures (orbx.h)
  s->cap = (nfeatures + nfeatures / 8 + 64 + 31) / 32 * 32;
  orbx_params p;
  orbx_default_params(&p);
  p.nfeatures = nfeatures;
  int rc = ORBX_OK;
  for (int k = 0; k < 2 && !rc; ++k) {
    if (cudaStreamCreateWithFlags(&s->st[k], cudaStreamNonBlocking) != cudaSuccess) rc = ORBX_E_CUDA;
    // handle 0 also serves run_device, whose launch sets are `chunk` frames
    if (!rc) rc = orbx_detector_create(&p, device, width, height, k == 0 ? s->chunk : s->host_chunk, &s->det[k]);
    if (!rc) rc = orbx_matcher_create(device, cross_check, &s->mat[k]);
    if (!rc) rc = orbx_detector_set_stream(s->det[k], s->st[k]);
    if (!rc) rc = orbx_matcher_set_stream(s->mat[k], s->st[k]);
  }
  if (!rc && cudaStreamCreateWithFlags(&s->copy, cudaStreamNonBlocking) != cudaSuccess) rc = ORBX_E_CUDA;
  if (!rc && cudaStreamCreateWithFlags(&s->d2h, cudaStreamNonBlocking) != cudaSuccess) rc = ORBX_E_CUDA;
  const size_t T = (size_t)max_frames, cap = (size_t)s->cap;
  if (!rc) rc = dev_alloc(&s->frames, T * height * width);
  if (!rc) rc = dev_alloc(&s->kps, T * cap);
  if (!rc) rc = dev_alloc(&s->desc, T * cap * 32);
  if (!rc) rc = dev_alloc(&s->n, T);
  if (!rc) rc = dev_alloc(&s->knn, T * cap * 2);
  if (!rc) rc = dev_alloc(&s->n_per_q, T * cap);
  if (!rc) rc = dev_alloc(&s->good, T * cap);
  if (!rc) rc = dev_alloc(&s->n_good, T);
  if (!rc) rc = dev_alloc(&s->n_k1, T);
  if (!rc && cudaDeviceSynchronize() != cudaSuccess) rc = ORBX_E_CUDA;  // the memsets ran on the legacy stream
  if (rc) {
    if (rc == ORBX_E_CUDA) set_error("orbx_sequence_create: CUDA error %s", cudaGetErrorString(cudaGetLastError()));
    orbx_sequence_destroy(s);
    return rc;
  }
  *out = s;
  return ORBX_OK;
}

int orbx_sequence_capacity(const orbx_sequence* s) { return s ? s->cap : 0; }

int orbx_sequence_set_timing(orbx_sequence* s, int enable) {
  if (!s) {
    set_error("orbx_sequence_set_timing: null handle");
    return ORBX_E_BADARG;
  }
  ORBX_CUDA(cudaSetDevice(s->device));
  if (enable && !s->t_begin) {
    ORBX_CUDA(cudaEventCreate(&s->t_begin));
    ORBX_CUDA(cudaEventCreate(&s->t_end));
  }
  s->timing = enable != 0;
  s->have_begin = false;
  s->passes = 0;
  return ORBX_OK;
}

int orbx_sequence_elapsed(orbx_sequence* s, float* ms, int* passes, int reset) {
  if (!s || !ms) {
    set_error("orbx_sequence_elapsed: bad argument");
    return ORBX_E_BADARG;
  }
  *ms = 0.0f;
  if (passes) *passes = s->passes;
  if (!s->timing || !s->have_begin || s->passes == 0) return ORBX_OK;
  ORBX_CUDA(cudaSetDevice(s->device));
  ORBX_CUDA(cudaEventSynchronize(s->t_end));
  ORBX_CUDA(cudaEventElapsedTime(ms, s->t_begin, s->t_end));
  if (reset) {
    s->have_begin = false;
    s->passes = 0;
  }
  return ORBX_OK;
}

int orbx_sequence_set_reference_filter(orbx_sequence* s, int enable, float ratio) {
  if (!s || !(ratio > 0.0f)) {
    set_error("orbx_sequence_set_reference_filter: bad argument");
    return ORBX_E_BADARG;
  }
  s->reference_filter = enable ? 1 : 0;
  s->filter_ratio = ratio;
  return ORBX_OK;
}

int orbx_sequence_sync(orbx_sequence* s) {
  if (!s) {
    set_error("orbx_sequence_sync: null handle");
    return ORBX_E_BADARG;
  }
  ORBX_CUDA(cudaSetDevice(s->device));
  ORBX_CUDA(cudaStreamSynchronize(s->copy));
  ORBX_CUDA(cudaStreamSynchronize(s->st[0]));
  ORBX_CUDA(cudaStreamSynchronize(s->st[1]));
  ORBX_CUDA(cudaStreamSynchronize(s->d2h));
  s->pipelining = false;  // idle now: the next host pass starts with the ramp schedule
  // a frame with more keypoints than the row capacity (tie floods) was truncated by the asynchronous batch calls: the
  // detectors park that in a device flag, reported here once as ORBX_E_CAPACITY
  int rc = ORBX_OK;
  for (int k = 0; k < 2; ++k) {
    const int r = orbx_detector_sync(s->det[k]);
    if (r && !rc) rc = r;
  }
  return rc;
}

int orbx_sequence_results(orbx_sequence* s, orbx_keypoint** kps, uint8_t** desc, int** n, orbx_dmatch** good,
                          int** n_good) {
  if (!s) {
    set_error("orbx_sequence_results: null handle");
    return ORBX_E_BADARG;
  }
  if (kps) *kps = s->kps;
  if (desc) *desc = s->desc;
  if (n) *n = s->n;
  if (good) *good = s->good;
  if (n_good) *n_good = s->n_good;
  return ORBX_OK;
}

int orbx_sequence_run_device(orbx_sequence* s, const uint8_t* frames_dev, size_t pitch, size_t frame_stride, int T) {
  if (!s || !frames_dev || T < 1 || T > s->max_frames || pitch < (size_t)s->w || frame_stride < pitch * s->h) {
    set_error("orbx_sequence_run_device: bad argument");
    return ORBX_E_BADARG;
  }
  if (!orbx::is_device_ptr(frames_dev)) {
    set_error("orbx_sequence_run_device: frames must be device memory (orbx_sequence_run_host takes host frames)");
    return ORBX_E_BADARG;
  }
  ORBX_CUDA(cudaSetDevice(s->device));
  begin_pass(s);
  if (int rc = join_all(s)) return rc;
  for (int a = 0; a < T; a += s->chunk) {
    const int b = std::min(a + s->chunk, T);
    if (int rc = extract(s, 0, frames_dev, pitch, frame_stride, a, b, s->chunk)) return rc;
    if (int rc = match(s, 0, a, b)) return rc;
  }
  return finish_pass(s);
}

int orbx_sequence_run_host(orbx_sequence* s, const uint8_t* frames, int T, orbx_keypoint* kps, uint8_t* desc, int* n,
                           orbx_dmatch* good, int* n_good) {
  if (!s || !frames || T < 1 || T > s->max_frames) {
    set_error("orbx_sequence_run_host: bad argument");
    return ORBX_E_BADARG;
  }
  ORBX_CUDA(cudaSetDevice(s->device));
  const size_t fb = (size_t)s->h * s->w, cap = (size_t)s->cap;
  // the second buffer set, on the first host pass (a handle that only ever runs device passes never pays for it)
  if (!s->have_spare) {
    const size_t Tm = (size_t)s->max_frames;
    int rc = dev_alloc(&s->spare.frames, Tm * fb);
    if (!rc) rc = dev_alloc(&s->spare.kps, Tm * cap);
    if (!rc) rc = dev_alloc(&s->spare.desc, Tm * cap * 32);
    if (!rc) rc = dev_alloc(&s->spare.n, Tm);
    if (!rc) rc = dev_alloc(&s->spare.knn, Tm * cap * 2);
    if (!rc) rc = dev_alloc(&s->spare.n_per_q, Tm * cap);
    if (!rc) rc = dev_alloc(&s->spare.good, Tm * cap);
    if (!rc) rc = dev_alloc(&s->spare.n_good, Tm);
    if (!rc) rc = dev_alloc(&s->spare.n_k1, Tm);
    if (!rc && cudaDeviceSynchronize() != cudaSuccess) rc = ORBX_E_CUDA;  // the memsets ran on the legacy stream
    if (rc) {
      void* part[9] = {s->spare.frames, s->spare.kps, s->spare.desc, s->spare.n, s->spare.knn,
                       s->spare.n_per_q, s->spare.good, s->spare.n_good, s->spare.n_k1};
      for (void* q : part) cudaFree(q);
      s->spare = orbx_sequence::Set();
      return rc;
    }
    s->have_spare = true;
  }
  // The call before was run_host too and nothing joined since: no join.  This pass takes the buffer set the last pass
  // did NOT use, so a chunk [a, b) only waits for what the pass before the last (same set) still does with these frames:
  //   upload        after that pass extracted every chunk that overlaps [a, b)                  (device frame staging)
  //   extraction    after that pass downloaded every chunk that overlaps [a, b)       (kps / desc / n / good / n_good)
  //                 and matched every chunk that reads a descriptor row of [a, b): chunk j matches the pairs (t-1, t)
  //                 for t in [a_j, b_j), so it reads frames a_j - 1 .. b_j - 1
  // (all long finished in practice).  Detector / matcher scratch is per compute stream and the streams are in order.
  // Its uploads are hidden under the last pass, so instead of the start-up ramp over two compute streams it takes
  // launch sets of `chunk` frames on one compute stream, like a device pass (measured on B200, 1000 frames 1241x376:
  // 11.7 ms per pass in the steady state against 12.0 ms with four 250-frame stages alternating between two streams).
  // After a sync the GPU is idle and the next pass starts with the ramp again (orbx_sequence_sync clears `pipelining`).
  const bool pipelined = s->pipelining;
  s->pipelining = false;  // set again when the whole pass has been queued: a call that fails half-way ends the pipelining
  const auto bounds = pipelined ? even_schedule(T, s->chunk) : chunk_schedule(T, s->host_chunk);
  const size_t nc = bounds.size();
  begin_pass(s);
  if (pipelined) {
    if (int rc = mark_begin(s)) return rc;
  } else if (int rc = join_all(s)) {
    return rc;
  }
  swap_sets(s);
  const orbx_sequence::ChunkEvents last = s->hist[1];  // same buffer set as this pass
  orbx_sequence::ChunkEvents cur;
  cur.extracted.resize(nc), cur.matched.resize(nc), cur.done.resize(nc);
  cur.bounds = bounds;
  std::vector<cudaEvent_t> uploaded(nc);
  for (size_t i = 0; i < bounds.size(); ++i) {
    const int a = bounds[i].first, b = bounds[i].second;
    for (size_t j = 0; j < last.bounds.size(); ++j)
      if (last.bounds[j].first < b && last.bounds[j].second > a)
        ORBX_CUDA(cudaStreamWaitEvent(s->copy, last.extracted[j], 0));
    ORBX_CUDA(cudaMemcpyAsync(s->frames + a * fb, frames + a * fb, (size_t)(b - a) * fb, cudaMemcpyHostToDevice, s->copy));
    if (int rc = take_event(s, &uploaded[i])) return rc;
    ORBX_CUDA(cudaEventRecord(uploaded[i], s->copy));
  }
  cudaEvent_t prev_extracted = nullptr;
  for (size_t i = 0; i < bounds.size(); ++i) {
    const int a = bounds[i].first, b = bounds[i].second, k = pipelined ? 0 : (int)(s->chunk_seq++ & 1);
    cudaStream_t st = s->st[k];
    ORBX_CUDA(cudaStreamWaitEvent(st, uploaded[i], 0));
    for (size_t j = 0; j < last.bounds.size(); ++j) {
      const int aj = last.bounds[j].first, bj = last.bounds[j].second;
      if (aj < b && bj > a) ORBX_CUDA(cudaStreamWaitEvent(st, last.done[j], 0));
      if (aj <= b && bj > a) ORBX_CUDA(cudaStreamWaitEvent(st, last.matched[j], 0));
    }
    if (int rc = extract(s, k, s->frames, (size_t)s->w, fb, a, b, pipelined ? s->chunk : s->host_chunk)) return rc;
    cudaEvent_t extracted;
    if (int rc = take_event(s, &extracted)) return rc;
    ORBX_CUDA(cudaEventRecord(extracted, st));
    cur.extracted[i] = extracted;
    // keypoints and descriptors of this chunk go back while it is being matched ...
    ORBX_CUDA(cudaStreamWaitEvent(s->d2h, extracted, 0));
    if (n) ORBX_CUDA(cudaMemcpyAsync(n + a, s->n + a, (size_t)(b - a) * sizeof(int), cudaMemcpyDeviceToHost, s->d2h));
    if (kps)
      ORBX_CUDA(cudaMemcpyAsync(kps + a * cap, s->kps + a * cap, (size_t)(b - a) * cap * sizeof(orbx_keypoint),
                                cudaMemcpyDeviceToHost, s->d2h));
    if (desc)
      ORBX_CUDA(cudaMemcpyAsync(desc + a * cap * 32, s->desc + a * cap * 32, (size_t)(b - a) * cap * 32,
                                cudaMemcpyDeviceToHost, s->d2h));
    // descriptors of frame a-1 come from the other stream
    if (prev_extracted) ORBX_CUDA(cudaStreamWaitEvent(st, prev_extracted, 0));
    prev_extracted = extracted;
    if (int rc = match(s, k, a, b)) return rc;
    // ... and its matches while the next chunk computes
    if (int rc = take_event(s, &cur.matched[i])) return rc;
    ORBX_CUDA(cudaEventRecord(cur.matched[i], st));
    ORBX_CUDA(cudaStreamWaitEvent(s->d2h, cur.matched[i], 0));
    if (good)
      ORBX_CUDA(cudaMemcpyAsync(good + a * cap, s->good + a * cap, (size_t)(b - a) * cap * sizeof(orbx_dmatch),
                                cudaMemcpyDeviceToHost, s->d2h));
    if (n_good)
      ORBX_CUDA(cudaMemcpyAsync(n_good + a, s->n_good + a, (size_t)(b - a) * sizeof(int), cudaMemcpyDeviceToHost, s->d2h));
    if (int rc = take_event(s, &cur.done[i])) return rc;
    ORBX_CUDA(cudaEventRecord(cur.done[i], s->d2h));
  }
  if (int rc = finish_pass(s)) return rc;
  s->hist[1] = s->hist[0];
  s->hist[0] = cur;
  s->pipelining = true;
  return ORBX_OK;
}

int orbx_sequence_wait_pass(orbx_sequence* s, int back) {
  if (!s || back < 0 || back > 2) {
    set_error("orbx_sequence_wait_pass: bad argument (back = 0, 1 or 2 passes before the most recent one)");
    return ORBX_E_BADARG;
  }
  if (s->pass_seq <= (unsigned long long)back) return ORBX_OK;  // no such pass was ever issued
  ORBX_CUDA(cudaSetDevice(s->device));
  ORBX_CUDA(cudaEventSynchronize(s->pass_done[(s->pass_seq - 1 - back) & 3]));
  if (back == 0) s->pipelining = false;  // everything issued has finished: idle
  return ORBX_OK;
}

}  // extern "C"
