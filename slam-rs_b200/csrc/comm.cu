// liborbx — multi-GPU entry of the C ABI: query-sharded map-to-frame matching over the GPUs of one box.
//
// /root/reference/src/mapping/map.rs:72-118 matches the descriptors of the visible map points against the frame's
// (match_descriptors at :107).  At BASELINE configs[3] sizes (1 M map descriptors, 5000 frame descriptors, 5e9 Hamming
// pairs per frame) the north_star shards the QUERY rows across the GPUs with the train set resident on every GPU:
//
//   orbx_comm_set_train     train set -> device 0, then ONE ncclBroadcast over NVLink / NVSwitch (once per map update)
//   orbx_comm_knn_match2 /  the query rows (host or device-0 memory) are cut into contiguous slices; device i's Hamming
//   orbx_comm_match         kernel reads its slice of the queries from device 0 and stores its DMatch records and counts
//                           STRAIGHT INTO device 0's result buffer through peer-mapped pointers (plain st.global over
//                           NVLink from the kernel's epilogue): no gather collective, no staging, no padding
//
// Single process, one host thread, one stream per device ("single process, ncclCommInitAll").  NCCL is
// bound at run time (dlopen of libnccl.so.2) so that liborbx carries no link-time dependency for single-GPU users.
#include <dlfcn.h>

#include <algorithm>
#include <condition_variable>
#include <functional>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>

#include "common.cuh"

namespace {

using orbx::set_error;

// the five NCCL entry points used, resolved from libnccl.so.2 on first use
struct Nccl {
  void* lib = nullptr;
  int (*CommInitAll)(void** comms, int ndev, const int* devlist) = nullptr;
  int (*CommDestroy)(void* comm) = nullptr;
  int (*GroupStart)() = nullptr;
  int (*GroupEnd)() = nullptr;
  int (*Broadcast)(const void* send, void* recv, size_t count, int datatype, int root, void* comm, cudaStream_t s) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
  bool ok = false;
};

Nccl& nccl() {
  static Nccl n = [] {
    Nccl x;
    x.lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!x.lib) x.lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!x.lib) return x;
    x.CommInitAll = (decltype(x.CommInitAll))dlsym(x.lib, "ncclCommInitAll");
    x.CommDestroy = (decltype(x.CommDestroy))dlsym(x.lib, "ncclCommDestroy");
    x.GroupStart = (decltype(x.GroupStart))dlsym(x.lib, "ncclGroupStart");
    x.GroupEnd = (decltype(x.GroupEnd))dlsym(x.lib, "ncclGroupEnd");
    x.Broadcast = (decltype(x.Broadcast))dlsym(x.lib, "ncclBroadcast");
    x.GetErrorString = (decltype(x.GetErrorString))dlsym(x.lib, "ncclGetErrorString");
    x.ok = x.CommInitAll && x.CommDestroy && x.GroupStart && x.GroupEnd && x.Broadcast;
    return x;
  }();
  return n;
}
constexpr int kNcclUint8 = 1;  // ncclUint8 (nccl.h: ncclInt8 = 0, ncclUint8 = 1)

#define ORBX_NCCL(call)                                                                                      \
  do {                                                                                                       \
    int r_ = (call);                                                                                         \
    if (r_ != 0) {                                                                                           \
      set_error("%s:%d: %s -> NCCL error %d (%s)", __FILE__, __LINE__, #call, r_,                            \
                nccl().GetErrorString ? nccl().GetErrorString(r_) : "?");                                    \
      return ORBX_E_CUDA;                                                                                    \
    }                                                                                                        \
  } while (0)

}  // namespace

// One host thread per rank >= 1 issues that rank's launches: from a single thread the ~6 runtime calls per rank are serial
// (25 us each rank: at 8 GPUs the last rank started 0.18 ms after the first - as long as half of its kernel time).
struct RankWorker {
  std::thread th;
  std::mutex mu;
  std::condition_variable cv;
  std::function<int()> job;
  bool pending = false, done = true, stop = false;
  int rc = 0;
  std::string err;
  void loop(int device) {
    cudaSetDevice(device);
    std::unique_lock<std::mutex> lk(mu);
    for (;;) {
      cv.wait(lk, [&] { return pending || stop; });
      if (stop) return;
      pending = false;
      lk.unlock();
      const int r = job();
      std::string e = r ? orbx_last_error() : "";
      lk.lock();
      rc = r;
      err = std::move(e);
      done = true;
      cv.notify_all();
    }
  }
  void post(std::function<int()> j) {
    {
      std::lock_guard<std::mutex> lk(mu);
      job = std::move(j);
      pending = true;
      done = false;
    }
    cv.notify_all();
  }
  int wait() {
    std::unique_lock<std::mutex> lk(mu);
    cv.wait(lk, [&] { return done; });
    return rc;
  }
  void shutdown() {
    {
      std::lock_guard<std::mutex> lk(mu);
      stop = true;
    }
    cv.notify_all();
    if (th.joinable()) th.join();
  }
};

struct orbx_comm {
  int ndev = 0;
  std::vector<std::unique_ptr<RankWorker>> workers;  // [ndev], entry 0 unused (rank 0 is issued by the calling thread)
  bool virtual_ranks = false;  // the same device listed more than once: ranks share a GPU, no NCCL (tests)
  std::vector<int> dev;
  std::vector<void*> nccl_comm;
  std::vector<orbx_matcher*> m;
  std::vector<cudaStream_t> st;
  std::vector<uint8_t*> train;  // per rank: resident train set
  std::vector<cudaEvent_t> done;
  cudaEvent_t ready = nullptr;  // on rank 0's stream: queries staged / previous call finished
  size_t train_cap = 0;
  int nt = 0;
  // rank-0 staging for host-pointer calls
  uint8_t* q_stage = nullptr;
  orbx_dmatch* out_stage = nullptr;
  int* cnt_stage = nullptr;
  size_t q_cap = 0, out_cap = 0, cnt_cap = 0;
  // timing of the last call (events on rank 0's stream)
  cudaEvent_t t0 = nullptr, t1 = nullptr;
  bool timing = false;
  double ms = 0;
  int calls = 0;
};

namespace {

int grow(orbx_comm* c, void** p, size_t* cap, size_t bytes) {
  if (bytes <= *cap) return ORBX_OK;
  ORBX_CUDA(cudaSetDevice(c->dev[0]));
  ORBX_CUDA(cudaStreamSynchronize(c->st[0]));
  cudaFree(*p);
  *p = nullptr;
  *cap = 0;
  const size_t want = orbx::align_up(bytes + bytes / 4, 256);
  if (cudaMalloc(p, want) != cudaSuccess) {
    cudaGetLastError();
    set_error("orbx_comm: cudaMalloc(%zu) failed", want);
    return ORBX_E_NOMEM;
  }
  *cap = want;
  return ORBX_OK;
}

// contiguous slice of rank r: rows [lo, hi), 8-row aligned boundaries so that slices stay 256-byte aligned
void slice(int nq, int ndev, int r, int* lo, int* hi) {
  const int per = (int)orbx::align_up((size_t)orbx::div_up(nq, ndev), 8);
  *lo = std::min(nq, r * per);
  *hi = std::min(nq, (r + 1) * per);
}

// the resident train set changed: every rank expands its copy once for the tensor-core kernel (in stream order behind
// the broadcast), so that the per-frame calls only expand their query slice
int cache_train(orbx_comm* c) {
  // a slice goes to the tensor-core kernel when slice rows x train rows >= 2^27: below 2^27 / 8192 train rows no
  // sensible slice does, and the expanded copy (320 B per row) would be dead weight
  if (c->nt < 16384) return ORBX_OK;
  for (int r = 0; r < c->ndev; ++r)
    if (int rc = orbx::matcher_cache_train(c->m[r], c->train[r], c->nt)) return rc;
  return ORBX_OK;
}

int run_sharded(orbx_comm* c, int K, const uint8_t* q, int nq, orbx_dmatch* out, int* counts_or_n) {
  if (!c || nq < 0 || (nq > 0 && (!q || !out))) {
    set_error("orbx_comm match: bad argument");
    return ORBX_E_BADARG;
  }
  const bool cnt_host = counts_or_n && !orbx::is_device_ptr(counts_or_n);
  if (nq == 0 || c->nt == 0) {  // matcher.rs:29-31: an empty side is an empty result
    if (K == 1 && counts_or_n) {
      if (cnt_host)
        *counts_or_n = 0;
      else
        ORBX_CUDA(cudaMemsetAsync(counts_or_n, 0, sizeof(int), c->st[0]));
    } else if (K == 2 && counts_or_n && nq > 0) {
      if (cnt_host)
        std::fill(counts_or_n, counts_or_n + nq, 0);
      else
        ORBX_CUDA(cudaMemsetAsync(counts_or_n, 0, sizeof(int) * nq, c->st[0]));
    }
    return ORBX_OK;
  }
  ORBX_CUDA(cudaSetDevice(c->dev[0]));
  cudaStream_t s0 = c->st[0];
  if (c->timing) ORBX_CUDA(cudaEventRecord(c->t0, s0));
  // queries and results live on rank 0 (staged there for host pointers)
  const uint8_t* dq = q;
  if (!orbx::is_device_ptr(q)) {
    if (int rc = grow(c, (void**)&c->q_stage, &c->q_cap, (size_t)nq * 32)) return rc;
    ORBX_CUDA(cudaMemcpyAsync(c->q_stage, q, (size_t)nq * 32, cudaMemcpyHostToDevice, s0));
    dq = c->q_stage;
  }
  const bool out_host = !orbx::is_device_ptr(out);
  orbx_dmatch* dout = out;
  if (out_host) {
    if (int rc = grow(c, (void**)&c->out_stage, &c->out_cap, (size_t)nq * K * sizeof(orbx_dmatch))) return rc;
    dout = c->out_stage;
  }
  int* dcnt = nullptr;
  if (K == 2 && counts_or_n) {
    dcnt = counts_or_n;
    if (cnt_host) {
      if (int rc = grow(c, (void**)&c->cnt_stage, &c->cnt_cap, (size_t)nq * sizeof(int))) return rc;
      dcnt = c->cnt_stage;
    }
  }
  ORBX_CUDA(cudaEventRecord(c->ready, s0));
  // rank r reads its query slice from rank 0's memory and writes its records / counts there (peer-mapped pointers);
  // ranks >= 1 are issued by their own host threads, rank 0 by this one, all at the same time
  auto issue = [c, K, dq, dout, dcnt, nq](int r) -> int {
    int lo, hi;
    slice(nq, c->ndev, r, &lo, &hi);
    if (hi <= lo) return ORBX_OK;
    ORBX_CUDA(cudaSetDevice(c->dev[r]));
    if (r > 0) ORBX_CUDA(cudaStreamWaitEvent(c->st[r], c->ready, 0));
    if (int rc = orbx::match_slice(c->m[r], K, dq + (size_t)lo * 32, hi - lo, c->train[r], c->nt, dout + (size_t)lo * K,
                                   dcnt ? dcnt + lo : nullptr, lo))
      return rc;
    if (r > 0) ORBX_CUDA(cudaEventRecord(c->done[r], c->st[r]));
    return ORBX_OK;
  };
  for (int r = 1; r < c->ndev; ++r) c->workers[r]->post([issue, r] { return issue(r); });
  int rc_all = issue(0);
  for (int r = 1; r < c->ndev; ++r) {
    const int rc = c->workers[r]->wait();
    if (rc && !rc_all) {
      set_error("%s", c->workers[r]->err.c_str());
      rc_all = rc;
    }
  }
  if (rc_all) return rc_all;
  ORBX_CUDA(cudaSetDevice(c->dev[0]));
  for (int r = 1; r < c->ndev; ++r) {
    int lo, hi;
    slice(nq, c->ndev, r, &lo, &hi);
    if (hi > lo) ORBX_CUDA(cudaStreamWaitEvent(s0, c->done[r], 0));
  }
  if (c->timing) {
    ORBX_CUDA(cudaEventRecord(c->t1, s0));
    ++c->calls;
  }
  if (out_host) ORBX_CUDA(cudaMemcpyAsync(out, dout, (size_t)nq * K * sizeof(orbx_dmatch), cudaMemcpyDeviceToHost, s0));
  if (K == 2 && counts_or_n && cnt_host)
    ORBX_CUDA(cudaMemcpyAsync(counts_or_n, dcnt, (size_t)nq * sizeof(int), cudaMemcpyDeviceToHost, s0));
  if (K == 1 && counts_or_n) {
    if (cnt_host)
      *counts_or_n = nq;
    else
      ORBX_CUDA(cudaMemcpyAsync(counts_or_n, &nq, sizeof(int), cudaMemcpyHostToDevice, s0));
  }
  if (out_host || cnt_host || (K == 1 && counts_or_n && !cnt_host) || c->timing) {
    ORBX_CUDA(cudaStreamSynchronize(s0));
    if (c->timing) {
      float t = 0;
      ORBX_CUDA(cudaEventElapsedTime(&t, c->t0, c->t1));
      c->ms += t;
    }
  }
  return ORBX_OK;
}

}  // namespace

extern "C" {

void orbx_comm_destroy(orbx_comm* c) {
  if (!c) return;
  for (auto& w : c->workers)
    if (w) w->shutdown();
  for (int r = 0; r < (int)c->dev.size(); ++r) {
    cudaSetDevice(c->dev[r]);
    if (r < (int)c->st.size() && c->st[r]) cudaStreamSynchronize(c->st[r]);
    if (r < (int)c->m.size() && c->m[r]) orbx_matcher_destroy(c->m[r]);
    if (r < (int)c->train.size()) cudaFree(c->train[r]);
    if (r < (int)c->done.size() && c->done[r]) cudaEventDestroy(c->done[r]);
    if (r < (int)c->st.size() && c->st[r]) cudaStreamDestroy(c->st[r]);
    if (r < (int)c->nccl_comm.size() && c->nccl_comm[r]) nccl().CommDestroy(c->nccl_comm[r]);
  }
  if (!c->dev.empty()) cudaSetDevice(c->dev[0]);
  if (c->ready) cudaEventDestroy(c->ready);
  if (c->t0) cudaEventDestroy(c->t0);
  if (c->t1) cudaEventDestroy(c->t1);
  cudaFree(c->q_stage);
  cudaFree(c->out_stage);
  cudaFree(c->cnt_stage);
  delete c;
}

int orbx_comm_create(const int* devices, int ndev, orbx_comm** out) {
  if (!out || ndev < 1 || ndev > 64) {
    set_error("orbx_comm_create: bad argument");
    return ORBX_E_BADARG;
  }
  *out = nullptr;
  int have = 0;
  if (cudaGetDeviceCount(&have) != cudaSuccess || have < 1) {
    cudaGetLastError();
    set_error("orbx_comm_create: no CUDA device (there is no CPU fallback)");
    return ORBX_E_CUDA;
  }
  auto* c = new orbx_comm();
  c->ndev = ndev;
  for (int r = 0; r < ndev; ++r) {
    const int d = devices ? devices[r] : r;
    if (d < 0 || d >= have) {
      set_error("orbx_comm_create: device %d out of range (%d visible)", d, have);
      delete c;
      return ORBX_E_BADARG;
    }
    for (int k = 0; k < r; ++k)
      if (c->dev[k] == d) c->virtual_ranks = true;
    c->dev.push_back(d);
  }
  c->m.assign(ndev, nullptr);
  c->st.assign(ndev, nullptr);
  c->train.assign(ndev, nullptr);
  c->done.assign(ndev, nullptr);
  c->nccl_comm.assign(ndev, nullptr);
  int rc = ORBX_OK;
  for (int r = 0; r < ndev && !rc; ++r) {
    if (cudaSetDevice(c->dev[r]) != cudaSuccess) rc = ORBX_E_CUDA;
    if (!rc && cudaStreamCreateWithFlags(&c->st[r], cudaStreamNonBlocking) != cudaSuccess) rc = ORBX_E_CUDA;
    if (!rc && cudaEventCreateWithFlags(&c->done[r], cudaEventDisableTiming) != cudaSuccess) rc = ORBX_E_CUDA;
    if (!rc) rc = orbx_matcher_create(c->dev[r], 0, &c->m[r]);
    if (!rc) rc = orbx_matcher_set_stream(c->m[r], c->st[r]);
    // rank r stores into (and loads from) rank 0's memory
    if (!rc && r > 0 && c->dev[r] != c->dev[0]) {
      int can = 0;
      cudaDeviceCanAccessPeer(&can, c->dev[r], c->dev[0]);
      if (!can) {
        set_error("orbx_comm_create: device %d has no peer access to device %d (NVLink / P2P required)", c->dev[r], c->dev[0]);
        orbx_comm_destroy(c);
        return ORBX_E_CUDA;
      }
      cudaError_t e = cudaDeviceEnablePeerAccess(c->dev[0], 0);
      if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) rc = ORBX_E_CUDA;
      cudaGetLastError();
    }
  }
  if (!rc) {
    if (cudaSetDevice(c->dev[0]) != cudaSuccess) rc = ORBX_E_CUDA;
    if (!rc && cudaEventCreateWithFlags(&c->ready, cudaEventDisableTiming) != cudaSuccess) rc = ORBX_E_CUDA;
    if (!rc && cudaEventCreate(&c->t0) != cudaSuccess) rc = ORBX_E_CUDA;
    if (!rc && cudaEventCreate(&c->t1) != cudaSuccess) rc = ORBX_E_CUDA;
  }
  if (rc) {
    if (rc == ORBX_E_CUDA) set_error("orbx_comm_create: CUDA error %s", cudaGetErrorString(cudaGetLastError()));
    orbx_comm_destroy(c);
    return rc;
  }
  c->workers.resize(ndev);
  for (int r = 1; r < ndev; ++r) {
    c->workers[r].reset(new RankWorker());
    RankWorker* w = c->workers[r].get();
    const int d = c->dev[r];
    w->th = std::thread([w, d] { w->loop(d); });
  }
  if (ndev > 1 && !c->virtual_ranks) {
    if (!nccl().ok) {
      set_error("orbx_comm_create: libnccl.so.2 not found (needed for more than one GPU)");
      orbx_comm_destroy(c);
      return ORBX_E_CUDA;
    }
    const int r_ = nccl().CommInitAll(c->nccl_comm.data(), ndev, c->dev.data());
    if (r_ != 0) {
      set_error("orbx_comm_create: ncclCommInitAll -> %d (%s)", r_, nccl().GetErrorString ? nccl().GetErrorString(r_) : "?");
      for (auto& p : c->nccl_comm) p = nullptr;
      orbx_comm_destroy(c);
      return ORBX_E_CUDA;
    }
  }
  *out = c;
  return ORBX_OK;
}

int orbx_comm_size(const orbx_comm* c) { return c ? c->ndev : 0; }

int orbx_comm_set_train(orbx_comm* c, const uint8_t* train, int nt) {
  if (!c || nt < 0 || (nt > 0 && !train)) {
    set_error("orbx_comm_set_train: bad argument");
    return ORBX_E_BADARG;
  }
  const size_t bytes = (size_t)nt * 32;
  if (bytes > c->train_cap) {
    for (int r = 0; r < c->ndev; ++r) {
      ORBX_CUDA(cudaSetDevice(c->dev[r]));
      ORBX_CUDA(cudaStreamSynchronize(c->st[r]));
      cudaFree(c->train[r]);
      c->train[r] = nullptr;
      if (cudaMalloc(&c->train[r], std::max<size_t>(bytes, 256)) != cudaSuccess) {
        cudaGetLastError();
        set_error("orbx_comm_set_train: cudaMalloc(%zu) failed on device %d", bytes, c->dev[r]);
        c->train_cap = 0;
        return ORBX_E_NOMEM;
      }
    }
    c->train_cap = bytes;
  }
  c->nt = nt;
  for (int r = 0; r < c->ndev; ++r) orbx::matcher_cache_train(c->m[r], nullptr, 0);
  if (nt == 0) return ORBX_OK;
  // to rank 0 (host or device source), then one broadcast: every other rank receives over NVLink
  ORBX_CUDA(cudaSetDevice(c->dev[0]));
  ORBX_CUDA(cudaMemcpyAsync(c->train[0], train, bytes, cudaMemcpyDefault, c->st[0]));
  if (c->ndev == 1) return cache_train(c);
  if (c->virtual_ranks) {
    ORBX_CUDA(cudaEventRecord(c->ready, c->st[0]));
    for (int r = 1; r < c->ndev; ++r) {
      ORBX_CUDA(cudaSetDevice(c->dev[r]));
      ORBX_CUDA(cudaStreamWaitEvent(c->st[r], c->ready, 0));
      ORBX_CUDA(cudaMemcpyAsync(c->train[r], c->train[0], bytes, cudaMemcpyDefault, c->st[r]));
    }
    return cache_train(c);
  }
  ORBX_NCCL(nccl().GroupStart());
  for (int r = 0; r < c->ndev; ++r)
    ORBX_NCCL(nccl().Broadcast(c->train[0], c->train[r], bytes, kNcclUint8, 0, c->nccl_comm[r], c->st[r]));
  ORBX_NCCL(nccl().GroupEnd());
  return cache_train(c);
}

int orbx_comm_knn_match2(orbx_comm* c, const uint8_t* q, int nq, orbx_dmatch* out, int* n_per_q) {
  return run_sharded(c, 2, q, nq, out, n_per_q);
}

int orbx_comm_match(orbx_comm* c, const uint8_t* q, int nq, orbx_dmatch* out, int* n) {
  return run_sharded(c, 1, q, nq, out, n);
}

int orbx_comm_sync(orbx_comm* c) {
  if (!c) {
    set_error("orbx_comm_sync: null handle");
    return ORBX_E_BADARG;
  }
  for (int r = c->ndev - 1; r >= 0; --r) {
    ORBX_CUDA(cudaSetDevice(c->dev[r]));
    ORBX_CUDA(cudaStreamSynchronize(c->st[r]));
  }
  return ORBX_OK;
}

int orbx_comm_set_timing(orbx_comm* c, int enable) {
  if (!c) return ORBX_E_BADARG;
  c->timing = enable != 0;
  c->ms = 0;
  c->calls = 0;
  return orbx_matcher_set_timing(c->m[0], enable);
}

int orbx_comm_timing(orbx_comm* c, float* ms, int* calls, float* kernel_ms_rank0, int reset) {
  if (!c) return ORBX_E_BADARG;
  if (ms) *ms = (float)c->ms;
  if (calls) *calls = c->calls;
  if (kernel_ms_rank0) {
    int ln = 0;
    if (int rc = orbx_matcher_timing(c->m[0], kernel_ms_rank0, &ln, reset)) return rc;
  }
  if (reset) c->ms = 0, c->calls = 0;
  return ORBX_OK;
}

}  // extern "C"
