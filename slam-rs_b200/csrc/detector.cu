// OrbDetector handle: geometry (S0), device scratch, and the batched S1-S9 pipeline.
//
// Replaces OrbDetector::{new, detect, detect_and_compute, max_features}
// (/root/reference/src/feature/detector.rs:15-44), i.e. cv::ORB::create + cv::ORB_Impl::detectAndCompute.
// Geometry per SURVEY.md Appendix A S0: level scale s_l = f32(pow((double)1.2f, l)), level size
// rint(f32(W) * (1/s_l)), per-level quotas from the geometric series in fp32.
//
// HBM layout for a batch of B frames (all buffers owned by the handle, sized at create time for
// max_w x max_h x max_batch):
//   pyr / blr : level-major, then frame-major pitched u8 planes, pitch = align(w_l, 128) so that every
//               row starts on a 128-byte line (coalesced 32-bit staging loads, TMA-legal strides)
//   cand, key : per level [B][cand_cap_l] u32 — FAST candidates and their Harris keys; cand_cap_l is the
//               NMS worst case (one survivor per 2x2 block of the culled interior), so it cannot overflow
//   sel       : per level [B][sel_cap_l] u64 — sorted survivors of retainBest(N_l)
//   spill     : u64 arena for tie floods (a level whose retainBest keeps more than sel_cap_l / the shared-memory sort
//               holds): one frame's worst case + 8192 entries per further frame of the batch
//   counters  : cand_count[B][8], hist[B][8][256], flags — zeroed with one memset per batch
#include <math.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "detector.cuh"

using namespace orbx;

struct orbx_detector {
  orbx_params p;
  int device = 0, max_w = 0, max_h = 0, max_batch = 0;
  cudaStream_t own_stream = nullptr, stream = nullptr;
  // allocations
  uint8_t *pyr = nullptr, *blr = nullptr;
  uint32_t *cand = nullptr, *key = nullptr, *counters = nullptr, *tables = nullptr, *kp_lvl = nullptr;
  unsigned long long *sel = nullptr, *spill = nullptr;
  size_t spill_cap = 0;   // entries of the tie-flood arena
  int* sticky = nullptr;  // [2] device words that survive the per-batch reset (DevGeom::sticky)
  size_t out_rows = 0;    // rows of out_kps / out_desc / kp_lvl (grown on demand: B x cap of the largest call)
  orbx_keypoint* out_kps = nullptr;
  uint8_t* out_desc = nullptr;
  int* out_n = nullptr;  // [max_batch] counts + [1] flags
  uint8_t* bgr_stage = nullptr;  // host BGR frames land here before the ingest kernel (allocated on first use)
  size_t bgr_stage_bytes = 0;
  size_t pyr_bytes = 0, cand_words = 0, sel_words = 0, counter_words = 0, table_words = 0;
  int out_cap = 0, sort_cap = 0;
  // geometry of the current image size
  int cur_w = 0, cur_h = 0, cur_batch = 0;
  int lw[kMaxLevels], lh[kMaxLevels], lpitch[kMaxLevels], quota[kMaxLevels], cand_cap[kMaxLevels], sel_cap[kMaxLevels];
  float lscale[kMaxLevels];
  size_t tab_x[kMaxLevels], tab_y[kMaxLevels];  // word offsets into tables
  size_t tab_fast = 0;                          // FAST tile table (level | x0 << 4 | y0 << 16)
  int rs_smem_w[kMaxLevels], rs_smem_h[kMaxLevels];
  DevGeom geom;
  size_t lvl_off[kMaxLevels];  // byte offset of level l inside pyr / blr for the last batch
  std::vector<int> host_n;
  EventTimer timer;
  // CUDA-graph replay of the single-frame host call (orbx_detector_set_graph): upload, the ~13 dependent launches and
  // the download of one (w, h, cap, descriptors?) configuration captured once and replayed with one graph launch
  struct GraphEntry {
    int w, h, cap, want_desc;
    cudaGraphExec_t exec;
  };
  bool use_graph = false;
  std::vector<GraphEntry> graphs;
  uint8_t* pin_img = nullptr;  // pinned staging: the frame, tightly packed
  uint8_t* pin_out = nullptr;  // pinned staging: n, flags, cap keypoints, cap descriptors
  size_t pin_img_bytes = 0, pin_out_bytes = 0;
  void drop_graphs() {
    for (auto& g : graphs) cudaGraphExecDestroy(g.exec);
    graphs.clear();
  }
};

namespace {

int rint_i(float v) { return (int)nearbyintf(v); }

void level_geometry(int w, int h, int n, int* lw, int* lh, float* ls) {
  const float sf = 1.2f;
  for (int l = 0; l < n; ++l) {
    const float scale = (float)pow((double)sf, (double)l);
    const float inv = 1.0f / scale;
    ls[l] = scale;
    lw[l] = rint_i((float)w * inv);
    lh[l] = rint_i((float)h * inv);
  }
}

void level_quotas(int N, int n, int* q) {
  const float sf = 1.2f;
  const float factor = (float)(1.0 / (double)sf);
  float nd = (float)N * (1.0f - factor) / (1.0f - (float)pow((double)factor, (double)n));
  int sum = 0;
  for (int l = 0; l < n - 1; ++l) {
    q[l] = rint_i(nd);
    sum += q[l];
    nd = nd * factor;
  }
  q[n - 1] = std::max(N - sum, 0);
}

// cv::resize INTER_LINEAR_EXACT coefficients (Appendix A S1): word = ofs << 16 | c1
void resize_table(int S, int D, uint32_t* tab) {
  const double scale = 1.0 / ((double)D / (double)S);
  for (int d = 0; d < D; ++d) {
    const double fv = scale * (d + 0.5) - 0.5;
    const int iv = (int)floor(fv);
    uint32_t ofs = 0, c1 = 0;
    if (iv >= 0 && S > 1) {
      if (iv < S - 1) {
        ofs = (uint32_t)iv;
        c1 = (uint32_t)nearbyint((fv - iv) * 256.0);
      } else {
        ofs = (uint32_t)(S - 1);
      }
    }
    tab[d] = (ofs << 16) | c1;
  }
}

int cand_capacity(int w, int h) {
  const int iw = w - 2 * kEdge, ih = h - 2 * kEdge;
  if (iw <= 0 || ih <= 0) return 1;
  return std::max(1, ((iw + 1) / 2) * ((ih + 1) / 2));
}

int next_pow2(int v) {
  int p = 1;
  while (p < v) p <<= 1;
  return p;
}

template <typename T>
int dev_alloc(T** p, size_t bytes) {
  cudaError_t e = cudaMalloc((void**)p, std::max<size_t>(bytes, 256));
  if (e != cudaSuccess) {
    set_error("cudaMalloc(%zu) failed: %s", bytes, cudaGetErrorString(e));
    return ORBX_E_NOMEM;
  }
  return ORBX_OK;
}

// fill per-size geometry (host side) for (w, h); returns false when the size is unsupported
void compute_size_geometry(orbx_detector* d, int w, int h) {
  const int n = d->p.nlevels;
  level_geometry(w, h, n, d->lw, d->lh, d->lscale);
  level_quotas(d->p.nfeatures, n, d->quota);
  for (int l = 0; l < n; ++l) {
    d->lpitch[l] = (int)align_up((size_t)d->lw[l], 128);
    d->cand_cap[l] = cand_capacity(d->lw[l], d->lh[l]);
    d->sel_cap[l] = std::min(d->cand_cap[l], std::max(2 * d->quota[l], d->quota[l] + 256));
  }
}

}  // namespace

extern "C" {

int orbx_detector_create(const orbx_params* params, int device, int max_w, int max_h, int max_batch,
                         orbx_detector** out) {
  if (!out) {
    set_error("orbx_detector_create: out is NULL");
    return ORBX_E_BADARG;
  }
  *out = nullptr;
  orbx_params p;
  orbx_default_params(&p);
  if (params) p = *params;
  orbx_params def;
  orbx_default_params(&def);
  if (p.nfeatures < 0 || p.fast_threshold < 1 || p.fast_threshold > 254 || p.scale_factor != def.scale_factor ||
      p.nlevels != def.nlevels || p.edge_threshold != def.edge_threshold || p.first_level != 0 || p.wta_k != 2 ||
      p.score_type != ORBX_HARRIS_SCORE || p.patch_size != def.patch_size) {
    set_error("orbx_detector_create: only nfeatures and fast_threshold may differ from ORB::create_def()");
    return ORBX_E_BADARG;
  }
  if (max_w < 16 || max_h < 16 || max_w > 4095 || max_h > 4095 || max_batch < 1) {
    set_error("orbx_detector_create: max_w/max_h must be in [16, 4095], max_batch >= 1");
    return ORBX_E_BADARG;
  }
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || device < 0 || device >= ndev) {
    cudaGetLastError();
    set_error("orbx_detector_create: CUDA device %d unavailable (%s); liborbx has no CPU fallback", device,
              e != cudaSuccess ? cudaGetErrorString(e) : "index out of range");
    return ORBX_E_CUDA;
  }
  ORBX_CUDA(cudaSetDevice(device));
  orbx_detector* d = new orbx_detector();
  d->p = p;
  d->device = device;
  d->max_w = max_w;
  d->max_h = max_h;
  d->max_batch = max_batch;
  memset(&d->geom, 0, sizeof(d->geom));
  compute_size_geometry(d, max_w, max_h);
  const int n = p.nlevels;
  const size_t B = (size_t)max_batch;
  d->pyr_bytes = 0;
  d->cand_words = d->sel_words = d->table_words = 0;
  d->out_cap = 0;
  int max_sel = 1;
  for (int l = 0; l < n; ++l) {
    d->pyr_bytes += B * d->lh[l] * d->lpitch[l];
    d->cand_words += B * d->cand_cap[l];
    d->sel_words += B * d->sel_cap[l];
    d->table_words += (size_t)d->lw[l] + d->lh[l];
    d->table_words += (size_t)div_up(d->lw[l], kFastTileW) * div_up(d->lh[l], kFastTileH);  // >= FAST tiles
    d->out_cap += d->sel_cap[l];
    max_sel = std::max(max_sel, d->sel_cap[l]);
  }
  d->sort_cap = std::min(next_pow2(max_sel), 16384);
  d->counter_words = B * kMaxLevels * (1 + 256 + 1 + 1) + 16;
  size_t worst = 0;  // every level of one frame flooded (sorted in power-of-two regions)
  for (int l = 0; l < n; ++l) worst += (size_t)next_pow2(d->cand_cap[l]);
  d->spill_cap = std::min<size_t>(worst + (B - 1) * std::min<size_t>(worst, 8192), 0x7FFFFFFFu);
  d->out_rows = B * d->out_cap;
  int rc = ORBX_OK;
  if (!rc) rc = dev_alloc(&d->pyr, d->pyr_bytes + 256);
  if (!rc) rc = dev_alloc(&d->blr, d->pyr_bytes + 256);
  if (!rc) rc = dev_alloc(&d->cand, d->cand_words * 4);
  if (!rc) rc = dev_alloc(&d->key, d->cand_words * 4);
  if (!rc) rc = dev_alloc(&d->sel, d->sel_words * 8);
  if (!rc) rc = dev_alloc(&d->spill, d->spill_cap * 8);
  if (!rc) rc = dev_alloc(&d->sticky, 2 * sizeof(int));
  if (!rc && cudaMemset(d->sticky, 0, 2 * sizeof(int)) != cudaSuccess) rc = ORBX_E_CUDA;
  if (!rc) rc = dev_alloc(&d->counters, d->counter_words * 4);
  if (!rc) rc = dev_alloc(&d->tables, d->table_words * 4);
  if (!rc) rc = dev_alloc(&d->kp_lvl, B * d->out_cap * 4);
  if (!rc) rc = dev_alloc(&d->out_kps, B * d->out_cap * sizeof(orbx_keypoint));
  if (!rc) rc = dev_alloc(&d->out_desc, B * d->out_cap * 32);
  if (!rc) rc = dev_alloc(&d->out_n, (B + 1) * sizeof(int));
  if (!rc) {
    e = cudaStreamCreateWithFlags(&d->own_stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) {
      set_error("cudaStreamCreate: %s", cudaGetErrorString(e));
      rc = ORBX_E_CUDA;
    }
  }
  if (rc) {
    orbx_detector_destroy(d);
    return rc;
  }
  d->stream = d->own_stream;
  d->host_n.resize(B + 1);
  d->cur_w = d->cur_h = 0;
  *out = d;
  return ORBX_OK;
}

void orbx_detector_destroy(orbx_detector* d) {
  if (!d) return;
  cudaSetDevice(d->device);
  if (d->stream) cudaStreamSynchronize(d->stream);
  d->timer.destroy();
  d->drop_graphs();
  if (d->pin_img) cudaFreeHost(d->pin_img);
  if (d->pin_out) cudaFreeHost(d->pin_out);
  cudaFree(d->pyr);
  cudaFree(d->blr);
  cudaFree(d->cand);
  cudaFree(d->key);
  cudaFree(d->sel);
  cudaFree(d->spill);
  cudaFree(d->sticky);
  cudaFree(d->counters);
  cudaFree(d->tables);
  cudaFree(d->kp_lvl);
  cudaFree(d->out_kps);
  cudaFree(d->out_desc);
  cudaFree(d->out_n);
  cudaFree(d->bgr_stage);
  if (d->own_stream) cudaStreamDestroy(d->own_stream);
  delete d;
}

int orbx_detector_max_features(const orbx_detector* d) { return d ? d->p.nfeatures : ORBX_E_BADARG; }

int orbx_detector_sync(orbx_detector* d) {
  if (!d) return ORBX_E_BADARG;
  ORBX_CUDA(cudaSetDevice(d->device));
  // overflow of asynchronous calls since the last sync (kernels cannot return an error code): read and clear
  int st[2] = {0, 0};
  ORBX_CUDA(cudaMemcpyAsync(st, d->sticky, sizeof(st), cudaMemcpyDeviceToHost, d->stream));
  ORBX_CUDA(cudaStreamSynchronize(d->stream));
  if (st[0] || st[1]) {
    ORBX_CUDA(cudaMemsetAsync(d->sticky, 0, sizeof(st), d->stream));
    ORBX_CUDA(cudaStreamSynchronize(d->stream));
    if (st[0])
      set_error("orbx detect: %d tied keypoints on one level exhausted the handle's tie arena (smaller batches hold more)",
                st[0]);
    else
      set_error("orbx detect: a frame produced %d keypoints, more than the row capacity of the asynchronous call "
                "(rows were truncated to cap; re-run with cap >= %d)", st[1], st[1]);
    return ORBX_E_CAPACITY;
  }
  return ORBX_OK;
}

int orbx_detector_set_stream(orbx_detector* d, void* stream) {
  if (!d) return ORBX_E_BADARG;
  d->stream = stream ? (cudaStream_t)stream : d->own_stream;
  d->drop_graphs();  // captured for the previous stream's device context; cheap to rebuild
  return ORBX_OK;
}

int orbx_detector_set_graph(orbx_detector* d, int enable) {
  if (!d) return ORBX_E_BADARG;
  d->use_graph = enable != 0;
  if (!d->use_graph) d->drop_graphs();
  return ORBX_OK;
}

int orbx_detector_set_timing(orbx_detector* d, int enable) {
  if (!d) return ORBX_E_BADARG;
  d->timer.enabled = enable != 0;
  return ORBX_OK;
}

int orbx_detector_timing(orbx_detector* d, float* ms_per_stage, int* launches_per_stage, int reset) {
  if (!d) return ORBX_E_BADARG;
  ORBX_CUDA(cudaSetDevice(d->device));
  d->timer.collect();
  for (int i = 0; i < ORBX_STAGE_COUNT; ++i) {
    if (ms_per_stage) ms_per_stage[i] = (float)d->timer.ms[i];
    if (launches_per_stage) launches_per_stage[i] = d->timer.launches[i];
  }
  if (reset) d->timer.reset();
  return ORBX_OK;
}

}  // extern "C"

#include "detector_run.inc"
