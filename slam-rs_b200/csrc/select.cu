// S3-S7 — retainBest(2N_l) by FAST score, Harris response, retainBest(N_l) by Harris, intensity-
// centroid orientation, keypoint assembly.
//
// Replaces KeyPointsFilter::retainBest, HarrisResponses, ICAngles and the pt*=scale loop of
// cv::ORB_Impl::computeKeyPoints (reached from /root/reference/src/feature/detector.rs:22,32).
// Arithmetic per SURVEY.md Appendix A S3-S7: both selections keep *all* entries >= the n-th best
// (ties kept); Harris sums are int32 and the final expression is non-fused fp32; the angle is
// cv::fastAtan2's polynomial in non-fused fp32.
#include "detector.cuh"

namespace orbx {

// order-preserving float -> uint key; every real response maps to a key > 0 (0 = "not selected")
__device__ __forceinline__ uint32_t float_key(float f) {
  const uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float key_float(uint32_t k) {
  return __uint_as_float((k & 0x80000000u) ? (k & 0x7FFFFFFFu) : ~k);
}

// ------------------------------------------------------------------------------------------------
// S3 + S4: threshold from the score histogram, Harris for the survivors
// ------------------------------------------------------------------------------------------------
// Harris response from the 9x9 window held as 9 rows x 3 aligned words (bytes `off`.. of each row)
__device__ __forceinline__ float harris_from_words(const unsigned* wds, unsigned off) {
  // rolling rows of horizontal difference / smoothing (7 columns each)
  int hd[3][7], hs[3][7];
  int a = 0, b = 0, c = 0;
#pragma unroll
  for (int r = 0; r < 9; ++r) {
    const unsigned w0 = wds[r * 3], w1 = wds[r * 3 + 1], w2 = wds[r * 3 + 2];
    const unsigned a0 = __funnelshift_r(w0, w1, 8 * off), a1 = __funnelshift_r(w1, w2, 8 * off), a2 = w2 >> (8 * off);
    int v[9];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      v[i] = (int)((a0 >> (8 * i)) & 0xFFu);
      v[4 + i] = (int)((a1 >> (8 * i)) & 0xFFu);
    }
    v[8] = (int)(a2 & 0xFFu);
    const int slot = r % 3;
#pragma unroll
    for (int i = 0; i < 7; ++i) {
      hd[slot][i] = v[i + 2] - v[i];
      hs[slot][i] = v[i] + 2 * v[i + 1] + v[i + 2];
    }
    if (r >= 2) {
      const int s0 = (r - 2) % 3, s1 = (r - 1) % 3, s2 = slot;
#pragma unroll
      for (int i = 0; i < 7; ++i) {
        const int Ix = hd[s0][i] + 2 * hd[s1][i] + hd[s2][i];
        const int Iy = hs[s2][i] - hs[s0][i];
        a += Ix * Ix;
        b += Iy * Iy;
        c += Ix * Iy;
      }
    }
  }
  const float scale = 1.0f / (4.0f * 7.0f * 255.0f);
  const float s4 = __fmul_rn(__fmul_rn(__fmul_rn(scale, scale), scale), scale);
  const float fa = (float)a, fb = (float)b, fc = (float)c;
  const float t1 = __fsub_rn(__fmul_rn(fa, fb), __fmul_rn(fc, fc));
  const float sum = __fadd_rn(fa, fb);
  const float t2 = __fmul_rn(__fmul_rn(0.04f, sum), sum);
  return __fmul_rn(__fsub_rn(t1, t2), s4);
}

constexpr int kHarrisBlocksPerLevel = 1;
constexpr int kHarrisWarps = 8;
constexpr int kHarrisSlab = 27;  // 27 words per candidate; odd stride = conflict-free per-thread reads
constexpr int kHarrisScan = 128;  // candidates scanned per warp iteration (4 independent loads per lane)
constexpr int kHarrisQueue = 32 + kHarrisScan;

// Largest bin b >= 1 of a 256-bin histogram with sum_{i >= b} hist[i] >= rem (0 if there is none), and the
// count strictly above it.  Block-wide (>= 256 threads), replaces a 256-step serial scan by thread 0.
// scratch: 10 words of shared memory.  Ends with a __syncthreads().
__device__ __forceinline__ void find_bin_desc(const uint32_t* hist, uint32_t rem, uint32_t* scratch, int& bin,
                                              uint32_t& above) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  uint32_t v = 0, incl = 0;
  if (tid < 256) {
    v = hist[255 - tid];  // position j = tid <-> bin 255 - j, so a prefix over j is a suffix over bins
    incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t n = __shfl_up_sync(0xFFFFFFFFu, incl, o);
      if (lane >= o) incl += n;
    }
    if (lane == 31) scratch[warp] = incl;
  }
  __syncthreads();
  if (tid < 256) {
    uint32_t pre = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w)
      if (w < warp) pre += scratch[w];
    incl += pre;
    if (tid == 254) {  // bin 1: default answer = "not reached": bin 0, everything in bins 1..255 above it
      scratch[8] = 0;
      scratch[9] = incl;
    }
  }
  __syncthreads();
  if (tid < 255 && incl >= rem && incl - v < rem) {
    scratch[8] = 255 - tid;
    scratch[9] = incl - v;
  }
  __syncthreads();
  bin = (int)scratch[8];
  above = scratch[9];
}

// Per-warp body of harris_kernel (ALIGNED: the level's base, pitch and frame stride are 4-byte aligned, so window
// words can be fetched with cp.async; otherwise they are assembled from byte loads).
template <bool ALIGNED>
__device__ __forceinline__ void harris_warp_loop(const uint32_t* __restrict__ cand, uint32_t* __restrict__ key,
                                                 const uint8_t* __restrict__ img, int pitch, int thr, uint32_t begin,
                                                 uint32_t end, uint32_t* q, unsigned* win0, int lane) {
  const unsigned lt = (1u << lane) - 1u;

  // Software pipeline per warp: scan (4 coalesced loads in flight per lane) until 32 survivors are queued,
  // issue their window fetch into slab `buf` (27 cp.async per lane, no wait), then compute the batch fetched one
  // round earlier from the other slab.  One issue site and one compute site keep the register count down.
  uint32_t base = begin, qn = 0, qhead = 0;
  int pend_n = 0, buf = 0;
  uint32_t pend_j = 0;
  unsigned pend_boff = 0;
  for (;;) {
    while (qn < 32 && base < end) {
      uint32_t cw4[kHarrisScan / 32];
#pragma unroll
      for (int k = 0; k < kHarrisScan / 32; ++k) {
        const uint32_t i = base + 32 * k + lane;
        cw4[k] = i < end ? cand[i] : 0u;
      }
#pragma unroll
      for (int k = 0; k < kHarrisScan / 32; ++k) {
        const uint32_t i = base + 32 * k + lane;
        const bool pass = i < end && (int)(cw4[k] >> 24) >= thr;
        if (i < end && !pass) key[i] = 0;
        const unsigned bal = __ballot_sync(0xFFFFFFFFu, pass);
        if (pass) q[qhead + qn + __popc(bal & lt)] = i;
        qn += __popc(bal);
      }
      base += kHarrisScan;
      __syncwarp();
    }
    const int take = (int)min(qn, 32u);
    uint32_t j = 0, cw = 0;
    unsigned off0 = 0;  // byte offset of the window's top-left pixel from img
    if (take) {
      // candidates q[qhead .. qhead + take): cooperative window fetch (word i = lane + 32t -> candidate i/27,
      // row (i%27)/3, aligned word (i%27)%3)
      if (lane < take) {
        j = q[qhead + lane];
        cw = cand[j];
        off0 = (unsigned)(((int)((cw >> 12) & 0xFFF) - 4) * pitch + (int)(cw & 0xFFF) - 4);
      }
      unsigned* win = win0 + buf * (32 * kHarrisSlab);
      const unsigned win_sa = (unsigned)__cvta_generic_to_shared(win);
#pragma unroll
      for (int t = 0; t < 27; ++t) {
        const int i = lane + 32 * t;
        const int cidx = i / 27, wd = i - cidx * 27;
        const int wrow = wd / 3, wword = wd - wrow * 3;
        const unsigned o = __shfl_sync(0xFFFFFFFFu, off0, cidx);
        if (cidx < take) {
          if (ALIGNED) {
            // img, pitch and the frame stride are 4-byte aligned: the aligned word holding the pixel
            const uint8_t* src = img + ((o & ~3u) + (unsigned)(wrow * pitch + 4 * wword));
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(win_sa + 4 * i), "l"(src));
          } else {
            const uint8_t* b = img + (o + (unsigned)(wrow * pitch + 4 * wword));
            win[i] = (unsigned)__ldg(b) | ((unsigned)__ldg(b + 1) << 8) | ((unsigned)__ldg(b + 2) << 16) |
                     ((unsigned)__ldg(b + 3) << 24);
          }
        }
      }
      asm volatile("cp.async.commit_group;\n" ::);
      qhead += take;
      qn -= take;
    }
    if (pend_n) {
      if (take)
        asm volatile("cp.async.wait_group 1;\n" ::: "memory");
      else
        asm volatile("cp.async.wait_group 0;\n" ::: "memory");
      __syncwarp();
      if (lane < pend_n) {
        key[pend_j] = float_key(harris_from_words(win0 + (buf ^ 1) * (32 * kHarrisSlab) + lane * kHarrisSlab, pend_boff));
      }
      __syncwarp();
    }
    if (!take) break;  // input exhausted, queue empty, pipeline drained
    pend_n = take;
    pend_j = j;
    pend_boff = ALIGNED ? (off0 & 3u) : 0u;
    buf ^= 1;
    // keep the queue short: move the (< 32) leftovers to the front once the head has advanced
    if (qn < 32 && qhead) {
      uint32_t tmp = 0;
      if (lane < qn) tmp = q[qhead + lane];
      __syncwarp();
      if (lane < qn) q[lane] = tmp;
      qhead = 0;
      __syncwarp();
    }
  }
}

// S3 threshold from the score histogram (parallel suffix scan), then Harris on the survivors.  After the
// threshold every warp works alone on its slice of the candidate list: ballot-compacts survivors into a
// warp-private queue and, 32 at a time, fetches the 9x9 windows with cp.async (word i = lane + 32t ->
// candidate i/27, row (i%27)/3, aligned word (i%27)%3: ~11 cache lines per load instruction, all 27 in
// flight) into its shared-memory slab, then runs one candidate per thread out of shared memory.
__global__ void __launch_bounds__(kHarrisWarps * 32, 2) harris_kernel(DevGeom g) {
  __shared__ uint32_t s_hist[256];
  __shared__ uint32_t s_scratch[10];
  __shared__ uint32_t s_queue[kHarrisWarps][kHarrisQueue];
  extern __shared__ unsigned s_win_dyn[];  // [kHarrisWarps][2][32 * kHarrisSlab] window slabs (54 KB: opt-in size)
  unsigned(*s_win)[2][32 * kHarrisSlab] = reinterpret_cast<unsigned(*)[2][32 * kHarrisSlab]>(s_win_dyn);
  const int l = blockIdx.y, frame = blockIdx.z;
  const DevLevel& L = g.lv[l];
  const size_t fl = (size_t)frame * kMaxLevels + l;
  const uint32_t count = min(g.cand_count[fl], (uint32_t)L.cand_cap);
  if (count == 0) return;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  // S3: thr = largest score s with #(score >= s) >= 2*N_l; everything if the list is short
  const uint32_t n_keep = 2u * (uint32_t)L.quota;
  int thr = 0;
  if (n_keep == 0) {
    thr = 256;  // retainBest(0): nothing survives
  } else if (count > n_keep) {
    s_hist[tid] = g.hist[fl * 256 + tid];
    __syncthreads();
    uint32_t above;
    find_bin_desc(s_hist, n_keep, s_scratch, thr, above);
  }
  const uint32_t* cand = L.cand + (size_t)frame * L.cand_cap;
  uint32_t* key = L.key + (size_t)frame * L.cand_cap;
  const uint8_t* img = L.img + (size_t)frame * L.frame_stride;
  const int pitch = L.pitch;
  const bool aligned = L.aligned4 != 0;

  // slice of this warp (multiple of kHarrisScan candidates)
  const uint32_t nwarps = gridDim.x * kHarrisWarps;
  const uint32_t per_warp = ((count + nwarps - 1) / nwarps + kHarrisScan - 1u) / kHarrisScan * kHarrisScan;
  const uint32_t begin = (blockIdx.x * kHarrisWarps + warp) * per_warp;
  const uint32_t end = min(count, begin + per_warp);
  if (begin >= end) return;
  if (aligned)
    harris_warp_loop<true>(cand, key, img, pitch, thr, begin, end, s_queue[warp], s_win[warp][0], lane);
  else
    harris_warp_loop<false>(cand, key, img, pitch, thr, begin, end, s_queue[warp], s_win[warp][0], lane);
}

int launch_harris(const DevGeom& g, cudaStream_t s) {
  if (g.batch == 0) return ORBX_OK;
  // one block per (level, frame) amortises the threshold scan and keeps the warp pipelines long; small batches
  // get more blocks per level so that the grid still fills the 148 SMs (2 resident blocks each)
  const int lf = g.nlevels * g.batch;
  const int bpl = lf >= 1184 ? kHarrisBlocksPerLevel : (lf >= 592 ? 2 * kHarrisBlocksPerLevel : 4 * kHarrisBlocksPerLevel);
  dim3 grid(bpl, g.nlevels, g.batch);
  constexpr int kSmem = kHarrisWarps * 2 * 32 * kHarrisSlab * (int)sizeof(unsigned);
  ORBX_CUDA(cudaFuncSetAttribute(harris_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
  harris_kernel<<<grid, 256, kSmem, s>>>(g);
  ORBX_LAUNCH_CHECK();
  return ORBX_OK;
}

// ------------------------------------------------------------------------------------------------
// S5: radix select of the N_l-th largest Harris key, compaction, bitonic sort (response desc, y, x)
// ------------------------------------------------------------------------------------------------
constexpr int kSelThreads = 512;

__global__ void __launch_bounds__(kSelThreads) select_kernel(DevGeom g, int sort_cap) {
  extern __shared__ __align__(16) unsigned long long s_sort[];  // sort_cap entries
  __shared__ uint32_t s_hist[256];
  __shared__ uint32_t s_scratch[10];
  __shared__ uint32_t s_nvalid, s_nsel;
  const int l = blockIdx.x, frame = blockIdx.y;
  const DevLevel& L = g.lv[l];
  const size_t fl = (size_t)frame * kMaxLevels + l;
  const uint32_t count = min(g.cand_count[fl], (uint32_t)L.cand_cap);
  const int tid = threadIdx.x;
  if (count == 0 || L.quota == 0) {
    if (tid == 0) {
      g.sel_count[fl] = 0;
      g.sel_off[fl] = kSelInline;
    }
    return;
  }
  const uint32_t* cand = L.cand + (size_t)frame * L.cand_cap;
  const uint32_t* key = L.key + (size_t)frame * L.cand_cap;

  if (tid == 0) {
    s_nvalid = 0;
    s_nsel = 0;
  }
  __syncthreads();
  {
    uint32_t nv = 0;
    for (uint32_t i = tid; i < count; i += kSelThreads) nv += key[i] != 0;
    for (int o = 16; o > 0; o >>= 1) nv += __shfl_xor_sync(0xFFFFFFFFu, nv, o);
    if ((tid & 31) == 0 && nv) atomicAdd(&s_nvalid, nv);
  }
  __syncthreads();
  const uint32_t nvalid = s_nvalid;
  uint32_t thr = 1;  // keep every valid entry
  if (nvalid > (uint32_t)L.quota) {
    // MSB-first 8-bit radix select of the quota-th largest key (bin search = parallel suffix scan)
    uint32_t prefix = 0, remaining = (uint32_t)L.quota;
    for (int pass = 0; pass < 4; ++pass) {
      const int shift = 24 - 8 * pass;
      if (tid < 256) s_hist[tid] = 0;
      __syncthreads();
      const uint32_t mask = pass == 0 ? 0u : (0xFFFFFFFFu << (shift + 8));
      for (uint32_t i = tid; i < count; i += kSelThreads) {
        const uint32_t k = key[i];
        if (k != 0 && (k & mask) == prefix) atomicAdd(&s_hist[(k >> shift) & 0xFF], 1u);
      }
      __syncthreads();
      int b;
      uint32_t above;
      find_bin_desc(s_hist, remaining, s_scratch, b, above);
      prefix |= (uint32_t)b << shift;
      remaining -= above;
    }
    thr = prefix;
  }
  // compaction into the sort buffer
  for (uint32_t i = tid; i < count; i += kSelThreads) {
    const uint32_t k = key[i];
    if (k != 0 && k >= thr) {
      const uint32_t slot = atomicAdd(&s_nsel, 1u);
      if (slot < (uint32_t)sort_cap)
        s_sort[slot] = ((unsigned long long)k << 32) | (unsigned long long)(0xFFFFFFu - (cand[i] & 0xFFFFFFu));
    }
  }
  __syncthreads();
  uint32_t nsel = s_nsel;
  unsigned long long* buf = s_sort;  // where the selection is sorted: shared memory, or the spill arena
  uint32_t off = kSelInline;
  uint32_t n2 = 1;
  while (n2 < nsel) n2 <<= 1;
  if (nsel > (uint32_t)L.sel_cap || nsel > (uint32_t)sort_cap) {
    // Tie flood: OpenCV's retainBest keeps EVERY entry >= the n-th best, so a level of equal responses (checkerboards,
    // binary images) selects far more than N_l.  Such a selection is compacted again and sorted in the handle's global
    // spill arena instead of shared memory; orient_kernel reads it from there (sel_off).
    __syncthreads();  // everyone has read s_nsel
    if (tid == 0) s_scratch[0] = atomicAdd(reinterpret_cast<unsigned*>(g.flags + 1), n2);
    __syncthreads();
    const uint32_t base = s_scratch[0];
    if (base > g.spill_cap || n2 > g.spill_cap - base) {
      // arena exhausted (many flooded frames in one batch): truncate and report ORBX_E_CAPACITY
      if (tid == 0) {
        atomicMax(g.flags, (int)nsel);
        atomicMax(g.sticky, (int)nsel);
      }
      nsel = min(nsel, (uint32_t)min(L.sel_cap, sort_cap));
      n2 = 1;
      while (n2 < nsel) n2 <<= 1;
    } else {
      buf = g.spill + base;
      off = base;
      if (tid == 0) s_nsel = 0;
      __syncthreads();
      for (uint32_t i = tid; i < count; i += kSelThreads) {
        const uint32_t k = key[i];
        if (k != 0 && k >= thr)
          buf[atomicAdd(&s_nsel, 1u)] = ((unsigned long long)k << 32) | (unsigned long long)(0xFFFFFFu - (cand[i] & 0xFFFFFFu));
      }
    }
  }
  for (uint32_t i = nsel + tid; i < n2; i += kSelThreads) buf[i] = 0ull;
  __syncthreads();
  if (off == kSelInline && n2 <= (uint32_t)kSelThreads) {
    // The usual case (a level's selection fits one element per thread): bitonic sort with the element in a register -
    // partner exchange by warp shuffle while the partner is in the same warp (j < 32), through shared memory otherwise:
    // 10 instead of 45 block barriers for 512 keys.  Keys are unique (position in the low word), so the result is the
    // same permutation as the generic loop's below.
    unsigned long long v = (uint32_t)tid < n2 ? s_sort[tid] : 0ull;
    for (uint32_t k = 2; k <= n2; k <<= 1) {
      for (uint32_t j = k >> 1; j > 0; j >>= 1) {
        unsigned long long o;
        if (j >= 32) {
          __syncthreads();  // everyone has read its partner of the stage before
          if ((uint32_t)tid < n2) s_sort[tid] = v;
          __syncthreads();
          o = (uint32_t)tid < n2 ? s_sort[tid ^ j] : 0ull;
        } else {
          o = __shfl_xor_sync(0xFFFFFFFFu, v, (int)j);
        }
        // descending runs where (i & k) == 0: there the lower index of a pair keeps the larger key
        const bool keep_max = (((uint32_t)tid & j) == 0) == (((uint32_t)tid & k) == 0);
        v = keep_max ? max(v, o) : min(v, o);
      }
    }
    unsigned long long* sel = L.sel + (size_t)frame * L.sel_cap;
    if ((uint32_t)tid < nsel) sel[tid] = v;
    if (tid == 0) {
      g.sel_count[fl] = nsel;
      g.sel_off[fl] = off;
    }
    return;
  }
  // bitonic sort, descending (block-private region: __syncthreads orders the global-memory passes too)
  for (uint32_t k = 2; k <= n2; k <<= 1) {
    for (uint32_t j = k >> 1; j > 0; j >>= 1) {
      for (uint32_t i = tid; i < n2; i += kSelThreads) {
        const uint32_t p = i ^ j;
        if (p > i) {
          const unsigned long long a = buf[i], b = buf[p];
          const bool desc = (i & k) == 0;
          if (desc ? (a < b) : (a > b)) {
            buf[i] = b;
            buf[p] = a;
          }
        }
      }
      __syncthreads();
    }
  }
  if (off == kSelInline) {
    unsigned long long* sel = L.sel + (size_t)frame * L.sel_cap;
    for (uint32_t i = tid; i < nsel; i += kSelThreads) sel[i] = s_sort[i];
  }
  if (tid == 0) {
    g.sel_count[fl] = nsel;
    g.sel_off[fl] = off;
  }
}

int launch_select(const DevGeom& g, int sort_cap, cudaStream_t s) {
  if (g.batch == 0) return ORBX_OK;
  const size_t smem = (size_t)sort_cap * sizeof(unsigned long long);
  // the opt-in is per device (a process may drive several GPUs through the `device` argument): set it on every launch
  if (smem > 48 * 1024)
    ORBX_CUDA(cudaFuncSetAttribute(select_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid(g.nlevels, g.batch);
  select_kernel<<<grid, kSelThreads, smem, s>>>(g, sort_cap);
  ORBX_LAUNCH_CHECK();
  return ORBX_OK;
}

// ------------------------------------------------------------------------------------------------
// S6 + S7: one warp per keypoint — IC moments over the radius-15 disc, fastAtan2, keypoint record
// ------------------------------------------------------------------------------------------------
__constant__ int c_umax[16] = {15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3};

__device__ __forceinline__ float fast_atan2_deg(float y, float x) {
  const float s = (float)(180.0 / 3.14159265358979323846);
  const float P1 = __fmul_rn(0.9997878412794807f, s), P3 = __fmul_rn(-0.3258083974640975f, s);
  const float P5 = __fmul_rn(0.1555786518463281f, s), P7 = __fmul_rn(-0.04432655554792128f, s);
  const float eps = 2.220446049250313e-16f;
  const float ax = fabsf(x), ay = fabsf(y);
  float a, c, c2;
  if (ax >= ay) {
    c = __fdiv_rn(ay, __fadd_rn(ax, eps));
    c2 = __fmul_rn(c, c);
    a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(P7, c2), P5), c2), P3), c2), P1), c);
  } else {
    c = __fdiv_rn(ax, __fadd_rn(ay, eps));
    c2 = __fmul_rn(c, c);
    a = __fsub_rn(90.0f,
                  __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(P7, c2), P5), c2), P3), c2), P1), c));
  }
  if (x < 0.0f) a = __fsub_rn(180.0f, a);
  if (y < 0.0f) a = __fsub_rn(360.0f, a);
  return a;
}

constexpr int kOrientWarps = 8;
constexpr int kOrientPerWarp = 16;

__device__ __forceinline__ int dp4a_us(unsigned a, unsigned b_signed_bytes, int c) {
  int d;
  asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b_signed_bytes), "r"(c));
  return d;
}

// One warp per keypoint.  The 31 x 36-byte patch (nine aligned 32-bit words per row) is fetched with
// row-contiguous lanes (word i = lane + 32t -> row i/9, word i%9: ~4 rows per load instruction instead
// of 32 scattered lines) and transposed through a per-warp shared-memory slab; then lane = patch row
// v = lane-15 funnel-shifts its nine words onto the u = -15 origin and takes the two row sums
// (sum u*p and sum p over |u| <= umax[|v|]) with 16 DP4A against per-lane coefficient words that are
// built once per warp; the warp totals are two integer warp reductions.
__global__ void __launch_bounds__(kOrientWarps * 32) orient_kernel(DevGeom g, OutArgs o) {
  const int frame = blockIdx.y;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // level-major prefix over the per-level selections of this frame
  uint32_t pre[kMaxLevels + 1];
  pre[0] = 0;
#pragma unroll
  for (int l = 0; l < kMaxLevels; ++l)
    pre[l + 1] = pre[l] + (l < g.nlevels ? g.sel_count[(size_t)frame * kMaxLevels + l] : 0u);
  const uint32_t total = pre[kMaxLevels];
  const uint32_t limit = min(total, (uint32_t)min(o.cap, o.lvl_cap));
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    // synchronous callers get the needed size (ORBX_E_CAPACITY); on the asynchronous path nothing on the host reads
    // n before using it as a row count, so it is clamped and the overflow is parked for orbx_detector_sync
    o.n[frame] = o.clamp_n ? (int)limit : (int)total;
    if (o.clamp_n && total > limit) atomicMax(g.sticky + 1, (int)total);
  }

  // per-lane coefficient words: byte b of word k <-> u = -15 + 4k + b
  const int v = min(lane, 30) - 15;
  const int um = lane < 31 ? c_umax[abs(v)] : -1;
  unsigned cu[8], cm[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    unsigned wu = 0, wm = 0;
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      const int u = -15 + 4 * k + b;
      if (abs(u) <= um) {
        wu |= ((unsigned)u & 0xFFu) << (8 * b);
        wm |= 1u << (8 * b);
      }
    }
    cu[k] = wu;
    cm[k] = wm;
  }

  __shared__ unsigned s_patch[kOrientWarps][2][31 * 9 + 1];  // two slabs: keypoint it+1 is fetched while it computes
  // staging geometry of this lane: word i = lane + 32t of the 31 x 9 patch
  int st_row[9], st_word[9];
#pragma unroll
  for (int t = 0; t < 9; ++t) {
    const int i = lane + 32 * t;
    st_row[t] = i / 9;
    st_word[t] = i - st_row[t] * 9;
  }

  // keypoint `it` of this warp: look it up and start the fetch of its patch (cp.async, one group per keypoint;
  // an empty group when there is nothing to fetch, so that "all but the newest group" always means "mine")
  struct Item {
    unsigned long long e;
    uint32_t j;
    int l, x, y;
    unsigned off;
    bool valid, aligned;
  };
  auto fetch = [&](int it) {
    Item k;
    k.j = (blockIdx.x * kOrientPerWarp + it) * kOrientWarps + warp;
    k.valid = it < kOrientPerWarp && k.j < limit;
    k.l = 0, k.x = 0, k.y = 0, k.off = 0, k.e = 0, k.aligned = false;
    if (k.valid) {
      uint32_t first = 0;
#pragma unroll
      for (int i = 1; i < kMaxLevels; ++i)
        if (k.j >= pre[i]) {
          k.l = i;
          first = pre[i];
        }
      const DevLevel& L = g.lv[k.l];
      const uint32_t so = g.sel_off[(size_t)frame * kMaxLevels + k.l];
      k.e = (so == kSelInline ? L.sel + (size_t)frame * L.sel_cap : g.spill + so)[k.j - first];
      const uint32_t pos = 0xFFFFFFu - (uint32_t)(k.e & 0xFFFFFFu);
      k.x = (int)(pos & 0xFFF), k.y = (int)(pos >> 12);
      k.aligned = L.aligned4 != 0;
      if (k.aligned) {
        // all rows share the alignment of row 0 (pitch and stride are multiples of 4)
        const int pitch = L.pitch;
        const uint8_t* seg0 = L.img + (size_t)frame * L.frame_stride + (size_t)(k.y - 15) * pitch + (k.x - 15);
        k.off = (unsigned)((uintptr_t)seg0 & 3);
        const uint8_t* base = seg0 - k.off;
        const unsigned sa = (unsigned)__cvta_generic_to_shared(&s_patch[warp][it & 1][lane]);
#pragma unroll
        for (int t = 0; t < 9; ++t)
          if (t < 8 || lane < 279 - 256)
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(sa + 128 * t),
                         "l"(base + (size_t)st_row[t] * pitch + 4 * st_word[t]));
      }
    }
    asm volatile("cp.async.commit_group;\n" ::);
    return k;
  };

  Item cur = fetch(0);
#pragma unroll 1
  for (int it = 0; it < kOrientPerWarp; ++it) {
    const Item nxt = fetch(it + 1);
    if (!cur.valid) return;  // j grows with it: nothing further either
    asm volatile("cp.async.wait_group 1;\n" ::: "memory");
    __syncwarp();
    const DevLevel& L = g.lv[cur.l];
    const int x = cur.x, y = cur.y, l = cur.l;
    unsigned a[8];
    if (cur.aligned) {
      unsigned w[9];
#pragma unroll
      for (int k = 0; k < 9; ++k) w[k] = s_patch[warp][it & 1][min(lane, 30) * 9 + k];
#pragma unroll
      for (int k = 0; k < 8; ++k) a[k] = __funnelshift_r(w[k], w[k + 1], 8 * cur.off);
    } else {
      const uint8_t* seg = L.img + (size_t)frame * L.frame_stride + (size_t)(y + v) * L.pitch + (x - 15);
#pragma unroll
      for (int k = 0; k < 8; ++k)
        a[k] = (unsigned)__ldg(seg + 4 * k) | ((unsigned)__ldg(seg + 4 * k + 1) << 8) |
               ((unsigned)__ldg(seg + 4 * k + 2) << 16) | ((unsigned)__ldg(seg + 4 * k + 3) << 24);
    }
    int ru = 0, rs = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      ru = dp4a_us(a[k], cu[k], ru);
      rs = dp4a_us(a[k], cm[k], rs);
    }
    const int m10 = __reduce_add_sync(0xFFFFFFFFu, ru);
    const int m01 = __reduce_add_sync(0xFFFFFFFFu, v * rs);
    if (lane == 0) {
      const float response = key_float((uint32_t)(cur.e >> 32));
      orbx_keypoint kp;
      kp.x = __fmul_rn((float)x, L.scale);
      kp.y = __fmul_rn((float)y, L.scale);
      kp.size = __fmul_rn(31.0f, L.scale);
      kp.angle = fast_atan2_deg((float)m01, (float)m10);
      kp.response = response;
      kp.octave = l;
      kp.class_id = -1;
      o.kps[(size_t)frame * o.cap + cur.j] = kp;
      o.kp_lvl[(size_t)frame * o.lvl_cap + cur.j] = ((uint32_t)l << 24) | ((uint32_t)y << 12) | (uint32_t)x;
    }
    __syncwarp();  // slab it & 1 is free for keypoint it + 2
    cur = nxt;
  }
}

int launch_orient(const DevGeom& g, const OutArgs& o, cudaStream_t s) {
  if (g.batch == 0) return ORBX_OK;
  const int lim = o.cap < o.lvl_cap ? o.cap : o.lvl_cap;
  dim3 grid(div_up(lim > 0 ? lim : 1, kOrientWarps * kOrientPerWarp), g.batch);
  orient_kernel<<<grid, kOrientWarps * 32, 0, s>>>(g, o);
  ORBX_LAUNCH_CHECK();
  return ORBX_OK;
}

}  // namespace orbx
