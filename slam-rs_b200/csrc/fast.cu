// S2 (+ the histogram that S3 needs) — FAST-9/16 corner score, strict 3x3 NMS, 31-px border cull.
//
// Replaces cv::FAST(threshold=20, nonmaxSuppression=true) + KeyPointsFilter::runByImageBorder as run
// per level by cv::ORB_Impl::detectAndCompute (reached from /root/reference/src/feature/detector.rs:22,32).
// Arithmetic per SURVEY.md Appendix A S2/S3: corner <=> 9 contiguous ring pixels all > v+t or all
// < v-t; score = max over the 16 arcs of the min |contrast| over the arc, minus 1; keep strict local
// maxima of the 3x3 score neighbourhood; keep 31 <= x < w-31, 31 <= y < h-31.
//
// One launch covers every level of every frame (tile table in DevGeom).  A block owns a 62x30 interior
// tile: it stages the 70x38 source footprint in shared memory with aligned 32-bit loads, computes the
// 64x32 score region (tile + 1 px NMS ring; 8 pixels per thread), then emits NMS survivors as packed
// (score<<24 | y<<12 | x) words: block-local compaction in shared memory, one global atomicAdd per
// block to reserve list space, and a per-(frame,level) 256-bin score histogram for retainBest(2N_l).
// Only pixels that can survive the border cull (and their NMS ring) are ever scored.
#include "detector.cuh"

namespace orbx {

constexpr int kScoreW = kFastTileW + 2;  // 64
constexpr int kScoreH = kFastTileH + 2;  // 32
constexpr int kSrcW = kScoreW + 6;       // 70
constexpr int kSrcH = kScoreH + 6;       // 38
constexpr int kSrcPitch = 80;            // 70 + up to 3 alignment bytes, rounded to words

// ring offsets in the shared-memory tile (dy*kSrcPitch + dx); folds to immediates after unrolling
__host__ __device__ constexpr int ring_off(int k) {
  constexpr int dx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
  constexpr int dy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};
  return dy[k] * kSrcPitch + dx[k];
}

// max over the 16 cyclic 9-arcs of the min over the arc (sliding-window min via 3-input min: DPX VIMNMX3)
__device__ __forceinline__ int arc9_maxmin(const int (&d)[16]) {
  int t3[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) t3[k] = __vimin3_s32(d[k], d[(k + 1) & 15], d[(k + 2) & 15]);
  int best = -65536;
#pragma unroll
  for (int k = 0; k < 16; k += 2) {
    const int a = __vimin3_s32(t3[k], t3[(k + 3) & 15], t3[(k + 6) & 15]);
    const int b = __vimin3_s32(t3[k + 1], t3[(k + 4) & 15], t3[(k + 7) & 15]);
    best = __vimax3_s32(best, a, b);
  }
  return best;
}

// 16-bit cyclic mask has a run of >= 9 set bits
__device__ __forceinline__ bool has_run9(unsigned m) {
  unsigned x = m | (m << 16);
  x &= x >> 1;  // runs of 2
  x &= x >> 2;  // runs of 4
  x &= x >> 4;  // runs of 8
  x &= (m | (m << 16)) >> 8;  // runs of 9
  return (x & 0xFFFFu) != 0;
}

__device__ __forceinline__ int fast_score(const uint8_t* c, int t) {
  const int v = *c;
  const int hi = v + t, lo = v - t;
  // quick reject: every opposite ring pair must contain a brighter (resp. darker) pixel
  const int p0 = c[ring_off(0)], p8 = c[ring_off(8)];
  bool br = max(p0, p8) > hi, dk = min(p0, p8) < lo;
  if (!(br | dk)) return 0;
  const int p4 = c[ring_off(4)], p12 = c[ring_off(12)];
  br &= max(p4, p12) > hi;
  dk &= min(p4, p12) < lo;
  if (!(br | dk)) return 0;
  int r[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) r[k] = c[ring_off(k)];
  unsigned mb = 0, md = 0;
#pragma unroll
  for (int k = 0; k < 16; ++k) {
    mb |= (r[k] > hi ? 1u : 0u) << k;
    md |= (r[k] < lo ? 1u : 0u) << k;
  }
  const bool cb = br && has_run9(mb);
  const bool cd = dk && has_run9(md);
  if (!(cb | cd)) return 0;
  int m = 0;
  if (cd) {
    int d[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) d[k] = v - r[k];
    m = arc9_maxmin(d);
  }
  if (cb) {
    int d[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) d[k] = r[k] - v;
    m = max(m, arc9_maxmin(d));
  }
  return m - 1;  // m > t >= 1 here
}

__global__ void __launch_bounds__(256) fast_nms_kernel(DevGeom g) {
  __shared__ __align__(16) uint8_t s_src[kSrcH * kSrcPitch];
  __shared__ uint8_t s_score[kScoreH][kScoreW];
  __shared__ uint32_t s_list[(kFastTileW / 2 + 1) * (kFastTileH / 2 + 1) + 32];
  __shared__ uint32_t s_hist[256];
  __shared__ uint32_t s_n, s_base;

  // tile -> (level, tx, ty)
  const int tile = blockIdx.x;
  int l = 0;
#pragma unroll
  for (int i = 1; i < kMaxLevels; ++i)
    if (i < g.nlevels && tile >= g.lv[i].fast_tile_begin) l = i;
  const DevLevel& L = g.lv[l];
  const int frame = blockIdx.y;
  const int lt = tile - L.fast_tile_begin;
  const int ty = lt / L.fast_tiles_x, tx = lt - ty * L.fast_tiles_x;
  const int x0 = kEdge + tx * kFastTileW;  // first interior pixel of the tile
  const int y0 = kEdge + ty * kFastTileH;
  const int sx0 = x0 - 4, sy0 = y0 - 4;    // source footprint origin (ring 3 + NMS 1)
  const int bx = sx0 & ~3;
  const int shift = sx0 - bx;               // 0..3
  const uint8_t* img = L.img + (size_t)frame * L.frame_stride;
  const int tid = threadIdx.x;

  s_hist[tid] = 0;
  if (tid == 0) s_n = 0;

  if (L.aligned4) {
    constexpr int wpr = kSrcPitch / 4;  // 20 words cover shift + 70 bytes
    for (int i = tid; i < wpr * kSrcH; i += 256) {
      const int r = i / wpr, c = i - r * wpr;
      const int gy = min(sy0 + r, L.h - 1);
      const int gx = min(bx + c * 4, (L.pitch - 4));  // clamp inside the row pitch
      const uint32_t v = __ldg(reinterpret_cast<const uint32_t*>(img + (size_t)gy * L.pitch + gx));
      *reinterpret_cast<uint32_t*>(s_src + r * kSrcPitch + c * 4) = v;
    }
  } else {
    for (int i = tid; i < kSrcPitch * kSrcH; i += 256) {
      const int r = i / kSrcPitch, c = i - r * kSrcPitch;
      const int gy = min(sy0 + r, L.h - 1);
      const int gx = min(bx + c, L.w - 1);
      s_src[r * kSrcPitch + c] = __ldg(img + (size_t)gy * L.pitch + gx);
    }
  }
  __syncthreads();

  // score region: (x0-1 .. x0+62) x (y0-1 .. y0+30); valid where 30 <= x <= w-31, 30 <= y <= h-31
  {
    const int cx = tid & 63;
    const int gx = x0 - 1 + cx;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int cy = (tid >> 6) + k * 4;
      const int gy = y0 - 1 + cy;
      int sc = 0;
      if (gx <= L.w - kEdge && gy <= L.h - kEdge)
        sc = fast_score(s_src + (cy + 3) * kSrcPitch + shift + cx + 3, g.fast_threshold);
      s_score[cy][cx] = (uint8_t)sc;
    }
  }
  __syncthreads();

  // NMS + border cull over the 62x30 interior
  for (int i = tid; i < kFastTileW * kFastTileH; i += 256) {
    const int iy = i / kFastTileW, ix = i - iy * kFastTileW;
    const int gx = x0 + ix, gy = y0 + iy;
    const int s = s_score[iy + 1][ix + 1];
    if (s == 0 || gx >= L.w - kEdge || gy >= L.h - kEdge) continue;
    const bool keep = s > s_score[iy][ix] && s > s_score[iy][ix + 1] && s > s_score[iy][ix + 2] &&
                      s > s_score[iy + 1][ix] && s > s_score[iy + 1][ix + 2] && s > s_score[iy + 2][ix] &&
                      s > s_score[iy + 2][ix + 1] && s > s_score[iy + 2][ix + 2];
    if (keep) {
      const uint32_t slot = atomicAdd(&s_n, 1u);
      s_list[slot] = pack_cand(gx, gy, s);
      atomicAdd(&s_hist[s], 1u);
    }
  }
  __syncthreads();
  const uint32_t n = s_n;
  if (n == 0) return;
  const size_t fl = (size_t)frame * kMaxLevels + l;
  if (tid == 0) s_base = atomicAdd(&g.cand_count[fl], n);
  if (s_hist[tid]) atomicAdd(&g.hist[fl * 256 + tid], s_hist[tid]);
  __syncthreads();
  uint32_t* dst = L.cand + (size_t)frame * L.cand_cap;
  for (uint32_t i = tid; i < n; i += 256) {
    const uint32_t at = s_base + i;
    if (at < (uint32_t)L.cand_cap) dst[at] = s_list[i];
  }
}

int launch_fast(const DevGeom& g, cudaStream_t s) {
  if (g.fast_tiles_total == 0 || g.batch == 0) return ORBX_OK;
  dim3 grid(g.fast_tiles_total, g.batch);
  fast_nms_kernel<<<grid, 256, 0, s>>>(g);
  ORBX_LAUNCH_CHECK();
  return ORBX_OK;
}

}  // namespace orbx
