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
constexpr int kSrcH = kScoreH + 6;       // 38
constexpr int kSrcPitch = 80;            // tile row pitch in bytes (72 used: 4 + 64 + 4)

// ring offsets in the shared-memory tile (dy*kSrcPitch + dx); folds to immediates after unrolling
__host__ __device__ constexpr int ring_off(int k) {
  constexpr int dx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
  constexpr int dy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};
  return dy[k] * kSrcPitch + dx[k];
}

// ---- packed 16x2 arithmetic: two pixels per instruction (VIMNMX3.S16x2 / VIADD.16x2 are native) ----
__device__ __forceinline__ unsigned min3x2(unsigned a, unsigned b, unsigned c) { return __vimin3_s16x2(a, b, c); }
__device__ __forceinline__ unsigned max3x2(unsigned a, unsigned b, unsigned c) { return __vimax3_s16x2(a, b, c); }

// D[k] = v - ring_k + 256 per 16-bit lane.  Returns m = max(dark, bright) per lane (signed 16x2), where
//   dark   = max over the 16 cyclic 9-arcs of min over the arc of (v - ring)
//   bright = max over arcs of min of (ring - v) = -(min over arcs of max over the arc of (v - ring))
// Sliding-window min/max over 9 = 3-input ops over windows of 3, then over 3 windows of 3.
__device__ __forceinline__ unsigned fast_m16x2(const unsigned (&D)[16]) {
  unsigned lo3[16], hi3[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) {
    lo3[k] = min3x2(D[k], D[(k + 1) & 15], D[(k + 2) & 15]);
    hi3[k] = max3x2(D[k], D[(k + 1) & 15], D[(k + 2) & 15]);
  }
  unsigned dark = 0u, brightn = 0x7FFF7FFFu;
#pragma unroll
  for (int k = 0; k < 16; k += 2) {
    const unsigned a = min3x2(lo3[k], lo3[(k + 3) & 15], lo3[(k + 6) & 15]);
    const unsigned b = min3x2(lo3[k + 1], lo3[(k + 4) & 15], lo3[(k + 7) & 15]);
    dark = max3x2(dark, a, b);
    const unsigned c = max3x2(hi3[k], hi3[(k + 3) & 15], hi3[(k + 6) & 15]);
    const unsigned d = max3x2(hi3[k + 1], hi3[(k + 4) & 15], hi3[(k + 7) & 15]);
    brightn = min3x2(brightn, c, d);
  }
  // dark - 256 and 256 - brightn per lane
  const unsigned a = __vsub2(dark, 0x01000100u);
  const unsigned b = __vsub2(0x01000100u, brightn);
  return __vmaxs2(a, b);
}

// four horizontally adjacent bytes starting DX columns from the thread's first pixel, taken from the
// three aligned words A (cols -4..-1), B (0..3), C (4..7) of one tile row
template <int DX>
__device__ __forceinline__ unsigned window4(unsigned A, unsigned B, unsigned C) {
  if (DX == 0) return B;
  if (DX < 0) return __funnelshift_r(A, B, 8 * (4 + DX));
  return __funnelshift_r(B, C, 8 * DX);
}
__device__ __forceinline__ unsigned lo_pair(unsigned w) { return __byte_perm(w, 0u, 0x4140); }  // (b0, b1) -> 16x2
__device__ __forceinline__ unsigned hi_pair(unsigned w) { return __byte_perm(w, 0u, 0x4342); }  // (b2, b3)

// scores of 4 adjacent pixels; `row` points at the word holding the 4 centre pixels (tile row pitch in words)
__device__ __forceinline__ unsigned fast_score4(const unsigned* row, int t) {
  constexpr int P = kSrcPitch / 4;
  unsigned Dl[16], Dh[16];
  unsigned vl, vh;
  {
    const unsigned B = row[0];
    vl = lo_pair(B) + 0x01000100u;
    vh = hi_pair(B) + 0x01000100u;
  }
#define ORBX_RING(K, DX, DY)                                                         \
  {                                                                                  \
    const unsigned* r_ = row + (DY) * P;                                             \
    const unsigned w_ = window4<DX>(r_[-1], r_[0], r_[1]);                           \
    Dl[K] = vl - lo_pair(w_);                                                        \
    Dh[K] = vh - hi_pair(w_);                                                        \
  }
  // compass points first: cheap necessary condition, evaluated for the whole warp
  ORBX_RING(0, 0, 3)
  ORBX_RING(8, 0, -3)
  ORBX_RING(4, 3, 0)
  ORBX_RING(12, -3, 0)
  {
    const unsigned thi = 0x01000100u + (unsigned)t * 0x00010001u;  // 256 + t
    const unsigned tlo = 0x01000100u - (unsigned)t * 0x00010001u;  // 256 - t
    unsigned possible = 0;
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      const unsigned* D = half ? Dh : Dl;
      const unsigned pd = __vmins2(__vmaxs2(D[0], D[8]), __vmaxs2(D[4], D[12]));  // dark needs pd > 256+t
      const unsigned pb = __vmaxs2(__vmins2(D[0], D[8]), __vmins2(D[4], D[12]));  // bright needs pb < 256-t
      possible |= __vcmpgts2(pd, thi) | __vcmpgts2(tlo, pb);
    }
    if (!__any_sync(0xFFFFFFFFu, possible != 0)) return 0u;
  }
  ORBX_RING(1, 1, 3)
  ORBX_RING(2, 2, 2)
  ORBX_RING(3, 3, 1)
  ORBX_RING(5, 3, -1)
  ORBX_RING(6, 2, -2)
  ORBX_RING(7, 1, -3)
  ORBX_RING(9, -1, -3)
  ORBX_RING(10, -2, -2)
  ORBX_RING(11, -3, -1)
  ORBX_RING(13, -3, 1)
  ORBX_RING(14, -2, 2)
  ORBX_RING(15, -1, 3)
#undef ORBX_RING
  const unsigned ml = fast_m16x2(Dl), mh = fast_m16x2(Dh);
  // score = m - 1 where m > t, else 0
  const int m0 = (short)(ml & 0xFFFF), m1 = (short)(ml >> 16), m2 = (short)(mh & 0xFFFF), m3 = (short)(mh >> 16);
  const unsigned s0 = m0 > t ? (unsigned)(m0 - 1) : 0u, s1 = m1 > t ? (unsigned)(m1 - 1) : 0u;
  const unsigned s2 = m2 > t ? (unsigned)(m2 - 1) : 0u, s3 = m3 > t ? (unsigned)(m3 - 1) : 0u;
  return s0 | (s1 << 8) | (s2 << 16) | (s3 << 24);
}

__global__ void __launch_bounds__(256) fast_nms_kernel(DevGeom g) {
  __shared__ __align__(16) unsigned s_src[kSrcH * (kSrcPitch / 4)];
  __shared__ __align__(16) uint8_t s_score[kScoreH][kScoreW];
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
  // tile column c <-> image x = x0 - 5 + c (score column 0 = x0 - 1 sits at tile column 4, word aligned);
  // tile row r <-> image y = y0 - 4 + r
  const int gx0 = x0 - 5, gy0 = y0 - 4;
  const uint8_t* img = L.img + (size_t)frame * L.frame_stride;
  const int tid = threadIdx.x;

  s_hist[tid] = 0;
  if (tid == 0) s_n = 0;

  constexpr int wpr = kSrcPitch / 4;  // 20 words per tile row, 18 used
  if (L.aligned4) {
    const int sh = 8 * (gx0 & 3);
    const int ga = gx0 & ~3;
    for (int i = tid; i < 18 * kSrcH; i += 256) {
      const int r = i / 18, c = i - r * 18;
      const int gy = min(gy0 + r, L.h - 1);
      const uint8_t* p = img + (size_t)gy * L.pitch;
      const int a0 = min(ga + c * 4, L.pitch - 4), a1 = min(ga + c * 4 + 4, L.pitch - 4);
      const unsigned w0 = __ldg(reinterpret_cast<const unsigned*>(p + a0));
      const unsigned w1 = __ldg(reinterpret_cast<const unsigned*>(p + a1));
      s_src[r * wpr + c] = __funnelshift_r(w0, w1, sh);
    }
  } else {
    uint8_t* sb = reinterpret_cast<uint8_t*>(s_src);
    for (int i = tid; i < 72 * kSrcH; i += 256) {
      const int r = i / 72, c = i - r * 72;
      const int gy = min(gy0 + r, L.h - 1);
      const int gx = min(gx0 + c, L.w - 1);
      sb[r * kSrcPitch + c] = __ldg(img + (size_t)gy * L.pitch + gx);
    }
  }
  __syncthreads();

  // score region: columns x0-1 .. x0+62, rows y0-1 .. y0+30; 4 pixels per thread, 2 rows per thread
  {
    const int c4 = tid & 15;          // group of 4 score columns
    const int gx = x0 - 1 + c4 * 4;   // image x of the first of the 4 pixels
#pragma unroll 1
    for (int k = 0; k < 2; ++k) {
      const int cy = (tid >> 4) + k * 16;
      const int gy = y0 - 1 + cy;
      unsigned sc = 0;
      // warp-uniform skip of rows/columns that are entirely outside the scored region is not needed:
      // out-of-range pixels are masked below (their tile bytes are clamped copies, never out of bounds)
      sc = fast_score4(s_src + (cy + 3) * wpr + c4 + 1, g.fast_threshold);
      if (gy > L.h - kEdge) sc = 0;
      const int over = gx + 3 - (L.w - kEdge);  // pixels with x > w-31 are not scored
      if (over > 0) sc = over >= 4 ? 0u : (sc & (0xFFFFFFFFu >> (8 * over)));
      *reinterpret_cast<unsigned*>(&s_score[cy][c4 * 4]) = sc;
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
