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

// ---- packed 16x2 arithmetic: two pixels per instruction (VIMNMX3.S16x2 is native, ALU pipe) ----
__device__ __forceinline__ unsigned min3x2(unsigned a, unsigned b, unsigned c) { return __vimin3_s16x2(a, b, c); }
__device__ __forceinline__ unsigned max3x2(unsigned a, unsigned b, unsigned c) { return __vimax3_s16x2(a, b, c); }

// R[k] = raw ring pixel k of two adjacent centres (one per 16-bit lane, 0..255).  Returns per lane
//   lo = max over the 16 cyclic 9-arcs of (min over the arc)   -> bright contrast = lo - v
//   hi = min over the 16 cyclic 9-arcs of (max over the arc)   -> dark   contrast = v - hi
// Sliding-window min/max over 9 = 3-input ops over windows of 3, then over 3 windows of 3 (80 ops / 2 px).
__device__ __forceinline__ void fast_arcs16x2(const unsigned (&R)[16], unsigned& lo, unsigned& hi) {
  unsigned lo3[16], hi3[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) {
    lo3[k] = min3x2(R[k], R[(k + 1) & 15], R[(k + 2) & 15]);
    hi3[k] = max3x2(R[k], R[(k + 1) & 15], R[(k + 2) & 15]);
  }
  unsigned l = 0u, h = 0x7FFF7FFFu;
#pragma unroll
  for (int k = 0; k < 16; k += 2) {
    const unsigned a = min3x2(lo3[k], lo3[(k + 3) & 15], lo3[(k + 6) & 15]);
    const unsigned b = min3x2(lo3[k + 1], lo3[(k + 4) & 15], lo3[(k + 7) & 15]);
    l = max3x2(l, a, b);
    const unsigned c = max3x2(hi3[k], hi3[(k + 3) & 15], hi3[(k + 6) & 15]);
    const unsigned d = max3x2(hi3[k + 1], hi3[(k + 4) & 15], hi3[(k + 7) & 15]);
    h = min3x2(h, c, d);
  }
  lo = l;
  hi = h;
}

// Shared-memory tiles hold the source footprint widened to 16 bits, twice: copy E pairs pixels
// (2j, 2j+1), copy O pairs (2j+1, 2j+2), so every ring read of two adjacent centres is one aligned
// 32-bit word already in 16x2 form (no PRMT / funnel shift in the hot loop).
constexpr int kPairPitch = 36;  // words (= pixel pairs) per tile row; 72 pixels

// scores (shifted: 0 = not a corner, s' = m - t >= 1 for corners, i.e. score = s' + t - 1) of 4 adjacent
// pixels whose first pixel is tile column tc (multiple of 4) of tile row tr
__device__ __forceinline__ void fast_score4(const unsigned* sE, const unsigned* sO, int tr, int tc, int t, unsigned& sl,
                                            unsigned& sh) {
  const int pe = tc >> 1;  // pair index of (tc, tc+1) in E; O pair (tc+1, tc+2) has the same index
  unsigned Rl[16], Rh[16];
  // ring offset (DX, DY): pixels tc+DX, tc+DX+1 (lo reg) and tc+DX+2, tc+DX+3 (hi reg)
#define ORBX_RING(K, DX, DY)                                                                  \
  {                                                                                           \
    const unsigned* r_ = (((DX) & 1) ? sO : sE) + (tr + (DY)) * kPairPitch + pe;               \
    constexpr int o_ = ((DX) & 1) ? ((DX) - 1) / 2 : (DX) / 2;                                 \
    Rl[K] = r_[o_];                                                                           \
    Rh[K] = r_[o_ + 1];                                                                       \
  }
  ORBX_RING(0, 0, 3)
  ORBX_RING(8, 0, -3)
  ORBX_RING(4, 3, 0)
  ORBX_RING(12, -3, 0)
  const unsigned* c_ = sE + tr * kPairPitch + pe;
  const unsigned vl = c_[0], vh = c_[1];
  // contrasts are kept biased by +256 per lane so that plain 32-bit adds never borrow across lanes
  constexpr unsigned kBias = 0x01000100u;
  const unsigned tk = kBias + (unsigned)t * 0x00010001u;  // 256 + t
  {
    // compass points first: any 9-arc holds one pixel of {0,8} and one of {4,12} — a necessary
    // condition evaluated for the whole warp (big win on smooth images, ~10 % overhead on dense texture)
    unsigned possible = 0;
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      const unsigned* R = half ? Rh : Rl;
      const unsigned v = half ? vh : vl;
      const unsigned pb = __vmins2(__vmaxs2(R[0], R[8]), __vmaxs2(R[4], R[12]));  // bright needs pb > v + t
      const unsigned pd = __vmaxs2(__vmins2(R[0], R[8]), __vmins2(R[4], R[12]));  // dark needs pd < v - t
      const unsigned mb = pb + kBias - v, md = v + kBias - pd;  // 256 + contrast, lanes in 1..511
      possible |= __vmaxs2(__vmaxs2(mb, md), tk) - tk;          // > 0 in some lane?
    }
    if (!__any_sync(0xFFFFFFFFu, possible != 0)) return;  // sl, sh stay 0
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
  unsigned lo, hi;
  fast_arcs16x2(Rl, lo, hi);
  sl = __vmaxs2(__vmaxs2(lo + kBias - vl, vl + kBias - hi), tk) - tk;  // max(m - t, 0) per lane
  fast_arcs16x2(Rh, lo, hi);
  sh = __vmaxs2(__vmaxs2(lo + kBias - vh, vh + kBias - hi), tk) - tk;
}

constexpr int kScPitch = 34;  // words per row of the 16x2 score tile: pad pair, 32 pairs, pad pair

__global__ void __launch_bounds__(256) fast_nms_kernel(DevGeom g) {
  __shared__ __align__(16) unsigned s_E[kSrcH * kPairPitch];
  __shared__ __align__(16) unsigned s_O[kSrcH * kPairPitch];
  __shared__ __align__(16) unsigned s_sc[kScoreH * kScPitch];
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
  const int Lw = L.w, Lh = L.h, pitch = L.pitch;
  const uint8_t* img = L.img + (size_t)frame * L.frame_stride;
  const int tid = threadIdx.x;
  // rows of the score region that can matter: image rows up to h - 31 (NMS ring of the last interior row)
  const int rows_needed = min(kScoreH, Lh - kEdge + 1 - (y0 - 1));  // score rows 0 .. rows_needed-1
  const int src_rows = min(kSrcH, rows_needed + 6);

  s_hist[tid] = 0;
  if (tid == 0) s_n = 0;

  // stage the 72 x 38 footprint, widened to the two 16x2 copies, straight from global memory:
  // thread -> tile word column c = tid % 18 (fixed), rows tid/18, +14, +28
  if (tid < 252) {
    const int c = tid % 18, r0 = tid / 18;
    unsigned w[3], nx[3];  // w = pixels 4c..4c+3 of the tile row, byte 0 of nx = pixel 4c+4
    if (L.aligned4) {
      const int sh = 8 * (gx0 & 3);
      const int ga = (gx0 & ~3) + c * 4;
      const uint8_t* c0 = img + min(ga, pitch - 4);
      const uint8_t* c1 = img + min(ga + 4, pitch - 4);
#pragma unroll
      for (int it = 0; it < 3; ++it) {
        const int r = r0 + 14 * it;
        if (r < src_rows) {
          const unsigned roff = (unsigned)(min(gy0 + r, Lh - 1) * pitch);
          const unsigned w0 = __ldg(reinterpret_cast<const unsigned*>(c0 + roff));
          const unsigned w1 = __ldg(reinterpret_cast<const unsigned*>(c1 + roff));
          w[it] = __funnelshift_r(w0, w1, sh);
          nx[it] = w1 >> sh;
        }
      }
    } else {
      const int gx = gx0 + 4 * c;
      const int b0 = min(gx, Lw - 1), b1 = min(gx + 1, Lw - 1), b2 = min(gx + 2, Lw - 1), b3 = min(gx + 3, Lw - 1),
                b4 = min(gx + 4, Lw - 1);
      for (int it = 0; it < 3; ++it) {
        const int r = r0 + 14 * it;
        if (r < src_rows) {
          const uint8_t* p = img + (size_t)min(gy0 + r, Lh - 1) * pitch;
          w[it] = (unsigned)__ldg(p + b0) | ((unsigned)__ldg(p + b1) << 8) | ((unsigned)__ldg(p + b2) << 16) |
                  ((unsigned)__ldg(p + b3) << 24);
          nx[it] = __ldg(p + b4);
        }
      }
    }
    unsigned* e = s_E + r0 * kPairPitch + 2 * c;
    unsigned* o = s_O + r0 * kPairPitch + 2 * c;
#pragma unroll
    for (int it = 0; it < 3; ++it) {
      if (r0 + 14 * it < src_rows) {
        e[it * 14 * kPairPitch] = __byte_perm(w[it], 0u, 0x4140);                        // (p0, p1)
        e[it * 14 * kPairPitch + 1] = __byte_perm(w[it], 0u, 0x4342);                    // (p2, p3)
        o[it * 14 * kPairPitch] = __byte_perm(w[it], 0u, 0x4241);                        // (p1, p2)
        o[it * 14 * kPairPitch + 1] = __byte_perm(w[it], nx[it], 0x4443) & 0x00FF00FFu;  // (p3, p4)
      }
    }
  }
  __syncthreads();

  // score region: columns x0-1 .. x0+62, rows y0-1 .. y0+30; 4 pixels per thread, 2 rows per thread
  const int c4 = tid & 15;               // group of 4 score columns
  const int gxs = x0 - 1 + c4 * 4;       // image x of the first of the 4 pixels
  const int over = gxs + 3 - (Lw - kEdge);  // pixels with x > w-31 are not scored
#pragma unroll 1
  for (int k = 0; k < 2; ++k) {
    const int cy = (tid >> 4) + k * 16;
    unsigned sl = 0, sh = 0;
    // a warp covers score rows 2w, 2w+1 (+16): skip the ones below the last row that can matter
    if ((cy & ~1) < rows_needed) fast_score4(s_E, s_O, cy + 3, 4 + 4 * c4, g.fast_threshold, sl, sh);
    if (cy >= rows_needed || over >= 4) sl = sh = 0;
    if (over >= 2) sh = 0;
    if (over == 3) sl &= 0xFFFFu;
    if (over == 1) sh &= 0xFFFFu;
    s_sc[cy * kScPitch + 1 + 2 * c4] = sl;
    s_sc[cy * kScPitch + 2 + 2 * c4] = sh;
  }
  __syncthreads();

  // strict 3x3 NMS in the same 16x2 form: horizontal 3-max of the rows above and below plus the two
  // side neighbours, then one compare per pixel; border cull; emit score = s' + t - 1
  const int sadd = g.fast_threshold - 1;
  unsigned colmask0 = 0, colmask1 = 0;  // bit 15 = even pixel of the pair, bit 31 = odd pixel
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int col = 4 * c4 + j;
    if (col >= 1 && col <= kFastTileW && gxs + j < Lw - kEdge) {
      if (j < 2)
        colmask0 |= 0x8000u << (16 * (j & 1));
      else
        colmask1 |= 0x8000u << (16 * (j & 1));
    }
  }
#pragma unroll 1
  for (int k = 0; k < 2; ++k) {
    const int cy = (tid >> 4) + k * 16;
    const int gy = y0 - 1 + cy;
    if (cy < 1 || cy > kFastTileH || gy >= Lh - kEdge) continue;
    const unsigned* base = s_sc + cy * kScPitch + 2 * c4;  // pairs (x-2,x-1) (x,x+1) (x+2,x+3) (x+4,x+5)
    unsigned nb0, nb1, m0, m1;
    {
      const uint2 a = *reinterpret_cast<const uint2*>(base - kScPitch), b = *reinterpret_cast<const uint2*>(base - kScPitch + 2);
      const unsigned l0 = __byte_perm(a.x, a.y, 0x5432), l1 = __byte_perm(a.y, b.x, 0x5432), l2 = __byte_perm(b.x, b.y, 0x5432);
      nb0 = max3x2(l0, a.y, l1);
      nb1 = max3x2(l1, b.x, l2);
    }
    {
      const uint2 a = *reinterpret_cast<const uint2*>(base + kScPitch), b = *reinterpret_cast<const uint2*>(base + kScPitch + 2);
      const unsigned l0 = __byte_perm(a.x, a.y, 0x5432), l1 = __byte_perm(a.y, b.x, 0x5432), l2 = __byte_perm(b.x, b.y, 0x5432);
      nb0 = max3x2(nb0, max3x2(l0, a.y, l1), 0u);
      nb1 = max3x2(nb1, max3x2(l1, b.x, l2), 0u);
    }
    {
      const uint2 a = *reinterpret_cast<const uint2*>(base), b = *reinterpret_cast<const uint2*>(base + 2);
      const unsigned l0 = __byte_perm(a.x, a.y, 0x5432), l1 = __byte_perm(a.y, b.x, 0x5432), l2 = __byte_perm(b.x, b.y, 0x5432);
      nb0 = max3x2(nb0, l0, l1);
      nb1 = max3x2(nb1, l1, l2);
      m0 = a.y;
      m1 = b.x;
    }
    // keep bit per lane: bit 15 / 31 of (nb | 0x8000) - m is clear iff nb < m (no borrow leaves a lane);
    // colmask has the same layout (interior columns 1..62, x < w - 31)
    unsigned k0 = ~((nb0 | 0x80008000u) - m0) & colmask0;
    unsigned k1 = ~((nb1 | 0x80008000u) - m1) & colmask1;
    // survivors are rare (~3 % of pixels): a short divergent loop over the set bits
    unsigned kb = ((k0 >> 15) & 1u) | ((k0 >> 30) & 2u) | ((k1 >> 13) & 4u) | ((k1 >> 28) & 8u);
    while (kb) {
      const int j = __ffs(kb) - 1;
      kb &= kb - 1;
      const unsigned mm = j < 2 ? m0 : m1;
      const int sv = (int)((j & 1) ? (mm >> 16) : (mm & 0xFFFFu)) + sadd;
      const uint32_t slot = atomicAdd(&s_n, 1u);
      s_list[slot] = pack_cand(gxs + j, gy, sv);
      atomicAdd(&s_hist[sv], 1u);
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
