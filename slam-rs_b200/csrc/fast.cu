// S2 (+ the histogram that S3 needs) — FAST-9/16 corner score, strict 3x3 NMS, 31-px border cull.
//
// Replaces cv::FAST(threshold=20, nonmaxSuppression=true) + KeyPointsFilter::runByImageBorder as run
// per level by cv::ORB_Impl::detectAndCompute (reached from /root/reference/src/feature/detector.rs:22,32).
// Arithmetic per SURVEY.md Appendix A S2/S3: corner <=> 9 contiguous ring pixels all > v+t or all
// < v-t; score = max over the 16 arcs of the min |contrast| over the arc, minus 1; keep strict local
// maxima of the 3x3 score neighbourhood; keep 31 <= x < w-31, 31 <= y < h-31.
//
// One launch covers every level of every frame (tile table in DevGeom).  A block owns a 62x30 interior
// tile and its 64x32 score region (tile + 1 px NMS ring):
//   1. stage the 72x38 source footprint in shared memory, widened to 16x2 pixel pairs (plain + inverted);
//   2. pretest every pixel pair on the 4 compass points (a necessary condition for any 9-arc) and
//      compact the (pair, polarity) work items that pass into a block-wide queue;
//   3. run the 9-of-16 arc network (40 packed 3-input min/max) only on queued items, all lanes busy;
//   4. strict 3x3 NMS on the 16x2 score tile, border cull, emit survivors as packed
//      (score<<24 | y<<12 | x) words: block-local compaction, one global atomicAdd per block to reserve
//      list space, and a per-(frame,level) 256-bin score histogram for retainBest(2N_l).
// Only pixels that can survive the border cull (and their NMS ring) are ever scored.
#include "detector.cuh"

namespace orbx {

constexpr int kScoreH = kFastTileH + 2;  // 32 score rows (score columns: 64 = 32 pixel pairs)
constexpr int kSrcH = kScoreH + 6;       // 38 source rows
constexpr int kPairPitch = 36;           // words (= pixel pairs) per tile row: 72 pixels = 4 + 64 + 4
constexpr int kPlane = 1376;             // words per plane (38 x 36 = 1368, padded: 2 * kPlane % 32 == 0)
constexpr int kScoreOff = 4 * kPlane;    // score tile (32 x 36 words) sits behind the four planes
constexpr int kQueueCap = 2 * kScoreH * 32;  // every pair, both polarities

// ---- packed 16x2 arithmetic: two pixels per instruction (VIMNMX3.S16x2 is native, ALU pipe) ----
__device__ __forceinline__ unsigned min3x2(unsigned a, unsigned b, unsigned c) { return __vimin3_s16x2(a, b, c); }
__device__ __forceinline__ unsigned max3x2(unsigned a, unsigned b, unsigned c) { return __vimax3_s16x2(a, b, c); }

// Shared-memory planes hold the source footprint widened to 16 bits, as pixel pairs: plane E pairs tile
// pixels (2p, 2p+1), plane O pairs (2p+1, 2p+2), so every ring read of two adjacent centres is one aligned
// 32-bit word already in 16x2 form.  Planes 2 and 3 hold 255 - pixel: on them the bright-arc network
// (max over the 16 arcs of the min over the arc) returns 255 - (min over arcs of the max), i.e. the dark
// contrast comes out of the same 40 instructions: score' = arcs(plane) - centre(plane) - t, either polarity.
//
// B points at the centre pair of a work item in plane E (bright) or plane ~E (dark).
__device__ __forceinline__ unsigned fast_arc_score(const unsigned* B, unsigned tk) {
  unsigned R[16];
#define ORBX_RING(K, DX, DY) \
  R[K] = B[(((DX) & 1) ? kPlane + ((DX) - 1) / 2 : (DX) / 2) + (DY) * kPairPitch];
  ORBX_RING(0, 0, 3)
  ORBX_RING(1, 1, 3)
  ORBX_RING(2, 2, 2)
  ORBX_RING(3, 3, 1)
  ORBX_RING(4, 3, 0)
  ORBX_RING(5, 3, -1)
  ORBX_RING(6, 2, -2)
  ORBX_RING(7, 1, -3)
  ORBX_RING(8, 0, -3)
  ORBX_RING(9, -1, -3)
  ORBX_RING(10, -2, -2)
  ORBX_RING(11, -3, -1)
  ORBX_RING(12, -3, 0)
  ORBX_RING(13, -3, 1)
  ORBX_RING(14, -2, 2)
  ORBX_RING(15, -1, 3)
#undef ORBX_RING
  const unsigned v = B[0];
  // sliding min over 9 = 3-input min over windows of 3, then over 3 windows of 3; max over the 16 arcs
  unsigned lo3[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) lo3[k] = min3x2(R[k], R[(k + 1) & 15], R[(k + 2) & 15]);
  unsigned l = 0u;
#pragma unroll
  for (int k = 0; k < 16; k += 2) {
    const unsigned a = min3x2(lo3[k], lo3[(k + 3) & 15], lo3[(k + 6) & 15]);
    const unsigned b = min3x2(lo3[k + 1], lo3[(k + 4) & 15], lo3[(k + 7) & 15]);
    l = max3x2(l, a, b);
  }
  // contrasts are kept biased by +256 per lane so that plain 32-bit adds never borrow across lanes:
  // tk = 256 + t per lane; result = max(contrast - t, 0) per lane (0 = not a corner of this polarity)
  constexpr unsigned kBias = 0x01000100u;
  return __vmaxs2(l + kBias - v, tk) - tk;
}

__global__ void __launch_bounds__(256, 6) fast_nms_kernel(DevGeom g) {
  // planes E, O, ~E, ~O, then the 16x2 score tile (row pitch kPairPitch, word (cy, pe) = score pair pe - 2)
  __shared__ __align__(16) unsigned s_mem[4 * kPlane + kScoreH * kPairPitch];
  __shared__ unsigned s_q[kQueueCap];  // work items: plane byte offset | score byte offset << 16
  uint32_t* const s_list = s_q;        // survivor list of step 4 (<= 31 x 15 strict maxima) reuses the queue
  __shared__ uint32_t s_hist[256];
  __shared__ uint32_t s_n, s_base, s_qn;

  // tile -> (level, first interior pixel)
  const uint32_t tt = __ldg(g.fast_tab + blockIdx.x);
  const int l = tt & 7;
  const int x0 = (tt >> 4) & 0xFFF, y0 = tt >> 16;
  const DevLevel& L = g.lv[l];
  const int frame = blockIdx.y;
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
  if (tid == 0) s_n = 0, s_qn = 0;

  // 1. stage the 72 x 38 footprint, widened to the 16x2 planes, straight from global memory:
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
    unsigned* e = s_mem + r0 * kPairPitch + 2 * c;
#pragma unroll
    for (int it = 0; it < 3; ++it) {
      if (r0 + 14 * it < src_rows) {
        const unsigned e0 = __byte_perm(w[it], 0u, 0x4140);                        // (p0, p1)
        const unsigned e1 = __byte_perm(w[it], 0u, 0x4342);                        // (p2, p3)
        const unsigned o0 = __byte_perm(w[it], 0u, 0x4241);                        // (p1, p2)
        const unsigned o1 = __byte_perm(w[it], nx[it], 0x4443) & 0x00FF00FFu;      // (p3, p4)
        unsigned* q = e + it * 14 * kPairPitch;
        *reinterpret_cast<uint2*>(q) = make_uint2(e0, e1);
        *reinterpret_cast<uint2*>(q + kPlane) = make_uint2(o0, o1);
        *reinterpret_cast<uint2*>(q + 2 * kPlane) = make_uint2(e0 ^ 0x00FF00FFu, e1 ^ 0x00FF00FFu);
        *reinterpret_cast<uint2*>(q + 3 * kPlane) = make_uint2(o0 ^ 0x00FF00FFu, o1 ^ 0x00FF00FFu);
      }
    }
  }
  __syncthreads();

  // 2. compass pretest: any 9-arc holds one pixel of {0, 8} and one of {4, 12}.  Warp w owns score rows
  // 4w .. 4w+3, lane j owns score pair j (columns 2j, 2j+1; tile pair pe = j + 2).  Items that pass go to the
  // block-wide queue (ballot-compacted: one shared atomic per warp and row); the pair's score word is zeroed.
  const int lane = tid & 31, warp = tid >> 5;
  const unsigned lt = (1u << lane) - 1u;
  constexpr unsigned kBias = 0x01000100u;
  const unsigned tk = kBias + (unsigned)g.fast_threshold * 0x00010001u;  // 256 + t per lane
  // pixels with x > w - 31 are never needed (x = w - 31 is the NMS ring of the last kept column)
  const bool col_ok = x0 - 1 + 2 * lane <= Lw - kEdge;
  if (4 * warp < rows_needed) {  // warp-uniform
    const int off0 = (4 * warp + 3) * kPairPitch + lane + 2;  // centre pair of row 4w in plane E
    const unsigned* B0 = s_mem + off0;
    unsigned mb[4], md[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const unsigned* B = B0 + i * kPairPitch;
      const unsigned v = B[0], a = B[3 * kPairPitch], b = B[-3 * kPairPitch], c = B[kPlane + 1], d = B[kPlane - 2];
      const unsigned pb = __vmins2(__vmaxs2(a, b), __vmaxs2(c, d));  // bright needs pb > v + t
      const unsigned pd = __vmaxs2(__vmins2(a, b), __vmins2(c, d));  // dark needs pd < v - t
      const bool ok = col_ok && 4 * warp + i < rows_needed;
      mb[i] = __ballot_sync(0xFFFFFFFFu, ok && __vmaxs2(pb + kBias - v, tk) != tk);
      md[i] = __ballot_sync(0xFFFFFFFFu, ok && __vmaxs2(v + kBias - pd, tk) != tk);
      s_mem[kScoreOff + off0 + (i - 3) * kPairPitch] = 0u;
    }
    int total = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) total += __popc(mb[i]) + __popc(md[i]);
    unsigned at = 0;
    if (lane == 0 && total)  // one reservation per warp (plain PTX: no compiler-side re-aggregation)
      asm volatile("atom.shared.add.u32 %0, [%1], %2;"
                   : "=r"(at)
                   : "r"((unsigned)__cvta_generic_to_shared(&s_qn)), "r"((unsigned)total)
                   : "memory");
    at = __shfl_sync(0xFFFFFFFFu, at, 0);
    // item = plane byte offset of the centre pair | byte offset of its score word << 16
    const unsigned item0 = (unsigned)(off0 * 4) | ((unsigned)((kScoreOff + off0 - 3 * kPairPitch) * 4) << 16);
    constexpr unsigned kRowStep = (unsigned)(kPairPitch * 4) * 0x00010001u;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      if ((mb[i] >> lane) & 1u) s_q[at + __popc(mb[i] & lt)] = item0 + i * kRowStep;
      at += __popc(mb[i]);
      if ((md[i] >> lane) & 1u) s_q[at + __popc(md[i] & lt)] = item0 + i * kRowStep + 2 * kPlane * 4;
      at += __popc(md[i]);
    }
  }
  __syncthreads();

  // 3. arc network on the queued items; a pixel cannot be a bright and a dark corner at once (2 x 9 > 16),
  // so OR-ing the two polarities' results into the score word is the per-lane maximum
  const unsigned nq = s_qn;
  if (nq == 0) return;  // block-uniform: no corner in this tile
#pragma unroll 1
  for (unsigned i = tid; i < nq; i += 256) {
    const unsigned item = s_q[i];
    const unsigned res =
        fast_arc_score(reinterpret_cast<const unsigned*>(reinterpret_cast<const char*>(s_mem) + (item & 0xFFFFu)), tk);
    if (res) atomicOr(reinterpret_cast<unsigned*>(reinterpret_cast<char*>(s_mem) + (item >> 16)), res);
  }
  __syncthreads();

  // 4. strict 3x3 NMS in 16x2 form.  Warp w owns output rows 4w .. 4w+3 (clipped to the interior rows
  // 1 .. 30 and to y < h - 31), lane j the pair (2j, 2j+1); rows roll through registers.  Per row:
  // c = the pair, hs = max(left, right neighbours), h3 = max(hs, c); neighbourhood of row cy =
  // max(h3[cy-1], h3[cy+1], hs[cy]).  Border cull; emit score = s' + t - 1.
  const int sadd = g.fast_threshold - 1;
  const int gxs = x0 - 1 + 2 * lane;  // image x of the pair's first pixel (score column 2 * lane)
  unsigned colmask = 0;               // bit 15 = first pixel may be emitted, bit 31 = second pixel
  if (lane >= 1 && gxs < Lw - kEdge) colmask |= 0x8000u;
  if (lane <= 30 && gxs + 1 < Lw - kEdge) colmask |= 0x80000000u;
  const int cy_lo = max(4 * warp, 1);
  const int cy_hi = min(min(4 * warp + 3, kFastTileH), Lh - kEdge - y0);  // gy = y0 - 1 + cy < Lh - kEdge
  if (cy_lo <= cy_hi) {
    const unsigned* S = s_mem + kScoreOff + lane + 2;
    unsigned c[6], h3[6], hs[6];  // rows 4w-1 .. 4w+4 (rows outside the tile only feed rows that are not emitted)
#pragma unroll
    for (int k = 0; k < 6; ++k) {
      const unsigned* p = S + min(max(4 * warp - 1 + k, 0), kScoreH - 1) * kPairPitch;
      const unsigned pl = p[-1], pc = p[0], pr = p[1];
      c[k] = pc;
      hs[k] = __vmaxs2(__byte_perm(pl, pc, 0x5432), __byte_perm(pc, pr, 0x5432));
      h3[k] = __vmaxs2(hs[k], pc);
    }
    // keep bit per lane: bit 15 / 31 of (nb | 0x8000) - m is clear iff nb < m (no borrow leaves a lane);
    // row i of this warp parks its two bits at 15 - i and 31 - i
    unsigned kacc = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const unsigned nb = max3x2(h3[i], h3[i + 2], hs[i + 1]);
      kacc |= (~((nb | 0x80008000u) - c[i + 1]) & colmask) >> i;
    }
    // rows outside cy_lo .. cy_hi (tile ring, image border) are dropped here: i in [ilo, ihi]
    {
      const int ilo = cy_lo - 4 * warp, ihi = cy_hi - 4 * warp;
      const unsigned rows16 = ((2u << (ihi - ilo)) - 1u) << (15 - ihi);
      kacc &= rows16 * 0x00010001u;
    }
    // survivors are rare (~3 % of pixels): a short divergent loop over the set bits, after all rows
    while (kacc) {
      const int bit = __ffs(kacc) - 1;
      kacc &= kacc - 1;
      const int hi = bit >> 4;
      const int cy = 4 * warp + 15 + 16 * hi - bit;
      const unsigned m = S[cy * kPairPitch];
      const int sv = (int)(hi ? (m >> 16) : (m & 0xFFFFu)) + sadd;
      const uint32_t slot = atomicAdd(&s_n, 1u);
      s_list[slot] = pack_cand(gxs + hi, y0 - 1 + cy, sv);
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
