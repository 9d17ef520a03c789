// S2 (+ the histogram that S3 needs) — FAST-9/16 corner score, strict 3x3 NMS, 31-px border cull.
//
// Replaces cv::FAST(threshold=20, nonmaxSuppression=true) + KeyPointsFilter::runByImageBorder as run
// per level by cv::ORB_Impl::detectAndCompute (reached from /root/reference/src/feature/detector.rs:22,32).
// Arithmetic per SURVEY.md Appendix A S2/S3: corner <=> 9 contiguous ring pixels all > v+t or all
// < v-t; score = max over the 16 arcs of the min |contrast| over the arc, minus 1; keep strict local
// maxima of the 3x3 score neighbourhood; keep 31 <= x < w-31, 31 <= y < h-31.
//
// One PERSISTENT launch covers every level of every frame: each block walks the (frame, tile) list with a stride of
// the grid.  A tile is a 62x30 interior region with its 64x32 score region (tile + 1 px NMS ring):
//   0. the 80x38-byte source box of tile i+1 is fetched by the TMA unit (cp.async.bulk.tensor.3d on a per-level
//      tensor map, one thread issues, completion on an mbarrier) into the other of two raw buffers while tile i is
//      processed - no per-thread address arithmetic, no alignment cases, no exposed load latency;
//   1. the raw bytes are widened to two planes of 16x2 pixel pairs (E: pixels (2p, 2p+1), O: (2p+1, 2p+2)) so that
//      every ring read of two adjacent centres is one aligned 32-bit word already in packed form;
//   2. every pixel pair is pretested on the 4 compass points (a necessary condition for any 9-arc); the pairs that
//      pass are compacted into a block-wide queue, bright items from the front, dark items from the back;
//   3. the 9-of-16 arc network (40 packed 3-input min/max) runs only on queued items, all lanes busy; the dark
//      polarity is the same network with min and max exchanged (no inverted copy of the tile is kept);
//   4. strict 3x3 NMS on the 16x2 score tile, border cull, survivors emitted as packed (score<<24 | y<<12 | x)
//      words: block-local compaction, one global atomicAdd per tile to reserve list space, and a per-(frame,level)
//      256-bin score histogram for retainBest(2N_l).
// Only pixels that can survive the border cull (and their NMS ring) are ever scored.
#include <cuda.h>
#include <string.h>

#include <algorithm>

#include "detector.cuh"

namespace orbx {

constexpr int kScoreH = kFastTileH + 2;  // 32 score rows (score columns: 64 = 32 pixel pairs)
constexpr int kSrcH = kScoreH + 6;       // 38 source rows
constexpr int kRawPitch = 80;            // bytes per raw tile row: 72 needed (4 + 64 + 4), padded to the TMA's 16-byte rule
constexpr int kRawBytes = kRawPitch * kSrcH;  // 3040 bytes land per box
constexpr int kRawBuf = 3072;                 // buffer stride: TMA destinations are 128-byte aligned
constexpr int kPairPitch = 36;           // words (= pixel pairs) per plane row
constexpr int kPlane = 1376;             // words per plane (38 x 36 = 1368, padded)
constexpr int kScoreOff = 2 * kPlane;    // score tile (32 x 36 words) sits behind the two planes
constexpr int kQueueCap = 2 * kScoreH * 32;  // every pair, both polarities

// ---- packed 16x2 arithmetic: two pixels per instruction (VIMNMX3.S16x2 is native, ALU pipe) ----
__device__ __forceinline__ unsigned min3x2(unsigned a, unsigned b, unsigned c) { return __vimin3_s16x2(a, b, c); }
__device__ __forceinline__ unsigned max3x2(unsigned a, unsigned b, unsigned c) { return __vimax3_s16x2(a, b, c); }

// B points at the centre pair of a work item in plane E.  BRIGHT: max over the 16 arcs of the min over the arc, minus
// the centre; DARK: the centre minus the min over the arcs of the max over the arc.  Result per 16-bit lane:
// max(contrast - t, 0) (0 = not a corner of this polarity).
template <bool BRIGHT>
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
  // sliding extremum over 9 = 3-input op over windows of 3, then over 3 windows of 3; then over the 16 arcs
  unsigned w3[16];
#pragma unroll
  for (int k = 0; k < 16; ++k)
    w3[k] = BRIGHT ? min3x2(R[k], R[(k + 1) & 15], R[(k + 2) & 15]) : max3x2(R[k], R[(k + 1) & 15], R[(k + 2) & 15]);
  unsigned l = BRIGHT ? 0u : 0x7FFF7FFFu;
#pragma unroll
  for (int k = 0; k < 16; k += 2) {
    if (BRIGHT) {
      const unsigned a = min3x2(w3[k], w3[(k + 3) & 15], w3[(k + 6) & 15]);
      const unsigned b = min3x2(w3[k + 1], w3[(k + 4) & 15], w3[(k + 7) & 15]);
      l = max3x2(l, a, b);
    } else {
      const unsigned a = max3x2(w3[k], w3[(k + 3) & 15], w3[(k + 6) & 15]);
      const unsigned b = max3x2(w3[k + 1], w3[(k + 4) & 15], w3[(k + 7) & 15]);
      l = min3x2(l, a, b);
    }
  }
  // contrasts are kept biased by +256 per lane so that plain 32-bit adds never borrow across lanes:
  // tk = 256 + t per lane; result = max(contrast - t, 0) per lane
  constexpr unsigned kBias = 0x01000100u;
  return __vmaxs2(BRIGHT ? l + kBias - v : v + kBias - l, tk) - tk;
}

// per-level tensor maps of the un-blurred pyramid: dims (x, y, frame), box 80 x 38 x 1 bytes
struct FastMaps {
  CUtensorMap lv[kMaxLevels];
};

__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(256, 6)
fast_nms_kernel(const __grid_constant__ DevGeom g, const __grid_constant__ FastMaps maps) {
  __shared__ __align__(128) uint8_t s_raw[2][kRawBuf];  // TMA destinations (double buffer)
  // planes E, O, then the 16x2 score tile (row pitch kPairPitch, word (cy, pe) = score pair pe - 2)
  __shared__ __align__(16) unsigned s_mem[2 * kPlane + kScoreH * kPairPitch];
  __shared__ unsigned s_q[kQueueCap];  // work items: plane byte offset | score byte offset << 16
  uint32_t* const s_list = s_q;        // survivor list of step 4 (<= 31 x 15 strict maxima) reuses the queue
  __shared__ uint32_t s_hist[256];
  __shared__ __align__(8) unsigned long long s_qn;  // queued items: bright count (low word), dark count (high word)
  __shared__ __align__(8) uint64_t s_bar[2];
  __shared__ uint32_t s_n, s_base;

  const int tid = threadIdx.x;
  const int lane = tid & 31, warp = tid >> 5;
  const long total = (long)g.fast_tiles_total * g.batch;
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_addr(&s_bar[0])));
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_addr(&s_bar[1])));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  // tile t -> (frame, level, first interior pixel); issue its box fetch into raw buffer b
  auto issue = [&](long t, int b) {
    const int frame = (int)(t / g.fast_tiles_total);
    const uint32_t tt = __ldg(g.fast_tab + (int)(t - (long)frame * g.fast_tiles_total));
    const int l = tt & 7;
    const int x0 = (tt >> 4) & 0xFFF, y0 = tt >> 16;
    const uint32_t bar = smem_addr(&s_bar[b]);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(kRawBytes) : "memory");
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(
            smem_addr(s_raw[b])),
        "l"(reinterpret_cast<uint64_t>(&maps.lv[l])), "r"(x0 - 5), "r"(y0 - 4), "r"(frame), "r"(bar)
        : "memory");
  };
  if (tid == 0 && blockIdx.x < total) issue(blockIdx.x, 0);

  constexpr unsigned kBias = 0x01000100u;
  const unsigned tk = kBias + (unsigned)g.fast_threshold * 0x00010001u;  // 256 + t per lane
  const unsigned lt = (1u << lane) - 1u;
  const int sadd = g.fast_threshold - 1;

  int it = 0;
  for (long t = blockIdx.x; t < total; t += gridDim.x, ++it) {
    const int buf = it & 1;
    // the raw buffer about to be refilled was last read in step 1 of the previous iteration (barriers since then)
    if (tid == 0 && t + gridDim.x < total) issue(t + gridDim.x, buf ^ 1);

    const int frame = (int)(t / g.fast_tiles_total);
    const uint32_t tt = __ldg(g.fast_tab + (int)(t - (long)frame * g.fast_tiles_total));
    const int l = tt & 7;
    const int x0 = (tt >> 4) & 0xFFF, y0 = tt >> 16;
    const DevLevel& L = g.lv[l];
    // tile column c <-> image x = x0 - 5 + c (score column 0 = x0 - 1 sits at tile column 4, word aligned);
    // tile row r <-> image y = y0 - 4 + r
    const int Lw = L.w, Lh = L.h;
    // rows of the score region that can matter: image rows up to h - 31 (NMS ring of the last interior row)
    const int rows_needed = min(kScoreH, Lh - kEdge + 1 - (y0 - 1));  // score rows 0 .. rows_needed-1
    const int src_rows = min(kSrcH, rows_needed + 6);

    s_hist[tid] = 0;
    if (tid == 0) s_n = 0, s_qn = 0ull;

    // 0. wait for this tile's box (phase parity = number of earlier uses of the buffer)
    {
      const uint32_t bar = smem_addr(&s_bar[buf]), parity = (it >> 1) & 1;
      uint32_t done;
      do {
        asm volatile(
            "{\n\t"
            ".reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t"
            "}"
            : "=r"(done)
            : "r"(bar), "r"(parity)
            : "memory");
      } while (!done);
    }

    // 1. widen the 72 x 38 raw bytes to the two 16x2 planes: thread -> word column c = tid % 18, rows tid/18, +14, +28
    if (tid < 252) {
      const int c = tid % 18, r0 = tid / 18;
      const unsigned* raw = reinterpret_cast<const unsigned*>(s_raw[buf]) + r0 * (kRawPitch / 4) + c;
      unsigned* e = s_mem + r0 * kPairPitch + 2 * c;
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        if (r0 + 14 * k < src_rows) {
          const unsigned w = raw[k * 14 * (kRawPitch / 4)], nx = raw[k * 14 * (kRawPitch / 4) + 1];
          const unsigned e0 = __byte_perm(w, 0u, 0x4140);                    // (p0, p1)
          const unsigned e1 = __byte_perm(w, 0u, 0x4342);                    // (p2, p3)
          const unsigned o0 = __byte_perm(w, 0u, 0x4241);                    // (p1, p2)
          const unsigned o1 = __byte_perm(w, nx, 0x4443) & 0x00FF00FFu;      // (p3, p4)
          unsigned* q = e + k * 14 * kPairPitch;
          *reinterpret_cast<uint2*>(q) = make_uint2(e0, e1);
          *reinterpret_cast<uint2*>(q + kPlane) = make_uint2(o0, o1);
        }
      }
    }
    __syncthreads();

    // 2. compass pretest: any 9-arc holds one pixel of {0, 8} and one of {4, 12}.  Warp w owns score rows
    // 4w .. 4w+3, lane j owns score pair j (columns 2j, 2j+1; tile pair pe = j + 2).  Items that pass go to the
    // block-wide queue - bright from the front, dark from the back (ballot-compacted: one shared 64-bit atomic per
    // warp); the pair's score word is zeroed.
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
      unsigned nb = 0, nd = 0;
#pragma unroll
      for (int i = 0; i < 4; ++i) nb += __popc(mb[i]), nd += __popc(md[i]);
      unsigned long long at = 0;
      if (lane == 0 && (nb | nd))  // one reservation per warp (plain PTX: no compiler-side re-aggregation)
        asm volatile("atom.shared.add.u64 %0, [%1], %2;"
                     : "=l"(at)
                     : "r"(smem_addr(&s_qn)), "l"((unsigned long long)nb | ((unsigned long long)nd << 32))
                     : "memory");
      at = __shfl_sync(0xFFFFFFFFu, at, 0);
      unsigned ab = (unsigned)at, ad = kQueueCap - 1 - (unsigned)(at >> 32);
      // item = plane byte offset of the centre pair | byte offset of its score word << 16
      const unsigned item0 = (unsigned)(off0 * 4) | ((unsigned)((kScoreOff + off0 - 3 * kPairPitch) * 4) << 16);
      constexpr unsigned kRowStep = (unsigned)(kPairPitch * 4) * 0x00010001u;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        if ((mb[i] >> lane) & 1u) s_q[ab + __popc(mb[i] & lt)] = item0 + i * kRowStep;
        ab += __popc(mb[i]);
        if ((md[i] >> lane) & 1u) s_q[ad - __popc(md[i] & lt)] = item0 + i * kRowStep;
        ad -= __popc(md[i]);
      }
    }
    __syncthreads();

    // 3. arc network on the queued items; a pixel cannot be a bright and a dark corner at once (2 x 9 > 16),
    // so OR-ing the two polarities' results into the score word is the per-lane maximum.  Bright items first, then
    // dark ones: only the warp that straddles the boundary runs both code paths.
    const unsigned nqb = (unsigned)s_qn, nqd = (unsigned)(s_qn >> 32);
    if (nqb + nqd != 0) {  // block-uniform
#pragma unroll 1
      for (unsigned i = tid; i < nqb + nqd; i += 256) {
        unsigned res, item;
        if (i < nqb) {
          item = s_q[i];
          res = fast_arc_score<true>(
              reinterpret_cast<const unsigned*>(reinterpret_cast<const char*>(s_mem) + (item & 0xFFFFu)), tk);
        } else {
          item = s_q[kQueueCap - 1 - (i - nqb)];
          res = fast_arc_score<false>(
              reinterpret_cast<const unsigned*>(reinterpret_cast<const char*>(s_mem) + (item & 0xFFFFu)), tk);
        }
        if (res) atomicOr(reinterpret_cast<unsigned*>(reinterpret_cast<char*>(s_mem) + (item >> 16)), res);
      }
      __syncthreads();

      // 4. strict 3x3 NMS in 16x2 form.  Warp w owns output rows 4w .. 4w+3 (clipped to the interior rows
      // 1 .. 30 and to y < h - 31), lane j the pair (2j, 2j+1); rows roll through registers.  Per row:
      // c = the pair, hs = max(left, right neighbours), h3 = max(hs, c); neighbourhood of row cy =
      // max(h3[cy-1], h3[cy+1], hs[cy]).  Border cull; emit score = s' + t - 1.
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
      if (n != 0) {  // block-uniform
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
    }
    __syncthreads();  // the planes, the queue and the counters are rewritten by the next tile
  }
}

// cuTensorMapEncodeTiled through the runtime's driver entry point (liborbx does not link libcuda)
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_tiled() {
  static EncodeTiledFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      p = nullptr;
    return (EncodeTiledFn)p;
  }();
  return fn;
}

// true if level 0 at (img, pitch, stride) can be described by a tensor map (16-byte rule on base and strides)
bool fast_tma_legal(const uint8_t* img, size_t pitch, size_t stride) {
  return ((((uintptr_t)img) | pitch | stride) & 15) == 0;
}

int launch_fast(const DevGeom& g, cudaStream_t s) {
  if (g.fast_tiles_total == 0 || g.batch == 0) return ORBX_OK;
  EncodeTiledFn enc = encode_tiled();
  if (!enc) {
    set_error("orbx detect: cuTensorMapEncodeTiled is not available from this driver");
    return ORBX_E_CUDA;
  }
  FastMaps maps;
  memset(&maps, 0, sizeof(maps));
  for (int l = 0; l < g.nlevels; ++l) {
    const DevLevel& L = g.lv[l];
    if (L.fast_tiles_x == 0) continue;  // level too small to hold a keypoint: no tile ever names it
    const cuuint64_t dims[3] = {(cuuint64_t)L.w, (cuuint64_t)L.h, (cuuint64_t)g.batch};
    // a one-frame batch may come with frame_stride == 0 from single-image calls: any legal value describes it
    const cuuint64_t fs = g.batch > 1 ? (cuuint64_t)L.frame_stride : (cuuint64_t)L.pitch * (cuuint64_t)L.h;
    const cuuint64_t strides[2] = {(cuuint64_t)L.pitch, fs};
    const cuuint32_t box[3] = {(cuuint32_t)kRawPitch, (cuuint32_t)kSrcH, 1u};
    const cuuint32_t estr[3] = {1u, 1u, 1u};
    const CUresult r = enc(&maps.lv[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<uint8_t*>(L.img), dims, strides, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
      set_error("orbx detect: cuTensorMapEncodeTiled(level %d: %d x %d, pitch %d) -> %d", l, L.w, L.h, L.pitch, (int)r);
      return ORBX_E_CUDA;
    }
  }
  // persistent: six resident blocks per SM, every block strides over the (frame, tile) list
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const long total = (long)g.fast_tiles_total * g.batch;
  const int grid = (int)std::min<long>(total, (long)sms * 6);
  fast_nms_kernel<<<grid, 256, 0, s>>>(g, maps);
  ORBX_LAUNCH_CHECK();
  return ORBX_OK;
}

}  // namespace orbx
