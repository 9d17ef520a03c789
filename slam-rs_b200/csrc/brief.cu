// S8 + S9 — 7x7 sigma=2 Gaussian blur of every level and steered rBRIEF-256 descriptors.
//
// Replaces GaussianBlur(level, 7x7, 2, 2, BORDER_REFLECT_101) and computeOrbDescriptors of
// cv::ORB_Impl::detectAndCompute (reached from /root/reference/src/feature/detector.rs:32).
// Arithmetic per SURVEY.md Appendix A S8/S9: separable fp32 filter with FMA accumulation (row pass
// sequential, column pass over symmetric pairs), round-half-even to u8; pattern points rotated in
// non-fused fp32 by (float)cos/sin((double)angle), rounded half-even, compared on the blurred level,
// bits packed LSB first.
#include "detector.cuh"

namespace orbx {

__constant__ float c_gauss[7] = {0x1.1f5f62p-4f, 0x1.0c70fcp-3f, 0x1.869472p-3f, 0x1.ba95c0p-3f,
                                 0x1.869472p-3f, 0x1.0c70fcp-3f, 0x1.1f5f62p-4f};

// rBRIEF bit_pattern_31_ (first 256 pairs): ax, ay, bx, by  — SURVEY.md Appendix B
__constant__ char4 c_pattern[256] = {
#include "brief_pattern.inc"
};

__device__ __forceinline__ int reflect101(int i, int n) {
  if (i < 0) i = -i;
  if (i >= n) i = 2 * n - 2 - i;
  return min(max(i, 0), n - 1);
}

// Blur tile = 128 x 35 output pixels owned by ONE warp (4 independent warps per block, no block barrier):
// the warp stages the 136 x 41 byte source footprint in its private shared-memory slab - every aligned word
// that lies inside the row's allocation goes global -> shared with cp.async (all rows in flight at once),
// the reflect-101 columns of border tiles are then patched shared -> shared by lane = row - and every lane
// walks down its 4-pixel column: per source row 3 LDS.32 -> 10 bytes -> fp32 (PRMT into the mantissa of 2^23,
// one FADD), the 7-tap row pass for 4 pixels (28 FMA-pipe ops), and - from a 7-row register window - the
// symmetric column pass, round-half-even by the 1.5*2^23 magic add, and one coalesced 32-bit store.
// 41 = 6 + 5 * 7 source rows: a full tile runs the 7-way unrolled steady loop without per-row bounds checks.
// FMA-pipe / issue bound at ~19 fp32 ops per pixel; HBM traffic is read P + write P.
constexpr int kBlurWarps = 4;
constexpr int kBlurSrcRows = kBlurTileH + 6;       // 41
constexpr int kBlurSrcWords = kBlurTileW / 4 + 2;  // 34 words: image columns x0-4 .. x0+131
static_assert(kBlurTileH % 7 == 0, "the steady loop is unrolled over the 7-row register window");

__device__ __forceinline__ float byte_to_float(unsigned w, int i) {
  return __fsub_rn(__uint_as_float(__byte_perm(w, 0x4B000000u, 0x7650 + i)), 8388608.0f);
}

// row pass (OpenCV order) of the 4 pixels of this lane: acc = k0*p[x-3]; acc = fma(k_j, p[x-3+j], acc), j = 1..6
__device__ __forceinline__ void blur_row_pass(const unsigned* p, float (&r)[4], float k0, float k1, float k2, float k3) {
  const unsigned w0 = p[0], w1 = p[1], w2 = p[2];
  float f[10];
  f[0] = byte_to_float(w0, 1);
  f[1] = byte_to_float(w0, 2);
  f[2] = byte_to_float(w0, 3);
#pragma unroll
  for (int i = 0; i < 4; ++i) f[3 + i] = byte_to_float(w1, i);
  f[7] = byte_to_float(w2, 0);
  f[8] = byte_to_float(w2, 1);
  f[9] = byte_to_float(w2, 2);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float acc = __fmul_rn(k0, f[i]);
    acc = __fmaf_rn(k1, f[i + 1], acc);
    acc = __fmaf_rn(k2, f[i + 2], acc);
    acc = __fmaf_rn(k3, f[i + 3], acc);
    acc = __fmaf_rn(k2, f[i + 4], acc);
    acc = __fmaf_rn(k1, f[i + 5], acc);
    acc = __fmaf_rn(k0, f[i + 6], acc);
    r[i] = acc;
  }
}

// the walk down the slab: 6 rows of row pass, then `nout` steps of (row pass of source row 6 + o, column pass,
// store of output row o).  FULL: nout == kBlurTileH, no bounds checks.
template <bool FULL>
__device__ __forceinline__ void blur_walk(const unsigned* p, uint8_t* out, int blur_pitch, int nout, bool store) {
  const float k0 = c_gauss[0], k1 = c_gauss[1], k2 = c_gauss[2], k3 = c_gauss[3];
  float R[7][4];
#pragma unroll
  for (int u = 0; u < 6; ++u) blur_row_pass(p + u * kBlurSrcWords, R[u], k0, k1, k2, k3);
  p += 6 * kBlurSrcWords;
#pragma unroll 1
  for (int ob = 0; ob < (FULL ? kBlurTileH : nout); ob += 7) {
#pragma unroll
    for (int u = 0; u < 7; ++u) {
      if (FULL || ob + u < nout) {
        constexpr int kDummy = 0;
        (void)kDummy;
        const int s = (6 + u) % 7;  // window slot of source row 6 + ob + u (ob is a multiple of 7)
        blur_row_pass(p + u * kBlurSrcWords, R[s], k0, k1, k2, k3);
        // column pass over window rows s-6 .. s (centre s-3, indices mod 7), symmetric pairs
        unsigned b[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          float acc = __fmul_rn(k3, R[(s + 4) % 7][i]);
          acc = __fmaf_rn(k2, __fadd_rn(R[(s + 5) % 7][i], R[(s + 3) % 7][i]), acc);
          acc = __fmaf_rn(k1, __fadd_rn(R[(s + 6) % 7][i], R[(s + 2) % 7][i]), acc);
          acc = __fmaf_rn(k0, __fadd_rn(R[s][i], R[(s + 1) % 7][i]), acc);
          // round half to even: acc in [0, 255.01) so the low byte of acc + 1.5*2^23 is rint(acc)
          b[i] = __float_as_uint(__fadd_rn(acc, 12582912.0f));
        }
        if (store) {
          const unsigned lo = __byte_perm(b[0], b[1], 0x0040), hi = __byte_perm(b[2], b[3], 0x0040);
          *reinterpret_cast<unsigned*>(out) = __byte_perm(lo, hi, 0x5410);
        }
        out += blur_pitch;
      }
    }
    p += 7 * kBlurSrcWords;
  }
}

__global__ void __launch_bounds__(kBlurWarps * 32, 6) blur7_kernel(DevGeom g) {
  __shared__ unsigned s_src[kBlurWarps][kBlurSrcRows * kBlurSrcWords];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int tile = blockIdx.x * kBlurWarps + warp;
  if (tile >= g.blur_tiles_total) return;
  int l = 0;
#pragma unroll
  for (int i = 1; i < kMaxLevels; ++i)
    if (i < g.nlevels && tile >= g.lv[i].blur_tile_begin) l = i;
  const DevLevel& L = g.lv[l];
  const int frame = blockIdx.y;
  const int lt = tile - L.blur_tile_begin;
  const int ty = lt / L.blur_tiles_x, tx = lt - ty * L.blur_tiles_x;
  const int x0 = tx * kBlurTileW, y0 = ty * kBlurTileH;
  const int w = L.w, h = L.h, pitch = L.pitch, blur_pitch = L.blur_pitch;
  const uint8_t* img = L.img + (size_t)frame * L.frame_stride;
  uint8_t* out = L.blur + (size_t)frame * L.blur_stride + (size_t)y0 * blur_pitch + x0 + 4 * lane;
  unsigned* slab = s_src[warp];

  // staging: lane owns word column `lane` (image x0-4+4*lane ..+3); lanes 0,1 also own columns 32,33
  const int nout = min(kBlurTileH, h - y0);  // output rows inside the image
  const int nrows = nout + 6;
  const int xa = x0 - 4 + 4 * lane, xb = xa + 128;
  if (L.aligned4 && w >= 16) {
    // words inside the row's allocation (they may run past w into the pitch padding: those bytes are either
    // patched below or only feed pad columns); rows are reflected here, columns afterwards
    const bool fa = xa >= 0 && xa + 3 < pitch && xa < w;
    const bool fb = lane < 2 && xb + 3 < pitch && xb < w;
    const unsigned sa = (unsigned)__cvta_generic_to_shared(slab + lane);
    if (y0 >= 3 && y0 + nout + 3 <= h) {  // no vertical reflection: one pointer walks down
      const uint8_t* ga = img + ((ptrdiff_t)(y0 - 3) * pitch + xa);
      unsigned d = sa;
      for (int r = 0; r < nrows; ++r) {
        if (fa) asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(d), "l"(ga));
        if (fb) asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(d + 128), "l"(ga + 128));
        ga += pitch;
        d += kBlurSrcWords * 4;
      }
    } else {
      for (int r = 0; r < nrows; ++r) {
        const uint8_t* ga = img + (ptrdiff_t)(reflect101(y0 - 3 + r, h) * pitch + xa);  // xa may be -4
        const unsigned d = sa + r * (kBlurSrcWords * 4);
        if (fa) asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(d), "l"(ga));
        if (fb) asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(d + 128), "l"(ga + 128));
      }
    }
    asm volatile("cp.async.commit_group;\n" ::);
    asm volatile("cp.async.wait_group 0;\n" ::: "memory");
    __syncwarp();
    // reflect-101 columns of border tiles, shared -> shared, lane = row
    const bool left = x0 == 0, right = x0 + kBlurTileW + 3 >= w;
    if (left || right) {
      for (int r = lane; r < nrows; r += 32) {
        unsigned* row = slab + r * kBlurSrcWords;
        // x = -4..-1 mirror x = 4, 3, 2, 1: byte 0 of word 2, bytes 3, 2, 1 of word 1
        if (left) row[0] = __byte_perm(row[1], row[2], 0x1234);
        if (right) {
          uint8_t* rb = reinterpret_cast<uint8_t*>(row);
          const int last = min(w + 2, x0 + kBlurTileW + 3);  // last column any in-image output reads
          for (int x = w; x <= last; ++x) rb[x - (x0 - 4)] = rb[(2 * w - 2 - x) - (x0 - 4)];
        }
      }
      __syncwarp();
    }
  } else {
    // unaligned external level 0 (or a level narrower than 16 px): every word assembled from reflected bytes
    int ca[4], cb[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      ca[k] = reflect101(xa + k, w);
      cb[k] = reflect101(xb + k, w);
    }
    for (int r = 0; r < nrows; ++r) {
      const uint8_t* row = img + (size_t)reflect101(y0 - 3 + r, h) * pitch;
      slab[r * kBlurSrcWords + lane] = (unsigned)__ldg(row + ca[0]) | ((unsigned)__ldg(row + ca[1]) << 8) |
                                       ((unsigned)__ldg(row + ca[2]) << 16) | ((unsigned)__ldg(row + ca[3]) << 24);
      if (lane < 2)
        slab[r * kBlurSrcWords + 32 + lane] = (unsigned)__ldg(row + cb[0]) | ((unsigned)__ldg(row + cb[1]) << 8) |
                                              ((unsigned)__ldg(row + cb[2]) << 16) |
                                              ((unsigned)__ldg(row + cb[3]) << 24);
    }
    __syncwarp();
  }

  const bool store = x0 + 4 * lane < w;  // pad columns inside the 128-byte aligned pitch may receive don't-care bytes
  if (nout == kBlurTileH)
    blur_walk<true>(slab + lane, out, blur_pitch, nout, store);
  else
    blur_walk<false>(slab + lane, out, blur_pitch, nout, store);
}

int launch_blur(const DevGeom& g, cudaStream_t s) {
  if (g.blur_tiles_total == 0 || g.batch == 0) return ORBX_OK;
  dim3 grid(div_up(g.blur_tiles_total, kBlurWarps), g.batch);
  blur7_kernel<<<grid, kBlurWarps * 32, 0, s>>>(g);
  ORBX_LAUNCH_CHECK();
  return ORBX_OK;
}

// one warp per keypoint; lane i evaluates pairs 32j+i, the ballot of (a < b) is descriptor word j.
// The 39 x 39 neighbourhood (|offset| <= 19 after rotation) is first copied into a per-warp shared-memory
// slab with row-contiguous aligned 32-bit loads (~3 rows per load instruction), so the 512 scattered
// samples hit shared memory instead of issuing 16 fully divergent global loads per lane.
constexpr int kBriefWarps = 8;
constexpr int kBriefPerWarp = 8;

// cvRound for |x| < 2^22 without the XU-pipe F2I: the fp32 add rounds half to even at integer granularity
__device__ __forceinline__ int rint_small(float x) {
  return (int)__float_as_uint(__fadd_rn(x, 12582912.0f)) - 0x4B400000;
}
// same rounding, bias left in: 0x4B400000 + cvRound(x) (callers fold the bias into an address)
__device__ __forceinline__ unsigned rint_raw(float x) { return __float_as_uint(__fadd_rn(x, 12582912.0f)); }
__device__ __forceinline__ unsigned lds_u8(unsigned shared_addr) {
  unsigned v;
  asm volatile("ld.shared.u8 %0, [%1];\n" : "=r"(v) : "r"(shared_addr));
  return v;
}
constexpr int kBriefRows = 39;
constexpr int kBriefWords = 11;                              // 44 bytes per row
constexpr int kBriefLoads = (kBriefRows * kBriefWords + 31) / 32;  // 14

__global__ void __launch_bounds__(kBriefWarps * 32) brief_kernel(DevGeom g, OutArgs o) {
  // pattern as fp32 (I2F and F2I are quarter-rate XU-pipe conversions: keep them out of the inner loop)
  __shared__ float4 s_pat[256];
  // two slabs per warp: the neighbourhood of keypoint it+1 is fetched (cp.async) while keypoint it is sampled
  __shared__ unsigned s_patch[kBriefWarps][2][kBriefRows * kBriefWords + 3];
  {
    const char4 c = c_pattern[threadIdx.x];
    s_pat[threadIdx.x] = make_float4((float)c.x, (float)c.y, (float)c.z, (float)c.w);
  }
  __syncthreads();
  const int frame = blockIdx.y;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int n = min(o.n[frame], min(o.cap, o.lvl_cap));
  int st_row[kBriefLoads], st_col[kBriefLoads];  // row and byte column of staged word i = lane + 32t
#pragma unroll
  for (int t = 0; t < kBriefLoads; ++t) {
    const int i = lane + 32 * t;
    st_row[t] = i / kBriefWords;
    st_col[t] = 4 * (i - st_row[t] * kBriefWords);
  }
  // lane k < kBriefPerWarp prepares keypoint k of this warp: position word and (float)cos/sin((double)theta).
  // The double-precision sin/cos (~150 instructions) is evaluated once per warp for all of its keypoints
  // instead of once per keypoint by all 32 lanes.
  uint32_t lv_mine = 0;
  float a_mine = 0.f, b_mine = 0.f;
  {
    const int jm = (blockIdx.x * kBriefPerWarp + min(lane, kBriefPerWarp - 1)) * kBriefWarps + warp;
    if (lane < kBriefPerWarp && jm < n) {
      lv_mine = o.kp_lvl[(size_t)frame * o.lvl_cap + jm];
      const float angle = o.kps[(size_t)frame * o.cap + jm].angle;
      const float th = __fmul_rn(angle, (float)(3.14159265358979323846 / 180.0));
      a_mine = (float)cos((double)th);
      b_mine = (float)sin((double)th);
    }
  }
  // start the fetch of keypoint `it` (one cp.async group per keypoint, empty when there is none); returns the
  // byte offset of the neighbourhood's first pixel inside its first aligned word
  auto fetch = [&](int it) -> int {
    int off = 0;
    const int j = (blockIdx.x * kBriefPerWarp + it) * kBriefWarps + warp;
    if (it < kBriefPerWarp && j < n) {
      const uint32_t lv = __shfl_sync(0xFFFFFFFFu, lv_mine, it);
      const int l = (int)(lv >> 24), cx = (int)(lv & 0xFFF), cy = (int)((lv >> 12) & 0xFFF);
      const DevLevel& L = g.lv[l];
      const int bp = L.blur_pitch;
      // blurred planes are internal: base, pitch and frame stride are multiples of 128
      const int o0 = (cy - 19) * bp + (cx - 19);
      off = o0 & 3;
      const uint8_t* base = L.blur + (size_t)frame * L.blur_stride + (o0 - off);
      const unsigned sa = (unsigned)__cvta_generic_to_shared(&s_patch[warp][it & 1][lane]);
#pragma unroll
      for (int t = 0; t < kBriefLoads; ++t)
        if (lane + 32 * t < kBriefRows * kBriefWords)
          asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(sa + 128 * t),
                       "l"(base + (st_row[t] * bp + st_col[t])));
    }
    asm volatile("cp.async.commit_group;\n" ::);
    return off;
  };

  int off = fetch(0);
#pragma unroll 1
  for (int it = 0; it < kBriefPerWarp; ++it) {
    const int j = (blockIdx.x * kBriefPerWarp + it) * kBriefWarps + warp;
    if (j >= n) return;  // warp-uniform; j grows with it
    const int off_next = fetch(it + 1);
    const float a = __shfl_sync(0xFFFFFFFFu, a_mine, it), b = __shfl_sync(0xFFFFFFFFu, b_mine, it);
    asm volatile("cp.async.wait_group 1;\n" ::: "memory");
    __syncwarp();
    // sample address = centre + iy * 44 + ix with ix, iy = cvRound(rotated offset).  The magic-add rounding
    // leaves 0x4B400000 + i in the register; the two biases are folded into the (32-bit, wrapping) base address.
    const unsigned centre = (unsigned)__cvta_generic_to_shared(&s_patch[warp][it & 1][0]) + 19 * (kBriefWords * 4) + 19 +
                            (unsigned)off - 0x4B400000u * (unsigned)(kBriefWords * 4 + 1);
    uint32_t word = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const float4 p = s_pat[k * 32 + lane];
      const float ax = p.x, ay = p.y, bx = p.z, by = p.w;
      const unsigned iax = rint_raw(__fsub_rn(__fmul_rn(ax, a), __fmul_rn(ay, b)));
      const unsigned iay = rint_raw(__fadd_rn(__fmul_rn(ax, b), __fmul_rn(ay, a)));
      const unsigned ibx = rint_raw(__fsub_rn(__fmul_rn(bx, a), __fmul_rn(by, b)));
      const unsigned iby = rint_raw(__fadd_rn(__fmul_rn(bx, b), __fmul_rn(by, a)));
      const unsigned va = lds_u8(centre + iay * (kBriefWords * 4) + iax);
      const unsigned vb = lds_u8(centre + iby * (kBriefWords * 4) + ibx);
      const uint32_t bal = __ballot_sync(0xFFFFFFFFu, va < vb);
      if (lane == k) word = bal;
    }
    if (lane < 8) reinterpret_cast<uint32_t*>(o.desc + ((size_t)frame * o.cap + j) * 32)[lane] = word;
    __syncwarp();  // slab it & 1 is free for keypoint it + 2
    off = off_next;
  }
}

int launch_brief(const DevGeom& g, const OutArgs& o, cudaStream_t s) {
  if (g.batch == 0 || !o.desc) return ORBX_OK;
  const int lim = o.cap < o.lvl_cap ? o.cap : o.lvl_cap;
  dim3 grid(div_up(lim > 0 ? lim : 1, kBriefWarps * kBriefPerWarp), g.batch);
  brief_kernel<<<grid, kBriefWarps * 32, 0, s>>>(g, o);
  ORBX_LAUNCH_CHECK();
  return ORBX_OK;
}

}  // namespace orbx
