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

constexpr int kBlurSrcW = kBlurTileW + 6;  // 70
constexpr int kBlurSrcH = kBlurTileH + 6;  // 38
constexpr int kBlurSrcPitch = 72;

__global__ void __launch_bounds__(256) blur7_kernel(DevGeom g) {
  __shared__ uint8_t s_src[kBlurSrcH][kBlurSrcPitch];
  __shared__ float s_row[kBlurSrcH][kBlurTileW + 1];
  const int tile = blockIdx.x;
  int l = 0;
#pragma unroll
  for (int i = 1; i < kMaxLevels; ++i)
    if (i < g.nlevels && tile >= g.lv[i].blur_tile_begin) l = i;
  const DevLevel& L = g.lv[l];
  const int frame = blockIdx.y;
  const int lt = tile - L.blur_tile_begin;
  const int ty = lt / L.blur_tiles_x, tx = lt - ty * L.blur_tiles_x;
  const int x0 = tx * kBlurTileW, y0 = ty * kBlurTileH;
  const uint8_t* img = L.img + (size_t)frame * L.frame_stride;
  const int tid = threadIdx.x;

  for (int i = tid; i < kBlurSrcW * kBlurSrcH; i += 256) {
    const int r = i / kBlurSrcW, c = i - r * kBlurSrcW;
    const int gy = reflect101(y0 - 3 + r, L.h);
    const int gx = reflect101(x0 - 3 + c, L.w);
    s_src[r][c] = __ldg(img + (size_t)gy * L.pitch + gx);
  }
  __syncthreads();
  // row pass: acc = k0*p[x-3]; acc = fma(k_j, p[x-3+j], acc), j = 1..6
  for (int i = tid; i < kBlurTileW * kBlurSrcH; i += 256) {
    const int r = i / kBlurTileW, c = i - r * kBlurTileW;
    float acc = __fmul_rn(c_gauss[0], (float)s_src[r][c]);
#pragma unroll
    for (int j = 1; j < 7; ++j) acc = __fmaf_rn(c_gauss[j], (float)s_src[r][c + j], acc);
    s_row[r][c] = acc;
  }
  __syncthreads();
  // column pass: acc = k3*r[y]; acc = fma(k_{3+j}, r[y+j] + r[y-j], acc), j = 1..3
  uint8_t* out = L.blur + (size_t)frame * L.blur_stride;
  for (int i = tid; i < kBlurTileW * kBlurTileH; i += 256) {
    const int r = i / kBlurTileW, c = i - r * kBlurTileW;
    const int gx = x0 + c, gy = y0 + r;
    if (gx >= L.w || gy >= L.h) continue;
    float acc = __fmul_rn(c_gauss[3], s_row[r + 3][c]);
#pragma unroll
    for (int j = 1; j <= 3; ++j)
      acc = __fmaf_rn(c_gauss[3 + j], __fadd_rn(s_row[r + 3 + j][c], s_row[r + 3 - j][c]), acc);
    const int v = __float2int_rn(acc);
    out[(size_t)gy * L.blur_pitch + gx] = (uint8_t)min(max(v, 0), 255);
  }
}

int launch_blur(const DevGeom& g, cudaStream_t s) {
  if (g.blur_tiles_total == 0 || g.batch == 0) return ORBX_OK;
  dim3 grid(g.blur_tiles_total, g.batch);
  blur7_kernel<<<grid, 256, 0, s>>>(g);
  ORBX_LAUNCH_CHECK();
  return ORBX_OK;
}

// one warp per keypoint; lane i evaluates pairs 32j+i, the ballot of (a < b) is descriptor word j
constexpr int kBriefWarps = 8;
constexpr int kBriefPerWarp = 4;

__global__ void __launch_bounds__(kBriefWarps * 32) brief_kernel(DevGeom g, OutArgs o) {
  __shared__ char4 s_pat[256];
  s_pat[threadIdx.x] = c_pattern[threadIdx.x];
  __syncthreads();
  const int frame = blockIdx.y;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int n = min(o.n[frame], min(o.cap, o.lvl_cap));
  for (int it = 0; it < kBriefPerWarp; ++it) {
    const int j = (blockIdx.x * kBriefPerWarp + it) * kBriefWarps + warp;
    if (j >= n) return;
    const uint32_t lv = o.kp_lvl[(size_t)frame * o.lvl_cap + j];
    const int l = (int)(lv >> 24), cx = (int)(lv & 0xFFF), cy = (int)((lv >> 12) & 0xFFF);
    const DevLevel& L = g.lv[l];
    const float angle = o.kps[(size_t)frame * o.cap + j].angle;
    const float th = __fmul_rn(angle, (float)(3.14159265358979323846 / 180.0));
    const float a = (float)cos((double)th), b = (float)sin((double)th);
    const uint8_t* img = L.blur + (size_t)frame * L.blur_stride + (size_t)cy * L.blur_pitch + cx;
    uint32_t word = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const char4 p = s_pat[k * 32 + lane];
      const float ax = (float)p.x, ay = (float)p.y, bx = (float)p.z, by = (float)p.w;
      const int iax = __float2int_rn(__fsub_rn(__fmul_rn(ax, a), __fmul_rn(ay, b)));
      const int iay = __float2int_rn(__fadd_rn(__fmul_rn(ax, b), __fmul_rn(ay, a)));
      const int ibx = __float2int_rn(__fsub_rn(__fmul_rn(bx, a), __fmul_rn(by, b)));
      const int iby = __float2int_rn(__fadd_rn(__fmul_rn(bx, b), __fmul_rn(by, a)));
      const int va = __ldg(img + iay * L.blur_pitch + iax);
      const int vb = __ldg(img + iby * L.blur_pitch + ibx);
      const uint32_t bal = __ballot_sync(0xFFFFFFFFu, va < vb);
      if (lane == k) word = bal;
    }
    if (lane < 8) reinterpret_cast<uint32_t*>(o.desc + ((size_t)frame * o.cap + j) * 32)[lane] = word;
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
