// S1 — pyramid level l = resize(level l-1, INTER_LINEAR_EXACT): 8.8 fixed-point bilinear chain.
//
// Replaces cv::resize inside cv::ORB_Impl::detectAndCompute (reached from
// /root/reference/src/feature/detector.rs:22,32).  Arithmetic per SURVEY.md Appendix A S1:
//   H = c0x*p[ofs] + c1x*p[ofs+1];  out = (c0y*H0 + c1y*H1 + 32768) >> 16, weights sum to 256.
// The per-axis (ofs, c1) tables are computed on the host in double (detector.cu) exactly as OpenCV
// does.  One launch per level over the whole batch.  A block of 128 threads owns a 128 x 32 output tile:
//   stage : the source footprint (<= 42 rows x 192 bytes) goes global -> shared with cp.async (16-byte
//           chunks when the plane is 16-byte aligned, else 4-byte, else byte loads);
//   pass 1: thread = output column; it walks down the source rows and writes
//           H = c0x*p[ofs] + c1x*p[ofs+1] (<= 65280) to shared memory: each source row is interpolated
//           once, 2 LDS.U8 + 2 IMAD + 1 STS per value with immediate offsets;
//   pass 2: thread = 4 adjacent columns x 8 rows; two LDS.128 of H, 8 IMAD, 3 PRMT, one 32-bit store.
// HBM-bound by design (reads px(l-1), writes px(l)).
#include "detector.cuh"

namespace orbx {

constexpr int kResizeThreads = 128;
constexpr int kResizeSrcPitch = 192;  // bytes per staged source row (128 * 1.2 + 2 + 15 alignment slack <= 176)
constexpr int kResizeMaxRows = 44;    // source rows per tile (32 * 1.2 + 2 <= 41)

__global__ void __launch_bounds__(kResizeThreads) resize_exact_kernel(ResizeArgs a) {
  __shared__ __align__(16) uint8_t s_src[kResizeMaxRows * kResizeSrcPitch];
  __shared__ __align__(16) uint32_t s_h[kResizeMaxRows * kResizeTileW];
  const int frame = blockIdx.z;
  const int ox0 = blockIdx.x * kResizeTileW;
  const int oy0 = blockIdx.y * kResizeTileH;
  const int ox1 = min(ox0 + kResizeTileW, a.dw) - 1;
  const int oy1 = min(oy0 + kResizeTileH, a.dh) - 1;
  const int sx0 = (int)(__ldg(a.xtab + ox0) >> 16);
  const int sx1 = min((int)(__ldg(a.xtab + ox1) >> 16) + 1, a.sw - 1);
  const int sy0 = (int)(__ldg(a.ytab + oy0) >> 16);
  const int sy1 = min((int)(__ldg(a.ytab + oy1) >> 16) + 1, a.sh - 1);
  const int rows = sy1 - sy0 + 1;
  const int bx = sx0 & ~15;          // staging origin
  const int cols = sx1 - bx + 1;     // bytes needed per row (<= kResizeSrcPitch, checked on the host)
  const uint8_t* src = a.src + (size_t)frame * a.src_stride + (size_t)sy0 * a.spitch + bx;
  const int tid = threadIdx.x;

  // stage
  {
    const unsigned sbase = (unsigned)__cvta_generic_to_shared(s_src);
    if (a.src_align >= 16) {
      const int cpr = (cols + 15) >> 4;
      for (int i = tid; i < cpr * rows; i += kResizeThreads) {
        const int r = i / cpr, c = i - r * cpr;
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(sbase + r * kResizeSrcPitch + c * 16),
                     "l"(src + (size_t)r * a.spitch + c * 16));
      }
    } else if (a.src_align >= 4) {
      const int cpr = (cols + 3) >> 2;
      for (int i = tid; i < cpr * rows; i += kResizeThreads) {
        const int r = i / cpr, c = i - r * cpr;
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(sbase + r * kResizeSrcPitch + c * 4),
                     "l"(src + (size_t)r * a.spitch + c * 4));
      }
    } else {
      for (int i = tid; i < cols * rows; i += kResizeThreads) {
        const int r = i / cols, c = i - r * cols;
        s_src[r * kResizeSrcPitch + c] = __ldg(src + (size_t)r * a.spitch + c);
      }
    }
    asm volatile("cp.async.commit_group;\n" ::);
    asm volatile("cp.async.wait_group 0;\n" ::: "memory");
  }
  __syncthreads();

  // pass 1
  {
    const int x = min(ox0 + tid, a.dw - 1);
    const uint32_t xt = __ldg(a.xtab + x);
    const int xo = (int)(xt >> 16);
    const int c1x = (int)(xt & 0xFFFF), c0x = 256 - c1x;
    const uint8_t* p0 = s_src + (xo - bx);
    const uint8_t* p1 = s_src + (min(xo + 1, a.sw - 1) - bx);
    uint32_t* hp = s_h + tid;
    int r = 0;
    for (; r + 8 <= rows; r += 8) {
#pragma unroll
      for (int u = 0; u < 8; ++u)
        hp[(r + u) * kResizeTileW] =
            (uint32_t)(c0x * (int)p0[(r + u) * kResizeSrcPitch] + c1x * (int)p1[(r + u) * kResizeSrcPitch]);
    }
    for (; r < rows; ++r)
      hp[r * kResizeTileW] = (uint32_t)(c0x * (int)p0[r * kResizeSrcPitch] + c1x * (int)p1[r * kResizeSrcPitch]);
  }
  __syncthreads();

  // pass 2
  const int cg = tid & 31, rg = tid >> 5;  // 32 column groups x 4 row groups of 8 rows
  const int x = ox0 + cg * 4;
  if (x >= a.dw) return;
  uint8_t* dst = a.dst + (size_t)frame * a.dst_stride + x;
#pragma unroll 2
  for (int ry = 0; ry < kResizeTileH / 4; ++ry) {
    const int y = oy0 + rg * (kResizeTileH / 4) + ry;
    if (y > oy1) break;
    const uint32_t yt = __ldg(a.ytab + y);
    const int yo = (int)(yt >> 16);
    const int c1y = (int)(yt & 0xFFFF), c0y = 256 - c1y;
    const uint4 h0 = *reinterpret_cast<const uint4*>(s_h + (yo - sy0) * kResizeTileW + cg * 4);
    const uint4 h1 = *reinterpret_cast<const uint4*>(s_h + (min(yo + 1, a.sh - 1) - sy0) * kResizeTileW + cg * 4);
    // (c0y*H0 + c1y*H1 + 32768) >> 16 is <= 255 by construction: the result is byte 2 of the sum
    const uint32_t v0 = (uint32_t)(c0y * (int)h0.x + c1y * (int)h1.x + 32768);
    const uint32_t v1 = (uint32_t)(c0y * (int)h0.y + c1y * (int)h1.y + 32768);
    const uint32_t v2 = (uint32_t)(c0y * (int)h0.z + c1y * (int)h1.z + 32768);
    const uint32_t v3 = (uint32_t)(c0y * (int)h0.w + c1y * (int)h1.w + 32768);
    // dpitch is a multiple of 128, so the whole word is ours even when it runs past dw
    *reinterpret_cast<uint32_t*>(dst + (size_t)y * a.dpitch) =
        __byte_perm(__byte_perm(v0, v1, 0x0062), __byte_perm(v2, v3, 0x0062), 0x5410);
  }
}

int launch_resize(const ResizeArgs& a, int batch, cudaStream_t s) {
  if (a.smem_w + 16 > kResizeSrcPitch || a.smem_h > kResizeMaxRows) {
    set_error("orbx resize: source footprint %d x %d of a tile exceeds the staged %d x %d", a.smem_w, a.smem_h,
              kResizeSrcPitch, kResizeMaxRows);
    return ORBX_E_BADARG;
  }
  dim3 grid(div_up(a.dw, kResizeTileW), div_up(a.dh, kResizeTileH), batch);
  resize_exact_kernel<<<grid, kResizeThreads, 0, s>>>(a);
  ORBX_LAUNCH_CHECK();
  return ORBX_OK;
}

}  // namespace orbx
