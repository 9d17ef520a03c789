// S1 — pyramid level l = resize(level l-1, INTER_LINEAR_EXACT): 8.8 fixed-point bilinear chain.
//
// Replaces cv::resize inside cv::ORB_Impl::detectAndCompute (reached from
// /root/reference/src/feature/detector.rs:22,32).  Arithmetic per SURVEY.md Appendix A S1:
//   H = c0x*p[ofs] + c1x*p[ofs+1];  out = (c0y*H0 + c1y*H1 + 32768) >> 16, weights sum to 256.
// The per-axis (ofs, c1) tables are computed on the host in double (detector.cu) exactly as OpenCV
// does.  One launch per level over the whole batch.  A block of 128 threads owns a 128 x 32 output tile:
//   pass 1 (horizontal): thread = output column; it walks down the ~40 source rows the tile touches and
//           writes H = c0x*p[ofs] + c1x*p[ofs+1] (<= 65280) to shared memory - each source row is
//           interpolated once, the two byte loads of a warp fall into one or two 32-byte sectors;
//   pass 2 (vertical): thread = 4 adjacent columns x 8 rows; two LDS.128 of H, 8 IMAD, one 32-bit store.
// HBM-bound by design (reads px(l-1), writes px(l)); ~15 instructions per output pixel.
#include "detector.cuh"

namespace orbx {

constexpr int kResizeThreads = 128;

__global__ void __launch_bounds__(kResizeThreads) resize_exact_kernel(ResizeArgs a) {
  extern __shared__ __align__(16) uint32_t s_h[];  // [source rows of the tile][128]
  const int frame = blockIdx.z;
  const int ox0 = blockIdx.x * kResizeTileW;
  const int oy0 = blockIdx.y * kResizeTileH;
  const int oy1 = min(oy0 + kResizeTileH, a.dh) - 1;
  const int sy0 = (int)(__ldg(a.ytab + oy0) >> 16);
  const int sy1 = min((int)(__ldg(a.ytab + oy1) >> 16) + 1, a.sh - 1);
  const int rows = sy1 - sy0 + 1;
  const uint8_t* src = a.src + (size_t)frame * a.src_stride + (size_t)sy0 * a.spitch;
  const int tid = threadIdx.x;

  // pass 1
  {
    const int x = min(ox0 + tid, a.dw - 1);
    const uint32_t xt = __ldg(a.xtab + x);
    const int xo = (int)(xt >> 16);
    const int c1x = (int)(xt & 0xFFFF), c0x = 256 - c1x;
    const uint8_t* p0 = src + xo;
    const uint8_t* p1 = src + min(xo + 1, a.sw - 1);
    const int sp = a.spitch;
    int r = 0;
    for (; r + 4 <= rows; r += 4) {
      int v0[4], v1[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        v0[u] = __ldg(p0 + (size_t)(r + u) * sp);
        v1[u] = __ldg(p1 + (size_t)(r + u) * sp);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) s_h[(r + u) * kResizeTileW + tid] = (uint32_t)(c0x * v0[u] + c1x * v1[u]);
    }
    for (; r < rows; ++r)
      s_h[r * kResizeTileW + tid] = (uint32_t)(c0x * (int)__ldg(p0 + (size_t)r * sp) + c1x * (int)__ldg(p1 + (size_t)r * sp));
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
    const uint32_t v0 = (uint32_t)(c0y * (int)h0.x + c1y * (int)h1.x + 32768) >> 16;  // <= 255 by construction
    const uint32_t v1 = (uint32_t)(c0y * (int)h0.y + c1y * (int)h1.y + 32768) >> 16;
    const uint32_t v2 = (uint32_t)(c0y * (int)h0.z + c1y * (int)h1.z + 32768) >> 16;
    const uint32_t v3 = (uint32_t)(c0y * (int)h0.w + c1y * (int)h1.w + 32768) >> 16;
    // dpitch is a multiple of 128, so the whole word is ours even when it runs past dw
    *reinterpret_cast<uint32_t*>(dst + (size_t)y * a.dpitch) = v0 | (v1 << 8) | (v2 << 16) | (v3 << 24);
  }
}

int launch_resize(const ResizeArgs& a, int batch, cudaStream_t s) {
  dim3 grid(div_up(a.dw, kResizeTileW), div_up(a.dh, kResizeTileH), batch);
  dim3 block(kResizeThreads);
  const size_t smem = (size_t)a.smem_h * kResizeTileW * sizeof(uint32_t);
  if (smem > 48 * 1024) {
    cudaFuncSetAttribute(resize_exact_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  }
  resize_exact_kernel<<<grid, block, smem, s>>>(a);
  ORBX_LAUNCH_CHECK();
  return ORBX_OK;
}

}  // namespace orbx
