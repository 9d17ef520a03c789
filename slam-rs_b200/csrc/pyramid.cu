// S1 — pyramid level l = resize(level l-1, INTER_LINEAR_EXACT): 8.8 fixed-point bilinear chain.
//
// Replaces cv::resize inside cv::ORB_Impl::detectAndCompute (reached from
// /root/reference/src/feature/detector.rs:22,32).  Arithmetic per SURVEY.md Appendix A S1:
//   H = c0x*p[ofs] + c1x*p[ofs+1];  out = (c0y*H0 + c1y*H1 + 32768) >> 16, weights sum to 256.
// The per-axis (ofs, c1) tables are computed on the host in double (detector.cu) exactly as OpenCV
// does.  One launch per level over the whole batch; the source footprint of a 128x16 output tile is
// staged in shared memory with aligned 32-bit loads, then each thread produces 4 adjacent output
// pixels of 2 rows and stores them as one 32-bit word per row.  HBM-bound: reads px(l-1), writes px(l).
#include "detector.cuh"

namespace orbx {

__global__ void __launch_bounds__(256) resize_exact_kernel(ResizeArgs a) {
  extern __shared__ __align__(16) uint8_t s_src[];
  const int frame = blockIdx.z;
  const int ox0 = blockIdx.x * kResizeTileW;
  const int oy0 = blockIdx.y * kResizeTileH;
  const int ox1 = min(ox0 + kResizeTileW, a.dw) - 1;
  const int oy1 = min(oy0 + kResizeTileH, a.dh) - 1;
  const int sx0 = (int)(__ldg(a.xtab + ox0) >> 16);
  const int sy0 = (int)(__ldg(a.ytab + oy0) >> 16);
  const int sx1 = min((int)(__ldg(a.xtab + ox1) >> 16) + 1, a.sw - 1);
  const int sy1 = min((int)(__ldg(a.ytab + oy1) >> 16) + 1, a.sh - 1);
  const int bx = sx0 & ~3;  // aligned staging origin
  const int cols = sx1 - bx + 1;
  const int rows = sy1 - sy0 + 1;
  const int spitch = a.smem_w;  // multiple of 4
  const uint8_t* src = a.src + (size_t)frame * a.src_stride;

  const int tid = threadIdx.y * blockDim.x + threadIdx.x;
  if (a.src_aligned4) {
    const int wpr = (cols + 3) >> 2;
    for (int i = tid; i < wpr * rows; i += 256) {
      const int r = i / wpr, c = i - r * wpr;
      // the last word of a row may run past sw but stays inside the padded pitch
      const uint32_t v = __ldg(reinterpret_cast<const uint32_t*>(src + (size_t)(sy0 + r) * a.spitch + bx) + c);
      *reinterpret_cast<uint32_t*>(s_src + r * spitch + c * 4) = v;
    }
  } else {
    for (int i = tid; i < cols * rows; i += 256) {
      const int r = i / cols, c = i - r * cols;
      s_src[r * spitch + c] = __ldg(src + (size_t)(sy0 + r) * a.spitch + bx + c);
    }
  }
  __syncthreads();

  const int x = ox0 + threadIdx.x * 4;
  if (x >= a.dw) return;
  uint32_t xt[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) xt[k] = __ldg(a.xtab + min(x + k, a.dw - 1));
  uint8_t* dst = a.dst + (size_t)frame * a.dst_stride;
#pragma unroll
  for (int ry = 0; ry < 2; ++ry) {
    const int y = oy0 + threadIdx.y * 2 + ry;
    if (y >= a.dh) break;
    const uint32_t yt = __ldg(a.ytab + y);
    const int yo = (int)(yt >> 16);
    const int c1y = (int)(yt & 0xFFFF), c0y = 256 - c1y;
    const uint8_t* r0 = s_src + (yo - sy0) * spitch - bx;
    const uint8_t* r1 = s_src + (min(yo + 1, a.sh - 1) - sy0) * spitch - bx;
    uint32_t packed = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int xo = (int)(xt[k] >> 16);
      const int c1x = (int)(xt[k] & 0xFFFF), c0x = 256 - c1x;
      const int xo1 = min(xo + 1, a.sw - 1);
      const int h0 = c0x * r0[xo] + c1x * r0[xo1];
      const int h1 = c0x * r1[xo] + c1x * r1[xo1];
      int v = (c0y * h0 + c1y * h1 + 32768) >> 16;
      v = min(max(v, 0), 255);
      packed |= (uint32_t)v << (8 * k);
    }
    uint8_t* o = dst + (size_t)y * a.dpitch + x;
    if (x + 3 < a.dpitch) {
      *reinterpret_cast<uint32_t*>(o) = packed;  // dpitch is a multiple of 128: pad bytes are ours
    } else {
      for (int k = 0; k < 4 && x + k < a.dw; ++k) o[k] = (uint8_t)(packed >> (8 * k));
    }
  }
}

int launch_resize(const ResizeArgs& a, int batch, cudaStream_t s) {
  dim3 grid(div_up(a.dw, kResizeTileW), div_up(a.dh, kResizeTileH), batch);
  dim3 block(32, 8);
  const size_t smem = (size_t)a.smem_w * a.smem_h;
  if (smem > 48 * 1024) {
    cudaFuncSetAttribute(resize_exact_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  }
  resize_exact_kernel<<<grid, block, smem, s>>>(a);
  ORBX_LAUNCH_CHECK();
  return ORBX_OK;
}

}  // namespace orbx
