// S1 — pyramid level l = resize(level l-1, INTER_LINEAR_EXACT): 8.8 fixed-point bilinear chain.
//
// Replaces cv::resize inside cv::ORB_Impl::detectAndCompute (reached from
// /root/reference/src/feature/detector.rs:22,32).  Arithmetic per SURVEY.md Appendix A S1:
//   H = c0x*p[ofs] + c1x*p[ofs+1];  out = (c0y*H0 + c1y*H1 + 32768) >> 16, weights sum to 256.
// Nothing is rounded before the final shift, so out = (sum_ij cy_i*cx_j*p_ij + 32768) >> 16 and the two
// 1-D passes commute exactly; the column (y) pass runs first here because it works on packed bytes.
// The per-axis (ofs, c1) tables are computed on the host in double (detector.cu) exactly as OpenCV
// does.  One launch per level over the whole batch.  A block of 128 threads owns a 128 x 32 output tile:
//   stage : the source footprint (<= 44 rows x 192 bytes) goes global -> shared with cp.async (16-byte
//           chunks when the plane is 16-byte aligned, else 4-byte, else byte loads);
//   pass 1: thread = source word (4 pixels) x output row: V = c0y*row0 + c1y*row1 on two 16x2 lanes
//           (even / odd pixels; V <= 65280 fits a lane), stored as two u16 planes;
//   pass 2: thread = 4 adjacent output columns x 8 rows: per pixel 2 LDS.U16 + 2 IMAD, byte 2 of the sum is
//           the result, one 32-bit store per row.
// HBM-bound by design (reads px(l-1), writes px(l)).
#include "detector.cuh"

namespace orbx {

constexpr int kResizeThreads = 128;
constexpr int kResizeSrcPitch = 192;  // bytes per staged source row (128 * 1.2 + 2 + 15 alignment slack <= 176)
constexpr int kResizeMaxRows = 44;    // source rows per tile (32 * 1.2 + 2 <= 41)
constexpr int kResizeSrcWords = kResizeSrcPitch / 4;
constexpr int kResizeVPitch = 100;    // words per V row: even-pixel plane [0, 48), odd-pixel plane [50, 98)
constexpr int kResizeVOdd = 50;

__global__ void __launch_bounds__(kResizeThreads) resize_exact_kernel(ResizeArgs a) {
  __shared__ __align__(16) uint32_t s_src[kResizeMaxRows * kResizeSrcWords];
  __shared__ __align__(16) uint32_t s_v[kResizeTileH * kResizeVPitch];
  __shared__ uint32_t s_y[kResizeTileH];  // word offset of source row r0 | word offset of r1 << 16
  __shared__ uint32_t s_cy[kResizeTileH];
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
  const int nrows_out = oy1 - oy0 + 1;

  // stage
  {
    const unsigned sbase = (unsigned)__cvta_generic_to_shared(s_src);
    if (a.src_align >= 16) {
      // 16 threads per row (12 chunks of 16 bytes at most), 8 rows per sweep
      const int c = tid & 15, r0 = tid >> 4;
      if (c * 16 < cols) {
        const uint8_t* g = src + (size_t)r0 * a.spitch + c * 16;
        unsigned d = sbase + r0 * kResizeSrcPitch + c * 16;
        for (int r = r0; r < rows; r += 8) {
          asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(d), "l"(g));
          g += (size_t)8 * a.spitch;
          d += 8 * kResizeSrcPitch;
        }
      }
    } else if (a.src_align >= 4) {
      // 64 threads per row (48 words at most), 2 rows per sweep
      const int c = tid & 63, r0 = tid >> 6;
      if (c * 4 < cols) {
        const uint8_t* g = src + (size_t)r0 * a.spitch + c * 4;
        unsigned d = sbase + r0 * kResizeSrcPitch + c * 4;
        for (int r = r0; r < rows; r += 2) {
          asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(d), "l"(g));
          g += (size_t)2 * a.spitch;
          d += 2 * kResizeSrcPitch;
        }
      }
    } else {
      uint8_t* sb = reinterpret_cast<uint8_t*>(s_src);
      for (int i = tid; i < cols * rows; i += kResizeThreads) {
        const int r = i / cols, c = i - r * cols;
        sb[r * kResizeSrcPitch + c] = __ldg(src + (size_t)r * a.spitch + c);
      }
    }
    asm volatile("cp.async.commit_group;\n" ::);
    if (tid < nrows_out) {
      const uint32_t yt = __ldg(a.ytab + oy0 + tid);
      const int yo = (int)(yt >> 16);
      s_y[tid] = (uint32_t)((yo - sy0) * kResizeSrcWords) |
                 ((uint32_t)((min(yo + 1, a.sh - 1) - sy0) * kResizeSrcWords) << 16);
      s_cy[tid] = yt & 0xFFFFu;
    }
    asm volatile("cp.async.wait_group 0;\n" ::: "memory");
  }
  __syncthreads();

  // pass 1 (columns): thread = source word wc, output rows rsub, rsub + nsub, ...
  {
    const int nw = (cols + 3) >> 2;  // <= 48
    const int rsub = tid / nw, wc = tid - rsub * nw;
    const int nsub = kResizeThreads / nw;
    if (rsub < nsub) {
      const uint32_t* sp = s_src + wc;
      uint32_t* vp = s_v + wc;
      for (int ry = rsub; ry < nrows_out; ry += nsub) {
        const uint32_t yo = s_y[ry];
        const uint32_t c1y = s_cy[ry], c0y = 256u - c1y;
        const uint32_t w0 = sp[yo & 0xFFFFu], w1 = sp[yo >> 16];
        // lanes hold <= 255 * 256, so the 32-bit products never carry from the low into the high lane
        vp[ry * kResizeVPitch] = __byte_perm(w0, 0u, 0x4240) * c0y + __byte_perm(w1, 0u, 0x4240) * c1y;
        vp[ry * kResizeVPitch + kResizeVOdd] = __byte_perm(w0, 0u, 0x4341) * c0y + __byte_perm(w1, 0u, 0x4341) * c1y;
      }
    }
  }
  __syncthreads();

  // pass 2 (rows): thread = 4 adjacent output columns x 8 rows
  const int cg = tid & 31, rg = tid >> 5;  // 32 column groups x 4 row groups of 8 rows
  const int x = ox0 + cg * 4;
  if (x >= a.dw) return;
  unsigned a0[4], a1[4];  // byte offsets of V[xo], V[xo + 1] inside a V row
  uint32_t c1x[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const uint32_t xt = __ldg(a.xtab + min(x + k, a.dw - 1));
    const int xo = (int)(xt >> 16);
    c1x[k] = xt & 0xFFFFu;
    const int p0 = xo - bx, p1 = min(xo + 1, a.sw - 1) - bx;
    a0[k] = (unsigned)(((p0 & 1) ? kResizeVOdd * 4 : 0) + (p0 >> 1) * 2);
    a1[k] = (unsigned)(((p1 & 1) ? kResizeVOdd * 4 : 0) + (p1 >> 1) * 2);
  }
  uint8_t* dst = a.dst + (size_t)frame * a.dst_stride + (size_t)(oy0 + rg * 8) * a.dpitch + x;
  const uint8_t* vb = reinterpret_cast<const uint8_t*>(s_v) + rg * 8 * (kResizeVPitch * 4);
  const int nr = min(8, nrows_out - rg * 8);
#pragma unroll 2
  for (int u = 0; u < nr; ++u) {
    uint32_t v[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const uint32_t V0 = *reinterpret_cast<const uint16_t*>(vb + a0[k]);
      const uint32_t V1 = *reinterpret_cast<const uint16_t*>(vb + a1[k]);
      // (c0x*V0 + c1x*V1 + 32768) >> 16 is <= 255 by construction: the result is byte 2 of the sum
      v[k] = (256u - c1x[k]) * V0 + c1x[k] * V1 + 32768u;
    }
    // dpitch is a multiple of 128, so the whole word is ours even when it runs past dw
    *reinterpret_cast<uint32_t*>(dst) =
        __byte_perm(__byte_perm(v[0], v[1], 0x0062), __byte_perm(v[2], v[3], 0x0062), 0x5410);
    dst += a.dpitch;
    vb += kResizeVPitch * 4;
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
  if (debug_sync() && cudaDeviceSynchronize() != cudaSuccess) {
    set_error("resize debug: src=%p dst=%p sstride=%zu dstride=%zu sw=%d sh=%d sp=%d dw=%d dh=%d dp=%d align=%d batch=%d xtab=%p ytab=%p",
              (const void*)a.src, (void*)a.dst, a.src_stride, a.dst_stride, a.sw, a.sh, a.spitch, a.dw, a.dh, a.dpitch,
              a.src_align, batch, (const void*)a.xtab, (const void*)a.ytab);
    return ORBX_E_CUDA;
  }
  ORBX_LAUNCH_CHECK();
  return ORBX_OK;
}

}  // namespace orbx
