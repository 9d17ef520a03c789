// Frame ingest — BGR (8UC3) -> gray (8UC1), the step immediately in front of S1.
//
// The reference converts every decoded frame with imgproc::cvt_color(COLOR_BGR2GRAY)
// (/root/reference/examples/visual_odometry.rs:93-99, src/main.rs:43-49), and cv::ORB itself converts a
// non-8UC1 input the same way before building the pyramid (reached from src/feature/detector.rs:22,32).
// Arithmetic pinned against cv2 4.13 (oracle/restated.py::bgr_to_gray): 15-bit fixed point,
//   gray = (B*3735 + G*19235 + R*9798 + 16384) >> 15.
// HBM-bound streaming kernel: each thread converts 4 pixels (12 source bytes as three aligned words when the
// row is 4-byte aligned) and stores one 32-bit word into the detector's pitched level-0 plane.
#include "detector.cuh"

namespace orbx {

__device__ __forceinline__ unsigned gray15(unsigned b, unsigned g, unsigned r) {
  return (b * 3735u + g * 19235u + r * 9798u + 16384u) >> 15;
}

__global__ void __launch_bounds__(256) bgr_to_gray_kernel(const uint8_t* __restrict__ src, size_t spitch,
                                                          size_t sstride, uint8_t* __restrict__ dst, int dpitch,
                                                          size_t dstride, int w, int h, int aligned4) {
  const int x4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
  const int y = blockIdx.y;
  if (x4 >= w) return;
  const uint8_t* s = src + (size_t)blockIdx.z * sstride + (size_t)y * spitch + (size_t)x4 * 3;
  uint8_t* d = dst + (size_t)blockIdx.z * dstride + (size_t)y * dpitch + x4;
  unsigned out;
  if (aligned4 && x4 + 3 < w) {
    const unsigned w0 = __ldg(reinterpret_cast<const unsigned*>(s));      // B0 G0 R0 B1
    const unsigned w1 = __ldg(reinterpret_cast<const unsigned*>(s) + 1);  // G1 R1 B2 G2
    const unsigned w2 = __ldg(reinterpret_cast<const unsigned*>(s) + 2);  // R2 B3 G3 R3
    const unsigned g0 = gray15(w0 & 0xFF, (w0 >> 8) & 0xFF, (w0 >> 16) & 0xFF);
    const unsigned g1 = gray15(w0 >> 24, w1 & 0xFF, (w1 >> 8) & 0xFF);
    const unsigned g2 = gray15((w1 >> 16) & 0xFF, w1 >> 24, w2 & 0xFF);
    const unsigned g3 = gray15((w2 >> 8) & 0xFF, (w2 >> 16) & 0xFF, w2 >> 24);
    out = g0 | (g1 << 8) | (g2 << 16) | (g3 << 24);
  } else {
    out = 0;
    for (int k = 0; k < 4 && x4 + k < w; ++k)
      out |= gray15(__ldg(s + 3 * k), __ldg(s + 3 * k + 1), __ldg(s + 3 * k + 2)) << (8 * k);
  }
  // dpitch is a multiple of 128: the whole word belongs to this row even when it runs past w
  *reinterpret_cast<unsigned*>(d) = out;
}

// Repack of device-resident 8UC1 frames whose base / pitch / frame stride is not 4-byte aligned (e.g. tightly packed
// 1241-byte rows, which is how frames arrive from the host with a plain contiguous DMA) into the detector's pitched
// level-0 plane, so that every later stage takes its aligned path.  Thread = 16 output bytes: the five aligned source
// words that cover them, four funnel shifts, one 128-bit store.  Streaming: reads and writes each level-0 byte once.
__global__ void __launch_bounds__(256) repack_gray_kernel(const uint8_t* __restrict__ src, size_t spitch, size_t sstride,
                                                          const uint8_t* __restrict__ src_end, uint8_t* __restrict__ dst,
                                                          int dpitch, size_t dstride, int w) {
  const int x16 = (blockIdx.x * blockDim.x + threadIdx.x) * 16;
  if (x16 >= w) return;
  const uint8_t* s = src + (size_t)blockIdx.z * sstride + (size_t)blockIdx.y * spitch + x16;
  uint8_t* d = dst + (size_t)blockIdx.z * dstride + (size_t)blockIdx.y * dpitch + x16;
  const uintptr_t sa = (uintptr_t)s;
  const unsigned* a = reinterpret_cast<const unsigned*>(sa & ~(uintptr_t)3);
  uint4 out;
  if (reinterpret_cast<const uint8_t*>(a + 5) <= src_end) {
    const unsigned sh = 8 * (unsigned)(sa & 3);
    const unsigned w0 = __ldg(a), w1 = __ldg(a + 1), w2 = __ldg(a + 2), w3 = __ldg(a + 3), w4 = __ldg(a + 4);
    out = make_uint4(__funnelshift_r(w0, w1, sh), __funnelshift_r(w1, w2, sh), __funnelshift_r(w2, w3, sh),
                     __funnelshift_r(w3, w4, sh));
  } else {  // the last words of the buffer: never read past its end
    unsigned o[4] = {0, 0, 0, 0};
    for (int k = 0; k < 16 && s + k < src_end; ++k) o[k >> 2] |= (unsigned)__ldg(s + k) << (8 * (k & 3));
    out = make_uint4(o[0], o[1], o[2], o[3]);
  }
  // dpitch is a multiple of 128 and the plane is 256-byte aligned: the 16 bytes belong to this row even past w
  *reinterpret_cast<uint4*>(d) = out;
}

int launch_repack_gray(const uint8_t* src, size_t spitch, size_t sstride, uint8_t* dst, int dpitch, size_t dstride, int w,
                       int h, int batch, cudaStream_t s) {
  if (batch == 0) return ORBX_OK;
  const uint8_t* end = src + (size_t)(batch - 1) * sstride + (size_t)(h - 1) * spitch + w;
  dim3 grid(div_up(div_up(w, 16), 128), h, batch);
  repack_gray_kernel<<<grid, 128, 0, s>>>(src, spitch, sstride, end, dst, dpitch, dstride, w);
  ORBX_LAUNCH_CHECK();
  return ORBX_OK;
}

int launch_bgr_to_gray(const uint8_t* src, size_t spitch, size_t sstride, uint8_t* dst, int dpitch, size_t dstride,
                       int w, int h, int batch, cudaStream_t s) {
  if (batch == 0) return ORBX_OK;
  const int aligned4 = ((((uintptr_t)src) | spitch | sstride) & 3) == 0;
  dim3 grid(div_up(div_up(w, 4), 256), h, batch);
  bgr_to_gray_kernel<<<grid, 256, 0, s>>>(src, spitch, sstride, dst, dpitch, dstride, w, h, aligned4);
  ORBX_LAUNCH_CHECK();
  return ORBX_OK;
}

}  // namespace orbx
