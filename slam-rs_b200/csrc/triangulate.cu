// liborbx — SURVEY.md §8 f4: mapping::Triangulator::triangulate (src/mapping/triangulation.rs:92-158) and the
// caller's transform_points_to_world (examples/point_cloud.rs:522-544) on the device.
//
// Per correspondence i (one thread): the 4x4 DLT system cv::triangulatePoints builds (f64, products and
// differences unfused), its right singular vector of the smallest singular value by one-sided Jacobi in f64,
// rounded to the CV_32F homogeneous point the reference reads (triangulation.rs:121-125), then the reference's
// own f64 arithmetic: the |w| < 1e-10 test, the division, the two cheirality tests (nalgebra's column-order
// accumulation, unfused).  A second single-block kernel compacts the survivors in input order (id = i) and
// attaches descriptor row i when i < desc_rows (:141-150).
#include "common.cuh"
#include "jacobi.cuh"

using orbx::div_up;
using orbx::g_launches;
using orbx::is_device_ptr;
using orbx::null_vector4;
using orbx::set_error;

struct orbx_triangulator {
  int device = 0;
  int max_points = 0;
  double fx = 0, fy = 0, cx = 0, cy = 0;
  double min_parallax_deg = 1.0;   // stored only, like the reference (triangulation.rs:54-56, 138)
  double max_reproj_error = 4.0;
  cudaStream_t own_stream = nullptr, stream = nullptr;
  float *pts1 = nullptr, *pts2 = nullptr;  // [max_points][2] staging of host inputs
  uint8_t* desc_in = nullptr;              // [max_points][32]
  double* xyz_all = nullptr;               // [max_points][3]
  float* p4 = nullptr;                     // [max_points][4]
  uint8_t* keep = nullptr;                 // [max_points]
  int32_t* ids = nullptr;                  // [max_points]
  double* xyz = nullptr;                   // [max_points][3]
  uint8_t* desc_out = nullptr;             // [max_points][32]
  uint8_t* has_desc = nullptr;             // [max_points]
  int* count = nullptr;                    // [1]
  int* count_host = nullptr;               // pinned
};

namespace orbx {

struct TriCams {
  double P1[12], P2[12];  // K [R | t], row major 3x4
  double R1[9], t1[3], R2[9], t2[3];
};

// r * p + t, third row only (triangulation.rs:186-197): ((r20*x + r21*y) + r22*z) + t2
__device__ __forceinline__ double cam_z(const double* R, const double* t, double x, double y, double z) {
  return __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(R[6], x), __dmul_rn(R[7], y)), __dmul_rn(R[8], z)), t[2]);
}

__device__ __forceinline__ void triangulate_point(const TriCams& cams, float2 q1, float2 q2, float* p4_out,
                                                  double* xyz_out, uint8_t* keep_out) {
  double a[4][4];
  {
    const double x1 = (double)q1.x, y1 = (double)q1.y, x2 = (double)q2.x, y2 = (double)q2.y;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      a[0][k] = __dsub_rn(__dmul_rn(x1, cams.P1[8 + k]), cams.P1[k]);
      a[1][k] = __dsub_rn(__dmul_rn(y1, cams.P1[8 + k]), cams.P1[4 + k]);
      a[2][k] = __dsub_rn(__dmul_rn(x2, cams.P2[8 + k]), cams.P2[k]);
      a[3][k] = __dsub_rn(__dmul_rn(y2, cams.P2[8 + k]), cams.P2[4 + k]);
    }
  }
  double h[4];
  null_vector4(a, h);
  const float fx = (float)h[0], fy = (float)h[1], fz = (float)h[2], fw = (float)h[3];
  if (p4_out) *reinterpret_cast<float4*>(p4_out) = make_float4(fx, fy, fz, fw);

  // triangulation.rs:121-136 in f64 on the f32 values
  const double w = (double)fw;
  bool ok = !(fabs(w) < 1e-10);
  const double X = __ddiv_rn((double)fx, w), Y = __ddiv_rn((double)fy, w), Z = __ddiv_rn((double)fz, w);
  ok = ok && cam_z(cams.R1, cams.t1, X, Y, Z) > 0.0 && cam_z(cams.R2, cams.t2, X, Y, Z) > 0.0;
  xyz_out[0] = X;
  xyz_out[1] = Y;
  xyz_out[2] = Z;
  *keep_out = ok ? 1 : 0;
}

__global__ void __launch_bounds__(128) triangulate_kernel(const float* __restrict__ pts1, const float* __restrict__ pts2,
                                                         int n, TriCams cams, float* __restrict__ p4,
                                                         double* __restrict__ xyz_all, uint8_t* __restrict__ keep) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  triangulate_point(cams, reinterpret_cast<const float2*>(pts1)[i], reinterpret_cast<const float2*>(pts2)[i],
                    p4 ? p4 + 4 * (size_t)i : nullptr, xyz_all + 3 * (size_t)i, keep + i);
}

// K [R | t] on the device, same unfused arithmetic as build_projection below
__device__ __forceinline__ void device_projection(double fx, double fy, double cx, double cy, const double* R,
                                                  const double* t, double* P) {
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const double r0 = j < 3 ? R[j] : t[0], r1 = j < 3 ? R[3 + j] : t[1], r2 = j < 3 ? R[6 + j] : t[2];
    P[j] = __dadd_rn(__dmul_rn(fx, r0), __dmul_rn(cx, r2));
    P[4 + j] = __dadd_rn(__dmul_rn(fy, r1), __dmul_rn(cy, r2));
    P[8 + j] = r2;
  }
}

// batched form: problem p = blockIdx.y with poses read from device memory (NULL first pose = identity / zero);
// rows at p * max_rows
__global__ void __launch_bounds__(128) triangulate_batch_kernel(const float* __restrict__ pts1, const float* __restrict__ pts2,
                                                               const int* __restrict__ n_dev, int max_rows,
                                                               const double* __restrict__ R1, const double* __restrict__ t1,
                                                               const double* __restrict__ R2, const double* __restrict__ t2,
                                                               double fx, double fy, double cx, double cy,
                                                               double* __restrict__ xyz_all, uint8_t* __restrict__ keep) {
  __shared__ TriCams cams;
  const int p = blockIdx.y;
  const int n = min(n_dev[p], max_rows);
  if (blockIdx.x * blockDim.x >= n) return;
  if (threadIdx.x < 9) {
    cams.R1[threadIdx.x] = R1 ? R1[9 * p + threadIdx.x] : (threadIdx.x % 4 == 0 ? 1.0 : 0.0);
    cams.R2[threadIdx.x] = R2[9 * p + threadIdx.x];
  }
  if (threadIdx.x < 3) {
    cams.t1[threadIdx.x] = t1 ? t1[3 * p + threadIdx.x] : 0.0;
    cams.t2[threadIdx.x] = t2[3 * p + threadIdx.x];
  }
  __syncthreads();
  if (threadIdx.x == 0) device_projection(fx, fy, cx, cy, cams.R1, cams.t1, cams.P1);
  if (threadIdx.x == 32) device_projection(fx, fy, cx, cy, cams.R2, cams.t2, cams.P2);
  __syncthreads();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const size_t at = (size_t)p * max_rows + i;
  triangulate_point(cams, reinterpret_cast<const float2*>(pts1)[at], reinterpret_cast<const float2*>(pts2)[at], nullptr,
                    xyz_all + 3 * at, keep + at);
}

// ordered compaction (one block): survivors keep their input order, id = input index (triangulation.rs:140)
__global__ void __launch_bounds__(1024) triangulate_compact_kernel(const double* __restrict__ xyz_all,
                                                                  const uint8_t* __restrict__ keep, int n,
                                                                  const uint8_t* __restrict__ desc_in, int desc_rows,
                                                                  int32_t* __restrict__ ids, double* __restrict__ xyz,
                                                                  uint8_t* __restrict__ desc_out,
                                                                  uint8_t* __restrict__ has_desc, int* __restrict__ count,
                                                                  const int* __restrict__ n_dev, int max_rows) {
  __shared__ int s_warp[32];
  __shared__ int s_base;
  // batched form: problem blockIdx.x owns rows [blockIdx.x * max_rows, +n_dev[blockIdx.x]) of every array
  if (n_dev) {
    const size_t off = (size_t)blockIdx.x * max_rows;
    n = min(n_dev[blockIdx.x], max_rows);
    xyz_all += 3 * off, keep += off, ids += off, xyz += 3 * off, count += blockIdx.x;
    if (desc_out) desc_out += 32 * off;
    if (has_desc) has_desc += off;
  }
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) s_base = 0;
  __syncthreads();
  for (int b = 0; b < n; b += 1024) {
    const int i = b + threadIdx.x;
    const bool k = i < n && keep[i];
    const int base = s_base;  // written before the barrier that ended the previous round
    const unsigned bal = __ballot_sync(0xFFFFFFFFu, k);
    if (lane == 0) s_warp[warp] = __popc(bal);
    __syncthreads();
    if (warp == 0) {
      const int c = s_warp[lane];
      int incl = c;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
        if (lane >= o) incl += t;
      }
      s_warp[lane] = incl - c;
      if (lane == 31) s_base = base + incl;
    }
    __syncthreads();
    if (k) {
      const int at = base + s_warp[warp] + __popc(bal & ((1u << lane) - 1u));
      ids[at] = i;
      xyz[3 * (size_t)at] = xyz_all[3 * (size_t)i];
      xyz[3 * (size_t)at + 1] = xyz_all[3 * (size_t)i + 1];
      xyz[3 * (size_t)at + 2] = xyz_all[3 * (size_t)i + 2];
      const bool hd = desc_in && i < desc_rows;
      if (has_desc) has_desc[at] = hd ? 1 : 0;
      if (desc_out) {
        uint4 d0 = make_uint4(0, 0, 0, 0), d1 = d0;
        if (hd) {
          d0 = __ldg(reinterpret_cast<const uint4*>(desc_in + (size_t)i * 32));
          d1 = __ldg(reinterpret_cast<const uint4*>(desc_in + (size_t)i * 32) + 1);
        }
        reinterpret_cast<uint4*>(desc_out + (size_t)at * 32)[0] = d0;
        reinterpret_cast<uint4*>(desc_out + (size_t)at * 32)[1] = d1;
      }
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) *count = s_base;
}

// point_cloud.rs:522-544: p_world = r_ctw * p + t_ctw with r_ctw = R^T, t_ctw = -r_ctw * t (both precomputed on
// the host in nalgebra's order); rows of r_ctw in rc[0..8], t_ctw in rc[9..11]
struct WorldXf {
  double rc[12];
};
__global__ void to_world_kernel(double* xyz, int n, WorldXf xf) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double x = xyz[3 * (size_t)i], y = xyz[3 * (size_t)i + 1], z = xyz[3 * (size_t)i + 2];
#pragma unroll
  for (int r = 0; r < 3; ++r)
    xyz[3 * (size_t)i + r] = __dadd_rn(
        __dadd_rn(__dadd_rn(__dmul_rn(xf.rc[3 * r], x), __dmul_rn(xf.rc[3 * r + 1], y)), __dmul_rn(xf.rc[3 * r + 2], z)),
        xf.rc[9 + r]);
}

// K * [R | t] the way cv::gemm's small-matrix path sums it (k ascending, unfused); the zero entries of K add
// exact zeros (triangulation.rs:161-183, camera.rs:39-52)
static void build_projection(double fx, double fy, double cx, double cy, const double* R, const double* t, double* P) {
  for (int j = 0; j < 4; ++j) {
    const volatile double r0 = j < 3 ? R[j] : t[0], r1 = j < 3 ? R[3 + j] : t[1], r2 = j < 3 ? R[6 + j] : t[2];
    volatile double a = fx * r0, b = cx * r2;
    P[j] = a + b;
    volatile double c = fy * r1, d = cy * r2;
    P[4 + j] = c + d;
    P[8 + j] = r2;
  }
}

template <class T>
static int tri_alloc(T** p, size_t n) {
  cudaError_t e = cudaMalloc((void**)p, (n ? n : 1) * sizeof(T));
  if (e != cudaSuccess) {
    set_error("orbx_triangulator: cudaMalloc of %zu bytes failed: %s", n * sizeof(T), cudaGetErrorString(e));
    return ORBX_E_CUDA;
  }
  return ORBX_OK;
}

}  // namespace orbx

using namespace orbx;

extern "C" {

int orbx_triangulator_create(int device, int max_points, double fx, double fy, double cx, double cy,
                             orbx_triangulator** out) {
  if (!out || max_points <= 0) {
    set_error("orbx_triangulator_create: bad argument");
    return ORBX_E_BADARG;
  }
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) {
    cudaGetLastError();
    set_error("orbx_triangulator_create: no CUDA device %d (there is no CPU fallback)", device);
    return ORBX_E_CUDA;
  }
  ORBX_CUDA(cudaSetDevice(device));
  auto* t = new orbx_triangulator();
  t->device = device;
  t->max_points = max_points;
  t->fx = fx, t->fy = fy, t->cx = cx, t->cy = cy;
  const size_t n = (size_t)max_points;
  int rc = ORBX_OK;
  if (cudaStreamCreateWithFlags(&t->own_stream, cudaStreamNonBlocking) != cudaSuccess) rc = ORBX_E_CUDA;
  t->stream = t->own_stream;
  if (!rc) rc = tri_alloc(&t->pts1, n * 2);
  if (!rc) rc = tri_alloc(&t->pts2, n * 2);
  if (!rc) rc = tri_alloc(&t->desc_in, n * 32);
  if (!rc) rc = tri_alloc(&t->xyz_all, n * 3);
  if (!rc) rc = tri_alloc(&t->p4, n * 4);
  if (!rc) rc = tri_alloc(&t->keep, n);
  if (!rc) rc = tri_alloc(&t->ids, n);
  if (!rc) rc = tri_alloc(&t->xyz, n * 3);
  if (!rc) rc = tri_alloc(&t->desc_out, n * 32);
  if (!rc) rc = tri_alloc(&t->has_desc, n);
  if (!rc) rc = tri_alloc(&t->count, 1);
  if (!rc && cudaMallocHost((void**)&t->count_host, sizeof(int)) != cudaSuccess) rc = ORBX_E_CUDA;
  if (rc) {
    orbx_triangulator_destroy(t);
    return rc;
  }
  *out = t;
  return ORBX_OK;
}

void orbx_triangulator_destroy(orbx_triangulator* t) {
  if (!t) return;
  cudaSetDevice(t->device);
  cudaFree(t->pts1), cudaFree(t->pts2), cudaFree(t->desc_in), cudaFree(t->xyz_all), cudaFree(t->p4);
  cudaFree(t->keep), cudaFree(t->ids), cudaFree(t->xyz), cudaFree(t->desc_out), cudaFree(t->has_desc);
  cudaFree(t->count);
  if (t->count_host) cudaFreeHost(t->count_host);
  if (t->own_stream) cudaStreamDestroy(t->own_stream);
  delete t;
}

int orbx_triangulator_set_thresholds(orbx_triangulator* t, double min_parallax_deg, double max_reproj_error) {
  if (!t) {
    set_error("orbx_triangulator_set_thresholds: bad argument");
    return ORBX_E_BADARG;
  }
  t->min_parallax_deg = min_parallax_deg;
  t->max_reproj_error = max_reproj_error;
  return ORBX_OK;
}

int orbx_triangulator_get_thresholds(const orbx_triangulator* t, double* min_parallax_deg, double* max_reproj_error) {
  if (!t) {
    set_error("orbx_triangulator_get_thresholds: bad argument");
    return ORBX_E_BADARG;
  }
  if (min_parallax_deg) *min_parallax_deg = t->min_parallax_deg;
  if (max_reproj_error) *max_reproj_error = t->max_reproj_error;
  return ORBX_OK;
}

int orbx_triangulate(orbx_triangulator* t, const double* R1, const double* t1, const double* R2, const double* t2,
                     const float* pts1, int n1, const float* pts2, int n2, const uint8_t* desc, int desc_rows,
                     int32_t* ids, double* xyz, uint8_t* out_desc, uint8_t* has_desc, float* points4d, int cap,
                     int* n_out) {
  if (!t || !R1 || !t1 || !R2 || !t2 || n1 < 0 || n2 < 0 || !n_out || cap < 0 || desc_rows < 0) {
    set_error("orbx_triangulate: bad argument");
    return ORBX_E_BADARG;
  }
  *n_out = 0;
  if (n1 != n2) {  // triangulation.rs:100-102
    set_error("Point arrays must have the same length");
    return ORBX_E_BADARG;
  }
  const int n = n1;
  if (n == 0) return ORBX_OK;  // :104-106
  if (!pts1 || !pts2 || !ids || !xyz || (desc_rows > 0 && !desc)) {
    set_error("orbx_triangulate: bad argument");
    return ORBX_E_BADARG;
  }
  if (n > t->max_points) {
    set_error("orbx_triangulate: %d points exceed max_points %d", n, t->max_points);
    return ORBX_E_CAPACITY;
  }
  ORBX_CUDA(cudaSetDevice(t->device));
  TriCams cams;
  build_projection(t->fx, t->fy, t->cx, t->cy, R1, t1, cams.P1);
  build_projection(t->fx, t->fy, t->cx, t->cy, R2, t2, cams.P2);
  for (int i = 0; i < 9; ++i) cams.R1[i] = R1[i], cams.R2[i] = R2[i];
  for (int i = 0; i < 3; ++i) cams.t1[i] = t1[i], cams.t2[i] = t2[i];

  const float *d1 = pts1, *d2 = pts2;
  if (!is_device_ptr(pts1)) {
    ORBX_CUDA(cudaMemcpyAsync(t->pts1, pts1, (size_t)n * 8, cudaMemcpyHostToDevice, t->stream));
    d1 = t->pts1;
  }
  if (!is_device_ptr(pts2)) {
    ORBX_CUDA(cudaMemcpyAsync(t->pts2, pts2, (size_t)n * 8, cudaMemcpyHostToDevice, t->stream));
    d2 = t->pts2;
  }
  const int drows = desc ? (desc_rows < n ? desc_rows : n) : 0;
  const uint8_t* dd = desc;
  if (drows > 0 && !is_device_ptr(desc)) {
    ORBX_CUDA(cudaMemcpyAsync(t->desc_in, desc, (size_t)drows * 32, cudaMemcpyHostToDevice, t->stream));
    dd = t->desc_in;
  }
  const bool ids_dev = is_device_ptr(ids), xyz_dev = is_device_ptr(xyz);
  const bool od_dev = out_desc && is_device_ptr(out_desc), hd_dev = has_desc && is_device_ptr(has_desc);
  if ((ids_dev || xyz_dev || od_dev || hd_dev) && cap < n) {
    set_error("orbx_triangulate: device outputs need cap >= n (%d < %d)", cap, n);
    return ORBX_E_CAPACITY;
  }
  triangulate_kernel<<<div_up(n, 128), 128, 0, t->stream>>>(d1, d2, n, cams, t->p4, t->xyz_all, t->keep);
  ORBX_LAUNCH_CHECK();
  triangulate_compact_kernel<<<1, 1024, 0, t->stream>>>(t->xyz_all, t->keep, n, drows > 0 ? dd : nullptr, drows,
                                                       ids_dev ? ids : t->ids, xyz_dev ? xyz : t->xyz,
                                                       out_desc ? (od_dev ? out_desc : t->desc_out) : nullptr,
                                                       has_desc ? (hd_dev ? has_desc : t->has_desc) : nullptr, t->count,
                                                       nullptr, 0);
  ORBX_LAUNCH_CHECK();
  ORBX_CUDA(cudaMemcpyAsync(t->count_host, t->count, sizeof(int), cudaMemcpyDeviceToHost, t->stream));
  if (points4d)
    ORBX_CUDA(cudaMemcpyAsync(points4d, t->p4, (size_t)n * 16, cudaMemcpyDefault, t->stream));
  ORBX_CUDA(cudaStreamSynchronize(t->stream));
  const int m = *t->count_host;
  *n_out = m;
  if (m > cap) {
    set_error("orbx_triangulate: %d points triangulated, cap %d", m, cap);
    return ORBX_E_CAPACITY;
  }
  if (!ids_dev) ORBX_CUDA(cudaMemcpyAsync(ids, t->ids, (size_t)m * 4, cudaMemcpyDeviceToHost, t->stream));
  if (!xyz_dev) ORBX_CUDA(cudaMemcpyAsync(xyz, t->xyz, (size_t)m * 24, cudaMemcpyDeviceToHost, t->stream));
  if (out_desc && !od_dev)
    ORBX_CUDA(cudaMemcpyAsync(out_desc, t->desc_out, (size_t)m * 32, cudaMemcpyDeviceToHost, t->stream));
  if (has_desc && !hd_dev)
    ORBX_CUDA(cudaMemcpyAsync(has_desc, t->has_desc, (size_t)m, cudaMemcpyDeviceToHost, t->stream));
  ORBX_CUDA(cudaStreamSynchronize(t->stream));
  return ORBX_OK;
}

int orbx_triangulator_set_stream(orbx_triangulator* t, void* stream) {
  if (!t) return ORBX_E_BADARG;
  t->stream = stream ? (cudaStream_t)stream : t->own_stream;
  return ORBX_OK;
}

int orbx_triangulate_batch(orbx_triangulator* t, const double* R1, const double* t1, const double* R2, const double* t2,
                           const float* pts1, const float* pts2, const int* n_dev, int npairs, int max_rows, int32_t* ids,
                           double* xyz, int* n_out) {
  if (!t || npairs < 0 || max_rows < 0) {
    set_error("orbx_triangulate_batch: bad argument");
    return ORBX_E_BADARG;
  }
  if (npairs == 0 || max_rows == 0) return ORBX_OK;
  if ((R1 && !is_device_ptr(R1)) || (t1 && !is_device_ptr(t1)) || !is_device_ptr(R2) || !is_device_ptr(t2) ||
      !is_device_ptr(pts1) || !is_device_ptr(pts2) || !is_device_ptr(n_dev) || !is_device_ptr(ids) || !is_device_ptr(xyz) ||
      !is_device_ptr(n_out)) {
    set_error("orbx_triangulate_batch: device pointers required");
    return ORBX_E_BADARG;
  }
  if ((size_t)npairs * max_rows > (size_t)t->max_points) {
    set_error("orbx_triangulate_batch: %d x %d rows exceed max_points %d", npairs, max_rows, t->max_points);
    return ORBX_E_CAPACITY;
  }
  ORBX_CUDA(cudaSetDevice(t->device));
  triangulate_batch_kernel<<<dim3(div_up(max_rows, 128), npairs), 128, 0, t->stream>>>(
      pts1, pts2, n_dev, max_rows, R1, t1, R2, t2, t->fx, t->fy, t->cx, t->cy, t->xyz_all, t->keep);
  ORBX_LAUNCH_CHECK();
  triangulate_compact_kernel<<<npairs, 1024, 0, t->stream>>>(t->xyz_all, t->keep, 0, nullptr, 0, ids, xyz, nullptr, nullptr,
                                                             n_out, n_dev, max_rows);
  ORBX_LAUNCH_CHECK();
  return ORBX_OK;
}

int orbx_transform_points_to_world(orbx_triangulator* t, const double* R, const double* tv, double* xyz, int n) {
  if (!t || !R || !tv || n < 0 || (n > 0 && !xyz)) {
    set_error("orbx_transform_points_to_world: bad argument");
    return ORBX_E_BADARG;
  }
  if (n == 0) return ORBX_OK;
  ORBX_CUDA(cudaSetDevice(t->device));
  WorldXf xf;
  // r_ctw = R^T; t_ctw = (-r_ctw) * t, columns accumulated left to right (nalgebra gemv), unfused
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) xf.rc[3 * r + c] = R[3 * c + r];
  for (int r = 0; r < 3; ++r) {
    volatile double a = -xf.rc[3 * r] * tv[0], b = -xf.rc[3 * r + 1] * tv[1], c = -xf.rc[3 * r + 2] * tv[2];
    volatile double ab = a + b;
    xf.rc[9 + r] = ab + c;
  }
  double* d = xyz;
  const bool dev = is_device_ptr(xyz);
  if (!dev) {
    if (n > t->max_points) {
      set_error("orbx_transform_points_to_world: %d points exceed max_points %d", n, t->max_points);
      return ORBX_E_CAPACITY;
    }
    ORBX_CUDA(cudaMemcpyAsync(t->xyz_all, xyz, (size_t)n * 24, cudaMemcpyHostToDevice, t->stream));
    d = t->xyz_all;
  }
  to_world_kernel<<<div_up(n, 256), 256, 0, t->stream>>>(d, n, xf);
  ORBX_LAUNCH_CHECK();
  if (!dev) ORBX_CUDA(cudaMemcpyAsync(xyz, d, (size_t)n * 24, cudaMemcpyDeviceToHost, t->stream));
  ORBX_CUDA(cudaStreamSynchronize(t->stream));
  return ORBX_OK;
}

}  // extern "C"
