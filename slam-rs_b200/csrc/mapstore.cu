// Device-resident map-descriptor store with projection cull.
//
// Replaces the hot part of mapping::Map (/root/reference/src/mapping/map.rs): `add_points` (:31-37),
// `project_point` (:50-69), `find_matches` (:72-118) together with `descriptors_to_mat` (:155-177),
// `update_observations` (:121-127) and `prune_outliers` (:130-137).  The reference clones every visible
// descriptor into a Vec<Vec<u8>> and rebuilds an OpenCV Mat byte by byte (N x 32 FFI calls per frame) before
// it can call the matcher; here positions, descriptors and observation counts live in HBM, the visibility test
// runs as one kernel over the whole map, the visible descriptors are gathered on the device (32 B per row,
// two 128-bit moves) and handed to the same Hamming kernel the FeatureMatcher uses.
//
// Arithmetic of project_point, restated: p_cam_i = ((r_i0*X + r_i1*Y) + r_i2*Z) + t_i in f64 without fused
// multiply-adds (Rust/nalgebra do not contract), visible iff Z_cam > 0 and 0 <= x < 4000 and 0 <= y < 3000
// with x = fx*(Xc/Zc) + cx, y = fy*(Yc/Zc) + cy (map.rs:58-66, :90), and only points that carry a descriptor
// take part (:91).  Point id == slot index (ids are handed out sequentially and never reused, :33-35).
// The reference iterates a HashMap, so its row order is unspecified; this store uses ascending id, which
// cannot change the result set (every query row is matched independently, min_dist is global).
#include <algorithm>

#include "common.cuh"

struct orbx_map {
  int device = 0;
  int capacity = 0;
  int used = 0;       // slots handed out so far == next_id
  int alive = 0;      // Map::size()
  int min_observations = 2;
  double fx = 0, fy = 0, cx = 0, cy = 0;
  cudaStream_t stream = nullptr;
  double* pos = nullptr;     // [capacity][3]
  uint8_t* desc = nullptr;   // [capacity][32]
  uint8_t* flags = nullptr;  // bit 0: alive, bit 1: has descriptor
  int* obs = nullptr;        // [capacity]
  // per-call scratch
  int* vis_slot = nullptr;       // [capacity]
  uint8_t* vis_desc = nullptr;   // [capacity][32]
  int* block_sums = nullptr;     // [blocks + 2]
  orbx_dmatch *raw = nullptr, *good = nullptr;  // [capacity]
  int* pairs = nullptr;          // [capacity][2]
  double* pose = nullptr;        // R (9, row major) + t (3)
};

namespace orbx {

constexpr int kMapBlock = 1024;
enum { MAP_ALIVE = 1, MAP_HAS_DESC = 2 };

struct MapCam {
  double fx, fy, cx, cy;
};

// map.rs:50-69 + the bounds test of :90; returns visibility and the projection
__device__ __forceinline__ bool project_visible(const double* p, const double* rt, MapCam c, double& x, double& y) {
  const double X = p[0], Y = p[1], Z = p[2];
  const double xc = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(rt[0], X), __dmul_rn(rt[1], Y)), __dmul_rn(rt[2], Z)), rt[9]);
  const double yc = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(rt[3], X), __dmul_rn(rt[4], Y)), __dmul_rn(rt[5], Z)), rt[10]);
  const double zc = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(rt[6], X), __dmul_rn(rt[7], Y)), __dmul_rn(rt[8], Z)), rt[11]);
  x = y = 0.0;
  if (!(zc > 0.0)) return false;  // map.rs:61 (p_cam.z <= 0 -> None; NaN is not visible either)
  x = __dadd_rn(__dmul_rn(c.fx, __ddiv_rn(xc, zc)), c.cx);
  y = __dadd_rn(__dmul_rn(c.fy, __ddiv_rn(yc, zc)), c.cy);
  return x >= 0.0 && x < 4000.0 && y >= 0.0 && y < 3000.0;
}

// pass 1: visibility flag per slot -> per-block counts.  pass 2 (after the scan of the block sums): ordered
// scatter of the slot indices and gather of the descriptors.
template <int PASS>
__global__ void __launch_bounds__(kMapBlock) map_cull_kernel(const double* pos, const uint8_t* desc, const uint8_t* flags,
                                                            int used, const double* rt, MapCam cam, int* block_sums,
                                                            int* vis_slot, uint8_t* vis_desc) {
  __shared__ int s_warp[32];
  __shared__ double s_rt[12];
  if (threadIdx.x < 12) s_rt[threadIdx.x] = rt[threadIdx.x];
  __syncthreads();
  const int i = blockIdx.x * kMapBlock + threadIdx.x;
  bool vis = false;
  if (i < used && (flags[i] & (MAP_ALIVE | MAP_HAS_DESC)) == (MAP_ALIVE | MAP_HAS_DESC)) {
    double x, y;
    vis = project_visible(pos + (size_t)i * 3, s_rt, cam, x, y);
  }
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned bal = __ballot_sync(0xFFFFFFFFu, vis);
  if (lane == 0) s_warp[warp] = __popc(bal);
  __syncthreads();
  if (warp == 0) {
    int v = s_warp[lane], incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int n = __shfl_up_sync(0xFFFFFFFFu, incl, o);
      if (lane >= o) incl += n;
    }
    s_warp[lane] = incl - v;
    if (PASS == 1 && lane == 31) block_sums[blockIdx.x] = incl;
  }
  if (PASS == 1) return;
  __syncthreads();
  if (vis) {
    const int at = block_sums[blockIdx.x] + s_warp[warp] + __popc(bal & ((1u << lane) - 1u));
    vis_slot[at] = i;
    const uint4* s = reinterpret_cast<const uint4*>(desc + (size_t)i * 32);
    uint4* d = reinterpret_cast<uint4*>(vis_desc + (size_t)at * 32);
    d[0] = __ldg(s);
    d[1] = __ldg(s + 1);
  }
}

// exclusive scan of the block sums in place (one block); total -> sums[n]
__global__ void __launch_bounds__(1024) map_scan_kernel(int* sums, int n) {
  __shared__ int s_part[1024];
  const int per = (n + 1023) / 1024;
  const int b = threadIdx.x * per, e = min(n, b + per);
  int acc = 0;
  for (int i = b; i < e; ++i) acc += sums[i];
  s_part[threadIdx.x] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    int run = 0;
    for (int i = 0; i < 1024; ++i) {
      const int v = s_part[i];
      s_part[i] = run;
      run += v;
    }
    sums[n] = run;
  }
  __syncthreads();
  int run = s_part[threadIdx.x];
  for (int i = b; i < e; ++i) {
    const int v = sums[i];
    sums[i] = run;
    run += v;
  }
}

// (map_id, kp_idx) = (visible_points[query_idx], train_idx)  — map.rs:111-115
__global__ void map_pairs_kernel(const orbx_dmatch* good, int n, const int* vis_slot, int* pairs) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  pairs[2 * i] = vis_slot[good[i].query_idx];
  pairs[2 * i + 1] = good[i].train_idx;
}

__global__ void map_project_kernel(const double* pos, const uint8_t* flags, int used, const double* rt, MapCam cam,
                                   double* xy, uint8_t* vis) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= used) return;
  double x, y;
  const bool v = project_visible(pos + (size_t)i * 3, rt, cam, x, y);
  xy[2 * i] = x;
  xy[2 * i + 1] = y;
  vis[i] = (uint8_t)((v ? 1 : 0) | ((flags[i] & MAP_ALIVE) ? 2 : 0) | ((flags[i] & MAP_HAS_DESC) ? 4 : 0));
}

__global__ void map_observe_kernel(const int* pairs, int n, int used, const uint8_t* flags, int* obs) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int id = pairs[2 * i];
  if (id >= 0 && id < used && (flags[id] & MAP_ALIVE)) atomicAdd(&obs[id], 1);  // map.rs:122-126
}

__global__ void map_prune_kernel(uint8_t* flags, const int* obs, int used, int min_obs, int* removed) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  bool kill = false;
  if (i < used && (flags[i] & MAP_ALIVE) && obs[i] < min_obs) {  // map.rs:133-134 retain(observations >= min)
    flags[i] &= (uint8_t)~MAP_ALIVE;
    kill = true;
  }
  const unsigned bal = __ballot_sync(0xFFFFFFFFu, kill);
  if ((threadIdx.x & 31) == 0 && bal) atomicAdd(removed, __popc(bal));
}

__global__ void map_count_kernel(const uint8_t* flags, const int* obs, int used, int min_obs, int* count) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const bool ok = i < used && (flags[i] & MAP_ALIVE) && obs[i] >= min_obs;
  const unsigned bal = __ballot_sync(0xFFFFFFFFu, ok);
  if ((threadIdx.x & 31) == 0 && bal) atomicAdd(count, __popc(bal));
}

}  // namespace orbx

using namespace orbx;

namespace {

template <typename T>
int map_alloc(T** p, size_t count) {
  cudaError_t e = cudaMalloc((void**)p, std::max<size_t>(count * sizeof(T), 256));
  if (e != cudaSuccess) {
    set_error("orbx_map: cudaMalloc(%zu) failed: %s", count * sizeof(T), cudaGetErrorString(e));
    return ORBX_E_NOMEM;
  }
  return ORBX_OK;
}

int upload_pose(orbx_map* m, const double* R, const double* t) {
  double h[12];
  for (int i = 0; i < 9; ++i) h[i] = R[i];
  for (int i = 0; i < 3; ++i) h[9 + i] = t[i];
  ORBX_CUDA(cudaMemcpyAsync(m->pose, h, sizeof(h), cudaMemcpyHostToDevice, m->stream));
  ORBX_CUDA(cudaStreamSynchronize(m->stream));  // h is a stack buffer
  return ORBX_OK;
}

}  // namespace

extern "C" {

int orbx_map_create(int device, int capacity, double fx, double fy, double cx, double cy, orbx_map** out) {
  if (!out || capacity <= 0) {
    set_error("orbx_map_create: bad argument");
    return ORBX_E_BADARG;
  }
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || device < 0 || device >= ndev) {
    cudaGetLastError();
    set_error("orbx_map_create: CUDA device %d unavailable (%s); liborbx has no CPU fallback", device,
              e != cudaSuccess ? cudaGetErrorString(e) : "index out of range");
    return ORBX_E_CUDA;
  }
  ORBX_CUDA(cudaSetDevice(device));
  orbx_map* m = new orbx_map();
  m->device = device;
  m->capacity = capacity;
  m->fx = fx;
  m->fy = fy;
  m->cx = cx;
  m->cy = cy;
  const size_t cap = (size_t)capacity;
  const size_t blocks = (cap + kMapBlock - 1) / kMapBlock;
  int rc = ORBX_OK;
  if ((e = cudaStreamCreateWithFlags(&m->stream, cudaStreamNonBlocking)) != cudaSuccess) rc = ORBX_E_CUDA;
  if (!rc) rc = map_alloc(&m->pos, cap * 3);
  if (!rc) rc = map_alloc(&m->desc, cap * 32);
  if (!rc) rc = map_alloc(&m->flags, cap);
  if (!rc) rc = map_alloc(&m->obs, cap);
  if (!rc) rc = map_alloc(&m->vis_slot, cap);
  if (!rc) rc = map_alloc(&m->vis_desc, cap * 32);
  if (!rc) rc = map_alloc(&m->block_sums, blocks + 2);
  if (!rc) rc = map_alloc(&m->raw, cap);
  if (!rc) rc = map_alloc(&m->good, cap);
  if (!rc) rc = map_alloc(&m->pairs, cap * 2);
  if (!rc) rc = map_alloc(&m->pose, 12);
  if (!rc && (cudaMemset(m->flags, 0, cap) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess)) rc = ORBX_E_CUDA;
  if (rc) {
    orbx_map_destroy(m);
    return rc;
  }
  *out = m;
  return ORBX_OK;
}

void orbx_map_destroy(orbx_map* m) {
  if (!m) return;
  cudaSetDevice(m->device);
  if (m->stream) cudaStreamSynchronize(m->stream);
  cudaFree(m->pos);
  cudaFree(m->desc);
  cudaFree(m->flags);
  cudaFree(m->obs);
  cudaFree(m->vis_slot);
  cudaFree(m->vis_desc);
  cudaFree(m->block_sums);
  cudaFree(m->raw);
  cudaFree(m->good);
  cudaFree(m->pairs);
  cudaFree(m->pose);
  if (m->stream) cudaStreamDestroy(m->stream);
  delete m;
}

int orbx_map_size(const orbx_map* m) { return m ? m->alive : ORBX_E_BADARG; }

int orbx_map_clear(orbx_map* m) {
  if (!m) return ORBX_E_BADARG;
  ORBX_CUDA(cudaSetDevice(m->device));
  ORBX_CUDA(cudaMemsetAsync(m->flags, 0, (size_t)m->capacity, m->stream));
  ORBX_CUDA(cudaStreamSynchronize(m->stream));
  m->used = m->alive = 0;  // map.rs:147-150: points.clear(); next_id = 0
  return ORBX_OK;
}

int orbx_map_add_points(orbx_map* m, const double* xyz, const uint8_t* desc, const uint8_t* has_desc,
                        const int* observations, int n, int64_t* first_id) {
  if (!m || n < 0 || (n > 0 && !xyz)) {
    set_error("orbx_map_add_points: bad argument");
    return ORBX_E_BADARG;
  }
  if (first_id) *first_id = m->used;
  if (n == 0) return ORBX_OK;
  if (m->used + n > m->capacity) {
    set_error("orbx_map_add_points: %d + %d points exceed the capacity %d", m->used, n, m->capacity);
    return ORBX_E_CAPACITY;
  }
  ORBX_CUDA(cudaSetDevice(m->device));
  const size_t at = (size_t)m->used;
  ORBX_CUDA(cudaMemcpyAsync(m->pos + at * 3, xyz, (size_t)n * 24, cudaMemcpyDefault, m->stream));
  std::vector<uint8_t> fl((size_t)n);
  for (int i = 0; i < n; ++i)
    fl[i] = (uint8_t)(MAP_ALIVE | ((desc && (!has_desc || has_desc[i])) ? MAP_HAS_DESC : 0));
  if (desc)
    ORBX_CUDA(cudaMemcpyAsync(m->desc + at * 32, desc, (size_t)n * 32, cudaMemcpyDefault, m->stream));
  else
    ORBX_CUDA(cudaMemsetAsync(m->desc + at * 32, 0, (size_t)n * 32, m->stream));
  ORBX_CUDA(cudaMemcpyAsync(m->flags + at, fl.data(), (size_t)n, cudaMemcpyHostToDevice, m->stream));
  std::vector<int> ob;
  if (observations) {
    ORBX_CUDA(cudaMemcpyAsync(m->obs + at, observations, (size_t)n * 4, cudaMemcpyDefault, m->stream));
  } else {
    ob.assign((size_t)n, 1);  // MapPoint::new / with_descriptor start at observations = 1 (triangulation.rs:29,39)
    ORBX_CUDA(cudaMemcpyAsync(m->obs + at, ob.data(), (size_t)n * 4, cudaMemcpyHostToDevice, m->stream));
  }
  ORBX_CUDA(cudaStreamSynchronize(m->stream));
  m->used += n;
  m->alive += n;
  return ORBX_OK;
}

int orbx_map_project(orbx_map* m, const double* R, const double* t, double* xy, uint8_t* vis) {
  if (!m || !R || !t || !xy || !vis) {
    set_error("orbx_map_project: bad argument");
    return ORBX_E_BADARG;
  }
  if (m->used == 0) return ORBX_OK;
  ORBX_CUDA(cudaSetDevice(m->device));
  int rc = upload_pose(m, R, t);
  if (rc) return rc;
  double* dxy = nullptr;
  uint8_t* dvis = nullptr;
  if ((rc = map_alloc(&dxy, (size_t)m->used * 2))) return rc;
  if ((rc = map_alloc(&dvis, (size_t)m->used))) {
    cudaFree(dxy);
    return rc;
  }
  const MapCam cam{m->fx, m->fy, m->cx, m->cy};
  map_project_kernel<<<div_up(m->used, 256), 256, 0, m->stream>>>(m->pos, m->flags, m->used, m->pose, cam, dxy, dvis);
  g_launches.fetch_add(1);
  cudaError_t e = cudaMemcpyAsync(xy, dxy, (size_t)m->used * 16, cudaMemcpyDeviceToHost, m->stream);
  if (e == cudaSuccess) e = cudaMemcpyAsync(vis, dvis, (size_t)m->used, cudaMemcpyDeviceToHost, m->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(m->stream);
  cudaFree(dxy);
  cudaFree(dvis);
  if (e != cudaSuccess) {
    set_error("orbx_map_project: %s", cudaGetErrorString(e));
    return ORBX_E_CUDA;
  }
  return ORBX_OK;
}

int orbx_map_find_matches(orbx_map* m, orbx_matcher* matcher, const double* R, const double* t,
                          const uint8_t* frame_desc, int n_frame, int* pairs, int cap, int* n_pairs, int* n_visible) {
  if (!m || !matcher || !R || !t || n_frame < 0 || (n_frame > 0 && !frame_desc) || !n_pairs || cap < 0 ||
      (cap > 0 && !pairs)) {
    set_error("orbx_map_find_matches: bad argument");
    return ORBX_E_BADARG;
  }
  *n_pairs = 0;
  if (n_visible) *n_visible = 0;
  if (m->used == 0) return ORBX_OK;
  ORBX_CUDA(cudaSetDevice(m->device));
  int rc = upload_pose(m, R, t);
  if (rc) return rc;
  const MapCam cam{m->fx, m->fy, m->cx, m->cy};
  const int blocks = div_up(m->used, kMapBlock);
  map_cull_kernel<1><<<blocks, kMapBlock, 0, m->stream>>>(m->pos, m->desc, m->flags, m->used, m->pose, cam, m->block_sums,
                                                        m->vis_slot, m->vis_desc);
  ORBX_LAUNCH_CHECK();
  map_scan_kernel<<<1, 1024, 0, m->stream>>>(m->block_sums, blocks);
  ORBX_LAUNCH_CHECK();
  map_cull_kernel<2><<<blocks, kMapBlock, 0, m->stream>>>(m->pos, m->desc, m->flags, m->used, m->pose, cam, m->block_sums,
                                                        m->vis_slot, m->vis_desc);
  ORBX_LAUNCH_CHECK();
  int nvis = 0;
  ORBX_CUDA(cudaMemcpyAsync(&nvis, m->block_sums + blocks, sizeof(int), cudaMemcpyDeviceToHost, m->stream));
  ORBX_CUDA(cudaStreamSynchronize(m->stream));
  if (n_visible) *n_visible = nvis;
  if (nvis == 0 || n_frame == 0) return ORBX_OK;  // map.rs:100-102 and the empty guard of matcher.rs:29-31
  // map.rs:107-108: match_descriptors(map_desc, descriptors) then filter_good_matches(.., 2.0)
  int nraw = 0, ngood = 0;
  if ((rc = orbx_match(matcher, m->vis_desc, nvis, frame_desc, n_frame, m->raw, &nraw))) return rc;
  if ((rc = orbx_filter_good_matches(matcher, m->raw, nraw, 2.0f, m->good, &ngood))) return rc;
  if ((rc = orbx_matcher_sync(matcher))) return rc;
  *n_pairs = ngood;
  if (ngood == 0) return ORBX_OK;
  map_pairs_kernel<<<div_up(ngood, 256), 256, 0, m->stream>>>(m->good, ngood, m->vis_slot, m->pairs);
  ORBX_LAUNCH_CHECK();
  const int take = std::min(ngood, cap);
  if (take)
    ORBX_CUDA(cudaMemcpyAsync(pairs, m->pairs, (size_t)take * 8, cudaMemcpyDefault, m->stream));
  ORBX_CUDA(cudaStreamSynchronize(m->stream));
  if (ngood > cap) {
    set_error("orbx_map_find_matches: %d matches exceed cap=%d", ngood, cap);
    return ORBX_E_CAPACITY;
  }
  return ORBX_OK;
}

int orbx_map_update_observations(orbx_map* m, const int* pairs, int n) {
  if (!m || n < 0 || (n > 0 && !pairs)) {
    set_error("orbx_map_update_observations: bad argument");
    return ORBX_E_BADARG;
  }
  if (n == 0) return ORBX_OK;
  if (n > m->capacity) {
    set_error("orbx_map_update_observations: n exceeds the capacity");
    return ORBX_E_BADARG;
  }
  ORBX_CUDA(cudaSetDevice(m->device));
  ORBX_CUDA(cudaMemcpyAsync(m->pairs, pairs, (size_t)n * 8, cudaMemcpyDefault, m->stream));
  map_observe_kernel<<<div_up(n, 256), 256, 0, m->stream>>>(m->pairs, n, m->used, m->flags, m->obs);
  ORBX_LAUNCH_CHECK();
  ORBX_CUDA(cudaStreamSynchronize(m->stream));
  return ORBX_OK;
}

static int count_into(orbx_map* m, bool prune, int* result) {
  ORBX_CUDA(cudaSetDevice(m->device));
  int zero = 0;
  ORBX_CUDA(cudaMemcpyAsync(m->block_sums, &zero, sizeof(int), cudaMemcpyHostToDevice, m->stream));
  if (m->used) {
    if (prune)
      map_prune_kernel<<<div_up(m->used, 256), 256, 0, m->stream>>>(m->flags, m->obs, m->used, m->min_observations,
                                                                  m->block_sums);
    else
      map_count_kernel<<<div_up(m->used, 256), 256, 0, m->stream>>>(m->flags, m->obs, m->used, m->min_observations,
                                                                  m->block_sums);
    ORBX_LAUNCH_CHECK();
  }
  ORBX_CUDA(cudaMemcpyAsync(result, m->block_sums, sizeof(int), cudaMemcpyDeviceToHost, m->stream));
  ORBX_CUDA(cudaStreamSynchronize(m->stream));
  return ORBX_OK;
}

int orbx_map_prune_outliers(orbx_map* m, int* removed) {
  if (!m || !removed) return ORBX_E_BADARG;
  int rc = count_into(m, true, removed);
  if (!rc) m->alive -= *removed;
  return rc;
}

int orbx_map_stable_count(orbx_map* m, int* count) {
  if (!m || !count) return ORBX_E_BADARG;
  return count_into(m, false, count);
}

int orbx_map_read_observations(orbx_map* m, int* obs, uint8_t* alive, int n) {
  if (!m || n < 0 || n > m->used) return ORBX_E_BADARG;
  if (n == 0) return ORBX_OK;
  ORBX_CUDA(cudaSetDevice(m->device));
  std::vector<uint8_t> fl((size_t)n);
  if (obs) ORBX_CUDA(cudaMemcpyAsync(obs, m->obs, (size_t)n * 4, cudaMemcpyDeviceToHost, m->stream));
  ORBX_CUDA(cudaMemcpyAsync(fl.data(), m->flags, (size_t)n, cudaMemcpyDeviceToHost, m->stream));
  ORBX_CUDA(cudaStreamSynchronize(m->stream));
  if (alive)
    for (int i = 0; i < n; ++i) alive[i] = fl[i] & MAP_ALIVE;
  return ORBX_OK;
}

}  // extern "C"
