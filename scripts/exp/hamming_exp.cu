// Experiment harness (not product): Hamming k-best kernel variants timed side by side on one GPU.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o hamming_exp hamming_exp.cu && ./hamming_exp
// Variants: NPOP = POPCs per 256-bit pair (8 plain, 6/5/4 with carry-save adders in LOP3),
//           FUSED = forward best-2 and reverse (train->query) best-1 from ONE distance evaluation
//                   (warp CREDUX.MIN of the reverse key, per-warp column minima in smem, RED.MIN to global).
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e_)); exit(1);} } while (0)

constexpr int kQPB = 128, kTile = 256;

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
  unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N)); }

__device__ __forceinline__ unsigned xor3(unsigned a, unsigned b, unsigned c) {
  unsigned r; asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r;
}
__device__ __forceinline__ unsigned maj3(unsigned a, unsigned b, unsigned c) {
  unsigned r; asm("lop3.b32 %0, %1, %2, %3, 0xE8;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r;
}

template <int NPOP>
__device__ __forceinline__ unsigned dist256(const uint4& qa, const uint4& qb, const uint4& x, const uint4& y) {
  const unsigned x0 = qa.x ^ x.x, x1 = qa.y ^ x.y, x2 = qa.z ^ x.z, x3 = qa.w ^ x.w;
  const unsigned x4 = qb.x ^ y.x, x5 = qb.y ^ y.y, x6 = qb.z ^ y.z, x7 = qb.w ^ y.w;
  if (NPOP == 8) {
    return __popc(x0) + __popc(x1) + __popc(x2) + __popc(x3) + __popc(x4) + __popc(x5) + __popc(x6) + __popc(x7);
  }
  const unsigned s1 = xor3(x0, x1, x2), c1 = maj3(x0, x1, x2);
  const unsigned s2 = xor3(x3, x4, x5), c2 = maj3(x3, x4, x5);
  if (NPOP == 6) {
    return (__popc(s1) + __popc(s2) + __popc(x6)) + __popc(x7) + 2 * (__popc(c1) + __popc(c2));
  }
  const unsigned s3 = xor3(s1, s2, x6), c3 = maj3(s1, s2, x6);
  if (NPOP == 5) {
    return (__popc(s3) + __popc(x7)) + 2 * (__popc(c1) + __popc(c2) + __popc(c3));
  }
  const unsigned t = xor3(c1, c2, c3), f = maj3(c1, c2, c3);
  return (__popc(s3) + __popc(x7)) + 2 * __popc(t) + 4 * __popc(f);
}

template <int K>
__device__ __forceinline__ void best_update(unsigned& b1, unsigned& b2, unsigned key) {
  if (K == 1) { b1 = min(b1, key); }
  else { unsigned hi = max(b1, key); b1 = min(b1, key); b2 = min(b2, hi); }
}

struct Args {
  const uint8_t* q; const uint8_t* t; size_t q_stride, t_stride; int nq, nt;
  uint2* out;        // [prob][nq] (b1, b2)
  unsigned* rev;     // [prob][nt] packed reverse key (dist<<16 | query), pre-set to 0xFFFFFFFF (FUSED only)
};

template <int K, int NPOP, bool FUSED, int UNROLL>
__global__ void __launch_bounds__(kQPB) ham_kernel(Args a) {
  __shared__ uint4 tile[2][kTile * 2];
  __shared__ unsigned s_col[FUSED ? kQPB / 32 : 1][FUSED ? kTile : 1];
  const int prob = blockIdx.z;
  const int nq = a.nq, nt = a.nt;
  const int q0 = blockIdx.x * kQPB;
  const int qi = q0 + threadIdx.x;
  const bool active = qi < nq;
  const uint4* qp = reinterpret_cast<const uint4*>(a.q + (size_t)prob * a.q_stride);
  const uint4* tp = reinterpret_cast<const uint4*>(a.t + (size_t)prob * a.t_stride);
  uint4 qa = make_uint4(0, 0, 0, 0), qb = qa;
  if (active) { qa = __ldg(qp + (size_t)qi * 2); qb = __ldg(qp + (size_t)qi * 2 + 1); }
  unsigned b1 = 0xFFFFFFFFu, b2 = 0xFFFFFFFFu;
  const unsigned qoff = active ? (unsigned)qi : 0x7F000000u;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ntiles = (nt + kTile - 1) / kTile;
  auto issue = [&](int tl, int bufi) {
    const int base = tl * kTile;
    const int rows = min(kTile, nt - base);
    const uint4* src = tp + (size_t)base * 2;
    for (int i = threadIdx.x; i < rows * 2; i += kQPB) cp_async16(&tile[bufi][i], src + i);
    cp_async_commit();
  };
  if (ntiles > 0) issue(0, 0);
  for (int tl = 0; tl < ntiles; ++tl) {
    const int cur = tl & 1;
    if (tl + 1 < ntiles) { issue(tl + 1, cur ^ 1); cp_async_wait<1>(); } else { cp_async_wait<0>(); }
    __syncthreads();
    const int base_local = tl * kTile;
    const int rows = min(kTile, nt - base_local);
    const uint4* tt = tile[cur];
    int j = 0;
#pragma unroll 1
    for (; j + UNROLL <= rows; j += UNROLL) {
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) {
        uint4 x = tt[(j + u) * 2], y = tt[(j + u) * 2 + 1];
        unsigned d = dist256<NPOP>(qa, qb, x, y);
        best_update<K>(b1, b2, (d << 16) + (unsigned)(base_local + j + u));
        if (FUSED) {
          unsigned r = __reduce_min_sync(0xFFFFFFFFu, (d << 16) + qoff);
          if (lane == 0) s_col[warp][j + u] = r;
        }
      }
    }
    for (; j < rows; ++j) {
      uint4 x = tt[j * 2], y = tt[j * 2 + 1];
      unsigned d = dist256<NPOP>(qa, qb, x, y);
      best_update<K>(b1, b2, (d << 16) + (unsigned)(base_local + j));
      if (FUSED) {
        unsigned r = __reduce_min_sync(0xFFFFFFFFu, (d << 16) + qoff);
        if (lane == 0) s_col[warp][j] = r;
      }
    }
    __syncthreads();
    if (FUSED) {
      unsigned* rv = a.rev + (size_t)prob * nt + base_local;
      for (int c = threadIdx.x; c < rows; c += kQPB) {
        unsigned m = min(min(s_col[0][c], s_col[1][c]), min(s_col[2][c], s_col[3][c]));
        atomicMin(rv + c, m);
      }
      // next tile's s_col writes happen after the next __syncthreads at the loop top
    }
  }
  if (!active) return;
  a.out[(size_t)prob * nq + qi] = make_uint2(b1, b2);
}

// ---------------- pipe microbenchmarks -----------------
template <int MODE>
__global__ void pipe_bench(unsigned* out, int iters, unsigned seed) {
  unsigned v[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = seed * (threadIdx.x + 1) + i * 0x9E3779B9u + blockIdx.x;
  unsigned acc = 0;
  for (int it = 0; it < iters; ++it) {
    if (MODE == 0) {  // 8 POPC
#pragma unroll
      for (int i = 0; i < 8; ++i) { unsigned p = __popc(v[i]); v[i] = v[i] + p + 0x01010101u; }
    } else if (MODE == 1) {  // 8 CREDUX.MIN
#pragma unroll
      for (int i = 0; i < 8; ++i) { unsigned p = __reduce_min_sync(0xFFFFFFFFu, v[i]); v[i] = v[i] + p + 0x01010101u; }
    } else if (MODE == 2) {  // 4 POPC + 4 CREDUX interleaved
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        unsigned p = (i & 1) ? __reduce_min_sync(0xFFFFFFFFu, v[i]) : __popc(v[i]);
        v[i] = v[i] + p + 0x01010101u;
      }
    } else if (MODE == 3) {  // 8 POPC + 2 CREDUX
#pragma unroll
      for (int i = 0; i < 8; ++i) { unsigned p = __popc(v[i]); v[i] = v[i] + p + 0x01010101u; }
      acc += __reduce_min_sync(0xFFFFFFFFu, v[0]) + __reduce_min_sync(0xFFFFFFFFu, v[1]);
    } else if (MODE == 4) {  // 8 POPC + 16 LOP3 (dependent chains on other regs)
#pragma unroll
      for (int i = 0; i < 8; ++i) { unsigned p = __popc(v[i]); v[i] = v[i] + p + 0x01010101u; }
#pragma unroll
      for (int i = 0; i < 16; ++i) acc = xor3(acc, v[i & 7], 0x5bd1e995u + i) ;
    } else if (MODE == 5) {  // 4 POPC + 24 LOP3/VIMNMX
#pragma unroll
      for (int i = 0; i < 4; ++i) { unsigned p = __popc(v[i]); v[i] = v[i] + p + 0x01010101u; }
#pragma unroll
      for (int i = 0; i < 24; ++i) acc = maj3(acc, v[i & 7], 0x5bd1e995u + i);
    }
  }
  unsigned r = acc;
#pragma unroll
  for (int i = 0; i < 8; ++i) r ^= v[i];
  if (r == 0x12345678u) out[0] = r;
}

template <int MODE>
static void run_pipe(const char* name, double ops_per_iter) {
  unsigned* d; CK(cudaMalloc(&d, 4));
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int iters = 1 << 13, threads = 256, blocks = 148 * 8;
  float best = 1e30f;
  for (int rep = 0; rep < 4; ++rep) {
    cudaEventRecord(e0);
    pipe_bench<MODE><<<blocks, threads>>>(d, iters, 12345u + rep);
    cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    if (rep > 0 && ms < best) best = ms;
  }
  double lanes_per_clk_sm = (double)blocks * threads * iters * ops_per_iter / (best * 1e-3) / 148.0 / 1.965e9;
  printf("pipe %-34s %8.3f ms  -> %6.2f lane-ops/clk/SM (counting %g ops/iter)\n", name, best, lanes_per_clk_sm, ops_per_iter);
  cudaFree(d);
}

// ---------------- driver -----------------
static std::vector<uint2> g_ref_out; static std::vector<unsigned> g_ref_rev;

template <int K, int NPOP, bool FUSED, int UNROLL>
static float run_variant(const char* name, Args a, int nprob, bool is_ref, bool check_rev) {
  dim3 grid((a.nq + kQPB - 1) / kQPB, 1, nprob);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e30f;
  for (int rep = 0; rep < 4; ++rep) {
    if (FUSED) CK(cudaMemset(a.rev, 0xFF, (size_t)nprob * a.nt * 4));
    cudaEventRecord(e0);
    ham_kernel<K, NPOP, FUSED, UNROLL><<<grid, kQPB>>>(a);
    cudaEventRecord(e1); CK(cudaEventSynchronize(e1)); CK(cudaGetLastError());
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    if (rep > 0 && ms < best) best = ms;
  }
  std::vector<uint2> h((size_t)nprob * a.nq);
  CK(cudaMemcpy(h.data(), a.out, h.size() * sizeof(uint2), cudaMemcpyDeviceToHost));
  const char* verdict = "";
  if (is_ref) { g_ref_out = h; verdict = "(ref)"; }
  else {
    size_t bad = 0;
    for (size_t i = 0; i < h.size(); ++i) if (h[i].x != g_ref_out[i].x || (K == 2 && h[i].y != g_ref_out[i].y)) ++bad;
    verdict = bad ? "MISMATCH" : "ok";
  }
  const char* rv = "";
  if (FUSED) {
    std::vector<unsigned> r((size_t)nprob * a.nt);
    CK(cudaMemcpy(r.data(), a.rev, r.size() * 4, cudaMemcpyDeviceToHost));
    if (check_rev && !g_ref_rev.empty()) {
      size_t bad = 0; for (size_t i = 0; i < r.size(); ++i) if (r[i] != g_ref_rev[i]) ++bad;
      rv = bad ? " rev MISMATCH" : " rev ok";
    }
  }
  double pairs = (double)nprob * a.nq * a.nt;
  printf("%-28s %8.3f ms  %8.1f Gpairs/s  %s%s\n", name, best, pairs / (best * 1e-3) / 1e9, verdict, rv);
  return best;
}

int main(int argc, char** argv) {
  const int nprob = argc > 1 ? atoi(argv[1]) : 250, n = argc > 2 ? atoi(argv[2]) : 2000;
  run_pipe<0>("8 POPC", 8);
  run_pipe<1>("8 CREDUX.MIN", 8);
  run_pipe<2>("4 POPC + 4 CREDUX (count 8)", 8);
  run_pipe<3>("8 POPC + 2 CREDUX (count POPC)", 8);
  run_pipe<4>("8 POPC + 16 LOP3 (count POPC)", 8);
  run_pipe<5>("4 POPC + 24 LOP3 (count LOP3)", 24);

  size_t bytes = (size_t)(nprob + 1) * n * 32;
  std::vector<uint8_t> h(bytes);
  uint64_t s = 88172645463325252ull;
  for (size_t i = 0; i < bytes; ++i) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; h[i] = (uint8_t)(s >> 32); }
  // make consecutive "frames" correlated: frame p+1 row i = frame p row (i*7+3)%n with a few flipped bits for half the rows
  for (int p = 1; p <= nprob; ++p)
    for (int i = 0; i < n; i += 2) {
      const uint8_t* src = &h[((size_t)(p - 1) * n + ((size_t)i * 7 + 3) % n) * 32];
      uint8_t* dst = &h[((size_t)p * n + i) * 32];
      for (int b = 0; b < 32; ++b) dst[b] = src[b] ^ (uint8_t)((dst[b] & dst[(b + 1) & 31] & dst[(b + 2) & 31]) & 0x11);
    }
  uint8_t* d; CK(cudaMalloc(&d, bytes)); CK(cudaMemcpy(d, h.data(), bytes, cudaMemcpyHostToDevice));
  Args a{}; a.q = d; a.t = d + (size_t)n * 32; a.q_stride = a.t_stride = (size_t)n * 32; a.nq = a.nt = n;
  CK(cudaMalloc(&a.out, (size_t)nprob * n * sizeof(uint2)));
  CK(cudaMalloc(&a.rev, (size_t)nprob * n * 4));

  // reverse reference: k=1 with roles swapped -> packed key (dist<<16 | query idx)
  {
    Args r = a; r.q = a.t; r.t = a.q;
    run_variant<1, 8, false, 4>("k1 reverse pass (8 POPC)", r, nprob, true, false);
    g_ref_rev.resize((size_t)nprob * n);
    for (size_t i = 0; i < g_ref_rev.size(); ++i) g_ref_rev[i] = g_ref_out[i].x;
    run_variant<1, 5, false, 4>("k1 reverse pass (5 POPC)", r, nprob, false, false);
    run_variant<1, 4, false, 4>("k1 reverse pass (4 POPC)", r, nprob, false, false);
  }
  run_variant<2, 8, false, 4>("k2 8POPC u4 (current)", a, nprob, true, false);
  run_variant<2, 8, false, 8>("k2 8POPC u8", a, nprob, false, false);
  run_variant<2, 6, false, 4>("k2 6POPC u4", a, nprob, false, false);
  run_variant<2, 5, false, 4>("k2 5POPC u4", a, nprob, false, false);
  run_variant<2, 5, false, 8>("k2 5POPC u8", a, nprob, false, false);
  run_variant<2, 4, false, 4>("k2 4POPC u4", a, nprob, false, false);
  run_variant<2, 4, false, 8>("k2 4POPC u8", a, nprob, false, false);
  run_variant<2, 8, true, 4>("k2+rev fused 8POPC u4", a, nprob, false, true);
  run_variant<2, 6, true, 4>("k2+rev fused 6POPC u4", a, nprob, false, true);
  run_variant<2, 5, true, 4>("k2+rev fused 5POPC u4", a, nprob, false, true);
  run_variant<2, 4, true, 4>("k2+rev fused 4POPC u4", a, nprob, false, true);
  run_variant<2, 4, true, 8>("k2+rev fused 4POPC u8", a, nprob, false, true);
  return 0;
}
