// tcgen05.mma issue-rate probe (B200, sm_100a): what does ONE dispatch of a 128 x N x 32-byte MMA cost, per kind, layout and
// operand source, with 1 or 148 CTAs running?  One thread per CTA issues `tiles` groups of 9 MMAs (the Hamming matcher's
// K loop) into two alternating TMEM accumulators, one commit per group; it waits for group i-1 before it issues group
// i+1 (the matcher's double buffering) or only at the end (depth = all).  Cycles from %clock64, wall time from
// %globaltimer: their ratio is the SM clock the run really had.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_rate mma_rate.cu && ./mma_rate
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

#define CK(x)                                                                                   \
  do {                                                                                          \
    cudaError_t e_ = (x);                                                                       \
    if (e_ != cudaSuccess) {                                                                    \
      printf("%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e_));                         \
      return 1;                                                                                 \
    }                                                                                           \
  } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t done;
  do {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t"
        "}"
        : "=r"(done)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
  } while (!done);
}
__device__ __forceinline__ void mma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// K-major descriptors: no swizzle (8 x 16 B core matrices, LBO 128 B, SBO = 8-row group pitch) or 128-byte swizzle atoms
__device__ __forceinline__ uint64_t desc_none(uint32_t saddr, int sbo) {
  return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)((128 >> 4) & 0x3FFF) << 16) | ((uint64_t)((sbo >> 4) & 0x3FFF) << 32) |
         ((uint64_t)1 << 46);
}
__device__ __forceinline__ uint64_t desc_sw128(uint32_t saddr) {
  return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)1 << 16) | ((uint64_t)((1024 >> 4) & 0x3FFF) << 32) | ((uint64_t)1 << 46) |
         ((uint64_t)2 << 61);
}

template <int KIND>  // 0 = i8, 1 = f8f6f4 (e4m3), 2 = f16 (bf16)
__device__ __forceinline__ void mma_ss(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  if (KIND == 0)
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(d),
                 "l"(a), "l"(b), "r"(idesc), "r"(acc)
                 : "memory");
  else if (KIND == 1)
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}" ::"r"(d),
                 "l"(a), "l"(b), "r"(idesc), "r"(acc)
                 : "memory");
  else
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d),
                 "l"(a), "l"(b), "r"(idesc), "r"(acc)
                 : "memory");
}
// the same, called by a whole converged warp: operands are warp-uniform, one elected lane issues
template <int KIND>
__device__ __forceinline__ void mma_ss_elect(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  if (KIND == 0)
    asm volatile(
        "{\n\t.reg .pred p, e;\n\telect.sync _|e, 0xffffffff;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "@e tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(d),
        "l"(a), "l"(b), "r"(idesc), "r"(acc)
        : "memory");
  else if (KIND == 1)
    asm volatile(
        "{\n\t.reg .pred p, e;\n\telect.sync _|e, 0xffffffff;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "@e tcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}" ::"r"(d),
        "l"(a), "l"(b), "r"(idesc), "r"(acc)
        : "memory");
  else
    asm volatile(
        "{\n\t.reg .pred p, e;\n\telect.sync _|e, 0xffffffff;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "@e tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d),
        "l"(a), "l"(b), "r"(idesc), "r"(acc)
        : "memory");
}
__device__ __forceinline__ void mma_commit_elect(uint64_t* bar) {
  asm volatile(
      "{\n\t.reg .pred e;\n\telect.sync _|e, 0xffffffff;\n\t"
      "@e tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}" ::"r"(smem_u32(bar))
      : "memory");
}
// A from TMEM (the .ts form): a = TMEM address of the A tile
template <int KIND>
__device__ __forceinline__ void mma_ts(uint32_t d, uint32_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  if (KIND == 0)
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d),
                 "r"(a), "l"(b), "r"(idesc), "r"(acc)
                 : "memory");
  else if (KIND == 1)
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f8f6f4 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d),
                 "r"(a), "l"(b), "r"(idesc), "r"(acc)
                 : "memory");
  else
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d),
                 "r"(a), "l"(b), "r"(idesc), "r"(acc)
                 : "memory");
}

struct Result {
  long long cycles;
  unsigned long long ns;
};

constexpr int kRowBytes = 320;                 // the matcher's expanded rows: 10 K steps of 32 bytes
constexpr int kGroup = 8 * kRowBytes;          // no-swizzle 8-row group pitch
constexpr int kSmem = (128 + 256) * kRowBytes + 1024;

// layout: 0 = no swizzle, 1 = 128-byte swizzle; src: 0 = A from shared memory, 1 = A from TMEM; depth: groups in flight
template <int KIND>
__global__ void __launch_bounds__(128, 1) rate_kernel(int N, int layout, int src, int tiles, int depth, int kper, Result* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bars[3];
  __shared__ uint32_t s_tmem;
  for (int i = threadIdx.x; i < kSmem / 4; i += blockDim.x)
    reinterpret_cast<uint32_t*>(smem)[i] = KIND == 2 ? 0x3F803F80u : 0x01FF01FFu * (1 + (i & 1));  // small finite values
  if (threadIdx.x == 0) {
    mbar_init(bars, 1);
    mbar_init(bars + 1, 1);
    mbar_init(bars + 2, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = s_tmem;
  if (src == 2 && threadIdx.x < 32) {
    // the whole warp walks the loop (every value below is warp-uniform), one elected lane issues
    const uint32_t cfmt = KIND == 0 ? 2u : 1u, abfmt = KIND == 1 ? 0u : 1u;
    const uint32_t idesc = (cfmt << 4) | (abfmt << 7) | (abfmt << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    const uint32_t sa = smem_u32(smem), sb = sa + 128 * kRowBytes;
    uint32_t ph[2] = {0, 0};
    long long c0 = clock64();
    unsigned long long t0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (int t = 0; t < tiles; ++t) {
      const int buf = t & 1;
      if (t >= depth) {
        const int wb = (t - depth) & 1;
        mbar_wait(bars + wb, ph[wb] & 1);
        ++ph[wb];
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      }
      const uint32_t d = tmem + buf * 256;
#pragma unroll
      for (int k = 0; k < 9; ++k)
        if (k < kper) mma_ss_elect<KIND>(d, desc_none(sa + k * 256, kGroup), desc_none(sb + k * 256, kGroup), idesc, k > 0);
      mma_commit_elect(bars + buf);
    }
    mma_commit_elect(bars + 2);
    mbar_wait(bars + 2, 0);
    long long c1 = clock64();
    unsigned long long t1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
    if (threadIdx.x == 0) {
      out[blockIdx.x].cycles = c1 - c0;
      out[blockIdx.x].ns = t1 - t0;
    }
  } else if (src < 2 && threadIdx.x == 0) {
    const uint32_t cfmt = KIND == 0 ? 2u : 1u, abfmt = KIND == 1 ? 0u : 1u;
    const uint32_t idesc = (cfmt << 4) | (abfmt << 7) | (abfmt << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    const uint32_t sa = smem_u32(smem), sb = sa + 128 * kRowBytes;
    // accumulators: two buffers of N columns; the TMEM A tile (src 1) sits behind them: 128 lanes x 8 columns per K step
    const uint32_t ta = tmem + 384;
    uint32_t ph[2] = {0, 0};
    long long c0 = clock64();
    unsigned long long t0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (int t = 0; t < tiles; ++t) {
      const int buf = t & 1;
      if (t >= depth) {  // wait for group t - depth (same buffer when depth == 2)
        const int wb = (t - depth) & 1;
        mbar_wait(bars + wb, ph[wb] & 1);
        ++ph[wb];
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      }
      const uint32_t d = tmem + (src ? (N == 128 ? buf * 128 : 0) : buf * 256);  // src 1: the A tile needs columns too
      for (int k = 0; k < kper; ++k) {
        const uint64_t bd = layout ? desc_sw128(sb + (k >> 2) * 32768 + (k & 3) * 32) : desc_none(sb + k * 256, kGroup);
        if (src)
          mma_ts<KIND>(d, ta + k * 8, bd, idesc, k > 0);
        else
          mma_ss<KIND>(d, layout ? desc_sw128(sa + (k >> 2) * 16384 + (k & 3) * 32) : desc_none(sa + k * 256, kGroup), bd, idesc,
                       k > 0);
      }
      mma_commit(bars + buf);
    }
    mma_commit(bars + 2);  // a commit tracks every MMA issued before it: one wait covers the tail
    mbar_wait(bars + 2, 0);
    long long c1 = clock64();
    unsigned long long t1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
    out[blockIdx.x].cycles = c1 - c0;
    out[blockIdx.x].ns = t1 - t0;
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512));
}

// TMEM read rate: `nwarps` warps (warp w reads lanes 32 (w % 4) ..+31) read 256 columns with tcgen05.ld 32x32b.x32, `iters`
// times; reported per 128 lanes x 256 columns x 4 B = 128 KB (one accumulator tile of the matcher)
template <int inflight>
__global__ void __launch_bounds__(512, 1) ldtm_kernel(int iters, Result* out, int* sink) {
  __shared__ uint32_t s_tmem;
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = s_tmem;
  const int warp = threadIdx.x >> 5;
  const uint32_t tbase = tmem + ((uint32_t)(32 * (warp & 3)) << 16) + (warp >> 2) * (256 / (blockDim.x / 128));
  const int ncols = 256 / (blockDim.x / 128);  // columns per warp: the warps that share a lane quarter split the 256 columns
  int acc = 0;
  __syncthreads();
  const long long c0 = clock64();
  for (int it = 0; it < iters; ++it) {
    for (int c = 0; c < ncols; c += 32 * inflight) {
      int v[inflight][32];
#pragma unroll
      for (int k = 0; k < inflight; ++k)
        asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
            "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
            "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
            : "=r"(v[k][0]), "=r"(v[k][1]), "=r"(v[k][2]), "=r"(v[k][3]), "=r"(v[k][4]), "=r"(v[k][5]), "=r"(v[k][6]), "=r"(v[k][7]),
              "=r"(v[k][8]), "=r"(v[k][9]), "=r"(v[k][10]), "=r"(v[k][11]), "=r"(v[k][12]), "=r"(v[k][13]), "=r"(v[k][14]),
              "=r"(v[k][15]), "=r"(v[k][16]), "=r"(v[k][17]), "=r"(v[k][18]), "=r"(v[k][19]), "=r"(v[k][20]), "=r"(v[k][21]),
              "=r"(v[k][22]), "=r"(v[k][23]), "=r"(v[k][24]), "=r"(v[k][25]), "=r"(v[k][26]), "=r"(v[k][27]), "=r"(v[k][28]),
              "=r"(v[k][29]), "=r"(v[k][30]), "=r"(v[k][31])
            : "r"(tbase + c + 32 * k)
            : "memory");
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
      for (int k = 0; k < inflight; ++k) acc ^= v[k][0] ^ v[k][31];
    }
  }
  __syncthreads();
  const long long c1 = clock64();
  if (threadIdx.x == 0) out[blockIdx.x].cycles = c1 - c0;
  if (acc == 0x12345678) *sink = acc;
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512));
}

// cta_group::2: a cluster of two CTAs, CTA 0's warp 0 issues M = 256 MMAs (A tile and B half of each CTA in its own
// shared memory), multicast commit to both CTAs; cycles per MMA as seen by the issuer
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1) rate2_kernel(int N, int tiles, int depth, int kper, Result* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bars[3];
  __shared__ uint32_t s_tmem;
  uint32_t rank;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
  for (int i = threadIdx.x; i < kSmem / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x01FF01FFu * (1 + (i & 1));
  if (threadIdx.x == 0) {
    mbar_init(bars, 1);
    mbar_init(bars + 1, 1);
    mbar_init(bars + 2, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::);
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = s_tmem;
  if (rank == 0 && threadIdx.x < 32) {
    const uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(256 >> 4) << 24);
    const uint32_t sa = smem_u32(smem), sb = sa + 128 * kRowBytes;
    uint32_t ph[2] = {0, 0};
    long long c0 = clock64();
    unsigned long long t0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (int t = 0; t < tiles; ++t) {
      const int buf = t & 1;
      if (t >= depth) {
        const int wb = (t - depth) & 1;
        mbar_wait(bars + wb, ph[wb] & 1);
        ++ph[wb];
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      }
      const uint32_t d = tmem + buf * 256;
#pragma unroll
      for (int k = 0; k < 9; ++k)
        if (k < kper)
          asm volatile(
              "{\n\t.reg .pred p, e;\n\telect.sync _|e, 0xffffffff;\n\tsetp.ne.b32 p, %4, 0;\n\t"
              "@e tcgen05.mma.cta_group::2.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(d),
              "l"(desc_none(sa + k * 256, kGroup)), "l"(desc_none(sb + k * 256, kGroup)), "r"(idesc), "r"((uint32_t)(k > 0))
              : "memory");
      asm volatile(
          "{\n\t.reg .pred e;\n\t.reg .b16 m;\n\tmov.b16 m, 3;\n\telect.sync _|e, 0xffffffff;\n\t"
          "@e tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], m;\n\t}" ::"r"(
              smem_u32(bars + buf))
          : "memory");
    }
    asm volatile(
        "{\n\t.reg .pred e;\n\t.reg .b16 m;\n\tmov.b16 m, 3;\n\telect.sync _|e, 0xffffffff;\n\t"
        "@e tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], m;\n\t}" ::"r"(
            smem_u32(bars + 2))
        : "memory");
    mbar_wait(bars + 2, 0);
    long long c1 = clock64();
    unsigned long long t1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
    if (threadIdx.x == 0) {
      out[blockIdx.x >> 1].cycles = c1 - c0;
      out[blockIdx.x >> 1].ns = t1 - t0;
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512));
}

template <int KIND>
int run(const char* kind, int grid, int N, int layout, int src, int tiles, int depth, int kper, Result* d_out) {
  CK(cudaFuncSetAttribute(rate_kernel<KIND>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
  for (int rep = 0; rep < 2; ++rep) rate_kernel<KIND><<<grid, 128, kSmem>>>(N, layout, src, tiles, depth, kper, d_out);
  CK(cudaDeviceSynchronize());
  Result h[148];
  CK(cudaMemcpy(h, d_out, sizeof(Result) * grid, cudaMemcpyDeviceToHost));
  long long cmax = 0;
  unsigned long long nmax = 0;
  for (int i = 0; i < grid; ++i) {
    if (h[i].cycles > cmax) cmax = h[i].cycles;
    if (h[i].ns > nmax) nmax = h[i].ns;
  }
  const double per = (double)cmax / ((double)tiles * kper);
  printf("%-5s grid %3d N %3d %-7s A:%-26s depth %4d k/commit %d : %7.1f cycles per MMA, %6.1f ns per group, clock %.0f MHz\n", kind, grid,
         N, layout ? "sw128" : "noswz", src == 2 ? "smem, warp issues (elect)" : src ? "tmem" : "smem", depth, kper, per, (double)nmax / tiles, 1e3 * cmax / (double)nmax);
  return 0;
}

int main() {
  setvbuf(stdout, nullptr, _IONBF, 0);
  Result* d_out;
  CK(cudaMalloc(&d_out, sizeof(Result) * 148));
  {
    int* sink;
    CK(cudaMalloc(&sink, 4));
    for (int grid : {1, 148})
      for (int threads : {128, 256, 512})
        for (int inflight : {1, 2}) {
          const int iters = 2000;
          for (int rep = 0; rep < 2; ++rep) {
            if (inflight == 1)
              ldtm_kernel<1><<<grid, threads>>>(iters, d_out, sink);
            else
              ldtm_kernel<2><<<grid, threads>>>(iters, d_out, sink);
          }
          CK(cudaDeviceSynchronize());
          Result h[148];
          CK(cudaMemcpy(h, d_out, sizeof(Result) * grid, cudaMemcpyDeviceToHost));
          long long cmax = 0;
          for (int i = 0; i < grid; ++i) cmax = h[i].cycles > cmax ? h[i].cycles : cmax;
          printf("LDTM 32x32b.x32: grid %3d, %2d warps, %d loads in flight per warp: %7.1f cycles per 128 KB tile (%.1f B/cycle/SM)\n", grid,
                 threads / 32, inflight, (double)cmax / iters, 131072.0 * iters / cmax);
        }
  }
  const int tiles = 2000;
  CK(cudaFuncSetAttribute(rate2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
  for (int grid : {2, 148})
    for (int N : {256, 128})
      for (int depth : {2, tiles}) {
        for (int rep = 0; rep < 2; ++rep) rate2_kernel<<<grid, 128, kSmem>>>(N, tiles, depth, 9, d_out);
        CK(cudaDeviceSynchronize());
        Result h[74];
        CK(cudaMemcpy(h, d_out, sizeof(Result) * (grid / 2), cudaMemcpyDeviceToHost));
        long long cmax = 0;
        unsigned long long nmax = 0;
        for (int i = 0; i < grid / 2; ++i) {
          cmax = h[i].cycles > cmax ? h[i].cycles : cmax;
          nmax = h[i].ns > nmax ? h[i].ns : nmax;
        }
        printf("i8 cta_group::2 (M 256) grid %3d N %3d depth %4d: %7.1f cycles per MMA, %6.1f ns per group of 9, clock %.0f MHz\n", grid, N,
               depth, (double)cmax / (tiles * 9.0), (double)nmax / tiles, 1e3 * cmax / (double)nmax);
      }
  for (int grid : {1, 148}) {
    for (int N : {256, 128}) {
      run<0>("i8", grid, N, 0, 0, tiles, 2, 9, d_out);
      run<0>("i8", grid, N, 0, 2, tiles, 2, 9, d_out);
      run<0>("i8", grid, N, 0, 2, tiles, tiles, 9, d_out);
      run<2>("bf16", grid, N, 0, 2, tiles, 2, 9, d_out);
      run<0>("i8", grid, N, 0, 0, tiles, tiles, 9, d_out);
      run<0>("i8", grid, N, 1, 0, tiles, 2, 8, d_out);
      run<0>("i8", grid, N, 0, 1, tiles, 2, 9, d_out);
      run<1>("e4m3", grid, N, 0, 0, tiles, 2, 9, d_out);
      run<2>("bf16", grid, N, 0, 0, tiles, 2, 9, d_out);
      run<2>("bf16", grid, N, 1, 0, tiles, 2, 8, d_out);
    }
  }
  return 0;
}
