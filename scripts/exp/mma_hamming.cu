// Experiment: brute-force Hamming k=2 + reverse best-1 as an exact integer contraction on the
// 5th-generation tensor cores (tcgen05.mma kind::i8, accumulators in TMEM), measured against the POPC kernel.
//
//   d(q, t) = popcount(q ^ t) over 256 bits.  With bits encoded as +-a (query) and +-b (train):
//   sum_k A[q][k] * B[t][k] = a*b*(256 - 2 d).   a = 8, b = 16  ->  256 * (128 - d)
//   One more K-slice of 32 int8 carries the column index and a validity flag, so the accumulator IS the packed
//   search key:   acc = 256 * (128 - d) + (255 - (t mod 256))      (max acc <=> min d, ties -> lowest t)
//   invalid train rows (t >= nt) get -260096 from the flag bytes and never win.
//
// Build:  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o mma_hamming mma_hamming.cu
// Run:    ./mma_hamming [npairs] [rows]      (self-checks against a CPU loop, then times)
#include <cuda_runtime.h>
#include <limits.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <vector>

#define CK(x)                                                                              \
  do {                                                                                     \
    cudaError_t e_ = (x);                                                                  \
    if (e_ != cudaSuccess) {                                                               \
      printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__);      \
      exit(1);                                                                             \
    }                                                                                      \
  } while (0)

constexpr int kKBytes = 288;                 // K of one MMA chain: 256 data + 32 aux int8
constexpr int kRowBytes = 320;               // stored per row: 256 data + 32 aux (query role) + 32 aux (train role)
constexpr int kChunks = kRowBytes / 16;      // 20 chunks of 16 bytes
constexpr int kGroupBytes = kChunks * 128;   // 8 rows x 288 bytes, core-matrix major: [chunk][row & 7][16]
constexpr int kTileM = 128, kTileN = 256;
constexpr int kABytes = kTileM / 8 * kGroupBytes;  // 36864
constexpr int kBBytes = kTileN / 8 * kGroupBytes;  // 73728
constexpr int kStages = 2;
constexpr int kGroups = 4;                   // epilogue groups of four warps: TMEM buffer g >> 1, column half g & 1
constexpr int kSmemBytes = kABytes + kStages * kBBytes + kGroups * 4 * 128 * 4 + 128 * 8 * 8 + 256;
constexpr int kThreads = 64 + kGroups * 128; // warp 0 producer, warp 1 MMA, 16 epilogue warps
constexpr int kInvalid = -80000;             // anything below is "no valid column"

// ------------------------------------------------------------------------------------------------ expansion
// E layout per descriptor set: 8-row groups of kGroupBytes; row r, byte k at (r / 8) * kGroupBytes + (k / 16) * 128 +
// (r & 7) * 16 + (k & 15) - the canonical K-major no-swizzle UMMA layout (core matrix = 8 rows x 16 bytes), so a tile
// of any 8-aligned row range is one contiguous block: loaded with one cp.async.bulk, no tensor map.
// chunks 0-15: the 256 bits as +-16; chunks 16-17: aux of the QUERY role (16 x 1, 16 x 127); chunks 18-19: aux of the
// TRAIN role (three bytes summing to 255 - (row & 255), then 16 x (-128) on rows >= n).  One copy serves both roles:
// the descriptor of the last MMA of a chain points at chunks 16-17 for A and at chunks 18-19 for B.
__global__ void expand_kernel(const uint8_t* __restrict__ desc, size_t desc_stride, const int* __restrict__ n_dev,
                              int rows_padded, uint8_t* __restrict__ out, size_t out_stride) {
  const int set = blockIdx.y;
  const int n = n_dev[set];
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;  // 16-byte unit of the set, in storage order
  if (idx >= rows_padded * kChunks) return;
  const int grp = idx / (kChunks * 8), rem = idx - grp * (kChunks * 8);
  const int c = rem >> 3, row = grp * 8 + (rem & 7);
  uint4 o;
  if (c < 16) {
    o = make_uint4(0, 0, 0, 0);
    if (row < n) {
      const unsigned bits = *reinterpret_cast<const unsigned short*>(desc + set * desc_stride + (size_t)row * 32 + 2 * c);
      unsigned w[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const unsigned nib = (bits >> (4 * i)) & 0xFu;
        const unsigned x = (nib * 0x00204081u) & 0x01010101u;  // bit j -> byte j
        w[i] = 0xF0F0F0F0u ^ (x * 0xE0u);                      // -16 / +16
      }
      o = make_uint4(w[0], w[1], w[2], w[3]);
    }
  } else if (c == 16) {
    o = make_uint4(0x01010101u, 0x01010101u, 0x01010101u, 0x01010101u);
  } else if (c == 17) {
    o = make_uint4(0x7F7F7F7Fu, 0x7F7F7F7Fu, 0x7F7F7F7Fu, 0x7F7F7F7Fu);
  } else if (c == 18) {
    const int v = 255 - (row & 255);
    const int b0 = min(v, 127), b1 = min(v - b0, 127), b2 = v - b0 - b1;
    o = make_uint4((unsigned)b0 | ((unsigned)b1 << 8) | ((unsigned)b2 << 16), 0, 0, 0);
  } else {
    const unsigned f = row < n ? 0u : 0x80808080u;  // -128 x 16 on invalid rows
    o = make_uint4(f, f, f, f);
  }
  reinterpret_cast<uint4*>(out + set * out_stride)[idx] = o;
}

// ------------------------------------------------------------------------------------------------ PTX helpers
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
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
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ int g_lbo = 128, g_sbo = kGroupBytes;
__device__ __forceinline__ uint64_t make_desc_rt(uint32_t saddr, int lbo, int sbo) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  return d;
}
// timing experiment: 128-byte swizzled K-major atoms (8 rows x 128 B, SBO = 1024 B)
__device__ __forceinline__ uint64_t make_desc_sw128(uint32_t saddr) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)((1024 >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr) {
  // K-major, SWIZZLE_NONE: LBO = 128 B (next 16-byte K chunk), SBO = kGroupBytes (next 8-row group), version 1 (sm_100)
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((128 >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((kGroupBytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  return d;
}
// kind::i8: D s32 (c_format 2), A / B signed int8 (format 1), both K-major, N >> 3 at bit 17, M >> 4 at bit 24
constexpr uint32_t kIdesc = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(kTileN >> 3) << 17) | ((uint32_t)(kTileM >> 4) << 24);

__device__ __forceinline__ void mma_i8(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(kIdesc), "r"(accumulate)
      : "memory");
}
// rate probes (operands are garbage): the same tile shape issued as kind::f8f6f4 (e4m3 x e4m3 -> f32) and as
// kind::f16 (bf16 x bf16 -> f32, K = 16 elements = the same 32 bytes per row)
constexpr uint32_t kIdescF8 = (1u << 4) | ((uint32_t)(kTileN >> 3) << 17) | ((uint32_t)(kTileM >> 4) << 24);
constexpr uint32_t kIdescBf16 = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(kTileN >> 3) << 17) | ((uint32_t)(kTileM >> 4) << 24);
__device__ __forceinline__ void mma_f8(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(kIdescF8), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void mma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(kIdescBf16), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void mma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// converged-warp issue (see mma_rate.cu): warp-uniform operands, one elected lane issues
__device__ __forceinline__ void mma_i8_elect(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p, e;\n\telect.sync _|e, 0xffffffff;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "@e tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(kIdesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void mma_commit_elect(uint64_t* bar) {
  asm volatile(
      "{\n\t.reg .pred e;\n\telect.sync _|e, 0xffffffff;\n\t"
      "@e tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}" ::"r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ int uniform(int v) { return __shfl_sync(0xFFFFFFFFu, v, 0); }
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, int* v) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ int max3(int a, int b, int c) { return max(max(a, b), c); }

struct MmaArgs {
  const uint8_t* ea;  // expanded query sets  [problem][rows_padded / 8][kGroupBytes]
  const uint8_t* eb;  // expanded train sets
  size_t ea_stride, eb_stride;
  const int* nq_dev;
  const int* nt_dev;
  int max_rows, npairs, mtiles;
  unsigned long long* best;     // [problem][max_rows][2] packed min-order keys (dist << 32 | train), ~0 = none
  unsigned long long* rev;      // [problem][max_rows] reverse best-1 (dist << 32 | query), preset ~0; may be null
  int mode;                     // experiment switches: 2 = skip the best-2 math, 4 = skip the TMEM loads
};

// wait for the TMEM loads of this thread; the registers are listed so that nothing that reads them is scheduled earlier
__device__ __forceinline__ void tmem_wait32(int* v) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7]), "+r"(v[8]),
                 "+r"(v[9]), "+r"(v[10]), "+r"(v[11]), "+r"(v[12]), "+r"(v[13]), "+r"(v[14]), "+r"(v[15]), "+r"(v[16]),
                 "+r"(v[17]), "+r"(v[18]), "+r"(v[19]), "+r"(v[20]), "+r"(v[21]), "+r"(v[22]), "+r"(v[23]), "+r"(v[24]),
                 "+r"(v[25]), "+r"(v[26]), "+r"(v[27]), "+r"(v[28]), "+r"(v[29]), "+r"(v[30]), "+r"(v[31])
               :
               : "memory");
}

// One CTA per SM, persistent over work units (problem, 128-query tile).  Per unit the A tile is loaded once and the
// train set streams through a two-stage B ring; accumulators alternate between two 256-column TMEM buffers.  Sixteen
// epilogue warps in four groups: group g drains columns (g & 1) * 128 .. + 127 of buffer g >> 1, while the MMA of the
// next tile fills the other buffer.
__global__ void __launch_bounds__(kThreads, 1) hamming_mma_kernel(MmaArgs a) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sA = smem;
  uint8_t* sB = smem + kABytes;
  int* s_col = reinterpret_cast<int*>(smem + kABytes + kStages * kBBytes);               // [4 groups][4 quarters][128]
  unsigned long long* s_merge = reinterpret_cast<unsigned long long*>(s_col + kGroups * 4 * 128);  // [4][128][2]
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_merge + 128 * 8);
  uint64_t* a_full = bars + 0;
  uint64_t* a_empty = bars + 1;
  uint64_t* b_full = bars + 2;   // [2]
  uint64_t* b_empty = bars + 4;  // [2]
  uint64_t* t_full = bars + 6;   // [2]
  uint64_t* t_empty = bars + 8;  // [2]
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bars + 10);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    mbar_init(a_full, 1);
    mbar_init(a_empty, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(b_full + i, 1);
      mbar_init(b_empty + i, 1);
      mbar_init(t_full + i, 1);
      mbar_init(t_empty + i, 8);  // one arrival per epilogue warp of the two groups that share the buffer
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(s_tmem)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *s_tmem;

  const int nunits = a.npairs * a.mtiles;
  if (warp == 0) {
    // ===== producer: bulk copies global -> shared, completion on the full barriers
    if (lane == 0 && !(a.mode & 8)) {
      uint32_t ua = 0, ub = 0;  // units / B tiles issued so far
      for (int u = blockIdx.x; u < nunits; u += gridDim.x) {
        const int p = u / a.mtiles, mt = u - p * a.mtiles;
        const int nq = min(a.nq_dev[p], a.max_rows), nt = min(a.nt_dev[p], a.max_rows);
        if (mt * kTileM >= nq || nt <= 0) continue;
        mbar_wait(a_empty, (ua & 1) ^ 1);
        mbar_expect_tx(a_full, kABytes);
        bulk_g2s(sA, a.ea + p * a.ea_stride + (size_t)mt * kABytes, kABytes, a_full);
        ++ua;
        const int ntiles = (nt + kTileN - 1) / kTileN;
        for (int n = 0; n < ntiles; ++n, ++ub) {
          const int s = ub & 1;
          mbar_wait(b_empty + s, ((ub >> 1) & 1) ^ 1);
          mbar_expect_tx(b_full + s, kBBytes);
          bulk_g2s(sB + s * kBBytes, a.eb + p * a.eb_stride + (size_t)n * kBBytes, kBBytes, b_full + s);
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: converged warp + elect (mode & 128), else one thread (the round-2 product until the rate probe)
    if (a.mode & 128) {
      const uint32_t tm = (uint32_t)uniform((int)tmem);
      uint32_t ua = 0, ub = 0, tc0 = 0, tc1 = 0;
      for (int u = blockIdx.x; u < nunits; u += gridDim.x) {
        const int p = u / a.mtiles, mt = u - p * a.mtiles;
        const int nq = uniform(min(a.nq_dev[p], a.max_rows)), nt = uniform(min(a.nt_dev[p], a.max_rows));
        if (mt * kTileM >= nq || nt <= 0) continue;
        if (!(a.mode & 8)) mbar_wait(a_full, ua & 1);
        ++ua;
        const int ntiles = (nt + kTileN - 1) / kTileN;
        for (int n = 0; n < ntiles; ++n, ++ub) {
          const int s = ub & 1, buf = n & 1;
          const uint32_t tcnt = buf ? tc1 : tc0;
          if (!(a.mode & 256)) mbar_wait(t_empty + buf, (tcnt & 1) ^ 1);
          if (buf) ++tc1; else ++tc0;
          if (!(a.mode & 8)) mbar_wait(b_full + s, (ub >> 1) & 1);
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          const uint32_t sa = smem_u32(sA), sb = smem_u32(sB + s * kBBytes);
#pragma unroll
          for (int k = 0; k < kKBytes / 32; ++k)
            mma_i8_elect(tm + buf * kTileN, make_desc(sa + k * 256), make_desc(sb + k * 256 + (k == 8 ? 256 : 0)), k > 0);
          mma_commit_elect(b_empty + s);
          mma_commit_elect(t_full + buf);
        }
        mma_commit_elect(a_empty);
      }
      if ((a.mode & 256) && (a.mode & 8)) {  // free-running probe: nobody waited for the MMAs; a_full is unused without copies
        mma_commit_elect(a_full);
        mbar_wait(a_full, 0);
      }
    } else if (lane == 0) {
      uint32_t ua = 0, ub = 0;
      uint32_t tcount[2] = {0, 0};  // tiles issued into each TMEM buffer
      for (int u = blockIdx.x; u < nunits; u += gridDim.x) {
        const int p = u / a.mtiles, mt = u - p * a.mtiles;
        const int nq = min(a.nq_dev[p], a.max_rows), nt = min(a.nt_dev[p], a.max_rows);
        if (mt * kTileM >= nq || nt <= 0) continue;
        if (!(a.mode & 8)) mbar_wait(a_full, ua & 1);
        ++ua;
        const int ntiles = (nt + kTileN - 1) / kTileN;
        for (int n = 0; n < ntiles; ++n, ++ub) {
          const int s = ub & 1, buf = n & 1;
          mbar_wait(t_empty + buf, (tcount[buf] & 1) ^ 1);
          ++tcount[buf];
          if (!(a.mode & 8)) mbar_wait(b_full + s, (ub >> 1) & 1);
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          const uint32_t sa = smem_u32(sA), sb = smem_u32(sB + s * kBBytes);
#pragma unroll
          for (int k = 0; k < kKBytes / 32; ++k) {  // last step: aux slices, chunks 16-17 of A against chunks 18-19 of B
            if (a.mode & 16) {  // layout timing only (operands are garbage): SW128 atoms, 4 K-steps of 32 B per atom
              if (k < 8)
                mma_i8(tmem + buf * kTileN, make_desc_sw128(sa + (k >> 2) * 16384 + (k & 3) * 32),
                       make_desc_sw128(sb + (k >> 2) * 32768 + (k & 3) * 32), k > 0);
              else
                mma_i8(tmem + buf * kTileN, make_desc(sa), make_desc(sb), 1);
            } else if (a.mode & 32)
              mma_f8(tmem + buf * kTileN, make_desc(sa + k * 256), make_desc(sb + k * 256 + (k == 8 ? 256 : 0)), k > 0);
            else if (a.mode & 64)
              mma_bf16(tmem + buf * kTileN, make_desc(sa + k * 256), make_desc(sb + k * 256 + (k == 8 ? 256 : 0)), k > 0);
            else
              mma_i8(tmem + buf * kTileN, make_desc(sa + k * 256), make_desc(sb + k * 256 + (k == 8 ? 256 : 0)), k > 0);
          }
          mma_commit(b_empty + s);
          mma_commit(t_full + buf);
        }
        mma_commit(a_empty);
      }
    }
  } else {
    // ===== epilogue
    const int g = (warp - 2) >> 2, q = warp & 3;  // TMEM lanes 32q .. 32q+31 belong to warps with warp % 4 == q
    const int buf = g >> 1, half = g & 1;
    const int row = 32 * q + lane;                // row of the A tile
    int* col = s_col + g * 4 * 128;
    uint32_t tc = 0;  // tiles drained from this group's buffer
    for (int u = blockIdx.x; u < nunits && !(a.mode & 256); u += gridDim.x) {
      const int p = u / a.mtiles, mt = u - p * a.mtiles;
      const int nq = min(a.nq_dev[p], a.max_rows), nt = min(a.nt_dev[p], a.max_rows);
      if (mt * kTileM >= nq || nt <= 0) continue;
      const int qi = mt * kTileM + row;
      const bool active = qi < nq;
      const int rowbias = active ? 127 - row : -(1 << 30);
      unsigned long long b1 = ~0ull, b2 = ~0ull;  // running best-2 of this thread's column share, min order (dist << 32 | train)
      const int ntiles = (nt + kTileN - 1) / kTileN;
      for (int n = buf; n < ntiles; n += 2, ++tc) {
        mbar_wait(t_full + buf, tc & 1);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        if (a.mode & 512) {  // handshake only
          asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
          __syncwarp();
          if (lane == 0) mbar_arrive(t_empty + buf);
          continue;
        }
        int m1 = INT_MIN, m2 = INT_MIN;
        const uint32_t tbase = tmem + buf * kTileN + half * 128 + ((uint32_t)(32 * q) << 16);
        int va[32], vb[32];
        if (!(a.mode & 4)) tmem_ld32(tbase, va);
        if (!(a.mode & 1024)) tmem_wait32(va);
#pragma unroll
        for (int ch = 0; ch < 4; ++ch) {
          int* v = (ch & 1) ? vb : va;
          int* vn = (ch & 1) ? va : vb;
          if (ch < 3 && !(a.mode & 4)) tmem_ld32(tbase + 32 * (ch + 1), vn);  // in flight while this chunk is reduced
#pragma unroll
          for (int j = 0; j < 32; j += 2) {
            if (a.mode & 2048) break;
            const int x = v[j], y = v[j + 1];
            if (!(a.mode & 2)) {
              const int hi = max(x, y), lo = min(x, y);
              const int t = min(m1, hi);
              m1 = max(m1, hi);
              m2 = max3(m2, t, lo);
            } else {
              m1 ^= x + y;
            }
            if (a.rev) {
              const int rx = __reduce_max_sync(0xFFFFFFFFu, x * 128 + rowbias);
              const int ry = __reduce_max_sync(0xFFFFFFFFu, y * 128 + rowbias);
              if (lane == 0) *reinterpret_cast<int2*>(col + q * 128 + 32 * ch + j) = make_int2(rx, ry);
            }
          }
          if (ch < 3 && !(a.mode & 1024)) tmem_wait32(vn);
        }
        // this warp's share of the TMEM buffer is free for the MMA of tile n + 2
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncwarp();
        if (lane == 0) mbar_arrive(t_empty + buf);
        // tile-local keys -> global min-order keys
        const int n0 = n * kTileN + half * 128;  // first train row of this group's columns ... but idx is tile-relative:
        const int nt0 = n * kTileN;
        if (m1 > kInvalid) {
          const unsigned long long k1 =
              ((unsigned long long)(unsigned)(128 - (m1 >> 9)) << 32) | (unsigned long long)(nt0 + 255 - (m1 & 511));
          const unsigned long long hi = max(b1, k1);
          b1 = min(b1, k1);
          b2 = min(b2, hi);
        }
        if (m2 > kInvalid) {
          const unsigned long long k2 =
              ((unsigned long long)(unsigned)(128 - (m2 >> 9)) << 32) | (unsigned long long)(nt0 + 255 - (m2 & 511));
          const unsigned long long hi = max(b1, k2);
          b1 = min(b1, k2);
          b2 = min(b2, hi);
        }
        if (a.rev) {
          // column maxima over the 128 rows: four quarter results per column -> one RED.MIN per (unit, column)
          asm volatile("bar.sync %0, 128;" ::"r"(1 + g) : "memory");
          const int c = (warp - 2 - 4 * g) * 32 + lane;
          const int m = max(max(col[c], col[128 + c]), max(col[256 + c], col[384 + c]));
          if (n0 + c < nt && m > kInvalid * 128) {
            const unsigned d = (unsigned)(128 - (m >> 16));
            const unsigned r = (unsigned)(127 - (m & 127));
            atomicMin(a.rev + (size_t)p * a.max_rows + n0 + c,
                      ((unsigned long long)d << 32) | (unsigned long long)(mt * kTileM + r));
          }
          asm volatile("bar.sync %0, 128;" ::"r"(1 + g) : "memory");
        }
      }
      // merge the four groups' running best-2 (groups 1-3 -> shared -> group 0), then write
      if (g > 0) {
        s_merge[((g - 1) * 128 + row) * 2] = b1;
        s_merge[((g - 1) * 128 + row) * 2 + 1] = b2;
      }
      asm volatile("bar.sync 5, 512;" ::: "memory");
      if (g == 0) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          const unsigned long long o1 = s_merge[(k * 128 + row) * 2], o2 = s_merge[(k * 128 + row) * 2 + 1];
          const unsigned long long hi = max(b1, o1);
          b1 = min(b1, o1);
          b2 = min(min(b2, o2), hi);
        }
        if (active) {
          unsigned long long* o = a.best + ((size_t)p * a.max_rows + qi) * 2;
          o[0] = b1;
          o[1] = b2;
        }
      }
      asm volatile("bar.sync 5, 512;" ::: "memory");
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512));
}

// ================================================================================================ v3
// B-RESIDENT order: a unit is (problem, 256-row train tile); the B tile is loaded once and the query tiles stream
// through a three-stage A ring (41 KB per tile-MMA from L2 instead of 82 KB).  The best-2 of every (query tile,
// train tile) goes to a per-train-tile partial (merged by a later kernel), the reverse best-1 as before.
constexpr int kAStages = 3;
constexpr int kSmemV3 = kBBytes + kAStages * kABytes + kGroups * 4 * 128 * 4 + 2 * 128 * 2 * 8 + 256;

struct MmaArgsV3 {
  const uint8_t* ea;
  const uint8_t* eb;
  size_t ea_stride, eb_stride;
  const int* nq_dev;
  const int* nt_dev;
  int max_rows, npairs, ntiles_max;
  unsigned long long* part;  // [problem][ntiles_max][max_rows][2]
  unsigned long long* rev;
};

__global__ void __launch_bounds__(kThreads, 1) hamming_mma_kernel_v3(MmaArgsV3 a) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sB = smem;
  uint8_t* sA = smem + kBBytes;
  int* s_col = reinterpret_cast<int*>(smem + kBBytes + kAStages * kABytes);                        // [4 groups][4][128]
  unsigned long long* s_half = reinterpret_cast<unsigned long long*>(s_col + kGroups * 4 * 128);  // [2 buffers][128][2]
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_half + 2 * 128 * 2);
  uint64_t* b_full = bars + 0;
  uint64_t* b_empty = bars + 1;
  uint64_t* a_full = bars + 2;   // [3]
  uint64_t* a_empty = bars + 5;  // [3]
  uint64_t* t_full = bars + 8;   // [2]
  uint64_t* t_empty = bars + 10; // [2]
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bars + 12);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    mbar_init(b_full, 1);
    mbar_init(b_empty, 1);
    for (int i = 0; i < kAStages; ++i) mbar_init(a_full + i, 1), mbar_init(a_empty + i, 1);
    for (int i = 0; i < 2; ++i) mbar_init(t_full + i, 1), mbar_init(t_empty + i, 8);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(s_tmem)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *s_tmem;

  const int nunits = a.npairs * a.ntiles_max;
  if (warp == 0) {
    if (lane == 0) {
      uint32_t ub = 0, ua = 0;
      for (int u = blockIdx.x; u < nunits; u += gridDim.x) {
        const int p = u / a.ntiles_max, n = u - p * a.ntiles_max;
        const int nq = min(a.nq_dev[p], a.max_rows), nt = min(a.nt_dev[p], a.max_rows);
        if (n * kTileN >= nt || nq <= 0) continue;
        mbar_wait(b_empty, (ub & 1) ^ 1);
        mbar_expect_tx(b_full, kBBytes);
        bulk_g2s(sB, a.eb + p * a.eb_stride + (size_t)n * kBBytes, kBBytes, b_full);
        ++ub;
        const int mtiles = (nq + kTileM - 1) / kTileM;
        for (int mt = 0; mt < mtiles; ++mt, ++ua) {
          const int s = ua % kAStages;
          mbar_wait(a_empty + s, ((ua / kAStages) & 1) ^ 1);
          mbar_expect_tx(a_full + s, kABytes);
          bulk_g2s(sA + s * kABytes, a.ea + p * a.ea_stride + (size_t)mt * kABytes, kABytes, a_full + s);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      uint32_t ub = 0, ua = 0, tcount0 = 0, tcount1 = 0;
      for (int u = blockIdx.x; u < nunits; u += gridDim.x) {
        const int p = u / a.ntiles_max, n = u - p * a.ntiles_max;
        const int nq = min(a.nq_dev[p], a.max_rows), nt = min(a.nt_dev[p], a.max_rows);
        if (n * kTileN >= nt || nq <= 0) continue;
        mbar_wait(b_full, ub & 1);
        ++ub;
        const int mtiles = (nq + kTileM - 1) / kTileM;
        for (int mt = 0; mt < mtiles; ++mt, ++ua) {
          const int s = ua % kAStages, buf = mt & 1;
          const uint32_t tcnt = buf ? tcount1 : tcount0;
          mbar_wait(t_empty + buf, (tcnt & 1) ^ 1);
          if (buf) ++tcount1; else ++tcount0;
          mbar_wait(a_full + s, (ua / kAStages) & 1);
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          const uint32_t sa = smem_u32(sA + s * kABytes), sb = smem_u32(sB);
#pragma unroll
          for (int k = 0; k < kKBytes / 32; ++k)
            mma_i8(tmem + buf * kTileN, make_desc(sa + k * 256), make_desc(sb + k * 256 + (k == 8 ? 256 : 0)), k > 0);
          mma_commit(a_empty + s);
          mma_commit(t_full + buf);
        }
        mma_commit(b_empty);
      }
    }
  } else {
    const int g = (warp - 2) >> 2, q = warp & 3;
    const int buf = g >> 1, half = g & 1;
    const int row = 32 * q + lane;
    int* col = s_col + g * 4 * 128;
    unsigned long long* hx = s_half + buf * 256;
    uint32_t tc = 0;
    for (int u = blockIdx.x; u < nunits; u += gridDim.x) {
      const int p = u / a.ntiles_max, n = u - p * a.ntiles_max;
      const int nq = min(a.nq_dev[p], a.max_rows), nt = min(a.nt_dev[p], a.max_rows);
      if (n * kTileN >= nt || nq <= 0) continue;
      const int mtiles = (nq + kTileM - 1) / kTileM;
      const int nt0 = n * kTileN;
      for (int mt = buf; mt < mtiles; mt += 2, ++tc) {
        const int qi = mt * kTileM + row;
        const bool active = qi < nq;
        const int rowbias = active ? 127 - row : -(1 << 30);
        mbar_wait(t_full + buf, tc & 1);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        int m1 = INT_MIN, m2 = INT_MIN;
        const uint32_t tbase = tmem + buf * kTileN + half * 128 + ((uint32_t)(32 * q) << 16);
        int va[32], vb[32];
        tmem_ld32(tbase, va);
        tmem_wait32(va);
#pragma unroll
        for (int ch = 0; ch < 4; ++ch) {
          int* v = (ch & 1) ? vb : va;
          int* vn = (ch & 1) ? va : vb;
          if (ch < 3) tmem_ld32(tbase + 32 * (ch + 1), vn);
#pragma unroll
          for (int j = 0; j < 32; j += 2) {
            const int x = v[j], y = v[j + 1];
            const int hi = max(x, y), lo = min(x, y);
            const int t = min(m1, hi);
            m1 = max(m1, hi);
            m2 = max3(m2, t, lo);
            if (a.rev) {
              const int rx = __reduce_max_sync(0xFFFFFFFFu, x * 128 + rowbias);
              const int ry = __reduce_max_sync(0xFFFFFFFFu, y * 128 + rowbias);
              if (lane == 0) *reinterpret_cast<int2*>(col + q * 128 + 32 * ch + j) = make_int2(rx, ry);
            }
          }
          if (ch < 3) tmem_wait32(vn);
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncwarp();
        if (lane == 0) mbar_arrive(t_empty + buf);
        // the two column halves of this (query tile, train tile) meet in shared memory; half 0 writes the partial
        unsigned long long k1 = ~0ull, k2 = ~0ull;
        if (m1 > kInvalid) k1 = ((unsigned long long)(unsigned)(128 - (m1 >> 9)) << 32) | (unsigned long long)(nt0 + 255 - (m1 & 511));
        if (m2 > kInvalid) k2 = ((unsigned long long)(unsigned)(128 - (m2 >> 9)) << 32) | (unsigned long long)(nt0 + 255 - (m2 & 511));
        if (half == 1) {
          hx[2 * row] = k1;
          hx[2 * row + 1] = k2;
        }
        asm volatile("bar.sync %0, 256;" ::"r"(6 + buf) : "memory");
        if (half == 0) {
          const unsigned long long o1 = hx[2 * row], o2 = hx[2 * row + 1];
          const unsigned long long hi = max(k1, o1);
          k1 = min(k1, o1);
          k2 = min(min(k2, o2), hi);
          if (active) {
            unsigned long long* o = a.part + (((size_t)p * a.ntiles_max + n) * a.max_rows + qi) * 2;
            *reinterpret_cast<ulonglong2*>(o) = make_ulonglong2(k1, k2);
          }
        }
        if (a.rev) {
          asm volatile("bar.sync %0, 128;" ::"r"(1 + g) : "memory");
          const int c = (warp - 2 - 4 * g) * 32 + lane;
          const int n0 = nt0 + half * 128;
          const int m = max(max(col[c], col[128 + c]), max(col[256 + c], col[384 + c]));
          if (n0 + c < nt && m > kInvalid * 128) {
            const unsigned d = (unsigned)(128 - (m >> 16));
            const unsigned r = (unsigned)(127 - (m & 127));
            atomicMin(a.rev + (size_t)p * a.max_rows + n0 + c,
                      ((unsigned long long)d << 32) | (unsigned long long)(mt * kTileM + r));
          }
        }
        asm volatile("bar.sync %0, 256;" ::"r"(6 + buf) : "memory");  // hx / col are rewritten by the next tile of this buffer
      }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512));
}

// ================================================================================================ v4
// CTA PAIR (cta_group::2): a cluster of two CTAs owns a unit (problem, PAIR of 128-query tiles).  CTA r keeps the A tile
// of query tile 2 * pair + r and loads only rows 128 r .. 128 r + 127 of every 256-row train tile: one tcgen05.mma
// with M = 256 (issued by CTA 0) multiplies both A tiles with both B halves, read from the two CTAs' shared memory, and
// leaves rows of A tile r x all 256 columns in CTA r's TMEM.  41 KB instead of 82 KB from L2 per 128 x 256 tile - the
// one-CTA kernel pulls 148 x 82 KB per ~1800 cycles = the chip's L2 -> SM limit (~6.3 KB per cycle).
//   barriers (same offsets in both CTAs): a_full / b_full[2] filled by the CTA's own bulk copies; CTA 1's warp 1 relays
//   them to CTA 0 (remote arrive on a_peer / b_peer[2]); a_empty / b_empty[2] / t_full[2] arrive in BOTH CTAs by the
//   multicast commit of CTA 0; t_empty[2] lives in CTA 0 and counts the 16 epilogue warps of the two CTAs that share
//   the TMEM buffer.
constexpr int kBHalf = kBBytes / 2;
constexpr int kSmemV4 = kABytes + kStages * kBHalf + kGroups * 4 * 128 * 4 + 128 * 8 * 8 + 256;

__device__ __forceinline__ uint32_t cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// arrive on the barrier at the same offset in CTA `cta` of the cluster
__device__ __forceinline__ void mbar_arrive_remote(uint64_t* bar, uint32_t cta) {
  asm volatile(
      "{\n\t.reg .b32 ra;\n\tmapa.shared::cluster.u32 ra, %0, %1;\n\t"
      "mbarrier.arrive.release.cluster.shared::cluster.b64 _, [ra];\n\t}" ::"r"(smem_u32(bar)),
      "r"(cta)
      : "memory");
}
__device__ __forceinline__ void mbar_wait_cluster(uint64_t* bar, uint32_t parity) {
  uint32_t done;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
  } while (!done);
}
constexpr uint32_t kIdesc2 = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(kTileN >> 3) << 17) | ((uint32_t)(256 >> 4) << 24);
__device__ __forceinline__ void mma2_i8_elect(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p, e;\n\telect.sync _|e, 0xffffffff;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "@e tcgen05.mma.cta_group::2.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(kIdesc2), "r"(accumulate)
      : "memory");
}
// completion of everything issued so far arrives on the barrier at this offset in both CTAs of the pair
__device__ __forceinline__ void mma2_commit_elect(uint64_t* bar) {
  asm volatile(
      "{\n\t.reg .pred e;\n\t.reg .b16 m;\n\tmov.b16 m, 3;\n\telect.sync _|e, 0xffffffff;\n\t"
      "@e tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], m;\n\t}" ::"r"(
          smem_u32(bar))
      : "memory");
}

template <bool REV>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1) hamming_mma_kernel_v4(MmaArgs a) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sA = smem;
  uint8_t* sB = smem + kABytes;
  int* s_col = reinterpret_cast<int*>(smem + kABytes + kStages * kBHalf);
  unsigned long long* s_merge = reinterpret_cast<unsigned long long*>(s_col + kGroups * 4 * 128);
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_merge + 128 * 8);
  uint64_t* a_full = bars + 0;
  uint64_t* a_empty = bars + 1;
  uint64_t* b_full = bars + 2;   // [2]
  uint64_t* b_empty = bars + 4;  // [2]
  uint64_t* t_full = bars + 6;   // [2]
  uint64_t* t_empty = bars + 8;  // [2]  CTA 0's are used
  uint64_t* a_peer = bars + 10;  //      CTA 0's: CTA 1's A tile has landed
  uint64_t* b_peer = bars + 11;  // [2]  CTA 0's: CTA 1's B half has landed
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bars + 13);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_rank();
  if (threadIdx.x == 0) {
    mbar_init(a_full, 1);
    mbar_init(a_empty, 1);
    mbar_init(a_peer, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(b_full + i, 1);
      mbar_init(b_empty + i, 1);
      mbar_init(b_peer + i, 1);
      mbar_init(t_full + i, 1);
      mbar_init(t_empty + i, 16);  // the epilogue warps of both CTAs that drain this buffer
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(s_tmem)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  cluster_sync_all();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *s_tmem;

  const int mpairs = (a.mtiles + 1) / 2;
  const int nunits = a.npairs * mpairs;
  const int cid = blockIdx.x >> 1, nclusters = gridDim.x >> 1;
  // a unit is live for the PAIR (both CTAs walk the same tiles); rows of a query tile past nq are inactive
#define V4_UNIT                                                                                   \
  const int p = u / mpairs, mp = u - p * mpairs;                                                  \
  const int nq = uniform(min(a.nq_dev[p], a.max_rows)), nt = uniform(min(a.nt_dev[p], a.max_rows)); \
  if (2 * mp * kTileM >= nq || nt <= 0) continue;                                                 \
  const int mt = 2 * mp + (int)rank;                                                              \
  const int ntiles = (nt + kTileN - 1) / kTileN;

  if (warp == 0) {
    // ===== producer (each CTA: its A tile, its half of every B tile); the warp stays converged, lane 0 copies
    {
      uint32_t ua = 0, ub = 0;
      for (int u = cid; u < nunits; u += nclusters) {
        V4_UNIT
        mbar_wait(a_empty, (ua & 1) ^ 1);
        if (lane == 0) {
          mbar_expect_tx(a_full, kABytes);
          bulk_g2s(sA, a.ea + p * a.ea_stride + (size_t)mt * kABytes, kABytes, a_full);
        }
        ++ua;
        for (int n = 0; n < ntiles; ++n, ++ub) {
          const int s = ub & 1;
          mbar_wait(b_empty + s, ((ub >> 1) & 1) ^ 1);
          if (lane == 0) {
            mbar_expect_tx(b_full + s, kBHalf);
            bulk_g2s(sB + s * kBHalf, a.eb + p * a.eb_stride + (size_t)n * kBBytes + (size_t)rank * kBHalf, kBHalf, b_full + s);
          }
          __syncwarp();
        }
      }
    }
  } else if (warp == 1 && rank == 1) {
    // ===== relay: tell CTA 0 that this CTA's operands have landed
    uint32_t ua = 0, ub = 0;
    for (int u = cid; u < nunits; u += nclusters) {
      V4_UNIT
      (void)mt;
      mbar_wait(a_full, ua & 1);
      ++ua;
      if (lane == 0) mbar_arrive_remote(a_peer, 0);
      for (int n = 0; n < ntiles; ++n, ++ub) {
        const int s = ub & 1;
        mbar_wait(b_full + s, (ub >> 1) & 1);
        if (lane == 0) mbar_arrive_remote(b_peer + s, 0);
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (CTA 0): converged warp, one elected lane
    const uint32_t tm = (uint32_t)uniform((int)tmem);
    uint32_t ua = 0, ub = 0, tc0 = 0, tc1 = 0;
    for (int u = cid; u < nunits; u += nclusters) {
      V4_UNIT
      (void)mt;
      mbar_wait(a_full, ua & 1);
      mbar_wait_cluster(a_peer, ua & 1);
      ++ua;
      for (int n = 0; n < ntiles; ++n, ++ub) {
        const int s = ub & 1, buf = n & 1;
        const uint32_t tcnt = buf ? tc1 : tc0;
        mbar_wait_cluster(t_empty + buf, (tcnt & 1) ^ 1);
        if (buf) ++tc1; else ++tc0;
        mbar_wait(b_full + s, (ub >> 1) & 1);
        mbar_wait_cluster(b_peer + s, (ub >> 1) & 1);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t sa = smem_u32(sA), sb = smem_u32(sB + s * kBHalf);
#pragma unroll
        for (int k = 0; k < kKBytes / 32; ++k)
          mma2_i8_elect(tm + buf * kTileN, make_desc(sa + k * 256), make_desc(sb + k * 256 + (k == 8 ? 256 : 0)), k > 0);
        mma2_commit_elect(b_empty + s);
        mma2_commit_elect(t_full + buf);
      }
      mma2_commit_elect(a_empty);
    }
  } else {
    // ===== epilogue (both CTAs, each on its own TMEM): as v2, t_empty arrivals go to CTA 0
    const int g = (warp - 2) >> 2, q = warp & 3;
    const int buf = g >> 1, half = g & 1;
    const int row = 32 * q + lane;
    int* col = s_col + g * 4 * 128;
    uint32_t tc = 0;
    for (int u = cid; u < nunits; u += nclusters) {
      V4_UNIT
      const int qi = mt * kTileM + row;
      const bool active = qi < nq;
      const int rowbias = active ? 127 - row : -(1 << 30);
      unsigned long long b1 = ~0ull, b2 = ~0ull;
      for (int n = buf; n < ntiles; n += 2, ++tc) {
        mbar_wait(t_full + buf, tc & 1);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        int m1 = INT_MIN, m2 = INT_MIN;
        const uint32_t tbase = tmem + buf * kTileN + half * 128 + ((uint32_t)(32 * q) << 16);
        int va[32], vb[32];
        tmem_ld32(tbase, va);
        tmem_wait32(va);
#pragma unroll
        for (int ch = 0; ch < 4; ++ch) {
          int* v = (ch & 1) ? vb : va;
          int* vn = (ch & 1) ? va : vb;
          if (ch < 3) tmem_ld32(tbase + 32 * (ch + 1), vn);
#pragma unroll
          for (int j = 0; j < 32; j += 2) {
            const int x = v[j], y = v[j + 1];
            const int hi = max(x, y), lo = min(x, y);
            const int t = min(m1, hi);
            m1 = max(m1, hi);
            m2 = max3(m2, t, lo);
            if (REV) {
              const int rx = __reduce_max_sync(0xFFFFFFFFu, x * 128 + rowbias);
              const int ry = __reduce_max_sync(0xFFFFFFFFu, y * 128 + rowbias);
              if (lane == 0) *reinterpret_cast<int2*>(col + q * 128 + 32 * ch + j) = make_int2(rx, ry);
            }
          }
          if (ch < 3) tmem_wait32(vn);
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncwarp();
        if (lane == 0) mbar_arrive_remote(t_empty + buf, 0);
        const int nt0 = n * kTileN;
        if (m1 > kInvalid) {
          const unsigned long long k1 =
              ((unsigned long long)(unsigned)(128 - (m1 >> 9)) << 32) | (unsigned long long)(nt0 + 255 - (m1 & 511));
          const unsigned long long hi = max(b1, k1);
          b1 = min(b1, k1);
          b2 = min(b2, hi);
        }
        if (m2 > kInvalid) {
          const unsigned long long k2 =
              ((unsigned long long)(unsigned)(128 - (m2 >> 9)) << 32) | (unsigned long long)(nt0 + 255 - (m2 & 511));
          const unsigned long long hi = max(b1, k2);
          b1 = min(b1, k2);
          b2 = min(b2, hi);
        }
        if (REV) {
          asm volatile("bar.sync %0, 128;" ::"r"(1 + g) : "memory");
          const int c = (warp - 2 - 4 * g) * 32 + lane;
          const int n0 = nt0 + half * 128;
          const int m = max(max(col[c], col[128 + c]), max(col[256 + c], col[384 + c]));
          if (n0 + c < nt && m > kInvalid * 128) {
            const unsigned d = (unsigned)(128 - (m >> 16));
            const unsigned r = (unsigned)(127 - (m & 127));
            atomicMin(a.rev + (size_t)p * a.max_rows + n0 + c,
                      ((unsigned long long)d << 32) | (unsigned long long)(mt * kTileM + r));
          }
          asm volatile("bar.sync %0, 128;" ::"r"(1 + g) : "memory");
        }
      }
      if (g > 0) {
        s_merge[((g - 1) * 128 + row) * 2] = b1;
        s_merge[((g - 1) * 128 + row) * 2 + 1] = b2;
      }
      asm volatile("bar.sync 5, 512;" ::: "memory");
      if (g == 0) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          const unsigned long long o1 = s_merge[(k * 128 + row) * 2], o2 = s_merge[(k * 128 + row) * 2 + 1];
          const unsigned long long hi = max(b1, o1);
          b1 = min(b1, o1);
          b2 = min(min(b2, o2), hi);
        }
        if (active) {
          unsigned long long* o = a.best + ((size_t)p * a.max_rows + qi) * 2;
          o[0] = b1;
          o[1] = b2;
        }
      }
      asm volatile("bar.sync 5, 512;" ::: "memory");
    }
  }
#undef V4_UNIT
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  cluster_sync_all();
  if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512));
}

// ------------------------------------------------------------------------------------------------ host
static unsigned popc256(const uint8_t* a, const uint8_t* b) {
  unsigned d = 0;
  for (int i = 0; i < 32; ++i) d += __builtin_popcount((unsigned)(a[i] ^ b[i]));
  return d;
}

int main(int argc, char** argv) {
  const int npairs = argc > 1 ? atoi(argv[1]) : 999;
  const int rows = argc > 2 ? atoi(argv[2]) : 2000;
  const int cap = (rows + rows / 8 + 64 + 31) / 32 * 32;
  const int rows_padded = (cap + 255) / 256 * 256;
  const int nsets = npairs + 1;
  printf("npairs %d rows %d cap %d rows_padded %d smem %d\n", npairs, rows, cap, rows_padded, kSmemBytes);
  // descriptor sets: random, set s+1 holds noisy copies of half of set s (real matches), and a few exact duplicates
  std::vector<uint8_t> desc((size_t)nsets * cap * 32);
  std::vector<int> n(nsets);
  srand(1);
  for (int s = 0; s < nsets; ++s) {
    n[s] = rows - (s % 7) * 13;
    if (s == 3) n[s] = 257;
    if (s == 5) n[s] = 1;
    for (size_t i = 0; i < (size_t)cap * 32; ++i) desc[(size_t)s * cap * 32 + i] = (uint8_t)(rand() >> 7);
    if (s > 0)
      for (int r = 0; r < n[s] / 2 && r < n[s - 1]; ++r) {
        uint8_t* d = &desc[((size_t)s * cap + r) * 32];
        memcpy(d, &desc[((size_t)(s - 1) * cap + (r * 7) % n[s - 1]) * 32], 32);
        if (r % 5) d[rand() % 32] ^= (uint8_t)(1 << (rand() % 8));  // r % 5 == 0: exact duplicate (distance 0)
        if (r % 11 == 0 && r + 1 < n[s]) memcpy(d + 32, d, 32);      // tie between two train rows
      }
  }
  uint8_t *d_desc, *d_ea;
  int* d_n;
  unsigned long long* d_best;
  unsigned long long* d_rev;
  const size_t estride = (size_t)rows_padded / 8 * kGroupBytes;
  CK(cudaMalloc(&d_desc, desc.size()));
  CK(cudaMalloc(&d_ea, estride * nsets + kBBytes));
  CK(cudaMalloc(&d_n, nsets * sizeof(int)));
  CK(cudaMalloc(&d_best, (size_t)npairs * cap * 2 * sizeof(unsigned long long)));
  CK(cudaMalloc(&d_rev, (size_t)npairs * cap * sizeof(unsigned long long)));
  CK(cudaMemcpy(d_desc, desc.data(), desc.size(), cudaMemcpyHostToDevice));
  CK(cudaMemcpy(d_n, n.data(), nsets * sizeof(int), cudaMemcpyHostToDevice));
  CK(cudaMemset(d_best, 0xFF, (size_t)npairs * cap * 2 * sizeof(unsigned long long)));
  CK(cudaFuncSetAttribute(hamming_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
  int sms = 0;
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));

  MmaArgs a{};
  a.ea = d_ea;                      // pair p: query = set p, train = set p + 1
  a.eb = d_ea + estride;
  a.ea_stride = a.eb_stride = estride;
  a.nq_dev = d_n;
  a.nt_dev = d_n + 1;
  a.max_rows = cap;
  a.npairs = npairs;
  a.mtiles = (cap + kTileM - 1) / kTileM;
  a.best = d_best;
  a.rev = d_rev;
  cudaEvent_t e0, e1, e2;
  cudaEventCreate(&e0), cudaEventCreate(&e1), cudaEventCreate(&e2);
  for (int rep = 0; rep < 4; ++rep) {
    CK(cudaMemset(d_rev, 0xFF, (size_t)npairs * cap * sizeof(unsigned long long)));
    cudaEventRecord(e0);
    dim3 ge((rows_padded * kChunks + 255) / 256, nsets);
    expand_kernel<<<ge, 256>>>(d_desc, (size_t)cap * 32, d_n, rows_padded, d_ea, estride);
    cudaEventRecord(e1);
    hamming_mma_kernel<<<sms, kThreads, kSmemBytes>>>(a);
    cudaEventRecord(e2);
    CK(cudaDeviceSynchronize());
    float ms_e, ms_k;
    cudaEventElapsedTime(&ms_e, e0, e1);
    cudaEventElapsedTime(&ms_k, e1, e2);
    double pairs = 0;
    for (int p = 0; p < npairs; ++p) pairs += (double)n[p] * n[p + 1];
    printf("rep %d: expand %.3f ms, mma kernel %.3f ms, %.1f G pairs/s (kernel), %.1f G pairs/s (both)\n", rep, ms_e, ms_k,
           pairs / ms_k * 1e-6, pairs / (ms_k + ms_e) * 1e-6);
  }
  // check a few pairs against the CPU loop
  std::vector<unsigned long long> best((size_t)npairs * cap * 2);
  std::vector<unsigned long long> rev((size_t)npairs * cap);
  CK(cudaMemcpy(best.data(), d_best, best.size() * 8, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(rev.data(), d_rev, rev.size() * 8, cudaMemcpyDeviceToHost));
  long bad = 0, checked = 0;
  const int check_pairs[] = {0, 1, 2, 3, 4, 5, npairs / 2, npairs - 1};
  for (int p : check_pairs) {
    if (p < 0 || p >= npairs) continue;
    const int nq = n[p], nt = n[p + 1];
    std::vector<unsigned long long> r(nt, ~0ull);
    for (int i = 0; i < nq; ++i) {
      unsigned long long b1 = ~0ull, b2 = ~0ull;
      for (int j = 0; j < nt; ++j) {
        const unsigned d = popc256(&desc[((size_t)p * cap + i) * 32], &desc[((size_t)(p + 1) * cap + j) * 32]);
        const unsigned long long key = ((unsigned long long)d << 32) | (unsigned long long)j;
        const unsigned long long hi = std::max(b1, key);
        b1 = std::min(b1, key);
        b2 = std::min(b2, hi);
        const unsigned long long rk = ((unsigned long long)d << 32) | (unsigned long long)i;
        r[j] = std::min(r[j], rk);
      }
      const unsigned long long* o = &best[((size_t)p * cap + i) * 2];
      ++checked;
      if (o[0] != b1 || o[1] != b2) {
        if (bad < 10) printf("pair %d query %d: got %016llx %016llx want %016llx %016llx\n", p, i, o[0], o[1], b1, b2);
        ++bad;
      }
    }
    for (int j = 0; j < nt; ++j) {
      ++checked;
      if (rev[(size_t)p * cap + j] != r[j]) {
        if (bad < 20) printf("pair %d train %d: rev got %016llx want %016llx\n", p, j, rev[(size_t)p * cap + j], r[j]);
        ++bad;
      }
    }
  }
  printf("checked %ld results, %ld mismatches -> %s\n", checked, bad, bad ? "FAIL" : "OK (bit-exact)");
  // where the time goes: the same launch with parts of the epilogue switched off
  const struct { const char* name; bool rev; int mode; } exps[] = {
      {"full", true, 0}, {"no reverse (k=2 only)", false, 0}, {"no reverse, no best-2 math", false, 2},
      {"no reverse, no math, no TMEM loads (MMA + copies only)", false, 6}, {"MMA only: no copies either", false, 14}, {"MMA only, SW128 descriptors (layout timing, garbage data)", false, 30}, {"reverse only (no best-2 math)", true, 2},
      {"converged-warp issue: full", true, 128}, {"converged-warp issue: no reverse (k=2 only)", false, 128},
      {"converged-warp issue: no reverse, no best-2 math", false, 128 | 2},
      {"converged-warp issue: no reverse, no math, no TMEM loads", false, 128 | 6},
      {"converged-warp issue: MMA only (no copies)", false, 128 | 14},
      {"converged-warp issue: reverse only (no best-2 math)", true, 128 | 2},
      {"converged-warp issue: MMA only, no tcgen05.wait::ld", false, 128 | 14 | 1024},
      {"converged-warp issue: MMA only, no inner loop", false, 128 | 14 | 2048},
      {"converged-warp issue: MMA only, no wait::ld, no inner loop", false, 128 | 14 | 1024 | 2048},
      {"converged-warp issue: MMA only, epilogue = wait + arrive only", false, 128 | 14 | 512},
      {"converged-warp issue: MMA only, free running (no epilogue warps, no TMEM handshake)", false, 128 | 14 | 256},
      {"MMA only, issued as kind::f8f6f4 (rate probe, garbage data)", false, 14 | 32},
      {"MMA only, issued as kind::f16 bf16 (rate probe, garbage data)", false, 14 | 64}};
  for (auto& e : exps) {
    MmaArgs b = a;
    b.rev = e.rev ? d_rev : nullptr;
    b.mode = e.mode;
    hamming_mma_kernel<<<sms, kThreads, kSmemBytes>>>(b);
    cudaEventRecord(e1);
    hamming_mma_kernel<<<sms, kThreads, kSmemBytes>>>(b);
    cudaEventRecord(e2);
    CK(cudaDeviceSynchronize());
    float ms;
    cudaEventElapsedTime(&ms, e1, e2);
    printf("  %-58s %.3f ms\n", e.name, ms);
  }
  // ---- v3: B-resident order, per-train-tile partials
  {
    const int ntiles_max = rows_padded / kTileN;
    unsigned long long* d_part;
    CK(cudaMalloc(&d_part, (size_t)npairs * ntiles_max * cap * 2 * 8));
    CK(cudaFuncSetAttribute(hamming_mma_kernel_v3, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemV3));
    MmaArgsV3 b{};
    b.ea = d_ea, b.eb = d_ea + estride, b.ea_stride = b.eb_stride = estride;
    b.nq_dev = d_n, b.nt_dev = d_n + 1, b.max_rows = cap, b.npairs = npairs, b.ntiles_max = ntiles_max;
    b.part = d_part;
    for (int withrev = 1; withrev >= 0; --withrev) {
      b.rev = withrev ? d_rev : nullptr;
      for (int rep = 0; rep < 3; ++rep) {
        CK(cudaMemset(d_part, 0xFF, (size_t)npairs * ntiles_max * cap * 2 * 8));
        CK(cudaMemset(d_rev, 0xFF, (size_t)npairs * cap * sizeof(unsigned long long)));
        cudaEventRecord(e1);
        hamming_mma_kernel_v3<<<sms, kThreads, kSmemV3>>>(b);
        cudaEventRecord(e2);
        CK(cudaDeviceSynchronize());
        float ms;
        cudaEventElapsedTime(&ms, e1, e2);
        printf("v3 (B resident, smem %d)%s rep %d: %.3f ms\n", kSmemV3, withrev ? "" : " no reverse", rep, ms);
      }
      if (!withrev) break;
      std::vector<unsigned long long> part((size_t)npairs * ntiles_max * cap * 2), rev3((size_t)npairs * cap);
      CK(cudaMemcpy(part.data(), d_part, part.size() * 8, cudaMemcpyDeviceToHost));
      CK(cudaMemcpy(rev3.data(), d_rev, rev3.size() * 8, cudaMemcpyDeviceToHost));
      long bad3 = 0, chk3 = 0;
      for (int p : check_pairs) {
        if (p < 0 || p >= npairs) continue;
        const int nq = n[p], nt = n[p + 1];
        for (int i = 0; i < nq; ++i) {
          unsigned long long b1 = ~0ull, b2 = ~0ull;
          for (int t = 0; t < (nt + kTileN - 1) / kTileN; ++t)
            for (int k = 0; k < 2; ++k) {
              const unsigned long long key = part[(((size_t)p * ntiles_max + t) * cap + i) * 2 + k];
              const unsigned long long hi = std::max(b1, key);
              b1 = std::min(b1, key);
              b2 = std::min(b2, hi);
            }
          const unsigned long long* o = &best[((size_t)p * cap + i) * 2];  // v2 results, already checked against the CPU
          ++chk3;
          if (o[0] != b1 || o[1] != b2) {
            if (bad3 < 10) printf("v3 pair %d query %d: got %016llx %016llx want %016llx %016llx\n", p, i, b1, b2, o[0], o[1]);
            ++bad3;
          }
        }
        for (int j = 0; j < nt; ++j) {
          ++chk3;
          if (rev3[(size_t)p * cap + j] != rev[(size_t)p * cap + j]) ++bad3;
        }
      }
      printf("v3 checked %ld results, %ld mismatches -> %s\n", chk3, bad3, bad3 ? "FAIL" : "OK (bit-exact)");
    }
  }
  // ---- v4: CTA pairs (cta_group::2), each CTA loads half of every B tile
  if (!getenv("SKIP_V4")) {
    CK(cudaFuncSetAttribute(hamming_mma_kernel_v4<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemV4));
    CK(cudaFuncSetAttribute(hamming_mma_kernel_v4<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemV4));
    const int grid = sms & ~1;
    for (int withrev = 1; withrev >= 0; --withrev) {
      MmaArgs b = a;
      b.mode = 0;
      b.rev = withrev ? d_rev : nullptr;
      for (int rep = 0; rep < 3; ++rep) {
        CK(cudaMemset(d_best, 0xFF, (size_t)npairs * cap * 2 * sizeof(unsigned long long)));
        CK(cudaMemset(d_rev, 0xFF, (size_t)npairs * cap * sizeof(unsigned long long)));
        cudaEventRecord(e1);
        if (withrev)
          hamming_mma_kernel_v4<true><<<grid, kThreads, kSmemV4>>>(b);
        else
          hamming_mma_kernel_v4<false><<<grid, kThreads, kSmemV4>>>(b);
        cudaEventRecord(e2);
        CK(cudaDeviceSynchronize());
        float ms;
        cudaEventElapsedTime(&ms, e1, e2);
        printf("v4 (CTA pair, smem %d)%s rep %d: %.3f ms\n", kSmemV4, withrev ? "" : " no reverse", rep, ms);
      }
      std::vector<unsigned long long> best4(best.size()), rev4(rev.size());
      CK(cudaMemcpy(best4.data(), d_best, best4.size() * 8, cudaMemcpyDeviceToHost));
      CK(cudaMemcpy(rev4.data(), d_rev, rev4.size() * 8, cudaMemcpyDeviceToHost));
      long bad4 = 0;
      for (size_t i = 0; i < best.size(); ++i)
        if (best4[i] != best[i]) {
          if (bad4 < 8) printf("v4 best[%zu] (pair %zu row %zu k %zu): got %016llx want %016llx\n", i, i / (2 * cap), (i / 2) % cap, i & 1, best4[i], best[i]);
          ++bad4;
        }
      if (withrev)
        for (size_t i = 0; i < rev.size(); ++i)
          if (rev4[i] != rev[i]) {
            if (bad4 < 16) printf("v4 rev[%zu]: got %016llx want %016llx\n", i, rev4[i], rev[i]);
            ++bad4;
          }
      printf("v4 compared %zu + %zu results with v2, %ld mismatches -> %s\n", best.size(), withrev ? rev.size() : (size_t)0, bad4,
             bad4 ? "FAIL" : "OK (bit-exact)");
    }
  }
  return bad ? 1 : 0;
}
