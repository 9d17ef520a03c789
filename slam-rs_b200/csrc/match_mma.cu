// S10 on the 5th-generation tensor cores: brute-force NORM_HAMMING k-best as an exact integer contraction.
//
// Serves the same calls as hamming_knn_kernel in match.cu (cv::BFMatcher::knnMatchImpl behind
// /root/reference/src/feature/matcher.rs:34 and src/mapping/map.rs:107) for problems large enough to fill the chip; the
// POPC kernel stays for small ones.  VERDICT r1 item 5 asked for this formulation to be measured; it is bit-exact and
// ~3x faster than the POPC pipe allows (profiles/README.md), so it is the default for batched / large problems.
//
//   d(q, t) = popcount(q ^ t) over 256 bits.  Bits encoded as +-16 int8:  sum_k E[q][k] E[t][k] = 256 (256 - 2 d)
//   = 512 (128 - d).  One more K-slice of 32 int8 carries the train row's index inside its 256-row tile and a
//   validity flag, so the accumulator IS the packed search key:
//        acc = 512 (128 - d) + (255 - (t mod 256))        max acc  <=>  min d, ties -> lowest t  (OpenCV's order)
//   rows t >= nt get -260096 from the flag bytes and never win.  int8 x int8 -> int32 on tcgen05.mma kind::i8 is exact.
//
// expand_desc_kernel   32-byte descriptors -> 320-byte int8 rows in the canonical K-major no-swizzle UMMA layout
//                      (8-row x 16-byte core matrices), one copy for both roles: chunks 0-15 the bits, 16-17 the aux
//                      slice of the QUERY role, 18-19 the aux slice of the TRAIN role.  A tile of any 8-aligned row
//                      range is contiguous in global memory: loaded with ONE cp.async.bulk per tile, no tensor map.
// hamming_mma_kernel   persistent, one CTA per SM, warp-specialised: warp 0 = bulk-copy producer (mbarrier
//                      complete_tx), warp 1 = MMA issuer (one thread; 9 x tcgen05.mma 128x256x32 per tile into one of
//                      two 256-column TMEM accumulators, tcgen05.commit -> mbarriers), warps 2-17 = epilogue: four
//                      groups of four warps drain (buffer, column half) with tcgen05.ld while the next tile's MMA runs.
//                      Without the reverse search the launch runs as clusters of two CTAs that take the two query
//                      tiles of a pair and share every train tile (multicast bulk copy of half a tile each).
//                      Forward best-2 per query row = max-reduction of the keys in registers (2.5 VIMNMX per column);
//                      reverse best-1 per train row (cross-check) = CREDUX.MAX over the 32 rows of a warp, four
//                      quarter results per column combined in shared memory, one RED.MIN.U64 per (unit, column).
#include <limits.h>

#include <algorithm>
#include <cstdlib>

#include "match_mma.cuh"

namespace orbx {

namespace {

constexpr int kKBytes = 288;                 // K of one MMA chain: 256 data + 32 aux int8
constexpr int kChunks = kMmaRowBytes / 16;   // 20 chunks of 16 bytes stored per row
constexpr int kGroupBytes = kChunks * 128;   // 8 rows: [chunk][row & 7][16]
constexpr int kTileM = 128, kTileN = kMmaTileN;
constexpr int kABytes = kTileM / 8 * kGroupBytes;  // 40960
constexpr int kBBytes = kTileN / 8 * kGroupBytes;  // 81920
constexpr int kStages = 2;
constexpr int kGroups = 4;                   // epilogue groups of four warps: TMEM buffer g >> 1, column half g & 1
constexpr int kSmemBytes = kABytes + kStages * kBBytes + kGroups * 4 * 128 * 4 + 128 * 8 * 8 + 256;
constexpr int kThreads = 64 + kGroups * 128;
constexpr int kInvalid = -80000;             // accumulators below this belong to rows >= nt

// ------------------------------------------------------------------------------------------------ expansion
__global__ void __launch_bounds__(256) expand_desc_kernel(const uint8_t* __restrict__ desc, size_t desc_stride,
                                                          const int* __restrict__ n_dev, int n_fixed, int max_rows,
                                                          int rows_padded, uint8_t* __restrict__ out, size_t out_stride) {
  const int set = blockIdx.y;
  const int n = n_dev ? min(n_dev[set], max_rows) : n_fixed;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;  // 16-byte unit of the set, in storage order (coalesced stores)
  if (idx >= rows_padded * kChunks) return;
  // nothing past the last 256-row tile that holds a valid row is ever loaded
  const int grp = idx / (kChunks * 8), rem = idx - grp * (kChunks * 8);
  if (grp * 8 >= ((n + 255) & ~255)) return;
  const int c = rem >> 3, row = grp * 8 + (rem & 7);
  uint4 o;
  if (c < 16) {
    o = make_uint4(0, 0, 0, 0);
    if (row < n) {
      const unsigned bits =
          __ldg(reinterpret_cast<const unsigned short*>(desc + (size_t)set * desc_stride + (size_t)row * 32 + 2 * c));
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
  reinterpret_cast<uint4*>(out + (size_t)set * out_stride)[idx] = o;
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
// bulk asynchronous copy (TMA engine, SASS UBLKCP): completion is signalled on the mbarrier as transferred bytes
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
// shared-memory matrix descriptor: K-major, SWIZZLE_NONE, LBO = 128 B (next 16-byte K chunk), SBO = kGroupBytes (next
// 8-row group), descriptor version 1 (sm_100)
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((128 >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((kGroupBytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  return d;
}
// instruction descriptor of kind::i8: D s32 (c_format 2), A / B signed int8 (format 1), both K-major, N >> 3 at bit 17,
// M >> 4 at bit 24
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
__device__ __forceinline__ void mma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// The same two, called by a whole CONVERGED warp with warp-uniform operands; one elected lane issues.  This is what keeps
// the descriptors in uniform registers: a single thread inside `if (lane == 0)` makes the compiler wrap every UTCIMMA in
// an ELECT / R2UR.BROADCAST / BRA.U.ANY loop, 214 cycles per 128 x 256 x 32 MMA instead of the tensor pipe's 128
// (scripts/exp/mma_rate.cu, profiles/r02_mma_rate.log).
__device__ __forceinline__ void mma_i8_elect(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p, e;\n\t"
      "elect.sync _|e, 0xffffffff;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "@e tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(kIdesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void mma_commit_elect(uint64_t* bar) {
  asm volatile(
      "{\n\t"
      ".reg .pred e;\n\t"
      "elect.sync _|e, 0xffffffff;\n\t"
      "@e tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t"
      "}" ::"r"(smem_u32(bar))
      : "memory");
}
// lane 0's value, known to the compiler as warp-uniform
__device__ __forceinline__ int uniform(int v) { return __shfl_sync(0xFFFFFFFFu, v, 0); }

// ---- CTA pairs that share their B tiles (PAIR instances of the kernel)
__device__ __forceinline__ uint32_t cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// bulk copy global -> the same shared-memory offset of BOTH CTAs of the pair; every destination CTA's barrier at the
// offset of `bar` receives the bytes
__device__ __forceinline__ void bulk_g2s_pair(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile(
      "{\n\t"
      ".reg .b16 m;\n\t"
      "mov.b16 m, 3;\n\t"
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1], %2, [%3], m;\n\t"
      "}" ::"r"(smem_u32(dst)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
// completion of the MMAs issued so far arrives on the barrier at this offset in both CTAs of the pair
__device__ __forceinline__ void mma_commit_pair_elect(uint64_t* bar) {
  asm volatile(
      "{\n\t"
      ".reg .pred e;\n\t"
      ".reg .b16 m;\n\t"
      "mov.b16 m, 3;\n\t"
      "elect.sync _|e, 0xffffffff;\n\t"
      "@e tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], m;\n\t"
      "}" ::"r"(smem_u32(bar))
      : "memory");
}
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
// wait for this thread's TMEM loads; the registers are operands so that nothing that reads them is scheduled earlier
__device__ __forceinline__ void tmem_wait32(int* v) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7]), "+r"(v[8]),
                 "+r"(v[9]), "+r"(v[10]), "+r"(v[11]), "+r"(v[12]), "+r"(v[13]), "+r"(v[14]), "+r"(v[15]), "+r"(v[16]),
                 "+r"(v[17]), "+r"(v[18]), "+r"(v[19]), "+r"(v[20]), "+r"(v[21]), "+r"(v[22]), "+r"(v[23]), "+r"(v[24]),
                 "+r"(v[25]), "+r"(v[26]), "+r"(v[27]), "+r"(v[28]), "+r"(v[29]), "+r"(v[30]), "+r"(v[31])
               :
               : "memory");
}
__device__ __forceinline__ int max3(int a, int b, int c) { return max(max(a, b), c); }

// work unit u -> (problem p, train split sp, query tile mt), mt fastest: the CTAs that run at the same time share B tiles
struct Unit {
  int p, sp, mt, nq, nt, tile0, tile1;  // tiles [tile0, tile1) of the train set
  bool live;
};
// PAIR: a unit is a PAIR of query tiles, CTA `rank` of the cluster takes tile 2 * pair + rank; the unit is live for the
// pair (both CTAs walk the same train tiles, which they share), rows of a query tile past nq are simply inactive
template <bool PAIR>
__device__ __forceinline__ Unit decode_unit(const MmaMatchArgs& a, int u, int rank) {
  Unit w;
  const int per = PAIR ? (a.mtiles + 1) >> 1 : a.mtiles;
  const int m0 = u % per;
  w.mt = PAIR ? 2 * m0 + rank : m0;
  const int r = u / per;
  w.sp = r % a.nsplits;
  w.p = r / a.nsplits;
  w.nq = a.nq_dev ? min(a.nq_dev[w.p], a.max_rows) : a.nq_fixed;
  w.nt = a.nt_dev ? min(a.nt_dev[w.p], a.max_rows_t) : a.nt_fixed;
  const int ntiles = (w.nt + kTileN - 1) / kTileN;
  w.tile0 = min(ntiles, w.sp * a.tiles_per_split);
  w.tile1 = min(ntiles, w.tile0 + a.tiles_per_split);
  w.live = (PAIR ? 2 * m0 : m0) * kTileM < w.nq && w.tile0 < w.tile1;
  return w;
}

// REV: also produce the reverse (train -> query) best-1; a compile-time switch, so that the k-best-only search (map-to-frame)
// carries none of the reverse search's instructions or branches
// PAIR: launched as clusters of two CTAs that take the two query tiles of a pair and SHARE every B tile: each CTA loads
// one half (41 KB) and the bulk copy multicasts it into both CTAs' shared memory, so a 128 x 256 tile costs 41 KB
// instead of 82 KB of L2 -> SM traffic (the limit of the k-best-only search: 148 x 82 KB per 1800-cycle tile is the
// chip's LTS rate).  Everything else stays per CTA - own A tile, own MMAs (cta_group::1), own TMEM hand-off; only the
// B ring is coupled: a stage is refilled when BOTH CTAs have consumed it (their commits arrive on both b_empty).
template <bool REV, bool PAIR>
__global__ void __launch_bounds__(kThreads, 1) hamming_mma_kernel(MmaMatchArgs a) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sA = smem;
  uint8_t* sB = smem + kABytes;
  int* s_col = reinterpret_cast<int*>(smem + kABytes + kStages * kBBytes);                          // [4 groups][4 quarters][128]
  unsigned long long* s_merge = reinterpret_cast<unsigned long long*>(s_col + kGroups * 4 * 128);  // [3][128][2]
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
      mbar_init(b_empty + i, PAIR ? 2 : 1);  // PAIR: this CTA's and the peer's MMA warp
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
  if (PAIR) cluster_sync_all();  // the peer's barriers are initialised before anything arrives on them
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *s_tmem;

  const int rank = PAIR ? (int)cluster_rank() : 0;
  const int cid = PAIR ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int nworkers = PAIR ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  const int nunits = a.npairs * a.nsplits * (PAIR ? (a.mtiles + 1) >> 1 : a.mtiles);
  if (warp == 0) {
    // ===== producer: bulk copies global -> shared, completion on the full barriers
    if (lane == 0) {
      uint32_t ua = 0, ub = 0;  // units / B tiles issued so far
      for (int u = cid; u < nunits; u += nworkers) {
        const Unit w = decode_unit<PAIR>(a, u, rank);
        if (!w.live) continue;
        mbar_wait(a_empty, (ua & 1) ^ 1);
        mbar_expect_tx(a_full, kABytes);
        bulk_g2s(sA, a.ea + (size_t)w.p * a.ea_stride + (size_t)w.mt * kABytes, kABytes, a_full);
        ++ua;
        for (int n = w.tile0; n < w.tile1; ++n, ++ub) {
          const int s = ub & 1;
          mbar_wait(b_empty + s, ((ub >> 1) & 1) ^ 1);
          mbar_expect_tx(b_full + s, kBBytes);  // PAIR: half from this CTA's copy, half from the peer's
          const uint8_t* src = a.eb + (size_t)w.p * a.eb_stride + (size_t)n * kBBytes;
          if (PAIR)
            bulk_g2s_pair(sB + s * kBBytes + rank * (kBBytes / 2), src + rank * (kBBytes / 2), kBBytes / 2, b_full + s);
          else
            bulk_g2s(sB + s * kBBytes, src, kBBytes, b_full + s);
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: the whole warp walks the tile loop converged (every value in it is warp-uniform), one elected
    // lane issues the MMAs and commits
    {
      const uint32_t tm = (uint32_t)uniform((int)tmem);
      uint32_t ua = 0, ub = 0;
      uint32_t tcount0 = 0, tcount1 = 0;  // tiles issued into each TMEM buffer
      for (int u = cid; u < nunits; u += nworkers) {
        const Unit w = decode_unit<PAIR>(a, u, rank);
        const int live = uniform(w.live ? 1 : 0), tile0 = uniform(w.tile0), tile1 = uniform(w.tile1);
        if (!live) continue;
        mbar_wait(a_full, ua & 1);
        ++ua;
        for (int n = tile0; n < tile1; ++n, ++ub) {
          const int s = ub & 1, buf = (n - tile0) & 1;
          const uint32_t tcnt = buf ? tcount1 : tcount0;
          mbar_wait(t_empty + buf, (tcnt & 1) ^ 1);
          if (buf) ++tcount1; else ++tcount0;
          mbar_wait(b_full + s, (ub >> 1) & 1);
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          const uint32_t sa = smem_u32(sA), sb = smem_u32(sB + s * kBBytes);
#pragma unroll
          for (int k = 0; k < kKBytes / 32; ++k)  // last step: aux slices, chunks 16-17 of A against chunks 18-19 of B
            mma_i8_elect(tm + buf * kTileN, make_desc(sa + k * 256), make_desc(sb + k * 256 + (k == 8 ? 256 : 0)), k > 0);
          if (PAIR)
            mma_commit_pair_elect(b_empty + s);  // both CTAs' producers learn that this CTA is done with the stage
          else
            mma_commit_elect(b_empty + s);
          mma_commit_elect(t_full + buf);
        }
        mma_commit_elect(a_empty);
      }
    }
  } else {
    // ===== epilogue
    const int g = (warp - 2) >> 2, q = warp & 3;  // TMEM lanes 32q .. 32q+31 belong to warps with warp % 4 == q
    const int buf = g >> 1, half = g & 1;
    const int row = 32 * q + lane;                // row of the A tile
    int* col = s_col + g * 4 * 128;
    uint32_t tc = 0;  // tiles drained from this group's buffer
    for (int u = cid; u < nunits; u += nworkers) {
      const Unit w = decode_unit<PAIR>(a, u, rank);
      if (!w.live) continue;
      const int qi = w.mt * kTileM + row;
      const bool active = qi < w.nq;
      const int rowbias = active ? 127 - row : -(1 << 30);
      unsigned long long b1 = ~0ull, b2 = ~0ull;  // best-2 of this thread's column share, min order (dist << 32 | train)
      // k-best only (!REV): a tile of a LATER train range can only change (b1, b2) with a distance below that of the
      // K-th best so far (equal distances lose on the index), i.e. with an accumulator >= thr; a 32-column chunk in which
      // no lane of the warp sees one is skipped after a max3 chain (0.5 instead of 2.5 instructions per column).  On a
      // long train stream (map-to-frame) nearly every chunk is.
      int thr = active ? INT_MIN : INT_MAX;
      for (int n = w.tile0 + buf; n < w.tile1; n += 2, ++tc) {
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
          if (ch < 3) tmem_ld32(tbase + 32 * (ch + 1), vn);  // in flight while this chunk is reduced
          bool exact = true;
          if (!REV) {
            int gm = v[0];
#pragma unroll
            for (int j = 1; j < 31; j += 2) gm = max3(gm, v[j], v[j + 1]);
            gm = max(gm, v[31]);
            exact = __any_sync(0xFFFFFFFFu, gm >= thr);
          }
          if (exact) {
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
          }
          if (ch < 3) tmem_wait32(vn);
        }
        // this warp's share of the TMEM buffer is free for the MMA of the tile after next
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncwarp();
        if (lane == 0) mbar_arrive(t_empty + buf);
        // tile-local keys -> global min-order keys
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
        if (!REV && active) {
          const unsigned long long bk = a.K == 2 ? b2 : b1;
          thr = bk == ~0ull ? INT_MIN : 512 * (129 - (int)(bk >> 32));
        }
        if (REV) {
          // column maxima over the 128 rows: four quarter results per column -> one RED.MIN per (unit, column)
          asm volatile("bar.sync %0, 128;" ::"r"(1 + g) : "memory");
          const int c = (warp - 2 - 4 * g) * 32 + lane;
          const int n0 = nt0 + half * 128;
          const int m = max(max(col[c], col[128 + c]), max(col[256 + c], col[384 + c]));
          if (n0 + c < w.nt && m > kInvalid * 128) {
            const unsigned d = (unsigned)(128 - (m >> 16));
            const unsigned r = (unsigned)(127 - (m & 127));
            atomicMin(a.rev + (size_t)w.p * a.rev_rows + n0 + c,
                      ((unsigned long long)d << 32) | (unsigned long long)(w.mt * kTileM + r));
          }
          asm volatile("bar.sync %0, 128;" ::"r"(1 + g) : "memory");
        }
      }
      // merge the four groups' best-2 (groups 1-3 -> shared -> group 0), then write this (problem, split)'s partial
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
          unsigned long long* o = a.part + (((size_t)w.p * a.nsplits + w.sp) * a.max_rows + qi) * a.K;
          o[0] = b1;
          if (a.K == 2) o[1] = b2;
        }
      }
      asm volatile("bar.sync 5, 512;" ::: "memory");
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (PAIR) cluster_sync_all();  // nothing of the peer is in flight towards this CTA's shared memory any more
  if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512));
}

}  // namespace

// CTA pairs pay for the k-best-only search (map-to-frame: 0.950 -> 0.905 ms per 5000 x 1 M frame) when the query tiles pair
// up without much waste - an odd tile count leaves one CTA of the last pair idle; with the reverse search, whose epilogue
// is the bound, the coupled B ring costs 6 % (measured), so those launches stay single CTAs.
bool mma_use_pairs(int mtiles, bool reverse) {
  static const bool no_pairs = getenv("ORBX_MMA_NO_PAIRS") != nullptr;  // A/B switch for measurements
  return !reverse && !no_pairs && mtiles >= 2 && (mtiles % 2 == 0 || mtiles >= 9);
}

size_t mma_expanded_bytes(int rows_padded) { return (size_t)rows_padded / 8 * kGroupBytes; }

int launch_expand_desc(const uint8_t* desc, size_t desc_stride, const int* n_dev, int n_fixed, int max_rows, int nsets,
                       int rows_padded, uint8_t* out, size_t out_stride, cudaStream_t s) {
  if (nsets <= 0) return ORBX_OK;
  dim3 grid(div_up(rows_padded * kChunks, 256), nsets);
  expand_desc_kernel<<<grid, 256, 0, s>>>(desc, desc_stride, n_dev, n_fixed, max_rows, rows_padded, out, out_stride);
  ORBX_LAUNCH_CHECK();
  return ORBX_OK;
}

int launch_hamming_mma(const MmaMatchArgs& a, int num_sms, cudaStream_t s) {
  const bool pair = mma_use_pairs(a.mtiles, a.rev != nullptr) && num_sms >= 2;
  const int nunits = a.npairs * a.nsplits * (pair ? (a.mtiles + 1) / 2 : a.mtiles);
  if (nunits <= 0) return ORBX_OK;
  void (*kern)(MmaMatchArgs) =
      a.rev ? hamming_mma_kernel<true, false> : (pair ? hamming_mma_kernel<false, true> : hamming_mma_kernel<false, false>);
  ORBX_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(pair ? 2 * std::min(num_sms / 2, nunits) : std::min(num_sms, nunits));
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = kSmemBytes;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pair ? 1 : 0;
  ORBX_CUDA(cudaLaunchKernelEx(&cfg, kern, a));
  ORBX_LAUNCH_CHECK();
  return ORBX_OK;
}

}  // namespace orbx
