// S10 — brute-force NORM_HAMMING matcher (k=1 / k=2, cross-check, Lowe ratio, min-distance filter).
//
// Replaces cv::BFMatcher::knnMatchImpl -> cv::batchDistance reached from
// /root/reference/src/feature/matcher.rs:34 (train_match) and the pure-Rust post filter at
// matcher.rs:40-66.  Semantics: dist = popcount(q XOR t) over 256 bits;
// K best in ascending (distance, trainIdx) order, i.e. ties go to the lowest train index.
//
// Kernel shape: one thread owns one query descriptor (8 x u32 in registers); the train set is streamed
// through shared memory in 256-row tiles (cp.async double buffer) and every lane reads the same train
// row (LDS.128 broadcast, conflict free).  Per (query, train) pair: 8 LOP3(xor), then a carry-save
// adder tree (6 LOP3: three 3:2 compressors) folds the 8 words into 2 "ones" + 3 "twos" words so only
// 5 POPC are issued instead of 8 (POPC is the quarter-rate pipe: 16 lanes/clk/SM measured, LOP3 runs at
// 64), and a packed-key min/max update of the running best-2: key = dist << 16 | local_train_idx, so
// the lexicographic (distance, index) order is one unsigned compare and the lowest index wins ties.
// With cross-check the reverse (train -> query) best-1 comes out of the SAME distance evaluation: a
// warp-wide CREDUX.MIN of (dist << 16 | query lane) per train row (32 lanes/clk/SM, its own pipe),
// per-warp column minima in shared memory, one RED.MIN.U64 per (block, train row) to global memory.
// Integer work only; tensor cores are deliberately not used.
#include <cooperative_groups.h>
#include <cuda_pipeline.h>
#include <string.h>

#include <algorithm>
#include <cstdlib>
#include <mutex>
#include <vector>

#include "common.cuh"
#include "match_mma.cuh"

namespace orbx {

constexpr int kQPB = 128;        // queries (= threads) per block
constexpr int kTile = 256;       // train rows per shared-memory tile (8 KB)
constexpr int kMaxSplitRows = 65536;  // local train index must fit 16 bits

}  // namespace orbx

struct orbx_matcher {
  int device = 0;
  int cross_check = 0;
  int num_sms = 148;
  cudaStream_t own_stream = nullptr, stream = nullptr;
  // staging for host-pointer calls and partial results; grown on demand
  void* buf[10] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  size_t cap[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
  int mode = 0;  // 0 = auto (tensor-core kernel for large problems), 1 = POPC kernel only, 2 = tensor-core kernel always
  // resident train set whose expanded form in buf[B_EXPT] is current (orbx_comm_set_train / matcher_cache_train)
  const uint8_t* cached_t = nullptr;
  int cached_nt = 0;
  orbx::EventTimer timer;
};

namespace orbx {

enum { B_Q = 0, B_T, B_OUT, B_CNT, B_PART, B_REV, B_OUT2, B_CNT2, B_EXPQ, B_EXPT };

static int ensure(orbx_matcher* m, int which, size_t bytes) {
  if (bytes <= m->cap[which]) return ORBX_OK;
  if (m->buf[which]) {
    ORBX_CUDA(cudaStreamSynchronize(m->stream));
    ORBX_CUDA(cudaFree(m->buf[which]));
    m->buf[which] = nullptr;
    m->cap[which] = 0;
  }
  size_t want = align_up(bytes + bytes / 4, 256);
  cudaError_t e = cudaMalloc(&m->buf[which], want);
  if (e != cudaSuccess) {
    set_error("cudaMalloc(%zu) failed: %s", want, cudaGetErrorString(e));
    return ORBX_E_NOMEM;
  }
  m->cap[which] = want;
  return ORBX_OK;
}

// ------------------------------------------------------------------------------------------------
// Hamming k-best kernel
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
  unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;\n" ::"n"(N));
}

struct MatchArgs {
  const uint8_t* q;
  const uint8_t* t;
  size_t q_stride, t_stride;  // bytes between problems
  const int* nq_dev;          // per problem counts (device) or nullptr
  const int* nt_dev;
  int nq_fixed, nt_fixed;  // used when the pointers are null
  int max_rows;            // clamp for device counts; output row pitch
  int rows_per_split;      // multiple of kTile, <= kMaxSplitRows
  int nsplits;
  unsigned long long* part;  // [problem][split][max_rows][K] packed (dist<<32 | train_idx)
  orbx_dmatch* out;          // used when nsplits == 1: [problem][max_rows][K]
  int* n_per_q;              // optional [problem][max_rows] (K=2) ; for K=1: unused
  int* n_out;                // optional per problem count (K=1 batch): nq or 0
  unsigned long long* rev;   // FUSED only: [problem][max_rows_t] reverse best-1 (dist << 32 | query_idx), preset ~0
  int rev_rows;              // row pitch of rev (max train rows per problem)
  int q_index_base;          // added to query_idx in the records (query-sharded matching: this launch owns a slice)
};

__device__ __forceinline__ unsigned lop3_xor3(unsigned a, unsigned b, unsigned c) {
  unsigned r;
  asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
  return r;
}
__device__ __forceinline__ unsigned lop3_maj3(unsigned a, unsigned b, unsigned c) {
  unsigned r;
  asm("lop3.b32 %0, %1, %2, %3, 0xE8;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
  return r;
}

// popcount of the 256-bit XOR with 5 POPC: three carry-save 3:2 compressors (sum = a^b^c, carry = maj)
// leave ones = {s3, x7} and twos = {c1, c2, c3};  dist = pop(s3) + pop(x7) + 2 (pop(c1)+pop(c2)+pop(c3)).
__device__ __forceinline__ unsigned hamming256(const uint4& qa, const uint4& qb, const uint4& x, const uint4& y) {
  const unsigned x0 = qa.x ^ x.x, x1 = qa.y ^ x.y, x2 = qa.z ^ x.z, x3 = qa.w ^ x.w;
  const unsigned x4 = qb.x ^ y.x, x5 = qb.y ^ y.y, x6 = qb.z ^ y.z, x7 = qb.w ^ y.w;
  const unsigned s1 = lop3_xor3(x0, x1, x2), c1 = lop3_maj3(x0, x1, x2);
  const unsigned s2 = lop3_xor3(x3, x4, x5), c2 = lop3_maj3(x3, x4, x5);
  const unsigned s3 = lop3_xor3(s1, s2, x6), c3 = lop3_maj3(s1, s2, x6);
  return (__popc(s3) + __popc(x7)) + 2 * (__popc(c1) + __popc(c2) + __popc(c3));
}

template <int K>
__device__ __forceinline__ void best_update(unsigned& b1, unsigned& b2, unsigned key) {
  if (K == 1) {
    b1 = min(b1, key);
  } else {
    unsigned hi = max(b1, key);
    b1 = min(b1, key);
    b2 = min(b2, hi);
  }
}

// two keys at once (keys are unique: they carry the train row): 5 min/max instead of 6
__device__ __forceinline__ void best_update2(unsigned& b1, unsigned& b2, unsigned k0, unsigned k1) {
  const unsigned m = min(k0, k1), M = max(k0, k1);
  const unsigned t = max(b1, m);
  b1 = min(b1, m);
  b2 = __vimin3_u32(t, b2, M);
}

template <int K, bool FUSED>
__global__ void __launch_bounds__(kQPB) hamming_knn_kernel(MatchArgs a) {
  __shared__ uint4 tile[2][kTile * 2];
  __shared__ unsigned s_col[FUSED ? kQPB / 32 : 1][FUSED ? kTile : 1];
  const int prob = blockIdx.z;
  const int split = blockIdx.y;
  int nq = a.nq_dev ? min(a.nq_dev[prob], a.max_rows) : a.nq_fixed;
  int nt = a.nt_dev ? min(a.nt_dev[prob], a.max_rows) : a.nt_fixed;
  const int q0 = blockIdx.x * kQPB;
  if (q0 >= nq) return;
  const int qi = q0 + threadIdx.x;
  const bool active = qi < nq;

  const int t_begin = split * a.rows_per_split;
  const int t_end = min(nt, t_begin + a.rows_per_split);

  const uint4* qp = reinterpret_cast<const uint4*>(a.q + (size_t)prob * a.q_stride);
  const uint4* tp = reinterpret_cast<const uint4*>(a.t + (size_t)prob * a.t_stride);

  uint4 qa = make_uint4(0, 0, 0, 0), qb = qa;
  if (active) {
    qa = __ldg(qp + (size_t)qi * 2);
    qb = __ldg(qp + (size_t)qi * 2 + 1);
  }
  unsigned b1 = 0xFFFFFFFFu, b2 = 0xFFFFFFFFu;
  // reverse key of this thread inside its block: dist << 16 | lane-in-block; idle lanes never win
  const unsigned qoff = active ? threadIdx.x : 0x7F000000u;
  const int warp = threadIdx.x >> 5;
  const bool lane0 = (threadIdx.x & 31) == 0;

  const int ntiles = t_end > t_begin ? (t_end - t_begin + kTile - 1) / kTile : 0;
  auto issue = [&](int tl, int bufi) {
    const int base = t_begin + tl * kTile;
    const int rows = min(kTile, t_end - base);
    const uint4* src = tp + (size_t)base * 2;
    for (int i = threadIdx.x; i < rows * 2; i += kQPB) cp_async16(&tile[bufi][i], src + i);
    cp_async_commit();
  };
  if (ntiles > 0) issue(0, 0);
  for (int tl = 0; tl < ntiles; ++tl) {
    const int cur = tl & 1;
    if (tl + 1 < ntiles) {
      issue(tl + 1, cur ^ 1);
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();
    const int base_local = tl * kTile;
    const int rows = min(kTile, t_end - (t_begin + base_local));
    const uint4* tt = tile[cur];
    int j = 0;
#pragma unroll 1
    for (; j + 4 <= rows; j += 4) {
      unsigned key[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const uint4 x = tt[(j + u) * 2], y = tt[(j + u) * 2 + 1];
        const unsigned d = hamming256(qa, qb, x, y);
        key[u] = (d << 16) + (unsigned)(base_local + j + u);
        if (K == 1) best_update<1>(b1, b2, key[u]);
        if (FUSED) {
          const unsigned r = __reduce_min_sync(0xFFFFFFFFu, (d << 16) + qoff);
          if (lane0) s_col[warp][j + u] = r;
        }
      }
      if (K == 2) {
        best_update2(b1, b2, key[0], key[1]);
        best_update2(b1, b2, key[2], key[3]);
      }
    }
    for (; j < rows; ++j) {
      const uint4 x = tt[j * 2], y = tt[j * 2 + 1];
      const unsigned d = hamming256(qa, qb, x, y);
      best_update<K>(b1, b2, (d << 16) + (unsigned)(base_local + j));
      if (FUSED) {
        const unsigned r = __reduce_min_sync(0xFFFFFFFFu, (d << 16) + qoff);
        if (lane0) s_col[warp][j] = r;
      }
    }
    __syncthreads();
    if (FUSED) {
      // column minima of this block -> global reverse best-1 (lowest query index wins ties)
      unsigned long long* rv = a.rev + (size_t)prob * a.rev_rows + (t_begin + base_local);
      for (int c = threadIdx.x; c < rows; c += kQPB) {
        const unsigned m = min(min(s_col[0][c], s_col[1][c]), min(s_col[2][c], s_col[3][c]));
        atomicMin(rv + c, ((unsigned long long)(m >> 16) << 32) | (unsigned long long)(q0 + (int)(m & 0xFFFFu)));
      }
      // s_col is rewritten only after the __syncthreads at the top of the next tile
    }
  }
  if (!active) return;

  if (a.nsplits == 1) {
    // fused epilogue: DMatch records straight from registers
    orbx_dmatch* o = a.out + ((size_t)prob * a.max_rows + qi) * K;
    int cnt = 0;
    if (b1 != 0xFFFFFFFFu) {
      o[0] = orbx_dmatch{qi + a.q_index_base, (int)(b1 & 0xFFFF), 0, (float)(b1 >> 16)};
      cnt = 1;
    }
    if (K == 2 && b2 != 0xFFFFFFFFu) {
      o[1] = orbx_dmatch{qi + a.q_index_base, (int)(b2 & 0xFFFF), 0, (float)(b2 >> 16)};
      cnt = 2;
    }
    if (a.n_per_q) a.n_per_q[(size_t)prob * a.max_rows + qi] = cnt;
    if (a.n_out && qi == 0) a.n_out[prob] = (nt > 0) ? nq : 0;
  } else {
    unsigned long long* p = a.part + (((size_t)prob * a.nsplits + split) * a.max_rows + qi) * K;
    auto widen = [&](unsigned k) -> unsigned long long {
      if (k == 0xFFFFFFFFu) return ~0ull;
      return ((unsigned long long)(k >> 16) << 32) | (unsigned long long)(t_begin + (int)(k & 0xFFFF));
    };
    p[0] = widen(b1);
    if (K == 2) p[1] = widen(b2);
  }
}

// merge per-split partial best-K into DMatch records
template <int K>
__global__ void merge_splits_kernel(MatchArgs a) {
  const int prob = blockIdx.y;
  int nq = a.nq_dev ? min(a.nq_dev[prob], a.max_rows) : a.nq_fixed;
  int nt = a.nt_dev ? min(a.nt_dev[prob], a.max_rows) : a.nt_fixed;
  const int qi = blockIdx.x * blockDim.x + threadIdx.x;
  if (qi == 0 && a.n_out) a.n_out[prob] = (nt > 0) ? nq : 0;
  if (qi >= nq) return;
  unsigned long long b1 = ~0ull, b2 = ~0ull;
  const int used = min(a.nsplits, (nt + a.rows_per_split - 1) / a.rows_per_split);
  for (int s = 0; s < used; ++s) {
    const unsigned long long* p = a.part + (((size_t)prob * a.nsplits + s) * a.max_rows + qi) * K;
    for (int k = 0; k < K; ++k) {
      unsigned long long key = p[k];
      unsigned long long hi = max(b1, key);
      b1 = min(b1, key);
      b2 = min(b2, hi);
    }
  }
  orbx_dmatch* o = a.out + ((size_t)prob * a.max_rows + qi) * K;
  int cnt = 0;
  if (b1 != ~0ull) {
    o[0] = orbx_dmatch{qi + a.q_index_base, (int)(b1 & 0xFFFFFFFFu), 0, (float)(unsigned)(b1 >> 32)};
    cnt = 1;
  }
  if (K == 2 && b2 != ~0ull) {
    o[1] = orbx_dmatch{qi + a.q_index_base, (int)(b2 & 0xFFFFFFFFu), 0, (float)(unsigned)(b2 >> 32)};
    cnt = 2;
  }
  if (a.n_per_q) a.n_per_q[(size_t)prob * a.max_rows + qi] = cnt;
}

// ------------------------------------------------------------------------------------------------
// ordered compaction helpers (one block per problem; order preserved = ascending query index)
// ------------------------------------------------------------------------------------------------
constexpr int kCompactThreads = 1024;

// returns the exclusive offset of this thread's flag within the block, and the block total
__device__ __forceinline__ int block_excl_scan_flag(bool flag, int* warp_sums, int& total) {
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  unsigned bal = __ballot_sync(0xFFFFFFFFu, flag);
  int within = __popc(bal & ((1u << lane) - 1));
  if (lane == 0) warp_sums[warp] = __popc(bal);
  __syncthreads();
  if (warp == 0) {
    int v = (lane < (blockDim.x >> 5)) ? warp_sums[lane] : 0;
    int incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      int n = __shfl_up_sync(0xFFFFFFFFu, incl, o);
      if (lane >= o) incl += n;
    }
    warp_sums[lane] = incl - v;
    if (lane == 31) warp_sums[32] = incl;
  }
  __syncthreads();
  int off = warp_sums[warp] + within;
  total = warp_sums[32];
  __syncthreads();
  return off;
}

// mode 0: Lowe ratio on knn2 (in: [rows][2], cnt: [rows]);  mode 1: min-distance filter on k=1 list
// mode 2: cross-check of forward k=1 (in) against reverse k=1 (rev)
__global__ void __launch_bounds__(kCompactThreads)
compact_kernel(int mode, const orbx_dmatch* in, const int* cnt, const unsigned long long* rev, const int* n_dev,
               int n_fixed, int max_rows, float ratio, orbx_dmatch* out, int* n_out) {
  __shared__ int warp_sums[33];
  __shared__ float s_min[32];
  __shared__ float s_thr;
  const int prob = blockIdx.x;
  int n = n_dev ? min(n_dev[prob], max_rows) : n_fixed;
  const size_t rowbase = (size_t)prob * max_rows;
  const int stride_in = (mode == 0) ? 2 : 1;
  in += rowbase * stride_in;
  if (cnt) cnt += rowbase;
  if (rev) rev += rowbase;
  out += rowbase;

  if (mode == 1) {
    float mn = 3.402823466e+38f;
    for (int i = threadIdx.x; i < n; i += blockDim.x) mn = fminf(mn, in[i].distance);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mn = fminf(mn, __shfl_xor_sync(0xFFFFFFFFu, mn, o));
    if ((threadIdx.x & 31) == 0) s_min[threadIdx.x >> 5] = mn;
    __syncthreads();
    if (threadIdx.x == 0) {
      float m2 = s_min[0];
      for (int w = 1; w < (int)(blockDim.x >> 5); ++w) m2 = fminf(m2, s_min[w]);
      s_thr = fmaxf(__fmul_rn(ratio, m2), 30.0f);  // matcher.rs:58
    }
    __syncthreads();
  }
  int written = 0;
  for (int base = 0; base < n; base += blockDim.x) {
    const int i = base + threadIdx.x;
    bool keep = false;
    orbx_dmatch m;
    if (i < n) {
      m = in[(size_t)i * stride_in];
      if (mode == 0) {
        keep = cnt[i] >= 2 && m.distance < __fmul_rn(ratio, in[(size_t)i * 2 + 1].distance);
      } else if (mode == 1) {
        keep = m.distance < s_thr;
      } else {
        keep = (int)(unsigned)(rev[m.train_idx] & 0xFFFFFFFFull) == i;
      }
    }
    int total;
    int off = block_excl_scan_flag(keep, warp_sums, total);
    if (keep) out[written + off] = m;
    written += total;
  }
  if (threadIdx.x == 0) n_out[prob] = written;
}

// zero n_per_q for queries failing the mutual-nearest test (k=2 + cross-check composition)
__global__ void knn2_cross_mask_kernel(const orbx_dmatch* knn2, const unsigned long long* rev, int nq, int* n_per_q) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nq) return;
  if (n_per_q[i] == 0) return;
  const int j = knn2[(size_t)i * 2].train_idx;
  if ((int)(unsigned)(rev[j] & 0xFFFFFFFFull) != i) n_per_q[i] = 0;
}

// batched form: problem p = blockIdx.y; rows at p*max_rows
__global__ void knn2_cross_mask_batch_kernel(const orbx_dmatch* knn2, const unsigned long long* rev, const int* nq_dev,
                                             int max_rows, int* n_per_q) {
  const int prob = blockIdx.y;
  const int nq = min(nq_dev[prob], max_rows);
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nq) return;
  const size_t row = (size_t)prob * max_rows + i;
  if (n_per_q[row] == 0) return;
  const int j = knn2[row * 2].train_idx;
  if ((int)(unsigned)(rev[(size_t)prob * max_rows + j] & 0xFFFFFFFFull) != i) n_per_q[row] = 0;
}

// ------------------------------------------------------------------------------------------------
// host-side launch logic
// ------------------------------------------------------------------------------------------------
// fused = also produce the reverse (train -> query) best-1 into a.rev (preset to ~0 by the caller)
// The same search on the tensor cores (match_mma.cu): expand the descriptor sets to int8 rows, run the tcgen05 kernel
// into per-split partial keys, and let merge_splits_kernel write the DMatch records exactly as the split POPC path does.
template <int K>
static int launch_knn_mma(orbx_matcher* m, MatchArgs a, int npairs, bool fused, int rows_q, int rows_t) {
  const int rpq = (int)align_up((size_t)rows_q, kMmaTileN), rpt = (int)align_up((size_t)rows_t, kMmaTileN);
  const size_t eq = mma_expanded_bytes(rpq), et = mma_expanded_bytes(rpt);
  // frame-to-frame batches: problem p matches set p against set p + 1 of ONE array - expand the npairs + 1 sets once
  const bool chain = npairs > 1 && a.nq_dev && a.nt_dev && a.q_stride == a.t_stride && a.t == a.q + a.q_stride &&
                     a.nt_dev == a.nq_dev + 1;
  int rc;
  m->timer.begin(0, m->stream);
  MmaMatchArgs g{};
  if (chain) {
    if ((rc = ensure(m, B_EXPQ, eq * (npairs + 1) + 256))) return rc;
    uint8_t* e = (uint8_t*)m->buf[B_EXPQ];
    if ((rc = launch_expand_desc(a.q, a.q_stride, a.nq_dev, 0, a.max_rows, npairs + 1, rpq, e, eq, m->stream))) return rc;
    g.ea = e;
    g.eb = e + eq;
    g.ea_stride = g.eb_stride = eq;
  } else {
    if ((rc = ensure(m, B_EXPQ, eq * npairs + 256))) return rc;
    if (et * npairs + 256 > m->cap[B_EXPT]) m->cached_t = nullptr;  // the buffer is about to be replaced
    if ((rc = ensure(m, B_EXPT, et * npairs + 256))) return rc;
    if ((rc = launch_expand_desc(a.q, a.q_stride, a.nq_dev, a.nq_fixed, a.max_rows, npairs, rpq, (uint8_t*)m->buf[B_EXPQ], eq,
                                 m->stream)))
      return rc;
    const bool cached = npairs == 1 && !a.nt_dev && m->cached_t == a.t && m->cached_nt == a.nt_fixed && m->buf[B_EXPT];
    if (!cached) {
      m->cached_t = nullptr;
      if ((rc = launch_expand_desc(a.t, a.t_stride, a.nt_dev, a.nt_fixed, a.max_rows, npairs, rpt, (uint8_t*)m->buf[B_EXPT],
                                   et, m->stream)))
        return rc;
    }
    g.ea = (const uint8_t*)m->buf[B_EXPQ];
    g.eb = (const uint8_t*)m->buf[B_EXPT];
    g.ea_stride = eq;
    g.eb_stride = et;
  }
  g.nq_dev = a.nq_dev, g.nt_dev = a.nt_dev, g.nq_fixed = a.nq_fixed, g.nt_fixed = a.nt_fixed;
  g.max_rows = a.max_rows;
  g.max_rows_t = a.nt_dev ? a.max_rows : a.nt_fixed;
  g.npairs = npairs;
  g.mtiles = div_up(rows_q, 128);
  // A unit streams its whole train range and the persistent CTAs take units round-robin, so a pass lasts
  // ceil(units * splits / SMs) waves of (tiles_per_split + ~2 tiles of per-unit cost: A tile fetch, drain, partial store).
  // Few units (one large train set): pick the split count that minimises that, i.e. one that fills whole waves.
  const int ttiles = div_up(rows_t, kMmaTileN);
  const bool pair = mma_use_pairs(g.mtiles, fused) && m->num_sms >= 2;
  const long units = (long)npairs * (pair ? (g.mtiles + 1) / 2 : g.mtiles);
  const long workers = pair ? std::max(m->num_sms / 2, 1) : m->num_sms;
  int splits = 1;
  if (units < workers * 2) {
    const long smax = std::min<long>(ttiles, std::max<long>(1, workers * 16 / units));
    long best = -1;
    for (long sp = 1; sp <= smax; ++sp) {
      const long tps = div_up(ttiles, (int)sp), ns = div_up(ttiles, (int)tps);
      const long waves = (units * ns + workers - 1) / workers;
      const long cost = waves * (tps + 2);
      if (best < 0 || cost < best) best = cost, splits = (int)sp;
    }
  }
  g.tiles_per_split = div_up(ttiles, std::max(splits, 1));
  g.nsplits = div_up(ttiles, g.tiles_per_split);
  g.K = K;
  if ((rc = ensure(m, B_PART, (size_t)npairs * g.nsplits * a.max_rows * K * sizeof(unsigned long long)))) return rc;
  g.part = (unsigned long long*)m->buf[B_PART];
  g.rev = fused ? a.rev : nullptr;
  g.rev_rows = a.rev_rows;
  if ((rc = launch_hamming_mma(g, m->num_sms, m->stream))) return rc;
  a.part = g.part;
  a.nsplits = g.nsplits;
  a.rows_per_split = g.tiles_per_split * kMmaTileN;
  dim3 g2(div_up(rows_q, 256), npairs);
  merge_splits_kernel<K><<<g2, 256, 0, m->stream>>>(a);
  m->timer.end(m->stream);
  ORBX_LAUNCH_CHECK();
  return ORBX_OK;
}

template <int K>
static int launch_knn(orbx_matcher* m, MatchArgs a, int npairs, bool fused = false) {
  const int rows_q = a.nq_dev ? a.max_rows : a.nq_fixed;
  const int rows_t = a.nt_dev ? a.max_rows : a.nt_fixed;
  if (rows_q <= 0 || npairs <= 0) return ORBX_OK;
  // tensor-core path for problems that fill the chip (>= 2^27 descriptor pairs); the POPC kernel below for small ones.
  // Against a RESIDENT train set (matcher_cache_train: its expansion is already paid) the tensor cores win from 2^23
  // pairs on: 128 queries x 1 M rows take 0.67 ms on the POPC pipe.
  const bool resident = npairs == 1 && !a.nt_dev && m->cached_t == a.t && m->cached_nt == a.nt_fixed && m->buf[B_EXPT];
  const double pairs = (double)npairs * rows_q * rows_t;
  if (rows_t > 0 && (m->mode == 2 || (m->mode == 0 && (pairs >= 134217728.0 || (resident && pairs >= 8388608.0)))))
    return launch_knn_mma<K>(m, a, npairs, fused, rows_q, rows_t);
  const int qtiles = div_up(rows_q, kQPB);
  // about four waves of blocks (10 resident per SM) so that the last, partly filled wave costs little, but never
  // split below one tile
  const long target = (long)m->num_sms * 40;
  const long nblk = std::max<long>(1, (long)qtiles * npairs);
  // (grids that already hold a few hundred long blocks - frame-to-frame batches - gain nothing from splitting: measured)
  int splits = nblk >= (long)m->num_sms * 4
                   ? 1
                   : (int)std::min<long>(std::max<long>(1, (target + nblk - 1) / nblk), std::max(1, div_up(rows_t, kTile)));
  int rps = (int)align_up((size_t)div_up(std::max(rows_t, 1), splits), kTile);
  if (rps > kMaxSplitRows) rps = kMaxSplitRows;
  splits = std::max(1, div_up(std::max(rows_t, 1), rps));
  a.rows_per_split = rps;
  a.nsplits = splits;
  if (splits > 1) {
    int rc = ensure(m, B_PART, (size_t)npairs * splits * a.max_rows * K * sizeof(unsigned long long));
    if (rc) return rc;
    a.part = (unsigned long long*)m->buf[B_PART];
  }
  dim3 grid(qtiles, splits, npairs);
  m->timer.begin(0, m->stream);
  if (fused)
    hamming_knn_kernel<K, true><<<grid, kQPB, 0, m->stream>>>(a);
  else
    hamming_knn_kernel<K, false><<<grid, kQPB, 0, m->stream>>>(a);
  m->timer.end(m->stream);
  ORBX_LAUNCH_CHECK();
  if (splits > 1) {
    dim3 g2(div_up(rows_q, 256), npairs);
    merge_splits_kernel<K><<<g2, 256, 0, m->stream>>>(a);
    ORBX_LAUNCH_CHECK();
  }
  return ORBX_OK;
}


// pts1[i] = kp1[matches[i].query_idx].pt, pts2[i] = kp2[matches[i].train_idx].pt  (pose.rs:36-41); problem p =
// blockIdx.y.  Counts come from device arrays when given (batched form) or from the scalars.
__global__ void matched_points_kernel(const orbx_keypoint* kp1, const int* n1_dev, int n1, const orbx_keypoint* kp2,
                                      const int* n2_dev, int n2, size_t kp_stride, const orbx_dmatch* matches,
                                      const int* n_dev, int n, int max_rows, float2* pts1, float2* pts2, int* bad) {
  const int p = blockIdx.y;
  const int nn = min(n_dev ? n_dev[p] : n, max_rows);
  const unsigned c1 = (unsigned)(n1_dev ? n1_dev[p] : n1), c2 = (unsigned)(n2_dev ? n2_dev[p] : n2);
  const orbx_keypoint* a = kp1 + (size_t)p * kp_stride;
  const orbx_keypoint* b = kp2 + (size_t)p * kp_stride;
  const size_t base = (size_t)p * max_rows;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nn; i += gridDim.x * blockDim.x) {
    const orbx_dmatch m = matches[base + i];
    if ((unsigned)m.query_idx >= c1 || (unsigned)m.train_idx >= c2) {  // Vector::get(..)? fails in the reference
      atomicOr(bad, 1);
      continue;
    }
    const orbx_keypoint& ka = a[m.query_idx];
    const orbx_keypoint& kb = b[m.train_idx];
    pts1[base + i] = make_float2(ka.x, ka.y);
    pts2[base + i] = make_float2(kb.x, kb.y);
  }
}

// Query-sharded matching (csrc/comm.cu): rows [index_base, index_base + nq) of a larger query set against the whole
// train set, device pointers only, asynchronous on the matcher's stream.  `out` (nq x K records) and `counts` (K = 2:
// nq ints, may be null) may live on ANOTHER device: the epilogue stores them through peer-mapped pointers.
int match_slice(orbx_matcher* m, int K, const uint8_t* q, int nq, const uint8_t* t, int nt, orbx_dmatch* out, int* counts,
                int index_base) {
  if (nq <= 0) return ORBX_OK;
  ORBX_CUDA(cudaSetDevice(m->device));
  MatchArgs a{};
  a.q = q;
  a.t = t;
  a.nq_fixed = nq;
  a.nt_fixed = nt;
  a.max_rows = nq;
  a.out = out;
  a.n_per_q = K == 2 ? counts : nullptr;
  a.q_index_base = index_base;
  return K == 1 ? launch_knn<1>(m, a, 1) : launch_knn<2>(m, a, 1);
}

cudaStream_t matcher_stream(orbx_matcher* m) { return m->stream; }

// A train set that stays resident (the map of map-to-frame matching): expand it once, asynchronously on the matcher's
// stream; later single-problem calls with the same (pointer, rows) reuse the expanded copy.  The caller invalidates by
// calling this again (or with nt = 0) when the contents change.
int matcher_cache_train(orbx_matcher* m, const uint8_t* t, int nt) {
  m->cached_t = nullptr;
  m->cached_nt = 0;
  if (!t || nt <= 0) return ORBX_OK;
  ORBX_CUDA(cudaSetDevice(m->device));
  const int rpt = (int)align_up((size_t)nt, kMmaTileN);
  const size_t et = mma_expanded_bytes(rpt);
  if (int rc = ensure(m, B_EXPT, et + 256)) return rc;
  if (int rc = launch_expand_desc(t, 0, nullptr, nt, nt, 1, rpt, (uint8_t*)m->buf[B_EXPT], et, m->stream)) return rc;
  m->cached_t = t;
  m->cached_nt = nt;
  return ORBX_OK;
}

}  // namespace orbx

using namespace orbx;

extern "C" {

int orbx_matcher_create(int device, int cross_check, orbx_matcher** out) {
  if (!out) {
    set_error("orbx_matcher_create: out is NULL");
    return ORBX_E_BADARG;
  }
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || device < 0 || device >= ndev) {
    cudaGetLastError();
    set_error("orbx_matcher_create: CUDA device %d unavailable (%s); liborbx has no CPU fallback", device,
              e != cudaSuccess ? cudaGetErrorString(e) : "index out of range");
    return ORBX_E_CUDA;
  }
  ORBX_CUDA(cudaSetDevice(device));
  orbx_matcher* m = new orbx_matcher();
  m->device = device;
  m->cross_check = cross_check ? 1 : 0;
  cudaDeviceGetAttribute(&m->num_sms, cudaDevAttrMultiProcessorCount, device);
  e = cudaStreamCreateWithFlags(&m->own_stream, cudaStreamNonBlocking);
  if (e != cudaSuccess) {
    set_error("cudaStreamCreate: %s", cudaGetErrorString(e));
    delete m;
    return ORBX_E_CUDA;
  }
  m->stream = m->own_stream;
  if (const char* v = getenv("ORBX_MATCHER")) m->mode = !strcmp(v, "popc") ? 1 : (!strcmp(v, "mma") ? 2 : 0);
  *out = m;
  return ORBX_OK;
}

void orbx_matcher_destroy(orbx_matcher* m) {
  if (!m) return;
  cudaSetDevice(m->device);
  cudaStreamSynchronize(m->stream);
  m->timer.destroy();
  for (int i = 0; i < 10; ++i)
    if (m->buf[i]) cudaFree(m->buf[i]);
  if (m->own_stream) cudaStreamDestroy(m->own_stream);
  delete m;
}

int orbx_matcher_sync(orbx_matcher* m) {
  if (!m) return ORBX_E_BADARG;
  ORBX_CUDA(cudaSetDevice(m->device));
  ORBX_CUDA(cudaStreamSynchronize(m->stream));
  return ORBX_OK;
}

int orbx_matcher_set_stream(orbx_matcher* m, void* stream) {
  if (!m) return ORBX_E_BADARG;
  m->stream = stream ? (cudaStream_t)stream : m->own_stream;
  return ORBX_OK;
}

int orbx_matcher_set_mode(orbx_matcher* m, int mode) {
  if (!m || mode < 0 || mode > 2) {
    set_error("orbx_matcher_set_mode: bad argument");
    return ORBX_E_BADARG;
  }
  m->mode = mode;
  return ORBX_OK;
}

int orbx_matcher_set_timing(orbx_matcher* m, int enable) {
  if (!m) return ORBX_E_BADARG;
  m->timer.enabled = enable != 0;
  return ORBX_OK;
}

int orbx_matcher_timing(orbx_matcher* m, float* ms, int* launches, int reset) {
  if (!m) return ORBX_E_BADARG;
  ORBX_CUDA(cudaSetDevice(m->device));
  m->timer.collect();
  if (ms) *ms = (float)m->timer.ms[0];
  if (launches) *launches = m->timer.launches[0];
  if (reset) m->timer.reset();
  return ORBX_OK;
}

// stage a host array on the device if needed; returns the device pointer to use
static int stage_in(orbx_matcher* m, int which, const void* p, size_t bytes, const void** dev) {
  if (bytes == 0 || is_device_ptr(p)) {
    *dev = p;
    return ORBX_OK;
  }
  int rc = ensure(m, which, bytes);
  if (rc) return rc;
  ORBX_CUDA(cudaMemcpyAsync(m->buf[which], p, bytes, cudaMemcpyHostToDevice, m->stream));
  *dev = m->buf[which];
  return ORBX_OK;
}

static int set_count(orbx_matcher* m, int* n, int value) {
  if (!n) return ORBX_OK;
  if (is_device_ptr(n)) {
    ORBX_CUDA(cudaMemcpyAsync(n, &value, sizeof(int), cudaMemcpyHostToDevice, m->stream));
    ORBX_CUDA(cudaStreamSynchronize(m->stream));
  } else {
    *n = value;
  }
  return ORBX_OK;
}

// shared implementation of orbx_match / orbx_knn_match2 for one problem
static int match_one(orbx_matcher* m, int K, const uint8_t* q, int nq, const uint8_t* t, int nt,
                     orbx_dmatch* out, int* n_or_counts) {
  if (!m || nq < 0 || nt < 0 || (nq > 0 && !q) || (nt > 0 && !t)) {
    set_error("orbx match: bad argument");
    return ORBX_E_BADARG;
  }
  ORBX_CUDA(cudaSetDevice(m->device));
  // empty guard == matcher.rs:29-31
  if (nq == 0 || nt == 0) {
    if (K == 1) return set_count(m, n_or_counts, 0);
    if (nq > 0 && n_or_counts) {
      if (is_device_ptr(n_or_counts))
        ORBX_CUDA(cudaMemsetAsync(n_or_counts, 0, sizeof(int) * nq, m->stream));
      else
        for (int i = 0; i < nq; ++i) n_or_counts[i] = 0;
    }
    return ORBX_OK;
  }
  if (!out) {
    set_error("orbx match: out is NULL");
    return ORBX_E_BADARG;
  }
  const void *dq, *dt;
  int rc;
  if ((rc = stage_in(m, B_Q, q, (size_t)nq * 32, &dq))) return rc;
  if ((rc = stage_in(m, B_T, t, (size_t)nt * 32, &dt))) return rc;
  const bool out_dev = is_device_ptr(out);
  const bool cnt_dev = n_or_counts && is_device_ptr(n_or_counts);
  const bool cross = m->cross_check != 0;

  // forward pass
  orbx_dmatch* fwd;
  const bool direct = out_dev && !(K == 1 && cross);
  if (direct) {
    fwd = out;
  } else {
    if ((rc = ensure(m, B_OUT, (size_t)nq * K * sizeof(orbx_dmatch)))) return rc;
    fwd = (orbx_dmatch*)m->buf[B_OUT];
  }
  int* counts = nullptr;
  if (K == 2) {
    if (cnt_dev) {
      counts = n_or_counts;
    } else {
      if ((rc = ensure(m, B_CNT, (size_t)nq * sizeof(int)))) return rc;
      counts = (int*)m->buf[B_CNT];
    }
  }
  MatchArgs a{};
  a.q = (const uint8_t*)dq;
  a.t = (const uint8_t*)dt;
  a.nq_fixed = nq;
  a.nt_fixed = nt;
  a.max_rows = nq;
  a.out = fwd;
  a.n_per_q = counts;
  if (cross) {
    // reverse best-1 (train -> query) comes out of the same distance evaluation (fused kernel)
    if ((rc = ensure(m, B_REV, (size_t)nt * sizeof(unsigned long long)))) return rc;
    ORBX_CUDA(cudaMemsetAsync(m->buf[B_REV], 0xFF, (size_t)nt * sizeof(unsigned long long), m->stream));
    a.rev = (unsigned long long*)m->buf[B_REV];
    a.rev_rows = nt;
  }
  if (K == 1)
    rc = launch_knn<1>(m, a, 1, cross);
  else
    rc = launch_knn<2>(m, a, 1, cross);
  if (rc) return rc;

  int n_final = nq;
  if (cross) {
    if (K == 1) {
      if ((rc = ensure(m, B_OUT2, (size_t)nq * sizeof(orbx_dmatch)))) return rc;
      if ((rc = ensure(m, B_CNT2, sizeof(int)))) return rc;
      orbx_dmatch* dst = out_dev ? out : (orbx_dmatch*)m->buf[B_OUT2];
      compact_kernel<<<1, kCompactThreads, 0, m->stream>>>(2, fwd, nullptr, (const unsigned long long*)m->buf[B_REV], nullptr,
                                                          nq, nq, 0.f, dst, (int*)m->buf[B_CNT2]);
      ORBX_LAUNCH_CHECK();
      ORBX_CUDA(cudaMemcpyAsync(&n_final, m->buf[B_CNT2], sizeof(int), cudaMemcpyDeviceToHost, m->stream));
      ORBX_CUDA(cudaStreamSynchronize(m->stream));
      fwd = dst;
    } else {
      knn2_cross_mask_kernel<<<div_up(nq, 256), 256, 0, m->stream>>>(fwd, (const unsigned long long*)m->buf[B_REV], nq, counts);
      ORBX_LAUNCH_CHECK();
    }
  }
  if (!out_dev) {
    const size_t rows = (K == 1) ? (size_t)n_final : (size_t)nq * 2;
    if (rows)
      ORBX_CUDA(cudaMemcpyAsync(out, fwd, rows * sizeof(orbx_dmatch), cudaMemcpyDeviceToHost, m->stream));
  }
  if (K == 2 && n_or_counts && !cnt_dev)
    ORBX_CUDA(cudaMemcpyAsync(n_or_counts, counts, (size_t)nq * sizeof(int), cudaMemcpyDeviceToHost, m->stream));
  if (!out_dev || (n_or_counts && !cnt_dev)) ORBX_CUDA(cudaStreamSynchronize(m->stream));
  if (K == 1) return set_count(m, n_or_counts, n_final);
  return ORBX_OK;
}

int orbx_match(orbx_matcher* m, const uint8_t* q, int nq, const uint8_t* t, int nt, orbx_dmatch* out, int* n) {
  return match_one(m, 1, q, nq, t, nt, out, n);
}

int orbx_knn_match2(orbx_matcher* m, const uint8_t* q, int nq, const uint8_t* t, int nt, orbx_dmatch* out,
                    int* n_per_q) {
  return match_one(m, 2, q, nq, t, nt, out, n_per_q);
}

static int batch_args_ok(orbx_matcher* m, const void* q, const void* t, const void* out, int npairs, int max_rows) {
  if (!m || npairs < 0 || max_rows < 0) {
    set_error("orbx batch match: bad argument");
    return ORBX_E_BADARG;
  }
  if (npairs == 0 || max_rows == 0) return ORBX_OK;
  if (!is_device_ptr(q) || !is_device_ptr(t) || !is_device_ptr(out)) {
    set_error("orbx batch match: q, t and out must be device pointers");
    return ORBX_E_BADARG;
  }
  return ORBX_OK;
}

int orbx_knn_match2_batch(orbx_matcher* m, const uint8_t* q, size_t q_stride, const int* nq_dev, const uint8_t* t,
                          size_t t_stride, const int* nt_dev, int npairs, int max_rows, orbx_dmatch* out,
                          int* n_per_q) {
  int rc = batch_args_ok(m, q, t, out, npairs, max_rows);
  if (rc || npairs == 0 || max_rows == 0) return rc;
  if (!nq_dev || !nt_dev || !is_device_ptr(nq_dev) || !is_device_ptr(nt_dev)) {
    set_error("orbx_knn_match2_batch: counts must be device arrays");
    return ORBX_E_BADARG;
  }
  ORBX_CUDA(cudaSetDevice(m->device));
  MatchArgs a{};
  a.q = q;
  a.t = t;
  a.q_stride = q_stride;
  a.t_stride = t_stride;
  a.nq_dev = nq_dev;
  a.nt_dev = nt_dev;
  a.max_rows = max_rows;
  a.out = out;
  a.n_per_q = n_per_q;
  if (n_per_q)
    ORBX_CUDA(cudaMemsetAsync(n_per_q, 0, (size_t)npairs * max_rows * sizeof(int), m->stream));
  const bool cross = m->cross_check && n_per_q;
  if (cross) {
    // mutual-nearest composition: the reverse best-1 is produced by the same launch
    const size_t rev_bytes = (size_t)npairs * max_rows * sizeof(unsigned long long);
    if ((rc = ensure(m, B_REV, rev_bytes))) return rc;
    ORBX_CUDA(cudaMemsetAsync(m->buf[B_REV], 0xFF, rev_bytes, m->stream));
    a.rev = (unsigned long long*)m->buf[B_REV];
    a.rev_rows = max_rows;
  }
  if ((rc = launch_knn<2>(m, a, npairs, cross))) return rc;
  if (cross) {
    dim3 g(div_up(max_rows, 256), npairs);
    knn2_cross_mask_batch_kernel<<<g, 256, 0, m->stream>>>(out, (const unsigned long long*)m->buf[B_REV], nq_dev, max_rows,
                                                          n_per_q);
    ORBX_LAUNCH_CHECK();
  }
  return ORBX_OK;
}

int orbx_ratio_test_batch(orbx_matcher* m, const orbx_dmatch* knn2, const int* n_per_q, const int* nq_dev,
                          int npairs, int max_rows, float ratio, orbx_dmatch* out, int* n_out_dev) {
  if (!m || npairs < 0 || max_rows < 0) {
    set_error("orbx_ratio_test_batch: bad argument");
    return ORBX_E_BADARG;
  }
  if (npairs == 0) return ORBX_OK;
  if (!is_device_ptr(knn2) || !is_device_ptr(n_per_q) || !is_device_ptr(nq_dev) || !is_device_ptr(out) ||
      !is_device_ptr(n_out_dev)) {
    set_error("orbx_ratio_test_batch: device pointers required");
    return ORBX_E_BADARG;
  }
  ORBX_CUDA(cudaSetDevice(m->device));
  compact_kernel<<<npairs, kCompactThreads, 0, m->stream>>>(0, knn2, n_per_q, nullptr, nq_dev, 0, max_rows, ratio, out,
                                                           n_out_dev);
  ORBX_LAUNCH_CHECK();
  return ORBX_OK;
}

int orbx_match_batch(orbx_matcher* m, const uint8_t* q, size_t q_stride, const int* nq_dev, const uint8_t* t,
                     size_t t_stride, const int* nt_dev, int npairs, int max_rows, orbx_dmatch* out,
                     int* n_out_dev) {
  int rc = batch_args_ok(m, q, t, out, npairs, max_rows);
  if (rc || npairs == 0 || max_rows == 0) return rc;
  if (!nq_dev || !nt_dev || !is_device_ptr(nq_dev) || !is_device_ptr(nt_dev) || !n_out_dev ||
      !is_device_ptr(n_out_dev)) {
    set_error("orbx_match_batch: counts must be device arrays");
    return ORBX_E_BADARG;
  }
  ORBX_CUDA(cudaSetDevice(m->device));
  ORBX_CUDA(cudaMemsetAsync(n_out_dev, 0, (size_t)npairs * sizeof(int), m->stream));
  MatchArgs a{};
  a.q = q;
  a.t = t;
  a.q_stride = q_stride;
  a.t_stride = t_stride;
  a.nq_dev = nq_dev;
  a.nt_dev = nt_dev;
  a.max_rows = max_rows;
  a.out = out;
  a.n_out = n_out_dev;
  return launch_knn<1>(m, a, npairs);
}

// one-problem compaction for host or device arrays
static int compact_one(orbx_matcher* m, int mode, const orbx_dmatch* in, const int* cnt, int n, float ratio,
                       orbx_dmatch* out, int* n_out) {
  if (!m || n < 0 || !n_out) {
    set_error("orbx filter: bad argument");
    return ORBX_E_BADARG;
  }
  ORBX_CUDA(cudaSetDevice(m->device));
  if (n == 0) return set_count(m, n_out, 0);
  if (!in || !out || (mode == 0 && !cnt)) {
    set_error("orbx filter: NULL array");
    return ORBX_E_BADARG;
  }
  const size_t in_rows = (mode == 0) ? (size_t)n * 2 : (size_t)n;
  const void *din, *dcnt = nullptr;
  int rc;
  if ((rc = stage_in(m, B_OUT, in, in_rows * sizeof(orbx_dmatch), &din))) return rc;
  if (mode == 0 && (rc = stage_in(m, B_CNT, cnt, (size_t)n * sizeof(int), &dcnt))) return rc;
  const bool out_dev = is_device_ptr(out);
  orbx_dmatch* dout = out;
  if (!out_dev) {
    if ((rc = ensure(m, B_OUT2, (size_t)n * sizeof(orbx_dmatch)))) return rc;
    dout = (orbx_dmatch*)m->buf[B_OUT2];
  }
  if ((rc = ensure(m, B_CNT2, sizeof(int)))) return rc;
  compact_kernel<<<1, kCompactThreads, 0, m->stream>>>(mode, (const orbx_dmatch*)din, (const int*)dcnt, nullptr,
                                                      nullptr, n, n, ratio, dout, (int*)m->buf[B_CNT2]);
  ORBX_LAUNCH_CHECK();
  int cnt_h = 0;
  ORBX_CUDA(cudaMemcpyAsync(&cnt_h, m->buf[B_CNT2], sizeof(int), cudaMemcpyDeviceToHost, m->stream));
  ORBX_CUDA(cudaStreamSynchronize(m->stream));
  if (!out_dev && cnt_h)
    ORBX_CUDA(cudaMemcpy(out, dout, (size_t)cnt_h * sizeof(orbx_dmatch), cudaMemcpyDeviceToHost));
  return set_count(m, n_out, cnt_h);
}

int orbx_ratio_test(orbx_matcher* m, const orbx_dmatch* knn2, const int* n_per_q, int nq, float ratio,
                    orbx_dmatch* out, int* n) {
  return compact_one(m, 0, knn2, n_per_q, nq, ratio, out, n);
}

int orbx_filter_good_matches(orbx_matcher* m, const orbx_dmatch* in, int n, float ratio, orbx_dmatch* out,
                             int* n_out) {
  return compact_one(m, 1, in, nullptr, n, ratio, out, n_out);
}

int orbx_filter_good_matches_batch(orbx_matcher* m, const orbx_dmatch* in, const int* n_dev, int npairs,
                                   int max_rows, float ratio, orbx_dmatch* out, int* n_out_dev) {
  if (!m || npairs < 0 || max_rows < 0) {
    set_error("orbx_filter_good_matches_batch: bad argument");
    return ORBX_E_BADARG;
  }
  if (npairs == 0) return ORBX_OK;
  if (!is_device_ptr(in) || !is_device_ptr(out) || !is_device_ptr(n_dev) || !is_device_ptr(n_out_dev)) {
    set_error("orbx_filter_good_matches_batch: device pointers required");
    return ORBX_E_BADARG;
  }
  ORBX_CUDA(cudaSetDevice(m->device));
  compact_kernel<<<npairs, kCompactThreads, 0, m->stream>>>(1, in, nullptr, nullptr, n_dev, 0, max_rows, ratio, out,
                                                           n_out_dev);
  ORBX_LAUNCH_CHECK();
  return ORBX_OK;
}

// ------------------------------------------------------------------------------------------------
// Allocation-free k=2 search (no counterpart in the reference).
// Callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it) get
// the train-split Hamming search with its partial-key scratch in a file-scope device array instead of the handle's
// cudaMalloc'ed buffer.  The array costs kNoAllocScratchBytes (8 MiB) on every GPU that loads liborbx, until exit.
// ------------------------------------------------------------------------------------------------
}  // extern "C"

namespace orbx {
constexpr int kNoAllocMaxQueries = 8192;  // query rows per call
constexpr int kNoAllocMaxSplits = 64;     // train ranges per call (each <= 65536 rows)
constexpr size_t kNoAllocScratchWords = (size_t)kNoAllocMaxQueries * kNoAllocMaxSplits * 2;  // u64 keys: 8 MiB
// [split][query][2] packed (dist << 32 | train_idx).  In HOST code this name is only a stand-in for the symbol: the
// device address comes from cudaGetSymbolAddress (no_alloc_scratch below), once per device.
static __device__ __align__(256) unsigned long long g_noalloc_part[kNoAllocScratchWords];

// device address of g_noalloc_part and the SM count of `device` (which must be current), looked up once per device
static int no_alloc_scratch(int device, unsigned long long** ptr, int* sms) {
  constexpr int kMaxDev = 64;
  static unsigned long long* addr[kMaxDev] = {nullptr};
  static int nsm[kMaxDev] = {0};
  static std::mutex mu;
  if (device < 0 || device >= kMaxDev) {
    set_error("orbx_knn_match2_static: device %d out of range", device);
    return ORBX_E_BADARG;
  }
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[device]) {
    void* p = nullptr;
    ORBX_CUDA(cudaGetSymbolAddress(&p, g_noalloc_part));
    int n = 0;
    ORBX_CUDA(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device));
    addr[device] = (unsigned long long*)p;
    nsm[device] = n;
  }
  *ptr = addr[device];
  *sms = nsm[device];
  return ORBX_OK;
}
}  // namespace orbx

extern "C" {

int orbx_knn_match2_static(int device, const uint8_t* q, int nq, const uint8_t* t, int nt, orbx_dmatch* out, int* n_per_q,
                           void* stream) {
  if (nq < 0 || nt < 0 || (nq > 0 && (!q || !out || !n_per_q)) || (nt > 0 && nq > 0 && !t)) {
    set_error("orbx_knn_match2_static: bad argument");
    return ORBX_E_BADARG;
  }
  if (nq == 0) return ORBX_OK;
  if (nq > kNoAllocMaxQueries) {
    set_error("orbx_knn_match2_static: %d query rows exceed the static scratch (at most %d per call; use orbx_knn_match2)",
              nq, kNoAllocMaxQueries);
    return ORBX_E_CAPACITY;
  }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) {
    cudaGetLastError();
    set_error("orbx_knn_match2_static: CUDA device %d unavailable (there is no CPU fallback)", device);
    return ORBX_E_CUDA;
  }
  ORBX_CUDA(cudaSetDevice(device));
  if (!is_device_ptr(q) || !is_device_ptr(out) || !is_device_ptr(n_per_q) || (nt > 0 && !is_device_ptr(t))) {
    set_error("orbx_knn_match2_static: q, t, out and n_per_q must be device memory (staging would allocate)");
    return ORBX_E_BADARG;
  }
  cudaStream_t s = (cudaStream_t)stream;
  if (nt == 0) {  // matcher.rs:29-31: an empty side is an empty result
    ORBX_CUDA(cudaMemsetAsync(n_per_q, 0, sizeof(int) * (size_t)nq, s));
    return ORBX_OK;
  }
  unsigned long long* part = nullptr;
  int sms = 148;
  if (int rc = no_alloc_scratch(device, &part, &sms)) return rc;
  // the split rule of launch_knn: about four waves of blocks, ranges of whole 256-row tiles, at most 65536 rows each
  const int qtiles = div_up(nq, kQPB);
  int splits = qtiles >= sms * 4 ? 1 : std::min(std::max(1, div_up(sms * 40, qtiles)), std::max(1, div_up(nt, kTile)));
  splits = std::min(splits, kNoAllocMaxSplits);
  int rps = (int)align_up((size_t)div_up(nt, splits), kTile);
  if (rps > kMaxSplitRows) rps = kMaxSplitRows;
  splits = div_up(nt, rps);
  if (splits > kNoAllocMaxSplits) {
    set_error("orbx_knn_match2_static: %d train rows exceed the static scratch (at most %d per call; use orbx_knn_match2)",
              nt, kNoAllocMaxSplits * kMaxSplitRows);
    return ORBX_E_CAPACITY;
  }
  MatchArgs a{};
  a.q = q;
  a.t = t;
  a.nq_fixed = nq;
  a.nt_fixed = nt;
  a.max_rows = nq;
  a.out = out;
  a.n_per_q = n_per_q;
  a.rows_per_split = rps;
  a.nsplits = splits;
  a.part = part;  // every (split, active query) slot is written by the launch below before the merge reads it:
                  // nothing in the array has to start at zero, and nothing of an earlier call is read
  hamming_knn_kernel<2, false><<<dim3(qtiles, splits, 1), kQPB, 0, s>>>(a);
  ORBX_LAUNCH_CHECK();
  if (splits > 1) {
    merge_splits_kernel<2><<<dim3(div_up(nq, 256), 1), 256, 0, s>>>(a);
    ORBX_LAUNCH_CHECK();
  }
  return ORBX_OK;
}

// ------------------------------------------------------------------------------------------------
// SURVEY.md §8 f2 (first step): PoseEstimator::extract_matched_points
// ------------------------------------------------------------------------------------------------
int orbx_extract_matched_points(orbx_matcher* m, const orbx_keypoint* kp1, int n1, const orbx_keypoint* kp2, int n2,
                                const orbx_dmatch* matches, int n, float* pts1, float* pts2) {
  if (!m || n1 < 0 || n2 < 0 || n < 0 || (n > 0 && (!matches || !pts1 || !pts2 || !kp1 || !kp2))) {
    set_error("orbx_extract_matched_points: bad argument");
    return ORBX_E_BADARG;
  }
  if (n == 0) return ORBX_OK;
  ORBX_CUDA(cudaSetDevice(m->device));
  const void *d1, *d2, *dm;
  int rc;
  if ((rc = stage_in(m, B_Q, kp1, (size_t)n1 * sizeof(orbx_keypoint), &d1))) return rc;
  if ((rc = stage_in(m, B_T, kp2, (size_t)n2 * sizeof(orbx_keypoint), &d2))) return rc;
  if ((rc = stage_in(m, B_PART, matches, (size_t)n * sizeof(orbx_dmatch), &dm))) return rc;
  const bool o1 = is_device_ptr(pts1), o2 = is_device_ptr(pts2);
  if ((rc = ensure(m, B_CNT, sizeof(int)))) return rc;
  if (!o1 && (rc = ensure(m, B_OUT, (size_t)n * sizeof(float2)))) return rc;
  if (!o2 && (rc = ensure(m, B_OUT2, (size_t)n * sizeof(float2)))) return rc;
  float2* p1 = o1 ? (float2*)pts1 : (float2*)m->buf[B_OUT];
  float2* p2 = o2 ? (float2*)pts2 : (float2*)m->buf[B_OUT2];
  ORBX_CUDA(cudaMemsetAsync(m->buf[B_CNT], 0, sizeof(int), m->stream));
  matched_points_kernel<<<dim3(std::min(div_up(n, 256), 1024), 1), 256, 0, m->stream>>>(
      (const orbx_keypoint*)d1, nullptr, n1, (const orbx_keypoint*)d2, nullptr, n2, 0, (const orbx_dmatch*)dm, nullptr,
      n, n, p1, p2, (int*)m->buf[B_CNT]);
  ORBX_LAUNCH_CHECK();
  int bad = 0;
  ORBX_CUDA(cudaMemcpyAsync(&bad, m->buf[B_CNT], sizeof(int), cudaMemcpyDeviceToHost, m->stream));
  if (!o1) ORBX_CUDA(cudaMemcpyAsync(pts1, p1, (size_t)n * sizeof(float2), cudaMemcpyDeviceToHost, m->stream));
  if (!o2) ORBX_CUDA(cudaMemcpyAsync(pts2, p2, (size_t)n * sizeof(float2), cudaMemcpyDeviceToHost, m->stream));
  ORBX_CUDA(cudaStreamSynchronize(m->stream));
  if (bad) {
    set_error("orbx_extract_matched_points: a match index is outside its keypoint array");
    return ORBX_E_BADARG;
  }
  return ORBX_OK;
}

int orbx_extract_matched_points_batch(orbx_matcher* m, const orbx_keypoint* kp1, const int* n1_dev,
                                      const orbx_keypoint* kp2, const int* n2_dev, size_t kp_stride,
                                      const orbx_dmatch* matches, const int* n_dev, int npairs, int max_rows,
                                      float* pts1, float* pts2, int* bad_dev) {
  if (!m || npairs < 0 || max_rows < 0) {
    set_error("orbx_extract_matched_points_batch: bad argument");
    return ORBX_E_BADARG;
  }
  if (npairs == 0 || max_rows == 0) return ORBX_OK;
  if (!is_device_ptr(kp1) || !is_device_ptr(kp2) || !is_device_ptr(n1_dev) || !is_device_ptr(n2_dev) ||
      !is_device_ptr(matches) || !is_device_ptr(n_dev) || !is_device_ptr(pts1) || !is_device_ptr(pts2) ||
      !is_device_ptr(bad_dev)) {
    set_error("orbx_extract_matched_points_batch: device pointers required");
    return ORBX_E_BADARG;
  }
  ORBX_CUDA(cudaSetDevice(m->device));
  matched_points_kernel<<<dim3(std::min(div_up(max_rows, 256), 16), npairs), 256, 0, m->stream>>>(
      kp1, n1_dev, 0, kp2, n2_dev, 0, kp_stride, matches, n_dev, 0, max_rows, (float2*)pts1, (float2*)pts2, bad_dev);
  ORBX_LAUNCH_CHECK();
  return ORBX_OK;
}

// ------------------------------------------------------------------------------------------------
// POPC-pipe microbenchmark: the denominator of the S10 roofline
// ------------------------------------------------------------------------------------------------
}  // extern "C"

template <int MODE>
__global__ void popc_bench_kernel(unsigned* out, int iters, unsigned seed) {
  unsigned v[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = seed * (threadIdx.x + 1) + i * 0x9E3779B9u + blockIdx.x;
  unsigned acc = 0, b1 = ~0u, b2 = ~0u;
  for (int it = 0; it < iters; ++it) {
    if (MODE == 0) {
      // 8 independent POPC per iteration; operands change every iteration through a cheap add
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        unsigned p = __popc(v[i]);
        v[i] = v[i] + p + 0x01010101u;
      }
    } else {
      unsigned d = 0;
#pragma unroll
      for (int i = 0; i < 8; ++i) d += __popc(v[i] ^ (acc + i));
      unsigned key = (d << 16) | (unsigned)(it & 0xFFFF);
      unsigned hi = max(b1, key);
      b1 = min(b1, key);
      b2 = min(b2, hi);
      acc += 0x9E3779B9u;
    }
  }
  unsigned r = acc ^ b1 ^ b2;
#pragma unroll
  for (int i = 0; i < 8; ++i) r ^= v[i];
  if (r == 0x12345678u) out[0] = r;  // keep the work alive
}

extern "C" {

int orbx_popc_microbench(int device, int mode, double* popc_per_s) {
  if (!popc_per_s) return ORBX_E_BADARG;
  ORBX_CUDA(cudaSetDevice(device));
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
  unsigned* d = nullptr;
  ORBX_CUDA(cudaMalloc(&d, 4));
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  const int iters = 1 << 14, threads = 256, blocks = sms * 8;
  float best = 1e30f;
  for (int rep = 0; rep < 5; ++rep) {
    cudaEventRecord(e0);
    if (mode == 0)
      popc_bench_kernel<0><<<blocks, threads>>>(d, iters, 12345u + rep);
    else
      popc_bench_kernel<1><<<blocks, threads>>>(d, iters, 12345u + rep);
    g_launches.fetch_add(1);
    cudaEventRecord(e1);
    ORBX_CUDA(cudaEventSynchronize(e1));
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    if (rep > 0 && ms < best) best = ms;
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(d);
  *popc_per_s = (double)blocks * threads * (double)iters * 8.0 / (best * 1e-3);
  return ORBX_OK;
}

}  // extern "C"
