// liborbx — SURVEY.md §8 f2: odometry::PoseEstimator::compute_essential_matrix / recover_pose
// (src/odometry/pose.rs:47-114) on the device, batched over frame pairs.
//
// The reference calls cv::findEssentialMat(RANSAC, 0.999, 1.0, 1000) and cv::recoverPose.  What is reproduced here
// is that procedure, not an approximation of it (oracle/pose_restated.py states it and is pinned to cv2):
//   * points normalised with K in f64, threshold / ((fx + fy) / 2);
//   * the RANSAC index draws of cv::RNG(2^64 - 1) — the same 5-point subsets in the same order;
//   * every subset's five-point solutions (Nister: null space of the 5x9 system, ten cubic constraints, Gauss-Jordan,
//     det B(z) = 0 of degree 10) scored on all points with the Sampson error rounded to fp32;
//   * the sequential acceptance rule (strictly more inliers than the best so far) and RANSACUpdateNumIters — run as a
//     scan over hypotheses that were evaluated in parallel, so the winning hypothesis is the one OpenCV picks;
//   * recoverPose: SVD of E, the four (R, +-t) candidates, per-point triangulation with OpenCV's depth tests.
// Hypotheses are evaluated in waves (32, 64, 160, 744): the iteration budget usually falls below 100 after the first
// good model, and the kernels of later waves exit at once.  One thread solves one five-point problem in f64 (local
// arrays; the work is tiny next to extraction), one block counts the inliers of one hypothesis.
#include "common.cuh"
#include "jacobi.cuh"

using orbx::div_up;
using orbx::is_device_ptr;
using orbx::set_error;

namespace orbx {

constexpr int kPoseMaxIters = 1000;
constexpr int kPoseMaxModels = 10;
// Hypotheses are solved speculatively in waves (the iteration budget shrinks as soon as a good model is counted).  The
// solve kernel holds 87 threads per SM (2.6 KB of shared scratch each), so a wave of len x pairs solves costs
// ceil(len x pairs / 12.9 k) solve latencies: the first wave is short (the budget is ~5 at 95 % inliers).
constexpr int kPoseWaves = 5;
__constant__ int c_wave[kPoseWaves + 1] = {0, 8, 32, 96, 256, kPoseMaxIters};

struct PoseState {
  unsigned long long rng;
  int n;          // points of this problem (clamped to max_rows)
  int niters;     // current iteration budget
  int max_good;   // inliers of the best model so far
  int best_it, best_m;
  int iters_run;
  int status;     // ORBX_POSE_*
  int cand[4];    // recoverPose: points passing the depth tests per candidate
  int choice;
  double R1[9], R2[9], t[3];
};

struct PoseParams {
  double fx, fy, cx, cy;
  double prob, thr2f;  // thr2f = (double)(float)(thr * thr)
  int min_matches;
  int max_rows;        // row capacity per problem in every per-point array
};

// ORBX_HOST_TESTABLE_BEGIN (tests/test_pose_cpu.py compiles this section with g++ to check the solver without a GPU)
// monomial slot tables: linear [x y z 1] x linear -> quadratic [x2 xy xz x y2 yz y z2 z 1];
// quadratic x linear -> cubic in the column order x3 y3 x2y xy2 x2z x2 y2z y2 xyz xy | xz2 xz x yz2 yz y z3 z2 z 1
__constant__ unsigned char c_ll[4][4] = {{0, 1, 2, 3}, {1, 4, 5, 6}, {2, 5, 7, 8}, {3, 6, 8, 9}};
__constant__ unsigned char c_ql[10][4] = {{0, 2, 4, 5},    {2, 3, 8, 9},    {4, 8, 10, 11},   {5, 9, 11, 12}, {3, 1, 6, 7},
                                          {8, 6, 13, 14},  {9, 7, 14, 15},  {10, 13, 16, 17}, {11, 14, 17, 18},
                                          {12, 15, 18, 19}};

// ORBX_HOST_TESTABLE_SPLIT
// ------------------------------------------------------------------------------------------------ prepare
// normalise the points (cv::findEssentialMat: (x - cx) / fx in f64) and reset the per-problem state
__global__ void pose_prepare_kernel(const float* __restrict__ pts1, const float* __restrict__ pts2,
                                    const int* __restrict__ n_dev, int n_fixed, PoseParams pp, double* __restrict__ x1,
                                    double* __restrict__ x2, PoseState* __restrict__ st, bool reset_ransac) {
  const int p = blockIdx.y;
  const int n = min(n_dev ? n_dev[p] : n_fixed, pp.max_rows);
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    PoseState& s = st[p];
    s.n = n;
    if (reset_ransac) {
      s.rng = 0xFFFFFFFFFFFFFFFFull;
      s.niters = n >= pp.min_matches ? kPoseMaxIters : 0;
      s.max_good = 0;
      s.best_it = s.best_m = -1;
      s.iters_run = 0;
      s.status = n >= pp.min_matches ? ORBX_POSE_OK : ORBX_POSE_INSUFFICIENT_POINTS;
    }
    s.cand[0] = s.cand[1] = s.cand[2] = s.cand[3] = 0;
    s.choice = 0;
  }
  const size_t base = (size_t)p * pp.max_rows;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const float2 a = reinterpret_cast<const float2*>(pts1)[base + i];
    const float2 b = reinterpret_cast<const float2*>(pts2)[base + i];
    x1[2 * (base + i)] = __ddiv_rn(__dsub_rn((double)a.x, pp.cx), pp.fx);
    x1[2 * (base + i) + 1] = __ddiv_rn(__dsub_rn((double)a.y, pp.cy), pp.fy);
    x2[2 * (base + i)] = __ddiv_rn(__dsub_rn((double)b.x, pp.cx), pp.fx);
    x2[2 * (base + i) + 1] = __ddiv_rn(__dsub_rn((double)b.y, pp.cy), pp.fy);
  }
}

// ------------------------------------------------------------------------------------------------ RANSAC draws
// cv::RNG multiply-with-carry; getSubset: 5 indices, a repeated index is redrawn.  One thread per problem (the
// sequence is serial by construction); only the hypotheses of this wave are drawn.
__global__ void pose_sample_kernel(PoseState* __restrict__ st, int wave, int* __restrict__ samples) {
  const int p = blockIdx.x;
  if (threadIdx.x != 0) return;
  PoseState& s = st[p];
  const int w0 = c_wave[wave], w1 = c_wave[wave + 1];
  if (w0 >= s.niters) return;
  unsigned long long state = s.rng;
  const unsigned n = (unsigned)s.n;
  int* out = samples + ((size_t)p * kPoseMaxIters + w0) * 5;
  for (int it = w0; it < w1; ++it) {
    int idx[5];
    for (int i = 0; i < 5; ++i) {
      for (;;) {
        state = (unsigned long long)(unsigned)state * 4164903690ull + (state >> 32);
        const int v = (int)((unsigned)state % n);
        bool dup = false;
        for (int j = 0; j < i; ++j) dup |= idx[j] == v;
        if (!dup) {
          idx[i] = v;
          break;
        }
      }
      out[i] = idx[i];
    }
    out += 5;
  }
  s.rng = state;
}

// ------------------------------------------------------------------------------------------------ five-point
// ORBX_HOST_TESTABLE_RESUME
// The solver's large per-thread arrays live in a strided scratch area: element k of a thread is sm[k * SS].  On the
// device SS = 32 and sm points into shared memory at the thread's lane (conflict-free for any per-thread index; the
// arrays are indexed dynamically, so the alternative is local memory that falls out of L1); the host harness uses
// SS = 1.  Layout (doubles): M 10x20 at 0 (aliased first by the 5x9 epipolar system, later by the derivative table
// of the root finder), EEt 3x3x10 at 200, the null-space basis 4x9 at 290.
constexpr int kSolveScratch = 326;

template <int SS>
__device__ __forceinline__ double horner_s(const double* c, int deg, double z) {
  double v = c[deg * SS];
  for (int i = deg - 1; i >= 0; --i) v = v * z + c[i * SS];
  return v;
}

__device__ __forceinline__ double horner(const double* c, int deg, double z) { return horner_s<1>(c, deg, z); }

// real roots of a polynomial of degree <= 10 (ascending coefficients): between consecutive real roots of p' the
// polynomial is monotonic, so the roots of p^(k) bracket those of p^(k-1); every bracket with a sign change is
// bisected.  d: scratch for the derivative table (11 x 11, strided).  Returns the number of roots (ascending).
//
// The kernel is latency-bound (one solve per thread, dependent DFMA chains: profiles/r01_ncu_full_pose_v9.csv has 70 %
// of its stall samples in the Horner loop of the bisection), so each level is evaluated in independent chains: all
// bracket ends in one multi-point Horner pass, then the brackets with a sign change bisected four at a time.  Every
// bracket sees exactly the arithmetic of a one-at-a-time bisection; only the interleaving differs.
#ifndef ORBX_ROOT_LANES
#define ORBX_ROOT_LANES 4  // tests/test_pose_cpu.py builds the host harness with 1 as well and compares bit for bit
#endif
constexpr int kRootLanes = ORBX_ROOT_LANES;

template <int SS>
__device__ int real_roots(const double* coef, int deg, double* roots, double* d) {
  while (deg > 0 && coef[deg] == 0.0) --deg;
  if (deg <= 0) return 0;
#define D_(k, i) d[((k) * 11 + (i)) * SS]
  for (int i = 0; i <= deg; ++i) D_(0, i) = coef[i];
  for (int k = 1; k < deg; ++k)
    for (int i = 0; i <= deg - k; ++i) D_(k, i) = D_(k - 1, i + 1) * (double)(i + 1);
  double prev[10], cur[10];
  double ba[10], bb[10], bfa[10];  // brackets to bisect at this level: ends, value at the low end
  int bpos[10];                    // and where their root goes in cur
  int nprev = 1;
  prev[0] = -D_(deg - 1, 0) / D_(deg - 1, 1);  // the linear one
  for (int k = deg - 2; k >= 0; --k) {
    const int dg = deg - k;
    const double* c = &D_(k, 0);
    double mx = 0;  // Cauchy bound of this polynomial
    for (int i = 0; i < dg; ++i) mx = fmax(mx, fabs(c[i * SS] / c[dg * SS]));
    const double B = 1.0 + mx;
    // the polynomial at the bracket ends x_0 = -B, x_j = critical point j - 1 (clamped), x_(nprev + 1) = B
    double xs[12], fs[12];
    xs[0] = -B, xs[11] = B;
#pragma unroll
    for (int j = 1; j <= 10; ++j) {
      xs[j] = B;  // unused slots repeat the upper bound
      if (j <= nprev) xs[j] = fmin(fmax(prev[j - 1], -B), B);
    }
#pragma unroll
    for (int j = 0; j < 12; ++j) fs[j] = c[dg * SS];
    for (int i = dg - 1; i >= 0; --i) {
      const double ci = c[i * SS];
#pragma unroll
      for (int j = 0; j < 12; ++j) fs[j] = fs[j] * xs[j] + ci;
    }
    // scan the brackets in order: exact zeros at bracket ends are roots, sign changes are queued for bisection
    int ncur = 0, nb = 0;
    bool stop = false;
    double lo_last = xs[0], flo_last = fs[0];
#pragma unroll
    for (int j = 0; j < 11; ++j) {
      if (!stop && j <= nprev) {
        const double lo = xs[j], flo = fs[j], hi = xs[j + 1], fhi = fs[j + 1];
        if (flo == 0.0) {
          if (ncur == 0 || cur[ncur - 1] != lo) cur[ncur++] = lo;
        } else if ((flo < 0.0) != (fhi < 0.0) && fhi != 0.0) {
          ba[nb] = lo, bb[nb] = hi, bfa[nb] = flo, bpos[nb] = ncur;
          ++nb;
          cur[ncur++] = nan("");  // filled below; a bisected root never equals a later bracket end with value zero
        }
        lo_last = hi, flo_last = fhi;
        if (ncur >= dg) stop = true;
      }
    }
    if (flo_last == 0.0 && ncur < dg && (ncur == 0 || cur[ncur - 1] != lo_last)) cur[ncur++] = lo_last;
    // 34 bisection steps per bracket (2^-34 of the bracket is plenty for the critical points, which only serve as
    // brackets of the next level), kRootLanes brackets in flight
    for (int g = 0; g < nb; g += kRootLanes) {
      double a[kRootLanes], b[kRootLanes], fa[kRootLanes];
      bool live[kRootLanes];
#pragma unroll
      for (int u = 0; u < kRootLanes; ++u) {
        const int q = g + u < nb ? g + u : g;
        a[u] = ba[q], b[u] = bb[q], fa[u] = bfa[q];
        live[u] = g + u < nb;
      }
      for (int itr = 0; itr < 34; ++itr) {
        double m[kRootLanes], fm[kRootLanes];
        bool any = false;
#pragma unroll
        for (int u = 0; u < kRootLanes; ++u) {
          m[u] = 0.5 * (a[u] + b[u]);
          if (m[u] == a[u] || m[u] == b[u]) live[u] = false;
          any |= live[u];
          fm[u] = c[dg * SS];
        }
        if (!any) break;
        for (int i = dg - 1; i >= 0; --i) {
          const double ci = c[i * SS];
#pragma unroll
          for (int u = 0; u < kRootLanes; ++u) fm[u] = fm[u] * m[u] + ci;
        }
#pragma unroll
        for (int u = 0; u < kRootLanes; ++u) {
          if (!live[u]) continue;
          if (fm[u] == 0.0) {
            a[u] = b[u] = m[u];
            live[u] = false;
          } else if ((fm[u] < 0.0) == (fa[u] < 0.0)) {
            a[u] = m[u], fa[u] = fm[u];
          } else {
            b[u] = m[u];
          }
        }
      }
#pragma unroll
      for (int u = 0; u < kRootLanes; ++u)
        if (g + u < nb) ba[g + u] = a[u], bb[g + u] = b[u];
    }
    // the roots of the polynomial itself (k == 0) are polished by Newton steps inside the bracket down to rounding
    for (int q = 0; q < nb; ++q) {
      double a = ba[q], b = bb[q];
      if (k == 0 && a != b) {
        double x = 0.5 * (a + b);
        for (int itr = 0; itr < 6; ++itr) {
          double f = c[dg * SS], df = 0.0;
          for (int i = dg - 1; i >= 0; --i) {
            df = df * x + f;
            f = f * x + c[i * SS];
          }
          const double xn = x - f / df;
          if (!(xn >= a && xn <= b)) break;  // also catches df == 0
          const bool done = fabs(xn - x) <= 4.5e-16 * fabs(xn);
          x = xn;
          if (done) break;
        }
        a = b = x;
      }
      cur[bpos[q]] = 0.5 * (a + b);
    }
    nprev = ncur;
    for (int j = 0; j < ncur; ++j) prev[j] = cur[j];
  }
#undef D_
  for (int j = 0; j < nprev; ++j) roots[j] = prev[j];
  return nprev;
}

// x1, x2: the five normalised correspondences.  Writes up to 10 unit-norm essential matrices (row major).
template <int SS>
__device__ int five_point_solve(const double (&x1)[5][2], const double (&x2)[5][2], double* __restrict__ models,
                                double* __restrict__ sm) {
#define Q_(i, j) sm[((i) * 9 + (j)) * SS]
#define M_(i, j) sm[((i) * 20 + (j)) * SS]
#define EET_(i, j, m) sm[(200 + ((i) * 3 + (j)) * 10 + (m)) * SS]
#define EB_(k, j) sm[(290 + (k) * 9 + (j)) * SS]
  // ---- null space of the 5 x 9 epipolar system by Gauss-Jordan with full pivoting, then Gram-Schmidt
  for (int i = 0; i < 5; ++i) {
    const double a = x1[i][0], b = x1[i][1], c = x2[i][0], d = x2[i][1];
    Q_(i, 0) = c * a, Q_(i, 1) = c * b, Q_(i, 2) = c;
    Q_(i, 3) = d * a, Q_(i, 4) = d * b, Q_(i, 5) = d;
    Q_(i, 6) = a, Q_(i, 7) = b, Q_(i, 8) = 1.0;
  }
  int pc[5];
  unsigned pivots = 0;  // bit j: column j holds a pivot
  for (int r = 0; r < 5; ++r) {
    int br = r, bc = -1;
    double bv = 0;
    for (int i = r; i < 5; ++i)
      for (int j = 0; j < 9; ++j) {
        const double v = fabs(Q_(i, j));
        if (!((pivots >> j) & 1u) && v > bv) bv = v, br = i, bc = j;
      }
    if (bc < 0 || bv == 0.0) return 0;  // rank-deficient sample
    for (int j = 0; j < 9; ++j) {
      const double t = Q_(r, j);
      Q_(r, j) = Q_(br, j);
      Q_(br, j) = t;
    }
    pc[r] = bc;
    pivots |= 1u << bc;
    const double inv = 1.0 / Q_(r, bc);
    for (int j = 0; j < 9; ++j) Q_(r, j) *= inv;
    for (int i = 0; i < 5; ++i) {
      if (i == r) continue;
      const double f = Q_(i, bc);
      if (f != 0.0)
        for (int j = 0; j < 9; ++j) Q_(i, j) -= f * Q_(r, j);
    }
  }
  {
    int k = 0;
    for (int f = 0; f < 9; ++f) {
      if ((pivots >> f) & 1u) continue;
      for (int j = 0; j < 9; ++j) EB_(k, j) = 0.0;
      EB_(k, f) = 1.0;
      for (int r = 0; r < 5; ++r) EB_(k, pc[r]) = -Q_(r, f);
      ++k;
    }
    for (int a = 0; a < 4; ++a) {  // modified Gram-Schmidt
      for (int b = 0; b < a; ++b) {
        double dot = 0;
        for (int j = 0; j < 9; ++j) dot += EB_(a, j) * EB_(b, j);
        for (int j = 0; j < 9; ++j) EB_(a, j) -= dot * EB_(b, j);
      }
      double nn = 0;
      for (int j = 0; j < 9; ++j) nn += EB_(a, j) * EB_(a, j);
      const double inv = 1.0 / sqrt(nn);
      for (int j = 0; j < 9; ++j) EB_(a, j) *= inv;
    }
  }
  // ---- the ten cubic constraints on E = x E1 + y E2 + z E3 + E4: entry (r, c) is the linear form EB_(., 3r + c)
#define LIN_(r, c, k) EB_(k, 3 * (r) + (c))
  for (int i = 0; i < 10; ++i)
    for (int j = 0; j < 20; ++j) M_(i, j) = 0.0;
  for (int i = 0; i < 3; ++i)  // EEt (symmetric, quadratic forms)
    for (int j = i; j < 3; ++j) {
      double acc[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
      for (int k = 0; k < 3; ++k)
#pragma unroll
        for (int a = 0; a < 4; ++a) {
          const double la = LIN_(i, k, a);
#pragma unroll
          for (int b = 0; b < 4; ++b) acc[c_ll[a][b]] += la * LIN_(j, k, b);
        }
#pragma unroll
      for (int m = 0; m < 10; ++m) EET_(i, j, m) = EET_(j, i, m) = acc[m];
    }
  for (int m = 0; m < 10; ++m) {
    const double htr = 0.5 * (EET_(0, 0, m) + EET_(1, 1, m) + EET_(2, 2, m));
    EET_(0, 0, m) -= htr, EET_(1, 1, m) -= htr, EET_(2, 2, m) -= htr;
  }
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) {
      double row[20];
#pragma unroll
      for (int c = 0; c < 20; ++c) row[c] = 0.0;
      for (int k = 0; k < 3; ++k) {
        double lb[4];
#pragma unroll
        for (int b = 0; b < 4; ++b) lb[b] = LIN_(k, j, b);
#pragma unroll
        for (int m = 0; m < 10; ++m) {
          const double qv = EET_(i, k, m);
#pragma unroll
          for (int b = 0; b < 4; ++b) row[c_ql[m][b]] += qv * lb[b];
        }
      }
#pragma unroll
      for (int c = 0; c < 20; ++c) M_(1 + 3 * i + j, c) = row[c];
    }
  {  // det(E): e00 (e11 e22 - e12 e21) - e01 (e10 e22 - e12 e20) + e02 (e10 e21 - e11 e20)
    double row[20];
#pragma unroll
    for (int c = 0; c < 20; ++c) row[c] = 0.0;
    for (int t = 0; t < 3; ++t) {
      const int c0 = t == 0 ? 1 : 0, c1 = t == 2 ? 1 : 2;  // the two columns of the minor of e0t
      double mq[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
      for (int a = 0; a < 4; ++a) {
        const double p = LIN_(1, c0, a), q = LIN_(1, c1, a);
#pragma unroll
        for (int b = 0; b < 4; ++b) mq[c_ll[a][b]] += p * LIN_(2, c1, b) - q * LIN_(2, c0, b);
      }
      const double sgn = t == 1 ? -1.0 : 1.0;
      double lb[4];
#pragma unroll
      for (int b = 0; b < 4; ++b) lb[b] = sgn * LIN_(0, t, b);
#pragma unroll
      for (int m = 0; m < 10; ++m)
#pragma unroll
        for (int b = 0; b < 4; ++b) row[c_ql[m][b]] += mq[m] * lb[b];
    }
#pragma unroll
    for (int c = 0; c < 20; ++c) M_(0, c) = row[c];
  }
  // ---- Gauss-Jordan on the ten leading monomials (partial pivoting)
  for (int c = 0; c < 10; ++c) {
    int br = c;
    double bv = fabs(M_(c, c));
    for (int i = c + 1; i < 10; ++i) {
      const double v = fabs(M_(i, c));
      if (v > bv) bv = v, br = i;
    }
    if (bv == 0.0) return 0;
    if (br != c)
      for (int j = c; j < 20; ++j) {
        const double t = M_(c, j);
        M_(c, j) = M_(br, j);
        M_(br, j) = t;
      }
    const double inv = 1.0 / M_(c, c);
    for (int j = c; j < 20; ++j) M_(c, j) *= inv;
    for (int i = 0; i < 10; ++i) {
      if (i == c) continue;
      const double f = M_(i, c);
      if (f != 0.0)
        for (int j = c; j < 20; ++j) M_(i, j) -= f * M_(c, j);
    }
  }
  // ---- B(z) [x y 1]^T = 0: row pair (x2z, x2), (y2z, y2), (xyz, xy): upper row - z * lower row.  Tail columns:
  // 10 xz2, 11 xz, 12 x, 13 yz2, 14 yz, 15 y, 16 z3, 17 z2, 18 z, 19 1
  double bx[3][4], by[3][4], b1[3][5];
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    double e[10], f[10];
#pragma unroll
    for (int c = 0; c < 10; ++c) e[c] = M_(4 + 2 * r, 10 + c), f[c] = M_(5 + 2 * r, 10 + c);
    bx[r][0] = e[2], bx[r][1] = e[1] - f[2], bx[r][2] = e[0] - f[1], bx[r][3] = -f[0];
    by[r][0] = e[5], by[r][1] = e[4] - f[5], by[r][2] = e[3] - f[4], by[r][3] = -f[3];
    b1[r][0] = e[9], b1[r][1] = e[8] - f[9], b1[r][2] = e[7] - f[8], b1[r][3] = e[6] - f[7], b1[r][4] = -f[6];
  }
  // det B(z) = sum over rows r of b1[r] * cofactor_r, cofactor_r = +-(bx[s] by[t] - bx[t] by[s])
  double det[11];
#pragma unroll
  for (int i = 0; i < 11; ++i) det[i] = 0.0;
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    const int s = r == 0 ? 1 : 0, t = r == 2 ? 1 : 2;
    double cof[7] = {0, 0, 0, 0, 0, 0, 0};
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
      for (int b = 0; b < 4; ++b) cof[a + b] += bx[s][a] * by[t][b] - bx[t][a] * by[s][b];
    const double sgn = r == 1 ? -1.0 : 1.0;
#pragma unroll
    for (int a = 0; a < 5; ++a)
#pragma unroll
      for (int b = 0; b < 7; ++b) det[a + b] += sgn * b1[r][a] * cof[b];
  }
  {
    double mx = 0;
#pragma unroll
    for (int i = 0; i < 11; ++i) mx = fmax(mx, fabs(det[i]));
    if (!(mx > 0.0) || !isfinite(mx)) return 0;
    const double inv = 1.0 / mx;
#pragma unroll
    for (int i = 0; i < 11; ++i) det[i] *= inv;
  }
  double roots[10];
  const int nroots = real_roots<SS>(det, 10, roots, sm);  // the derivative table reuses M's storage
  int nm = 0;
  for (int k = 0; k < nroots && nm < kPoseMaxModels; ++k) {
    const double z = roots[k];
    double Bz[3][3];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      Bz[r][0] = horner(bx[r], 3, z);
      Bz[r][1] = horner(by[r], 3, z);
      Bz[r][2] = horner(b1[r], 4, z);
    }
    double best[3] = {0, 0, 0}, bestw = -1.0;
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
      for (int b = a + 1; b < 3; ++b) {
        const double v0 = Bz[a][1] * Bz[b][2] - Bz[a][2] * Bz[b][1];
        const double v1 = Bz[a][2] * Bz[b][0] - Bz[a][0] * Bz[b][2];
        const double v2 = Bz[a][0] * Bz[b][1] - Bz[a][1] * Bz[b][0];
        if (fabs(v2) > bestw) bestw = fabs(v2), best[0] = v0, best[1] = v1, best[2] = v2;
      }
    const double vn = sqrt(best[0] * best[0] + best[1] * best[1] + best[2] * best[2]);
    if (!(bestw > 1e-10 * vn)) continue;  // [x y 1] at infinity
    const double x = best[0] / best[2], y = best[1] / best[2];
    double E[9], nn = 0;
#pragma unroll
    for (int j = 0; j < 9; ++j) {
      E[j] = x * EB_(0, j) + y * EB_(1, j) + z * EB_(2, j) + EB_(3, j);
      nn += E[j] * E[j];
    }
    if (!(nn > 0.0) || !isfinite(nn)) continue;
    const double inv = 1.0 / sqrt(nn);
#pragma unroll
    for (int j = 0; j < 9; ++j) models[nm * 9 + j] = E[j] * inv;
    ++nm;
  }
  return nm;
#undef LIN_
#undef EB_
#undef EET_
#undef M_
#undef Q_
}

// ORBX_HOST_TESTABLE_END
__global__ void __launch_bounds__(32) pose_solve_kernel(const PoseState* __restrict__ st, int nproblems, int wave, PoseParams pp,
                                                       const double* __restrict__ x1, const double* __restrict__ x2,
                                                       const int* __restrict__ samples, double* __restrict__ models,
                                                       int* __restrict__ nmodels) {
  extern __shared__ double s_solve[];  // [kSolveScratch][32]
  // threads are laid over (pair, hypothesis of this wave): a warp holds 32 / len pairs in the short first wave
  const int len = c_wave[wave + 1] - c_wave[wave];
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  const int p = g / len;
  if (p >= nproblems) return;
  const int it = c_wave[wave] + g % len;
  if (it >= st[p].niters) return;
  const int* idx = samples + ((size_t)p * kPoseMaxIters + it) * 5;
  const size_t base = (size_t)p * pp.max_rows;
  double a[5][2], b[5][2];
  for (int i = 0; i < 5; ++i) {
    const size_t j = base + idx[i];
    a[i][0] = x1[2 * j], a[i][1] = x1[2 * j + 1];
    b[i][0] = x2[2 * j], b[i][1] = x2[2 * j + 1];
  }
  const size_t h = (size_t)p * kPoseMaxIters + it;
  nmodels[h] = five_point_solve<32>(a, b, models + h * (kPoseMaxModels * 9), s_solve + threadIdx.x);
}

// Sampson error of EMEstimatorCallback::computeError, unfused, rounded to fp32
__device__ __forceinline__ float sampson_f32(const double* E, double ax, double ay, double bx, double by) {
  const double e0 = __dadd_rn(__dadd_rn(__dmul_rn(E[0], ax), __dmul_rn(E[1], ay)), E[2]);
  const double e1 = __dadd_rn(__dadd_rn(__dmul_rn(E[3], ax), __dmul_rn(E[4], ay)), E[5]);
  const double e2 = __dadd_rn(__dadd_rn(__dmul_rn(E[6], ax), __dmul_rn(E[7], ay)), E[8]);
  const double t0 = __dadd_rn(__dadd_rn(__dmul_rn(E[0], bx), __dmul_rn(E[3], by)), E[6]);
  const double t1 = __dadd_rn(__dadd_rn(__dmul_rn(E[1], bx), __dmul_rn(E[4], by)), E[7]);
  const double num = __dadd_rn(__dadd_rn(__dmul_rn(bx, e0), __dmul_rn(by, e1)), e2);
  const double den = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(e0, e0), __dmul_rn(e1, e1)), __dmul_rn(t0, t0)), __dmul_rn(t1, t1));
  return (float)__ddiv_rn(__dmul_rn(num, num), den);
}

// sampson_f32(...) <= thr without the f64 division except next to the threshold
__device__ __forceinline__ bool sampson_inlier(const double* E, double ax, double ay, double bx, double by, float thr,
                                               double bound) {
  const double e0 = __dadd_rn(__dadd_rn(__dmul_rn(E[0], ax), __dmul_rn(E[1], ay)), E[2]);
  const double e1 = __dadd_rn(__dadd_rn(__dmul_rn(E[3], ax), __dmul_rn(E[4], ay)), E[5]);
  const double e2 = __dadd_rn(__dadd_rn(__dmul_rn(E[6], ax), __dmul_rn(E[7], ay)), E[8]);
  const double t0 = __dadd_rn(__dadd_rn(__dmul_rn(E[0], bx), __dmul_rn(E[3], by)), E[6]);
  const double t1 = __dadd_rn(__dadd_rn(__dmul_rn(E[1], bx), __dmul_rn(E[4], by)), E[7]);
  const double num = __dadd_rn(__dadd_rn(__dmul_rn(bx, e0), __dmul_rn(by, e1)), e2);
  const double den = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(e0, e0), __dmul_rn(e1, e1)), __dmul_rn(t0, t0)), __dmul_rn(t1, t1));
  const double n2 = __dmul_rn(num, num), lim = bound * den;
  if (n2 < lim * (1.0 - 1e-9)) return true;
  if (n2 > lim * (1.0 + 1e-9) || !(den > 0.0)) return den > 0.0 ? false : (float)__ddiv_rn(n2, den) <= thr;
  return (float)__ddiv_rn(n2, den) <= thr;
}

// blocks loop over the hypotheses of the wave: inlier count of each of a hypothesis' models over all points
__global__ void __launch_bounds__(128, 8) pose_count_kernel(const PoseState* __restrict__ st, int wave, PoseParams pp,
                                                        const double* __restrict__ x1, const double* __restrict__ x2,
                                                        const double* __restrict__ models, const int* __restrict__ nmodels,
                                                        int* __restrict__ counts) {
  __shared__ double s_E[kPoseMaxModels * 9];
  __shared__ int s_cnt[kPoseMaxModels];
  const int p = blockIdx.y;
  const int n = st[p].n;
  const int it_end = min(c_wave[wave + 1], st[p].niters);
  const float thr = (float)pp.thr2f;
  // (float)(num^2 / den) <= thr  <=>  num^2 / den < thr + half an fp32 ulp (thr's mantissa is odd or even: the tie
  // case is decided by the exact path); the division is only needed within 1e-9 of that bound
  const double bound = (double)thr + 0.5 * ((double)nextafterf(thr, 2.0f * thr) - (double)thr);
  for (int it = c_wave[wave] + blockIdx.x; it < it_end; it += gridDim.x) {
  const size_t h = (size_t)p * kPoseMaxIters + it;
  const int nm = nmodels[h];
  if (nm == 0) continue;
  __syncthreads();
  if (threadIdx.x < nm * 9) s_E[threadIdx.x] = models[h * (kPoseMaxModels * 9) + threadIdx.x];
  if (threadIdx.x < kPoseMaxModels) s_cnt[threadIdx.x] = 0;
  __syncthreads();
  const size_t base = (size_t)p * pp.max_rows;
  for (int m = 0; m < nm; ++m) {
    const double* E = s_E + 9 * m;
    int v = 0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      const double2 a = reinterpret_cast<const double2*>(x1)[base + i];
      const double2 b = reinterpret_cast<const double2*>(x2)[base + i];
      v += sampson_inlier(E, a.x, a.y, b.x, b.y, thr, bound);
    }
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
    if ((threadIdx.x & 31) == 0 && v) atomicAdd(&s_cnt[m], v);
  }
  __syncthreads();
  if (threadIdx.x < nm) counts[h * kPoseMaxModels + threadIdx.x] = s_cnt[threadIdx.x];
  }
}

// RANSACUpdateNumIters
__device__ int update_num_iters(double p, double ep, int model_points, int max_iters) {
  p = fmin(fmax(p, 0.0), 1.0);
  ep = fmin(fmax(ep, 0.0), 1.0);
  double num = fmax(1.0 - p, 2.2250738585072014e-308);
  double denom = 1.0 - pow(1.0 - ep, (double)model_points);
  if (denom < 2.2250738585072014e-308) return 0;
  num = log(num);
  denom = log(denom);
  return denom >= 0 || -num >= max_iters * (-denom) ? max_iters : (int)rint(num / denom);
}

// the sequential part of RANSACPointSetRegistrator::run over the hypotheses of one wave
__global__ void pose_scan_kernel(PoseState* __restrict__ st, int wave, PoseParams pp, const int* __restrict__ nmodels,
                                 const int* __restrict__ counts) {
  PoseState& s = st[blockIdx.x];
  if (threadIdx.x != 0) return;
  const int w1 = c_wave[wave + 1];
  int it = c_wave[wave];
  for (; it < w1 && it < s.niters; ++it) {
    const size_t h = (size_t)blockIdx.x * kPoseMaxIters + it;
    const int nm = nmodels[h];
    for (int m = 0; m < nm; ++m) {
      const int g = counts[h * kPoseMaxModels + m];
      if (g > max(s.max_good, 4)) {
        s.max_good = g;
        s.best_it = it;
        s.best_m = m;
        s.niters = update_num_iters(pp.prob, (double)(s.n - g) / s.n, 5, s.niters);
      }
    }
  }
  if (it > s.iters_run) s.iters_run = it;
}

// E and the inlier mask of the winning model
__global__ void pose_finish_kernel(PoseState* __restrict__ st, PoseParams pp, const double* __restrict__ x1,
                                   const double* __restrict__ x2, const double* __restrict__ models,
                                   double* __restrict__ E_out, uint8_t* __restrict__ mask, int* __restrict__ n_inliers,
                                   int* __restrict__ iters_out) {
  const int p = blockIdx.y;
  PoseState& s = st[p];
  const bool have = s.best_it >= 0;
  const double* E = models + (((size_t)p * kPoseMaxIters + max(s.best_it, 0)) * kPoseMaxModels + max(s.best_m, 0)) * 9;
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    for (int j = 0; j < 9; ++j) E_out[9 * p + j] = have ? E[j] : 0.0;
    if (n_inliers) n_inliers[p] = have ? s.max_good : 0;
    if (iters_out) iters_out[p] = s.iters_run;
    if (s.status == ORBX_POSE_OK && !have) s.status = ORBX_POSE_NO_MODEL;
  }
  if (!mask) return;
  const float thr = (float)pp.thr2f;
  const size_t base = (size_t)p * pp.max_rows;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < s.n; i += gridDim.x * blockDim.x)
    mask[base + i] = have && sampson_f32(E, x1[2 * (base + i)], x1[2 * (base + i) + 1], x2[2 * (base + i)],
                                         x2[2 * (base + i) + 1]) <= thr;
}

// ------------------------------------------------------------------------------------------------ recoverPose
// cv::decomposeEssentialMat: E = U S V^T, R1 = U W V^T, R2 = U W^T V^T, t = u3 (det U = det V = +1)
__global__ void pose_decompose_kernel(PoseState* __restrict__ st, const double* __restrict__ E_in, int nproblems) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= nproblems) return;
  double a[3][3], v[3][3];
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) a[r][c] = E_in[9 * p + 3 * r + c];
  jacobi_onesided<3, 3>(a, v);
  double nn[3];
  for (int c = 0; c < 3; ++c) nn[c] = a[0][c] * a[0][c] + a[1][c] * a[1][c] + a[2][c] * a[2][c];
  int o0 = 0, o1 = 1, o2 = 2;  // descending singular values
  if (nn[o0] < nn[o1]) { const int t = o0; o0 = o1; o1 = t; }
  if (nn[o1] < nn[o2]) { const int t = o1; o1 = o2; o2 = t; }
  if (nn[o0] < nn[o1]) { const int t = o0; o0 = o1; o1 = t; }
  double U[3][3], V[3][3];
  const double i0 = nn[o0] > 0 ? 1.0 / sqrt(nn[o0]) : 0.0, i1 = nn[o1] > 0 ? 1.0 / sqrt(nn[o1]) : 0.0;
  for (int r = 0; r < 3; ++r) {
    U[r][0] = (o0 == 0 ? a[r][0] : o0 == 1 ? a[r][1] : a[r][2]) * i0;
    U[r][1] = (o1 == 0 ? a[r][0] : o1 == 1 ? a[r][1] : a[r][2]) * i1;
    V[r][0] = o0 == 0 ? v[r][0] : o0 == 1 ? v[r][1] : v[r][2];
    V[r][1] = o1 == 0 ? v[r][0] : o1 == 1 ? v[r][1] : v[r][2];
  }
  // third columns complete right-handed bases (the third singular value of an essential matrix is 0)
  U[0][2] = U[1][0] * U[2][1] - U[2][0] * U[1][1];
  U[1][2] = U[2][0] * U[0][1] - U[0][0] * U[2][1];
  U[2][2] = U[0][0] * U[1][1] - U[1][0] * U[0][1];
  V[0][2] = V[1][0] * V[2][1] - V[2][0] * V[1][1];
  V[1][2] = V[2][0] * V[0][1] - V[0][0] * V[2][1];
  V[2][2] = V[0][0] * V[1][1] - V[1][0] * V[0][1];
  // U W = [-u2 u1 u3], U W^T = [u2 -u1 u3]
  PoseState& s = st[p];
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) {
      s.R1[3 * r + c] = -U[r][1] * V[c][0] + U[r][0] * V[c][1] + U[r][2] * V[c][2];
      s.R2[3 * r + c] = U[r][1] * V[c][0] - U[r][0] * V[c][1] + U[r][2] * V[c][2];
    }
  for (int r = 0; r < 3; ++r) s.t[r] = U[r][2];
}

// per point: triangulate against each of the four candidates (cv::triangulatePoints on f64 points) and apply the
// depth tests of cv::recoverPose (distanceThresh = 50); bit c of bits[i] = candidate c keeps the point
__global__ void __launch_bounds__(128) pose_cheirality_kernel(PoseState* __restrict__ st, PoseParams pp,
                                                             const double* __restrict__ x1, const double* __restrict__ x2,
                                                             uint8_t* __restrict__ bits) {
  __shared__ int s_cnt[4];
  const int p = blockIdx.y;
  PoseState& s = st[p];
  if (threadIdx.x < 4) s_cnt[threadIdx.x] = 0;
  __syncthreads();
  const size_t base = (size_t)p * pp.max_rows;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  unsigned keep = 0;
  if (i < s.n) {
    const double ax = x1[2 * (base + i)], ay = x1[2 * (base + i) + 1];
    const double bx = x2[2 * (base + i)], by = x2[2 * (base + i) + 1];
    // Candidates c and c + 2 share R and differ in the sign of t: their DLT systems differ by the sign of column 3,
    // every step of the Jacobi iteration is sign-symmetric in IEEE arithmetic, so the null vector of the second is
    // (q0, q1, q2, -q3) bit for bit - the point and its depth in the second view are the exact negatives.  Two SVDs
    // per point instead of four, same results.
    for (int c = 0; c < 2; ++c) {
      const double* R = c ? s.R2 : s.R1;
      double P[3][4];
      for (int r = 0; r < 3; ++r) {
        P[r][0] = R[3 * r], P[r][1] = R[3 * r + 1], P[r][2] = R[3 * r + 2];
        P[r][3] = s.t[r];
      }
      double a[4][4];
      // first view: P0 = [I | 0]
      a[0][0] = -1.0, a[0][1] = 0.0, a[0][2] = ax, a[0][3] = 0.0;
      a[1][0] = 0.0, a[1][1] = -1.0, a[1][2] = ay, a[1][3] = 0.0;
      for (int k = 0; k < 4; ++k) {
        a[2][k] = __dsub_rn(__dmul_rn(bx, P[2][k]), P[0][k]);
        a[3][k] = __dsub_rn(__dmul_rn(by, P[2][k]), P[1][k]);
      }
      double q[4];
      null_vector4(a, q);
      const double w = q[2] * q[3];
      const double X = q[0] / q[3], Y = q[1] / q[3], Z = q[2] / q[3];
      const double z2 = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(P[2][0], X), __dmul_rn(P[2][1], Y)), __dmul_rn(P[2][2], Z)), P[2][3]);
      const bool ok_pos = w > 0.0 && Z < 50.0 && z2 > 0.0 && z2 < 50.0;          // (R, t)
      const bool ok_neg = w < 0.0 && -Z < 50.0 && -z2 > 0.0 && -z2 < 50.0;       // (R, -t)
      keep |= (ok_pos ? 1u << c : 0u) | (ok_neg ? 1u << (c + 2) : 0u);
    }
    bits[base + i] = (uint8_t)keep;
  }
  for (int c = 0; c < 4; ++c) {
    const unsigned bal = __ballot_sync(0xFFFFFFFFu, (keep >> c) & 1u);
    if ((threadIdx.x & 31) == 0 && bal) atomicAdd(&s_cnt[c], __popc(bal));
  }
  __syncthreads();
  if (threadIdx.x < 4 && s_cnt[threadIdx.x]) atomicAdd(&s.cand[threadIdx.x], s_cnt[threadIdx.x]);
}

// the first maximum in OpenCV's order (R1, t), (R2, t), (R1, -t), (R2, -t); pose.rs:104: fewer than min_matches
// points in front of the cameras is the reference's "Too few inliers" error
__global__ void pose_select_kernel(PoseState* __restrict__ st, PoseParams pp, int nproblems, double* __restrict__ R_out,
                                   double* __restrict__ t_out, int* __restrict__ n_good, int* __restrict__ status,
                                   bool have_ransac) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= nproblems) return;
  PoseState& s = st[p];
  const int g0 = s.cand[0], g1 = s.cand[1], g2 = s.cand[2], g3 = s.cand[3];
  int k;
  if (g0 >= g1 && g0 >= g2 && g0 >= g3)
    k = 0;
  else if (g1 >= g0 && g1 >= g2 && g1 >= g3)
    k = 1;
  else if (g2 >= g0 && g2 >= g1 && g2 >= g3)
    k = 2;
  else
    k = 3;
  s.choice = k;
  const double* R = (k & 1) ? s.R2 : s.R1;
  const double sg = k >= 2 ? -1.0 : 1.0;
  for (int j = 0; j < 9; ++j) R_out[9 * p + j] = R[j];
  for (int j = 0; j < 3; ++j) t_out[3 * p + j] = sg * s.t[j];
  const int good = k == 0 ? g0 : k == 1 ? g1 : k == 2 ? g2 : g3;
  if (n_good) n_good[p] = good;
  int stt = have_ransac ? s.status : (s.n >= pp.min_matches ? ORBX_POSE_OK : ORBX_POSE_INSUFFICIENT_POINTS);
  if (stt == ORBX_POSE_OK && good < pp.min_matches) stt = ORBX_POSE_TOO_FEW_INLIERS;
  s.status = stt;
  if (status) status[p] = stt;
}

__global__ void pose_mask_kernel(const PoseState* __restrict__ st, PoseParams pp, const uint8_t* __restrict__ bits,
                                 uint8_t* __restrict__ mask) {
  const int p = blockIdx.y;
  const size_t base = (size_t)p * pp.max_rows;
  const int k = st[p].choice;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < st[p].n; i += gridDim.x * blockDim.x)
    mask[base + i] = (bits[base + i] >> k) & 1;
}

__global__ void pose_status_kernel(const PoseState* __restrict__ st, int nproblems, int* __restrict__ status) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p < nproblems) status[p] = st[p].status;
}

template <class T>
static int pose_alloc(T** p, size_t n) {
  cudaError_t e = cudaMalloc((void**)p, (n ? n : 1) * sizeof(T));
  if (e != cudaSuccess) {
    set_error("orbx_pose_estimator: cudaMalloc of %zu bytes failed: %s", n * sizeof(T), cudaGetErrorString(e));
    return ORBX_E_NOMEM;
  }
  return ORBX_OK;
}

}  // namespace orbx

using namespace orbx;

struct orbx_pose_estimator {
  int device = 0, max_points = 0, max_problems = 0;
  PoseParams pp{};
  cudaStream_t own_stream = nullptr, stream = nullptr;
  // work buffers sized max_problems x max_points (points) / max_problems x 1000 (hypotheses)
  double *x1 = nullptr, *x2 = nullptr;
  int *samples = nullptr, *nmodels = nullptr, *counts = nullptr;
  double* models = nullptr;
  PoseState* state = nullptr;
  uint8_t* bits = nullptr;
  // staging for the single-problem host API
  float *h_pts1 = nullptr, *h_pts2 = nullptr;
  double *d_E = nullptr, *d_R = nullptr, *d_t = nullptr;
  uint8_t* d_mask = nullptr;
  int* d_ints = nullptr;  // n_inliers, iters, n_good, status
};

namespace {

// RANSAC for `np` problems laid out with row capacity pp.max_rows; everything on pe->stream
int run_essential(orbx_pose_estimator* pe, const float* pts1, const float* pts2, const int* n_dev, int n_fixed, int np,
                  double* E_out, uint8_t* mask, int* n_inliers, int* iters) {
  cudaStream_t s = pe->stream;
  const PoseParams& pp = pe->pp;
  ORBX_CUDA(cudaMemsetAsync(pe->nmodels, 0, (size_t)np * kPoseMaxIters * sizeof(int), s));
  pose_prepare_kernel<<<dim3(std::max(1, std::min(div_up(pp.max_rows, 256), 64)), np), 256, 0, s>>>(
      pts1, pts2, n_dev, n_fixed, pp, pe->x1, pe->x2, pe->state, true);
  ORBX_LAUNCH_CHECK();
  const int wave[kPoseWaves + 1] = {0, 8, 32, 96, 256, kPoseMaxIters};
  constexpr int kSolveSmem = kSolveScratch * 32 * (int)sizeof(double);
  ORBX_CUDA(cudaFuncSetAttribute(pose_solve_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSolveSmem));
  for (int w = 0; w < kPoseWaves; ++w) {
    const int len = wave[w + 1] - wave[w];
    pose_sample_kernel<<<np, 32, 0, s>>>(pe->state, w, pe->samples);
    ORBX_LAUNCH_CHECK();
    pose_solve_kernel<<<div_up(len * np, 32), 32, kSolveSmem, s>>>(pe->state, np, w, pp, pe->x1, pe->x2, pe->samples, pe->models,
                                                                   pe->nmodels);
    ORBX_LAUNCH_CHECK();
    pose_count_kernel<<<dim3(std::min(len, 32), np), 128, 0, s>>>(pe->state, w, pp, pe->x1, pe->x2, pe->models, pe->nmodels, pe->counts);
    ORBX_LAUNCH_CHECK();
    pose_scan_kernel<<<np, 32, 0, s>>>(pe->state, w, pp, pe->nmodels, pe->counts);
    ORBX_LAUNCH_CHECK();
  }
  pose_finish_kernel<<<dim3(std::max(1, std::min(div_up(pp.max_rows, 256), 64)), np), 256, 0, s>>>(
      pe->state, pp, pe->x1, pe->x2, pe->models, E_out, mask, n_inliers, iters);
  ORBX_LAUNCH_CHECK();
  return ORBX_OK;
}

// recoverPose for `np` problems; prepare == true normalises the points first (stand-alone call)
int run_recover(orbx_pose_estimator* pe, const double* E, const float* pts1, const float* pts2, const int* n_dev,
                int n_fixed, int np, bool prepare, double* R, double* t, uint8_t* mask, int* n_good, int* status) {
  cudaStream_t s = pe->stream;
  const PoseParams& pp = pe->pp;
  const dim3 gpts(std::max(1, std::min(div_up(pp.max_rows, 256), 64)), np);
  pose_prepare_kernel<<<gpts, 256, 0, s>>>(pts1, pts2, n_dev, n_fixed, pp, pe->x1, pe->x2, pe->state, prepare);
  ORBX_LAUNCH_CHECK();
  pose_decompose_kernel<<<div_up(np, 64), 64, 0, s>>>(pe->state, E, np);
  ORBX_LAUNCH_CHECK();
  pose_cheirality_kernel<<<dim3(div_up(pp.max_rows, 128), np), 128, 0, s>>>(pe->state, pp, pe->x1, pe->x2, pe->bits);
  ORBX_LAUNCH_CHECK();
  pose_select_kernel<<<div_up(np, 64), 64, 0, s>>>(pe->state, pp, np, R, t, n_good, status, !prepare);
  ORBX_LAUNCH_CHECK();
  if (mask) {
    pose_mask_kernel<<<gpts, 256, 0, s>>>(pe->state, pp, pe->bits, mask);
    ORBX_LAUNCH_CHECK();
  }
  return ORBX_OK;
}

int stage_points(orbx_pose_estimator* pe, const float* pts1, const float* pts2, int n, const float** d1, const float** d2) {
  *d1 = pts1, *d2 = pts2;
  if (!is_device_ptr(pts1)) {
    ORBX_CUDA(cudaMemcpyAsync(pe->h_pts1, pts1, (size_t)n * 8, cudaMemcpyHostToDevice, pe->stream));
    *d1 = pe->h_pts1;
  }
  if (!is_device_ptr(pts2)) {
    ORBX_CUDA(cudaMemcpyAsync(pe->h_pts2, pts2, (size_t)n * 8, cudaMemcpyHostToDevice, pe->stream));
    *d2 = pe->h_pts2;
  }
  return ORBX_OK;
}

}  // namespace

extern "C" {

int orbx_pose_estimator_create(int device, int max_points, int max_problems, double fx, double fy, double cx, double cy,
                               orbx_pose_estimator** out) {
  if (!out || max_points < 8 || max_problems < 1) {
    set_error("orbx_pose_estimator_create: bad argument");
    return ORBX_E_BADARG;
  }
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) {
    cudaGetLastError();
    set_error("orbx_pose_estimator_create: no CUDA device %d (there is no CPU fallback)", device);
    return ORBX_E_CUDA;
  }
  ORBX_CUDA(cudaSetDevice(device));
  auto* pe = new orbx_pose_estimator();
  pe->device = device;
  pe->max_points = max_points;
  pe->max_problems = max_problems;
  pe->pp.fx = fx, pe->pp.fy = fy, pe->pp.cx = cx, pe->pp.cy = cy;
  pe->pp.prob = 0.999;                                  // pose.rs:69
  const double thr = 1.0 / ((fx + fy) / 2.0);           // pose.rs:70 threshold 1.0, normalised like findEssentialMat
  pe->pp.thr2f = (double)(float)(thr * thr);
  pe->pp.min_matches = 8;                               // pose.rs:21
  pe->pp.max_rows = max_points;
  const size_t P = (size_t)max_problems, N = (size_t)max_points;
  int rc = ORBX_OK;
  if (cudaStreamCreateWithFlags(&pe->own_stream, cudaStreamNonBlocking) != cudaSuccess) rc = ORBX_E_CUDA;
  pe->stream = pe->own_stream;
  if (!rc) rc = pose_alloc(&pe->x1, P * N * 2);
  if (!rc) rc = pose_alloc(&pe->x2, P * N * 2);
  if (!rc) rc = pose_alloc(&pe->samples, P * kPoseMaxIters * 5);
  if (!rc) rc = pose_alloc(&pe->nmodels, P * kPoseMaxIters);
  if (!rc) rc = pose_alloc(&pe->counts, P * kPoseMaxIters * kPoseMaxModels);
  if (!rc) rc = pose_alloc(&pe->models, P * kPoseMaxIters * kPoseMaxModels * 9);
  if (!rc) rc = pose_alloc(&pe->state, P);
  if (!rc) rc = pose_alloc(&pe->bits, P * N);
  if (!rc) rc = pose_alloc(&pe->h_pts1, N * 2);
  if (!rc) rc = pose_alloc(&pe->h_pts2, N * 2);
  if (!rc) rc = pose_alloc(&pe->d_E, 9);
  if (!rc) rc = pose_alloc(&pe->d_R, 9);
  if (!rc) rc = pose_alloc(&pe->d_t, 3);
  if (!rc) rc = pose_alloc(&pe->d_mask, N);
  if (!rc) rc = pose_alloc(&pe->d_ints, 4);
  if (rc) {
    orbx_pose_estimator_destroy(pe);
    return rc;
  }
  *out = pe;
  return ORBX_OK;
}

void orbx_pose_estimator_destroy(orbx_pose_estimator* pe) {
  if (!pe) return;
  cudaSetDevice(pe->device);
  if (pe->stream) cudaStreamSynchronize(pe->stream);
  cudaFree(pe->x1), cudaFree(pe->x2), cudaFree(pe->samples), cudaFree(pe->nmodels), cudaFree(pe->counts);
  cudaFree(pe->models), cudaFree(pe->state), cudaFree(pe->bits), cudaFree(pe->h_pts1), cudaFree(pe->h_pts2);
  cudaFree(pe->d_E), cudaFree(pe->d_R), cudaFree(pe->d_t), cudaFree(pe->d_mask), cudaFree(pe->d_ints);
  if (pe->own_stream) cudaStreamDestroy(pe->own_stream);
  delete pe;
}

int orbx_pose_estimator_set_stream(orbx_pose_estimator* pe, void* stream) {
  if (!pe) return ORBX_E_BADARG;
  pe->stream = stream ? (cudaStream_t)stream : pe->own_stream;
  return ORBX_OK;
}

int orbx_pose_estimator_sync(orbx_pose_estimator* pe) {
  if (!pe) return ORBX_E_BADARG;
  ORBX_CUDA(cudaSetDevice(pe->device));
  ORBX_CUDA(cudaStreamSynchronize(pe->stream));
  return ORBX_OK;
}

int orbx_compute_essential_matrix(orbx_pose_estimator* pe, const float* pts1, int n1, const float* pts2, int n2,
                                  double* E, uint8_t* mask, int* n_inliers, int* iterations) {
  if (!pe || n1 < 0 || n2 < 0 || !E) {
    set_error("orbx_compute_essential_matrix: bad argument");
    return ORBX_E_BADARG;
  }
  if (n1 < pe->pp.min_matches || n2 < pe->pp.min_matches) {  // pose.rs:52-59
    set_error("Insufficient points: %d (need %d)", n1, pe->pp.min_matches);
    return ORBX_E_DATA;
  }
  if (n1 != n2 || !pts1 || !pts2) {  // cv::findEssentialMat asserts equal sizes
    set_error("orbx_compute_essential_matrix: point arrays must have the same length (%d vs %d)", n1, n2);
    return ORBX_E_BADARG;
  }
  if (n1 > pe->max_points) {
    set_error("orbx_compute_essential_matrix: %d points exceed max_points %d", n1, pe->max_points);
    return ORBX_E_CAPACITY;
  }
  ORBX_CUDA(cudaSetDevice(pe->device));
  const float *d1, *d2;
  int rc = stage_points(pe, pts1, pts2, n1, &d1, &d2);
  if (rc) return rc;
  if ((rc = run_essential(pe, d1, d2, nullptr, n1, 1, pe->d_E, mask ? pe->d_mask : nullptr, pe->d_ints, pe->d_ints + 1)))
    return rc;
  pose_status_kernel<<<1, 32, 0, pe->stream>>>(pe->state, 1, pe->d_ints + 3);
  ORBX_LAUNCH_CHECK();
  int ints[4] = {0, 0, 0, 0};
  ORBX_CUDA(cudaMemcpyAsync(E, pe->d_E, 72, cudaMemcpyDefault, pe->stream));
  ORBX_CUDA(cudaMemcpyAsync(ints, pe->d_ints, 16, cudaMemcpyDeviceToHost, pe->stream));
  if (mask) ORBX_CUDA(cudaMemcpyAsync(mask, pe->d_mask, (size_t)n1, cudaMemcpyDefault, pe->stream));
  ORBX_CUDA(cudaStreamSynchronize(pe->stream));
  if (n_inliers) *n_inliers = ints[0];
  if (iterations) *iterations = ints[1];
  if (ints[3] == ORBX_POSE_NO_MODEL) {  // pose.rs:75-77
    set_error("Failed to compute essential matrix");
    return ORBX_E_DATA;
  }
  return ORBX_OK;
}

int orbx_recover_pose(orbx_pose_estimator* pe, const double* E, const float* pts1, int n1, const float* pts2, int n2,
                      double* R, double* t, uint8_t* mask, int* n_good) {
  if (!pe || !E || n1 < 0 || n1 != n2 || (n1 > 0 && (!pts1 || !pts2)) || !R || !t) {
    set_error("orbx_recover_pose: bad argument");
    return ORBX_E_BADARG;
  }
  if (n1 > pe->max_points) {
    set_error("orbx_recover_pose: %d points exceed max_points %d", n1, pe->max_points);
    return ORBX_E_CAPACITY;
  }
  ORBX_CUDA(cudaSetDevice(pe->device));
  const float *d1 = pts1, *d2 = pts2;
  int rc = ORBX_OK;
  if (n1 > 0 && (rc = stage_points(pe, pts1, pts2, n1, &d1, &d2))) return rc;
  const double* dE = E;
  if (!is_device_ptr(E)) {
    ORBX_CUDA(cudaMemcpyAsync(pe->d_E, E, 72, cudaMemcpyHostToDevice, pe->stream));
    dE = pe->d_E;
  }
  if ((rc = run_recover(pe, dE, d1, d2, nullptr, n1, 1, true, pe->d_R, pe->d_t, mask ? pe->d_mask : nullptr,
                        pe->d_ints + 2, pe->d_ints + 3)))
    return rc;
  int ints[4] = {0, 0, 0, 0};
  ORBX_CUDA(cudaMemcpyAsync(R, pe->d_R, 72, cudaMemcpyDefault, pe->stream));
  ORBX_CUDA(cudaMemcpyAsync(t, pe->d_t, 24, cudaMemcpyDefault, pe->stream));
  ORBX_CUDA(cudaMemcpyAsync(ints, pe->d_ints, 16, cudaMemcpyDeviceToHost, pe->stream));
  if (mask && n1 > 0) ORBX_CUDA(cudaMemcpyAsync(mask, pe->d_mask, (size_t)n1, cudaMemcpyDefault, pe->stream));
  ORBX_CUDA(cudaStreamSynchronize(pe->stream));
  if (n_good) *n_good = ints[2];
  if (ints[2] < pe->pp.min_matches) {  // pose.rs:104-106
    set_error("Too few inliers: %d", ints[2]);
    return ORBX_E_DATA;
  }
  return ORBX_OK;
}

int orbx_estimate_pose_batch(orbx_pose_estimator* pe, const float* pts1, const float* pts2, const int* n_dev, int npairs,
                             int max_rows, double* E, double* R, double* t, int* n_inliers, int* n_good, int* status,
                             uint8_t* ransac_mask, uint8_t* pose_mask) {
  if (!pe || npairs < 0 || max_rows < 0) {
    set_error("orbx_estimate_pose_batch: bad argument");
    return ORBX_E_BADARG;
  }
  if (npairs == 0) return ORBX_OK;
  if (!is_device_ptr(pts1) || !is_device_ptr(pts2) || !is_device_ptr(n_dev) || !is_device_ptr(E) || !is_device_ptr(R) ||
      !is_device_ptr(t) || !is_device_ptr(status) || (n_inliers && !is_device_ptr(n_inliers)) ||
      (n_good && !is_device_ptr(n_good)) || (ransac_mask && !is_device_ptr(ransac_mask)) ||
      (pose_mask && !is_device_ptr(pose_mask))) {
    set_error("orbx_estimate_pose_batch: device pointers required");
    return ORBX_E_BADARG;
  }
  if (npairs > pe->max_problems || (size_t)npairs * max_rows > (size_t)pe->max_problems * pe->max_points) {
    set_error("orbx_estimate_pose_batch: %d x %d exceeds the handle's %d x %d", npairs, max_rows, pe->max_problems,
              pe->max_points);
    return ORBX_E_CAPACITY;
  }
  ORBX_CUDA(cudaSetDevice(pe->device));
  const int saved = pe->pp.max_rows;
  pe->pp.max_rows = max_rows;
  int rc = run_essential(pe, pts1, pts2, n_dev, 0, npairs, E, ransac_mask, n_inliers, nullptr);
  if (!rc) rc = run_recover(pe, E, pts1, pts2, n_dev, 0, npairs, false, R, t, pose_mask, n_good, status);
  pe->pp.max_rows = saved;
  return rc;
}

}  // extern "C"
