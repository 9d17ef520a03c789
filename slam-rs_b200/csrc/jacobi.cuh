// One-sided (Hestenes) Jacobi SVD for the tiny f64 systems of the geometry kernels (triangulate.cu, pose.cu):
// column pairs of A are rotated until mutually orthogonal, V collects the rotations; afterwards column j of A is
// sigma_j * u_j and column j of V is v_j.  Converges in 4-6 sweeps on these sizes.
#pragma once

namespace orbx {

template <int R, int C>
__device__ __forceinline__ void jacobi_onesided(double (&a)[R][C], double (&v)[C][C]) {
#pragma unroll
  for (int r = 0; r < C; ++r)
#pragma unroll
    for (int c = 0; c < C; ++c) v[r][c] = r == c ? 1.0 : 0.0;
  for (int sweep = 0; sweep < 16; ++sweep) {
    bool rotated = false;
#pragma unroll
    for (int p = 0; p < C - 1; ++p) {
#pragma unroll
      for (int q = p + 1; q < C; ++q) {
        double alpha = 0, beta = 0, gamma = 0;
#pragma unroll
        for (int r = 0; r < R; ++r) {
          alpha += a[r][p] * a[r][p];
          beta += a[r][q] * a[r][q];
          gamma += a[r][p] * a[r][q];
        }
        if (fabs(gamma) <= 1e-15 * sqrt(alpha * beta) || gamma == 0.0) continue;
        rotated = true;
        const double zeta = (beta - alpha) / (2.0 * gamma);
        const double t = copysign(1.0, zeta) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
        const double c = 1.0 / sqrt(1.0 + t * t), s = c * t;
#pragma unroll
        for (int r = 0; r < R; ++r) {
          const double ap = a[r][p], aq = a[r][q];
          a[r][p] = c * ap - s * aq;
          a[r][q] = s * ap + c * aq;
        }
#pragma unroll
        for (int r = 0; r < C; ++r) {
          const double vp = v[r][p], vq = v[r][q];
          v[r][p] = c * vp - s * vq;
          v[r][q] = s * vp + c * vq;
        }
      }
    }
    if (!rotated) break;
  }
}

// unit right singular vector of the smallest singular value of a 4x4 system (a is destroyed)
__device__ __forceinline__ void null_vector4(double (&a)[4][4], double (&h)[4]) {
  double v[4][4];
  jacobi_onesided<4, 4>(a, v);
  int best = 0;
  double bn = 0;
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    double nn = 0;
#pragma unroll
    for (int r = 0; r < 4; ++r) nn += a[r][c] * a[r][c];
    if (c == 0 || nn < bn) bn = nn, best = c;
  }
#pragma unroll
  for (int r = 0; r < 4; ++r) h[r] = best == 0 ? v[r][0] : best == 1 ? v[r][1] : best == 2 ? v[r][2] : v[r][3];
  // LAPACK returns a unit vector; the rotations keep V orthonormal to rounding, renormalise anyway
  const double inv = 1.0 / sqrt(h[0] * h[0] + h[1] * h[1] + h[2] * h[2] + h[3] * h[3]);
#pragma unroll
  for (int r = 0; r < 4; ++r) h[r] *= inv;
}

}  // namespace orbx
