// eig_small.cuh -- largest real eigenvalue of a small GENERAL (non-symmetric) real matrix.
//
// The reference bounds the Lagrange remainder with r^2 = max Re lambda(q B), B = I + k_fb^T k_fb, from a general `torch.eig`
// of every item (safe_exploration/utils.py:175-185, 651-665) and asserts that all imaginary parts vanish.  On the CEM
// path q is symmetric and the kernels use the symmetric form C^T q C (Jacobi, reach_kernels.cuh); this routine is what
// keeps an entry point that is handed a non-symmetric shape matrix (test_utils.py:35-59 does that) on the reference's
// answer for n_s >= 3: reduction to Hessenberg form by Givens similarity rotations, then shifted QR iterations
// (Wilkinson shift from the trailing 2 x 2 block) with deflation.  A complex pair -- the reference's assertion -- and a
// failure to converge both return NaN, which the ellipsoid step reports through its NaN counter.
#pragma once
#include "common.cuh"

namespace se {

template <typename T>
SE_HD void givens(T a, T b, T& c, T& s) {  // [c s; -s c] [a; b] = [r; 0]
  if (b == T(0)) {
    c = T(1), s = T(0);
    return;
  }
  const T r = sqrt(a * a + b * b);
  c = a / r, s = b / r;
}

// n <= NMAX; `a` (row-major NMAX x NMAX storage, leading n x n used) is destroyed.
template <typename T, int NMAX>
__host__ __device__ __noinline__ T max_real_eig_general(T (&a)[NMAX][NMAX], int n) {
  const T nan = sqrt(T(-1));
  const T eps = sizeof(T) == 8 ? T(2.3e-16) : T(1.2e-7);
  // Hessenberg form: zero a[i][j], i >= j + 2, by rotations in the (i - 1, i) plane, applied as similarities
  for (int j = 0; j + 2 < n; ++j)
    for (int i = n - 1; i >= j + 2; --i) {
      T c, s;
      givens(a[i - 1][j], a[i][j], c, s);
      for (int k = 0; k < n; ++k) {  // rows i - 1, i
        const T x = a[i - 1][k], y = a[i][k];
        a[i - 1][k] = c * x + s * y, a[i][k] = -s * x + c * y;
      }
      for (int k = 0; k < n; ++k) {  // columns i - 1, i
        const T x = a[k][i - 1], y = a[k][i];
        a[k][i - 1] = c * x + s * y, a[k][i] = -s * x + c * y;
      }
      a[i][j] = T(0);
    }
  T best = nan;
  bool have = false;
  auto take = [&](T ev) {
    if (!have || ev > best) best = ev;
    have = true;
  };
  int hi = n - 1, iter = 0;
  T cs[NMAX], sn[NMAX];
  while (hi >= 0) {
    if (hi == 0) {
      take(a[0][0]);
      break;
    }
    int l = hi;
    while (l > 0) {
      const T sc = fabs(a[l][l]) + fabs(a[l - 1][l - 1]);
      if (fabs(a[l][l - 1]) <= eps * (sc > T(0) ? sc : T(1))) break;
      --l;
    }
    if (l == hi) {  // 1 x 1 block
      take(a[hi][hi]);
      --hi, iter = 0;
      continue;
    }
    const T p = a[hi - 1][hi - 1], q = a[hi - 1][hi], r = a[hi][hi - 1], t = a[hi][hi];
    const T half = T(0.5) * (p - t), disc = half * half + q * r;
    if (l == hi - 1) {  // isolated 2 x 2 block: closed form
      if (disc < T(0)) return nan;  // complex pair: the reference asserts
      take(T(0.5) * (p + t) + sqrt(disc));
      hi -= 2, iter = 0;
      continue;
    }
    if (++iter > 60) return nan;
    // Wilkinson shift: the eigenvalue of the trailing block closer to a[hi][hi]
    T mu = t;
    if (disc >= T(0)) {
      const T sq = sqrt(disc), e1 = T(0.5) * (p + t) + sq, e2 = T(0.5) * (p + t) - sq;
      mu = fabs(e1 - t) < fabs(e2 - t) ? e1 : e2;
    }
    if (iter == 15 || iter == 30) mu = t + fabs(r) + fabs(a[hi - 1][hi - 2]);  // exceptional shift
    for (int i = l; i <= hi; ++i) a[i][i] -= mu;
    for (int k = l; k < hi; ++k) {  // R = Q^T (A - mu I), rotations kept
      givens(a[k][k], a[k + 1][k], cs[k], sn[k]);
      for (int j = k; j <= hi; ++j) {
        const T x = a[k][j], y = a[k + 1][j];
        a[k][j] = cs[k] * x + sn[k] * y, a[k + 1][j] = -sn[k] * x + cs[k] * y;
      }
      a[k + 1][k] = T(0);
    }
    for (int k = l; k < hi; ++k) {  // A' = R Q + mu I
      const int last = k + 1 < hi ? k + 1 : hi;
      for (int i = l; i <= last; ++i) {
        const T x = a[i][k], y = a[i][k + 1];
        a[i][k] = cs[k] * x + sn[k] * y, a[i][k + 1] = -sn[k] * x + cs[k] * y;
      }
    }
    for (int i = l; i <= hi; ++i) a[i][i] += mu;
  }
  return best;
}

}  // namespace se
