// pyi_univariate.cuh -- the univariate fitting models of the windowed path
// (math/interpolate/univariate/*.hpp, selected by factory.hpp:27-62) as device
// functions on a window held in per-thread arrays.  Included by
// pyi_windowed.cu (compiled with -fmad=false: only the explicit fma() calls
// below fuse, mirroring the reference's std::fma).
//
// Each model is the pair compute_coefficients + interpolate_ / derivative_ of
// the reference on n nodes (xa, ya), evaluated at x; NaN where the reference
// returns Fill (fewer than min_size nodes, x outside the nodes, ...).  A fresh
// interpolator object per call is restated (the reference keeps work arrays
// between calls; for a fixed window size every entry it reads is rewritten).
//
// Two models depend on how Eigen vectorises the reference's expressions; the
// x86-64 default build (SSE2, packets of two doubles, 16-byte aligned
// buffers) is what is reproduced:
//   * Polynomial: the in-place divided-difference update
//     (polynomial.hpp:77-82) reads work_(ix-1+k) while writing work_(ix+k);
//     Eigen's linear-vectorised assignment peels to a 16-byte boundary, then
//     processes packets (both lanes loaded before the store).  From five nodes
//     on the table therefore differs from Newton's divided differences; the
//     reference's value is what is returned here.
//   * CSplinePeriodic: the two .sum() reductions (cspline_periodic.hpp:93-97,
//     112-113) add packet-wise (even / odd partial sums) when they have at
//     least two elements.
#pragma once

namespace pyi {

__device__ __forceinline__ double uni_nan() { return CUDART_NAN; }

// Eigen redux (Redux.h, LinearVectorizedTraversal / NoUnrolling) of n terms
// t(k) with SSE2 packets of two and no alignment peel.
template <typename Term>
__device__ __forceinline__ double eigen_sum_sse2(int n, Term t) {
  const int aligned = (n / 2) * 2;
  if (aligned == 0) {
    double r = t(0);
    for (int k = 1; k < n; ++k) r = r + t(k);
    return r;
  }
  const int aligned2 = (n / 4) * 4;
  double a0 = t(0), a1 = t(1);
  if (aligned > 2) {
    double b0 = t(2), b1 = t(3);
    for (int k = 4; k < aligned2; k += 4) {
      a0 = a0 + t(k);
      a1 = a1 + t(k + 1);
      b0 = b0 + t(k + 2);
      b1 = b1 + t(k + 3);
    }
    a0 = a0 + b0;
    a1 = a1 + b1;
    if (aligned > aligned2) {
      a0 = a0 + t(aligned2);
      a1 = a1 + t(aligned2 + 1);
    }
  }
  double r = a0 + a1;
  for (int k = aligned; k < n; ++k) r = r + t(k);
  return r;
}

// CSplineCoefficients + the Horner forms of cspline_base.hpp:32-41,83-142
__device__ __forceinline__ double cspline_eval(const double* xa, const double* ya, int n,
                                               const double* c2, double x, bool derivative) {
  int i0, i1;
  if (!bracket_search(xa, n, x, i0, i1)) return uni_nan();
  const double dx = xa[i1] - xa[i0];
  const double dy = ya[i1] - ya[i0];
  const double h = x - xa[i0];
  const double third = 1.0 / 3.0;
  const double c0 = c2[i0];
  const double c1 = c2[i1];
  const double b = (dy / dx) - dx * (c1 + 2.0 * c0) * third;
  const double d = (c1 - c0) / (3.0 * dx);
  if (derivative) return b + h * fma(3.0 * h, d, 2.0 * c0);
  return ya[i0] + h * fma(h, fma(h, d, c0), b);
}

// univariate/cspline.hpp:53-140 (natural boundary conditions)
template <int MAXW>
__device__ __forceinline__ double uni_cspline(const double* xa, const double* ya, int n, double x,
                                              bool derivative) {
  if (n < 4) return uni_nan();
  const int m = n - 2;
  double sub[MAXW], diag[MAXW], sup[MAXW], b[MAXW], c[MAXW], d[MAXW], c2[MAXW];
  for (int i = 0; i < m; ++i) {
    const double x1 = xa[i + 1];
    const double y1 = ya[i + 1];
    const double h0 = x1 - xa[i];
    const double h1 = xa[i + 2] - x1;
    const double y0 = y1 - ya[i];
    const double y2 = ya[i + 2] - y1;
    const double g0 = (h0 != 0.0) ? 1.0 / h0 : 0.0;
    const double g1 = (h1 != 0.0) ? 1.0 / h1 : 0.0;
    sub[i] = h0;
    diag[i] = 2.0 * (h0 + h1);
    sup[i] = h1;
    b[i] = 3.0 * fma(y2, g1, -y0 * g0);
  }
  double item = 1.0 / diag[0];
  c[0] = sup[0] * item;
  d[0] = b[0] * item;
  for (int i = 1; i < m - 1; ++i) {
    item = sub[i];
    const double inv_m = 1.0 / (diag[i] - item * c[i - 1]);
    c[i] = sup[i] * inv_m;
    d[i] = fma(-item, d[i - 1], b[i]) * inv_m;
  }
  item = sub[m - 1];
  d[m - 1] = (b[m - 1] - item * d[m - 2]) / (diag[m - 1] - item * c[m - 2]);
  c2[0] = 0.0;
  c2[n - 1] = 0.0;
  c2[m] = d[m - 1];
  for (int i = m - 2; i >= 0; --i) c2[i + 1] = fma(-c[i], c2[i + 2], d[i]);
  return cspline_eval(xa, ya, n, c2, x, derivative);
}

// univariate/cspline_not_a_knot.hpp:52-229
template <int MAXW>
__device__ __forceinline__ double uni_cspline_nak(const double* xa, const double* ya, int n,
                                                  double x, bool derivative) {
  if (n < 4) return uni_nan();
  const int m = n - 2;
  double h[MAXW], delta[MAXW], diag[MAXW], b[MAXW], c[MAXW], d[MAXW], c2[MAXW];
  const double eps = 2.220446049250313080847263336181640625e-16 * 10.0;
  bool dup = false;
  for (int i = 0; i < n - 1; ++i) {
    h[i] = xa[i + 1] - xa[i];
    dup = dup || fabs(h[i]) < eps;
  }
  if (dup) return uni_nan();
  for (int i = 0; i < n - 1; ++i) delta[i] = (ya[i + 1] - ya[i]) / h[i];
  const double alpha_0 = (h[0] + h[1]) / h[1];
  const double beta_0 = -h[0] / h[1];
  const double alpha_n = -h[m] / h[m - 1];
  const double beta_n = (h[m - 1] + h[m]) / h[m - 1];
  // reduced system rows 0 .. m-1; sub-diagonal of row i is h[i] except the last
  diag[0] = fma(h[0], alpha_0, 2.0 * (h[0] + h[1]));
  const double a01 = fma(h[0], beta_0, h[1]);
  b[0] = 3.0 * (delta[1] - delta[0]);
  for (int i = 2; i < m; ++i) {
    const int row = i - 1;
    diag[row] = 2.0 * (h[i - 1] + h[i]);
    b[row] = 3.0 * (delta[i] - delta[i - 1]);
  }
  const int last = m - 1;
  const double sub_last = fma(h[m], alpha_n, h[m - 1]);  // last > 0 always (n >= 4)
  diag[last] = fma(h[m], beta_n, 2.0 * (h[m - 1] + h[m]));
  b[last] = 3.0 * (delta[m] - delta[m - 1]);
  double inv_diag = 1.0 / diag[0];
  c[0] = a01 * inv_diag;
  d[0] = b[0] * inv_diag;
  for (int i = 1; i < m - 1; ++i) {
    const double sub = h[i];
    inv_diag = 1.0 / fma(-sub, c[i - 1], diag[i]);
    c[i] = h[i + 1] * inv_diag;
    d[i] = fma(-sub, d[i - 1], b[i]) * inv_diag;
  }
  inv_diag = 1.0 / fma(-sub_last, c[last - 1], diag[last]);
  d[last] = fma(-sub_last, d[last - 1], b[last]) * inv_diag;
  c2[m] = d[last];
  for (int i = last - 1; i >= 0; --i) c2[i + 1] = fma(-c[i], c2[i + 2], d[i]);
  c2[0] = fma(alpha_0, c2[1], beta_0 * c2[2]);
  c2[n - 1] = fma(alpha_n, c2[n - 3], beta_n * c2[n - 2]);
  return cspline_eval(xa, ya, n, c2, x, derivative);
}

// univariate/cspline_periodic.hpp:51-194 (Sherman-Morrison cyclic solve)
template <int MAXW>
__device__ __forceinline__ double uni_cspline_periodic(const double* xa, const double* ya, int n,
                                                       double x, bool derivative) {
  if (n < 4) return uni_nan();
  const int s = n - 1;  // system size
  // cyclic tridiagonal matrix: diag[i] = A(i,i), off[i] = A(i,i+1) = A(i+1,i)
  // for i < s-1, corner = A(s-1,0) = A(0,s-1)
  double diag[MAXW], off[MAXW], b[MAXW], alpha[MAXW], gamma[MAXW], delta[MAXW], z[MAXW], c2[MAXW];
  for (int i = 0; i < n - 2; ++i) {
    const double h0 = xa[i + 1] - xa[i];
    const double h1 = xa[i + 2] - xa[i + 1];
    const double y0 = ya[i + 1] - ya[i];
    const double y1 = ya[i + 2] - ya[i + 1];
    const double g0 = (h0 != 0.0) ? 1.0 / h0 : 0.0;
    const double g1 = (h1 != 0.0) ? 1.0 / h1 : 0.0;
    off[i] = h1;
    diag[i] = 2.0 * (h0 + h1);
    b[i] = 3.0 * fma(y1, g1, -y0 * g0);
  }
  double corner;
  {
    const int i = n - 2;
    const double h0 = xa[i + 1] - xa[i];
    const double h1 = xa[1] - xa[0];
    const double y0 = ya[i + 1] - ya[i];
    const double y1 = ya[1] - ya[0];
    const double g0 = (h0 != 0.0) ? 1.0 / h0 : 0.0;
    const double g1 = (h1 != 0.0) ? 1.0 / h1 : 0.0;
    corner = h1;
    diag[i] = 2.0 * (h0 + h1);
    b[i] = 3.0 * fma(y1, g1, -y0 * g0);
  }
  // n == 4 (s == 3): A(0, s-1) written by the wrap-around is also A(0, 2) =
  // A(s-3, s-1)... the entries the solver reads are named below.
  const int sm1 = s - 1, sm2 = s - 2, sm3 = s - 3;
  // A(i, i-1) = off[i-1]; A(i, i+1) = off[i]; A(sm1, 0) = corner.
  // For s == 3 the loop wrote A(2,1) = A(1,2) = off[1] and the wrap-around
  // A(2,0) = A(0,2) = corner; A(0,1) = A(1,0) = off[0].
  alpha[0] = diag[0];
  gamma[0] = off[0] / alpha[0];
  delta[0] = corner / alpha[0];
  for (int i = 1; i < sm2; ++i) {
    alpha[i] = diag[i] - off[i - 1] * gamma[i - 1];
    gamma[i] = off[i] / alpha[i];
    delta[i] = -delta[i - 1] * off[i - 1] / alpha[i];
  }
  const double sum_delta =
      eigen_sum_sse2(sm2, [&](int k) { return alpha[k] * delta[k] * delta[k]; });
  alpha[sm2] = diag[sm2] - off[sm3] * gamma[sm3];
  gamma[sm2] = (off[sm2] - off[sm3] * delta[sm3]) / alpha[sm2];
  alpha[sm1] = diag[sm1] - sum_delta - alpha[sm2] * gamma[sm2] * gamma[sm2];
  z[0] = b[0];
  for (int i = 1; i < sm1; ++i) z[i] = b[i] - z[i - 1] * gamma[i - 1];
  const double sum_z = eigen_sum_sse2(sm2, [&](int k) { return delta[k] * z[k]; });
  z[sm1] = b[sm1] - sum_z - gamma[sm2] * z[sm2];
  // c = z / alpha; back substitution into x[0 .. s) = c2[1 .. n)
  double* xs = c2 + 1;
  xs[sm1] = z[sm1] / alpha[sm1];
  xs[sm2] = z[sm2] / alpha[sm2] - gamma[sm2] * xs[sm1];
  for (int i = sm3; i >= 0; --i) {
    xs[i] = z[i] / alpha[i] - gamma[i] * xs[i + 1] - delta[i] * xs[sm1];
  }
  c2[0] = c2[n - 1];
  return cspline_eval(xa, ya, n, c2, x, derivative);
}

// univariate/linear.hpp:29-53
__device__ __forceinline__ double uni_linear(const double* xa, const double* ya, int n, double x,
                                             bool derivative) {
  if (n < 2) return uni_nan();
  int i0, i1;
  if (!bracket_search(xa, n, x, i0, i1)) return uni_nan();
  const double dx = xa[i1] - xa[i0];
  const double dy = ya[i1] - ya[i0];
  if (derivative) return dy / dx;
  return fma(dy / dx, x - xa[i0], ya[i0]);
}

// univariate/akima.hpp:64-184 and akima_periodic.hpp:27-37.  The slopes live
// in m[2 .. n]; boundary_condition() is called with size - 1 (akima.hpp:86), so
// besides the unused m[0], m[1], m[n+1] it REPLACES the last real slope m[n]
// (by 2 m[n-1] - m[n-2], or by the first slope m[2] in the periodic variant),
// which the node slopes s[n-3] .. s[n-1] then read.  Reproduced as is.
template <int MAXW>
__device__ __forceinline__ double uni_akima(const double* xa, const double* ya, int n, double x,
                                            bool derivative, bool periodic) {
  if (n < 5) return uni_nan();
  double m[MAXW + 4], s[MAXW];
  for (int i = 0; i < n - 1; ++i) m[i + 2] = (ya[i + 1] - ya[i]) / (xa[i + 1] - xa[i]);
  m[n] = periodic ? m[2] : 2.0 * m[n - 1] - m[n - 2];
  for (int i = 2; i < n - 2; ++i) {
    const double w1 = fabs(m[i + 3] - m[i + 2]);
    const double w2 = fabs(m[i + 1] - m[i]);
    const double den = w1 + w2;
    s[i] = (den != 0.0) ? fma(w1, m[i + 1], w2 * m[i + 2]) / den : (m[i + 1] + m[i + 2]) * 0.5;
  }
  s[0] = m[2];
  s[1] = (m[2] + m[3]) * 0.5;
  s[n - 2] = (m[n - 1] + m[n]) * 0.5;
  s[n - 1] = m[n];
  int i0, i1;
  if (!bracket_search(xa, n, x, i0, i1)) return uni_nan();
  const double dx = xa[i1] - xa[i0];
  const double h = x - xa[i0];
  const double bi = s[i0];
  const double mi = m[i0 + 2];
  const double ci = (3.0 * mi - 2.0 * s[i0] - s[i1]) / dx;
  const double di = (s[i0] + s[i1] - 2.0 * mi) / (dx * dx);
  if (derivative) return bi + h * fma(3.0 * h, di, 2.0 * ci);
  return ya[i0] + h * fma(h, fma(h, di, ci), bi);
}

// univariate/steffen.hpp:46-176
__device__ __forceinline__ double steffen_sign(double y) {  // copysign(1, y) of steffen.hpp:36-39
  return y < 0.0 ? -1.0 : 1.0;
}
template <int MAXW>
__device__ __forceinline__ double uni_steffen(const double* xa, const double* ya, int n, double x,
                                              bool derivative) {
  if (n < 3) return uni_nan();
  double yp[MAXW];
  yp[0] = (ya[1] - ya[0]) / (xa[1] - xa[0]);
  for (int i = 1; i < n - 1; ++i) {
    const double hi = xa[i + 1] - xa[i];
    const double him1 = xa[i] - xa[i - 1];
    const double si = (ya[i + 1] - ya[i]) / hi;
    const double sim1 = (ya[i] - ya[i - 1]) / him1;
    const double pi = (sim1 * hi + si * him1) / (him1 + hi);
    const double sign_sum = steffen_sign(sim1) + steffen_sign(si);
    // std::min({a, b, c}): first minimum under operator<
    double mn = fabs(sim1);
    const double a1 = fabs(si);
    if (a1 < mn) mn = a1;
    const double a2 = 0.5 * fabs(pi);
    if (a2 < mn) mn = a2;
    yp[i] = sign_sum * mn;
  }
  yp[n - 1] = (ya[n - 1] - ya[n - 2]) / (xa[n - 1] - xa[n - 2]);
  int i0, i1;
  if (!bracket_search(xa, n, x, i0, i1)) return uni_nan();
  const double h = x - xa[i0];
  const double hi = xa[i1] - xa[i0];
  const double si = (ya[i1] - ya[i0]) / hi;
  const double a = (yp[i0] + yp[i1] - 2.0 * si) / (hi * hi);
  const double b = (3.0 * si - 2.0 * yp[i0] - yp[i1]) / hi;
  const double c = yp[i0];
  if (derivative) return c + h * fma(3.0 * h, a, 2.0 * b);
  return ya[i0] + h * fma(h, fma(h, a, b), c);
}

// univariate/polynomial.hpp:58-166 with Eigen's SSE2 traversal of the aliased
// update (see the header comment)
template <int MAXW>
__device__ __forceinline__ double uni_polynomial(const double* xa, const double* ya, int n,
                                                 double x, bool derivative) {
  if (n < 3) return uni_nan();
  double work[MAXW];
  work[0] = ya[0];
  for (int k = 1; k < n; ++k) work[k] = (ya[k] - ya[k - 1]) / (xa[k] - xa[k - 1]);
  for (int ix = 2; ix < n; ++ix) {
    const int cnt = n - ix;
    int k = 0;
    const int peel = (ix & 1) < cnt ? (ix & 1) : cnt;  // first 16-byte aligned destination
    for (; k < peel; ++k) {
      work[ix + k] = (work[ix + k] - work[ix - 1 + k]) / (xa[ix + k] - xa[k]);
    }
    for (; k + 2 <= cnt; k += 2) {
      const double a0 = (work[ix + k] - work[ix - 1 + k]) / (xa[ix + k] - xa[k]);
      const double a1 = (work[ix + k + 1] - work[ix + k]) / (xa[ix + k + 1] - xa[k + 1]);
      work[ix + k] = a0;
      work[ix + k + 1] = a1;
    }
    for (; k < cnt; ++k) {
      work[ix + k] = (work[ix + k] - work[ix - 1 + k]) / (xa[ix + k] - xa[k]);
    }
  }
  int i0, i1;
  if (!bracket_search(xa, n, x, i0, i1)) return uni_nan();
  if (!derivative) {
    double y = work[n - 1];
    for (int ix = n - 1; ix > 0; --ix) y = fma(y, x - xa[ix - 1], work[ix - 1]);
    return y;
  }
  // taylor(), polynomial.hpp:104-128
  double c[MAXW], w[MAXW];
  for (int k = 0; k < n; ++k) {
    c[k] = 0.0;
    w[k] = 0.0;
  }
  w[n - 1] = 1.0;
  c[0] = work[0];
  for (int ix = n - 1; ix > 0; --ix) {
    const double dx = xa[n - 1 - ix] - x;
    w[ix - 1] = -w[ix] * dx;
    for (int jx = ix; jx < n - 1; ++jx) w[jx] = w[jx] - w[jx + 1] * dx;
    for (int jx = ix - 1; jx < n; ++jx) c[jx - ix + 1] = c[jx - ix + 1] + w[jx] * work[n - ix];
  }
  return c[1];
}

// Univariate<T>::operator() / derivative() for the model `method`
// (pyi_spatial_method >= PYI_S_AKIMA)
template <int MAXW>
__device__ __noinline__ double uni_eval(int method, const double* xa, const double* ya, int n,
                                        double x, bool derivative) {
  switch (method) {
    case PYI_S_AKIMA:
      return uni_akima<MAXW>(xa, ya, n, x, derivative, false);
    case PYI_S_AKIMA_PERIODIC:
      return uni_akima<MAXW>(xa, ya, n, x, derivative, true);
    case PYI_S_C_SPLINE:
      return uni_cspline<MAXW>(xa, ya, n, x, derivative);
    case PYI_S_C_SPLINE_NOT_A_KNOT:
      return uni_cspline_nak<MAXW>(xa, ya, n, x, derivative);
    case PYI_S_C_SPLINE_PERIODIC:
      return uni_cspline_periodic<MAXW>(xa, ya, n, x, derivative);
    case PYI_S_LINEAR:
      return uni_linear(xa, ya, n, x, derivative);
    case PYI_S_STEFFEN:
      return uni_steffen<MAXW>(xa, ya, n, x, derivative);
    default:
      return uni_polynomial<MAXW>(xa, ya, n, x, derivative);
  }
}

}  // namespace pyi
