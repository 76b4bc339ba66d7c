// pyi_device.cuh -- device-side axis lookups for the B200 (sm_100a) kernels.
//
// Index selection must be bit-exact with the reference's C++ kernel
// (math::Axis<T>::find_indexes, cxx/include/pyinterp/math/axis.hpp:720-778 and
// the containers of math/axis/container.hpp).  Everything that decides an index
// is therefore written with explicit round-to-nearest intrinsics (__dsub_rn,
// __dmul_rn, __dadd_rn): nvcc may not contract them into FMAs, which would
// change the rounding the reference's x86-64 build performs (it has no implicit
// FMA: no -march, CMakeLists.txt:139-141).  The file is also compiled with
// -fmad=false so that plain expressions keep IEEE semantics; explicit fma()
// calls mirror the reference's explicit std::fma.
#pragma once

#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>

namespace pyi {

// Flat device mirror of pyi_axis_desc (include/pyinterp_b200.h); `points`
// is a device pointer.  Passed by value inside __grid_constant__ parameters.
struct DAxis {
  int32_t kind;      // 0 regular, 1 irregular
  int32_t coord_i64; // 0 f64, 1 i64
  int32_t asc;
  int32_t periodic;  // is_periodic_
  int32_t n;         // size (host rejects axes >= 2^31 - 2)
  int32_t lut_bins;  // irregular f64: number of uniform bins of the lookup table (0: none)
  // f64
  double start, step, inv_step, period, inv_period, min, max;
  // i64
  long long istart, istep, istep2, iperiod, imin, imax;
  const void* points;     // irregular: n values
  const int32_t* lut;     // irregular f64: lut_bins + 1 lower-bound indexes
  double lut_min, lut_inv; // bin(c) = (c - lut_min) * lut_inv
  // fast-path constants (pyi_fast2d.cuh), computed on the host with the
  // reference's expressions
  double hi;            // fl(min + period)
  double seam_c1;       // periodic: normalize(node(0), min = node(n-1)), math/axis.hpp:294-301
  double seam_c1_f32;   // same against the float-rounded node(n-1) (float results)
  int32_t fast_ok;      // ascending f64 axis with moderate magnitudes and spacings, or an
                        // ascending regular i64 axis with |start|, span < 2^61 (then
                        // inv_step = 1.0 / step and ispan = (n - 1) * step)
  long long ispan;
  // tolerance-mode bilinear kernel (pyi_fast2d_tol.cuh): regular ascending f64
  // axis whose brackets can be verified from t = (c - node(k)) / dx alone.
  // 0: not eligible; 1: general (node(k), node(k+1) with the reference's two
  // roundings); 2: exact nodes (fma(k, step, start) is the reference's node for
  // every k and every spacing, the seam included, is exactly `step`).
  int32_t tol_ok;
  int32_t tol_xhi;         // periodic axes: exponent field (high word) bound of |x|, 2^28 steps
  double tol_off;          // -(start * inv_step) - 0.5: rint(c * inv_step + tol_off) = floor guess
  double tol_poff;         // -(min * inv_period) - 0.5, periodic axes
  // A DESCENDING regular f64 axis without period (e.g. latitude 90 -> -90) that
  // passed the same checks as fast_ok: the kernels that know about it bracket
  // node(k) > c > node(k+1) by index and swap the two corners (the reference's
  // pair is (k+1, k): lower node first, math/axis.hpp:759-766).  fast_ok stays 0,
  // so every kernel that does not test desc_ok keeps rejecting such an axis.
  int32_t desc_ok;
};

// ---------------------------------------------------------------------------
// scalar helpers (math.hpp)
// ---------------------------------------------------------------------------

// IEEE remainder(d, p), p > 0 finite.  Exact, like std::remainder
// (math.hpp:104-111 calls it).  Fast path: n = rint(d / p) estimated with the
// host-computed reciprocal, r = fma(-n, p, d) is exact for |n| < 2^48 (the
// true remainder for an n within one of the correct quotient is representable,
// Sterbenz), then n is corrected by at most one step and ties go to even n.
// Huge quotients, infinities and NaN take libdevice's exact remainder().
__device__ __forceinline__ double ieee_remainder(double d, double p, double inv_p) {
  const double q = __dmul_rn(d, inv_p);
  if (!(fabs(q) < 281474976710656.0)) {  // 2^48; also catches NaN / inf
    return remainder(d, p);
  }
  double n = rint(q);
  double r = fma(-n, p, d);
  const double half = 0.5 * p;
  if (fabs(r) > half) {  // estimate off by one
    const double s = r > 0.0 ? 1.0 : -1.0;
    n += s;
    r = __dsub_rn(r, s * p);
  } else if (fabs(r) == half) {  // exact tie: n must be even
    const double n2 = 0.5 * n;
    if (n2 != rint(n2)) {  // n odd
      r = -r;
    }
  }
  // std::remainder returns a zero with the sign of d
  if (r == 0.0) {
    r = copysign(0.0, d);
  }
  return r;
}

// math.hpp:104-111 + :120-123: remainder with the sign of the period, + min.
__device__ __forceinline__ double normalize_period(double x, double min, double period,
                                                   double inv_period) {
  double r = ieee_remainder(__dsub_rn(x, min), period, inv_period);
  if (r != 0.0 && (signbit(r) != signbit(period))) {
    r = __dadd_rn(r, period);
  }
  return __dadd_rn(r, min);
}

// math.hpp:93-97
__device__ __forceinline__ long long remainder_i64(long long x, long long y) {
  const long long r = x % y;
  return (r != 0 && (r ^ y) < 0) ? r + y : r;
}

// Axis::normalize_coordinate(c, min) axis.hpp:294-301 -- tests period_, not
// is_periodic_.
__device__ __forceinline__ double axis_normalize(const DAxis& a, double c, double min) {
  if (a.period != 0.0 && (c >= __dadd_rn(min, a.period) || c < min)) {
    if (a.period > 0.0) {
      return normalize_period(c, min, a.period, a.inv_period);
    }
    // negative period: keep libdevice semantics (never used by sane axes)
    double r = remainder(__dsub_rn(c, min), a.period);
    if (r != 0.0 && (signbit(r) != signbit(a.period))) r = __dadd_rn(r, a.period);
    return __dadd_rn(r, min);
  }
  return c;
}
__device__ __forceinline__ long long axis_normalize(const DAxis& a, long long c, long long min) {
  if (a.iperiod != 0 && (c >= min + a.iperiod || c < min)) {
    return remainder_i64(c - min, a.iperiod) + min;
  }
  return c;
}

// ---------------------------------------------------------------------------
// containers (math/axis/container.hpp)
// ---------------------------------------------------------------------------
// coordinate_value: regular = start + index * step as two roundings
// (container.hpp:480-483), irregular = points[index].
__device__ __forceinline__ double axis_coord_f64(const DAxis& a, int i) {
  if (a.kind == 0) {
    return __dadd_rn(a.start, __dmul_rn(static_cast<double>(i), a.step));
  }
  return __ldg(static_cast<const double*>(a.points) + i);
}
__device__ __forceinline__ long long axis_coord_i64(const DAxis& a, int i) {
  if (a.kind == 0) {
    return a.istart + static_cast<long long>(i) * a.istep;
  }
  return __ldg(static_cast<const long long*>(a.points) + i);
}

// Regular<floating>::find_index(c, bounded=false) container.hpp:550-564.
// static_cast<int64_t> of NaN / out-of-range is UB in the reference; its
// x86-64 build yields INT64_MIN, i.e. "negative" => -1.  Every out-of-range
// value therefore maps to -1 here as well.
__device__ __forceinline__ int regular_find_f64(const DAxis& a, double c) {
  const double r = round(__dmul_rn(__dsub_rn(c, a.start), a.inv_step));
  return (r >= 0.0 && r < static_cast<double>(a.n)) ? static_cast<int>(r) : -1;
}

// Signed 64-bit division with C++ truncation.  For |num|, |den| < 2^62 and a
// quotient below 2^40 the quotient estimated in double is off by at most one
// (relative error < 2^-51) and is corrected exactly with integer arithmetic;
// otherwise the (slow) emulated 64-bit division runs.
__device__ __forceinline__ long long div_trunc_i64(long long num, long long den) {
  const long long lim = 1LL << 62;
  if (num > -lim && num < lim && den > -lim && den < lim) {
    const long long un = num < 0 ? -num : num;
    const long long ud = den < 0 ? -den : den;
    const double qd = __ddiv_rn(static_cast<double>(un), static_cast<double>(ud));
    if (qd < 1099511627776.0) {  // 2^40
      long long uq = static_cast<long long>(qd);
      const long long ur = un - uq * ud;
      uq += (ur < 0) ? -1 : ((ur >= ud) ? 1 : 0);
      return ((num < 0) != (den < 0)) ? -uq : uq;
    }
  }
  return num / den;
}

// Regular<integral>::find_index + round container.hpp:588-615
__device__ __forceinline__ int regular_find_i64(const DAxis& a, long long c) {
  const long long num =
      static_cast<long long>(static_cast<unsigned long long>(c) -
                             static_cast<unsigned long long>(a.istart));
  const long long idx = num > 0 ? div_trunc_i64(num + a.istep2, a.istep)
                                : div_trunc_i64(num - a.istep2, a.istep);
  return (idx >= 0 && idx < a.n) ? static_cast<int>(idx) : -1;
}

// std::lower_bound on [lo, hi) with the axis comparator (a < b ascending,
// a > b descending), branch-free halving.
template <typename T>
__device__ __forceinline__ int lower_bound_range(const T* __restrict__ p, int lo, int hi, T c,
                                                 bool asc) {
  int len = hi - lo;
  while (len > 0) {
    const int half = len >> 1;
    const T v = __ldg(p + lo + half);
    const bool before = asc ? (v < c) : (v > c);
    lo = before ? lo + half + 1 : lo;
    len = before ? len - half - 1 : half;
  }
  return lo;
}

// Irregular::find_index_impl(c, bounded=false) container.hpp:382-404
template <typename T>
__device__ __forceinline__ int irregular_tail(const T* __restrict__ p, int n, int it, T c,
                                              bool asc) {
  if (it == 0) {
    const T v = __ldg(p);
    return (asc ? (c < v) : (c > v)) ? -1 : 0;
  }
  if (it == n) {
    const T v = __ldg(p + n - 1);
    return (asc ? (v < c) : (v > c)) ? -1 : n - 1;
  }
  const T lo = __ldg(p + it - 1);
  const T hi = __ldg(p + it);
  const T dl = c - lo;
  const T dh = hi - c;
  return (asc ? (dl < dh) : (dl > dh)) ? it - 1 : it;
}

__device__ __forceinline__ int irregular_find_f64(const DAxis& a, double c) {
  const double* p = static_cast<const double*>(a.points);
  int lo = 0, hi = a.n;
  if (a.lut_bins > 0) {
    // Uniform-bin table built on the host with this very expression:
    // bin() is monotone in c, lut[b] = #points whose bin is < b, so the
    // lower bound of c lies in [lut[bin(c)], lut[bin(c) + 1]].
    const double f = __dmul_rn(__dsub_rn(c, a.lut_min), a.lut_inv);
    if (f >= 0.0 && f < static_cast<double>(a.lut_bins)) {
      const int b = static_cast<int>(f);
      lo = __ldg(a.lut + b);
      hi = __ldg(a.lut + b + 1);
    } else if (f < 0.0) {
      hi = __ldg(a.lut);  // bin < 0: before every point whose bin >= 0
    } else if (f >= static_cast<double>(a.lut_bins)) {
      lo = __ldg(a.lut + a.lut_bins);
    }  // NaN never reaches this function
  }
  const int it = lower_bound_range<double>(p, lo, hi, c, a.asc != 0);
  return irregular_tail<double>(p, a.n, it, c, a.asc != 0);
}
__device__ __forceinline__ int irregular_find_i64(const DAxis& a, long long c) {
  const long long* p = static_cast<const long long*>(a.points);
  const int it = lower_bound_range<long long>(p, 0, a.n, c, a.asc != 0);
  return irregular_tail<long long>(p, a.n, it, c, a.asc != 0);
}

__device__ __forceinline__ int container_find(const DAxis& a, double c) {
  return a.kind == 0 ? regular_find_f64(a, c) : irregular_find_f64(a, c);
}
__device__ __forceinline__ int container_find(const DAxis& a, long long c) {
  return a.kind == 0 ? regular_find_i64(a, c) : irregular_find_i64(a, c);
}
__device__ __forceinline__ double axis_coord(const DAxis& a, int i, double) {
  return axis_coord_f64(a, i);
}
__device__ __forceinline__ long long axis_coord(const DAxis& a, int i, long long) {
  return axis_coord_i64(a, i);
}
__device__ __forceinline__ bool is_fill(double c) { return isnan(c); }
__device__ __forceinline__ bool is_fill(long long c) { return c == LLONG_MIN; }
__device__ __forceinline__ double axis_min(const DAxis& a, double) { return a.min; }
__device__ __forceinline__ long long axis_min(const DAxis& a, long long) { return a.imin; }

// ---------------------------------------------------------------------------
// Axis<T>::find_indexes(T) axis.hpp:720-778.  T = double or long long.
// ---------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ bool find_indexes(const DAxis& a, T c, int& o0, int& o1) {
  if (is_fill(c)) {
    return false;
  }
  c = axis_normalize(a, c, axis_min(a, T{}));
  int i0 = container_find(a, c);
  const int n = a.n;
  if (i0 < 0) {
    o0 = n - 1;
    o1 = 0;
    return a.periodic != 0;
  }
  const T node = axis_coord(a, i0, T{});
  if (c == node) {  // delta == 0, axis.hpp:748-757 (pair sorted by index)
    if (n == 1) {
      return false;
    }
    const int j = (i0 == n - 1) ? i0 - 1 : i0 + 1;
    o0 = min(i0, j);
    o1 = max(i0, j);
    return true;
  }
  int i1 = i0;
  const int up = a.asc ? 1 : -1;
  if (c < node) {  // delta < 0, axis.hpp:759-766 (pair sorted by coordinate)
    i0 -= up;
  } else {
    i1 += up;
  }
  if (a.periodic) {  // math::remainder(i, n) for i in [-1, n]
    o0 = i0 < 0 ? i0 + n : (i0 >= n ? i0 - n : i0);
    o1 = i1 < 0 ? i1 + n : (i1 >= n ? i1 - n : i1);
    return true;
  }
  o0 = i0;
  o1 = i1;
  return i0 >= 0 && i1 >= 0 && i0 < n && i1 < n;
}

}  // namespace pyi
