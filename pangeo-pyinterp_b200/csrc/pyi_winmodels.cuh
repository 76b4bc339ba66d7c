// pyi_winmodels.cuh -- K4m: the windowed fitting models other than bicubic on
// a 6 x 6 window of a float64 / float32 grid with regular ascending x / y axes,
// 2-D .. 4-D, in TOLERANCE value mode.  Included by pyi_windowed.cu after
// pyi_bicubicw.cuh, whose window selection, row loads and tags it shares.
//
// What is computed: windowed::pybind::{bi,tri,quadri}variate with the Spline
// evaluator (pybind/windowed/bivariate.hpp:46-76, math/interpolate/bivariate/
// spline.hpp:96-118): the univariate model along y on each of the six window
// rows, then the same model along x through the six row values.  On a uniform
// axis the models below do not depend on the spacing at all -- with
// A_k = v[k+1] - v[k] (node-to-node differences) and t the position in the
// centre cell, every one of them is the cubic Hermite form
//     s(t) = v2 + t (D2 + t ((3 A2 - 2 D2 - D3) + t (D2 + D3 - 2 A2)))
// (akima.hpp:176-183, steffen.hpp:160-175 with h = t dx), D2 / D3 being the
// model's derivative estimates at the two centre nodes times the spacing:
//   * akima / akima_periodic (akima.hpp:84-133): D = (w1 A_prev + w2 A_next) /
//     (w1 + w2) with w1 = |A_next+1 - A_next|, w2 = |A_prev - A_prev-1| (the
//     mean of the two when both weights vanish); the slope past the window's
//     last node is REPLACED by 2 A3 - A2 (by A0 in the periodic variant): the
//     reference calls boundary_condition(m, size - 1), akima.hpp:86 (reproduced
//     as in pyi_univariate.cuh);
//   * steffen (steffen.hpp:101-127): D = (sign(A_prev) + sign(A_next)) *
//     min(|A_prev|, |A_next|, |A_prev + A_next| / 4);
//   * the models that are linear in the values (linear / bilinear, c_spline,
//     c_spline_not_a_knot, c_spline_periodic) on 3-D / 4-D grids: two dot
//     products with the model's six cubics (pyi_host.cu: model_stencil_cubics);
//     2-D grids take the separable kernel (pyi_bicubics.cuh) instead;
//   * polynomial (polynomial.hpp:58-102), also linear in the values: dot
//     products with six quintics built from the reference's own (aliased)
//     divided-difference table and the axis step (pyi_host.cu: polynomial_stencil).
// No division by a spacing is left (the reference divides by the rounded node
// differences; axes whose max |coordinate| / |step| reaches 2^16 stay on the
// generic kernel, as for bicubic): values agree with the reference to rounding
// (bar 1e-12 / 1e-5), not bit for bit; PYI_VALUES_EXACT and
// PYINTERP_B200_BICUBIC=exact keep the bit-identical generic kernel (K6).
// Everything outside the kernel's reach (incomplete windows in shrink mode,
// out-of-bounds or non-finite coordinates, infinite values) is tagged and
// recomputed by the generic kernel in redo mode.
#pragma once

namespace pyi {

enum { WM_STENCIL = 0, WM_AKIMA = 1, WM_STEFFEN = 2, WM_POLY = 3 };

__device__ __forceinline__ double quintic(const double (&c)[6], double t) {
  return fma(fma(fma(fma(fma(c[5], t, c[4]), t, c[3]), t, c[2]), t, c[1]), t, c[0]);
}

__device__ __forceinline__ double hermite_unit(double v2, double a2, double d2, double d3, double t) {
  const double c = fma(-2.0, d2, fma(3.0, a2, -d3));
  const double d = fma(-2.0, a2, __dadd_rn(d2, d3));
  return fma(t, fma(t, fma(t, d, c), d2), v2);
}

__device__ __forceinline__ double akima_estimate(double w1, double w2, double prev, double next) {
  const double den = __dadd_rn(w1, w2);
  return den != 0.0 ? fma(w1, prev, __dmul_rn(w2, next)) / den : __dmul_rn(__dadd_rn(prev, next), 0.5);
}

__device__ __forceinline__ double akima6(const double (&v)[6], double t, bool periodic) {
  const double a0 = __dsub_rn(v[1], v[0]), a1 = __dsub_rn(v[2], v[1]), a2 = __dsub_rn(v[3], v[2]);
  const double a3 = __dsub_rn(v[4], v[3]);
  const double a4 = periodic ? a0 : fma(2.0, a3, -a2);
  const double d2 = akima_estimate(fabs(__dsub_rn(a3, a2)), fabs(__dsub_rn(a1, a0)), a1, a2);
  const double d3 = akima_estimate(fabs(__dsub_rn(a4, a3)), fabs(__dsub_rn(a2, a1)), a2, a3);
  return hermite_unit(v[2], a2, d2, d3, t);
}

__device__ __forceinline__ double steffen_estimate(double prev, double next) {
  const double sign_sum = __dadd_rn(prev < 0.0 ? -1.0 : 1.0, next < 0.0 ? -1.0 : 1.0);
  double mn = fabs(prev);
  const double b = fabs(next);
  if (b < mn) mn = b;
  const double c = __dmul_rn(0.25, fabs(__dadd_rn(prev, next)));
  if (c < mn) mn = c;
  return __dmul_rn(sign_sum, mn);
}

__device__ __forceinline__ double steffen6(const double (&v)[6], double t) {
  const double a1 = __dsub_rn(v[2], v[1]), a2 = __dsub_rn(v[3], v[2]), a3 = __dsub_rn(v[4], v[3]);
  return hermite_unit(v[2], a2, steffen_estimate(a1, a2), steffen_estimate(a2, a3), t);
}

template <int MODEL>
__device__ __forceinline__ double winmodel6(const WinNdParams& p, const double (&v)[6], const double (&w)[6],
                                            double t) {
  if (MODEL == WM_STENCIL || MODEL == WM_POLY) return dot6(w, v);
  if (MODEL == WM_AKIMA) return akima6(v, t, p.method == PYI_S_AKIMA_PERIODIC);
  return steffen6(v, t);
}

// One query point.  Returns false when it has to be recomputed by the generic
// kernel.  (Window selection and the centre bracket as bicubicw_point.)
template <int DT, int ND, typename ZT, typename R, int MODEL>
__device__ __forceinline__ bool winmodel_point(const WinNdParams& p, double x, double y, ZT z, double u,
                                               R& res) {
  const DAxis& ax = p.ax[0];
  const DAxis& ay = p.ax[1];
  int wx0 = 0, wy0 = 0, iz0 = 0, iz1 = 0, iu0 = 0, iu1 = 0;
  bool ok = sane(x) && sane(y) && (ax.period == 0.0 || ax.period > 0.0078125);
  bool is_nan = false;
  ok = window6_or_nan(ax, x, wx0, is_nan) && ok;
  ok = window6_or_nan(ay, y, wy0, is_nan) && ok;
  ok = ok && !ay.periodic;  // rows of the window are contiguous in memory
  if (ND >= 3 && !window_pair<ZT>(p.ax[2], z, iz0, iz1)) {
    ok = false;
    is_nan = true;
  }
  if (ND == 4 && !window_pair<double>(p.ax[3], u, iu0, iu1)) {
    ok = false;
    is_nan = true;
  }
  is_nan = is_nan && p.boundary == PYI_BOUNDARY_UNDEF && !p.bounds_error;
  if (!PYI_ANY(ok)) {
    res = RT<R>::nan();
    return is_nan;
  }
  wx0 = ok ? wx0 : 0;
  wy0 = ok ? wy0 : 0;
  const int nx = ax.n;
  const bool per = ax.periodic != 0 && ax.period != 0.0;
  const double half_p = __dmul_rn(ax.period, 0.5);
  const double lo = __dsub_rn(x, half_p);
  const double hi = __dadd_rn(x, half_p);
  double xc[2], yc[2];
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    const int i = wx0 + 2 + k;
    double raw = ok ? axis_coord_f64(ax, i >= nx ? i - nx : i) : 0.0;
    ok = ok && sane(raw);
    if (per && !(raw >= lo && raw < hi)) raw = norm_fast(raw, lo, ax.period, ax.inv_period);
    xc[k] = raw;
    yc[k] = ok ? axis_coord_f64(ay, wy0 + 2 + k) : 0.0;
  }
  // ascending centre brackets holding the query strictly after their first node
  ok = ok && xc[0] < x && x <= xc[1] && yc[0] < y && y <= yc[1];
  if (!PYI_ANY(ok)) {
    res = RT<R>::nan();
    return is_nan;
  }
  const double tx = ok ? __dsub_rn(x, xc[0]) / __dsub_rn(xc[1], xc[0]) : 0.5;
  const double ty = ok ? __dsub_rn(y, yc[0]) / __dsub_rn(yc[1], yc[0]) : 0.5;
  double wx[6], wy[6];
  if (MODEL == WM_STENCIL) {
#pragma unroll
    for (int j = 0; j < 6; ++j) {
      wx[j] = cubic4(p.cw[0][j + 1][0], p.cw[0][j + 1][1], p.cw[0][j + 1][2], p.cw[0][j + 1][3], tx);
      wy[j] = cubic4(p.cw[1][j + 1][0], p.cw[1][j + 1][1], p.cw[1][j + 1][2], p.cw[1][j + 1][3], ty);
    }
  }
  if (MODEL == WM_POLY) {
#pragma unroll
    for (int j = 0; j < 6; ++j) {
      wx[j] = quintic(p.pw[0][j], tx);
      wy[j] = quintic(p.pw[1][j], ty);
    }
  }
  const long long n1 = ay.n;
  const long long plane_cells = static_cast<long long>(nx) * n1;
  const long long n3 = ND == 4 ? p.ax[3].n : 1;
  constexpr int kPlanes = ND == 2 ? 1 : (ND == 3 ? 2 : 4);
  const void* data = ND == 2 ? p.data : p.planes;
  double f[kPlanes];
  bool has_nan = false, weird = false;
#pragma unroll
  for (int t = 0; t < kPlanes; ++t) {
    long long plane_offset = 0;
    if (ND >= 3) {
      const int pz = (ND == 4 ? (t >> 1) : t) ? iz1 : iz0;
      const int pu = ND == 4 ? ((t & 1) ? iu1 : iu0) : 0;
      plane_offset = (pz * n3 + pu) * plane_cells;
    }
    double r[6];
#pragma unroll
    for (int i = 0; i < 6; ++i) {
      const int ii = wx0 + i;
      const long long base = plane_offset + (ii >= nx ? ii - nx : ii) * n1 + wy0;
      double v[6];
      load_row6<DT>(data, p.dtype, base, p.cells, ok, v);
      // every value of the window counts for the NaN rule (cache.hpp:165-171),
      // also those a model does not read
      const double sum = __dadd_rn(__dadd_rn(__dadd_rn(v[0], v[1]), __dadd_rn(v[2], v[3])), __dadd_rn(v[4], v[5]));
      if (!isfinite(sum)) {
        bool nan_here = false;
#pragma unroll
        for (int j = 0; j < 6; ++j) nan_here = nan_here || v[j] != v[j];
        has_nan = has_nan || nan_here;
        weird = weird || !nan_here;
      }
      r[i] = winmodel6<MODEL>(p, v, wy, ty);
    }
    f[t] = winmodel6<MODEL>(p, r, wx, tx);
  }
  if (!ok) {
    res = RT<R>::nan();
    return is_nan;
  }
  if (has_nan) {
    res = RT<R>::nan();
    return true;
  }
  double worst = f[0];
#pragma unroll
  for (int t = 1; t < kPlanes; ++t) worst = __dadd_rn(worst, f[t]);
  // infinities, overflow: the generic code decides
  if (weird || !isfinite(worst)) return false;
  if (ND == 2) {
    res = static_cast<R>(f[0]);
    return true;
  }
  const ZT z0 = axis_coord(p.ax[2], iz0, ZT{});
  const ZT z1 = axis_coord(p.ax[2], iz1, ZT{});
  if (ND == 3) {
    res = static_cast<R>(axis_interp<ZT, double>(p.z_method, z, z0, z1, f[0], f[kPlanes - 1]));
    return true;
  }
  const double u0 = axis_coord_f64(p.ax[3], iu0);
  const double u1 = axis_coord_f64(p.ax[3], iu1);
  const R f0 = static_cast<R>(axis_interp<ZT, double>(p.z_method, z, z0, z1, f[0], f[kPlanes / 2]));
  const R f1 =
      static_cast<R>(axis_interp<ZT, double>(p.z_method, z, z0, z1, f[kPlanes > 1 ? 1 : 0], f[kPlanes - 1]));
  res = axis_interp<double, R>(p.u_method, u, u0, u1, f0, f1);
  return true;
}

template <int DT, int ND, typename ZT, typename R, int MODEL, int MINB>
__global__ void __launch_bounds__(128, MINB)
winmodel_kernel(const __grid_constant__ WinNdParams p) {
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  R* __restrict__ out = static_cast<R*>(p.out);
  const ZT* __restrict__ zs = static_cast<const ZT*>(p.z);
  for (long long k = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; k < p.n;
       k += stride) {
    R res;
    const ZT z = ND >= 3 ? __ldcs(zs + k) : ZT{};
    const double u = ND == 4 ? __ldcs(p.u + k) : 0.0;
    const bool done =
        winmodel_point<DT, ND, ZT, R, MODEL>(p, __ldcs(p.x + k), __ldcs(p.y + k), z, u, res);
    __stcs(out + k, done ? res : RedoTag<R>::value());
  }
}

// Which models the kernel serves, or -1.
static int winmodel_of(const WinNdParams& p) {
  switch (p.method) {
    case PYI_S_AKIMA:
    case PYI_S_AKIMA_PERIODIC:
      return WM_AKIMA;
    case PYI_S_STEFFEN:
      return WM_STEFFEN;
    case PYI_S_POLYNOMIAL:
      return WM_POLY;
    case PYI_S_LINEAR:
    case PYI_S_BILINEAR:
    case PYI_S_C_SPLINE:
    case PYI_S_C_SPLINE_NOT_A_KNOT:
    case PYI_S_C_SPLINE_PERIODIC:
      return p.ndim >= 3 ? WM_STENCIL : -1;  // 2-D: pyi_bicubics.cuh
    default:
      return -1;
  }
}

// float64 / float32 grids whose x / y axes are regular, ascending and close
// enough to the origin for uniform-spacing arithmetic (w_ok, cw_ok: pyi_host.cu),
// complete 6 x 6 windows; 3-D / 4-D grids need the plane-major copy.
static bool winmodels_eligible(const WinNdParams& p) {
  if (winmodel_of(p) < 0 || p.half_x != 3 || p.half_y != 3 || !p.w_ok || !p.cw_ok) return false;
  if (p.dtype != PYI_F64 && p.dtype != PYI_F32) return false;
  if (p.ndim < 2 || (p.ndim >= 3 && p.planes == nullptr)) return false;
  if (p.ax[0].kind != 0 || p.ax[1].kind != 0 || p.ax[1].periodic) return false;
  // 16-byte row loads of float64 grids
  const void* data = p.ndim == 2 ? p.data : p.planes;
  return p.dtype != PYI_F64 || (reinterpret_cast<uintptr_t>(data) & 15) == 0;
}

template <int ND, typename ZT, int MODEL, int MINB>
static void launch_winmodels_b(const WinNdParams& p, cudaStream_t stream) {
  const int blocks = grid_for(p.n, 128, MINB < 4 ? 8 : MINB);
  if (p.dtype == PYI_F64) {
    winmodel_kernel<PYI_F64, ND, ZT, double, MODEL, MINB><<<blocks, 128, 0, stream>>>(p);
  } else {
    winmodel_kernel<PYI_F32, ND, ZT, float, MODEL, MINB><<<blocks, 128, 0, stream>>>(p);
  }
}

template <int ND, typename ZT, int MODEL>
static void launch_winmodels_m(const WinNdParams& p, cudaStream_t stream) {
  // resident blocks per SM: 2-D (L2-resident grid, latency bound) 8 x 128 threads at <= 64
  // registers; 3-D / 4-D (DRAM bound) no cap.  PYINTERP_B200_M_MINB = 5 | 8 | 10 for experiments.
  if constexpr (ND == 2) {
    const char* e = getenv("PYINTERP_B200_M_MINB");
    const int minb = e ? atoi(e) : 8;
    if (minb == 5) return launch_winmodels_b<ND, ZT, MODEL, 5>(p, stream);
    if (minb == 10) return launch_winmodels_b<ND, ZT, MODEL, 10>(p, stream);
    return launch_winmodels_b<ND, ZT, MODEL, 8>(p, stream);
  } else {
    return launch_winmodels_b<ND, ZT, MODEL, 1>(p, stream);
  }
}

template <int ND, typename ZT>
static void launch_winmodels_z(const WinNdParams& p, cudaStream_t stream) {
  switch (winmodel_of(p)) {
    case WM_AKIMA: launch_winmodels_m<ND, ZT, WM_AKIMA>(p, stream); break;
    case WM_STEFFEN: launch_winmodels_m<ND, ZT, WM_STEFFEN>(p, stream); break;
    case WM_POLY: launch_winmodels_m<ND, ZT, WM_POLY>(p, stream); break;
    default:
      if constexpr (ND >= 3) launch_winmodels_m<ND, ZT, WM_STENCIL>(p, stream);
      break;
  }
}

// The fast pass; the caller follows it with the generic kernel in redo mode.
static cudaError_t launch_winmodels(const WinNdParams& p, cudaStream_t stream) {
  const bool zi = p.ndim >= 3 && p.ax[2].coord_i64 != 0;
  if (p.ndim == 2) {
    launch_winmodels_z<2, double>(p, stream);
  } else if (p.ndim == 3) {
    if (zi) launch_winmodels_z<3, long long>(p, stream);
    else launch_winmodels_z<3, double>(p, stream);
  } else {
    if (zi) launch_winmodels_z<4, long long>(p, stream);
    else launch_winmodels_z<4, double>(p, stream);
  }
  g_launches.fetch_add(1);
  return cudaGetLastError();
}

}  // namespace pyi
