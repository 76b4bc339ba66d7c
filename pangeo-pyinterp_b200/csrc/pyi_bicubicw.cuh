// pyi_bicubicw.cuh -- K4w: bicubic on a 6 x 6 window through PRECOMPUTED
// node-derivative weights, on 2-D, 3-D and 4-D grids.
//
// The reference's bicubic (math/interpolate/bivariate/bicubic.hpp:44-186) needs
// the first derivative of the natural cubic spline through the six window nodes
// at the two centre nodes -- ten tridiagonal solves per point
// (cspline.hpp:70-117).  Those derivatives are LINEAR in the node values:
//     d(node 2) = sum_k W[0][k] v[k],   d(node 3) = sum_k W[1][k] v[k]
// where W depends only on the six abscissae.  On a REGULAR axis every window
// has the same W: it is computed once per grid, by the very device code of the
// exact fast path (pyi_fastwin.cuh) applied to the six unit vectors, and sits
// in the kernel parameters.  A point then costs 120 FMAs for its ten
// derivative pairs instead of ~900 FP64 operations, plus the 16-term cell
// polynomial (bicubic.hpp:117-186, unchanged).  Grids with an irregular x or y
// axis keep the exact fast path: with very unequal spacings W has large
// cancelling entries and the rounding of the two orders can differ by more
// than the tolerance.
//
// Parity: the summation order differs from the reference's Thomas recurrences,
// so values agree to rounding (a few 1e-16 of the window's magnitude; the
// north-star tolerance is 1e-12 relative) instead of bit for bit.  Window
// selection, NaN handling and every case outside this kernel's reach stay
// bit-exact: points with an incomplete window, out-of-bounds or non-finite
// coordinates, or infinite values are tagged and recomputed by the generic
// kernel in redo mode (as in pyi_fastwin_nd.cuh).  PYINTERP_B200_BICUBIC=exact
// selects the bit-identical fast path instead.
//
// 3-D / 4-D grids read the plane-major copy ([z][u][x][y]).
// Included by pyi_windowed.cu after pyi_fastwin.cuh.
#pragma once

namespace pyi {

// W of one window: out[0..5] = d(node 2) weights, out[6..11] = d(node 3).
// Unusable windows (zero / mixed-sign / immoderate spacings) get NaN weights:
// the point then goes to the generic kernel.
__device__ __forceinline__ void spline6_weights(const double (&xa)[6], double* out) {
  bool ok = true, neg = false, pos = false;
#pragma unroll
  for (int k = 0; k < 5; ++k) {
    const double h = __dsub_rn(xa[k + 1], xa[k]);
    ok = ok && moderate_spacing(h);
    neg = neg || negative(h);
    pos = pos || !negative(h);
  }
  ok = ok && !(neg && pos);
  if (!ok) {
    for (int k = 0; k < 12; ++k) out[k] = CUDART_NAN;
    return;
  }
  double s[S6_COUNT];
  spline6_prepare(xa, s);
  for (int k = 0; k < 6; ++k) {
    double e[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
    e[k] = 1.0;
    double d2, d3;
    spline6_node_derivatives(s, e, d2, d3);
    out[k] = d2;
    out[6 + k] = d3;
  }
}

// Setup kernel (one thread): W of a window of spacing `step`.
__global__ void spline6_weights_kernel(double step, double* __restrict__ out) {
  double xa[6];
  for (int k = 0; k < 6; ++k) xa[k] = __dmul_rn(static_cast<double>(k), step);
  spline6_weights(xa, out);
}

cudaError_t launch_spline6_weights(double step, double* out12, cudaStream_t stream) {
  spline6_weights_kernel<<<1, 1, 0, stream>>>(step, out12);
  g_launches.fetch_add(1);
  return cudaGetLastError();
}

__device__ __forceinline__ double dot6(const double* w, const double (&v)[6]) {
  double s = __dmul_rn(w[0], v[0]);
#pragma unroll
  for (int k = 1; k < 6; ++k) s = fma(w[k], v[k], s);
  return s;
}

// Six consecutive grid values from element `base`.  float64 grids: aligned
// 16-byte loads covering the run (3 when base is even, 4 when odd) instead of
// six 8-byte ones; `cells` bounds the over-read of the odd case.
template <int DT>
__device__ __forceinline__ void load_row6(const void* data, int dtype, long long base,
                                          long long cells, bool ok, double (&v)[6]) {
  if (ok) PYI_CHECK_RANGE(base, 6, cells, "K4w window row");
  if (DT == PYI_F64) {
    const long long b0 = base & ~1LL;
    const bool odd = (base & 1LL) != 0;
    if (b0 + 8 <= cells) {
      const double2* q = reinterpret_cast<const double2*>(static_cast<const double*>(data) + b0);
      const double2 zero = make_double2(0.0, 0.0);
      const double2 t0 = ok ? __ldg(q) : zero;
      const double2 t1 = ok ? __ldg(q + 1) : zero;
      const double2 t2 = ok ? __ldg(q + 2) : zero;
      const double2 t3 = ok && odd ? __ldg(q + 3) : zero;
      v[0] = odd ? t0.y : t0.x;
      v[1] = odd ? t1.x : t0.y;
      v[2] = odd ? t1.y : t1.x;
      v[3] = odd ? t2.x : t1.y;
      v[4] = odd ? t2.y : t2.x;
      v[5] = odd ? t3.x : t2.y;
      return;
    }
  }
#pragma unroll
  for (int j = 0; j < 6; ++j) v[j] = ok ? load_value<DT, double>(data, dtype, base + j) : 0.0;
}

// window6 (pyi_fastwin.cuh) that also recognises the windows for which the
// generic code (window_indexes, pyi_windowed.cuh) certainly fails in the
// "undef" boundary mode -- coordinate not bracketed by the axis, or a forward
// bracket whose six-node window leaves a non-periodic axis: the result is then
// NaN (cache_loader.hpp:69-80) and nothing has to be recomputed.
__device__ __forceinline__ bool window6_or_nan(const DAxis& a, double c, int& w0, bool& is_nan) {
  int p0, p1;
  const int n = a.n;
  if (!find_indexes<double>(a, c, p0, p1)) {
    is_nan = is_nan || n != 1;  // singleton axes have their own rule (axis.hpp:823-832)
    return false;
  }
  if (a.periodic) {
    if (n < 6) return false;
    const bool fwd = (p1 == p0 + 1) || (p0 == n - 1 && p1 == 0);
    w0 = p0 - 2 < 0 ? p0 - 2 + n : p0 - 2;
    return fwd;
  }
  if (p1 != p0 + 1) return false;
  w0 = p0 - 2;
  if (w0 < 0 || p1 + 2 >= n) {
    is_nan = true;
    return false;
  }
  return true;
}

// One query point.  Returns false when it has to be recomputed by the generic
// kernel.
template <int DT, int ND, typename ZT, typename R>
__device__ __forceinline__ bool bicubicw_point(const WinNdParams& p, double x, double y, ZT z,
                                               double u, R& res) {
  const DAxis& ax = p.ax[0];
  const DAxis& ay = p.ax[1];
  int wx0 = 0, wy0 = 0, iz0 = 0, iz1 = 0, iu0 = 0, iu1 = 0;
  // norm_fast needs a positive period and quotients below 2^48
  bool ok = sane(x) && sane(y) && (ax.period == 0.0 || ax.period > 0.0078125);
  // certain NaN: only without bounds_error (an out-of-bounds coordinate must be
  // reported) and in the undef boundary mode (shrink keeps partial windows)
  bool is_nan = false;
  ok = window6_or_nan(ax, x, wx0, is_nan) && ok;
  ok = window6_or_nan(ay, y, wy0, is_nan) && ok;
  ok = ok && !ay.periodic;  // rows of the window are contiguous in memory
  if (ND >= 3 && !window_pair<ZT>(p.ax[2], z, iz0, iz1)) {
    ok = false;
    is_nan = true;  // window_pair only fails on a non-singleton axis
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
  // the centre bracket (cache_loader.hpp:110-133: x re-centred on the query)
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
  // the bracket holds c strictly in the window's own order (window_usable)
  ok = ok && (xc[0] < xc[1] ? (xc[0] < x && x <= xc[1]) : (xc[0] > x && x >= xc[1]));
  ok = ok && (yc[0] < yc[1] ? (yc[0] < y && y <= yc[1]) : (yc[0] > y && y >= yc[1]));
  if (!PYI_ANY(ok)) {
    res = RT<R>::nan();
    return is_nan;
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
    // Row by row: z and dz/dy at the columns 2, 3 (dot products with Wy), folded
    // at once into the four column derivatives d/dx at the rows 2, 3 (dot
    // products with Wx, accumulated in row order = dot6's order).
    const double* wy = p.w6[1];
    const double* wx = p.w6[0];
    double zz[4], zy_[4], zx_[4], zxy_[4];
#pragma unroll
    for (int i = 0; i < 6; ++i) {
      const int ii = wx0 + i;
      const long long base = plane_offset + (ii >= nx ? ii - nx : ii) * n1 + wy0;
      double v[6];
      load_row6<DT>(data, p.dtype, base, p.cells, ok, v);
      const double y2 = dot6(wy, v);
      const double y3 = dot6(wy + 6, v);
      if (DT != -1 && (y2 != y2 || y3 != y3)) {
        // rare: a NaN (-> NaN result, cache.hpp:165-171) or an infinity (-> generic)
        bool nan_here = false;
#pragma unroll
        for (int j = 0; j < 6; ++j) nan_here = nan_here || v[j] != v[j];
        has_nan = has_nan || nan_here;
        weird = weird || !nan_here;
      }
      if (i == 2) {
        zz[0] = v[2];
        zz[1] = v[3];
        zy_[0] = y2;
        zy_[1] = y3;
      }
      if (i == 3) {
        zz[2] = v[2];
        zz[3] = v[3];
        zy_[2] = y2;
        zy_[3] = y3;
      }
      // index 0: (i0,j0) 1: (i0,j1) 2: (i1,j0) 3: (i1,j1)
      if (i == 0) {
        zx_[0] = __dmul_rn(wx[0], v[2]);
        zx_[1] = __dmul_rn(wx[0], v[3]);
        zx_[2] = __dmul_rn(wx[6], v[2]);
        zx_[3] = __dmul_rn(wx[6], v[3]);
        zxy_[0] = __dmul_rn(wx[0], y2);
        zxy_[1] = __dmul_rn(wx[0], y3);
        zxy_[2] = __dmul_rn(wx[6], y2);
        zxy_[3] = __dmul_rn(wx[6], y3);
      } else {
        zx_[0] = fma(wx[i], v[2], zx_[0]);
        zx_[1] = fma(wx[i], v[3], zx_[1]);
        zx_[2] = fma(wx[6 + i], v[2], zx_[2]);
        zx_[3] = fma(wx[6 + i], v[3], zx_[3]);
        zxy_[0] = fma(wx[i], y2, zxy_[0]);
        zxy_[1] = fma(wx[i], y3, zxy_[1]);
        zxy_[2] = fma(wx[6 + i], y2, zxy_[2]);
        zxy_[3] = fma(wx[6 + i], y3, zxy_[3]);
      }
    }
    f[t] = bicubic_cell(x, y, xc[0], xc[1], yc[0], yc[1], zz, zx_, zy_, zxy_);
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
  // unusable windows (NaN weights), infinities, overflow: exact generic code
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
  // planes (z0,u0), (z0,u1), (z1,u0), (z1,u1) (quadrivariate.hpp:126-146)
  const double u0 = axis_coord_f64(p.ax[3], iu0);
  const double u1 = axis_coord_f64(p.ax[3], iu1);
  const R f0 = static_cast<R>(axis_interp<ZT, double>(p.z_method, z, z0, z1, f[0], f[kPlanes / 2]));
  const R f1 =
      static_cast<R>(axis_interp<ZT, double>(p.z_method, z, z0, z1, f[kPlanes > 1 ? 1 : 0], f[kPlanes - 1]));
  res = axis_interp<double, R>(p.u_method, u, u0, u1, f0, f1);
  return true;
}

template <int DT, int ND, typename ZT, typename R, int MINB>
__global__ void __launch_bounds__(128, MINB)
bicubicw_kernel(const __grid_constant__ WinNdParams p) {
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  R* __restrict__ out = static_cast<R*>(p.out);
  const ZT* __restrict__ zs = static_cast<const ZT*>(p.z);
  for (long long k = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; k < p.n;
       k += stride) {
    R res;
    const ZT z = ND >= 3 ? __ldcs(zs + k) : ZT{};
    const double u = ND == 4 ? __ldcs(p.u + k) : 0.0;
    const bool done =
        bicubicw_point<DT, ND, ZT, R>(p, __ldcs(p.x + k), __ldcs(p.y + k), z, u, res);
    __stcs(out + k, done ? res : RedoTag<R>::value());
  }
}

template <int ND, typename ZT, int MINB>
static cudaError_t launch_bicubicw_m(const WinNdParams& p, cudaStream_t stream) {
  const int blocks = grid_for(p.n, 128, MINB < 4 ? 8 : MINB);
  switch (p.dtype) {
    case PYI_F64:
      bicubicw_kernel<PYI_F64, ND, ZT, double, MINB><<<blocks, 128, 0, stream>>>(p);
      break;
    case PYI_F32:
      bicubicw_kernel<PYI_F32, ND, ZT, float, MINB><<<blocks, 128, 0, stream>>>(p);
      break;
    case PYI_I64:
    case PYI_U64:
      bicubicw_kernel<-1, ND, ZT, double, MINB><<<blocks, 128, 0, stream>>>(p);
      break;
    default:
      bicubicw_kernel<-1, ND, ZT, float, MINB><<<blocks, 128, 0, stream>>>(p);
      break;
  }
  g_launches.fetch_add(1);
  return cudaGetLastError();
}

template <int ND, typename ZT>
static cudaError_t launch_bicubicw_z(const WinNdParams& p, cudaStream_t stream) {
  // Resident blocks per SM (register cap).  2-D (L2-resident grid, latency
  // bound): 5 blocks x 128 threads at 96 registers measured best of {4, 5, 6, 8}
  // (12.3 / 13.2 / 12.5 / 10.8 Gpts/s).  3-D / 4-D (DRAM bound): no cap, the
  // compiler keeps more loads in flight per thread.  PYINTERP_B200_W_MINB=1|5
  // overrides for experiments.
  const char* e = getenv("PYINTERP_B200_W_MINB");
  const int minb = e ? atoi(e) : (ND == 2 ? 5 : 1);
  if (minb == 5) return launch_bicubicw_m<ND, ZT, 5>(p, stream);
  return launch_bicubicw_m<ND, ZT, 1>(p, stream);
}

// The weights pass; the caller follows it with the generic kernel in redo mode.
static cudaError_t launch_bicubicw(const WinNdParams& p, cudaStream_t stream) {
  if (p.ndim == 2) return launch_bicubicw_z<2, double>(p, stream);
  const bool zi = p.ax[2].coord_i64 != 0;
  if (p.ndim == 3) {
    return zi ? launch_bicubicw_z<3, long long>(p, stream) : launch_bicubicw_z<3, double>(p, stream);
  }
  return zi ? launch_bicubicw_z<4, long long>(p, stream) : launch_bicubicw_z<4, double>(p, stream);
}

}  // namespace pyi
