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
  ok = ok && window6(ax, x, wx0) && window6(ay, y, wy0);
  ok = ok && !ay.periodic;  // rows of the window are contiguous in memory
  if (ND >= 3) ok = ok && window_pair<ZT>(p.ax[2], z, iz0, iz1);
  if (ND == 4) ok = ok && window_pair<double>(p.ax[3], u, iu0, iu1);
  if (!PYI_ANY(ok)) return false;
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
  if (!PYI_ANY(ok)) return false;
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
    // rows: z and dz/dy at the columns 2, 3 of each of the six rows
    double c_z2[6], c_z3[6], c_y2[6], c_y3[6];
    {
      const double* wy = p.w6[1];
#pragma unroll
      for (int i = 0; i < 6; ++i) {
        const int ii = wx0 + i;
        const long long base = plane_offset + (ii >= nx ? ii - nx : ii) * n1 + wy0;
        double v[6];
#pragma unroll
        for (int j = 0; j < 6; ++j) v[j] = ok ? load_value<DT, double>(data, p.dtype, base + j) : 0.0;
        c_z2[i] = v[2];
        c_z3[i] = v[3];
        c_y2[i] = dot6(wy, v);
        c_y3[i] = dot6(wy + 6, v);
        if (DT != -1 && (c_y2[i] != c_y2[i] || c_y3[i] != c_y3[i])) {
          // rare: a NaN (-> NaN result, cache.hpp:165-171) or an infinity (-> generic)
          bool nan_here = false;
#pragma unroll
          for (int j = 0; j < 6; ++j) nan_here = nan_here || v[j] != v[j];
          has_nan = has_nan || nan_here;
          weird = weird || !nan_here;
        }
      }
    }
    // columns: d/dx of those four columns at the rows 2, 3
    const double* wx = p.w6[0];
    const double zz[4] = {c_z2[2], c_z3[2], c_z2[3], c_z3[3]};
    const double zy_[4] = {c_y2[2], c_y3[2], c_y2[3], c_y3[3]};
    const double zx_[4] = {dot6(wx, c_z2), dot6(wx, c_z3), dot6(wx + 6, c_z2), dot6(wx + 6, c_z3)};
    const double zxy_[4] = {dot6(wx, c_y2), dot6(wx, c_y3), dot6(wx + 6, c_y2), dot6(wx + 6, c_y3)};
    f[t] = bicubic_cell(x, y, xc[0], xc[1], yc[0], yc[1], zz, zx_, zy_, zxy_);
  }
  if (!ok) return false;
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

template <int DT, int ND, typename ZT, typename R>
__global__ void __launch_bounds__(128)
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

template <int ND, typename ZT>
static cudaError_t launch_bicubicw_z(const WinNdParams& p, cudaStream_t stream) {
  const int blocks = grid_for(p.n, 128, 8);
  switch (p.dtype) {
    case PYI_F64:
      bicubicw_kernel<PYI_F64, ND, ZT, double><<<blocks, 128, 0, stream>>>(p);
      break;
    case PYI_F32:
      bicubicw_kernel<PYI_F32, ND, ZT, float><<<blocks, 128, 0, stream>>>(p);
      break;
    case PYI_I64:
    case PYI_U64:
      bicubicw_kernel<-1, ND, ZT, double><<<blocks, 128, 0, stream>>>(p);
      break;
    default:
      bicubicw_kernel<-1, ND, ZT, float><<<blocks, 128, 0, stream>>>(p);
      break;
  }
  g_launches.fetch_add(1);
  return cudaGetLastError();
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
