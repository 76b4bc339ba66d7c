// pyi_windowed_nd.cuh -- K6: windowed interpolation on 1-D .. 4-D grids with
// every fitting model, included by pyi_windowed.cu after
// pyi_windowed.cuh (window_indexes, bracket_search, SplineAxis, bicubic_cell)
// and pyi_univariate.cuh.
//
// One thread = one query point of
//   windowed::pybind::univariate_single / univariate_derivative_single
//       pybind/windowed/univariate.hpp:47-120
//   windowed::pybind::bivariate_single      pybind/windowed/bivariate.hpp:46-76
//   windowed::pybind::trivariate_single     pybind/windowed/trivariate.hpp:66-123
//   windowed::pybind::quadrivariate_single  pybind/windowed/quadrivariate.hpp:50-148
// over the window loader math/interpolate/cache_loader.hpp:46-172.  Like K4 the
// kernel is stateless: every point loads its own window (the reference's cache
// miss).  The x / y windows have 2 * half nodes, the z / u "windows" are the
// bracketing pair (cache.hpp:116-124: half window 1).  Each z / u plane is
// interpolated in double by the spatial model, then combined with linear() /
// nearest() along z (in double) and along u (in the RESULT type: the
// reference casts the two z results first, quadrivariate.hpp:126-146).
#pragma once

namespace pyi {

// Univariate<T>::min_size of the model (akima.hpp:28, cspline*.hpp, linear.hpp:25,
// steffen.hpp:27, polynomial.hpp:26)
__device__ __forceinline__ int spline_min_size(int method) {
  switch (method) {
    case PYI_S_AKIMA:
    case PYI_S_AKIMA_PERIODIC: return 5;
    case PYI_S_C_SPLINE:
    case PYI_S_C_SPLINE_NOT_A_KNOT:
    case PYI_S_C_SPLINE_PERIODIC: return 4;
    case PYI_S_LINEAR: return 2;
    default: return 3;
  }
}

// BivariateBase / Spline evaluation of one x-y plane of the window.  The value
// at window node (i, j) is data[(wx[i] * sx + wy[j] * sy) + plane].  Returns
// the interpolated value (double); any NaN among the values sets any_nan.
template <int MAXW>
__device__ __forceinline__ double plane_eval(const WinNdParams& p, const int* wx, int nx,
                                             const int* wy, int ny, const double* xa,
                                             const double* ya, long long sx, long long sy,
                                             long long plane, double x, double y, bool& any_nan) {
  const int method = p.method;
  if (method >= PYI_S_AKIMA) {
    // bivariate/spline.hpp:62-84: along y on every row, then along x
    // BivariateBase::prepare: either side below the model's min_size ->
    // invalid -> NaN (interpolate/bivariate.hpp:56-59, spline.hpp:42-55)
    const int need = spline_min_size(method);
    if (nx < need || ny < need) {
      for (int i = 0; i < nx; ++i) {
        const long long base = wx[i] * sx + plane;
        for (int j = 0; j < ny; ++j) {
          any_nan |= isnan(load_value<-1, double>(p.data, p.dtype, base + wy[j] * sy));
        }
      }
      return CUDART_NAN;
    }
    double v[MAXW];
    for (int i = 0; i < nx; ++i) {
      double row[MAXW];
      const long long base = wx[i] * sx + plane;
      for (int j = 0; j < ny; ++j) {
        const double val = load_value<-1, double>(p.data, p.dtype, base + wy[j] * sy);
        row[j] = val;
        any_nan |= isnan(val);
      }
      v[i] = uni_eval<MAXW>(method, ya, row, ny, y, false);
    }
    return uni_eval<MAXW>(method, xa, v, nx, x, false);
  }
  // Bicubic / Bilinear on the window (K4's arithmetic)
  const bool sizes_ok = nx >= 2 && ny >= 2;
  int i0 = 0, i1 = 0, j0 = 0, j1 = 0;
  const bool found =
      sizes_ok && bracket_search(xa, nx, x, i0, i1) && bracket_search(ya, ny, y, j0, j1);
  const bool bicubic = method == PYI_S_BICUBIC;
  SplineAxis<MAXW> sya;
  if (bicubic && found) sya.prepare(ya, ny);
  double zc0[MAXW], zc1[MAXW], zyc0[MAXW], zyc1[MAXW];
  for (int i = 0; i < nx; ++i) {
    double row[MAXW];
    const long long base = wx[i] * sx + plane;
    for (int j = 0; j < ny; ++j) {
      const double val = load_value<-1, double>(p.data, p.dtype, base + wy[j] * sy);
      row[j] = val;
      any_nan |= isnan(val);
    }
    zc0[i] = row[j0];
    zc1[i] = row[j1];
    if (bicubic && found && !any_nan) {
      spline_two_nodes<MAXW>(sya, ya, row, j0, j1, zyc0[i], zyc1[i]);
    }
  }
  if (any_nan || !found) return CUDART_NAN;
  if (!bicubic) {
    const double t = (x - xa[i0]) / (xa[i1] - xa[i0]);
    const double u = (y - ya[j0]) / (ya[j1] - ya[j0]);
    const double tc = 1.0 - t;
    const double uc = 1.0 - u;
    const double bottom = fma(t, zc0[i1], tc * zc0[i0]);
    const double top = fma(t, zc1[i1], tc * zc1[i0]);
    return fma(u, top, uc * bottom);
  }
  SplineAxis<MAXW> sxa;
  sxa.prepare(xa, nx);
  double z[4] = {zc0[i0], zc1[i0], zc0[i1], zc1[i1]};
  double zy_[4] = {zyc0[i0], zyc1[i0], zyc0[i1], zyc1[i1]};
  double zx_[4], zxy_[4];
  spline_two_nodes<MAXW>(sxa, xa, zc0, i0, i1, zx_[0], zx_[2]);
  spline_two_nodes<MAXW>(sxa, xa, zc1, i0, i1, zx_[1], zx_[3]);
  spline_two_nodes<MAXW>(sxa, xa, zyc0, i0, i1, zxy_[0], zxy_[2]);
  spline_two_nodes<MAXW>(sxa, xa, zyc1, i0, i1, zxy_[1], zxy_[3]);
  return bicubic_cell(x, y, xa[i0], xa[i1], ya[j0], ya[j1], z, zx_, zy_, zxy_);
}

// One query point.  Returns -1 (value in res, NaN when it cannot be computed)
// or the first axis whose coordinate is truly out of bounds
// (cache_loader.hpp:66-80).
template <int MAXW, int ND, typename ZT, typename R>
__device__ __forceinline__ int windowed_nd_point(const WinNdParams& p, double x, double y, ZT z,
                                                 double u, R& res) {
  res = RT<R>::nan();
  int bx[MAXW], by[MAXW];
  int fx = 0, fy = 0, nx = 0, ny = 1, c0, c1;
  const DAxis& ax = p.ax[0];
  if (!window_indexes<MAXW>(ax, x, p.half_x, p.boundary, bx, fx, nx, c0, c1)) {
    return axis_contains(ax, x) ? -1 : 0;
  }
  by[0] = 0;
  if (ND >= 2) {
    if (!window_indexes<MAXW>(p.ax[1], y, p.half_y, p.boundary, by, fy, ny, c0, c1)) {
      return axis_contains(p.ax[1], y) ? -1 : 1;
    }
  }
  int iz[2] = {0, 0}, iu[2] = {0, 0};
  if (ND >= 3) {
    if (!window_pair<ZT>(p.ax[2], z, iz[0], iz[1])) {
      return axis_contains(p.ax[2], z) ? -1 : 2;
    }
  }
  if (ND >= 4) {
    if (!window_pair<double>(p.ax[3], u, iu[0], iu[1])) {
      return axis_contains(p.ax[3], u) ? -1 : 3;
    }
  }
  const int* wx = bx + fx;
  const int* wy = by + fy;
  // coordinates; x re-centred on the query on a periodic axis
  // (cache_loader.hpp:104-128, math.hpp:131-152)
  double xa[MAXW], ya[MAXW];
  const bool per = ax.periodic != 0 && ax.period != 0.0;
  const double half_p = ax.period * 0.5;
  const double lo = x - half_p;
  const double hi = x + half_p;
  for (int k = 0; k < nx; ++k) {
    double raw = axis_coord_f64(ax, wx[k]);
    if (per && !(raw >= lo && raw < hi)) {
      raw = normalize_period(raw, lo, ax.period, ax.inv_period);
    }
    xa[k] = raw;
  }
  if (ND >= 2) {
    for (int k = 0; k < ny; ++k) ya[k] = axis_coord_f64(p.ax[1], wy[k]);
  }
  // element strides of the C-ordered grid
  const long long n1 = ND >= 2 ? p.ax[1].n : 1;
  const long long n2 = ND >= 3 ? p.ax[2].n : 1;
  const long long n3 = ND >= 4 ? p.ax[3].n : 1;
  const long long sy = n2 * n3;
  const long long sx = n1 * sy;
  bool any_nan = false;
  if (ND == 1) {
    double row[MAXW];
    for (int k = 0; k < nx; ++k) {
      const double val = load_value<-1, double>(p.data, p.dtype, wx[k]);
      row[k] = val;
      any_nan |= isnan(val);
    }
    if (any_nan) return -1;
    res = static_cast<R>(uni_eval<MAXW>(p.method, xa, row, nx, x, p.derivative != 0));
    return -1;
  }
  if (ND == 2) {
    const double f = plane_eval<MAXW>(p, wx, nx, wy, ny, xa, ya, sx, sy, 0, x, y, any_nan);
    if (!any_nan) res = static_cast<R>(f);
    return -1;
  }
  if (ND == 3) {
    const double f0 =
        plane_eval<MAXW>(p, wx, nx, wy, ny, xa, ya, sx, sy, iz[0] * n3, x, y, any_nan);
    const double f1 =
        plane_eval<MAXW>(p, wx, nx, wy, ny, xa, ya, sx, sy, iz[1] * n3, x, y, any_nan);
    if (any_nan) return -1;
    const ZT z0 = axis_coord(p.ax[2], iz[0], ZT{});
    const ZT z1 = axis_coord(p.ax[2], iz[1], ZT{});
    res = static_cast<R>(axis_interp<ZT, double>(p.z_method, z, z0, z1, f0, f1));
    return -1;
  }
  // ND == 4: planes (z0,u0), (z0,u1), (z1,u0), (z1,u1)
  const double f00 =
      plane_eval<MAXW>(p, wx, nx, wy, ny, xa, ya, sx, sy, iz[0] * n3 + iu[0], x, y, any_nan);
  const double f01 =
      plane_eval<MAXW>(p, wx, nx, wy, ny, xa, ya, sx, sy, iz[0] * n3 + iu[1], x, y, any_nan);
  const double f10 =
      plane_eval<MAXW>(p, wx, nx, wy, ny, xa, ya, sx, sy, iz[1] * n3 + iu[0], x, y, any_nan);
  const double f11 =
      plane_eval<MAXW>(p, wx, nx, wy, ny, xa, ya, sx, sy, iz[1] * n3 + iu[1], x, y, any_nan);
  if (any_nan) return -1;
  const ZT z0 = axis_coord(p.ax[2], iz[0], ZT{});
  const ZT z1 = axis_coord(p.ax[2], iz[1], ZT{});
  const double u0 = axis_coord_f64(p.ax[3], iu[0]);
  const double u1 = axis_coord_f64(p.ax[3], iu[1]);
  const R f0 = static_cast<R>(axis_interp<ZT, double>(p.z_method, z, z0, z1, f00, f10));
  const R f1 = static_cast<R>(axis_interp<ZT, double>(p.z_method, z, z0, z1, f01, f11));
  res = axis_interp<double, R>(p.u_method, u, u0, u1, f0, f1);
  return -1;
}

template <int MAXW, int ND, typename ZT, typename R>
__global__ void __launch_bounds__(128)
windowed_nd_kernel(const __grid_constant__ WinNdParams p) {
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  unsigned long long key = kNoOob;
  R* __restrict__ out = static_cast<R*>(p.out);
  const ZT* __restrict__ zs = static_cast<const ZT*>(p.z);
  for (long long k = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; k < p.n;
       k += stride) {
    R res;
    const double y = ND >= 2 ? __ldg(p.y + k) : 0.0;
    const ZT z = ND >= 3 ? __ldg(zs + k) : ZT{};
    const double u = ND >= 4 ? __ldg(p.u + k) : 0.0;
    const int bad = windowed_nd_point<MAXW, ND, ZT, R>(p, __ldg(p.x + k), y, z, u, res);
    if (bad >= 0) {
      const unsigned long long kk =
          (static_cast<unsigned long long>(p.index_base + k) << 2) | bad;
      key = kk < key ? kk : key;
    }
    __stcs(out + k, res);
  }
  if (p.bounds_error && key != kNoOob) {
    atomicMin(p.oob_key, key);
  }
}

// One tagged point of the redo pass (not inlined: two call sites, rare path).
template <int MAXW, int ND, typename ZT, typename R>
__device__ __noinline__ void redo_point(const WinNdParams& p, long long k,
                                        unsigned long long& key) {
  R* out = static_cast<R*>(p.out);
  const ZT* zs = static_cast<const ZT*>(p.z);
  R res;
  const double y = ND >= 2 ? __ldg(p.y + k) : 0.0;
  const ZT z = ND >= 3 ? __ldg(zs + k) : ZT{};
  const double u = ND >= 4 ? __ldg(p.u + k) : 0.0;
  const int bad = windowed_nd_point<MAXW, ND, ZT, R>(p, __ldg(p.x + k), y, z, u, res);
  if (bad >= 0) {
    const unsigned long long kk = (static_cast<unsigned long long>(p.index_base + k) << 2) | bad;
    key = kk < key ? kk : key;
  }
  out[k] = res;
}

// Redo pass after a bicubic fast pass (pyi_bicubicw.cuh, pyi_fastwin_nd.cuh):
// only the points that pass tagged are computed.  Tagged points are sparse
// (window at an edge, out-of-bounds or non-finite input), so each warp scans
// its share of the output, queues the tagged indexes in shared memory and
// runs the generic code on 32 queued points at a time -- full warps instead of
// one live lane per tagged point.
template <int MAXW, int ND, typename ZT, typename R>
__global__ void __launch_bounds__(128)
windowed_nd_redo_kernel(const __grid_constant__ WinNdParams p) {
  __shared__ long long queue[4][64];
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  long long* q = queue[warp];
  unsigned queued = 0;
  unsigned long long key = kNoOob;
  R* __restrict__ out = static_cast<R*>(p.out);
  const ZT* __restrict__ zs = static_cast<const ZT*>(p.z);
  auto compute = [&](long long k) { redo_point<MAXW, ND, ZT, R>(p, k, key); };
  // the fast pass listed what it tagged (pyi_bicubics.cuh): no scan
  if (p.queue != nullptr && *p.queue <= p.queue_cap) {
    const unsigned total = *p.queue;
    const unsigned stride = gridDim.x * blockDim.x;
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) compute(p.queue[1 + i]);
    if (p.bounds_error && key != kNoOob) atomicMin(p.oob_key, key);
    return;
  }
  // the scan reads kScan x 32 outputs per step so that the loads overlap
  constexpr int kScan = 8;
  const long long warps = static_cast<long long>(gridDim.x) * 4;
  const long long first = (static_cast<long long>(blockIdx.x) * 4 + warp) * (32 * kScan);
  for (long long base = first; base < p.n; base += warps * (32 * kScan)) {
    R seen[kScan];
#pragma unroll
    for (int j = 0; j < kScan; ++j) {
      const long long k = base + j * 32 + lane;
      seen[j] = k < p.n ? __ldcs(out + k) : R(0);
    }
#pragma unroll
    for (int j = 0; j < kScan; ++j) {
      const long long k = base + j * 32 + lane;
      const bool tagged = RedoTag<R>::is(seen[j]);
      const unsigned mask = __ballot_sync(0xffffffffu, tagged);
      if (mask == 0) continue;
      if (tagged) q[queued + __popc(mask & ((1u << lane) - 1u))] = k;
      queued += __popc(mask);
      __syncwarp();
      if (queued >= 32) {
        const long long mine = q[lane];
        const long long moved = lane + 32 < queued ? q[lane + 32] : 0;
        __syncwarp();
        q[lane] = moved;
        queued -= 32;
        __syncwarp();
        compute(mine);
      }
    }
  }
  if (lane < queued) compute(q[lane]);
  if (p.bounds_error && key != kNoOob) {
    atomicMin(p.oob_key, key);
  }
}

template <int MAXW, int ND, typename ZT>
static cudaError_t launch_windowed_nd_r(const WinNdParams& p, cudaStream_t stream) {
  const int blocks = grid_for(p.n, 128, 8);
  const bool r64 = p.dtype == PYI_F64 || p.dtype == PYI_I64 || p.dtype == PYI_U64;
  if (p.redo) {
    // only the 6 x 6 bicubic fast passes (2-D .. 4-D) are followed by a redo pass
    if constexpr (MAXW == 6 && ND >= 2) {
      if (r64) {
        windowed_nd_redo_kernel<MAXW, ND, ZT, double><<<blocks, 128, 0, stream>>>(p);
      } else {
        windowed_nd_redo_kernel<MAXW, ND, ZT, float><<<blocks, 128, 0, stream>>>(p);
      }
      g_launches.fetch_add(1);
      return cudaGetLastError();
    } else {
      return cudaErrorInvalidValue;
    }
  }
  if (r64) {
    windowed_nd_kernel<MAXW, ND, ZT, double><<<blocks, 128, 0, stream>>>(p);
  } else {
    windowed_nd_kernel<MAXW, ND, ZT, float><<<blocks, 128, 0, stream>>>(p);
  }
  g_launches.fetch_add(1);
  return cudaGetLastError();
}

template <int MAXW>
cudaError_t launch_windowed_nd_w(const WinNdParams& p, cudaStream_t stream) {
  const bool zi = p.ndim >= 3 && p.ax[2].coord_i64 != 0;
  switch (p.ndim) {
    case 1: return launch_windowed_nd_r<MAXW, 1, double>(p, stream);
    case 2: return launch_windowed_nd_r<MAXW, 2, double>(p, stream);
    case 3:
      return zi ? launch_windowed_nd_r<MAXW, 3, long long>(p, stream)
                : launch_windowed_nd_r<MAXW, 3, double>(p, stream);
    case 4:
      return zi ? launch_windowed_nd_r<MAXW, 4, long long>(p, stream)
                : launch_windowed_nd_r<MAXW, 4, double>(p, stream);
    default: return cudaErrorInvalidValue;
  }
}

}  // namespace pyi
