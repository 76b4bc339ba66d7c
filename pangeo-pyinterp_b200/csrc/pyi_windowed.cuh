// pyi_windowed.cuh -- K4: windowed 2-D interpolation (bicubic / windowed
// bilinear), included by pyi_kernels.cu.  All arithmetic is double regardless
// of the grid dtype (pybind/windowed/bivariate.hpp:32-33); only the final value
// is cast to the result type.
//
// The kernel is stateless: every point (re)loads its own window, i.e. the
// reference's behaviour for a cache miss (cache_loader.hpp:188-212), which is
// what a uniformly random query stream produces on the CPU as well.
//
// Work saved without changing a single rounding:
//   * Bicubic::compute_coefficients (bicubic.hpp:57-87) differentiates the
//     whole window (3 * 2h splines); only the four corners of the cell that
//     brackets the query are read by interpolate_ (bicubic.hpp:89-186).  The
//     kernel solves the nx row splines (zy at columns j0, j1), two column
//     splines of za (zx) and two of zy (zxy): nx + 4 solves instead of
//     2 * ny + nx.  Each solve is the reference's own sequence of operations.
//   * The forward-elimination factors of the Thomas solve
//     (cspline.hpp:53-81) depend on the abscissae only; they are computed once
//     per axis and reused by every solve on that axis.
#pragma once

namespace pyi {

// Axis::find_indexes(c, half, boundary) axis.hpp:814-881 for the two boundary
// modes the windowed configuration exposes (undef, shrink).  buf holds 2*half
// entries; the window is buf[first .. first+count).
template <int MAXW>
__device__ __forceinline__ bool window_indexes(const DAxis& a, double c, int half, int boundary,
                                               int* buf, int& first, int& count, int& c0,
                                               int& c1) {
  int p0 = 0, p1 = 0;
  const int n = a.n;
  if (!find_indexes<double>(a, c, p0, p1)) {
    if (n != 1) {
      return false;
    }
    for (int k = 0; k < 2 * half; ++k) buf[k] = 0;  // singleton axis, axis.hpp:823-832
    first = 0;
    count = 2 * half;
    c0 = 0;
    c1 = 0;
    return true;
  }
  int front = half - 1, back = half;
  buf[front] = p0;
  buf[back] = p1;
  c0 = 0;
  c1 = 1;
  for (int shift = 1; shift < half; ++shift) {
    int before = p0 - shift;
    if (before < 0) {
      before = a.periodic ? static_cast<int>(remainder_i64(before, n)) : -1;
    }
    if (before >= 0) {
      buf[--front] = before;
      ++c0;
      ++c1;
    } else if (boundary != PYI_BOUNDARY_SHRINK) {
      return false;
    }
    int after = p1 + shift;
    if (after >= n) {
      after = a.periodic ? static_cast<int>(remainder_i64(after, n)) : -1;
    }
    if (after >= 0) {
      buf[++back] = after;
    } else if (boundary != PYI_BOUNDARY_SHRINK) {
      return false;
    }
  }
  first = front;
  count = back - front + 1;
  return true;
}

// BracketFinder::search bracket_finder.hpp:40-78
__device__ __forceinline__ bool bracket_search(const double* xa, int n, double x, int& o0,
                                               int& o1) {
  if (n == 0) return false;
  const bool asc = n == 1 || xa[0] <= xa[n - 1];
  int lo = 0, len = n;
  while (len > 0) {
    const int half = len >> 1;
    const double v = xa[lo + half];
    const bool before = asc ? (v < x) : (v > x);
    lo = before ? lo + half + 1 : lo;
    len = before ? len - half - 1 : half;
  }
  if (lo == n) return false;
  if (lo == 0) {
    o0 = 0;
    o1 = 1;
    return xa[0] == x && n > 1;
  }
  o0 = lo - 1;
  o1 = lo;
  return asc ? (xa[o0] <= x && x <= xa[o1]) : (xa[o0] >= x && x >= xa[o1]);
}

// Abscissa-only part of CSpline::compute_coefficients + the forward sweep of
// solve_symmetric_tridiagonal (cspline.hpp:83-140, 53-81).
template <int MAXW>
struct SplineAxis {
  double g[MAXW];      // 1 / h_k, h_k = xa[k+1] - xa[k] (0 when h_k == 0)
  double lower[MAXW];  // A(i, i-1) = h_i
  double cfac[MAXW];   // c_(i)
  double inv_m[MAXW];  // 1 / (A(i,i) - A(i,i-1) c_(i-1)); inv_m[0] = 1 / A(0,0)
  double last_den;     // A(m-1,m-1) - A(m-1,m-2) c_(m-2)
  int n;

  __device__ __forceinline__ void prepare(const double* xa, int n_) {
    n = n_;
    if (n < 4) return;  // CSpline::min_size, cspline.hpp:29
    const int m = n - 2;
    for (int k = 0; k < n - 1; ++k) {
      const double h = xa[k + 1] - xa[k];
      lower[k] = h;
      g[k] = (h != 0.0) ? 1.0 / h : 0.0;
    }
    // row i: lower = h_i, diag = 2 (h_i + h_{i+1}), upper = h_{i+1}
    double item = 1.0 / (2.0 * (lower[0] + lower[1]));
    inv_m[0] = item;
    cfac[0] = lower[1] * item;
    for (int i = 1; i < m - 1; ++i) {
      const double diag = 2.0 * (lower[i] + lower[i + 1]);
      const double im = 1.0 / (diag - lower[i] * cfac[i - 1]);
      inv_m[i] = im;
      cfac[i] = lower[i + 1] * im;
    }
    const double diag = 2.0 * (lower[m - 1] + lower[m]);
    last_den = diag - lower[m - 1] * cfac[m - 2];
  }

  // Natural-spline coefficients x_ for ordinates ya (c2[0] = c2[n-1] = 0).
  __device__ __forceinline__ void solve(const double* xa, const double* ya, double* c2) const {
    const int m = n - 2;
    double d[MAXW];
    double* x = c2 + 1;
    c2[0] = 0.0;
    c2[n - 1] = 0.0;
    double bi = 0.0;
    for (int i = 0; i < m; ++i) {
      const double y_i0 = ya[i + 1] - ya[i];
      const double y_i1 = ya[i + 2] - ya[i + 1];
      bi = 3.0 * fma(y_i1, g[i + 1], -y_i0 * g[i]);
      if (i == 0) {
        d[0] = bi * inv_m[0];
      } else if (i < m - 1) {
        d[i] = fma(-lower[i], d[i - 1], bi) * inv_m[i];
      }
    }
    d[m - 1] = (bi - lower[m - 1] * d[m - 2]) / last_den;
    x[m - 1] = d[m - 1];
    for (int i = m - 2; i >= 0; --i) {
      x[i] = fma(-cfac[i], x[i + 1], d[i]);
    }
    (void)xa;
  }
};

// CSplineBase::derivative_ at the node xa[k] (cspline_base.hpp:116-142 with
// CSplineCoefficients :32-41)
__device__ __forceinline__ double spline_node_derivative(const double* xa, const double* ya,
                                                         int n, const double* c2, int k) {
  int i0, i1;
  const double x = xa[k];
  if (!bracket_search(xa, n, x, i0, i1)) {
    return CUDART_NAN;
  }
  const double dx = xa[i1] - xa[i0];
  const double dy = ya[i1] - ya[i0];
  const double h = x - xa[i0];
  const double third = 1.0 / 3.0;
  const double c0 = c2[i0];
  const double c1 = c2[i1];
  const double b = (dy / dx) - dx * (c1 + 2.0 * c0) * third;
  const double dd = (c1 - c0) / (3.0 * dx);
  return b + h * fma(3.0 * h, dd, 2.0 * c0);
}

// derivative of the natural spline through (xa, ya) at nodes ka and kb
template <int MAXW>
__device__ __forceinline__ void spline_two_nodes(const SplineAxis<MAXW>& sa, const double* xa,
                                                 const double* ya, int ka, int kb, double& da,
                                                 double& db) {
  if (sa.n < 4) {  // Univariate::derivative: compute_coefficients false -> NaN
    da = CUDART_NAN;
    db = CUDART_NAN;
    return;
  }
  double c2[MAXW];
  sa.solve(xa, ya, c2);
  da = spline_node_derivative(xa, ya, sa.n, c2, ka);
  db = spline_node_derivative(xa, ya, sa.n, c2, kb);
}

// Bicubic::interpolate_ bicubic.hpp:89-186 on the bracketing cell
__device__ __forceinline__ double bicubic_cell(double x, double y, double x0, double x1,
                                               double y0, double y1, const double z[4],
                                               const double zx_[4], const double zy_[4],
                                               const double zxy_[4]) {
  // index 0: (i0,j0) 1: (i0,j1) 2: (i1,j0) 3: (i1,j1)
  const double dx = x1 - x0;
  const double dy = y1 - y0;
  const double dxdy = dx * dy;
  const double t = (x - x0) / dx;
  const double u = (y - y0) / dy;
  const double z00 = z[0], z01 = z[1], z10 = z[2], z11 = z[3];
  const double zx00 = zx_[0] * dx, zx01 = zx_[1] * dx, zx10 = zx_[2] * dx, zx11 = zx_[3] * dx;
  const double zy00 = zy_[0] * dy, zy01 = zy_[1] * dy, zy10 = zy_[2] * dy, zy11 = zy_[3] * dy;
  const double zxy00 = zxy_[0] * dxdy, zxy01 = zxy_[1] * dxdy;
  const double zxy10 = zxy_[2] * dxdy, zxy11 = zxy_[3] * dxdy;
  const double t2 = t * t, t3 = t2 * t;
  const double u2 = u * u, u3 = u2 * u;
  const double p0 = z00 + u * zy00 + u2 * fma(3.0, z01 - z00, -2.0 * zy00 - zy01) +
                    u3 * (2.0 * (z00 - z01) + zy00 + zy01);
  const double p1 = zx00 + u * zxy00 + u2 * fma(3.0, zx01 - zx00, -2.0 * zxy00 - zxy01) +
                    u3 * (2.0 * (zx00 - zx01) + zxy00 + zxy01);
  const double a20 = fma(3.0, z10 - z00, -2.0 * zx00 - zx10);
  const double a21 = fma(3.0, zy10 - zy00, -2.0 * zxy00 - zxy10);
  const double a22 = 9.0 * (z00 - z01 - z10 + z11) + 6.0 * (zx00 - zx01 + zy00 - zy10) +
                     3.0 * (zx10 - zx11 + zy01 - zy11) + 4.0 * zxy00 + 2.0 * (zxy01 + zxy10) +
                     zxy11;
  const double a23 = 6.0 * (z01 - z00 + z10 - z11) + 4.0 * (zx01 - zx00) +
                     3.0 * (zy10 - zy00 - zy01 + zy11) + 2.0 * (zx11 - zx10 - zxy00 - zxy01) -
                     zxy10 - zxy11;
  const double p2 = a20 + u * a21 + u2 * a22 + u3 * a23;
  const double a30 = 2.0 * (z00 - z10) + zx00 + zx10;
  const double a31 = zxy00 + zxy10 + 2.0 * (zy00 - zy10);
  const double a32 = 6.0 * (z01 - z00 + z10 - z11) + 4.0 * (-zy00 + zy10) +
                     3.0 * (zx01 - zx00 - zx10 + zx11) + 2.0 * (zy11 - zy01 - zxy00 - zxy10) -
                     zxy01 - zxy11;
  const double a33 = 4.0 * (z00 - z01 - z10 + z11) +
                     2.0 * (zx00 - zx01 + zx10 - zx11 + zy00 + zy01 - zy10 - zy11) + zxy00 +
                     zxy01 + zxy10 + zxy11;
  const double p3 = a30 + u * a31 + u2 * a32 + u3 * a33;
  return p0 + t * p1 + t2 * p2 + t3 * p3;
}

// One query point of windowed::pybind::bivariate
// (pybind/windowed/bivariate.hpp:46-76 + cache_loader.hpp:46-172).
// Axis::contains axis.hpp:405-416
__device__ __forceinline__ bool axis_contains(const DAxis& a, double c) {
  return !isnan(c) && (a.periodic || (c >= a.min && c <= a.max));
}
__device__ __forceinline__ bool axis_contains(const DAxis& a, long long c) {
  return c != LLONG_MIN && (a.periodic || (c >= a.imin && c <= a.imax));
}

// Axis::find_indexes(c, 1, boundary) axis.hpp:814-881: the bracketing pair, or
// (0, 0) on a singleton axis
template <typename T>
__device__ __forceinline__ bool window_pair(const DAxis& a, T c, int& i0, int& i1) {
  if (find_indexes<T>(a, c, i0, i1)) return true;
  if (a.n != 1) return false;
  i0 = 0;
  i1 = 0;
  return true;
}

// The "recompute me" tags (quiet NaNs with a payload): results are either
// finite, infinite, the canonical NaN the kernels write, or the default NaN
// of an invalid operation -- never these.
template <typename R>
struct RedoTag;
template <>
struct RedoTag<double> {
  static __device__ __forceinline__ double value() {
    return __longlong_as_double(0x7ff8b200b200b200LL);
  }
  static __device__ __forceinline__ bool is(double v) {
    return __double_as_longlong(v) == 0x7ff8b200b200b200LL;
  }
};
template <>
struct RedoTag<float> {
  static __device__ __forceinline__ float value() { return __int_as_float(0x7fcb2000); }
  static __device__ __forceinline__ bool is(float v) { return __float_as_int(v) == 0x7fcb2000; }
};

// Returns -1 (value in res, possibly NaN) or the axis that is truly out of
// bounds (cache_loader.hpp:69-80).
template <int MAXW>
__device__ __forceinline__ int windowed_point(const WinParams& p, double x, double y,
                                              double& res) {
  res = CUDART_NAN;
  int bx[MAXW], by[MAXW];
  int fx, fy, nx, ny, cx0, cx1, cy0, cy1;
  const DAxis& ax = p.ax[0];
  const DAxis& ay = p.ax[1];
  if (!window_indexes<MAXW>(ax, x, p.half_x, p.boundary, bx, fx, nx, cx0, cx1)) {
    // Axis::contains axis.hpp:261-272
    const bool inside = !isnan(x) && (ax.periodic || (x >= ax.min && x <= ax.max));
    return inside ? -1 : 0;
  }
  if (!window_indexes<MAXW>(ay, y, p.half_y, p.boundary, by, fy, ny, cy0, cy1)) {
    const bool inside = !isnan(y) && (ay.periodic || (y >= ay.min && y <= ay.max));
    return inside ? -1 : 1;
  }
  const int* wx = bx + fx;
  const int* wy = by + fy;
  double xa[MAXW], ya[MAXW];
  // x coordinates re-centred on the query when the axis is periodic
  // (cache_loader.hpp:110-133, math.hpp:131-152)
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
  for (int k = 0; k < ny; ++k) {
    ya[k] = axis_coord_f64(ay, wy[k]);
  }
  // BivariateBase::prepare: fewer than 2 nodes on an axis -> invalid -> NaN
  // (interpolate/bivariate.hpp:56-59,133-138)
  const bool sizes_ok = nx >= 2 && ny >= 2;
  int i0 = 0, i1 = 0, j0 = 0, j1 = 0;
  const bool found = sizes_ok && bracket_search(xa, nx, x, i0, i1) &&
                     bracket_search(ya, ny, y, j0, j1);
  const bool bicubic = p.method == PYI_W_BICUBIC;

  SplineAxis<MAXW> sy;
  if (bicubic && found) sy.prepare(ya, ny);

  // load the window row by row; any NaN invalidates it (cache.hpp:165-171)
  double zc0[MAXW], zc1[MAXW];    // za(:, j0), za(:, j1)
  double zyc0[MAXW], zyc1[MAXW];  // zy(:, j0), zy(:, j1)
  bool any_nan = false;
  const long long n1 = ay.n;
  for (int i = 0; i < nx; ++i) {
    double row[MAXW];
    const long long base = wx[i] * n1;
    for (int j = 0; j < ny; ++j) {
      const double v = load_value<-1, double>(p.data, p.dtype, base + wy[j]);
      row[j] = v;
      any_nan |= isnan(v);
    }
    zc0[i] = row[j0];
    zc1[i] = row[j1];
    if (bicubic && found && !any_nan) {
      spline_two_nodes<MAXW>(sy, ya, row, j0, j1, zyc0[i], zyc1[i]);
    }
  }
  if (any_nan || !found) {
    return -1;  // NaN
  }
  if (!bicubic) {
    // windowed Bilinear::interpolate_ bilinear.hpp:34-83 (explicit fma)
    const double t = (x - xa[i0]) / (xa[i1] - xa[i0]);
    const double u = (y - ya[j0]) / (ya[j1] - ya[j0]);
    const double tc = 1.0 - t;
    const double uc = 1.0 - u;
    const double bottom = fma(t, zc0[i1], tc * zc0[i0]);
    const double top = fma(t, zc1[i1], tc * zc1[i0]);
    res = fma(u, top, uc * bottom);
    return -1;
  }
  SplineAxis<MAXW> sx;
  sx.prepare(xa, nx);
  double z[4] = {zc0[i0], zc1[i0], zc0[i1], zc1[i1]};
  double zy_[4] = {zyc0[i0], zyc1[i0], zyc0[i1], zyc1[i1]};
  double zx_[4], zxy_[4];
  spline_two_nodes<MAXW>(sx, xa, zc0, i0, i1, zx_[0], zx_[2]);
  spline_two_nodes<MAXW>(sx, xa, zc1, i0, i1, zx_[1], zx_[3]);
  spline_two_nodes<MAXW>(sx, xa, zyc0, i0, i1, zxy_[0], zxy_[2]);
  spline_two_nodes<MAXW>(sx, xa, zyc1, i0, i1, zxy_[1], zxy_[3]);
  res = bicubic_cell(x, y, xa[i0], xa[i1], ya[j0], ya[j1], z, zx_, zy_, zxy_);
  return -1;
}

#ifndef PYI_WINDOWED_DEVICE_ONLY  // kernel and launchers: pyi_windowed.cu only
template <int MAXW, typename R>
__global__ void __launch_bounds__(128)
windowed2d_kernel(const __grid_constant__ WinParams p) {
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  unsigned long long key = kNoOob;
  R* __restrict__ out = static_cast<R*>(p.out);
  for (long long k = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; k < p.n;
       k += stride) {
    double res;
    const int bad = windowed_point<MAXW>(p, __ldg(p.x + k), __ldg(p.y + k), res);
    if (bad >= 0) {
      const unsigned long long kk =
          (static_cast<unsigned long long>(p.index_base + k) << 2) | bad;
      key = kk < key ? kk : key;
    }
    __stcs(out + k, static_cast<R>(res));
  }
  if (p.bounds_error && key != kNoOob) {
    atomicMin(p.oob_key, key);
  }
}

template <int MAXW>
static cudaError_t launch_windowed_w(const WinParams& p, cudaStream_t stream) {
  const int blocks = grid_for(p.n, 128, 8);
  const bool r64 = p.dtype == PYI_F64 || p.dtype == PYI_I64 || p.dtype == PYI_U64;
  if (r64) {
    windowed2d_kernel<MAXW, double><<<blocks, 128, 0, stream>>>(p);
  } else {
    windowed2d_kernel<MAXW, float><<<blocks, 128, 0, stream>>>(p);
  }
  g_launches.fetch_add(1);
  return cudaGetLastError();
}

static cudaError_t launch_bicubic6_fast(const WinParams& p, cudaStream_t stream);  // pyi_fastwin.cuh

cudaError_t launch_windowed2d(const WinParams& p, cudaStream_t stream) {
  if (p.n <= 0) return cudaSuccess;
  if (p.method == PYI_W_BICUBIC && p.half_x == 3 && p.half_y == 3 && !force_generic()) {
    return launch_bicubic6_fast(p, stream);
  }
  const int w = 2 * (p.half_x > p.half_y ? p.half_x : p.half_y);
  if (w <= 6) return launch_windowed_w<6>(p, stream);
  if (w <= 12) return launch_windowed_w<12>(p, stream);
  if (w <= 32) return launch_windowed_w<32>(p, stream);
  return cudaErrorInvalidValue;
}

#endif  // PYI_WINDOWED_DEVICE_ONLY

}  // namespace pyi
