// pyi_fastwin_nd.cuh -- K6 fast path: bicubic (6 x 6 window) on 3-D / 4-D grids
// = windowed::pybind::trivariate_single / quadrivariate_single
// (pybind/windowed/trivariate.hpp:66-123, quadrivariate.hpp:50-148) with the
// bicubic fitting model: the K4 fast arithmetic (pyi_fastwin.cuh) on each of
// the 2 (z) or 4 (z, u) bracketing planes, then linear() / nearest() along z
// and u exactly as the generic K6 kernel combines them.
//
// Layout: the planes are read from a PLANE-MAJOR copy of the grid
// ([z][u][x][y], built once per grid by plane_major_kernel), so that the six
// rows of a window are six contiguous 48-byte runs like on a 2-D grid; in the
// caller's C order ([x][y][z][u]) the 36 values of one plane are 36 separate
// sectors and the kernel would be bound by the gather request rate.
//
// Points the fast arithmetic does not cover (incomplete windows at the edges,
// out-of-bounds coordinates, values outside the range the shared-reciprocal
// divisions are proven for) are tagged with a NaN payload no computation
// produces; the generic K6 kernel then runs in "redo" mode over the output and
// recomputes exactly those, including the bounds-error report.  A NaN inside a
// complete window gives NaN directly (cache.hpp:165-171).
//
// Included by pyi_windowed.cu after pyi_windowed.cuh and pyi_fastwin.cuh.
#pragma once

namespace pyi {

constexpr int kWinNdSlots = 42;  // K4's 38 + the four plane results

// Row pass of K4 (bicubic6_rows) on one plane of the plane-major copy; also
// reports whether a value outside the proven range was a NaN.
template <int DT>
__device__ __forceinline__ void bicubic6_rows_plane(const WinNdParams& p,
                                                    const double (&sy)[S6_COUNT], bool ok,
                                                    int wx0, int wy0, long long plane_offset,
                                                    double* __restrict__ sm, bool& fine,
                                                    bool& has_nan) {
  const int nx = p.ax[0].n;
  const long long n1 = p.ax[1].n;
  double nxt[6];
  {
    const long long base = plane_offset + (wx0 >= nx ? wx0 - nx : wx0) * n1 + wy0;
#pragma unroll
    for (int j = 0; j < 6; ++j) {
      nxt[j] = ok ? load_value<DT, double>(p.planes, p.dtype, base + j) : 0.0;
    }
  }
#pragma unroll 1
  for (int i0 = 0; i0 < 6; ++i0) {
    double cur[6];
#pragma unroll
    for (int j = 0; j < 6; ++j) cur[j] = nxt[j];
    if (i0 + 1 < 6) {
      const int i = wx0 + i0 + 1;
      const long long base = plane_offset + (i >= nx ? i - nx : i) * n1 + wy0;
#pragma unroll
      for (int j = 0; j < 6; ++j) {
        nxt[j] = ok ? load_value<DT, double>(p.planes, p.dtype, base + j) : 0.0;
      }
    }
    if (!row_values_ok<DT>(cur)) {
      fine = false;
#pragma unroll
      for (int j = 0; j < 6; ++j) has_nan = has_nan || cur[j] != cur[j];
    }
    double d2, d3;
    spline6_node_derivatives(sy, cur, d2, d3);
    double* row = sm + i0 * 4 * kWinThreads;
    row[0] = cur[2];
    row[kWinThreads] = cur[3];
    row[2 * kWinThreads] = d2;
    row[3 * kWinThreads] = d3;
  }
}

// One query point.  Returns false when the point has to be recomputed by the
// generic kernel.
template <int DT, int ND, typename ZT, typename R>
__device__ __forceinline__ bool bicubic6_nd_point(const WinNdParams& p, double x, double y, ZT z,
                                                  double u, R& res, double* __restrict__ sm) {
  const DAxis& ax = p.ax[0];
  const DAxis& ay = p.ax[1];
  int wx0 = 0, wy0 = 0, iz0 = 0, iz1 = 0, iu0 = 0, iu1 = 0;
  // norm_fast needs a positive period and quotients below 2^48
  bool ok = sane(x) && sane(y) && (ax.period == 0.0 || ax.period > 0.0078125);
  ok = ok && window6(ax, x, wx0) && window6(ay, y, wy0);
  ok = ok && !ay.periodic;  // rows of the window are contiguous in memory
  ok = ok && window_pair<ZT>(p.ax[2], z, iz0, iz1);
  if (ND == 4) ok = ok && window_pair<double>(p.ax[3], u, iu0, iu1);
  if (!PYI_ANY(ok)) return false;
  wx0 = ok ? wx0 : 0;
  wy0 = ok ? wy0 : 0;
  const int nx = ax.n;
  // coordinates (cache_loader.hpp:110-133): x re-centred on the query
  double xa[6], ya[6];
  const bool per = ax.periodic != 0 && ax.period != 0.0;
  const double half_p = __dmul_rn(ax.period, 0.5);
  const double lo = __dsub_rn(x, half_p);
  const double hi = __dadd_rn(x, half_p);
#pragma unroll
  for (int k = 0; k < 6; ++k) {
    const int i = wx0 + k;
    double raw = ok ? axis_coord_f64(ax, i >= nx ? i - nx : i) : 0.0;
    ok = ok && sane(raw);
    if (per && !(raw >= lo && raw < hi)) {
      raw = norm_fast(raw, lo, ax.period, ax.inv_period);
    }
    xa[k] = raw;
    ya[k] = ok ? axis_coord_f64(ay, wy0 + k) : 0.0;
  }
  ok = ok && window_usable(xa, x) && window_usable(ya, y);
  if (!PYI_ANY(ok)) return false;
#pragma unroll
  for (int k = 0; k < 6; ++k) sm[(32 + k) * kWinThreads] = xa[k];
  const double ya2 = ya[2], ya3 = ya[3];
  const bool uni_y = p.u6_ok[1] != 0 && PYI_ALL(!ok || spacing_is(ya, ay.step));
  const bool uni_x = p.u6_ok[0] != 0 && PYI_ALL(!ok || spacing_is(xa, ax.step));
  // non-uniform axes: the 24 spline constants are recomputed per plane (they
  // would not survive the plane loop in registers)
  const long long plane_cells = static_cast<long long>(nx) * ay.n;
  const long long n3 = ND == 4 ? p.ax[3].n : 1;
  constexpr int kPlanes = ND == 4 ? 4 : 2;
  bool fine = true, has_nan = false;
#pragma unroll 1
  for (int t = 0; t < kPlanes; ++t) {
    const int pz = (ND == 4 ? (t >> 1) : t) ? iz1 : iz0;
    const int pu = ND == 4 ? ((t & 1) ? iu1 : iu0) : 0;
    const long long plane_offset = (pz * n3 + pu) * plane_cells;
    if (uni_y) {
      bicubic6_rows_plane<DT>(p, p.u6[1], ok, wx0, wy0, plane_offset, sm, fine, has_nan);
    } else {
      double yb[6];
#pragma unroll
      for (int k = 0; k < 6; ++k) yb[k] = ok ? axis_coord_f64(ay, wy0 + k) : static_cast<double>(k);
      double sy[S6_COUNT];
      spline6_prepare(yb, sy);
      bicubic6_rows_plane<DT>(p, sy, ok, wx0, wy0, plane_offset, sm, fine, has_nan);
    }
    double xb[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) xb[k] = sm[(32 + k) * kWinThreads];
    if (uni_x) {
      bicubic6_columns(p.u6[0], sm);
    } else {
      double sx[S6_COUNT];
      spline6_prepare(xb, sx);
      bicubic6_columns(sx, sm);
    }
    // index 0: (i0,j0) 1: (i0,j1) 2: (i1,j0) 3: (i1,j1), i0 = j0 = 2
    const double zz[4] = {sm[8 * kWinThreads], sm[9 * kWinThreads], sm[12 * kWinThreads],
                          sm[13 * kWinThreads]};
    const double zy_[4] = {sm[10 * kWinThreads], sm[11 * kWinThreads], sm[14 * kWinThreads],
                           sm[15 * kWinThreads]};
    const double zx_[4] = {sm[24 * kWinThreads], sm[26 * kWinThreads], sm[25 * kWinThreads],
                           sm[27 * kWinThreads]};
    const double zxy_[4] = {sm[28 * kWinThreads], sm[30 * kWinThreads], sm[29 * kWinThreads],
                            sm[31 * kWinThreads]};
    sm[(38 + t) * kWinThreads] = bicubic_cell(x, y, xb[2], xb[3], ya2, ya3, zz, zx_, zy_, zxy_);
  }
  if (!ok) return false;
  if (!fine) {
    if (!has_nan) return false;
    res = RT<R>::nan();  // a NaN in a complete window (cache.hpp:165-171)
    return true;
  }
  const ZT z0 = axis_coord(p.ax[2], iz0, ZT{});
  const ZT z1 = axis_coord(p.ax[2], iz1, ZT{});
  if (ND == 3) {
    const double f0 = sm[38 * kWinThreads], f1 = sm[39 * kWinThreads];
    res = static_cast<R>(axis_interp<ZT, double>(p.z_method, z, z0, z1, f0, f1));
    return true;
  }
  // planes (z0,u0), (z0,u1), (z1,u0), (z1,u1) (quadrivariate.hpp:126-146)
  const double f00 = sm[38 * kWinThreads], f01 = sm[39 * kWinThreads];
  const double f10 = sm[40 * kWinThreads], f11 = sm[41 * kWinThreads];
  const double u0 = axis_coord_f64(p.ax[3], iu0);
  const double u1 = axis_coord_f64(p.ax[3], iu1);
  const R f0 = static_cast<R>(axis_interp<ZT, double>(p.z_method, z, z0, z1, f00, f10));
  const R f1 = static_cast<R>(axis_interp<ZT, double>(p.z_method, z, z0, z1, f01, f11));
  res = axis_interp<double, R>(p.u_method, u, u0, u1, f0, f1);
  return true;
}

template <int DT, int ND, typename ZT, typename R>
__global__ void __launch_bounds__(kWinThreads, 4)
bicubic6_nd_kernel(const __grid_constant__ WinNdParams p) {
  __shared__ double sm[kWinNdSlots * kWinThreads];
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  R* __restrict__ out = static_cast<R*>(p.out);
  const ZT* __restrict__ zs = static_cast<const ZT*>(p.z);
  for (long long k = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; k < p.n;
       k += stride) {
    R res;
    const double u = ND == 4 ? __ldcs(p.u + k) : 0.0;
    const bool done = bicubic6_nd_point<DT, ND, ZT, R>(p, __ldcs(p.x + k), __ldcs(p.y + k),
                                                       __ldcs(zs + k), u, res, sm + threadIdx.x);
    __stcs(out + k, done ? res : RedoTag<R>::value());
  }
}

// Plane-major copy: out[((iz * n3 + iu) * nx + ix) * ny + iy] = in[((ix * ny + iy) * n2 + iz) * n3 + iu]
template <typename E>
__global__ void __launch_bounds__(kThreads)
plane_major_kernel(const E* __restrict__ in, E* __restrict__ out, long long nx, long long ny,
                   long long n2, long long n3) {
  const long long cells = nx * ny * n2 * n3;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  const long long plane = nx * ny;
  for (long long k = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; k < cells;
       k += stride) {
    // k runs over the INPUT (coalesced reads, scattered 1..8-byte writes)
    const long long iu = k % n3;
    const long long r = k / n3;
    const long long iz = r % n2;
    const long long xy = r / n2;
    out[(iz * n3 + iu) * plane + xy] = in[k];
  }
}

cudaError_t launch_plane_major(const void* in, void* out, int elem_size, const long long* shape,
                               cudaStream_t stream) {
  const long long cells = shape[0] * shape[1] * shape[2] * shape[3];
  if (cells <= 0) return cudaSuccess;
  const int blocks = grid_for(cells, kThreads, 8);
  switch (elem_size) {
    case 1:
      plane_major_kernel<unsigned char><<<blocks, kThreads, 0, stream>>>(
          static_cast<const unsigned char*>(in), static_cast<unsigned char*>(out), shape[0],
          shape[1], shape[2], shape[3]);
      break;
    case 2:
      plane_major_kernel<unsigned short><<<blocks, kThreads, 0, stream>>>(
          static_cast<const unsigned short*>(in), static_cast<unsigned short*>(out), shape[0],
          shape[1], shape[2], shape[3]);
      break;
    case 4:
      plane_major_kernel<unsigned int><<<blocks, kThreads, 0, stream>>>(
          static_cast<const unsigned int*>(in), static_cast<unsigned int*>(out), shape[0],
          shape[1], shape[2], shape[3]);
      break;
    default:
      plane_major_kernel<unsigned long long><<<blocks, kThreads, 0, stream>>>(
          static_cast<const unsigned long long*>(in), static_cast<unsigned long long*>(out),
          shape[0], shape[1], shape[2], shape[3]);
      break;
  }
  g_launches.fetch_add(1);
  return cudaGetLastError();
}

template <int ND, typename ZT>
static cudaError_t launch_bicubic6_nd_z(const WinNdParams& p, cudaStream_t stream) {
  const int blocks = grid_for(p.n, kWinThreads, 4);
  switch (p.dtype) {
    case PYI_F64:
      bicubic6_nd_kernel<PYI_F64, ND, ZT, double><<<blocks, kWinThreads, 0, stream>>>(p);
      break;
    case PYI_F32:
      bicubic6_nd_kernel<PYI_F32, ND, ZT, float><<<blocks, kWinThreads, 0, stream>>>(p);
      break;
    case PYI_I64:
    case PYI_U64:
      bicubic6_nd_kernel<-1, ND, ZT, double><<<blocks, kWinThreads, 0, stream>>>(p);
      break;
    default:
      bicubic6_nd_kernel<-1, ND, ZT, float><<<blocks, kWinThreads, 0, stream>>>(p);
      break;
  }
  g_launches.fetch_add(1);
  return cudaGetLastError();
}

// The fast pass; the caller follows it with the generic kernel in redo mode.
static cudaError_t launch_bicubic6_nd_fast(const WinNdParams& p, cudaStream_t stream) {
  const bool zi = p.ax[2].coord_i64 != 0;
  if (p.ndim == 3) {
    return zi ? launch_bicubic6_nd_z<3, long long>(p, stream)
              : launch_bicubic6_nd_z<3, double>(p, stream);
  }
  return zi ? launch_bicubic6_nd_z<4, long long>(p, stream)
            : launch_bicubic6_nd_z<4, double>(p, stream);
}

}  // namespace pyi
