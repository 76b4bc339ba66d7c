// pyi_kernels.cu -- hand-written sm_100a kernels of the n-variate regular-grid
// interpolation path.  Compiled with -fmad=false: the only fused operations are
// the explicit fma() calls that mirror the reference's std::fma.
//
// Kernels:
//   geometric_kernel<ND,DT,R,ZI64>  K1-K3: axis lookups, 2^ND-corner gather and
//       the bilinear / idw / nearest evaluators of
//       cxx/include/pyinterp/pybind/geometric/{bi,tri,quadri}variate.hpp and
//       math/interpolate/geometric/{bivariate,multivariate}.hpp
//   windowed2d_kernel<MAXW>         K4: window build + bicubic / windowed
//       bilinear of pybind/windowed/bivariate.hpp, math/interpolate/
//       cache_loader.hpp, bivariate/bicubic.hpp, univariate/cspline.hpp
//   find_indexes_kernel / find_index_kernel   K5: Axis::find_indexes / find_index
//   pack_cells2d_kernel             one-off cell packing of a 2-D grid
#include "pyi_common.cuh"

namespace pyi {

std::atomic<long long> g_launches{0};
long long launch_count() { return g_launches.load(); }

int sm_count() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

int grid_for(long long n, int per_block, int waves) {
  const long long need = (n + per_block - 1) / per_block;
  const long long cap = static_cast<long long>(sm_count()) * waves;
  return static_cast<int>(need < 1 ? 1 : (need < cap ? need : cap));
}

// ---------------------------------------------------------------------------
// geometric evaluators (math/interpolate/geometric/bivariate.hpp)
// ---------------------------------------------------------------------------
// Boost.Geometry comparable_distance on spherical_equatorial<degree> points:
// comparable haversine on the unit sphere in the coordinate type.
template <typename R>
__device__ __forceinline__ R hav(R theta) {
  const R s = RT<R>::sin_(theta / R(2));
  return s * s;
}
template <typename R>
__device__ __forceinline__ R comparable_distance(R lon1d, R lat1d, R lon2d, R lat2d) {
  const R d2r = RT<R>::pi() / R(180);
  const R lon1 = lon1d * d2r;
  const R lat1 = lat1d * d2r;
  const R lon2 = lon2d * d2r;
  const R lat2 = lat2d * d2r;
  return hav<R>(lat2 - lat1) + RT<R>::cos_(lat1) * RT<R>::cos_(lat2) * hav<R>(lon2 - lon1);
}

// Bilinear::evaluate bivariate.hpp:110-132
template <typename R>
__device__ __forceinline__ R eval_bilinear(R x, R y, R x0, R y0, R x1, R y1, R q00, R q01, R q10,
                                           R q11) {
  const R dx = x1 - x0;
  const R dy = y1 - y0;
  const R t = (x - x0) / dx;
  const R u = (y - y0) / dy;
  const R omt = R(1) - t;
  const R omu = R(1) - u;
  return omt * (omu * q00 + u * q01) + t * (omu * q10 + u * q11);
}

// The four comparable distances of a query to the corners of its cell, in the
// corner order (x0,y0) (x0,y1) (x1,y0) (x1,y1).  Same operations on the same
// operands as four calls of comparable_distance; the seven distinct sines /
// cosines (out of sixteen) are evaluated once.
template <typename R>
__device__ __forceinline__ void corner_distances(R x, R y, R x0, R y0, R x1, R y1, R (&d2)[4]) {
  const R d2r = RT<R>::pi() / R(180);
  const R lon = x * d2r, lat = y * d2r;
  const R lat_c[2] = {y0 * d2r, y1 * d2r};
  const R lon_c[2] = {x0 * d2r, x1 * d2r};
  const R cos_lat = RT<R>::cos_(lat);
  R hav_lat[2], cos_c[2], hav_lon[2];
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    hav_lat[j] = hav<R>(lat_c[j] - lat);
    cos_c[j] = RT<R>::cos_(lat_c[j]);
    hav_lon[j] = hav<R>(lon_c[j] - lon);
  }
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int i = k >> 1, j = k & 1;
    d2[k] = hav_lat[j] + cos_lat * cos_c[j] * hav_lon[i];
  }
}

// InverseDistanceWeighting::evaluate bivariate.hpp:189-232
template <typename R>
__device__ __forceinline__ R eval_idw(int exp_, R x, R y, R x0, R y0, R x1, R y1, R q00, R q01,
                                      R q10, R q11) {
  const R cq[4] = {q00, q01, q10, q11};
  const R eps2 = RT<R>::eps() * RT<R>::eps();
  const double half_exp = exp_ * 0.5;
  R d2[4];
  corner_distances<R>(x, y, x0, y0, x1, y1, d2);
  R wsum = R(0);
  R vsum = R(0);
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    if (d2[k] <= eps2) {
      return cq[k];
    }
    // exponent 2 (the default, config.cpp:46): std::pow(d, 1.0) is d itself, no pow()
    const double dd = static_cast<double>(d2[k]);
    const R w = static_cast<R>(1 / (exp_ == 2 ? dd : pow(dd, half_exp)));
    wsum += w;
    vsum += cq[k] * w;
  }
  return vsum / wsum;
}

// Nearest::evaluate bivariate.hpp:276-302 (first of equal distances wins)
template <typename R>
__device__ __forceinline__ R eval_nearest(R x, R y, R x0, R y0, R x1, R y1, R q00, R q01, R q10,
                                          R q11) {
  R d2[4];
  corner_distances<R>(x, y, x0, y0, x1, y1, d2);
  R best = d2[0];
  R val = q00;
  if (d2[1] < best) { best = d2[1]; val = q01; }
  if (d2[2] < best) { best = d2[2]; val = q10; }
  if (d2[3] < best) { best = d2[3]; val = q11; }
  return val;
}

template <typename R>
__device__ __forceinline__ R eval_spatial(int method, int idw_exp, R x, R y, R x0, R y0, R x1,
                                          R y1, R q00, R q01, R q10, R q11) {
  if (method == PYI_BILINEAR) {
    return eval_bilinear<R>(x, y, x0, y0, x1, y1, q00, q01, q10, q11);
  }
  if (method == PYI_IDW) {
    return eval_idw<R>(idw_exp, x, y, x0, y0, x1, y1, q00, q01, q10, q11);
  }
  return eval_nearest<R>(x, y, x0, y0, x1, y1, q00, q01, q10, q11);
}


// ---------------------------------------------------------------------------
// K1-K3: geometric interpolation (pyi_geometric.cuh)
// ---------------------------------------------------------------------------
}  // namespace pyi
#include "pyi_geometric.cuh"
namespace pyi {

template <int ND, bool ZI64>
static cudaError_t launch_geometric_nd(const GeoParams& p, cudaStream_t stream) {
  const int blocks = grid_for(p.n, kThreads, 8);
  const bool r64 = p.dtype == PYI_F64 || p.dtype == PYI_I64 || p.dtype == PYI_U64;
  if (p.packed != nullptr) {
    // cell-packed copy (already in the result type): one aligned run per point
    // cells of 64 / 128 bytes are read by lane pairs / quads (one request per cell)
    if (r64) {
      if constexpr (ND >= 3) {
        geometric_group_kernel<ND, double, ZI64><<<blocks, kThreads, 0, stream>>>(p);
      } else {
        geometric_kernel<ND, -1, double, ZI64, true><<<blocks, kThreads, 0, stream>>>(p);
      }
    } else {
      if constexpr (ND >= 4) {
        geometric_group_kernel<ND, float, ZI64><<<blocks, kThreads, 0, stream>>>(p);
      } else {
        geometric_kernel<ND, -1, float, ZI64, true><<<blocks, kThreads, 0, stream>>>(p);
      }
    }
  } else if (p.dtype == PYI_F64) {
    geometric_kernel<ND, PYI_F64, double, ZI64, false><<<blocks, kThreads, 0, stream>>>(p);
  } else if (p.dtype == PYI_F32) {
    geometric_kernel<ND, PYI_F32, float, ZI64, false><<<blocks, kThreads, 0, stream>>>(p);
  } else if (r64) {
    geometric_kernel<ND, -1, double, ZI64, false><<<blocks, kThreads, 0, stream>>>(p);
  } else {
    geometric_kernel<ND, -1, float, ZI64, false><<<blocks, kThreads, 0, stream>>>(p);
  }
  g_launches.fetch_add(1);
  return cudaGetLastError();
}

// K1 fast path (pyi_fast2d.cuh)
static bool fast2d_eligible(const GeoParams& p);
static cudaError_t launch_fast2d(const GeoParams& p, cudaStream_t stream);
// K1 in tolerance value mode (pyi_fast2d_tol.cuh)
static bool tol2d_eligible(const GeoParams& p);
static cudaError_t launch_tol2d(const GeoParams& p, cudaStream_t stream);
// K2 / K3 fast path (pyi_fastnd.cuh)
static bool fastnd_eligible(const GeoParams& p);
static cudaError_t launch_fastnd(const GeoParams& p, cudaStream_t stream);

cudaError_t launch_geometric(const GeoParams& p, cudaStream_t stream) {
  if (p.n <= 0) return cudaSuccess;
  if (tol2d_eligible(p)) return launch_tol2d(p, stream);
  if (fast2d_eligible(p)) return launch_fast2d(p, stream);
  if (fastnd_eligible(p)) return launch_fastnd(p, stream);
  const bool zi64 = p.ndim >= 3 && p.ax[2].coord_i64;
  switch (p.ndim) {
    case 2: return launch_geometric_nd<2, false>(p, stream);
    case 3: return zi64 ? launch_geometric_nd<3, true>(p, stream)
                        : launch_geometric_nd<3, false>(p, stream);
    case 4: return zi64 ? launch_geometric_nd<4, true>(p, stream)
                        : launch_geometric_nd<4, false>(p, stream);
    default: return cudaErrorInvalidValue;
  }
}

// ---------------------------------------------------------------------------
// K5: Axis::find_indexes / find_index (pybind/axis.hpp:84-101, axis.hpp:319-333)
// ---------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(kThreads)
find_indexes_kernel(const __grid_constant__ FindParams p) {
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  const T* __restrict__ c = static_cast<const T*>(p.coords);
  longlong2* __restrict__ out = reinterpret_cast<longlong2*>(p.out);
  for (long long k = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; k < p.n;
       k += stride) {
    int i0 = -1, i1 = -1;
    const bool ok = find_indexes<T>(p.ax, __ldg(c + k), i0, i1);
    out[k] = ok ? make_longlong2(i0, i1) : make_longlong2(-1, -1);
  }
}

// Axis::find_index(c, bounded).  The reference's casts of out-of-range doubles
// to int64 are reproduced as its x86-64 build executes them: NaN and |v| >=
// 2^63 become INT64_MIN, i.e. "below the axis".
template <typename T>
__device__ __forceinline__ long long find_index_bounded(const DAxis& a, T c, bool bounded);
template <>
__device__ __forceinline__ long long find_index_bounded<double>(const DAxis& a, double c,
                                                                 bool bounded) {
  if (isnan(c)) return -1;
  c = axis_normalize(a, c, a.min);
  long long r;
  if (a.kind == 0) {
    const double v = round(__dmul_rn(__dsub_rn(c, a.start), a.inv_step));
    const bool low = !(v >= 0.0) || v >= 9223372036854775808.0;
    if (low) {
      r = bounded ? 0 : -1;
    } else if (v >= static_cast<double>(a.n)) {
      r = bounded ? a.n - 1 : -1;
    } else {
      r = static_cast<long long>(v);
    }
  } else {
    r = irregular_find_f64(a, c);
    if (r < 0 && bounded) {
      // container.hpp:390-397: below the first / beyond the last point
      const double first = __ldg(static_cast<const double*>(a.points));
      r = (a.asc ? (c < first) : (c > first)) ? 0 : a.n - 1;
    }
  }
  if (r == -1 && a.periodic) {  // axis.hpp:326-331
    return __dsub_rn(c, a.max) < __dsub_rn(__dadd_rn(a.min, a.period), c) ? a.n - 1 : 0;
  }
  return r;
}
template <>
__device__ __forceinline__ long long find_index_bounded<long long>(const DAxis& a, long long c,
                                                                    bool bounded) {
  if (c == LLONG_MIN) return -1;
  c = axis_normalize(a, c, a.imin);
  long long r;
  if (a.kind == 0) {
    const long long num = static_cast<long long>(static_cast<unsigned long long>(c) -
                                                 static_cast<unsigned long long>(a.istart));
    const long long idx = num > 0 ? div_trunc_i64(num + a.istep2, a.istep)
                                  : div_trunc_i64(num - a.istep2, a.istep);
    if (idx < 0) {
      r = bounded ? 0 : -1;
    } else if (idx >= a.n) {
      r = bounded ? a.n - 1 : -1;
    } else {
      r = idx;
    }
  } else {
    r = irregular_find_i64(a, c);
    if (r < 0 && bounded) {
      const long long first = __ldg(static_cast<const long long*>(a.points));
      r = (a.asc ? (c < first) : (c > first)) ? 0 : a.n - 1;
    }
  }
  if (r == -1 && a.periodic) {
    return (c - a.imax) < (a.imin + a.iperiod - c) ? a.n - 1 : 0;
  }
  return r;
}

template <typename T>
__global__ void __launch_bounds__(kThreads)
find_index_kernel(const __grid_constant__ FindParams p) {
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  const T* __restrict__ c = static_cast<const T*>(p.coords);
  for (long long k = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; k < p.n;
       k += stride) {
    p.out[k] = find_index_bounded<T>(p.ax, __ldg(c + k), p.bounded != 0);
  }
}

cudaError_t launch_find_indexes(const FindParams& p, cudaStream_t stream) {
  if (p.n <= 0) return cudaSuccess;
  const int blocks = grid_for(p.n, kThreads, 8);
  if (p.ax.coord_i64) {
    find_indexes_kernel<long long><<<blocks, kThreads, 0, stream>>>(p);
  } else {
    find_indexes_kernel<double><<<blocks, kThreads, 0, stream>>>(p);
  }
  g_launches.fetch_add(1);
  return cudaGetLastError();
}

cudaError_t launch_find_index(const FindParams& p, cudaStream_t stream) {
  if (p.n <= 0) return cudaSuccess;
  const int blocks = grid_for(p.n, kThreads, 8);
  if (p.ax.coord_i64) {
    find_index_kernel<long long><<<blocks, kThreads, 0, stream>>>(p);
  } else {
    find_index_kernel<double><<<blocks, kThreads, 0, stream>>>(p);
  }
  g_launches.fetch_add(1);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------
// one-off cell packing: the 2^ND corners of cell (ix, iy[, iz[, iu]]) contiguous
// in the result type, corner (a, b, c, d) at ((a*2 + b)*NZ + c)*NU + d: a 2-D
// f64 / 3-D f32 cell is exactly one 32-byte sector, a 4-D f32 cell one 64-byte
// line.  ix + 1 wraps (the periodic seam cell is materialised), the other axes
// clamp (their last cell is never addressed by a valid bracket).  One thread
// per output element: coalesced stores.
// ---------------------------------------------------------------------------
template <typename R>
__global__ void __launch_bounds__(kThreads)
pack_cells_kernel(const void* __restrict__ data, int dtype, int ndim, int nx, int ny, int nz, int nu,
                  R* __restrict__ cells) {
  const int corners = 1 << ndim;
  const long long total = static_cast<long long>(nx) * ny * nz * nu * corners;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long e = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; e < total;
       e += stride) {
    int corner = static_cast<int>(e & (corners - 1));
    long long c = e >> ndim;
    int d = 0, cc = 0;
    int iu = 0, iz = 0;
    if (ndim >= 4) {
      d = corner & 1;
      corner >>= 1;
      iu = static_cast<int>(c % nu);
      c /= nu;
    }
    if (ndim >= 3) {
      cc = corner & 1;
      corner >>= 1;
      iz = static_cast<int>(c % nz);
      c /= nz;
    }
    const int b = corner & 1;
    const int a = corner >> 1;
    const int iy = static_cast<int>(c % ny);
    const int ix = static_cast<int>(c / ny);
    const int jx = (ix + a == nx) ? 0 : ix + a;
    const int jy = (iy + b == ny) ? iy : iy + b;
    const int jz = (iz + cc == nz) ? iz : iz + cc;
    const int ju = (iu + d == nu) ? iu : iu + d;
    const long long off = ((static_cast<long long>(jx) * ny + jy) * nz + jz) * nu + ju;
    cells[e] = load_value<-1, R>(data, dtype, off);
  }
}

cudaError_t launch_pack_cells(const void* data, int dtype, int ndim, const int* shape, void* packed,
                              cudaStream_t stream) {
  const int nx = shape[0], ny = shape[1];
  const int nz = ndim >= 3 ? shape[2] : 1;
  const int nu = ndim >= 4 ? shape[3] : 1;
  const long long total = (static_cast<long long>(nx) * ny * nz * nu) << ndim;
  const int blocks = grid_for(total, kThreads, 8);
  const bool r64 = dtype == PYI_F64 || dtype == PYI_I64 || dtype == PYI_U64;
  if (r64) {
    pack_cells_kernel<double><<<blocks, kThreads, 0, stream>>>(data, dtype, ndim, nx, ny, nz, nu,
                                                               static_cast<double*>(packed));
  } else {
    pack_cells_kernel<float><<<blocks, kThreads, 0, stream>>>(data, dtype, ndim, nx, ny, nz, nu,
                                                              static_cast<float*>(packed));
  }
  g_launches.fetch_add(1);
  return cudaGetLastError();
}

}  // namespace pyi

#include "pyi_fast2d.cuh"
#include "pyi_fast2d_tol.cuh"
#include "pyi_fastnd.cuh"
