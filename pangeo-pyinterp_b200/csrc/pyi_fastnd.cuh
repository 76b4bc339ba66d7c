// pyi_fastnd.cuh -- K2 / K3 fast path: trivariate / quadrivariate with the
// bilinear / idw / nearest spatial evaluator and linear or nearest z / u
// interpolation, float or double results (every grid dtype), ascending axes,
// reading the cell-packed copy.  Included by pyi_kernels.cu after pyi_fast2d.cuh.
//
// What is computed is trivariate_single / quadrivariate_single
// (pybind/geometric/trivariate.hpp:47-123, quadrivariate.hpp:50-148) with
// geometric::trivariate / quadrivariate (math/interpolate/geometric/
// multivariate.hpp:133-153, 271-306), Bilinear::evaluate
// (geometric/bivariate.hpp:110-132) and linear / nearest
// (interpolate/linear.hpp:22-33, nearest.hpp:21-32): same roundings, same
// operation order.  As in pyi_fast2d.cuh the brackets are GUESSED with a floor
// and VERIFIED against the reference's own node values; a verified bracket is
// the reference's bracket, anything else is re-done by the generic code
// (geometric_eval) behind a non-inlined call.  The generic path costs ~1100
// instructions per 4-D point (ncu, profiles/), this one a few hundred, and the
// kernels are instruction-bound once the gather is one request per cell.
//
//   x  regular ascending f64, with or without period
//   y  regular or irregular ascending f64, no period
//   z  regular ascending i64 (TemporalAxis), or regular / irregular ascending f64
//   u  regular or irregular ascending f64
#pragma once

namespace pyi {

constexpr int kNdMaxIrregular = 256;  // z / u points staged in shared memory
constexpr int kNdMaxLut = 4 * kNdMaxIrregular + 1;

// Verified bracket on a regular ascending int64 axis (|start|, span < 2^61, so
// the wrapped difference below is in [0, span) only for coordinates inside the
// axis).  k = floor((z - start) / step) estimated in double and corrected by
// at most one step with exact integer arithmetic; d0 = z - node(k), exact.
// Regular<int64>::find_index rounds to the nearest node and Axis::find_indexes
// steps by the sign of z - node (container.hpp:588-615, axis.hpp:745-766): for
// an ascending axis that is the pair (k, k + 1) with node(k) <= z < node(k+1).
// The last node itself ((n-2, n-1), axis.hpp:748-757) is left to the generic code.
__device__ __forceinline__ void bracket_regular_i64(const DAxis& a, long long z, int& k,
                                                    long long& d0, bool& ok) {
  const long long num = static_cast<long long>(static_cast<unsigned long long>(z) -
                                               static_cast<unsigned long long>(a.istart));
  const bool inside = num >= 0 && num < a.ispan;
  const long long nn = inside ? num : 0;
  int kk = __double2int_rd(__dmul_rn(static_cast<double>(nn), a.inv_step));
  long long d = nn - static_cast<long long>(kk) * a.istep;
  if (d < 0) {
    kk -= 1;
    d += a.istep;
  } else if (d >= a.istep) {
    kk += 1;
    d -= a.istep;
  }
  ok = ok && inside && d >= 0 && d < a.istep &&
       static_cast<unsigned>(kk) < static_cast<unsigned>(a.n - 1);
  k = ok ? kk : 0;
  d0 = d;
}

// regular (ascending or descending) or irregular ascending f64 axis without period (uniform branch)
__device__ __forceinline__ RegBracket bracket_f64(const DAxis& a, const double* sp, const int* slut,
                                                  double c, bool& ok) {
  if (a.kind != 0) return bracket_irregular(a, sp, slut, c, ok);
  return bracket_regular<false, 2>(a, c, 0.0, ok);
}

// linear<Z, R> / nearest<Z, R> (interpolate/linear.hpp:22-33, nearest.hpp:21-32)
// on a verified bracket z0 <= z < z1.  f64 axis: dl = fl(z - z0), c1 = z1.
template <typename R>
__device__ __forceinline__ R div_ieee_r(R a, R b);
template <>
__device__ __forceinline__ float div_ieee_r<float>(float a, float b) { return __fdiv_rn(a, b); }
template <>
__device__ __forceinline__ double div_ieee_r<double>(double a, double b) { return __ddiv_rn(a, b); }

template <typename R>
__device__ __forceinline__ R axis_mix_f64(int method, double z, const RegBracket& b, R y0, R y1) {
  if (method == PYI_AXIS_NEAREST) {
    // |z0 - z| < |z1 - z| ? y0 : y1 ; |z0 - z| == dl exactly
    return b.dl < __dsub_rn(b.c1, z) ? y0 : y1;
  }
  const R dx = static_cast<R>(__dsub_rn(b.c1, b.c0));
  const R t = div_ieee_r<R>(static_cast<R>(b.dl), dx);
  return y0 + t * (y1 - y0);
}
// i64 axis: d0 = z - z0 and step = z1 - z0, exact
template <typename R>
__device__ __forceinline__ R axis_mix_i64(int method, long long d0, long long step, R y0, R y1) {
  if (method == PYI_AXIS_NEAREST) {
    return d0 < step - d0 ? y0 : y1;
  }
  const R dx = static_cast<R>(step);
  const R t = div_ieee_r<R>(static_cast<R>(d0), dx);
  return y0 + t * (y1 - y0);
}

// The generic code, out of line (reads the packed copy with its own loads).
template <int ND, bool ZI64, typename R>
__device__ __noinline__ R slow_point_nd(const GeoParams& p, double x, double y,
                                        typename ZOf<ZI64>::type z, double u, long long index) {
  R res;
  const int bad = geometric_eval<ND, -1, R, ZI64, true>(p, x, y, z, u, res);
  if (bad < 0) return res;
  if (p.bounds_error) {
    atomicMin(p.oob_key, (static_cast<unsigned long long>(p.index_base + index) << 2) | bad);
  }
  return RT<R>::nan();
}

__device__ __forceinline__ void stage_axis(const DAxis& a, double* sp, int* slut) {
  if (a.kind == 0 || a.coord_i64) return;
  const double* gp = static_cast<const double*>(a.points);
  for (int i = threadIdx.x; i < a.n; i += blockDim.x) sp[i] = __ldg(gp + i);
  for (int i = threadIdx.x; i <= a.lut_bins; i += blockDim.x) slut[i] = __ldg(a.lut + i);
}

// One point per thread and iteration; whole warps iterate together so that
// lane pairs can share the line requests of 64-byte cells (4-D).
template <int ND, bool ZI64, typename R, bool BIL, int MINB = 2>
__global__ void __launch_bounds__(kFastThreads, MINB)
bilinear_nd_fast_kernel(const __grid_constant__ GeoParams p) {
  using Z = typename ZOf<ZI64>::type;
  __shared__ double s_y[kFastMaxIrregular];
  __shared__ int s_ylut[kFastMaxLut];
  __shared__ double s_z[kNdMaxIrregular];
  __shared__ int s_zlut[kNdMaxLut];
  __shared__ double s_u[ND >= 4 ? kNdMaxIrregular : 1];
  __shared__ int s_ulut[ND >= 4 ? kNdMaxLut : 1];
  stage_axis(p.ax[1], s_y, s_ylut);
  stage_axis(p.ax[2], s_z, s_zlut);
  if (ND >= 4) stage_axis(p.ax[3], s_u, s_ulut);
  __syncthreads();

  const DAxis& ax = p.ax[0];
  const DAxis& ay = p.ax[1];
  const DAxis& az = p.ax[2];
  const bool per = ax.period != 0.0;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  const unsigned lane = threadIdx.x & 31u;
  R* __restrict__ out = static_cast<R*>(p.out);
  constexpr bool bilinear = BIL;  // idw / nearest (trigonometry) are compiled apart
  const Z* __restrict__ zs = static_cast<const Z*>(p.z);

  long long k = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  // software pipeline: the coordinates of the next iteration are in flight
  double x = 0.0, y = 0.0, u = 0.0;
  Z z = Z(0);
  if (k < p.n) {
    x = __ldcs(p.x + k);
    y = __ldcs(p.y + k);
    z = __ldcs(zs + k);
    if (ND >= 4) u = __ldcs(p.u + k);
  }
  while (k - lane < p.n) {  // warp-uniform bound
    const bool active = k < p.n;
    const long long kn = k + stride;
    double xn = 0.0, yn = 0.0, un = 0.0;
    Z zn = Z(0);
    if (kn < p.n) {
      xn = __ldcs(p.x + kn);
      yn = __ldcs(p.y + kn);
      zn = __ldcs(zs + kn);
      if (ND >= 4) un = __ldcs(p.u + kn);
    }
    // ---- brackets ---------------------------------------------------------
    bool ok = active && sane(x) && sane(y);
    double c = x;
    if (per) {
      const double cn = norm_fast(x, ax.min, ax.period, ax.inv_period);
      c = outside_period(ax, x) ? cn : x;
    }
    const RegBracket bx = per ? bracket_regular<true>(ax, c, sizeof(R) == 8 ? ax.seam_c1 : ax.seam_c1_f32, ok)
                              : bracket_regular<false>(ax, c, 0.0, ok);
    const RegBracket by = bracket_f64(ay, s_y, s_ylut, y, ok);
    RegBracket bz;
    int kz = 0;
    long long dz = 0;
    if (ZI64) {
      bracket_regular_i64(az, static_cast<long long>(z), kz, dz, ok);
    } else {
      ok = ok && sane(static_cast<double>(z));
      bz = bracket_f64(az, s_z, s_zlut, static_cast<double>(z), ok);
      kz = bz.k;
    }
    RegBracket bu;
    if (ND >= 4) {
      ok = ok && sane(u);
      bu = bracket_f64(p.ax[3], s_u, s_ulut, u, ok);
    }
    // x1 is normalised against the R-rounded x0 (bivariate.hpp:78-89); the
    // seam constant already is, any other cell needing it leaves the hot path
    const R x0 = static_cast<R>(bx.c0);
    const double x0d = static_cast<double>(x0);
    const double hi0 = per ? __dadd_rn(x0d, ax.period) : 0.0;
    if (per) ok = ok && (bx.k == ax.n - 1 || !(bx.c1 >= hi0 || bx.c1 < x0d));
    // ---- gather (in flight while the weights are computed) ------------------
    CellRef r;
    r.sx = r.sy = r.sz = r.su = false;
    r.cell = (static_cast<long long>(bx.k) * ay.n + by.k) * az.n + kz;
    if (ND >= 4) r.cell = r.cell * p.ax[3].n + bu.k;
    if (!ok) r.cell = 0;  // always a legal address
    // descending regular f64 axes (y, z, u): corner order of the reference's pair (k + 1, k)
    r.sy = by.swap;
    if (!ZI64) r.sz = bz.swap;
    if (ND >= 4) r.su = bu.swap;
    Corners<ND, R> q;
    if constexpr (Corners<ND, R>::BYTES >= 64) {  // 64- / 128-byte cells: one request per lane group
      geo_gather_group<ND, R>(p, r, q);
    } else {
      geo_gather_packed<ND, R>(p, r, q);
    }
    // ---- weights (Bilinear::evaluate, geometric/bivariate.hpp:110-132) -------
    double px = x;
    if (per) {
      const bool out2 = x >= hi0 || x < x0d;
      const double pn = norm_fast(x, x0d, ax.period, ax.inv_period);
      px = out2 ? pn : x;
    }
    const R y0 = static_cast<R>(by.c0);
    const R x1 = static_cast<R>(bx.c1);
    const R y1 = static_cast<R>(by.c1);
    const R pxr = static_cast<R>(px);
    const R pyr = static_cast<R>(y);
    constexpr int NU = Corners<ND, R>::NU;
    R sp[2][NU];
    if (bilinear) {
      const R dx = x1 - x0;
      const R dy = y1 - y0;
      const R t = div_ieee_r<R>(pxr - x0, dx);
      const R w = div_ieee_r<R>(pyr - y0, dy);
      const R omt = R(1) - t;
      const R omw = R(1) - w;
#pragma unroll
      for (int cz = 0; cz < 2; ++cz) {
#pragma unroll
        for (int d = 0; d < NU; ++d) {
          sp[cz][d] = omt * (omw * q.q[cz][d][0][0] + w * q.q[cz][d][0][1]) +
                      t * (omw * q.q[cz][d][1][0] + w * q.q[cz][d][1][1]);
        }
      }
    } else {  // idw / nearest on the same verified corners (geometric/bivariate.hpp:189-302)
#pragma unroll
      for (int cz = 0; cz < 2; ++cz) {
#pragma unroll
        for (int d = 0; d < NU; ++d) {
          sp[cz][d] = eval_spatial<R>(p.method, p.idw_exp, pxr, pyr, x0, y0, x1, y1, q.q[cz][d][0][0],
                                      q.q[cz][d][0][1], q.q[cz][d][1][0], q.q[cz][d][1][1]);
        }
      }
    }
    R along[NU];
#pragma unroll
    for (int d = 0; d < NU; ++d) {
      along[d] = ZI64 ? axis_mix_i64<R>(p.z_method, dz, az.istep, sp[0][d], sp[1][d])
                      : axis_mix_f64<R>(p.z_method, static_cast<double>(z), bz, sp[0][d], sp[1][d]);
    }
    R res = along[0];
    if (ND >= 4) res = axis_mix_f64<R>(p.u_method, u, bu, along[0], along[NU - 1]);
    if (!ok && active) {
      // a finite coordinate beyond the ends of an axis that does not wrap (or
      // a NaT / out-of-range time) cannot be bracketed: NaN; with bounds_error,
      // and for anything else, the generic code decides
      bool far = sane(x) && sane(y) && (beyond_ends(ax, c) || beyond_ends(ay, y));
      if (ZI64) {
        const long long num = static_cast<long long>(static_cast<unsigned long long>(z) -
                                                     static_cast<unsigned long long>(az.istart));
        far = far || num < 0 || num > az.ispan;
      } else {
        far = far || (sane(static_cast<double>(z)) && beyond_ends(az, static_cast<double>(z)));
      }
      if (ND >= 4) far = far || (sane(u) && beyond_ends(p.ax[3], u));
      res = (far && !p.bounds_error) ? RT<R>::nan() : slow_point_nd<ND, ZI64, R>(p, x, y, z, u, k);
    }
    if (active) __stcs(out + k, res);
    k = kn;
    x = xn;
    y = yn;
    z = zn;
    u = un;
  }
}

// Eligibility of the N-D fast path (everything else runs the generic kernels).
static bool axis_f64_fast(const DAxis& a, int max_irregular) {
  if (!(a.fast_ok || a.desc_ok) || a.coord_i64 || a.n < 2 || a.period != 0.0) return false;
  if (a.kind != 0 && (a.n > max_irregular || a.lut_bins <= 0 || a.lut_bins + 1 > 4 * max_irregular + 1)) {
    return false;
  }
  return true;
}

static bool fastnd_eligible(const GeoParams& p) {
  if (p.ndim < 3 || p.packed == nullptr) return false;
  const DAxis& ax = p.ax[0];
  if (!ax.fast_ok || ax.coord_i64 || ax.kind != 0 || ax.n < 2) return false;
  if (ax.period != 0.0 && !(ax.period > 0.0 && ax.step < ax.period)) return false;
  if (!axis_f64_fast(p.ax[1], kFastMaxIrregular)) return false;
  const DAxis& az = p.ax[2];
  if (az.coord_i64) {
    if (!az.fast_ok || az.kind != 0) return false;
  } else if (!axis_f64_fast(az, kNdMaxIrregular)) {
    return false;
  }
  if (p.ndim >= 4 && !axis_f64_fast(p.ax[3], kNdMaxIrregular)) return false;
  return true;
}

template <typename R, bool BIL>
static void launch_fastnd_r(const GeoParams& p, cudaStream_t stream) {
  const bool zi64 = p.ax[2].coord_i64 != 0;
  if constexpr (sizeof(R) == 4 && BIL) {
    // tuning knob for the benchmark configurations (float results, int64 time): PYI_ND_MINB = 2 .. 5
    const char* e = std::getenv("PYI_ND_MINB");
    // (round-2 sweep, 1e8 points: trivariate 32.9 / 32.1 / 31.4 / 30.9 and quadrivariate 29.4 / 26.9 /
    // 27.7 / 27.4 Gpts/s at 2 / 3 / 4 / 5 blocks: the DRAM request rate is the limit, not occupancy)
    if (e && zi64 && e[0] >= '3' && e[0] <= '5') {
      const int b = e[0] - '0';
      const int nb = grid_for(p.n, kFastThreads, b);
      if (p.ndim == 3) {
        if (b == 3) bilinear_nd_fast_kernel<3, true, R, BIL, 3><<<nb, kFastThreads, 0, stream>>>(p);
        else if (b == 4) bilinear_nd_fast_kernel<3, true, R, BIL, 4><<<nb, kFastThreads, 0, stream>>>(p);
        else bilinear_nd_fast_kernel<3, true, R, BIL, 5><<<nb, kFastThreads, 0, stream>>>(p);
      } else {
        if (b == 3) bilinear_nd_fast_kernel<4, true, R, BIL, 3><<<nb, kFastThreads, 0, stream>>>(p);
        else if (b == 4) bilinear_nd_fast_kernel<4, true, R, BIL, 4><<<nb, kFastThreads, 0, stream>>>(p);
        else bilinear_nd_fast_kernel<4, true, R, BIL, 5><<<nb, kFastThreads, 0, stream>>>(p);
      }
      return;
    }
  }
  const int blocks = grid_for(p.n, kFastThreads, 2);
  if (p.ndim == 3) {
    if (zi64) bilinear_nd_fast_kernel<3, true, R, BIL><<<blocks, kFastThreads, 0, stream>>>(p);
    else bilinear_nd_fast_kernel<3, false, R, BIL><<<blocks, kFastThreads, 0, stream>>>(p);
  } else {
    if (zi64) bilinear_nd_fast_kernel<4, true, R, BIL><<<blocks, kFastThreads, 0, stream>>>(p);
    else bilinear_nd_fast_kernel<4, false, R, BIL><<<blocks, kFastThreads, 0, stream>>>(p);
  }
}

static cudaError_t launch_fastnd(const GeoParams& p, cudaStream_t stream) {
  const bool r64 = p.dtype == PYI_F64 || p.dtype == PYI_I64 || p.dtype == PYI_U64;
  const bool bil = p.method == PYI_BILINEAR;
  if (r64) {
    if (bil) launch_fastnd_r<double, true>(p, stream);
    else launch_fastnd_r<double, false>(p, stream);
  } else {
    if (bil) launch_fastnd_r<float, true>(p, stream);
    else launch_fastnd_r<float, false>(p, stream);
  }
  g_launches.fetch_add(1);
  return cudaGetLastError();
}

}  // namespace pyi
