// pyi_fast2d_tol.cuh -- K1t: bivariate bilinear on a float64-result 2-D grid in
// TOLERANCE value mode (pyi_value_mode of include/pyinterp_b200.h).  Included by
// pyi_kernels.cu after pyi_fast2d.cuh, whose pieces it reuses (cell-packed
// gather, irregular bracket, generic redo).
//
// Same function as bivariate_single (pybind/geometric/bivariate.hpp:47-97) with
// Bilinear::evaluate (math/interpolate/geometric/bivariate.hpp:110-132).  Index
// selection stays bit-exact; the VALUE arithmetic is allowed to differ from the
// reference by rounding (north star: 1e-12 relative on float64):
//
//   * bracket: k = rint(c * inv_step + off) with off = -(start * inv_step) - 0.5
//     is a floor guess made with two additions of 1.5 * 2^52 (no 64-bit
//     conversions, which issue at a quarter of the FP64 rate).  The cell is then
//     VERIFIED from the nodes themselves: c0 = node(k) and c1 = node(k + 1) are
//     the reference's coordinate_value (start + k * step, two roundings,
//     container.hpp:480-483), and t = (c - c0) / (c1 - c0) lies strictly inside
//     (2^-21, 1 - 2^-21) only if c0 < c < c1, which is the reference's bracket
//     (math/axis.hpp:745-766, ascending axis).  Everything else -- a point on a
//     node, a wrong guess, NaN, a coordinate outside the axis -- leaves the hot
//     path: beyond the ends of a non-periodic axis the result is NaN, the rest is
//     queued (a device counter + index list per launch) and re-done by the generic
//     code in tol2d_redo_kernel -- the hot loop has no call.  When more than 2^20
//     points of a launch are queued (e.g. every query exactly on a node) the
//     bit-identical kernel of pyi_fast2d.cuh, launched behind the queue, recomputes
//     the whole launch; it returns at once otherwise.
//   * "exact nodes" axes (DAxis::tol_ok == 2, e.g. every multiple-of-2^-k grid
//     such as the 0.25 degree one): the host has checked that fma(k, step, start)
//     IS the reference's node for every k and that every spacing, the periodic
//     seam included, is exactly `step`; then c0 is one fma and t = (c - c0) *
//     inv_step, bit-identical to the reference's quotient when step is a power
//     of two.
//   * general regular axes: dx = c1 - c0 differs from `step` by the rounding of
//     the nodes (relative e = 1 - dx * inv_step, |e| < 2^-26 by the host's
//     gate), t = t0 + t0 * e with t0 = (c - c0) * inv_step: first-order exact,
//     e^2 < 2^-52.
//   * periodic x: the cell guess uses c = x - floor((x - min) / period) * period
//     in one fma; the weight's numerator repeats the reference's own roundings
//     (query normalised against the cell's x0, bivariate.hpp:83-88) so that t
//     matches the reference to an ulp on any grid spacing.
//   * evaluation: the reference's (1-t)((1-u) q00 + u q01) + t((1-u) q10 + u q11)
//     with the inner sums contracted into fmas (same structure, so infinities and
//     NaNs in the grid propagate as in the reference).
//
// Per point: ~27 FP64 instructions instead of ~72 (two correctly rounded
// divisions, unfused products), four points per thread with 32-byte streaming
// loads / stores.
#pragma once

namespace pyi {

// Irregular ascending y without period.  Shared-memory lookups share the L1TEX
// data pipe with the cell gather, and random 16-byte reads conflict on banks
// (ncu, round 2: 33 shared wavefronts per warp of points for three loads, 23 of
// them bank conflicts; the pipe 98 % busy), so a point reads as little as it can:
//   bin table   2 bytes: j = nodes below the bin, flags "a node inside" /
//               "several nodes inside" (bins: the host's uniform-bin table,
//               pyi_axis_host.hpp: build_lut, whose bin() is monotone in c,
//               merged 2^shift at a time when the axis has more than 1500 bins)
//   cell table  16 bytes: node(k) and 1 / (node(k+1) - node(k))
// k = (number of nodes <= c) - 1.  A bin without a node gives k = j - 1 at once;
// with one node inside, cell j is read first and cell j - 1 only by the lanes
// whose c lies below node(j); fuller bins are scanned.  Every decision is an
// exact comparison with a node value and bin() is the host's own expression, so
// the bracket is the reference's (container.hpp:382-404 + axis.hpp:745-766 on an
// ascending axis); c >= node(k) is still verified from the cell that was read.
constexpr int kTolMaxBins = 6000;
constexpr unsigned short kTolOne = 0x8000, kTolMany = 0x4000;

__device__ __forceinline__ void tol_irregular(const DAxis& a, const unsigned short* __restrict__ sbin,
                                              const double2* __restrict__ scell, int shift, double c,
                                              int& k, double& t, bool& safe) {
  const double f = __dmul_rn(__dsub_rn(c, a.lut_min), a.lut_inv);
  int bin = __double2int_rd(f);  // saturating; NaN -> 0 and fails the checks below
  const bool in_table = sane(c) && static_cast<unsigned>(bin) < static_cast<unsigned>(a.lut_bins);
  bin = in_table ? bin : 0;
  PYI_CHECK_RANGE(bin >> shift, 1, ((a.lut_bins - 1) >> shift) + 1, "K1t bin table");
  const unsigned e = sbin[bin >> shift];
  const int j = static_cast<int>(e & 0x3fffu);
  const bool inside = (e & kTolOne) != 0;
  int kk = inside ? j : j - 1;
  kk = min(max(kk, 0), a.n - 1);
  PYI_CHECK_RANGE(kk, 1, a.n, "K1t cell table");
  double2 cell = scell[kk];  // {node(kk), 1 / spacing}
  bool below = inside && c < cell.x;  // the node of the bin is above c: cell j - 1
  if (e & kTolMany) {  // several nodes in the bin: the last one that is <= c
    below = false;
    kk = j - 1;
    for (int m = j; m < a.n && scell[m].x <= c; ++m) kk = m;
    kk = min(max(kk, 0), a.n - 1);
    cell = scell[kk];  // (c >= lut_min = node(0) inside the table: kk >= 0 here)
  }
  if (below) {
    kk = j - 1;
    cell = scell[max(kk, 0)];
  }
  const bool ok = in_table && static_cast<unsigned>(kk) < static_cast<unsigned>(a.n - 1);
  const double dl = __dsub_rn(c, cell.x);
  k = ok ? kk : 0;
  t = __dmul_rn(dl, cell.y);
  safe = safe && ok && !negative(dl);
}

template <int PTS>
struct VecIO;
template <>
struct VecIO<2> {
  static __device__ __forceinline__ void load(const double* p, double (&v)[2]) {
    const double2 t = __ldcs(reinterpret_cast<const double2*>(p));
    v[0] = t.x;
    v[1] = t.y;
  }
  static __device__ __forceinline__ void store(double* p, const double (&v)[2]) {
    __stcs(reinterpret_cast<double2*>(p), make_double2(v[0], v[1]));
  }
};
template <>
struct VecIO<4> {
  static __device__ __forceinline__ void load(const double* p, double (&v)[4]) {
    asm volatile("ld.global.cs.v4.f64 {%0,%1,%2,%3}, [%4];"
                 : "=d"(v[0]), "=d"(v[1]), "=d"(v[2]), "=d"(v[3])
                 : "l"(p));
  }
  static __device__ __forceinline__ void store(double* p, const double (&v)[4]) {
    asm volatile("st.global.cs.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(v[0]), "d"(v[1]),
                 "d"(v[2]), "d"(v[3])
                 : "memory");
  }
};

// XM: 1 general / 2 exact nodes (x is always regular); YM: 0 irregular,
// 1 general regular, 2 exact-nodes regular.  PTS consecutive points per thread.
template <bool PER, int XM, int YM, int PTS, int MINB, bool PREFETCH>
__global__ void __launch_bounds__(kFastThreads, MINB)
bilinear2d_tol_kernel(const __grid_constant__ GeoParams p) {
  // irregular y: packed tables in dynamic shared memory (sized by the launcher)
  extern __shared__ __align__(16) unsigned char s_raw[];
  const int ny_ = p.ax[1].n;
  int shift = 0;
  double2* s_cell = reinterpret_cast<double2*>(s_raw);
  unsigned short* s_bin = nullptr;
  if (YM == 0) {
    while (((p.ax[1].lut_bins - 1) >> shift) + 1 > kTolMaxBins) ++shift;
    const int nb = ((p.ax[1].lut_bins - 1) >> shift) + 1;
    s_bin = reinterpret_cast<unsigned short*>(s_cell + ny_);
    const double* gp = static_cast<const double*>(p.ax[1].points);
    const int* lut = p.ax[1].lut;
    for (int i = threadIdx.x; i < ny_; i += blockDim.x) {
      const double a = __ldg(gp + i);
      // reciprocal spacing of cell i (the last entry is never used: cell n - 1 is not a bracket)
      const double inv = i + 1 < ny_ ? __ddiv_rn(1.0, __dsub_rn(__ldg(gp + i + 1), a)) : 0.0;
      s_cell[i] = make_double2(a, inv);
    }
    for (int i = threadIdx.x; i < nb; i += blockDim.x) {
      const int lo = __ldg(lut + (i << shift));
      const int last = min((i + 1) << shift, p.ax[1].lut_bins);
      const int count = __ldg(lut + last) - lo;
      s_bin[i] = static_cast<unsigned short>(lo | (count == 1 ? kTolOne : 0) | (count > 1 ? kTolMany : 0));
    }
    __syncthreads();
  }
  const DAxis& ax = p.ax[0];
  const DAxis& ay = p.ax[1];
  const unsigned groups = static_cast<unsigned>(p.n / PTS);  // the launcher keeps n below 2^32
  const unsigned stride = gridDim.x * blockDim.x;
  double* __restrict__ out = static_cast<double*>(p.out);
  // software pipeline: the coordinates of the next iteration are in flight while
  // this one is computed (ncu, round 2: 52 % of the stall samples sat on the
  // first use of x, the DRAM latency of the stream loads)
  unsigned g = blockIdx.x * blockDim.x + threadIdx.x;
  double x[PTS], y[PTS];
  if (g < groups) {
    VecIO<PTS>::load(p.x + static_cast<size_t>(g) * PTS, x);
    VecIO<PTS>::load(p.y + static_cast<size_t>(g) * PTS, y);
  }
  while (g < groups) {
    const size_t base = static_cast<size_t>(g) * PTS;
    const unsigned gn = g + stride;  // groups <= 2^30: no wrap
    double xn[PTS], yn[PTS];
    if (PREFETCH && gn < groups) {
      VecIO<PTS>::load(p.x + static_cast<size_t>(gn) * PTS, xn);
      VecIO<PTS>::load(p.y + static_cast<size_t>(gn) * PTS, yn);
    }
    double t[PTS], u[PTS], o[PTS];
    long long cell[PTS];
    bool safe[PTS];
    bool all_safe = true;
#pragma unroll
    for (int i = 0; i < PTS; ++i) {
      double c = x[i], jf = 0.0;
      bool s = true;
      if (PER) {
        // c = x - floor((x - min) / period) * period, one rounding: only the
        // cell guess depends on it
        int unused;
        jf = rint_magic(fma(x[i], ax.inv_period, ax.tol_poff), unused);
        c = fma(-jf, ax.period, x[i]);
        // |x| below 2^28 steps (and 2^40): c is then within 2^-24 of a step of
        // the reference's normalised coordinate; farther points are re-done
        s = static_cast<int>(static_cast<unsigned>(__double2hiint(x[i])) & 0x7ff00000u) < ax.tol_xhi;
      }
      int kx, ky;
      tol_regular<XM, PER>(ax, c, x[i], jf, kx, t[i], s);
      if (YM == 0) {
        tol_irregular(ay, s_bin, s_cell, shift, y[i], ky, u[i], s);
      } else {
        tol_regular<YM, false>(ay, y[i], y[i], 0.0, ky, u[i], s);
      }
      safe[i] = s;
      all_safe = all_safe && s;
      cell[i] = s ? static_cast<long long>(kx) * ay.n + ky : 0LL;  // cell 0 is always a legal address
    }
    Cell<double> q[PTS];
#pragma unroll
    for (int i = 0; i < PTS; ++i) {
      PYI_CHECK_RANGE(cell[i], 1, static_cast<long long>(ax.n) * ay.n, "K1t packed cell");
      q[i].load(p.packed, cell[i]);
    }
#pragma unroll
    for (int i = 0; i < PTS; ++i) {
      const double omt = __dsub_rn(1.0, t[i]);
      const double omu = __dsub_rn(1.0, u[i]);
      double a, b;
      if (YM != 0 && ay.desc_ok) {  // uniform: descending y, u is measured from node k + 1
        a = fma(u[i], q[i].v00, __dmul_rn(omu, q[i].v01));
        b = fma(u[i], q[i].v10, __dmul_rn(omu, q[i].v11));
      } else {
        a = fma(u[i], q[i].v01, __dmul_rn(omu, q[i].v00));
        b = fma(u[i], q[i].v11, __dmul_rn(omu, q[i].v10));
      }
      o[i] = fma(t[i], b, __dmul_rn(omt, a));
    }
    if (!all_safe) {  // rare: on a node, outside an axis, NaN
#pragma unroll
      for (int i = 0; i < PTS; ++i) {
        if (safe[i]) continue;
        // a finite coordinate beyond the ends of an axis that does not wrap is
        // NaN (see beyond_ends); with bounds_error the generic code reports it.
        // Everything else is queued for the redo kernel: no call in this loop.
        bool far = false;
        if (!p.bounds_error && sane(x[i]) && sane(y[i])) {
          far = beyond_ends(ay, y[i]) || (!PER && beyond_ends(ax, x[i]));
        }
        if (!far) {
          const unsigned slot = atomicAdd(p.queue, 1u);
          if (slot < p.queue_cap) p.queue[1 + slot] = static_cast<unsigned>(base + i);
        }
        o[i] = RT<double>::nan();
      }
    }
    VecIO<PTS>::store(out + base, o);
    g = gn;
    if (PREFETCH) {
#pragma unroll
      for (int i = 0; i < PTS; ++i) {
        x[i] = xn[i];
        y[i] = yn[i];
      }
    } else if (g < groups) {
      VecIO<PTS>::load(p.x + static_cast<size_t>(g) * PTS, x);
      VecIO<PTS>::load(p.y + static_cast<size_t>(g) * PTS, y);
    }
  }
}

// K1f: the same verified brackets for FLOAT32 results (float32 and small-integer
// grids), regular y.  Nothing is approximated here: once the bracket is the
// reference's, the value is its own float arithmetic (bivariate.hpp:78-96 casts
// the corner coordinates and the query to float; Bilinear::evaluate,
// geometric/bivariate.hpp:110-132, divides and sums in float) -- the operations
// of fast2d_point<float> (pyi_fast2d.cuh) on the same operands, so the results
// are BIT-IDENTICAL to the reference; only the search for the bracket is the
// cheap one.  Two points per thread, 16-byte loads, 8-byte stores.
template <bool PER, int XM, int YM, int MINB>
__global__ void __launch_bounds__(kFastThreads, MINB)
bilinear2d_tolf_kernel(const __grid_constant__ GeoParams p) {
  const DAxis& ax = p.ax[0];
  const DAxis& ay = p.ax[1];
  const bool ydesc = ay.desc_ok != 0;
  const unsigned groups = static_cast<unsigned>(p.n / 2);
  const unsigned stride = gridDim.x * blockDim.x;
  float* __restrict__ out = static_cast<float*>(p.out);
  for (unsigned g = blockIdx.x * blockDim.x + threadIdx.x; g < groups; g += stride) {
    const size_t base = static_cast<size_t>(g) * 2;
    double x[2], y[2];
    VecIO<2>::load(p.x + base, x);
    VecIO<2>::load(p.y + base, y);
    long long cell[2];
    bool safe[2];
    float tnum[2], unum[2], dx[2], dy[2];
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      double c = x[i], jf = 0.0;
      bool s = true;
      if (PER) {
        int unused;
        jf = rint_magic(fma(x[i], ax.inv_period, ax.tol_poff), unused);
        c = fma(-jf, ax.period, x[i]);
        s = static_cast<int>(static_cast<unsigned>(__double2hiint(x[i])) & 0x7ff00000u) < ax.tol_xhi;
      }
      int kx, ky;
      double td, ud, x0n, x1n, y0n, y1n;
      tol_regular_nodes<XM, PER>(ax, c, x[i], jf, kx, td, s, x0n, x1n);
      tol_regular_nodes<YM, false>(ay, y[i], y[i], 0.0, ky, ud, s, y0n, y1n);
      // bivariate.hpp:78-89 in float: x1 and the query are normalised against the
      // float-rounded x0 promoted back to double (see fast2d_point)
      const float x0 = static_cast<float>(x0n);
      double px = x[i];
      if (PER) {
        const double x0d = static_cast<double>(x0);
        const double hi0 = __dadd_rn(x0d, ax.period);
        const bool seam = kx == ax.n - 1;
        x1n = seam ? ax.seam_c1_f32 : x1n;
        // the seam constant is normalised already; any other cell needing it is re-done
        s = s && (seam || !(x1n >= hi0 || x1n < x0d));
        const bool out2 = x[i] >= hi0 || x[i] < x0d;
        const double pn = norm_fast(x[i], x0d, ax.period, ax.inv_period);
        px = out2 ? pn : x[i];
      }
      tnum[i] = __fsub_rn(static_cast<float>(px), x0);
      dx[i] = __fsub_rn(static_cast<float>(x1n), x0);
      const float y0 = static_cast<float>(y0n);
      unum[i] = __fsub_rn(static_cast<float>(y[i]), y0);
      dy[i] = __fsub_rn(static_cast<float>(y1n), y0);
      safe[i] = s;
      // (indexes clamped as well: the load below must be legal whatever an unsettled lane holds)
      const int kxc = min(max(kx, 0), ax.n - 1), kyc = min(max(ky, 0), ay.n - 1);
      cell[i] = s ? static_cast<long long>(kxc) * ay.n + kyc : 0LL;
    }
    Cell<float> q[2];
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      PYI_CHECK_RANGE(cell[i], 1, static_cast<long long>(ax.n) * ay.n, "K1f packed cell");
      q[i].load(p.packed, cell[i]);
    }
    float o[2];
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      // descending y: the reference's first corner is the lower node, k + 1
      const float v00 = ydesc ? q[i].v01 : q[i].v00, v01 = ydesc ? q[i].v00 : q[i].v01;
      const float v10 = ydesc ? q[i].v11 : q[i].v10, v11 = ydesc ? q[i].v10 : q[i].v11;
      const float t = __fdiv_rn(tnum[i], dx[i]);
      const float u = __fdiv_rn(unum[i], dy[i]);
      const float omt = __fsub_rn(1.0f, t);
      const float omu = __fsub_rn(1.0f, u);
      const float a = __fadd_rn(__fmul_rn(omu, v00), __fmul_rn(u, v01));
      const float b = __fadd_rn(__fmul_rn(omu, v10), __fmul_rn(u, v11));
      o[i] = __fadd_rn(__fmul_rn(omt, a), __fmul_rn(t, b));
    }
    if (!(safe[0] && safe[1])) {  // rare: on a node, outside an axis, NaN (as in the float64 kernel)
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        if (safe[i]) continue;
        bool far = false;
        if (!p.bounds_error && sane(x[i]) && sane(y[i])) {
          far = beyond_ends(ay, y[i]) || (!PER && beyond_ends(ax, x[i]));
        }
        if (!far) {
          const unsigned slot = atomicAdd(p.queue, 1u);
          if (slot < p.queue_cap) p.queue[1 + slot] = static_cast<unsigned>(base + i);
        }
        o[i] = RT<float>::nan();
      }
    }
    __stcs(reinterpret_cast<float2*>(out + base), make_float2(o[0], o[1]));
  }
}

// The points the hot kernel queued (on a node, NaN, far-away longitudes ...)
// and the n % pts tail, through the generic code.  Nothing to do when the queue
// overflowed: the exact kernel launched next recomputes every point.
template <typename R>
__global__ void __launch_bounds__(kFastThreads)
tol2d_redo_kernel(const __grid_constant__ GeoParams p, int pts) {
  const unsigned total = *p.queue;
  if (total > p.queue_cap) return;
  R* __restrict__ out = static_cast<R*>(p.out);
  const unsigned stride = gridDim.x * blockDim.x;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const long long k = p.queue[1 + i];
    out[k] = slow_point<R>(p, __ldg(p.x + k), __ldg(p.y + k), k);
  }
  if (blockIdx.x == 0 && threadIdx.x < static_cast<unsigned>(p.n % pts)) {
    const long long k = p.n - (p.n % pts) + threadIdx.x;
    out[k] = slow_point<R>(p, __ldg(p.x + k), __ldg(p.y + k), k);
  }
}

// PYINTERP_B200_BILINEAR = exact | tolerance: what PYI_VALUES_DEFAULT means
// (read per launch: the tests switch it)
static int bilinear_default_mode() {
  const char* e = std::getenv("PYINTERP_B200_BILINEAR");
  return (e != nullptr && e[0] == 'e') ? PYI_VALUES_EXACT : PYI_VALUES_TOLERANCE;
}

// Tuning variant: PYI_TOL_VARIANT = <points per thread: 2|4><min blocks per SM: 2..5>[n]
// ("n": no register prefetch); read per launch so that one process can sweep them
static void tol2d_variant(int& pts, int& minb, bool& prefetch) {
  // round-2 sweeps on three B200 boxes (profiles/r2_k1_sweep*.txt): the variants
  // within 5 % of the best change from box to box (25n: 188 / 165 Gpts/s, 24: 146 /
  // 184); 2 points x 3 blocks with register prefetch gave 181 / 181
  pts = 2;
  minb = 3;
  prefetch = true;
  const char* e = std::getenv("PYI_TOL_VARIANT");
  if (e && (e[0] == '2' || e[0] == '4') && e[1] >= '2' && e[1] <= '5') {
    pts = e[0] - '0';
    minb = e[1] - '0';
    prefetch = e[2] != 'n';
  }
}

static bool tol2d_eligible(const GeoParams& p) {
  const int mode = p.value_mode == PYI_VALUES_DEFAULT ? bilinear_default_mode() : p.value_mode;
  if (mode != PYI_VALUES_TOLERANCE) return false;
  // small launches are launch bound (three launches and a memset here, one for
  // the bit-identical kernel): below 2^22 points DEFAULT means exact
  // (PYINTERP_B200_BILINEAR=tolerance forces the tolerance kernel at any size)
  const char* forced = std::getenv("PYINTERP_B200_BILINEAR");
  if (p.value_mode == PYI_VALUES_DEFAULT && p.n < (1LL << 22) && !(forced != nullptr && forced[0] == 't')) {
    return false;
  }
  if (p.method != PYI_BILINEAR || !fast2d_eligible(p)) return false;
  if (!p.ax[0].tol_ok) return false;
  if (p.ax[1].kind == 0 && !p.ax[1].tol_ok) return false;
  const bool r64 = p.dtype == PYI_F64 || p.dtype == PYI_I64 || p.dtype == PYI_U64;
  if (!r64) {
    // float32 results (K1f, bit-identical): regular y, 16-byte coordinate pairs, 8-byte result pairs
    if (p.ax[1].kind != 0) return false;
    const uintptr_t in = reinterpret_cast<uintptr_t>(p.x) | reinterpret_cast<uintptr_t>(p.y);
    return (in & 15) == 0 && (reinterpret_cast<uintptr_t>(p.out) & 7) == 0;
  }
  const uintptr_t align = reinterpret_cast<uintptr_t>(p.x) | reinterpret_cast<uintptr_t>(p.y) |
                          reinterpret_cast<uintptr_t>(p.out);
  return (align & 31) == 0;
}

template <bool PER, int XM, int YM>
static void launch_tol2d_v(const GeoParams& p, cudaStream_t stream) {
  int pts, minb;
  bool prefetch;
  tol2d_variant(pts, minb, prefetch);
  // the tuning variants exist for the benchmark configurations only (periodic
  // exact-node x; exact-node or irregular y); every other axis combination is
  // compiled once, with the default variant
  constexpr bool kSweep = PER && XM == 2 && YM != 1;
  if (!kSweep) {
    pts = 2;
    minb = 4;
    prefetch = false;
  }
  const int blocks = grid_for((p.n + pts - 1) / pts, kFastThreads, minb);
  // irregular y: cell table (16 bytes per node) + bin table (2 bytes per bin): at most 28 KB
  size_t smem = 0;
  if (YM == 0) {
    int shift = 0;
    while (((p.ax[1].lut_bins - 1) >> shift) + 1 > kTolMaxBins) ++shift;
    smem = static_cast<size_t>(p.ax[1].n) * 16 +
           static_cast<size_t>(((p.ax[1].lut_bins - 1) >> shift) + 1) * sizeof(unsigned short);
  }
  if constexpr (!kSweep) {
    bilinear2d_tol_kernel<PER, XM, YM, 2, 4, false><<<blocks, kFastThreads, smem, stream>>>(p);
    return;
  } else {
#define PYI_LAUNCH(P, B)                                                                        \
  do {                                                                                          \
    if (prefetch) bilinear2d_tol_kernel<PER, XM, YM, P, B, true><<<blocks, kFastThreads, smem, stream>>>(p); \
    else bilinear2d_tol_kernel<PER, XM, YM, P, B, false><<<blocks, kFastThreads, smem, stream>>>(p);         \
  } while (0)
  if (pts == 2) {
    switch (minb) {
      case 2: PYI_LAUNCH(2, 2); break;
      case 3: PYI_LAUNCH(2, 3); break;
      case 4: PYI_LAUNCH(2, 4); break;
      default: PYI_LAUNCH(2, 5); break;
    }
  } else {
    switch (minb) {
      case 2: PYI_LAUNCH(4, 2); break;
      case 3: PYI_LAUNCH(4, 3); break;
      default: PYI_LAUNCH(4, 4); break;
    }
  }
#undef PYI_LAUNCH
  }
}

template <bool PER>
static void launch_tol2d_p(const GeoParams& p, cudaStream_t stream) {
  const int xm = p.ax[0].tol_ok;
  const int ym = p.ax[1].kind != 0 ? 0 : p.ax[1].tol_ok;
  if (xm == 2) {
    if (ym == 2) launch_tol2d_v<PER, 2, 2>(p, stream);
    else if (ym == 1) launch_tol2d_v<PER, 2, 1>(p, stream);
    else launch_tol2d_v<PER, 2, 0>(p, stream);
  } else {
    if (ym == 2) launch_tol2d_v<PER, 1, 2>(p, stream);
    else if (ym == 1) launch_tol2d_v<PER, 1, 1>(p, stream);
    else launch_tol2d_v<PER, 1, 0>(p, stream);
  }
}

template <bool PER>
static void launch_tolf_p(const GeoParams& p, cudaStream_t stream) {
  const int blocks = grid_for((p.n + 1) / 2, kFastThreads, 4);
  const int xm = p.ax[0].tol_ok, ym = p.ax[1].tol_ok;
  if (xm == 2) {
    if (ym == 2) bilinear2d_tolf_kernel<PER, 2, 2, 4><<<blocks, kFastThreads, 0, stream>>>(p);
    else bilinear2d_tolf_kernel<PER, 2, 1, 4><<<blocks, kFastThreads, 0, stream>>>(p);
  } else {
    if (ym == 2) bilinear2d_tolf_kernel<PER, 1, 2, 4><<<blocks, kFastThreads, 0, stream>>>(p);
    else bilinear2d_tolf_kernel<PER, 1, 1, 4><<<blocks, kFastThreads, 0, stream>>>(p);
  }
}

static cudaError_t launch_tol2d(const GeoParams& p, cudaStream_t stream) {
  const bool r64 = p.dtype == PYI_F64 || p.dtype == PYI_I64 || p.dtype == PYI_U64;
  const long long chunk = 1LL << 31;  // multiple of 4: keeps the 32-byte alignment of the slices
  int pts, minb;
  bool prefetch;
  tol2d_variant(pts, minb, prefetch);
  for (long long begin = 0; begin < p.n; begin += chunk) {
    GeoParams q = p;
    q.n = p.n - begin < chunk ? p.n - begin : chunk;
    q.index_base = p.index_base + begin;
    q.x = p.x + begin;
    q.y = p.y + begin;
    q.out = static_cast<char*>(p.out) + begin * (r64 ? 8 : 4);
    // the queue: the caller's (host pipeline: one per stream slot), else a
    // per-launch one from the stream-ordered allocator, so that concurrent
    // calls on other streams never share it
    const bool own = p.queue == nullptr;
    cudaError_t e = cudaSuccess;
    if (own) {
      e = cudaMallocAsync(reinterpret_cast<void**>(&q.queue), (kTolQueue + 1) * sizeof(unsigned), stream);
      if (e != cudaSuccess) return e;
      q.queue_cap = kTolQueue;
    }
    e = cudaMemsetAsync(q.queue, 0, sizeof(unsigned), stream);
    if (e == cudaSuccess && r64) {
      if (p.ax[0].period != 0.0) launch_tol2d_p<true>(q, stream);
      else launch_tol2d_p<false>(q, stream);
      // kernels compiled once use the default points per thread
      const bool sweep = p.ax[0].period != 0.0 && p.ax[0].tol_ok == 2 && !(p.ax[1].kind == 0 && p.ax[1].tol_ok == 1);
      tol2d_redo_kernel<double><<<sm_count(), kFastThreads, 0, stream>>>(q, sweep ? pts : 2);
      g_launches.fetch_add(2);
      e = cudaGetLastError();
    } else if (e == cudaSuccess) {
      if (p.ax[0].period != 0.0) launch_tolf_p<true>(q, stream);
      else launch_tolf_p<false>(q, stream);
      tol2d_redo_kernel<float><<<sm_count(), kFastThreads, 0, stream>>>(q, 2);
      g_launches.fetch_add(2);
      e = cudaGetLastError();
    }
    // the bit-identical kernel behind the queue: returns at once unless it overflowed
    if (e == cudaSuccess) {
      q.run_if_overflow = q.queue;
      e = r64 ? launch_fast2d_r<double>(q, stream) : launch_fast2d_r<float>(q, stream);
    }
    if (own) {
      const cudaError_t e2 = cudaFreeAsync(q.queue, stream);
      if (e == cudaSuccess) e = e2;
    }
    if (e != cudaSuccess) return e;
  }
  return cudaSuccess;
}

}  // namespace pyi
