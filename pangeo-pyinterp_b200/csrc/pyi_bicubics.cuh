// pyi_bicubics.cuh -- K4s: bicubic on a 6 x 6 window of a 2-D float64 or float32
// grid with regular axes, in TOLERANCE value mode, as ONE separable 6 x 6 stencil whose
// rows are loaded by lane quads.  Included by pyi_windowed.cu after
// pyi_bicubicw.cuh (whose redo pass and tags it shares).
//
// What is computed: windowed::pybind::bivariate with the Bicubic evaluator
// (pybind/windowed/bivariate.hpp:46-76, math/interpolate/bivariate/bicubic.hpp:
// 57-186).  The reference fits natural cubic splines through the six window
// nodes to get dz/dx, dz/dy and d2z/dxdy at the four corners of the bracketing
// cell, then evaluates the bicubic Hermite patch of those sixteen numbers
// (bicubic.hpp:104-185 is GSL's coefficient form of it).  Two facts make this a
// plain stencil on a regular axis:
//   * the spline derivative at the two centre nodes is LINEAR in the six node
//     values, with weights W2[k], W3[k] that depend only on the (uniform)
//     spacing -- the weights K4w already precomputes per grid (pyi_bicubicw.cuh);
//   * the Hermite patch is linear in its corner data and separable, so
//         f(x, y) = sum_i sum_j Cx_i(t) Cy_j(u) z[i][j],     i, j = 0..5
//     with  C_k(t) = [k==2] h00(t) + [k==3] h01(t) + dx (h10(t) W2[k] + h11(t) W3[k])
//     and h00 = 1 - 3t^2 + 2t^3, h01 = 3t^2 - 2t^3, h10 = t - 2t^2 + t^3,
//     h11 = t^3 - t^2 the cubic Hermite basis.  Each C_k is a cubic in t whose
//     four coefficients are computed once per grid on the host (pyi_host.cu).
// 42 fused multiply-adds replace ~280 FP64 operations per point (ten dot
// products + GSL's 16-coefficient form), and the summation order differs from
// the reference's: values agree to rounding (1e-15 of the window's magnitude;
// bar 1e-12), not bit for bit -- this is the tolerance mode; the bit-identical
// kernel (pyi_fastwin.cuh) serves PYI_VALUES_EXACT.
//
// Memory (ncu, round 2: K4w was bound by the L1TEX data pipe at 83 %: 24
// scattered 16-byte loads per point, every lane its own cache line): here a
// QUAD of lanes loads one 64-byte aligned run per window row (four adjacent
// 16-byte pieces covering the six values), so a load instruction touches 8 - 11
// cache lines instead of 32.  Each lane accumulates its two columns over the
// six rows with the row weights, applies its two column weights once at the
// end, and a 4 x 4 transpose-reduce over the quad (three shuffles for four
// points) gives the values.  A warp first brackets 32 points (one per lane: the
// verified-bracket scheme of pyi_common.cuh), then processes them in four
// groups of eight.  The column-weight cubics a lane can need live in registers
// (shared-memory lookups and shuffles share the L1TEX data pipe with the
// gathers: the first version of this kernel saturated it at 97 %).
//
// Everything outside the kernel's reach is tagged and recomputed by the generic
// kernel in redo mode: points on a node or not bracketed with margin, windows
// with infinities, coordinates the generic code must report (bounds_error).
// Windows that leave a non-periodic axis are NaN in the "undef" boundary mode
// (cache_loader.hpp:69-80), as are windows holding a NaN (cache.hpp:165-171).
#pragma once

namespace pyi {

constexpr int kBcsThreads = 128;
enum { BCS_COMPUTE = 0, BCS_NAN = 1, BCS_REDO = 2, BCS_IDLE = 3 };

__device__ __forceinline__ double shfl_xor_f64(double v, int mask) {
  return __shfl_xor_sync(0xffffffffu, v, mask);
}

// cubic c0 + c1 t + c2 t^2 + c3 t^3
__device__ __forceinline__ double cubic4(double c0, double c1, double c2, double c3, double t) {
  return fma(fma(fma(c3, t, c2), t, c1), t, c0);
}

// Two adjacent grid values, loaded as one aligned piece (16 bytes of float64,
// 8 bytes of float32); the arithmetic is float64 for both (the reference's
// window holds doubles, cache.hpp:46-60, and only the result is cast).
// (loaded with asm volatile: __ldg is speculatable for the compiler, which may then drop the
// "window (0, 0) for a lane that is not computed" select in front of it, pyi_fast2d.cuh)
template <typename T>
struct BcsPair;
template <>
struct BcsPair<double> {
  using type = double2;
  static __device__ __forceinline__ double2 load(const double* p) {
    double2 v;
    asm volatile("ld.global.nc.v2.f64 {%0,%1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p));
    return v;
  }
};
template <>
struct BcsPair<float> {
  using type = float2;
  static __device__ __forceinline__ float2 load(const float* p) {
    float2 v;
    asm volatile("ld.global.nc.v2.f32 {%0,%1}, [%2];" : "=f"(v.x), "=f"(v.y) : "l"(p));
    return v;
  }
};

// T: grid (and result) type, double or float.
// XM / YM: tol_ok of the axes (1 general, 2 exact nodes); PER: x has a period;
// EVEN: ny is even, so every row of a window starts on the parity of wy0.
template <typename T, bool PER, int XM, int YM, bool EVEN, int MINB>
__global__ void __launch_bounds__(kBcsThreads, MINB)
bicubic_sep_kernel(const __grid_constant__ WinNdParams p) {
  using V2 = typename BcsPair<T>::type;
  const DAxis& ax = p.ax[0];
  const DAxis& ay = p.ax[1];
  const int nx = ax.n, ny = ay.n;
  const int cells = static_cast<int>(p.cells);  // < 2^31 (bicubics_eligible)
  const bool wraps = PER && ax.periodic != 0;
  const bool quiet_nan = p.boundary == PYI_BOUNDARY_UNDEF && !p.bounds_error;
  const unsigned lane = threadIdx.x & 31u;
  const int piece = static_cast<int>(lane & 3u);
  const T* __restrict__ data = static_cast<const T*>(p.data);
  T* __restrict__ out = static_cast<T*>(p.out);
  // This lane always holds columns 2 piece - par and 2 piece + 1 - par of a row
  // (par = parity of the row's first element): the cubics of the three column
  // weights it can need, j = 2 piece - 1, 2 piece, 2 piece + 1 (table entry
  // j + 1; entries 0 and 7 are zero), live in registers for the whole kernel.
  double cy[3][4];
#pragma unroll
  for (int e = 0; e < 3; ++e) {
#pragma unroll
    for (int c = 0; c < 4; ++c) cy[e][c] = p.cw[1][2 * piece + e][c];
  }
  const long long step = static_cast<long long>(gridDim.x) * (kBcsThreads / 32) * 32;
  for (long long base = (static_cast<long long>(blockIdx.x) * (kBcsThreads / 32) + (threadIdx.x >> 5)) * 32;
       base < p.n; base += step) {
    const long long k = base + lane;
    const bool active = k < p.n;
    double x = 0.0, y = 0.0;
    if (active) {
      x = __ldcs(p.x + k);
      y = __ldcs(p.y + k);
    }
    // ---- phase 1: one point per lane, verified brackets ---------------------
    bool safe = active;
    double c = x, jf = 0.0;
    if (PER) {
      int unused;
      jf = rint_magic(fma(x, ax.inv_period, ax.tol_poff), unused);
      c = fma(-jf, ax.period, x);
      safe = safe && static_cast<int>(static_cast<unsigned>(__double2hiint(x)) & 0x7ff00000u) < ax.tol_xhi;
    }
    int kx, ky;
    double tx, ty;
    tol_regular<XM, PER>(ax, c, x, jf, kx, tx, safe);
    tol_regular<YM, false, false>(ay, y, y, 0.0, ky, ty, safe);
    int wx0 = kx - 2;
    const int wy0 = ky - 2;
    bool inside = ky >= 2 && ky + 3 < ny;
    if (wraps) {
      wx0 = wx0 < 0 ? wx0 + nx : wx0;
    } else {
      inside = inside && kx >= 2 && kx + 3 < nx;
    }
    int status = BCS_IDLE;
    if (active) {
      if (safe) {
        // a verified bracket whose window leaves an axis: NaN in the undef mode
        status = inside ? BCS_COMPUTE : (quiet_nan ? BCS_NAN : BCS_REDO);
        // the aligned 64-byte run of the window's last row may extend up to two
        // elements past the six values: not across the end of the array
        const bool last_row = wraps ? (wx0 + 5 >= nx - 1) : (kx + 3 == nx - 1);
        if (inside && last_row && wy0 + 8 > ny) status = BCS_REDO;
      } else {
        // a finite coordinate beyond the ends of an axis that does not wrap
        // cannot be bracketed (beyond_ends, pyi_fast2d.cuh): NaN as well
        const bool far = sane(x) && sane(y) &&
                         ((y < ay.min || y > ay.max) || (!PER && (x < ax.min || x > ax.max)));
        status = (far && quiet_nan) ? BCS_NAN : BCS_REDO;
      }
    }
    // first element of the window, or -1 when the point is not computed here
    const int first = status == BCS_COMPUTE ? wx0 * ny + wy0 : -1;
    // ---- phase 2: four groups of eight points, one quad of lanes per point ----
    // Two groups per iteration of a rolled loop (register budget); the sums of
    // the four lanes of a quad are reduced with a 4 x 4 transpose-reduce: after
    // round 1 (partner piece ^ 1) a lane holds, for each pair of groups, the half
    // sum of the group of its own parity; after round 2 (partner piece ^ 2) lane
    // `piece` of quad q holds the value of group `piece`, i.e. of point 8 piece + q.
    const bool odd1 = (piece & 1) != 0;
    double half0 = 0.0, half1 = 0.0, held = 0.0;
#pragma unroll 1
    for (int s = 0; s < 4; ++s) {
      {
        double part;
        const int src = 8 * s + static_cast<int>(lane >> 2);
        const int gfirst = __shfl_sync(0xffffffffu, first, src);
        const double gtx = __shfl_sync(0xffffffffu, tx, src);
        const double gty = __shfl_sync(0xffffffffu, ty, src);
        // lanes of a quad whose point is not computed walk window (0, 0): legal addresses
        const int elem0 = gfirst < 0 ? 0 : gfirst;
        // The 16-byte pieces are aligned: a row starting on an odd element is
        // shifted by one.  With an even ny every row has the parity of its first
        // element; with an odd ny the rows alternate and the values of the two
        // parities are accumulated apart (a*: parity of row 0, b*: the other).
        const int par0 = elem0 & 1;
        int idx = elem0 - par0 + 2 * piece;
        double a0 = 0.0, a1 = 0.0, b0 = 0.0, b1 = 0.0;
  #pragma unroll
        for (int i = 0; i < 6; ++i) {
          const double cx = cubic4(p.cw[0][i + 1][0], p.cw[0][i + 1][1], p.cw[0][i + 1][2], p.cw[0][i + 1][3], gtx);
          if (EVEN || (i & 1) == 0) {
            PYI_CHECK_RANGE(idx, 2, cells, "K4s row piece");
            const V2 v = BcsPair<T>::load(data + idx);
            a0 = fma(cx, static_cast<double>(v.x), a0);
            a1 = fma(cx, static_cast<double>(v.y), a1);
          } else {
            // odd ny: rows 1, 3, 5 start on the other parity (a wrap keeps the
            // alternation: cells = nx ny has the parity of nx ... handled below)
            // (the shifted piece of a row that restarts at element 0 of the array, wy0 = 0
            // on a wrapped row, lies one past the unwrapped index: wrap the piece itself)
            int shifted = idx + (par0 ? 1 : -1);
            if (PER) shifted = shifted >= cells ? shifted - cells : shifted;
            PYI_CHECK_RANGE(shifted, 2, cells, "K4s shifted row piece");
            const V2 v = BcsPair<T>::load(data + shifted);
            b0 = fma(cx, static_cast<double>(v.x), b0);
            b1 = fma(cx, static_cast<double>(v.y), b1);
          }
          // next row; past the last row a periodic x continues at row 0
          idx += ny;
          if (PER) idx = idx >= cells ? idx - cells : idx;
        }
        // column weights, applied once per parity; values outside the six columns
        // (the run covers eight) are dropped HERE, after the accumulation: a NaN
        // next to the window must not reach the sum
        const double wm = cubic4(cy[0][0], cy[0][1], cy[0][2], cy[0][3], gty);  // column 2 piece - 1
        const double wc = cubic4(cy[1][0], cy[1][1], cy[1][2], cy[1][3], gty);  // column 2 piece
        const double wp = cubic4(cy[2][0], cy[2][1], cy[2][2], cy[2][3], gty);  // column 2 piece + 1
        {
          const int j0 = 2 * piece - par0;  // -1 .. 6
          a0 = static_cast<unsigned>(j0) < 6u ? a0 : 0.0;
          a1 = static_cast<unsigned>(j0 + 1) < 6u ? a1 : 0.0;
          part = fma(par0 ? wc : wp, a1, __dmul_rn(par0 ? wm : wc, a0));
        }
        if (!EVEN) {
          const int j1 = 2 * piece - (par0 ^ 1);
          b0 = static_cast<unsigned>(j1) < 6u ? b0 : 0.0;
          b1 = static_cast<unsigned>(j1 + 1) < 6u ? b1 : 0.0;
          part = fma(par0 ? wp : wc, b1, fma(par0 ? wc : wm, b0, part));
        }

        // round 1 of the reduction, once per pair of groups
        if ((s & 1) == 0) {
          held = part;
        } else {
          const double recv = shfl_xor_f64(odd1 ? held : part, 1);
          const double sum = __dadd_rn(odd1 ? part : held, recv);  // group (s - 1) + (piece & 1)
          if (s == 1) half0 = sum; else half1 = sum;
        }
      }
    }
    const bool odd2 = (piece & 2) != 0;
    const double r2 = shfl_xor_f64(odd2 ? half0 : half1, 2);
    const double total = __dadd_rn(odd2 ? half1 : half0, r2);  // group `piece`, point 8 piece + quad
    // lane L owns point L = 8 s + q: its value sits in lane 4 q + s
    const double mine = __shfl_sync(0xffffffffu, total, 4 * static_cast<int>(lane & 7u) + static_cast<int>(lane >> 3));
    T res = static_cast<T>(mine);  // ResultType of the grid (grid_dispatch.hpp:25-30)
    if (status != BCS_COMPUTE) {
      res = status == BCS_NAN ? RT<T>::nan() : RedoTag<T>::value();
    } else if (!isfinite(mine)) {
      // rare: a window with a NaN (-> NaN, cache.hpp:165-171) or an infinity /
      // an overflow (-> the generic code decides)
      bool has_nan = false;
      int row = wx0;
#pragma unroll 1
      for (int i = 0; i < 6; ++i) {
        const T* q = data + static_cast<long long>(row) * ny + wy0;
#pragma unroll 1
        PYI_CHECK_RANGE(q - data, 6, cells, "K4s NaN rescan");
        for (int j = 0; j < 6; ++j) has_nan = has_nan || isnan(__ldg(q + j));
        row = (PER && row + 1 == nx) ? 0 : row + 1;
      }
      res = has_nan ? RT<T>::nan() : RedoTag<T>::value();
    }
    if (active) {
      __stcs(out + k, res);
      // tagged points are also listed, so that the redo pass need not scan the output
      if (p.queue != nullptr && RedoTag<T>::is(res)) {
        const unsigned slot = atomicAdd(p.queue, 1u);
        if (slot < p.queue_cap) p.queue[1 + slot] = static_cast<unsigned>(k);
      }
    }
  }
}

static bool bicubics_eligible(const WinNdParams& p) {
  if (p.ndim != 2 || (p.dtype != PYI_F64 && p.dtype != PYI_F32) || !p.w_ok || !p.cw_ok) return false;
  const DAxis& ax = p.ax[0];
  const DAxis& ay = p.ax[1];
  if (!ax.fast_ok || !ay.fast_ok || !ax.tol_ok || !ay.tol_ok) return false;  // ascending regular f64
  if (ax.kind != 0 || ay.kind != 0 || ay.period != 0.0 || ax.n < 6 || ay.n < 6) return false;
  if (ax.period != 0.0 && !(ax.period > 0.0 && ax.step < ax.period)) return false;
  if (p.cells >= (1LL << 31) - 16) return false;  // 32-bit element indexes
  // odd ny: the rows of a window alternate parity; a periodic wrap keeps the
  // alternation only when nx is even
  if ((ay.n & 1) != 0 && ax.period != 0.0 && (ax.n & 1) != 0) return false;
  // aligned pieces of two elements
  return (reinterpret_cast<uintptr_t>(p.data) & (p.dtype == PYI_F64 ? 15 : 7)) == 0;
}

template <bool PER, int XM, int YM, bool EVEN>
static void launch_bicubics_e(const WinNdParams& p, cudaStream_t stream) {
  if (p.dtype == PYI_F32) {  // float32 grids: the default variant only
    const int blocks32 = grid_for((p.n + 31) / 32, kBcsThreads / 32, 6);
    bicubic_sep_kernel<float, PER, XM, YM, EVEN, 6><<<blocks32, kBcsThreads, 0, stream>>>(p);
    return;
  }
  // resident blocks per SM (round-2 sweep on B200: 22.7 / 25.1 / 23.4 Gpts/s at 4 / 6 / 8);
  // PYINTERP_B200_S_MINB = 4 | 6 | 8 for experiments
  int minb = 6;
  const char* e = getenv("PYINTERP_B200_S_MINB");
  if (e && e[0] >= '4' && e[0] <= '8') minb = e[0] - '0';
  const int blocks = grid_for((p.n + 31) / 32, kBcsThreads / 32, minb);
  switch (minb) {
    case 4: bicubic_sep_kernel<double, PER, XM, YM, EVEN, 4><<<blocks, kBcsThreads, 0, stream>>>(p); break;
    case 6: bicubic_sep_kernel<double, PER, XM, YM, EVEN, 6><<<blocks, kBcsThreads, 0, stream>>>(p); break;
    default: bicubic_sep_kernel<double, PER, XM, YM, EVEN, 8><<<blocks, kBcsThreads, 0, stream>>>(p); break;
  }
}

template <bool PER, int XM, int YM>
static void launch_bicubics_v(const WinNdParams& p, cudaStream_t stream) {
  if ((p.ax[1].n & 1) == 0) launch_bicubics_e<PER, XM, YM, true>(p, stream);
  else launch_bicubics_e<PER, XM, YM, false>(p, stream);
}

template <bool PER>
static void launch_bicubics_p(const WinNdParams& p, cudaStream_t stream) {
  const int xm = p.ax[0].tol_ok, ym = p.ax[1].tol_ok;
  if (xm == 2) {
    if (ym == 2) launch_bicubics_v<PER, 2, 2>(p, stream);
    else launch_bicubics_v<PER, 2, 1>(p, stream);
  } else {
    if (ym == 2) launch_bicubics_v<PER, 1, 2>(p, stream);
    else launch_bicubics_v<PER, 1, 1>(p, stream);
  }
}

// The separable pass; the caller follows it with the generic kernel in redo mode.
static cudaError_t launch_bicubics(const WinNdParams& p, cudaStream_t stream) {
  if (p.ax[0].period != 0.0) launch_bicubics_p<true>(p, stream);
  else launch_bicubics_p<false>(p, stream);
  g_launches.fetch_add(1);
  return cudaGetLastError();
}

}  // namespace pyi
