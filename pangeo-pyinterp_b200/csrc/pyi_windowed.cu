// pyi_windowed.cu -- translation unit of the windowed path: K4 windowed 2-D
// bicubic / bilinear (generic kernel pyi_windowed.cuh, 6 x 6 bicubic fast path
// pyi_fastwin.cuh) and K6 windowed 1-D .. 4-D with every fitting model
// (pyi_univariate.cuh, pyi_windowed_nd.cuh).
#include "pyi_common.cuh"

namespace pyi {
// Debug knob for the parity tests: PYINTERP_B200_GENERIC=1 routes windowed calls
// through the generic kernel so the fast path can be compared with it.
// PYINTERP_B200_BICUBIC=exact keeps the bit-identical bicubic fast paths
// (pyi_fastwin*.cuh) instead of the precomputed-weights kernel (pyi_bicubicw.cuh).
static bool exact_bicubic(const WinNdParams& p) {
  if (p.value_mode != PYI_VALUES_DEFAULT) return p.value_mode == PYI_VALUES_EXACT;
  const char* e = getenv("PYINTERP_B200_BICUBIC");  // per launch: the tests switch it
  return e != nullptr && e[0] == 'e';
}
static bool force_generic() {
  const char* e = getenv("PYINTERP_B200_GENERIC");
  return e != nullptr && e[0] == '1';
}
}  // namespace pyi
#include "pyi_windowed.cuh"
#include "pyi_fastwin.cuh"
#include "pyi_fastwin_nd.cuh"
#include "pyi_bicubicw.cuh"
#include "pyi_bicubics.cuh"
#include "pyi_winmodels.cuh"

namespace pyi {
// K6 (pyi_windowed_nd.cuh) is instantiated per window capacity in
// pyi_windowed_nd.cu, one object file each (PYI_ND_MAXW = 6, 12, 32)
template <int MAXW>
cudaError_t launch_windowed_nd_w(const WinNdParams& p, cudaStream_t stream);
extern template cudaError_t launch_windowed_nd_w<6>(const WinNdParams&, cudaStream_t);
extern template cudaError_t launch_windowed_nd_w<12>(const WinNdParams&, cudaStream_t);
extern template cudaError_t launch_windowed_nd_w<32>(const WinNdParams&, cudaStream_t);

cudaError_t launch_windowed_nd(const WinNdParams& p, cudaStream_t stream) {
  if (p.n <= 0) return cudaSuccess;
  // bicubic on a 6 x 6 window, 2-D .. 4-D: precomputed-weights pass, then the
  // generic kernel recomputes the points that pass tagged
  // `linear` / `bilinear` and the three `c_spline` models on a 6 x 6 window are
  // linear in the window values and cubic on a cell: in tolerance mode the
  // separable stencil (K4s) serves them with their own cubics (pyi_host.cu:
  // model_stencil_cubics); everything else about a point is the bicubic case
  const char* which = getenv("PYINTERP_B200_BICUBIC");
  const bool weights_only = which != nullptr && which[0] == 'w';
  const bool stencil_model = p.ndim == 2 && p.half_x == 3 && p.half_y == 3 && p.cw_ok &&
                             (p.method == PYI_S_LINEAR || p.method == PYI_S_BILINEAR ||
                              p.method == PYI_S_C_SPLINE || p.method == PYI_S_C_SPLINE_NOT_A_KNOT ||
                              p.method == PYI_S_C_SPLINE_PERIODIC) &&
                             !weights_only && bicubics_eligible(p);
  if (p.ndim >= 2 && (p.method == PYI_S_BICUBIC || stencil_model) && p.half_x == 3 && p.half_y == 3 &&
      p.w_ok && (p.ndim == 2 || p.planes) && !force_generic() && !exact_bicubic(p)) {
    // 2-D float64 / float32 grids with ascending axes: the separable stencil (K4s);
    // PYINTERP_B200_BICUBIC=weights keeps K4w for comparison
    const bool separable = bicubics_eligible(p) && !weights_only;
    WinNdParams q = p;
    q.queue = nullptr;
    bool own = false;
    cudaError_t e = cudaSuccess;
    if (separable && p.n < (1LL << 32)) {
      // the list of tagged points: the caller's (host pipeline slots) or one
      // from the stream-ordered allocator
      q.queue = p.queue;
      if (q.queue == nullptr) {
        e = cudaMallocAsync(reinterpret_cast<void**>(&q.queue), (kTolQueue + 1) * sizeof(unsigned), stream);
        if (e != cudaSuccess) return e;
        own = true;
      }
      q.queue_cap = kTolQueue;
      e = cudaMemsetAsync(q.queue, 0, sizeof(unsigned), stream);
    }
    if (e == cudaSuccess) e = separable ? launch_bicubics(q, stream) : launch_bicubicw(q, stream);
    if (e == cudaSuccess) {
      q.redo = 1;
      e = launch_windowed_nd_w<6>(q, stream);
    }
    if (own) {
      const cudaError_t e2 = cudaFreeAsync(q.queue, stream);
      if (e == cudaSuccess) e = e2;
    }
    return e;
  }
  // akima / akima_periodic / steffen (2-D .. 4-D) and the linear-in-values models on
  // 3-D / 4-D grids, 6 x 6 window, tolerance mode: K4m, then the redo pass (scanning)
  if (!weights_only && !force_generic() && !exact_bicubic(p) && winmodels_eligible(p)) {
    cudaError_t e = launch_winmodels(p, stream);
    if (e != cudaSuccess) return e;
    WinNdParams q = p;
    q.redo = 1;
    q.queue = nullptr;
    return launch_windowed_nd_w<6>(q, stream);
  }
  // 2-D bicubic / bilinear keep K4 and its fast path
  if (p.ndim == 2 && p.method <= PYI_S_BILINEAR) {
    WinParams q{};
    q.ax[0] = p.ax[0];
    q.ax[1] = p.ax[1];
    q.data = p.data;
    q.dtype = p.dtype;
    q.method = p.method;
    q.half_x = p.half_x;
    q.half_y = p.half_y;
    q.boundary = p.boundary;
    q.bounds_error = p.bounds_error;
    q.n = p.n;
    q.index_base = p.index_base;
    q.x = p.x;
    q.y = p.y;
    q.out = p.out;
    q.oob_key = p.oob_key;
    for (int a = 0; a < 2; ++a) {
      q.u6_ok[a] = p.u6_ok[a];
      for (int k = 0; k < 24; ++k) q.u6[a][k] = p.u6[a][k];
    }
    return launch_windowed2d(q, stream);
  }
  // 3-D / 4-D bicubic on a 6 x 6 window: fast pass over the plane-major copy,
  // then the generic kernel recomputes the points that pass tagged
  if (p.ndim >= 3 && p.method == PYI_S_BICUBIC && p.half_x == 3 && p.half_y == 3 && p.planes &&
      !force_generic()) {
    cudaError_t e = launch_bicubic6_nd_fast(p, stream);
    if (e != cudaSuccess) return e;
    WinNdParams q = p;
    q.redo = 1;
    q.queue = nullptr;  // this fast pass tags its points in the output only: the redo pass scans
    return launch_windowed_nd_w<6>(q, stream);
  }
  const int hy = p.ndim >= 2 ? p.half_y : 1;
  const int w = 2 * (p.half_x > hy ? p.half_x : hy);
  if (w <= 6) return launch_windowed_nd_w<6>(p, stream);
  if (w <= 12) return launch_windowed_nd_w<12>(p, stream);
  if (w <= 32) return launch_windowed_nd_w<32>(p, stream);
  return cudaErrorInvalidValue;
}

}  // namespace pyi
