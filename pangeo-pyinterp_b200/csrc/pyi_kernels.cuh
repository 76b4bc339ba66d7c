// pyi_kernels.cuh -- kernel parameter blocks and launch prototypes shared by
// pyi_kernels.cu (device code, -fmad=false) and pyi_host.cu (C-ABI).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "pyi_device.cuh"

namespace pyi {

// out-of-bounds report: key = (point index << 2) | axis, reduced with atomicMin
constexpr unsigned long long kNoOob = ~0ULL;

struct GeoParams {
  DAxis ax[4];
  const void* data;   // grid values, C order [x][y][z][u]
  const void* packed; // cell-packed copy (2^ndim corners per cell, result type) or nullptr
  int32_t dtype;      // pyi_dtype
  int32_t ndim;
  int32_t method, idw_exp, z_method, u_method;
  int32_t bounds_error;
  int32_t value_mode;  // pyi_value_mode as given by the caller (DEFAULT resolved at launch)
  long long n;
  long long index_base;  // global index of point 0 (chunked host calls)
  const double* x;
  const double* y;
  const void* z;  // double or int64 according to ax[2].coord_i64
  const double* u;
  void* out;      // double (f64/i64/u64 grids) or float
  unsigned long long* oob_key;
  // tolerance-mode bilinear (pyi_fast2d_tol.cuh): queue[0] counts the points the
  // hot kernel could not settle, queue[1 ..] holds the first queue_cap of them.
  // The exact kernel launched behind it only runs when the queue overflowed.
  unsigned* queue;
  unsigned queue_cap;
  // set only on the bit-identical kernel launched BEHIND the tolerance kernel:
  // it runs when *run_if_overflow > queue_cap and returns at once otherwise
  const unsigned* run_if_overflow;
};

struct WinParams {
  DAxis ax[2];
  const void* data;
  int32_t dtype;
  int32_t method;  // pyi_windowed_method
  int32_t half_x, half_y;
  int32_t boundary;
  int32_t bounds_error;
  long long n;
  long long index_base;
  const double* x;
  const double* y;
  void* out;
  unsigned long long* oob_key;
  // 24 natural-spline constants of a uniformly spaced six-node window on x / y
  // (pyi_fastwin.cuh: spline6_uniform_kernel), valid when u6_ok
  double u6[2][24];
  int32_t u6_ok[2];
};

// windowed path on 1-D .. 4-D grids, every fitting model (pyi_windowed_nd.cuh)
struct WinNdParams {
  DAxis ax[4];
  const void* data;
  int32_t dtype;
  int32_t ndim;
  int32_t method;  // pyi_spatial_method
  int32_t half_x, half_y;
  int32_t boundary;
  int32_t bounds_error;
  int32_t z_method, u_method;
  int32_t derivative;
  int32_t value_mode;  // pyi_value_mode as given by the caller
  long long n;
  long long index_base;
  const double* x;
  const double* y;
  const void* z;  // double or int64 according to ax[2].coord_i64
  const double* u;
  void* out;
  unsigned long long* oob_key;
  double u6[2][24];  // forwarded to K4 (2-D bicubic)
  int32_t u6_ok[2];
  const void* planes;  // plane-major copy [z][u][x][y] (3-D / 4-D bicubic fast path) or null
  int32_t redo;        // generic kernel: recompute only the points tagged by the fast pass
  // node-derivative weights of the six-node natural spline of the (uniform)
  // x / y axes (pyi_bicubicw.cuh), valid when w_ok
  int32_t w_ok;
  double w6[2][12];
  long long cells;  // elements of the whole grid (bounds the vector loads)
  // separable bicubic stencil (pyi_bicubics.cuh): cubic coefficients of the six
  // window weights C_j(t) of the x / y axis, entry j + 1 (entries 0 and 7 are
  // zero), valid when cw_ok
  int32_t cw_ok;
  double cw[2][8][4];
  // `polynomial` on a 6 x 6 window (pyi_winmodels.cuh): the six window weights of the x / y
  // axis as quintics in the cell coordinate (pyi_host.cu: polynomial_stencil), valid when cw_ok
  double pw[2][6][6];
  // separable pass: queue[0] counts the points it tagged for the redo pass,
  // queue[1 ..] holds the first queue_cap of them; the redo pass scans the
  // output for tags only when there is no queue or it overflowed
  unsigned* queue;
  unsigned queue_cap;
};

static_assert(sizeof(WinNdParams) <= 4000, "kernel parameters are limited to 4 KB");

struct FindParams {
  DAxis ax;
  long long n;
  const void* coords;
  long long* out;  // pairs (n x 2) or single indexes (n)
  int32_t bounded;
};

// capacity of the tolerance kernel's queue of unsettled points (entries; the
// buffer holds one counter more)
constexpr unsigned kTolQueue = 1u << 20;
cudaError_t launch_geometric(const GeoParams& p, cudaStream_t stream);
cudaError_t launch_windowed2d(const WinParams& p, cudaStream_t stream);
cudaError_t launch_windowed_nd(const WinNdParams& p, cudaStream_t stream);
// the 24 constants of a uniformly spaced six-node spline (device buffer of 24 doubles)
cudaError_t launch_spline6_uniform(double step, double* out24, cudaStream_t stream);
cudaError_t launch_find_indexes(const FindParams& p, cudaStream_t stream);
cudaError_t launch_find_index(const FindParams& p, cudaStream_t stream);
// dateutils::convert on the device (pyi_datetime.cu); bad_key = (index << 2) | op
cudaError_t launch_datetime_convert(const long long* in, long long* out, long long n,
                                    long long index_base, int src, int dst,
                                    unsigned long long* bad_key, cudaStream_t stream);
// fill::gauss_seidel on a device-resident row-major (nx, ny) grid (pyi_fill.cu)
cudaError_t fill_gauss_seidel(void* grid, int dtype, long long nx, long long ny, int first_guess,
                              int periodic, unsigned long long max_iterations, double epsilon,
                              double relaxation, unsigned long long* iterations, double* residual,
                              cudaStream_t stream);
cudaError_t fill_multigrid(void* grid, int dtype, long long nx, long long ny, int first_guess,
                           int periodic, unsigned long long max_iterations, double epsilon,
                           unsigned pre_smooth, unsigned post_smooth,
                           unsigned long long* iterations, double* residual, cudaStream_t stream);
// fill::loess on a device-resident row-major (nx, ny) grid (pyi_loess.cu): `result`
// receives the processed copy, `in` is left untouched; dtype PYI_F32 / PYI_F64
cudaError_t fill_loess(const void* in, void* result, int dtype, long long nx, long long ny,
                       unsigned half_x, unsigned half_y, int value_type, int first_guess, int periodic,
                       unsigned max_iterations, double epsilon, cudaStream_t stream);
// the same with its scratch in a file-scope device array (pyi_loess.cu): allocates nothing,
// at most PYI_LOESS_STATIC_MAX_BYTES of grid
cudaError_t fill_loess_static(const void* in, void* result, int dtype, long long nx, long long ny,
                              unsigned half_x, unsigned half_y, int value_type, int first_guess, int periodic,
                              unsigned max_iterations, double epsilon, cudaStream_t stream);
// six-node spline derivative weights of an axis (pyi_bicubicw.cuh)
cudaError_t launch_spline6_weights(double step, double* out12, cudaStream_t stream);
// plane-major copy of a 3-D / 4-D grid: [x][y][z][u] -> [z][u][x][y]
cudaError_t launch_plane_major(const void* in, void* out, int elem_size, const long long* shape,
                               cudaStream_t stream);
// cell packing (2-D .. 4-D): the 2^ndim corners of every cell contiguous, in the
// result type; x + 1 wraps (periodic seam cell), the other axes clamp.
cudaError_t launch_pack_cells(const void* data, int dtype, int ndim, const int* shape, void* packed,
                              cudaStream_t stream);
// number of kernels launched so far by this library (all threads)
long long launch_count();

}  // namespace pyi
