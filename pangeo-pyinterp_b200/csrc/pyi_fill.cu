// pyi_fill.cu -- K8: filling of the undefined (NaN) cells of a 2-D grid on the
// device, upstream of the interpolation path.
//
// Replaces fill::gauss_seidel (cxx/include/pyinterp/fill/gauss_seidel.hpp:
// 26-213): red-black relaxation of the discrete Laplace equation at the masked
// cells.  The update of a cell of one colour reads its four neighbours, all of
// the other colour, so a colour sweep is order independent and the device
// result is BIT-IDENTICAL to the reference's serial or threaded loop; the
// maximum correction of an iteration is an order independent maximum as well.
// Colour order: the reference returns std::max(calculate(0), calculate(1))
// (gauss_seidel.hpp:103,136) whose argument evaluation order is unspecified;
// g++, the compiler of the reference's Linux builds, evaluates right to left,
// so an iteration sweeps the BLACK cells ((ix + iy) odd) first, then the red.
//
// Layout: the grid stays in its caller's row-major (x, y) layout, one uint8
// mask plane beside it.  One thread per cell of the swept colour; one launch
// per colour.  The convergence test of gauss_seidel.hpp:199-207 runs on the
// device (last block of the second sweep) and turns the sweeps queued behind it
// into no-ops, so the host only looks at the state every few iterations and
// the iteration count / final residual are exactly the reference's.
//
// Compiled with -fmad=false: the reference's expression has no fused operation.
#include <mutex>

#include "pyi_common.cuh"

namespace pyi {

namespace {

struct FillState {
  unsigned long long max_bits;    // max |correction| of the running iteration (double bits)
  unsigned long long iterations;  // completed iterations
  double residual;                // max |correction| of the last completed iteration
  unsigned int ticket;            // blocks of the second sweep that have finished
  int done;                       // residual < epsilon reached
  int has_nan;                    // the grid holds at least one undefined cell
  int pad;
};

// fill/boundary.hpp:25-48
__device__ __forceinline__ long long periodic_index(long long index, long long size) {
  return (index + size) % size;
}
__device__ __forceinline__ long long reflective_index(long long index, long long size) {
  if (size == 1) return 0;
  const long long period = 2 * (size - 1);
  long long normalized = index % period;
  if (normalized < 0) normalized += period;
  return normalized < size ? normalized : period - normalized;
}
__device__ __forceinline__ long long neighbour(long long index, long long size, bool periodic) {
  // interior cells need neither form
  if (index >= 0 && index < size) return index;
  return periodic ? periodic_index(index, size) : reflective_index(index, size);
}

// mask = isnan(grid) (gauss_seidel.hpp:178-181); with first_guess == zero the
// masked cells are cleared in the same pass (gauss_seidel.hpp:183-185)
template <typename T>
__global__ void __launch_bounds__(kThreads)
fill_mask_kernel(T* __restrict__ grid, unsigned char* __restrict__ mask, long long n, int zero,
                 FillState* st) {
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  bool any = false;
  for (long long k = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; k < n;
       k += stride) {
    const T v = grid[k];
    const bool m = v != v;
    mask[k] = m ? 1 : 0;
    if (m && zero) grid[k] = T(0);
    any |= m;
  }
  if (PYI_ANY(any) && (threadIdx.x & 31) == 0) st->has_nan = 1;
}

// set_zonal_average (fill/helpers.hpp:27-75): for every y, the defined values
// are summed along x IN ORDER (boost::accumulators sum / count -- see the
// oracle header for the restated Boost algorithm) and the masked cells of the
// band take sum / count, or zero when the band is empty.  32 bands per block:
// a 64 x 32 tile is staged through shared memory with coalesced loads, then
// one thread per band accumulates its 64 values sequentially.
constexpr int kBandCols = 32;
constexpr int kBandRows = 64;

template <typename T>
__global__ void __launch_bounds__(kThreads)
fill_zonal_kernel(T* __restrict__ grid, const unsigned char* __restrict__ mask, long long nx,
                  long long ny) {
  __shared__ T tile[kBandRows][kBandCols + 1];
  __shared__ unsigned char mtile[kBandRows][kBandCols + 1];
  __shared__ T guess[kBandCols];
  const int tx = threadIdx.x & (kBandCols - 1);
  const int ty = threadIdx.x / kBandCols;  // 0..7
  const long long iy0 = static_cast<long long>(blockIdx.x) * kBandCols;
  const long long iy = iy0 + tx;
  T sum = T(0);
  unsigned long long count = 0;
  for (long long x0 = 0; x0 < nx; x0 += kBandRows) {
    for (int r = ty; r < kBandRows; r += kThreads / kBandCols) {
      const long long ix = x0 + r;
      const bool in = ix < nx && iy < ny;
      tile[r][tx] = in ? grid[ix * ny + iy] : T(0);
      mtile[r][tx] = in ? mask[ix * ny + iy] : 1;
    }
    __syncthreads();
    if (ty == 0) {
#pragma unroll 8
      for (int r = 0; r < kBandRows; ++r) {
        if (!mtile[r][tx]) {
          sum += tile[r][tx];
          ++count;
        }
      }
    }
    __syncthreads();
  }
  if (ty == 0) guess[tx] = count ? sum / static_cast<T>(count) : T(0);
  __syncthreads();
  if (iy < ny) {
    const T g = guess[tx];
    for (long long ix = ty; ix < nx; ix += kThreads / kBandCols) {
      if (mask[ix * ny + iy]) grid[ix * ny + iy] = g;
    }
  }
}

__device__ __forceinline__ double block_max(double v) {
  __shared__ double part[kThreads / 32];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double w = __shfl_xor_sync(0xffffffffu, v, o);
    v = w > v ? w : v;
  }
  if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x < 32) {
    v = threadIdx.x < kThreads / 32 ? part[threadIdx.x] : 0.0;
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) {
      const double w = __shfl_xor_sync(0xffffffffu, v, o);
      v = w > v ? w : v;
    }
  }
  return v;  // valid in thread 0
}

// Update of one masked cell (gauss_seidel.hpp:50-64); returns |correction|.
// `rhs` is null for the homogeneous problem: the reference then adds a zero
// matrix, kept here as "+ T(0)" so that a -0.0 sum turns into +0.0 as there.
template <typename T, typename IDX>
__device__ __forceinline__ double relax_cell(T* grid, const T* __restrict__ rhs, IDX nx, IDX ny,
                                             IDX ix, IDX iy, int periodic, T relaxation) {
  const IDX at = ix * ny + iy;
  const IDX ix0 = static_cast<IDX>(neighbour(static_cast<long long>(ix) - 1, nx, periodic));
  const IDX ix1 = static_cast<IDX>(neighbour(static_cast<long long>(ix) + 1, nx, periodic));
  const IDX iy0 = static_cast<IDX>(neighbour(static_cast<long long>(iy) - 1, ny, false));
  const IDX iy1 = static_cast<IDX>(neighbour(static_cast<long long>(iy) + 1, ny, false));
  const T cell = grid[at];
  const T f = rhs ? rhs[at] : T(0);
  const T residual =
      (T(0.25) * (grid[ix0 * ny + iy] + grid[ix1 * ny + iy] + grid[ix * ny + iy0] +
                  grid[ix * ny + iy1] + f) -
       cell) *
      relaxation;
  grid[at] = cell + residual;
  return fabs(static_cast<double>(residual));
}

// One colour of gauss_seidel_core (gauss_seidel.hpp:26-91).  The four
// neighbours of a cell have the other colour except across the periodic seam
// of an ODD x size, where rows 0 and nx - 1 share their colour: the serial
// reference updates (0, iy) from the old (nx - 1, iy), then (nx - 1, iy) from
// the new (0, iy).  The thread of row 0 therefore owns both cells and updates
// them in that order; every other pair of same-colour cells is independent.
template <typename T, typename IDX>
__global__ void __launch_bounds__(kThreads)
fill_sweep_kernel(T* grid, const T* __restrict__ rhs, const unsigned char* __restrict__ mask,
                  IDX nx, IDX ny, IDX hy, int periodic, T relaxation, int colour, T epsilon,
                  int phase, FillState* st) {
  if (*reinterpret_cast<volatile int*>(&st->done)) return;
  const IDX total = nx * hy;
  const IDX stride = static_cast<IDX>(gridDim.x) * blockDim.x;
  const bool seam = periodic && (nx & 1) && nx >= 3;
  double worst = 0.0;
  for (IDX k = static_cast<IDX>(blockIdx.x) * blockDim.x + threadIdx.x; k < total; k += stride) {
    const IDX ix = k / hy;
    const IDX iy = 2 * (k - ix * hy) + ((ix + colour) & 1);
    if (iy >= ny) continue;
    if (seam && ix == nx - 1) continue;
    if (mask[ix * ny + iy]) {
      const double a = relax_cell<T, IDX>(grid, rhs, nx, ny, ix, iy, periodic, relaxation);
      if (a > worst) worst = a;  // std::max: a NaN never replaces the maximum
    }
    if (seam && ix == 0 && mask[(nx - 1) * ny + iy]) {
      const double a = relax_cell<T, IDX>(grid, rhs, nx, ny, nx - 1, iy, periodic, relaxation);
      if (a > worst) worst = a;
    }
  }
  // phase 0: no bookkeeping; 1: first sweep of an iteration; 2: second sweep,
  // whose last block closes the iteration
  if (!phase) return;
  worst = block_max(worst);
  __shared__ bool last;
  if (threadIdx.x == 0) {
    if (worst > 0.0) {
      atomicMax(&st->max_bits, static_cast<unsigned long long>(__double_as_longlong(worst)));
    }
    last = false;
    if (phase == 2) {
      __threadfence();
      last = atomicAdd(&st->ticket, 1u) == gridDim.x - 1;
    }
    if (last) {
      // end of the iteration (gauss_seidel.hpp:199-207)
      __threadfence();
      const double r = __longlong_as_double(static_cast<long long>(atomicMax(&st->max_bits, 0ull)));
      st->iterations += 1;
      st->residual = r;
      st->max_bits = 0;
      st->ticket = 0;
      if (static_cast<T>(r) < epsilon) st->done = 1;
      __threadfence();
    }
  }
}

// Grow-only scratch of the fill entry points (mask plane, state).
struct FillScratch {
  int device{-1};
  unsigned char* mask{nullptr};
  size_t mask_bytes{0};
  FillState* state{nullptr};
  FillState* host_state{nullptr};  // pinned
  std::mutex mu;
  void release() {
    cudaFree(mask);
    cudaFree(state);
    cudaFreeHost(host_state);
    mask = nullptr;
    state = nullptr;
    host_state = nullptr;
    mask_bytes = 0;
  }
  ~FillScratch() { release(); }
  cudaError_t ensure(size_t cells) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev != device) {
      release();
      device = dev;
    }
    if (!state) {
      if ((e = cudaMalloc(reinterpret_cast<void**>(&state), sizeof(FillState))) != cudaSuccess) return e;
      if ((e = cudaMallocHost(reinterpret_cast<void**>(&host_state), sizeof(FillState))) != cudaSuccess) return e;
    }
    if (cells > mask_bytes) {
      cudaFree(mask);
      mask = nullptr;
      mask_bytes = 0;
      if ((e = cudaMalloc(reinterpret_cast<void**>(&mask), cells)) != cudaSuccess) return e;
      mask_bytes = cells;
    }
    return cudaSuccess;
  }
};

FillScratch& scratch() {
  static FillScratch s;
  return s;
}

template <typename T>
cudaError_t sweep(T* grid, const T* rhs, const unsigned char* mask, long long nx, long long ny,
                  int periodic, T relaxation, int colour, T epsilon, int phase, FillState* st,
                  cudaStream_t stream) {
  const long long hy = (ny + 1) / 2;
  const long long total = nx * hy;
  if (total <= 0) return cudaSuccess;
  const int blocks = grid_for(total, kThreads, 8);
  if (nx * ny < (1LL << 31)) {
    fill_sweep_kernel<T, unsigned><<<blocks, kThreads, 0, stream>>>(
        grid, rhs, mask, static_cast<unsigned>(nx), static_cast<unsigned>(ny),
        static_cast<unsigned>(hy), periodic, relaxation, colour, epsilon, phase, st);
  } else {
    fill_sweep_kernel<T, long long><<<blocks, kThreads, 0, stream>>>(
        grid, rhs, mask, nx, ny, hy, periodic, relaxation, colour, epsilon, phase, st);
  }
  g_launches.fetch_add(1);
  return cudaGetLastError();
}

template <typename T>
cudaError_t first_guess_pass(T* grid, FillScratch& sc, long long nx, long long ny, int first_guess,
                             cudaStream_t stream, bool* has_nan) {
  const long long n = nx * ny;
  cudaError_t e = cudaMemsetAsync(sc.state, 0, sizeof(FillState), stream);
  if (e != cudaSuccess) return e;
  fill_mask_kernel<T><<<grid_for(n, kThreads, 8), kThreads, 0, stream>>>(
      grid, sc.mask, n, first_guess == PYI_FIRST_GUESS_ZERO, sc.state);
  g_launches.fetch_add(1);
  if ((e = cudaGetLastError()) != cudaSuccess) return e;
  if ((e = cudaMemcpyAsync(sc.host_state, sc.state, sizeof(FillState), cudaMemcpyDeviceToHost,
                           stream)) != cudaSuccess) return e;
  if ((e = cudaStreamSynchronize(stream)) != cudaSuccess) return e;
  *has_nan = sc.host_state->has_nan != 0;
  if (*has_nan && first_guess == PYI_FIRST_GUESS_ZONAL_AVERAGE) {
    const int blocks = static_cast<int>((ny + kBandCols - 1) / kBandCols);
    fill_zonal_kernel<T><<<blocks, kThreads, 0, stream>>>(grid, sc.mask, nx, ny);
    g_launches.fetch_add(1);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
  }
  return cudaSuccess;
}

template <typename T>
cudaError_t gauss_seidel_typed(T* grid, long long nx, long long ny, int first_guess, int periodic,
                               unsigned long long max_iterations, double epsilon,
                               double relaxation, unsigned long long* iterations,
                               double* residual, cudaStream_t stream) {
  FillScratch& sc = scratch();
  std::lock_guard<std::mutex> lock(sc.mu);
  cudaError_t e = sc.ensure(static_cast<size_t>(nx * ny));
  if (e != cudaSuccess) return e;
  bool has_nan = false;
  if ((e = first_guess_pass<T>(grid, sc, nx, ny, first_guess, stream, &has_nan)) != cudaSuccess) {
    return e;
  }
  *iterations = 0;
  *residual = 0.0;
  if (!has_nan) return cudaSuccess;  // gauss_seidel.hpp:175-177
  // Iterations are queued in growing batches; the device stops by itself, the
  // host reads the state between batches.
  unsigned long long queued = 0;
  unsigned long long batch = 8;
  while (queued < max_iterations) {
    const unsigned long long todo = std::min(batch, max_iterations - queued);
    for (unsigned long long it = 0; it < todo; ++it) {
      for (int colour = 1; colour >= 0; --colour) {
        e = sweep<T>(grid, nullptr, sc.mask, nx, ny, periodic, static_cast<T>(relaxation), colour,
                     static_cast<T>(epsilon), colour == 1 ? 1 : 2, sc.state, stream);
        if (e != cudaSuccess) return e;
      }
    }
    queued += todo;
    if ((e = cudaMemcpyAsync(sc.host_state, sc.state, sizeof(FillState), cudaMemcpyDeviceToHost,
                             stream)) != cudaSuccess) return e;
    if ((e = cudaStreamSynchronize(stream)) != cudaSuccess) return e;
    if (sc.host_state->done) break;
    batch = std::min<unsigned long long>(batch * 2, 256);
  }
  if (queued > 0) {
    *iterations = sc.host_state->iterations;
    *residual = sc.host_state->residual;
  }
  return cudaSuccess;
}

}  // namespace

cudaError_t fill_gauss_seidel(void* grid, int dtype, long long nx, long long ny, int first_guess,
                              int periodic, unsigned long long max_iterations, double epsilon,
                              double relaxation, unsigned long long* iterations, double* residual,
                              cudaStream_t stream) {
  if (dtype == PYI_F32) {
    return gauss_seidel_typed<float>(static_cast<float*>(grid), nx, ny, first_guess, periodic,
                                     max_iterations, epsilon, relaxation, iterations, residual,
                                     stream);
  }
  return gauss_seidel_typed<double>(static_cast<double*>(grid), nx, ny, first_guess, periodic,
                                    max_iterations, epsilon, relaxation, iterations, residual,
                                    stream);
}

}  // namespace pyi
