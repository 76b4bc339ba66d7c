// pyi_loess.cu -- K9: fill::loess on the device.
//
// Replaces fill::loess (cxx/include/pyinterp/fill/loess.hpp:207-253): every
// processed cell becomes the tri-cube weighted mean of the defined values in
// a (2 nx + 1) x (2 ny + 1) window around it (loess_point, :72-103), in a
// single pass or iterated from a first guess until the largest change falls
// below epsilon.
//
// Bit parity: a cell's value is a serial sum over its window in a fixed order
// (x outer, y inner), so one thread per cell reproduces the reference's
// operations exactly -- same expressions, no fused operation (the file is
// compiled with -fmad=false), IEEE division and square root.  The window
// indexes are the reference's frame_index (fill/helpers.hpp:79-110): modulo on
// a periodic x, mirrored elsewhere; note that the DISTANCE uses the wrapped /
// mirrored index itself (static_cast<T>(wx - ix), loess.hpp:88-89), so a
// wrapped neighbour is far away and weighs nothing -- reproduced as it is.
// The two reductions: compute_zonal_average (:134-145) is an Eigen sum over a
// select() expression, which the reference's build evaluates as a plain
// sequential loop in storage order (established against the reference build on
// grids of every size, tests/test_oracle_vs_ref.py) -- a chain that cannot be
// reassociated, added here by ONE thread while the block stages the terms
// (4 ms per 1e6 cells, once per call, iterative mode only);
// compute_max_difference (:161-172) is a maximum, order independent.
#include <algorithm>
#include <cstring>
#include <mutex>
#include <vector>

#include "pyi_common.cuh"

namespace pyi {

namespace {

// math::remainder for integers (math.hpp:93-97)
__device__ __forceinline__ long long floor_mod(long long x, long long y) {
  const long long r = x % y;
  return (r != 0 && (r ^ y) < 0) ? r + y : r;
}

// one entry of frame_index (fill/helpers.hpp:79-110)
__device__ __forceinline__ long long frame_at(long long idx, long long size, bool is_angle) {
  if (idx >= 0 && idx < size) return idx;
  if (is_angle) return floor_mod(idx, size);
  if (size == 1) return 0;
  const long long where = floor_mod(idx, (size - 1) * 2);
  return where >= size ? size - 2 - floor_mod(where, size) : floor_mod(where, size);
}

template <typename T>
__device__ __forceinline__ T sqrt_rn(T v);
template <>
__device__ __forceinline__ double sqrt_rn<double>(double v) { return __dsqrt_rn(v); }
template <>
__device__ __forceinline__ float sqrt_rn<float>(float v) { return __fsqrt_rn(v); }

template <typename T>
__device__ __forceinline__ T div_ieee(T a, T b);
template <>
__device__ __forceinline__ double div_ieee<double>(double a, double b) { return __ddiv_rn(a, b); }
template <>
__device__ __forceinline__ float div_ieee<float>(float a, float b) { return __fdiv_rn(a, b); }

// tricube_weight (loess.hpp:27-35)
template <typename T>
__device__ __forceinline__ T tricube(T distance) {
  if (distance > T(1)) return T(0);
  const T d3 = distance * distance * distance;
  const T tmp = T(1) - d3;
  return tmp * tmp * tmp;
}

// should_process (loess.hpp:38-52); value_type: 0 undefined, 1 defined, 2 all
template <typename T>
__device__ __forceinline__ bool should_process(T value, int value_type) {
  const bool undefined = isnan(value);
  return value_type == 2 || (value_type == 1 ? !undefined : undefined);
}

// loess_single_pass (loess.hpp:182-198): one thread per cell.
template <typename T>
__global__ void __launch_bounds__(kThreads)
loess_pass_kernel(const T* __restrict__ values, const T* __restrict__ validity, T* __restrict__ result,
                  long long nx, long long ny, long long hx, long long hy, int periodic, int value_type) {
  const long long cells = nx * ny;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  const T nx_inv = div_ieee<T>(T(1), static_cast<T>(hx));
  const T ny_inv = div_ieee<T>(T(1), static_cast<T>(hy));
  for (long long cell = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; cell < cells;
       cell += stride) {
    const long long ix = cell / ny, iy = cell - ix * ny;
    const T current = values[cell];
    if (!should_process<T>(validity[cell], value_type)) {
      result[cell] = current;
      continue;
    }
    T weighted_sum = T(0), weight_sum = T(0);
    for (long long kx = -hx; kx <= hx; ++kx) {
      const long long wx = frame_at(ix + kx, nx, periodic != 0);
      const T dx = static_cast<T>(wx - ix) * nx_inv;
      const T dx2 = dx * dx;
      const T* __restrict__ row = values + wx * ny;
      for (long long ky = -hy; ky <= hy; ++ky) {
        const long long wy = frame_at(iy + ky, ny, false);
        const T zi = row[wy];
        if (!isnan(zi)) {
          const T dy = static_cast<T>(wy - iy) * ny_inv;
          const T weight = tricube<T>(sqrt_rn<T>(dx2 + dy * dy));
          weighted_sum += weight * zi;
          weight_sum += weight;
        }
      }
    }
    result[cell] = weight_sum != T(0) ? div_ieee<T>(weighted_sum, weight_sum) : current;
  }
}

// compute_zonal_average (loess.hpp:134-145): sequential sum in storage order of
// select(isFinite, value, 0), divided by the count; 0 when nothing is finite.
// One block: every thread stages a tile of terms in shared memory (coalesced
// loads), thread 0 adds them in order -- the chain of additions is the
// reference's, only the loads run ahead of it.
constexpr int kMeanTile = 4096;
template <typename T>
__global__ void __launch_bounds__(kThreads)
loess_mean_kernel(const T* __restrict__ data, long long cells, T* __restrict__ fill_value) {
  __shared__ T terms[kMeanTile];
  __shared__ int finite[kThreads];
  long long count = 0;
  T sum = T(0);
  for (long long base = 0; base < cells; base += kMeanTile) {
    const int n = static_cast<int>(cells - base < kMeanTile ? cells - base : kMeanTile);
    int mine = 0;
    for (int i = threadIdx.x; i < n; i += kThreads) {
      const T v = data[base + i];
      const bool ok = isfinite(v);
      terms[i] = ok ? v : T(0);
      mine += ok ? 1 : 0;
    }
    finite[threadIdx.x] = mine;
    __syncthreads();
    if (threadIdx.x == 0) {
      for (int i = 0; i < n; ++i) sum = (base == 0 && i == 0) ? terms[0] : sum + terms[i];
      for (int i = 0; i < kThreads; ++i) count += finite[i];
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) *fill_value = count == 0 ? T(0) : div_ieee<T>(sum, static_cast<T>(count));
}

// apply_first_guess (loess.hpp:148-157): NaN -> fill value (fill_value == nullptr: zero)
template <typename T>
__global__ void __launch_bounds__(kThreads)
loess_guess_kernel(T* __restrict__ data, long long cells, const T* __restrict__ fill_value) {
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  const T fill = fill_value ? *fill_value : T(0);
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < cells; i += stride) {
    if (isnan(data[i])) data[i] = fill;
  }
}

// compute_max_difference (loess.hpp:161-172): max over the cells where
// |current - previous| is not NaN, 0 when there is none; bits of a
// non-negative float order like unsigned integers.
template <typename T>
__global__ void __launch_bounds__(kThreads)
loess_maxdiff_kernel(const T* __restrict__ current, const T* __restrict__ previous, long long cells,
                     unsigned long long* __restrict__ max_bits) {
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  double best = 0.0;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < cells; i += stride) {
    const T d = fabs(current[i] - previous[i]);
    if (!isnan(d)) best = fmax(best, static_cast<double>(d));
  }
  for (int o = 16; o > 0; o >>= 1) best = fmax(best, __shfl_xor_sync(0xffffffffu, best, o));
  if ((threadIdx.x & 31) == 0 && best > 0.0) {
    atomicMax(max_bits, static_cast<unsigned long long>(__double_as_longlong(best)));
  }
}

template <typename T>
cudaError_t loess_typed(const T* in, T* result, long long nx, long long ny, unsigned hx, unsigned hy,
                        int value_type, int first_guess, int periodic, unsigned max_iterations,
                        double epsilon, cudaStream_t stream) {
  const long long cells = nx * ny;
  const size_t bytes = static_cast<size_t>(cells) * sizeof(T);
  const int blocks = grid_for(cells, kThreads, 8);
  cudaError_t e = cudaMemcpyAsync(result, in, bytes, cudaMemcpyDeviceToDevice, stream);
  if (e != cudaSuccess) return e;
  auto pass = [&](const T* values, const T* validity) {
    loess_pass_kernel<T><<<blocks, kThreads, 0, stream>>>(values, validity, result, nx, ny, hx, hy, periodic,
                                                          value_type);
    g_launches.fetch_add(1);
    return cudaGetLastError();
  };
  if (max_iterations == 1) return pass(in, in);  // single pass (loess.hpp:213-216)
  // iterative filling (loess.hpp:218-250)
  char* scratch = nullptr;
  const size_t padded = (bytes + 15) & ~static_cast<size_t>(15);  // the two scalars behind the copy stay aligned
  e = cudaMallocAsync(reinterpret_cast<void**>(&scratch), padded + 64, stream);
  if (e != cudaSuccess) return e;
  T* previous = reinterpret_cast<T*>(scratch);
  T* fill_value = reinterpret_cast<T*>(scratch + padded);
  unsigned long long* max_bits = reinterpret_cast<unsigned long long*>(scratch + padded + 16);
  auto finish = [&](cudaError_t status) {
    const cudaError_t freed = cudaFreeAsync(scratch, stream);
    return status != cudaSuccess ? status : freed;
  };
  if (first_guess == PYI_FIRST_GUESS_ZONAL_AVERAGE) {
    loess_mean_kernel<T><<<1, kThreads, 0, stream>>>(result, cells, fill_value);
    loess_guess_kernel<T><<<blocks, kThreads, 0, stream>>>(result, cells, fill_value);
    g_launches.fetch_add(2);
  } else {
    loess_guess_kernel<T><<<blocks, kThreads, 0, stream>>>(result, cells, nullptr);
    g_launches.fetch_add(1);
  }
  e = cudaGetLastError();
  if (e != cudaSuccess || max_iterations == 0) return finish(e);
  for (unsigned iter = 0; iter < max_iterations; ++iter) {
    e = cudaMemcpyAsync(previous, result, bytes, cudaMemcpyDeviceToDevice, stream);
    if (e == cudaSuccess) e = cudaMemsetAsync(max_bits, 0, sizeof(unsigned long long), stream);
    if (e == cudaSuccess) e = pass(previous, in);
    if (e != cudaSuccess) return finish(e);
    loess_maxdiff_kernel<T><<<blocks, kThreads, 0, stream>>>(result, previous, cells, max_bits);
    g_launches.fetch_add(1);
    unsigned long long bits = 0;
    e = cudaMemcpyAsync(&bits, max_bits, sizeof(bits), cudaMemcpyDeviceToHost, stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
    if (e != cudaSuccess) return finish(e);
    double diff;
    std::memcpy(&diff, &bits, sizeof(diff));
    // the reference compares in T: max |difference| < static_cast<T>(epsilon)
    if (static_cast<T>(diff) < static_cast<T>(epsilon)) break;
  }
  return finish(cudaSuccess);
}

// ---- the allocation-free variant (pyi_fill_loess_static) -----------------------
// The iterative filling needs a copy of the previous iterate and two scalars;
// fill_loess takes them from cudaMallocAsync.  Callers that may not allocate
// get the same kernels over a file-scope array instead: kLoessStaticBytes for
// the copy plus 256 bytes for the scalars, resident on every GPU that loads the
// library until the process exits.  All calls share it.
constexpr size_t kLoessStaticBytes = PYI_LOESS_STATIC_MAX_BYTES;
static __device__ __align__(256) unsigned char g_loess_static[kLoessStaticBytes + 256];

constexpr int kLoessStaticDevices = 64;
std::mutex g_loess_static_mutex;
unsigned char* g_loess_static_ptr[kLoessStaticDevices] = {};

// The device address of g_loess_static on the CURRENT device (in host code the
// array's name is only a handle for the runtime): looked up once per device.
cudaError_t loess_static_address(unsigned char** ptr) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= kLoessStaticDevices) return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lock(g_loess_static_mutex);
  if (g_loess_static_ptr[dev] == nullptr) {
    void* address = nullptr;
    e = cudaGetSymbolAddress(&address, g_loess_static);
    if (e != cudaSuccess) return e;
    g_loess_static_ptr[dev] = static_cast<unsigned char*>(address);
  }
  *ptr = g_loess_static_ptr[dev];
  return cudaSuccess;
}

// loess_typed with its scratch in g_loess_static: same kernels, same order of
// launches, hence the same bits.  Nothing in the array has to survive between
// calls and nothing is read before it is written: `previous` is overwritten by
// the copy that opens every iteration, the fill value by loess_mean_kernel, and
// the running maximum is cleared at the start of every iteration.
template <typename T>
cudaError_t loess_static_typed(const T* in, T* result, long long nx, long long ny, unsigned hx, unsigned hy,
                               int value_type, int first_guess, int periodic, unsigned max_iterations,
                               double epsilon, cudaStream_t stream) {
  const long long cells = nx * ny;
  const size_t bytes = static_cast<size_t>(cells) * sizeof(T);
  if (bytes > kLoessStaticBytes) return cudaErrorInvalidValue;  // checked by the caller as well
  unsigned char* scratch = nullptr;
  cudaError_t e = loess_static_address(&scratch);
  if (e != cudaSuccess) return e;
  T* previous = reinterpret_cast<T*>(scratch);
  T* fill_value = reinterpret_cast<T*>(scratch + kLoessStaticBytes);
  unsigned long long* max_bits = reinterpret_cast<unsigned long long*>(scratch + kLoessStaticBytes + 16);
  const int blocks = grid_for(cells, kThreads, 8);
  e = cudaMemcpyAsync(result, in, bytes, cudaMemcpyDeviceToDevice, stream);
  if (e != cudaSuccess) return e;
  auto pass = [&](const T* values, const T* validity) {
    loess_pass_kernel<T><<<blocks, kThreads, 0, stream>>>(values, validity, result, nx, ny, hx, hy, periodic,
                                                          value_type);
    g_launches.fetch_add(1);
    return cudaGetLastError();
  };
  if (max_iterations == 1) return pass(in, in);
  if (first_guess == PYI_FIRST_GUESS_ZONAL_AVERAGE) {
    loess_mean_kernel<T><<<1, kThreads, 0, stream>>>(result, cells, fill_value);
    loess_guess_kernel<T><<<blocks, kThreads, 0, stream>>>(result, cells, fill_value);
    g_launches.fetch_add(2);
  } else {
    loess_guess_kernel<T><<<blocks, kThreads, 0, stream>>>(result, cells, nullptr);
    g_launches.fetch_add(1);
  }
  e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  for (unsigned iter = 0; iter < max_iterations; ++iter) {
    e = cudaMemcpyAsync(previous, result, bytes, cudaMemcpyDeviceToDevice, stream);
    if (e == cudaSuccess) e = cudaMemsetAsync(max_bits, 0, sizeof(unsigned long long), stream);
    if (e == cudaSuccess) e = pass(previous, in);
    if (e != cudaSuccess) return e;
    loess_maxdiff_kernel<T><<<blocks, kThreads, 0, stream>>>(result, previous, cells, max_bits);
    g_launches.fetch_add(1);
    unsigned long long bits = 0;
    e = cudaMemcpyAsync(&bits, max_bits, sizeof(bits), cudaMemcpyDeviceToHost, stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
    if (e != cudaSuccess) return e;
    double diff;
    std::memcpy(&diff, &bits, sizeof(diff));
    if (static_cast<T>(diff) < static_cast<T>(epsilon)) break;
  }
  return cudaSuccess;
}

}  // namespace

cudaError_t fill_loess(const void* in, void* result, int dtype, long long nx, long long ny, unsigned half_x,
                       unsigned half_y, int value_type, int first_guess, int periodic,
                       unsigned max_iterations, double epsilon, cudaStream_t stream) {
  if (dtype == PYI_F32) {
    return loess_typed<float>(static_cast<const float*>(in), static_cast<float*>(result), nx, ny, half_x,
                              half_y, value_type, first_guess, periodic, max_iterations, epsilon, stream);
  }
  return loess_typed<double>(static_cast<const double*>(in), static_cast<double*>(result), nx, ny, half_x,
                             half_y, value_type, first_guess, periodic, max_iterations, epsilon, stream);
}

cudaError_t fill_loess_static(const void* in, void* result, int dtype, long long nx, long long ny,
                              unsigned half_x, unsigned half_y, int value_type, int first_guess, int periodic,
                              unsigned max_iterations, double epsilon, cudaStream_t stream) {
  if (dtype == PYI_F32) {
    return loess_static_typed<float>(static_cast<const float*>(in), static_cast<float*>(result), nx, ny,
                                     half_x, half_y, value_type, first_guess, periodic, max_iterations,
                                     epsilon, stream);
  }
  return loess_static_typed<double>(static_cast<const double*>(in), static_cast<double*>(result), nx, ny,
                                    half_x, half_y, value_type, first_guess, periodic, max_iterations,
                                    epsilon, stream);
}

}  // namespace pyi
