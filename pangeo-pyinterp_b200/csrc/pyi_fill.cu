// pyi_fill.cu -- K8: filling of the undefined (NaN) cells of a 2-D grid on the
// device, upstream of the interpolation path.
//
// Replaces fill::gauss_seidel (cxx/include/pyinterp/fill/gauss_seidel.hpp:
// 26-213) and fill::multigrid (cxx/include/pyinterp/fill/multi_grid.hpp:
// 30-334): red-black relaxation of the discrete Laplace equation at the masked
// cells, alone or as the smoother of a V-cycle.  The update of a cell of one
// colour reads its four neighbours, all of the other colour, so a colour sweep
// is order independent and the device result is BIT-IDENTICAL to the
// reference's serial or threaded loop; residuals, restrictions and the
// convergence maxima are pointwise or order independent as well, and the one
// accumulating step (prolong_grid's "+=") is evaluated per fine cell in the
// reference's loop order.
// Colour order: the reference returns std::max(calculate(0), calculate(1))
// (gauss_seidel.hpp:103,136) whose argument evaluation order is unspecified;
// g++, the compiler of the reference's Linux builds, evaluates right to left,
// so an iteration sweeps the BLACK cells ((ix + iy) odd) first, then the red.
//
// Layout: the grid stays in its caller's row-major (x, y) layout, one uint8
// mask plane beside it, plus one list of masked cells per colour (built once
// per call: mask and lists never change).  A sweep touches the masked cells
// only, one thread per cell, one launch per colour; levels with few masked
// cells run all their sweeps in ONE single-block launch.  The convergence test
// of gauss_seidel.hpp:199-207 runs on the device (every block of an iteration's
// first sweep tests the previous iteration's maximum) and turns the sweeps
// queued behind it into no-ops, so the host only looks at the state every few
// iterations and the iteration count / final residual are exactly the
// reference's.
//
// Compiled with -fmad=false: the reference's expressions have no fused
// operation.
#include <algorithm>
#include <cstring>
#include <limits>
#include <mutex>
#include <vector>

#include "pyi_common.cuh"

namespace pyi {

namespace {

struct FillState {
  // max |correction| of iteration i in slot i % 3 (bits of a non-negative
  // double); slot 0 also serves the multigrid convergence test
  unsigned long long max_bits[3];
  unsigned long long iterations;  // completed Gauss-Seidel iterations
  double residual;                // max |correction| of the last completed iteration
  int done;                       // residual < epsilon reached
  int pad;
};

constexpr int kMaxLevels = 40;
// Levels with at most this many masked cells are smoothed by one block.
constexpr unsigned kBlockSmoothLimit = 4096;
constexpr int kBlockSmoothThreads = 1024;

// Neighbour index of fill/boundary.hpp:25-48 for the only arguments the fill
// methods ever pass: an in-range index plus or minus one.  With
// size >= 1 the reference's formulas reduce to
//   periodic_index(-1, n) = n - 1          periodic_index(n, n) = 0
//   reflective_index(-1, n) = min(1, n - 1) reflective_index(n, n) = max(n - 2, 0)
// (n == 1: everything maps to 0), which avoids their 64-bit modulo.
__device__ __forceinline__ unsigned neighbour(long long index, unsigned size, bool periodic) {
  if (index < 0) return periodic ? size - 1 : (size > 1 ? 1u : 0u);
  if (index >= static_cast<long long>(size)) return periodic || size < 2 ? 0u : size - 2;
  return static_cast<unsigned>(index);
}

// The two cell lists of a level (colour 0: red, 1: black) and their counters.
struct Lists {
  unsigned* list[2];
  unsigned* counts;
};

// Appends cell k to the list of its colour (colour < 0: not masked).  Must be
// reached by every thread of the block: one atomic per block and colour, the
// cells of a block stay in index order inside their range.
__device__ __forceinline__ void append_cell(int colour, unsigned k, const Lists& out) {
  __shared__ unsigned warp_count[2][kThreads / 32];
  __shared__ unsigned base[2];
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned b0 = __ballot_sync(0xffffffffu, colour == 0);
  const unsigned b1 = __ballot_sync(0xffffffffu, colour == 1);
  if (lane == 0) {
    warp_count[0][warp] = __popc(b0);
    warp_count[1][warp] = __popc(b1);
  }
  __syncthreads();
  if (threadIdx.x < 2) {
    unsigned total = 0;
    for (int w = 0; w < kThreads / 32; ++w) {
      const unsigned c = warp_count[threadIdx.x][w];
      warp_count[threadIdx.x][w] = total;
      total += c;
    }
    base[threadIdx.x] = total ? atomicAdd(&out.counts[threadIdx.x], total) : 0u;
  }
  __syncthreads();
  if (colour >= 0) {
    const unsigned b = colour ? b1 : b0;
    out.list[colour][base[colour] + warp_count[colour][warp] + __popc(b & ((1u << lane) - 1u))] =
        k;
  }
  __syncthreads();
}

// mask = isnan(grid) (gauss_seidel.hpp:178-181) and the per-colour cell lists;
// with first_guess == zero the masked cells are cleared in the same pass
// (gauss_seidel.hpp:183-185)
template <typename T>
__global__ void __launch_bounds__(kThreads)
fill_mask_kernel(T* __restrict__ grid, unsigned char* __restrict__ mask, unsigned n, unsigned ny,
                 int zero, Lists out) {
  const unsigned stride = gridDim.x * blockDim.x;
  for (unsigned base = blockIdx.x * blockDim.x; base < n; base += stride) {
    const unsigned k = base + threadIdx.x;
    int colour = -1;
    if (k < n) {
      const T v = grid[k];
      const bool m = v != v;
      mask[k] = m ? 1 : 0;
      if (m) {
        if (zero) grid[k] = T(0);
        const unsigned ix = k / ny;
        colour = static_cast<int>((ix + (k - ix * ny)) & 1u);
      }
    }
    append_cell(colour, k, out);
  }
}

// restrict_mask (multi_grid.hpp:101-128): a coarse cell is masked when any of
// its 2 x 2 fine cells is
__global__ void __launch_bounds__(kThreads)
fill_restrict_mask_kernel(const unsigned char* __restrict__ fine, unsigned fnx, unsigned fny,
                          unsigned char* __restrict__ coarse, unsigned cnx, unsigned cny,
                          int periodic, Lists out) {
  const unsigned n = cnx * cny;
  const unsigned stride = gridDim.x * blockDim.x;
  for (unsigned base = blockIdx.x * blockDim.x; base < n; base += stride) {
    const unsigned k = base + threadIdx.x;
    int colour = -1;
    if (k < n) {
      const unsigned ix = k / cny, iy = k - ix * cny;
      const unsigned jx0 = 2 * ix, jy0 = 2 * iy;
      const unsigned jx1 = neighbour(static_cast<long long>(jx0) + 1, fnx, periodic);
      const unsigned jy1 = neighbour(static_cast<long long>(jy0) + 1, fny, false);
      const bool m = fine[jx0 * fny + jy0] || fine[jx1 * fny + jy0] || fine[jx0 * fny + jy1] ||
                     fine[jx1 * fny + jy1];
      coarse[k] = m ? 1 : 0;
      if (m) colour = static_cast<int>((ix + iy) & 1u);
    }
    append_cell(colour, k, out);
  }
}

// set_zonal_average (fill/helpers.hpp:27-75): for every y, the defined values
// are summed along x IN ORDER (boost::accumulators sum / count -- see the
// oracle's shim header for the restated Boost algorithm) and the masked cells
// of the band take sum / count, or zero when the band is empty.  32 bands per
// block: a 64 x 32 tile is staged through shared memory with coalesced loads,
// then one thread per band accumulates its 64 values sequentially.
constexpr int kBandCols = 32;
constexpr int kBandRows = 64;

template <typename T>
__global__ void __launch_bounds__(kThreads)
fill_zonal_kernel(T* __restrict__ grid, const unsigned char* __restrict__ mask, unsigned nx,
                  unsigned ny) {
  __shared__ T tile[kBandRows][kBandCols + 1];
  __shared__ unsigned char mtile[kBandRows][kBandCols + 1];
  __shared__ T guess[kBandCols];
  const unsigned tx = threadIdx.x & (kBandCols - 1);
  const unsigned ty = threadIdx.x / kBandCols;  // 0..7
  const unsigned iy = blockIdx.x * kBandCols + tx;
  T sum = T(0);
  unsigned long long count = 0;
  for (unsigned x0 = 0; x0 < nx; x0 += kBandRows) {
    for (unsigned r = ty; r < kBandRows; r += kThreads / kBandCols) {
      const unsigned ix = x0 + r;
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
    for (unsigned ix = ty; ix < nx; ix += kThreads / kBandCols) {
      if (mask[ix * ny + iy]) grid[ix * ny + iy] = g;
    }
  }
}

__device__ __forceinline__ double block_max(double v) {
  __shared__ double part[32];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double w = __shfl_xor_sync(0xffffffffu, v, o);
    v = w > v ? w : v;
  }
  if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x < 32) {
    v = threadIdx.x < (blockDim.x >> 5) ? part[threadIdx.x] : 0.0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const double w = __shfl_xor_sync(0xffffffffu, v, o);
      v = w > v ? w : v;
    }
  }
  return v;  // valid in thread 0
}

// Geometry of one grid level
struct Geo {
  unsigned nx, ny;
  int periodic;
};

// Update of one masked cell (gauss_seidel.hpp:50-64); returns |correction|.
// `rhs` is null for the homogeneous problem: the reference then adds a zero
// matrix, kept here as "+ T(0)" so that a -0.0 sum turns into +0.0 as there.
template <typename T>
__device__ __forceinline__ double relax_cell(T* grid, const T* rhs, Geo g,
                                             unsigned ix, unsigned iy, T relaxation) {
  const unsigned at = ix * g.ny + iy;
  const unsigned ix0 = neighbour(static_cast<long long>(ix) - 1, g.nx, g.periodic);
  const unsigned ix1 = neighbour(static_cast<long long>(ix) + 1, g.nx, g.periodic);
  const unsigned iy0 = neighbour(static_cast<long long>(iy) - 1, g.ny, false);
  const unsigned iy1 = neighbour(static_cast<long long>(iy) + 1, g.ny, false);
  const T cell = grid[at];
  const T f = rhs ? rhs[at] : T(0);
  const T residual =
      (T(0.25) * (grid[ix0 * g.ny + iy] + grid[ix1 * g.ny + iy] + grid[ix * g.ny + iy0] +
                  grid[ix * g.ny + iy1] + f) -
       cell) *
      relaxation;
  grid[at] = cell + residual;
  return fabs(static_cast<double>(residual));
}

// One listed cell of a colour sweep; returns the largest |correction| applied
// (std::max semantics of gauss_seidel.hpp:64: a NaN never replaces the
// maximum).  The four neighbours of a cell have the other colour except
// across the periodic seam of an ODD x size, where rows 0 and nx - 1 share
// their colour: the serial reference updates (0, iy) from the old
// (nx - 1, iy), then (nx - 1, iy) from the new (0, iy).  The thread of a
// masked (0, iy) therefore owns both cells and updates them in that order;
// every other pair of same-colour cells is independent.
template <typename T>
__device__ __forceinline__ double relax_listed(T* grid, const T* rhs,
                                               const unsigned char* __restrict__ mask, Geo g,
                                               unsigned at, T relaxation) {
  const unsigned ix = at / g.ny, iy = at - ix * g.ny;
  const bool seam = g.periodic && (g.nx & 1u) && g.nx >= 3;
  if (seam && ix == g.nx - 1 && mask[iy]) return 0.0;  // owned by the thread of (0, iy)
  double worst = 0.0;
  double a = relax_cell<T>(grid, rhs, g, ix, iy, relaxation);
  if (a > worst) worst = a;
  if (seam && ix == 0 && mask[(g.nx - 1) * g.ny + iy]) {
    a = relax_cell<T>(grid, rhs, g, g.nx - 1, iy, relaxation);
    if (a > worst) worst = a;
  }
  return worst;
}

// One colour of gauss_seidel_core (gauss_seidel.hpp:26-91) over its cell list.
// phase 0: no bookkeeping (multigrid smoothing); 1 / 2: first / second sweep
// of Gauss-Seidel iteration `it`.  The convergence test of
// gauss_seidel.hpp:199-207 for iteration it - 1 is evaluated by every block of
// the first sweep of iteration `it` from that iteration's (final) maximum:
// when it holds, block 0 records the result and nothing is swept any more.
template <typename T>
__global__ void __launch_bounds__(kThreads)
fill_sweep_kernel(T* grid, const T* __restrict__ rhs, const unsigned char* __restrict__ mask,
                  const unsigned* __restrict__ list, unsigned count, Geo g, T relaxation,
                  T epsilon, int phase, unsigned long long it, FillState* st) {
  if (phase) {
    if (*reinterpret_cast<volatile int*>(&st->done)) return;
    if (phase == 1) {
      if (it > 0) {
        const double r = __longlong_as_double(static_cast<long long>(
            *reinterpret_cast<volatile unsigned long long*>(&st->max_bits[(it - 1) % 3])));
        if (static_cast<T>(r) < epsilon) {
          if (blockIdx.x == 0 && threadIdx.x == 0) {
            st->iterations = it;
            st->residual = r;
            st->done = 1;
          }
          return;
        }
      }
      if (blockIdx.x == 0 && threadIdx.x == 0) st->max_bits[(it + 1) % 3] = 0;
    }
  }
  const unsigned stride = gridDim.x * blockDim.x;
  double worst = 0.0;
  for (unsigned k = blockIdx.x * blockDim.x + threadIdx.x; k < count; k += stride) {
    const double a = relax_listed<T>(grid, rhs, mask, g, list[k], relaxation);
    if (a > worst) worst = a;
  }
  if (!phase) return;
  worst = block_max(worst);
  if (threadIdx.x == 0 && worst > 0.0) {
    atomicMax(&st->max_bits[it % 3],
              static_cast<unsigned long long>(__double_as_longlong(worst)));
  }
}

// Closes a batch of `it` queued iterations: records the last one for the host.
template <typename T>
__global__ void fill_close_kernel(T epsilon, unsigned long long it, FillState* st) {
  if (st->done || it == 0) return;
  const double r = __longlong_as_double(static_cast<long long>(st->max_bits[(it - 1) % 3]));
  st->iterations = it;
  st->residual = r;
  if (static_cast<T>(r) < epsilon) st->done = 1;
}

// `sweeps` full iterations (black, then red) of a level with few masked cells
// by ONE block: smooth (multi_grid.hpp:199-208) and the coarsest-level solve
// (multi_grid.hpp:228-232).
template <typename T>
__device__ __forceinline__ void smooth_in_block(T* grid, const T* rhs,
                                                const unsigned char* __restrict__ mask,
                                                const unsigned* __restrict__ black,
                                                unsigned n_black,
                                                const unsigned* __restrict__ red, unsigned n_red,
                                                Geo g, T relaxation, unsigned sweeps) {
  for (unsigned s = 0; s < sweeps; ++s) {
    for (unsigned k = threadIdx.x; k < n_black; k += blockDim.x) {
      relax_listed<T>(grid, rhs, mask, g, black[k], relaxation);
    }
    __syncthreads();
    for (unsigned k = threadIdx.x; k < n_red; k += blockDim.x) {
      relax_listed<T>(grid, rhs, mask, g, red[k], relaxation);
    }
    __syncthreads();
  }
}

template <typename T>
__global__ void __launch_bounds__(kBlockSmoothThreads)
fill_smooth_block_kernel(T* grid, const T* __restrict__ rhs,
                         const unsigned char* __restrict__ mask,
                         const unsigned* __restrict__ black, unsigned n_black,
                         const unsigned* __restrict__ red, unsigned n_red, Geo g, T relaxation,
                         unsigned sweeps) {
  smooth_in_block<T>(grid, rhs, mask, black, n_black, red, n_red, g, relaxation, sweeps);
}

// compute_residual (multi_grid.hpp:30-56): r = f - A u at the masked cells, 0
// elsewhere.  track != 0: the maximum |r| also goes to st->max_bits (the
// convergence test of multi_grid.hpp:323-327).
template <typename T>
__device__ __forceinline__ double residual_cell(T* res, const T* grid, const T* rhs,
                                                const unsigned char* __restrict__ mask, Geo g,
                                                unsigned k) {
  T r = T(0);
  if (mask[k]) {
    const unsigned ix = k / g.ny, iy = k - ix * g.ny;
    const unsigned ix0 = neighbour(static_cast<long long>(ix) - 1, g.nx, g.periodic);
    const unsigned ix1 = neighbour(static_cast<long long>(ix) + 1, g.nx, g.periodic);
    const unsigned iy0 = neighbour(static_cast<long long>(iy) - 1, g.ny, false);
    const unsigned iy1 = neighbour(static_cast<long long>(iy) + 1, g.ny, false);
    const T au = grid[k] - T(0.25) * (grid[ix0 * g.ny + iy] + grid[ix1 * g.ny + iy] +
                                      grid[ix * g.ny + iy0] + grid[ix * g.ny + iy1]);
    r = (rhs ? rhs[k] : T(0)) - au;
  }
  res[k] = r;
  return fabs(static_cast<double>(r));
}

template <typename T>
__global__ void __launch_bounds__(kThreads)
fill_residual_kernel(T* __restrict__ res, const T* __restrict__ grid, const T* __restrict__ rhs,
                     const unsigned char* __restrict__ mask, Geo g, int track, FillState* st) {
  const unsigned n = g.nx * g.ny;
  const unsigned stride = gridDim.x * blockDim.x;
  double worst = 0.0;
  for (unsigned k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += stride) {
    const double a = residual_cell<T>(res, grid, rhs, mask, g, k);
    if (a > worst) worst = a;
  }
  if (!track) return;
  worst = block_max(worst);
  if (threadIdx.x == 0 && worst > 0.0) {
    atomicMax(&st->max_bits[0], static_cast<unsigned long long>(__double_as_longlong(worst)));
  }
}

// restrict_grid (multi_grid.hpp:65-93): 2 x 2 box average of the fine residual
// into the coarse right-hand side; the coarse error starts at zero
// (multi_grid.hpp:255).
template <typename T>
__device__ __forceinline__ void restrict_cell(T* coarse_rhs, T* coarse_err, const T* fine, Geo f,
                                              unsigned cny, unsigned k) {
  const unsigned ix = k / cny, iy = k - ix * cny;
  const unsigned jx0 = 2 * ix, jy0 = 2 * iy;
  const unsigned jx1 = neighbour(static_cast<long long>(jx0) + 1, f.nx, f.periodic);
  const unsigned jy1 = neighbour(static_cast<long long>(jy0) + 1, f.ny, false);
  coarse_rhs[k] = T(0.25) * (fine[jx0 * f.ny + jy0] + fine[jx1 * f.ny + jy0] +
                             fine[jx0 * f.ny + jy1] + fine[jx1 * f.ny + jy1]);
  coarse_err[k] = T(0);
}

template <typename T>
__global__ void __launch_bounds__(kThreads)
fill_restrict_kernel(T* __restrict__ coarse_rhs, T* __restrict__ coarse_err,
                     const T* __restrict__ fine, Geo f, unsigned cnx, unsigned cny) {
  const unsigned n = cnx * cny;
  const unsigned stride = gridDim.x * blockDim.x;
  for (unsigned k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += stride) {
    restrict_cell<T>(coarse_rhs, coarse_err, fine, f, cny, k);
  }
}

// prolong_grid (multi_grid.hpp:137-186) + the masked correction
// (multi_grid.hpp:266).  The reference scatters "+=" from every coarse cell
// (ix, iy), in loop order, to the fine cells (2ix | nbr(2ix + 1)) x
// (2iy | nbr(2iy + 1)); here every masked fine cell gathers its contributions
// in that same order.  Fine sizes are > 4 here (multi_grid.hpp:228), so along
// an axis a fine index receives at most two contributions: the regular one
// (as 2i, or as 2i + 1) and, for an odd fine size, the folded 2i + 1 == size
// of the last coarse cell (periodic: onto 0, reflective: onto size - 2).
struct Contribution {
  unsigned coarse[2];  // coarse indexes, ascending
  int role[2];         // 0: as 2i, 1: as nbr(2i + 1)
  int n;
};
__device__ __forceinline__ Contribution contributions(unsigned fine, unsigned fine_size,
                                                      unsigned coarse_size, bool periodic) {
  Contribution c;
  c.coarse[0] = fine >> 1;  // regular: 2i == fine, or 2i + 1 == fine < fine_size
  c.role[0] = static_cast<int>(fine & 1u);
  c.n = 1;
  c.coarse[1] = 0;
  c.role[1] = 1;
  if (fine_size & 1u) {
    const unsigned last = coarse_size - 1;  // 2 * last + 1 == fine_size: folded
    const unsigned target = periodic ? 0u : fine_size - 2;
    if (fine == target) {  // last > fine >> 1 for every size > 4
      c.coarse[1] = last;
      c.n = 2;
    }
  }
  return c;
}

template <typename T>
__device__ __forceinline__ void prolong_cell(T* grid, const T* coarse, Geo f, unsigned cnx,
                                             unsigned cny, unsigned at) {
  const unsigned fx = at / f.ny, fy = at - fx * f.ny;
  const Contribution cx = contributions(fx, f.nx, cnx, f.periodic);
  const Contribution cy = contributions(fy, f.ny, cny, false);
  T acc = T(0);
  for (int a = 0; a < cx.n; ++a) {
    const unsigned ix = cx.coarse[a];
    const unsigned kx1 = neighbour(static_cast<long long>(ix) + 1, cnx, f.periodic);
    for (int b = 0; b < cy.n; ++b) {
      const unsigned iy = cy.coarse[b];
      const unsigned ky1 = neighbour(static_cast<long long>(iy) + 1, cny, false);
      const T c00 = coarse[ix * cny + iy];
      T v;
      if (cx.role[a] == 0 && cy.role[b] == 0) {
        v = c00;
      } else if (cx.role[a] == 1 && cy.role[b] == 0) {
        v = (c00 + coarse[kx1 * cny + iy]) * T(0.5);
      } else if (cx.role[a] == 0) {
        v = (c00 + coarse[ix * cny + ky1]) * T(0.5);
      } else {
        v = (c00 + coarse[kx1 * cny + iy] + coarse[ix * cny + ky1] + coarse[kx1 * cny + ky1]) *
            T(0.25);
      }
      acc += v;
    }
  }
  grid[at] = grid[at] + acc;
}

template <typename T>
__global__ void __launch_bounds__(kThreads)
fill_prolong_kernel(T* __restrict__ grid, const unsigned* __restrict__ list0, unsigned count0,
                    const unsigned* __restrict__ list1, unsigned count1,
                    const T* __restrict__ coarse, Geo f, unsigned cnx, unsigned cny) {
  const unsigned stride = gridDim.x * blockDim.x;
  const unsigned count = count0 + count1;
  for (unsigned k = blockIdx.x * blockDim.x + threadIdx.x; k < count; k += stride) {
    prolong_cell<T>(grid, coarse, f, cnx, cny, k < count0 ? list0[k] : list1[k - count0]);
  }
}

// The lower part of a V-cycle -- every level of at most kTailCells cells, down
// to the coarsest -- in ONE single-block launch: the same steps in the same
// order as v_cycle (multi_grid.hpp:218-272), separated by block barriers
// instead of kernel boundaries.  The error and right-hand side of the levels
// BELOW the first one are produced and consumed inside this launch
// (restriction writes every cell, prolongation reads them): they live in
// shared memory, so the 50-iteration coarse solve runs at shared-memory
// latency.
constexpr unsigned kTailCells = 8192;
constexpr int kMaxTail = 16;
constexpr size_t kTailSmemLimit = 200 * 1024;

template <typename T>
struct TailLevel {
  T* grid;  // global arrays (level 0 of the tail; the others when smem is too small)
  T* rhs;
  const unsigned char* mask;
  const unsigned* list[2];
  unsigned count[2];
  Geo geo;
  unsigned smem_offset;  // elements: grid at smem + offset, rhs at smem + offset + cells
};
template <typename T>
struct TailArgs {
  TailLevel<T> lv[kMaxTail];
  int n;
  int in_smem;  // levels >= 1 (and their residual plane) are in shared memory
  unsigned pre, post;
  unsigned res_offset;
  T relaxation;
  T* res;
};

template <typename T>
__global__ void __launch_bounds__(kBlockSmoothThreads)
fill_vcycle_tail_kernel(const __grid_constant__ TailArgs<T> a) {
  extern __shared__ __align__(16) unsigned char tail_smem[];
  T* sm = reinterpret_cast<T*>(tail_smem);
  auto grid_of = [&](int l) -> T* {
    return a.in_smem && l > 0 ? sm + a.lv[l].smem_offset : a.lv[l].grid;
  };
  auto rhs_of = [&](int l) -> T* {
    const TailLevel<T>& v = a.lv[l];
    return a.in_smem && l > 0 ? sm + v.smem_offset + v.geo.nx * v.geo.ny : v.rhs;
  };
  auto res_of = [&](int l) -> T* { return a.in_smem && l > 0 ? sm + a.res_offset : a.res; };
  const int last = a.n - 1;
  for (int l = 0; l < last; ++l) {
    const TailLevel<T>& f = a.lv[l];
    const TailLevel<T>& c = a.lv[l + 1];
    T* grid = grid_of(l);
    T* rhs = rhs_of(l);
    T* res = res_of(l);
    smooth_in_block<T>(grid, rhs, f.mask, f.list[1], f.count[1], f.list[0], f.count[0], f.geo,
                       a.relaxation, a.pre);
    const unsigned fine_cells = f.geo.nx * f.geo.ny;
    for (unsigned k = threadIdx.x; k < fine_cells; k += blockDim.x) {
      residual_cell<T>(res, grid, rhs, f.mask, f.geo, k);
    }
    __syncthreads();
    T* coarse_err = grid_of(l + 1);
    T* coarse_rhs = rhs_of(l + 1);
    const unsigned coarse_cells = c.geo.nx * c.geo.ny;
    for (unsigned k = threadIdx.x; k < coarse_cells; k += blockDim.x) {
      restrict_cell<T>(coarse_rhs, coarse_err, res, f.geo, c.geo.ny, k);
    }
    __syncthreads();
  }
  {
    const TailLevel<T>& c = a.lv[last];
    smooth_in_block<T>(grid_of(last), rhs_of(last), c.mask, c.list[1], c.count[1], c.list[0],
                       c.count[0], c.geo, T(1.0), 50);
  }
  for (int l = last - 1; l >= 0; --l) {
    const TailLevel<T>& f = a.lv[l];
    const TailLevel<T>& c = a.lv[l + 1];
    T* grid = grid_of(l);
    const T* coarse = grid_of(l + 1);
    const unsigned count = f.count[0] + f.count[1];
    for (unsigned k = threadIdx.x; k < count; k += blockDim.x) {
      prolong_cell<T>(grid, coarse, f.geo, c.geo.nx, c.geo.ny,
                      k < f.count[0] ? f.list[0][k] : f.list[1][k - f.count[0]]);
    }
    __syncthreads();
    smooth_in_block<T>(grid, rhs_of(l), f.mask, f.list[1], f.count[1], f.list[0], f.count[0],
                       f.geo, a.relaxation, a.post);
  }
}

// ---------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------

// Grow-only device arena of the fill entry points + a pinned mirror of the
// state and of the list counters.
struct FillScratch {
  int device{-1};
  char* dev{nullptr};
  size_t dev_bytes{0};
  size_t used{0};
  char* pinned{nullptr};  // FillState + 2 * kMaxLevels counters
  std::mutex mu;
  static constexpr size_t kPinnedBytes = sizeof(FillState) + 2 * kMaxLevels * sizeof(unsigned);
  void release() {
    cudaFree(dev);
    cudaFreeHost(pinned);
    dev = nullptr;
    pinned = nullptr;
    dev_bytes = 0;
  }
  ~FillScratch() { release(); }
  cudaError_t ensure(size_t bytes) {
    int current = 0;
    cudaError_t e = cudaGetDevice(&current);
    if (e != cudaSuccess) return e;
    if (current != device) {
      release();
      device = current;
    }
    if (!pinned && (e = cudaMallocHost(reinterpret_cast<void**>(&pinned), kPinnedBytes)) != cudaSuccess) {
      return e;
    }
    if (bytes > dev_bytes) {
      cudaFree(dev);
      dev = nullptr;
      dev_bytes = 0;
      if ((e = cudaMalloc(reinterpret_cast<void**>(&dev), bytes)) != cudaSuccess) return e;
      dev_bytes = bytes;
    }
    used = 0;
    return cudaSuccess;
  }
  static size_t padded(size_t bytes) { return (bytes + 255) & ~static_cast<size_t>(255); }
  template <typename P>
  P* take(size_t count) {
    P* p = reinterpret_cast<P*>(dev + used);
    used += padded(count * sizeof(P));
    return p;
  }
};

FillScratch& scratch() {
  static FillScratch s;
  return s;
}

// One grid level: the caller's grid (level 0) or a coarse error equation.
template <typename T>
struct Level {
  Geo geo{};
  unsigned cells{0};
  T* grid{nullptr};  // level 0: the caller's grid; coarser: the error
  T* rhs{nullptr};   // level 0: none (homogeneous problem)
  unsigned char* mask{nullptr};
  unsigned* list[2]{nullptr, nullptr};
  unsigned count[2]{0, 0};
};

template <typename T>
size_t level_bytes(unsigned cells, bool coarse) {
  size_t b = FillScratch::padded(cells) + 2 * FillScratch::padded((cells / 2 + 1) * sizeof(unsigned));
  if (coarse) b += 2 * FillScratch::padded(cells * sizeof(T));
  return b;
}

template <typename T>
void carve_level(FillScratch& sc, Level<T>& lv, bool coarse) {
  lv.mask = sc.take<unsigned char>(lv.cells);
  lv.list[0] = sc.take<unsigned>(lv.cells / 2 + 1);
  lv.list[1] = sc.take<unsigned>(lv.cells / 2 + 1);
  if (coarse) {
    lv.grid = sc.take<T>(lv.cells);
    lv.rhs = sc.take<T>(lv.cells);
  }
}

inline int blocks_for(unsigned work) {
  return std::max(1, grid_for(static_cast<long long>(work), kThreads, 8));
}

#define FILL_CHECK(call)                 \
  do {                                   \
    cudaError_t e__ = (call);            \
    if (e__ != cudaSuccess) return e__;  \
  } while (0)
#define FILL_LAUNCHED()           \
  do {                            \
    g_launches.fetch_add(1);      \
    FILL_CHECK(cudaGetLastError()); \
  } while (0)

// One colour sweep of a level.
template <typename T>
cudaError_t sweep(const Level<T>& lv, int colour, T relaxation, T epsilon, int phase,
                  unsigned long long it, FillState* st, cudaStream_t stream) {
  fill_sweep_kernel<T><<<blocks_for(lv.count[colour]), kThreads, 0, stream>>>(
      lv.grid, lv.rhs, lv.mask, lv.list[colour], lv.count[colour], lv.geo, relaxation, epsilon,
      phase, it, st);
  FILL_LAUNCHED();
  return cudaSuccess;
}

// smooth (multi_grid.hpp:199-208): `sweeps` iterations, black then red.
template <typename T>
cudaError_t smooth(const Level<T>& lv, T relaxation, unsigned sweeps, FillState* st,
                   cudaStream_t stream) {
  if (sweeps == 0 || lv.count[0] + lv.count[1] == 0) return cudaSuccess;
  if (lv.count[0] + lv.count[1] <= kBlockSmoothLimit) {
    fill_smooth_block_kernel<T><<<1, kBlockSmoothThreads, 0, stream>>>(
        lv.grid, lv.rhs, lv.mask, lv.list[1], lv.count[1], lv.list[0], lv.count[0], lv.geo,
        relaxation, sweeps);
    FILL_LAUNCHED();
    return cudaSuccess;
  }
  for (unsigned s = 0; s < sweeps; ++s) {
    FILL_CHECK(sweep<T>(lv, 1, relaxation, T(0), 0, 0, st, stream));
    FILL_CHECK(sweep<T>(lv, 0, relaxation, T(0), 0, 0, st, stream));
  }
  return cudaSuccess;
}

// Shared opening of both methods: state, mask + lists of level 0 (and of the
// coarse levels), counters back to the host, first guess.
template <typename T>
cudaError_t prepare(FillScratch& sc, std::vector<Level<T>>& levels, FillState* st,
                    unsigned* counts, int first_guess, cudaStream_t stream, bool* has_nan) {
  FILL_CHECK(cudaMemsetAsync(sc.dev, 0, FillScratch::padded(FillScratch::kPinnedBytes), stream));
  Level<T>& l0 = levels[0];
  fill_mask_kernel<T><<<blocks_for(l0.cells), kThreads, 0, stream>>>(
      l0.grid, l0.mask, l0.cells, l0.geo.ny, first_guess == PYI_FIRST_GUESS_ZERO,
      Lists{{l0.list[0], l0.list[1]}, counts});
  FILL_LAUNCHED();
  for (size_t l = 1; l < levels.size(); ++l) {
    const Level<T>& f = levels[l - 1];
    Level<T>& c = levels[l];
    fill_restrict_mask_kernel<<<blocks_for(c.cells), kThreads, 0, stream>>>(
        f.mask, f.geo.nx, f.geo.ny, c.mask, c.geo.nx, c.geo.ny, c.geo.periodic,
        Lists{{c.list[0], c.list[1]}, counts + 2 * l});
    FILL_LAUNCHED();
  }
  FILL_CHECK(cudaMemcpyAsync(sc.pinned, sc.dev, FillScratch::kPinnedBytes, cudaMemcpyDeviceToHost,
                             stream));
  FILL_CHECK(cudaStreamSynchronize(stream));
  const unsigned* host_counts = reinterpret_cast<const unsigned*>(sc.pinned + sizeof(FillState));
  for (size_t l = 0; l < levels.size(); ++l) {
    levels[l].count[0] = host_counts[2 * l];
    levels[l].count[1] = host_counts[2 * l + 1];
  }
  *has_nan = l0.count[0] + l0.count[1] != 0;
  if (*has_nan && first_guess == PYI_FIRST_GUESS_ZONAL_AVERAGE) {
    const int blocks = static_cast<int>((l0.geo.ny + kBandCols - 1) / kBandCols);
    fill_zonal_kernel<T><<<blocks, kThreads, 0, stream>>>(l0.grid, l0.mask, l0.geo.nx, l0.geo.ny);
    FILL_LAUNCHED();
  }
  (void)st;
  return cudaSuccess;
}

template <typename T>
cudaError_t read_state(FillScratch& sc, FillState* st, cudaStream_t stream) {
  FILL_CHECK(cudaMemcpyAsync(sc.pinned, st, sizeof(FillState), cudaMemcpyDeviceToHost, stream));
  return cudaStreamSynchronize(stream);
}

// fill::gauss_seidel (gauss_seidel.hpp:166-211)
template <typename T>
cudaError_t gauss_seidel_typed(T* grid, unsigned nx, unsigned ny, int first_guess, int periodic,
                               unsigned long long max_iterations, double epsilon,
                               double relaxation, unsigned long long* iterations,
                               double* residual, cudaStream_t stream) {
  FillScratch& sc = scratch();
  std::lock_guard<std::mutex> lock(sc.mu);
  std::vector<Level<T>> levels(1);
  Level<T>& lv = levels[0];
  lv.geo = Geo{nx, ny, periodic};
  lv.cells = nx * ny;
  lv.grid = grid;
  FILL_CHECK(sc.ensure(FillScratch::padded(FillScratch::kPinnedBytes) + level_bytes<T>(lv.cells, false)));
  char* head = sc.take<char>(FillScratch::kPinnedBytes);
  FillState* st = reinterpret_cast<FillState*>(head);
  unsigned* counts = reinterpret_cast<unsigned*>(head + sizeof(FillState));
  carve_level<T>(sc, lv, false);
  bool has_nan = false;
  FILL_CHECK(prepare<T>(sc, levels, st, counts, first_guess, stream, &has_nan));
  *iterations = 0;
  *residual = 0.0;
  if (!has_nan) return cudaSuccess;  // gauss_seidel.hpp:175-177
  // Iterations are queued in growing batches; the device stops by itself, the
  // host reads the state between batches.
  const FillState* host = reinterpret_cast<const FillState*>(sc.pinned);
  unsigned long long queued = 0;
  unsigned long long batch = 8;
  while (queued < max_iterations) {
    const unsigned long long todo = std::min(batch, max_iterations - queued);
    for (unsigned long long it = queued; it < queued + todo; ++it) {
      FILL_CHECK(sweep<T>(lv, 1, static_cast<T>(relaxation), static_cast<T>(epsilon), 1, it, st, stream));
      FILL_CHECK(sweep<T>(lv, 0, static_cast<T>(relaxation), static_cast<T>(epsilon), 2, it, st, stream));
    }
    queued += todo;
    fill_close_kernel<T><<<1, 1, 0, stream>>>(static_cast<T>(epsilon), queued, st);
    FILL_LAUNCHED();
    FILL_CHECK(read_state<T>(sc, st, stream));
    if (host->done) break;
    batch = std::min<unsigned long long>(batch * 2, 256);
  }
  if (queued > 0) {
    *iterations = host->iterations;
    *residual = host->residual;
  }
  return cudaSuccess;
}

// v_cycle (multi_grid.hpp:218-272) from level l down and back
template <typename T>
cudaError_t v_cycle(std::vector<Level<T>>& levels, size_t l, T* res, T relaxation, unsigned pre,
                    unsigned post, FillState* st, cudaStream_t stream) {
  Level<T>& lv = levels[l];
  if (lv.cells <= kTailCells && levels.size() - l <= static_cast<size_t>(kMaxTail)) {
    // this level and everything below it: one single-block launch
    TailArgs<T> a{};
    a.n = static_cast<int>(levels.size() - l);
    unsigned offset = 0;
    for (int i = 0; i < a.n; ++i) {
      const Level<T>& s = levels[l + i];
      a.lv[i] = TailLevel<T>{s.grid, s.rhs, s.mask, {s.list[0], s.list[1]}, {s.count[0], s.count[1]},
                             s.geo, offset};
      if (i > 0) offset += 2 * s.cells;
    }
    a.res_offset = offset;  // residual plane of the largest shared-memory level
    if (a.n > 1) offset += levels[l + 1].cells;
    const size_t smem = static_cast<size_t>(offset) * sizeof(T);
    a.in_smem = smem <= kTailSmemLimit;
    a.pre = pre;
    a.post = post;
    a.relaxation = relaxation;
    a.res = res;
    const size_t dynamic = a.in_smem ? smem : 0;
    if (dynamic > 48 * 1024) {
      FILL_CHECK(cudaFuncSetAttribute(fill_vcycle_tail_kernel<T>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      static_cast<int>(kTailSmemLimit)));
    }
    fill_vcycle_tail_kernel<T><<<1, kBlockSmoothThreads, dynamic, stream>>>(a);
    FILL_LAUNCHED();
    return cudaSuccess;
  }
  if (l + 1 == levels.size()) {
    // coarsest grid: kCoarseSolveIterations plain Gauss-Seidel iterations
    return smooth<T>(lv, T(1.0), 50, st, stream);
  }
  Level<T>& cv = levels[l + 1];
  FILL_CHECK(smooth<T>(lv, relaxation, pre, st, stream));
  fill_residual_kernel<T><<<blocks_for(lv.cells), kThreads, 0, stream>>>(res, lv.grid, lv.rhs,
                                                                         lv.mask, lv.geo, 0, st);
  FILL_LAUNCHED();
  fill_restrict_kernel<T><<<blocks_for(cv.cells), kThreads, 0, stream>>>(
      cv.rhs, cv.grid, res, lv.geo, cv.geo.nx, cv.geo.ny);
  FILL_LAUNCHED();
  FILL_CHECK(v_cycle<T>(levels, l + 1, res, relaxation, pre, post, st, stream));
  if (lv.count[0] + lv.count[1] != 0) {
    fill_prolong_kernel<T><<<blocks_for(lv.count[0] + lv.count[1]), kThreads, 0, stream>>>(
        lv.grid, lv.list[0], lv.count[0], lv.list[1], lv.count[1], cv.grid, lv.geo, cv.geo.nx,
        cv.geo.ny);
    FILL_LAUNCHED();
  }
  return smooth<T>(lv, relaxation, post, st, stream);
}

// fill::multigrid (multi_grid.hpp:276-334)
template <typename T>
cudaError_t multigrid_typed(T* grid, unsigned nx, unsigned ny, int first_guess, int periodic,
                            unsigned long long max_iterations, double epsilon, unsigned pre,
                            unsigned post, unsigned long long* iterations, double* residual,
                            cudaStream_t stream) {
  FillScratch& sc = scratch();
  std::lock_guard<std::mutex> lock(sc.mu);
  // level sizes: halve until one side is <= min_size = 4 (multi_grid.hpp:228,305)
  std::vector<Level<T>> levels;
  size_t bytes = FillScratch::padded(FillScratch::kPinnedBytes) +
                 FillScratch::padded(static_cast<size_t>(nx) * ny * sizeof(T));  // + residual plane
  for (unsigned x = nx, y = ny;;) {
    Level<T> lv;
    lv.geo = Geo{x, y, periodic};
    lv.cells = x * y;
    bytes += level_bytes<T>(lv.cells, !levels.empty());
    levels.push_back(lv);
    if (x <= 4 || y <= 4 || static_cast<int>(levels.size()) == kMaxLevels) break;
    x = (x + 1) / 2;
    y = (y + 1) / 2;
  }
  FILL_CHECK(sc.ensure(bytes));
  char* head = sc.take<char>(FillScratch::kPinnedBytes);
  FillState* st = reinterpret_cast<FillState*>(head);
  unsigned* counts = reinterpret_cast<unsigned*>(head + sizeof(FillState));
  T* res = sc.take<T>(static_cast<size_t>(nx) * ny);
  levels[0].grid = grid;
  for (size_t l = 0; l < levels.size(); ++l) carve_level<T>(sc, levels[l], l != 0);
  bool has_nan = false;
  FILL_CHECK(prepare<T>(sc, levels, st, counts, first_guess, stream, &has_nan));
  *iterations = 0;
  *residual = 0.0;
  if (!has_nan) return cudaSuccess;  // multi_grid.hpp:281-283
  const FillState* host = reinterpret_cast<const FillState*>(sc.pinned);
  double worst = static_cast<double>(std::numeric_limits<T>::max());  // multi_grid.hpp:311
  for (unsigned long long it = 0; it < max_iterations; ++it) {
    ++*iterations;
    FILL_CHECK(v_cycle<T>(levels, 0, res, T(1.0), pre, post, st, stream));
    // convergence: max |f - A u| over the grid (multi_grid.hpp:321-327)
    FILL_CHECK(cudaMemsetAsync(&st->max_bits[0], 0, sizeof(st->max_bits[0]), stream));
    fill_residual_kernel<T><<<blocks_for(levels[0].cells), kThreads, 0, stream>>>(
        res, levels[0].grid, levels[0].rhs, levels[0].mask, levels[0].geo, 1, st);
    FILL_LAUNCHED();
    FILL_CHECK(read_state<T>(sc, st, stream));
    std::memcpy(&worst, &host->max_bits[0], sizeof(worst));
    if (worst < epsilon) break;
  }
  *residual = worst;
  return cudaSuccess;
}

bool fill_shape_ok(long long nx, long long ny) {
  return nx > 0 && ny > 0 && nx <= 0x7fffffffLL && ny <= 0x7fffffffLL && nx * ny <= 0x7fffffffLL;
}

}  // namespace

cudaError_t fill_gauss_seidel(void* grid, int dtype, long long nx, long long ny, int first_guess,
                              int periodic, unsigned long long max_iterations, double epsilon,
                              double relaxation, unsigned long long* iterations, double* residual,
                              cudaStream_t stream) {
  if (!fill_shape_ok(nx, ny)) return cudaErrorInvalidValue;
  const unsigned ux = static_cast<unsigned>(nx), uy = static_cast<unsigned>(ny);
  if (dtype == PYI_F32) {
    return gauss_seidel_typed<float>(static_cast<float*>(grid), ux, uy, first_guess, periodic,
                                     max_iterations, epsilon, relaxation, iterations, residual,
                                     stream);
  }
  return gauss_seidel_typed<double>(static_cast<double*>(grid), ux, uy, first_guess, periodic,
                                    max_iterations, epsilon, relaxation, iterations, residual,
                                    stream);
}

cudaError_t fill_multigrid(void* grid, int dtype, long long nx, long long ny, int first_guess,
                           int periodic, unsigned long long max_iterations, double epsilon,
                           unsigned pre_smooth, unsigned post_smooth,
                           unsigned long long* iterations, double* residual,
                           cudaStream_t stream) {
  if (!fill_shape_ok(nx, ny)) return cudaErrorInvalidValue;
  const unsigned ux = static_cast<unsigned>(nx), uy = static_cast<unsigned>(ny);
  if (dtype == PYI_F32) {
    return multigrid_typed<float>(static_cast<float*>(grid), ux, uy, first_guess, periodic,
                                  max_iterations, epsilon, pre_smooth, post_smooth, iterations,
                                  residual, stream);
  }
  return multigrid_typed<double>(static_cast<double*>(grid), ux, uy, first_guess, periodic,
                                 max_iterations, epsilon, pre_smooth, post_smooth, iterations,
                                 residual, stream);
}

}  // namespace pyi
