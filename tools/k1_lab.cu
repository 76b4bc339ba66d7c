// k1_lab.cu -- laboratory for the K1 (bivariate bilinear, 0.25 degree grid) memory
// pipeline: the arithmetic of pyi_fast2d_tol.cuh (periodic exact-node x, exact-node
// y) without the generic redo, with every knob that shapes the memory traffic as a
// template parameter.  Not part of the product; it exists to choose the launch
// shape / load policies of the product kernel from measurements.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -o tools/k1_lab tools/k1_lab.cu
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x)                                                        \
  do {                                                               \
    cudaError_t e__ = (x);                                           \
    if (e__ != cudaSuccess) {                                        \
      std::fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e__)); \
      std::exit(1);                                                  \
    }                                                                \
  } while (0)

constexpr double kMagic = 6755399441055744.0;
constexpr int NX = 1440, NY = 720;

struct P {
  const double* x;
  const double* y;
  double* out;
  const double* packed;
  unsigned groups;
};

__device__ __forceinline__ double rint_magic(double v, int& lo) {
  const double m = __dadd_rn(v, kMagic);
  lo = __double2loint(m);
  return __dsub_rn(m, kMagic);
}

enum { LD_CS = 0, LD_CA = 1, LD_NC_NA = 2, LD_L2_256 = 3, LD_CG = 4 };
enum { ST_CS = 0, ST_WB = 1, ST_CG = 2, ST_WT = 3 };
enum { CELL_NC_NA = 0, CELL_EVICT_LAST = 1, CELL_CA = 2 };

template <int POL>
__device__ __forceinline__ double2 ld2(const double* p) {
  double2 v;
  if (POL == LD_CS) asm volatile("ld.global.cs.v2.f64 {%0,%1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p));
  else if (POL == LD_CA) asm volatile("ld.global.ca.v2.f64 {%0,%1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p));
  else if (POL == LD_NC_NA)
    asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0,%1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p));
  else if (POL == LD_L2_256)
    asm volatile("ld.global.nc.L1::no_allocate.L2::256B.v2.f64 {%0,%1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p));
  else asm volatile("ld.global.cg.v2.f64 {%0,%1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p));
  return v;
}
template <int POL>
__device__ __forceinline__ void st2(double* p, double a, double b) {
  if (POL == ST_CS) asm volatile("st.global.cs.v2.f64 [%0], {%1,%2};" ::"l"(p), "d"(a), "d"(b) : "memory");
  else if (POL == ST_WB) asm volatile("st.global.wb.v2.f64 [%0], {%1,%2};" ::"l"(p), "d"(a), "d"(b) : "memory");
  else if (POL == ST_CG) asm volatile("st.global.cg.v2.f64 [%0], {%1,%2};" ::"l"(p), "d"(a), "d"(b) : "memory");
  else asm volatile("st.global.wt.v2.f64 [%0], {%1,%2};" ::"l"(p), "d"(a), "d"(b) : "memory");
}
template <int POL>
__device__ __forceinline__ void ldcell(const double* p, unsigned long long policy, double (&v)[4]) {
  if (POL == CELL_NC_NA)
    asm volatile("ld.global.nc.L1::no_allocate.v4.f64 {%0,%1,%2,%3}, [%4];"
                 : "=d"(v[0]), "=d"(v[1]), "=d"(v[2]), "=d"(v[3]) : "l"(p));
  else if (POL == CELL_EVICT_LAST)
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.f64 {%0,%1,%2,%3}, [%4], %5;"
                 : "=d"(v[0]), "=d"(v[1]), "=d"(v[2]), "=d"(v[3]) : "l"(p), "l"(policy));
  else
    asm volatile("ld.global.ca.v4.f64 {%0,%1,%2,%3}, [%4];"
                 : "=d"(v[0]), "=d"(v[1]), "=d"(v[2]), "=d"(v[3]) : "l"(p));
}

static P g_p;
static unsigned long long g_n;
static double g_peak = 6556.8;

// one point: cell index and weights (x periodic 0..360 step 0.25, y -89.875 + k 0.25)
__device__ __forceinline__ void point(double x, double y, long long& cell, double& t, double& u, bool& safe) {
  int unused, kx, ky;
  const double jf = rint_magic(fma(x, 1.0 / 360.0, -0.5), unused);
  const double c = fma(-jf, 360.0, x);
  const double q5 = fma(c, 4.0, -0.5);
  const double kf = rint_magic(q5, kx);
  const double c0 = fma(kf, 0.25, 0.0);
  const double r = fma(-jf, 360.0, __dsub_rn(x, c0));
  t = __dmul_rn(__dsub_rn(__dadd_rn(r, c0), c0), 4.0);
  const double q5y = fma(y, 4.0, 359.5 - 0.5);
  const double kfy = rint_magic(q5y, ky);
  const double y0 = fma(kfy, 0.25, -89.875);
  u = __dmul_rn(__dsub_rn(y, y0), 4.0);
  const int th = __double2hiint(t), uh = __double2hiint(u);
  safe = th > 0x3ea00000 && th < 0x3fefffff && uh > 0x3ea00000 && uh < 0x3fefffff &&
         static_cast<unsigned>(kx) < NX && static_cast<unsigned>(ky) < NY - 1;
  cell = safe ? static_cast<long long>(kx) * NY + ky : 0LL;
}

template <int THREADS, int MINB, int PTS2, int PF, int LDP, int STP, int CELLP>
__global__ void __launch_bounds__(THREADS, MINB) k1(const P p) {
  // PTS2 = number of point PAIRS per thread and iteration (1 or 2)
  unsigned long long policy = 0;
  if (CELLP == CELL_EVICT_LAST) asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(policy));
  const unsigned stride = gridDim.x * THREADS * PTS2;
  unsigned g = (blockIdx.x * THREADS + threadIdx.x) * PTS2;
  double2 xv[PTS2], yv[PTS2];
  if (g < p.groups) {
#pragma unroll
    for (int k = 0; k < PTS2; ++k) {
      xv[k] = ld2<LDP>(p.x + 2ull * (g + k));
      yv[k] = ld2<LDP>(p.y + 2ull * (g + k));
    }
  }
  while (g < p.groups) {
    const unsigned gn = g + stride;
    double2 xn[PTS2], yn[PTS2];
    if (PF && gn < p.groups) {
#pragma unroll
      for (int k = 0; k < PTS2; ++k) {
        xn[k] = ld2<LDP>(p.x + 2ull * (gn + k));
        yn[k] = ld2<LDP>(p.y + 2ull * (gn + k));
      }
    }
    long long cell[2 * PTS2];
    double t[2 * PTS2], u[2 * PTS2];
    bool safe[2 * PTS2];
#pragma unroll
    for (int k = 0; k < PTS2; ++k) {
      point(xv[k].x, yv[k].x, cell[2 * k], t[2 * k], u[2 * k], safe[2 * k]);
      point(xv[k].y, yv[k].y, cell[2 * k + 1], t[2 * k + 1], u[2 * k + 1], safe[2 * k + 1]);
    }
    double q[2 * PTS2][4];
#pragma unroll
    for (int k = 0; k < 2 * PTS2; ++k) ldcell<CELLP>(p.packed + 4 * cell[k], policy, q[k]);
    double o[2 * PTS2];
#pragma unroll
    for (int k = 0; k < 2 * PTS2; ++k) {
      const double omt = __dsub_rn(1.0, t[k]), omu = __dsub_rn(1.0, u[k]);
      const double a = fma(u[k], q[k][1], __dmul_rn(omu, q[k][0]));
      const double b = fma(u[k], q[k][3], __dmul_rn(omu, q[k][2]));
      o[k] = safe[k] ? fma(t[k], b, __dmul_rn(omt, a)) : __longlong_as_double(0x7ff8000000000000LL);
    }
#pragma unroll
    for (int k = 0; k < PTS2; ++k) st2<STP>(p.out + 2ull * (g + k), o[2 * k], o[2 * k + 1]);
    g = gn;
    if (PF) {
#pragma unroll
      for (int k = 0; k < PTS2; ++k) {
        xv[k] = xn[k];
        yv[k] = yn[k];
      }
    } else if (g < p.groups) {
#pragma unroll
      for (int k = 0; k < PTS2; ++k) {
        xv[k] = ld2<LDP>(p.x + 2ull * (g + k));
        yv[k] = ld2<LDP>(p.y + 2ull * (g + k));
      }
    }
  }
}

__global__ void fill(double* x, double* y, unsigned long long n) {
  for (unsigned long long i = blockIdx.x * 256ull + threadIdx.x; i < n; i += gridDim.x * 256ull) {
    unsigned long long h = i * 0x9E3779B97F4A7C15ull;
    h ^= h >> 29;
    h *= 0xBF58476D1CE4E5B9ull;
    h ^= h >> 32;
    x[i] = (h & 0xffffffffull) * (720.0 / 4294967296.0) - 180.0;
    y[i] = (h >> 32) * (179.0 / 4294967296.0) - 89.5;
  }
}
__global__ void fillg(double* g, unsigned long long n) {
  for (unsigned long long i = blockIdx.x * 256ull + threadIdx.x; i < n; i += gridDim.x * 256ull) g[i] = (i % 1000) * 1e-3;
}


template <int THREADS, int MINB, int PTS2, int PF, int LDP, int STP, int CELLP>
void run(const char* name, int blocks_per_sm) {
  int sms = 148;
  const int blocks = sms * blocks_per_sm;
  cudaEvent_t a, b;
  CK(cudaEventCreate(&a));
  CK(cudaEventCreate(&b));
  for (int w = 0; w < 2; ++w) k1<THREADS, MINB, PTS2, PF, LDP, STP, CELLP><<<blocks, THREADS>>>(g_p);
  CK(cudaDeviceSynchronize());
  const int steps = 6;
  CK(cudaEventRecord(a));
  for (int s = 0; s < steps; ++s) k1<THREADS, MINB, PTS2, PF, LDP, STP, CELLP><<<blocks, THREADS>>>(g_p);
  CK(cudaEventRecord(b));
  CK(cudaDeviceSynchronize());
  float ms = 0;
  CK(cudaEventElapsedTime(&ms, a, b));
  ms /= steps;
  cudaFuncAttributes fa;
  CK(cudaFuncGetAttributes(&fa, k1<THREADS, MINB, PTS2, PF, LDP, STP, CELLP>));
  std::printf("%-44s thr=%3d blk/SM=%d regs=%3d  %7.3f ms %7.1f Gpts/s frac %.4f\n", name, THREADS, blocks_per_sm,
              fa.numRegs, ms, g_n / ms / 1e6, g_n * 24.0 / ms / 1e6 / g_peak);
  std::fflush(stdout);
}


// ---------------------------------------------------------------------------
// TMA (cp.async.bulk) staged streams: a persistent block walks tiles of
// THREADS * 2 points; one thread arms an mbarrier and issues two bulk copies
// (x tile, y tile) per stage, STAGES tiles ahead; every thread reads its pair
// from shared memory, gathers its cells from L2 and stores its results.
// ---------------------------------------------------------------------------
__device__ __forceinline__ unsigned smem_u32(const void* p) {
  return static_cast<unsigned>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, unsigned long long* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
      "@P1 bra DONE;\n"
      "bra WAIT_LOOP;\n"
      "DONE:\n"
      "}" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void bulk_s2g(void* dst, const void* src, unsigned bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_u32(src)),
               "r"(bytes)
               : "memory");
}

template <int THREADS, int MINB, int STAGES, int STP, bool TMA_STORE>
__global__ void __launch_bounds__(THREADS, MINB) k1_tma(const P p) {
  constexpr int TILE = THREADS * 2;  // points per tile
  __shared__ alignas(128) double sx[STAGES][TILE];
  __shared__ alignas(128) double sy[STAGES][TILE];
  __shared__ alignas(128) double so[TMA_STORE ? 2 : 1][TMA_STORE ? TILE : 2];
  __shared__ alignas(8) unsigned long long full[STAGES];
  const unsigned tiles = p.groups / THREADS;  // whole tiles only (the lab uses n % TILE == 0)
  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) mbar_init(&full[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  unsigned issue = blockIdx.x;  // next tile to request
  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s, issue += gridDim.x) {
      if (issue < tiles) {
        mbar_expect_tx(&full[s], TILE * 16);
        bulk_g2s(sx[s], p.x + static_cast<size_t>(issue) * TILE, TILE * 8, &full[s]);
        bulk_g2s(sy[s], p.y + static_cast<size_t>(issue) * TILE, TILE * 8, &full[s]);
      }
    }
  }
  int stage = 0;
  unsigned parity = 0, it = 0;
  for (unsigned tile = blockIdx.x; tile < tiles; tile += gridDim.x, ++it) {
    mbar_wait(&full[stage], parity);
    const double2 xv = *reinterpret_cast<const double2*>(&sx[stage][2 * threadIdx.x]);
    const double2 yv = *reinterpret_cast<const double2*>(&sy[stage][2 * threadIdx.x]);
    long long cell[2];
    double t[2], u[2];
    bool safe[2];
    point(xv.x, yv.x, cell[0], t[0], u[0], safe[0]);
    point(xv.y, yv.y, cell[1], t[1], u[1], safe[1]);
    double q[2][4];
    ldcell<CELL_NC_NA>(p.packed + 4 * cell[0], 0, q[0]);
    ldcell<CELL_NC_NA>(p.packed + 4 * cell[1], 0, q[1]);
    __syncthreads();  // every thread has read its pair: the stage can be refilled
    if (threadIdx.x == 0) {
      const unsigned next = tile + STAGES * gridDim.x;
      if (next < tiles) {
        mbar_expect_tx(&full[stage], TILE * 16);
        bulk_g2s(sx[stage], p.x + static_cast<size_t>(next) * TILE, TILE * 8, &full[stage]);
        bulk_g2s(sy[stage], p.y + static_cast<size_t>(next) * TILE, TILE * 8, &full[stage]);
      }
    }
    double o[2];
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const double omt = __dsub_rn(1.0, t[k]), omu = __dsub_rn(1.0, u[k]);
      const double a = fma(u[k], q[k][1], __dmul_rn(omu, q[k][0]));
      const double b = fma(u[k], q[k][3], __dmul_rn(omu, q[k][2]));
      o[k] = safe[k] ? fma(t[k], b, __dmul_rn(omt, a)) : __longlong_as_double(0x7ff8000000000000LL);
    }
    if (TMA_STORE) {
      const int ob = it & 1;
      // the bulk store that read this buffer two tiles ago must have finished reading it
      if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
      __syncthreads();
      *reinterpret_cast<double2*>(&so[ob][2 * threadIdx.x]) = make_double2(o[0], o[1]);
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncthreads();
      if (threadIdx.x == 0) {
        bulk_s2g(p.out + static_cast<size_t>(tile) * TILE, so[ob], TILE * 8);
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
    } else {
      st2<STP>(p.out + static_cast<size_t>(tile) * TILE + 2 * threadIdx.x, o[0], o[1]);
    }
    if (++stage == STAGES) {
      stage = 0;
      parity ^= 1;
    }
  }
  if (TMA_STORE && threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

template <int THREADS, int MINB, int STAGES, int STP, bool TMA_STORE>
void run_tma(const char* name, int blocks_per_sm) {
  const int blocks = 148 * blocks_per_sm;
  cudaEvent_t a, b;
  CK(cudaEventCreate(&a));
  CK(cudaEventCreate(&b));
  for (int w = 0; w < 2; ++w) k1_tma<THREADS, MINB, STAGES, STP, TMA_STORE><<<blocks, THREADS>>>(g_p);
  CK(cudaDeviceSynchronize());
  const int steps = 6;
  CK(cudaEventRecord(a));
  for (int s = 0; s < steps; ++s) k1_tma<THREADS, MINB, STAGES, STP, TMA_STORE><<<blocks, THREADS>>>(g_p);
  CK(cudaEventRecord(b));
  CK(cudaDeviceSynchronize());
  float ms = 0;
  CK(cudaEventElapsedTime(&ms, a, b));
  ms /= steps;
  cudaFuncAttributes fa;
  CK(cudaFuncGetAttributes(&fa, k1_tma<THREADS, MINB, STAGES, STP, TMA_STORE>));
  std::printf("%-44s thr=%3d blk/SM=%d regs=%3d  %7.3f ms %7.1f Gpts/s frac %.4f\n", name, THREADS, blocks_per_sm,
              fa.numRegs, ms, g_n / ms / 1e6, g_n * 24.0 / ms / 1e6 / g_peak);
  std::fflush(stdout);
}

int main(int argc, char** argv) {
  g_n = argc > 1 ? static_cast<unsigned long long>(std::atof(argv[1])) : 1000000000ull;
  double *x, *y, *out, *packed;
  CK(cudaMalloc(&x, g_n * 8));
  CK(cudaMalloc(&y, g_n * 8));
  CK(cudaMalloc(&out, g_n * 8));
  CK(cudaMalloc(&packed, 4ull * NX * NY * 8));
  fill<<<148 * 8, 256>>>(x, y, g_n);
  fillg<<<148 * 8, 256>>>(packed, 4ull * NX * NY);
  CK(cudaDeviceSynchronize());
  g_p = P{x, y, out, packed, static_cast<unsigned>(g_n / 2)};
  const bool all = argc > 2;
  // THREADS, MINB, PTS2, PF, LDP, STP, CELLP
  run<256, 3, 1, 0, LD_CS, ST_CS, CELL_NC_NA>("base 2pt nopf cs/cs", 3);
  run<256, 3, 1, 1, LD_CS, ST_CS, CELL_NC_NA>("2pt pf cs/cs", 3);
  run<256, 4, 1, 0, LD_CS, ST_CS, CELL_NC_NA>("2pt nopf 4blk", 4);
  run<256, 4, 1, 1, LD_CS, ST_CS, CELL_NC_NA>("2pt pf 4blk", 4);
  run<256, 5, 1, 0, LD_CS, ST_CS, CELL_NC_NA>("2pt nopf 5blk", 5);
  run<256, 5, 1, 1, LD_CS, ST_CS, CELL_NC_NA>("2pt pf 5blk", 5);
  run<256, 6, 1, 0, LD_CS, ST_CS, CELL_NC_NA>("2pt nopf 6blk", 6);
  run<128, 8, 1, 1, LD_CS, ST_CS, CELL_NC_NA>("2pt pf 128thr x8", 8);
  run<128, 10, 1, 0, LD_CS, ST_CS, CELL_NC_NA>("2pt nopf 128thr x10", 10);
  run<512, 2, 1, 1, LD_CS, ST_CS, CELL_NC_NA>("2pt pf 512thr x2", 2);
  run_tma<256, 3, 2, ST_CS, false>("tma 2 stages, st.cs", 3);
  run_tma<256, 3, 4, ST_CS, false>("tma 4 stages, st.cs", 3);
  run_tma<256, 4, 4, ST_CS, false>("tma 4 stages, st.cs 4blk", 4);
  run_tma<256, 4, 3, ST_CS, false>("tma 3 stages, st.cs 4blk", 4);
  run_tma<256, 5, 3, ST_CS, false>("tma 3 stages, st.cs 5blk", 5);
  run_tma<256, 6, 2, ST_CS, false>("tma 2 stages, st.cs 6blk", 6);
  run_tma<256, 3, 5, ST_CS, false>("tma 5 stages, st.cs 3blk", 3);
  run_tma<128, 8, 4, ST_CS, false>("tma 4 stages 128thr x8", 8);
  run_tma<128, 6, 4, ST_CS, false>("tma 4 stages 128thr x6", 6);
  run_tma<512, 2, 2, ST_CS, false>("tma 2 stages 512thr x2", 2);
  run_tma<256, 3, 4, ST_CS, true>("tma 4 stages, tma store 3blk", 3);
  run_tma<256, 4, 4, ST_CS, true>("tma 4 stages, tma store 4blk", 4);
  run_tma<256, 5, 3, ST_CS, true>("tma 3 stages, tma store 5blk", 5);
  run_tma<128, 8, 4, ST_CS, true>("tma 4 stages, tma store 128thr x8", 8);
  (void)all;
  return 0;
}
