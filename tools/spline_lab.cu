// spline_lab.cu -- the node derivatives of the natural spline through six nodes, computed
// by the generic code (SplineAxis<6> + spline_two_nodes, pyi_windowed.cuh) and by the fast
// path (spline6_prepare + spline6_node_derivatives, pyi_fastwin.cuh), on the window
// coordinates of a periodic axis start + k * (360 / nx) re-centred on the query.
// nvcc -O3 -std=c++17 -fmad=false -gencode arch=compute_100a,code=sm_100a -I../pangeo-pyinterp_b200/csrc -o spline_lab spline_lab.cu
#include <cstdio>
#include "pyi_common.cuh"
namespace pyi { std::atomic<long long> g_launches{0}; }
#define PYI_WINDOWED_DEVICE_ONLY
#include "pyi_windowed.cuh"
#include "pyi_fastwin.cuh"

using namespace pyi;

__device__ unsigned long long mixu(unsigned long long z) {
  z += 0x9e3779b97f4a7c15ULL;
  z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
  z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
  return z ^ (z >> 31);
}
__device__ double uni01(unsigned long long b) { return (b >> 11) * (1.0 / 9007199254740992.0); }

__global__ void lab(int nx, double start, unsigned long long* counts, double* dump) {
  const unsigned long long id = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x;
  const double step = 360.0 / nx;
  const int cell = static_cast<int>(id % nx);
  // query inside `cell`, base period
  const double x = start + (cell + 0.02 + 0.96 * uni01(mixu(id))) * step;
  const double lo = x - 180.0, hi = x + 180.0;
  double xa[6], ya[6];
  for (int k = 0; k < 6; ++k) {
    int i = cell - 2 + k;
    i = ((i % nx) + nx) % nx;
    double raw = __dadd_rn(start, __dmul_rn(static_cast<double>(i), step));
    if (!(raw >= lo && raw < hi)) raw = normalize_period(raw, lo, 360.0, 1.0 / 360.0);
    xa[k] = raw;
    ya[k] = (uni01(mixu(id * 7 + k + 1)) - 0.5) * 40.0;
  }
  SplineAxis<6> sa;
  sa.prepare(xa, 6);
  double g2, g3;
  spline_two_nodes<6>(sa, xa, ya, 2, 3, g2, g3);
  double v[S6_COUNT];
  spline6_prepare(xa, v);
  double f2, f3;
  spline6_node_derivatives(v, ya, f2, f3);
  if (g2 != f2 || g3 != f3) {
    const unsigned long long slot = atomicAdd(counts + cell, 1ULL);
    if (atomicAdd(counts + 63, 1ULL) < 2) {
      double* d = dump + 40 * (slot & 1);
      for (int k = 0; k < 6; ++k) { d[k] = xa[k]; d[6 + k] = ya[k]; }
      d[12] = g2; d[13] = f2; d[14] = g3; d[15] = f3; d[16] = cell;
      // constants of both
      for (int k = 0; k < 5; ++k) { d[17 + k] = sa.g[k]; d[22 + k] = v[S6_G + k]; }
      d[27] = sa.inv_m[0]; d[28] = v[S6_IM]; d[29] = sa.inv_m[1]; d[30] = v[S6_IM + 1];
      d[31] = sa.inv_m[2]; d[32] = v[S6_IM + 2]; d[33] = sa.last_den; d[34] = v[S6_LAST];
      d[35] = sa.cfac[0]; d[36] = v[S6_CF]; d[37] = sa.cfac[1]; d[38] = v[S6_CF + 1];
    }
  }
}

int main() {
  unsigned long long* counts;
  double* dump;
  cudaMallocManaged(&counts, 64 * sizeof(*counts));
  cudaMallocManaged(&dump, 80 * sizeof(double));
  const int nxs[] = {21, 7, 11, 24, 13, 17, 19, 23, 27, 31};
  for (int nx : nxs) {
    for (int i = 0; i < 64; ++i) counts[i] = 0;
    lab<<<2048, 256>>>(nx, 10.0, counts, dump);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("cuda error %s\n", cudaGetErrorString(e)); return 1; }
    unsigned long long total = 0;
    for (int c = 0; c < nx; ++c) total += counts[c];
    printf("nx=%d: %llu of %d windows differ; per cell:", nx, total, 2048 * 256);
    for (int c = 0; c < nx; ++c) if (counts[c]) printf(" [%d]=%llu", c, counts[c]);
    printf("\n");
    if (total) {
      const double* d = dump;
      printf("  cell %.0f xa = %.17g %.17g %.17g %.17g %.17g %.17g\n", d[16], d[0], d[1], d[2], d[3], d[4], d[5]);
      printf("  ya = %.17g %.17g %.17g %.17g %.17g %.17g\n", d[6], d[7], d[8], d[9], d[10], d[11]);
      printf("  node 2: generic %.17g fast %.17g | node 3: generic %.17g fast %.17g\n", d[12], d[13], d[14], d[15]);
      for (int k = 0; k < 5; ++k) printf("  g[%d] generic %.17g fast %.17g %s\n", k, d[17 + k], d[22 + k], d[17 + k] == d[22 + k] ? "" : "<<<");
      const char* names[] = {"inv_m0", "inv_m1", "inv_m2", "last_den", "cfac0", "cfac1"};
      for (int k = 0; k < 6; ++k) printf("  %s generic %.17g fast %.17g %s\n", names[k], d[27 + 2 * k], d[28 + 2 * k], d[27 + 2 * k] == d[28 + 2 * k] ? "" : "<<<");
    }
  }
  return 0;
}
