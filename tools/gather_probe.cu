// gather_probe.cu -- measurement tool (not product code): how fast can a B200
// serve random aligned 32-byte (or 64-byte) gathers from a table far larger
// than L2, as a function of the load flavour and of the loads in flight per
// thread?  Guides the 3-D / 4-D corner gather design.
//
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o gather_probe gather_probe.cu
//   ./gather_probe [table_GiB=12] [points=2e8]
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>

__device__ __forceinline__ unsigned long long mix(unsigned long long z) {
  z += 0x9e3779b97f4a7c15ULL;
  z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
  z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
  return z ^ (z >> 31);
}

// MODE 0: ld.global.nc.L1::no_allocate.v4 x (BYTES/16)
// MODE 1: ld.global.cg.v4
// MODE 2: ld.global.nc.L2::64B.v4     (prefetch-size hint)
// MODE 3: cp.async.cg 16 B -> shared, wait_all, read back
// MODE 4: ld.global.cs.v4 (evict-first)
// MODE 5: ld.global.nc.L1::no_allocate.v4.u64            (one 256-bit load per 32 B)
// MODE 6: ld.global.nc.L1::no_allocate.L2::64B.v4.u64
// MODE 7: ld.global.cs.v4.u64
// MODE 8: ld.global.nc.L1::no_allocate.L2::cache_hint.L2::64B.v4.u64 with an evict_first policy
// MODE 9: ld.global.cs.L2::64B.v4.u64
template <int MODE>
__device__ __forceinline__ void load32(const void* p, unsigned long long (&w)[4], unsigned long long pol) {
  if (MODE == 5) {
    asm volatile("ld.global.nc.L1::no_allocate.v4.u64 {%0,%1,%2,%3}, [%4];"
                 : "=l"(w[0]), "=l"(w[1]), "=l"(w[2]), "=l"(w[3]) : "l"(p));
  } else if (MODE == 6) {
    asm volatile("ld.global.nc.L1::no_allocate.L2::64B.v4.u64 {%0,%1,%2,%3}, [%4];"
                 : "=l"(w[0]), "=l"(w[1]), "=l"(w[2]), "=l"(w[3]) : "l"(p));
  } else if (MODE == 7) {
    asm volatile("ld.global.cs.v4.u64 {%0,%1,%2,%3}, [%4];"
                 : "=l"(w[0]), "=l"(w[1]), "=l"(w[2]), "=l"(w[3]) : "l"(p));
  } else if (MODE == 8) {
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.L2::64B.v4.u64 {%0,%1,%2,%3}, [%4], %5;"
                 : "=l"(w[0]), "=l"(w[1]), "=l"(w[2]), "=l"(w[3]) : "l"(p), "l"(pol));
  } else {
    asm volatile("ld.global.cs.L2::64B.v4.u64 {%0,%1,%2,%3}, [%4];"
                 : "=l"(w[0]), "=l"(w[1]), "=l"(w[2]), "=l"(w[3]) : "l"(p));
  }
}

template <int MODE, int P, int BYTES>
__global__ void __launch_bounds__(256) probe(const uint4* __restrict__ table, unsigned long long cells,
                                            long long n, float* __restrict__ out) {
  constexpr int Q = BYTES / 16;
  extern __shared__ uint4 smem[];
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x * P;
  for (long long base = (static_cast<long long>(blockIdx.x) * blockDim.x) * P + threadIdx.x; base < n;
       base += stride) {
    uint4 v[P][Q];
    unsigned long long pol = 0;
    if (MODE == 8) asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
#pragma unroll
    for (int j = 0; j < P; ++j) {
      const unsigned long long k = base + static_cast<long long>(j) * blockDim.x;
      const unsigned long long cell = mix(k) % cells;
      const uint4* src = table + cell * Q;
      if (MODE >= 5) {
#pragma unroll
        for (int q = 0; q < Q; q += 2) {
          unsigned long long w[4];
          load32<MODE>(src + q, w, pol);
          v[j][q] = make_uint4((unsigned)w[0], (unsigned)(w[0] >> 32), (unsigned)w[1], (unsigned)(w[1] >> 32));
          v[j][q + 1] = make_uint4((unsigned)w[2], (unsigned)(w[2] >> 32), (unsigned)w[3], (unsigned)(w[3] >> 32));
        }
        continue;
      }
#pragma unroll
      for (int q = 0; q < Q; ++q) {
        if (MODE == 0) {
          asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                       : "=r"(v[j][q].x), "=r"(v[j][q].y), "=r"(v[j][q].z), "=r"(v[j][q].w) : "l"(src + q));
        } else if (MODE == 1) {
          asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];"
                       : "=r"(v[j][q].x), "=r"(v[j][q].y), "=r"(v[j][q].z), "=r"(v[j][q].w) : "l"(src + q));
        } else if (MODE == 2) {
          asm volatile("ld.global.nc.L2::64B.v4.u32 {%0,%1,%2,%3}, [%4];"
                       : "=r"(v[j][q].x), "=r"(v[j][q].y), "=r"(v[j][q].z), "=r"(v[j][q].w) : "l"(src + q));
        } else if (MODE == 4) {
          asm volatile("ld.global.cs.v4.u32 {%0,%1,%2,%3}, [%4];"
                       : "=r"(v[j][q].x), "=r"(v[j][q].y), "=r"(v[j][q].z), "=r"(v[j][q].w) : "l"(src + q));
        } else {
          const unsigned dst = static_cast<unsigned>(__cvta_generic_to_shared(
              smem + (static_cast<size_t>(j) * Q + q) * blockDim.x + threadIdx.x));
          asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src + q));
        }
      }
    }
    if (MODE == 3) {
      asm volatile("cp.async.wait_all;" ::: "memory");
#pragma unroll
      for (int j = 0; j < P; ++j) {
#pragma unroll
        for (int q = 0; q < Q; ++q) v[j][q] = smem[(static_cast<size_t>(j) * Q + q) * blockDim.x + threadIdx.x];
      }
    }
#pragma unroll
    for (int j = 0; j < P; ++j) {
      unsigned acc = 0;
#pragma unroll
      for (int q = 0; q < Q; ++q) acc += v[j][q].x ^ v[j][q].y ^ v[j][q].z ^ v[j][q].w;
      const long long k = base + static_cast<long long>(j) * blockDim.x;
      if (k < n) __stcs(out + k, __uint_as_float(acc));
    }
  }
}

// Lane pairs read 64-byte cells together: in one instruction lanes 2k and 2k+1
// fetch the two 32-byte halves of the SAME cell (first the even lane's cell,
// then the odd lane's), and swap halves with shuffles.  16 distinct lines per
// warp instruction instead of 32.
template <int P>
__global__ void __launch_bounds__(256) probe_pair64(const uint4* __restrict__ table, unsigned long long cells,
                                                   long long n, float* __restrict__ out) {
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x * P;
  const int odd = threadIdx.x & 1;
  for (long long base = (static_cast<long long>(blockIdx.x) * blockDim.x) * P + threadIdx.x; base < n;
       base += stride) {
    unsigned long long a[P][4], b[P][4];
#pragma unroll
    for (int j = 0; j < P; ++j) {
      const unsigned long long k = base + static_cast<long long>(j) * blockDim.x;
      const unsigned long long mine = mix(k) % cells;
      const unsigned long long other = __shfl_xor_sync(0xffffffffu, mine, 1);
      const unsigned long long c_even = odd ? other : mine;
      const unsigned long long c_odd = odd ? mine : other;
      const char* pe = reinterpret_cast<const char*>(table) + c_even * 64 + odd * 32;
      const char* po = reinterpret_cast<const char*>(table) + c_odd * 64 + odd * 32;
      asm volatile("ld.global.nc.L1::no_allocate.L2::64B.v4.u64 {%0,%1,%2,%3}, [%4];"
                   : "=l"(a[j][0]), "=l"(a[j][1]), "=l"(a[j][2]), "=l"(a[j][3]) : "l"(pe));
      asm volatile("ld.global.nc.L1::no_allocate.L2::64B.v4.u64 {%0,%1,%2,%3}, [%4];"
                   : "=l"(b[j][0]), "=l"(b[j][1]), "=l"(b[j][2]), "=l"(b[j][3]) : "l"(po));
    }
#pragma unroll
    for (int j = 0; j < P; ++j) {
      // even lane keeps a (low half of its cell) and needs the odd lane's a (high half);
      // odd lane keeps b (high half of its cell) and needs the even lane's b (low half)
      unsigned long long acc = 0;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const unsigned long long send = odd ? a[j][q] : b[j][q];
        const unsigned long long got = __shfl_xor_sync(0xffffffffu, send, 1);
        acc ^= (odd ? b[j][q] : a[j][q]) ^ got;
      }
      const long long k = base + static_cast<long long>(j) * blockDim.x;
      if (k < n) __stcs(out + k, __uint_as_float(static_cast<unsigned>(acc ^ (acc >> 32))));
    }
  }
}

template <int P>
static void run_pair(const uint4* table, unsigned long long cells, long long n, float* out) {
  int sms = 148, occ = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, probe_pair64<P>, 256, 0);
  cudaEvent_t a, b;
  cudaEventCreate(&a);
  cudaEventCreate(&b);
  float best = 1e30f;
  const long long n256 = n / 256 * 256;
  for (int it = 0; it < 4; ++it) {
    cudaEventRecord(a);
    probe_pair64<P><<<sms * occ, 256>>>(table, cells, n256, out);
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    if (it > 0 && ms < best) best = ms;
  }
  std::printf("%-28s P=%d B= 64 blocks/SM=%d in-flight/SM=%5d  %7.2f Ggather/s  %7.1f GB/s useful\n",
              "lane-pair v4.u64 L2::64B", P, occ, occ * 256 * P, n256 / best / 1e6, n256 / best / 1e6 * 64);
}

template <int MODE, int P, int BYTES>
static void run(const char* name, const uint4* table, unsigned long long cells, long long n, float* out,
                int blocks_per_sm) {
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  const size_t shm = MODE == 3 ? static_cast<size_t>(P) * (BYTES / 16) * 256 * 16 : 0;
  if (shm > 48 * 1024) cudaFuncSetAttribute(probe<MODE, P, BYTES>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)shm);
  int occ = 0;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, probe<MODE, P, BYTES>, 256, shm);
  const int bps = occ < blocks_per_sm ? occ : blocks_per_sm;
  if (bps <= 0) { std::printf("%-28s P=%d B=%d  cannot launch\n", name, P, BYTES); return; }
  cudaEvent_t a, b;
  cudaEventCreate(&a);
  cudaEventCreate(&b);
  float best = 1e30f;
  for (int it = 0; it < 4; ++it) {
    cudaEventRecord(a);
    probe<MODE, P, BYTES><<<sms * bps, 256, shm>>>(table, cells, n, out);
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    if (it > 0 && ms < best) best = ms;
  }
  const cudaError_t e = cudaGetLastError();
  std::printf("%-28s P=%d B=%3d blocks/SM=%d in-flight/SM=%5d  %7.2f Ggather/s  %7.1f GB/s useful%s\n", name, P,
              BYTES, bps, bps * 256 * P, n / best / 1e6, n / best / 1e6 * BYTES,
              e == cudaSuccess ? "" : cudaGetErrorString(e));
}

int main(int argc, char** argv) {
  const double gib = argc > 1 ? std::atof(argv[1]) : 12.0;
  const long long n = argc > 2 ? static_cast<long long>(std::atof(argv[2])) : 200000000LL;
  const size_t bytes = static_cast<size_t>(gib * (1ULL << 30)) & ~static_cast<size_t>(127);
  uint4* table = nullptr;
  float* out = nullptr;
  if (cudaMalloc(&table, bytes) != cudaSuccess || cudaMalloc(&out, n * 4) != cudaSuccess) {
    std::printf("allocation failed\n");
    return 1;
  }
  cudaMemset(table, 1, bytes);
  std::printf("table %.1f GiB, %lld gathers per launch\n", gib, n);
  run_pair<1>(table, bytes / 64, n, out);
  run_pair<2>(table, bytes / 64, n, out);
  run_pair<4>(table, bytes / 64, n, out);
  if (argc > 3) return 0;
#define SWEEP(MODE, NAME, B)                                     \
  run<MODE, 1, B>(NAME, table, bytes / B, n, out, 8);            \
  run<MODE, 2, B>(NAME, table, bytes / B, n, out, 8);            \
  run<MODE, 4, B>(NAME, table, bytes / B, n, out, 8);            \
  run<MODE, 8, B>(NAME, table, bytes / B, n, out, 8);
  SWEEP(0, "ld.nc.L1::no_allocate", 32)
  SWEEP(1, "ld.cg", 32)
  SWEEP(2, "ld.nc.L2::64B", 32)
  SWEEP(4, "ld.cs", 32)
  SWEEP(3, "cp.async.cg->smem", 32)
  run<3, 16, 32>("cp.async.cg->smem", table, bytes / 32, n, out, 8);
  SWEEP(5, "v4.u64 nc", 32)
  SWEEP(6, "v4.u64 nc L2::64B", 32)
  SWEEP(7, "v4.u64 cs", 32)
  SWEEP(8, "v4.u64 nc evict_first 64B", 32)
  SWEEP(9, "v4.u64 cs L2::64B", 32)
  SWEEP(6, "v4.u64 nc L2::64B", 64)
  SWEEP(9, "v4.u64 cs L2::64B", 64)
  SWEEP(0, "ld.nc.L1::no_allocate", 64)
  SWEEP(3, "cp.async.cg->smem", 64)
  SWEEP(0, "ld.nc.L1::no_allocate", 128)
  cudaFree(table);
  cudaFree(out);
  return 0;
}
