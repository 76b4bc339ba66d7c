// pcie_probe.cu -- ceiling of the host-fed (end-to-end) path: copy-only probe of
// the traffic one bivariate call moves per point (16 B host->device, 8 B
// device->host), no kernel, on 1 / 2 / 4 / 8 GPUs concurrently from ONE process
// (one host thread and three streams per GPU, 8M-point chunks: the shape of
// pyi_host.cu's pipeline).  Variants of the page-locked pool: default
// cudaHostAlloc, write-combined inputs, and NUMA-local allocation (the
// allocating thread is pinned to the CPUs of the GPU's NUMA node first).
//   nvcc -O2 -std=c++17 -o tools/pcie_probe tools/pcie_probe.cu -lpthread
#include <cuda_runtime.h>
#include <sched.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <string>
#include <thread>
#include <vector>

#define CK(x)                                                                         \
  do {                                                                                \
    cudaError_t e__ = (x);                                                            \
    if (e__ != cudaSuccess) {                                                         \
      std::fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e__));                  \
      std::exit(1);                                                                   \
    }                                                                                 \
  } while (0)

static int gpu_numa_node(int dev) {
  char bus[64];
  if (cudaDeviceGetPCIBusId(bus, sizeof(bus), dev) != cudaSuccess) return -1;
  for (char* c = bus; *c; ++c) *c = static_cast<char>(std::tolower(*c));
  std::ifstream f(std::string("/sys/bus/pci/devices/") + bus + "/numa_node");
  int node = -1;
  if (f) f >> node;
  return node;
}

static bool pin_to_node(int node) {
  if (node < 0) return false;
  std::ifstream f("/sys/devices/system/node/node" + std::to_string(node) + "/cpulist");
  std::string list;
  if (!f || !(f >> list)) return false;
  cpu_set_t set;
  CPU_ZERO(&set);
  size_t pos = 0;
  while (pos < list.size()) {
    size_t end = list.find(',', pos);
    if (end == std::string::npos) end = list.size();
    const std::string part = list.substr(pos, end - pos);
    const size_t dash = part.find('-');
    const int a = std::atoi(part.substr(0, dash).c_str());
    const int b = dash == std::string::npos ? a : std::atoi(part.substr(dash + 1).c_str());
    for (int c = a; c <= b && c < CPU_SETSIZE; ++c) CPU_SET(c, &set);
    pos = end + 1;
  }
  return sched_setaffinity(0, sizeof(set), &set) == 0;
}

struct PerGpu {
  int dev;
  double *hx, *hy, *ho;
  char* d[3];
  cudaStream_t s[3];
};

int main(int argc, char** argv) {
  const long long n = argc > 1 ? static_cast<long long>(std::atof(argv[1])) : 50000000LL;  // points per GPU
  const long long chunk = 1LL << 23;
  int ngpu = 0;
  CK(cudaGetDeviceCount(&ngpu));
  std::printf("# pcie_probe: %lld points per GPU (16 B in + 8 B out each), %d GPU(s) visible, %u host threads\n",
              n, ngpu, std::thread::hardware_concurrency());
  for (int d = 0; d < ngpu; ++d) std::printf("# gpu %d: numa node %d\n", d, gpu_numa_node(d));
  const char* variants[] = {"default", "write-combined-inputs", "numa-local"};
  const char* modes[] = {"h2d+d2h", "h2d-only", "d2h-only"};
  for (int variant = 0; variant < 3; ++variant) {
    for (int g = 1; g <= ngpu; g *= 2) {
      std::vector<PerGpu> gpus(g);
      std::vector<std::thread> th;
      for (int k = 0; k < g; ++k) {
        th.emplace_back([&, k] {
          PerGpu& p = gpus[k];
          p.dev = k;
          CK(cudaSetDevice(k));
          if (variant == 2) pin_to_node(gpu_numa_node(k));
          const unsigned in_flags = variant == 1 ? cudaHostAllocWriteCombined : cudaHostAllocDefault;
          CK(cudaHostAlloc(reinterpret_cast<void**>(&p.hx), n * 8, in_flags));
          CK(cudaHostAlloc(reinterpret_cast<void**>(&p.hy), n * 8, in_flags));
          CK(cudaHostAlloc(reinterpret_cast<void**>(&p.ho), n * 8, cudaHostAllocDefault));
          std::memset(p.hx, 1, n * 8);
          std::memset(p.hy, 2, n * 8);
          std::memset(p.ho, 0, n * 8);
          for (int s = 0; s < 3; ++s) {
            CK(cudaStreamCreateWithFlags(&p.s[s], cudaStreamNonBlocking));
            CK(cudaMalloc(reinterpret_cast<void**>(&p.d[s]), chunk * 24));
          }
        });
      }
      for (auto& t : th) t.join();
      for (int mode = 0; mode < 3; ++mode) {
        double best = 1e30;
        for (int rep = 0; rep < 4; ++rep) {
          const auto t0 = std::chrono::steady_clock::now();
          std::vector<std::thread> run;
          for (int k = 0; k < g; ++k) {
            run.emplace_back([&, k] {
              PerGpu& p = gpus[k];
              CK(cudaSetDevice(p.dev));
              int s = 0;
              for (long long b = 0; b < n; b += chunk, s = (s + 1) % 3) {
                const long long c = std::min(chunk, n - b);
                if (mode != 2) {
                  CK(cudaMemcpyAsync(p.d[s], p.hx + b, c * 8, cudaMemcpyHostToDevice, p.s[s]));
                  CK(cudaMemcpyAsync(p.d[s] + chunk * 8, p.hy + b, c * 8, cudaMemcpyHostToDevice, p.s[s]));
                }
                if (mode != 1) {
                  CK(cudaMemcpyAsync(p.ho + b, p.d[s] + chunk * 16, c * 8, cudaMemcpyDeviceToHost, p.s[s]));
                }
              }
              for (int q = 0; q < 3; ++q) CK(cudaStreamSynchronize(p.s[q]));
            });
          }
          for (auto& t : run) t.join();
          const double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
          if (rep > 0 && dt < best) best = dt;
        }
        const double bytes = static_cast<double>(n) * g * (mode == 0 ? 24 : (mode == 1 ? 16 : 8));
        std::printf("%-22s gpus=%d %-9s %8.2f ms  %7.1f GB/s aggregate  %7.3f Gpts/s ceiling\n", variants[variant],
                    g, modes[mode], best * 1e3, bytes / best / 1e9,
                    mode == 0 ? static_cast<double>(n) * g / best / 1e9 : 0.0);
        std::fflush(stdout);
      }
      for (auto& p : gpus) {
        CK(cudaSetDevice(p.dev));
        for (int s = 0; s < 3; ++s) {
          cudaFree(p.d[s]);
          cudaStreamDestroy(p.s[s]);
        }
        cudaFreeHost(p.hx);
        cudaFreeHost(p.hy);
        cudaFreeHost(p.ho);
      }
    }
  }
  return 0;
}
