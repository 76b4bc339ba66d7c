// pyi_host.cu -- the C-ABI of include/pyinterp_b200.h: axis / grid handles,
// device residency, the host-buffer pipeline and the bounds-error protocol.
// No compute happens on the host: every entry point that produces indexes or
// values launches the sm_100a kernels of pyi_kernels.cu, and returns
// PYI_NO_DEVICE when no CUDA device is usable.
#include <cuda_runtime.h>
#include <emmintrin.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <charconv>
#include <cmath>
#include <condition_variable>
#include <cstdio>
#include <cstring>
#include <deque>
#include <functional>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/pyinterp_b200.h"
#include "pyi_axis_host.hpp"
#include "pyi_kernels.cuh"

namespace {

thread_local std::string g_error;

int fail(int status, const std::string& what) {
  g_error = what;
  return status;
}
int cuda_fail(cudaError_t e, const char* where) {
  g_error = std::string(where) + ": " + cudaGetErrorString(e);
  return e == cudaErrorMemoryAllocation ? PYI_NOMEM : PYI_CUDA_ERROR;
}
#define PYI_CUDA(call)                                 \
  do {                                                 \
    cudaError_t e__ = (call);                          \
    if (e__ != cudaSuccess) return cuda_fail(e__, #call); \
  } while (0)

int device_count_cached() {
  static int n = -1;
  if (n < 0) {
    int c = 0;
    if (cudaGetDeviceCount(&c) != cudaSuccess) {
      cudaGetLastError();
      c = 0;
    }
    n = c;
  }
  return n;
}

int require_device() {
  if (device_count_cached() <= 0) {
    return fail(PYI_NO_DEVICE,
                "no CUDA device: pyinterp_b200 runs on the GPU only (no CPU fallback)");
  }
  return PYI_OK;
}

size_t dtype_size(int dtype) {
  switch (dtype) {
    case PYI_I8: case PYI_U8: return 1;
    case PYI_I16: case PYI_U16: return 2;
    case PYI_I32: case PYI_U32: case PYI_F32: return 4;
    default: return 8;
  }
}
bool result_is_f64(int dtype) {  // pybind/grid_dispatch.hpp:25-30
  return dtype == PYI_F64 || dtype == PYI_I64 || dtype == PYI_U64;
}

// normalize_period(x, min, period) exactly as math.hpp:104-123 (this file is
// compiled with -ffp-contract=off: no fused operations on the host).
double host_normalize(double x, double min, double period) {
  if (period != 0.0 && (x >= min + period || x < min)) {
    double r = std::remainder(x - min, period);
    if (r != 0.0 && (std::signbit(r) != std::signbit(period))) r += period;
    return r + min;
  }
  return x;
}

// a - b exactly representable? (TwoSum residual of the subtraction)
bool exact_difference(double a, double b, double& diff) {
  const volatile double s = a - b;
  const volatile double bb = s - a;
  const volatile double err = (a - (s - bb)) + (-b - bb);
  diff = s;
  return err == 0.0;
}

// Constants of the tolerance-mode bilinear kernel (pyi_fast2d_tol.cuh) for a
// regular float64 axis that already passed the fast-path checks (ascending, or
// descending without period: step < 0, every expression below holds as it is).
void tol_constants(const pyi_axis_desc& d, pyi::DAxis& out) {
  out.tol_ok = 0;
  const double step = d.step.f, start = d.start.f;
  const double astep = std::fabs(step);  // negative on a descending axis (no period then)
  const double maxabs = std::max({std::fabs(d.min_value.f), std::fabs(d.max_value.f),
                                  std::fabs(d.min_value.f + d.period.f)});
  // |1 - dx / step| < 2^-26 for every spacing dx (the kernel's first-order
  // reciprocal), and cell numbers far below 2^31
  if (!(maxabs / astep < 67108864.0) || d.size >= (1LL << 30)) return;
  if (d.period.f != 0.0 && !(d.period.f >= 9.765625e-4)) return;  // |x| < 2^40 => |x / period| < 2^50
  const double last = start + static_cast<double>(d.size - 1) * step;
  const bool wraps = d.is_periodic != 0;
  if (wraps) {
    const volatile double e = 1.0 - (out.seam_c1 - last) * d.inv_step;
    if (!(std::fabs(e) < 1.4901161193847656e-08)) return;
  }
  {
    int ex = 0;
    std::frexp(astep, &ex);                   // |step| = m * 2^ex, m in [0.5, 1)
    const int bound = std::min(40, ex - 1 + 28);  // |x| < 2^bound <= 2^28 * step
    out.tol_xhi = (1023 + bound) << 20;
  }
  out.tol_off = -(start * d.inv_step) - 0.5;
  out.tol_poff = d.period.f != 0.0 ? -(d.min_value.f * (1.0 / d.period.f)) - 0.5 : 0.0;
  out.tol_ok = 1;
  // exact nodes: fma(k, step, start) is the reference's node (two roundings)
  // for every k, and every spacing is exactly `step`
  bool exact = true;
  double prev = start;
  for (int64_t k = 0; exact && k < d.size; ++k) {
    const volatile double prod = static_cast<double>(k) * step;
    const volatile double node = start + prod;
    exact = std::fma(static_cast<double>(k), step, start) == node;
    double diff;
    if (exact && k > 0) exact = exact_difference(node, prev, diff) && diff == step;
    prev = node;
  }
  if (exact && wraps) {
    double diff;
    exact = exact_difference(out.seam_c1, last, diff) && diff == step;
  }
  if (exact) out.tol_ok = 2;
}

// Per-axis constants of the bilinear fast path (pyi_fast2d.cuh).
void fast_constants(const pyi_axis* a, pyi::DAxis& out) {
  const pyi_axis_desc& d = a->d;
  out.hi = out.seam_c1 = out.seam_c1_f32 = 0.0;
  out.fast_ok = 0;
  out.tol_ok = 0;
  out.tol_xhi = 0;
  out.tol_off = out.tol_poff = 0.0;
  out.ispan = 0;
  out.desc_ok = 0;
  if (d.coord_type == PYI_COORD_I64) {
    // regular ascending int64 (temporal) axis without period: pyi_fastnd.cuh
    const long long lim = 1LL << 61;
    if (d.kind != PYI_AXIS_REGULAR || !d.is_ascending || d.size < 2 || d.period.i != 0) return;
    if (d.step.i <= 0 || d.start.i <= -lim || d.start.i >= lim) return;
    if (d.step.i >= lim / (d.size - 1)) return;
    out.ispan = (d.size - 1) * d.step.i;
    out.inv_step = 1.0 / static_cast<double>(d.step.i);
    out.fast_ok = 1;
    return;
  }
  if (d.coord_type != PYI_COORD_F64 || d.size < 2) return;
  const double lim = 1099511627776.0;  // 2^40, the kernel's sane() bound
  auto moderate = [&](double v) { return std::fabs(v) < lim; };
  auto spacing_ok = [](double h) { return h > 1e-30 && h < 1e30; };
  if (!d.is_ascending) {
    // descending regular axis without period: the same checks mirrored (min is
    // the last node, max the first); the kernels that test desc_ok serve it
    if (d.kind != PYI_AXIS_REGULAR || d.is_periodic || d.period.f != 0.0) return;
    const volatile double last = d.start.f + static_cast<double>(d.size - 1) * d.step.f;
    const volatile double q_last = (last - d.start.f) * d.inv_step;
    const bool ok = moderate(d.min_value.f) && moderate(d.max_value.f) && spacing_ok(-d.step.f) &&
                    d.max_value.f == d.start.f && d.min_value.f == last &&
                    std::round(q_last) == static_cast<double>(d.size - 1);
    if (!ok) return;
    out.hi = d.min_value.f;
    out.desc_ok = 1;
    tol_constants(d, out);
    return;
  }
  bool ok = moderate(d.min_value.f) && moderate(d.max_value.f) && moderate(d.period.f);
  if (d.kind == PYI_AXIS_REGULAR) {
    ok = ok && spacing_ok(d.step.f);
    const volatile double last = d.start.f + static_cast<double>(d.size - 1) * d.step.f;
    // the kernels resolve "beyond the ends" without the generic code: that
    // needs min / max to be the end nodes and the nearest-node rounding of the
    // last node to be exact (container.hpp:553-563 is monotone in c)
    const volatile double q_last = (last - d.start.f) * d.inv_step;
    ok = ok && d.min_value.f == d.start.f && d.max_value.f == last &&
         std::round(q_last) == static_cast<double>(d.size - 1);
    if (d.is_periodic && d.period.f > 0.0) {
      out.seam_c1 = host_normalize(d.start.f, last, d.period.f);
      const double last32 = static_cast<double>(static_cast<float>(last));
      out.seam_c1_f32 = host_normalize(d.start.f, last32, d.period.f);
      ok = ok && out.seam_c1 > last && out.seam_c1_f32 > last32;
    }
  } else {
    for (size_t k = 0; ok && k + 1 < a->pf.size(); ++k) ok = spacing_ok(a->pf[k + 1] - a->pf[k]);
    ok = ok && d.min_value.f == a->pf.front() && d.max_value.f == a->pf.back();
  }
  out.hi = d.min_value.f + d.period.f;
  out.fast_ok = ok ? 1 : 0;
  if (ok && d.kind == PYI_AXIS_REGULAR) tol_constants(d, out);
}

// Device mirror of an axis on the current device (uploaded once per device).
int device_axis(pyi_axis* a, pyi::DAxis& out) {
  const pyi_axis_desc& d = a->d;
  if (d.size >= 2147483646LL) {
    return fail(PYI_BAD_ARG, "axis too large for the device kernels");
  }
  std::memset(&out, 0, sizeof(out));
  out.kind = d.kind;
  out.coord_i64 = d.coord_type == PYI_COORD_I64;
  out.asc = d.is_ascending;
  out.periodic = d.is_periodic;
  out.n = static_cast<int32_t>(d.size);
  if (d.coord_type == PYI_COORD_F64) {
    out.start = d.start.f;
    out.step = d.step.f;
    out.inv_step = d.inv_step;
    out.period = d.period.f;
    out.inv_period = d.period.f != 0.0 ? 1.0 / d.period.f : 0.0;
    out.min = d.min_value.f;
    out.max = d.max_value.f;
  } else {
    out.istart = d.start.i;
    out.istep = d.step.i;
    out.istep2 = d.step_2;
    out.iperiod = d.period.i;
    out.imin = d.min_value.i;
    out.imax = d.max_value.i;
  }
  fast_constants(a, out);
  if (d.kind == PYI_AXIS_IRREGULAR) {
    int dev = 0;
    PYI_CUDA(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lock(a->mu);
    auto it = a->dev.find(dev);
    if (it == a->dev.end()) {
      pyi_axis::Dev m;
      const size_t bytes = static_cast<size_t>(d.size) * 8;
      PYI_CUDA(cudaMalloc(&m.points, bytes));
      PYI_CUDA(cudaMemcpy(m.points, d.points, bytes, cudaMemcpyHostToDevice));
      if (a->lut_bins > 0) {
        const size_t lb = a->lut.size() * sizeof(int32_t);
        PYI_CUDA(cudaMalloc(reinterpret_cast<void**>(&m.lut), lb));
        PYI_CUDA(cudaMemcpy(m.lut, a->lut.data(), lb, cudaMemcpyHostToDevice));
      }
      it = a->dev.emplace(dev, m).first;
    }
    out.points = it->second.points;
    out.lut = it->second.lut;
    out.lut_bins = it->second.lut ? a->lut_bins : 0;
    out.lut_min = a->lut_min;
    out.lut_inv = a->lut_inv;
  }
  return PYI_OK;
}

void free_axis_device(pyi_axis* a) {
  for (auto& kv : a->dev) {
    int cur = 0;
    cudaGetDevice(&cur);
    cudaSetDevice(kv.first);
    cudaFree(kv.second.points);
    cudaFree(kv.second.lut);
    cudaSetDevice(cur);
  }
  a->dev.clear();
}

bool host_pinned(const void* p) {
  cudaPointerAttributes attr{};
  if (cudaPointerGetAttributes(&attr, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return attr.type == cudaMemoryTypeHost;
}

// Staging threads of the host pipeline: a persistent pool shared by every call
// of the process (round 2: creating and joining 16 threads per 64 MB copy cost a
// quarter of a pageable call; eight device shards each spawning their own
// oversubscribed the host).  The caller of run() works on its own batch too, so
// a batch completes even when every worker is busy with other callers.  The
// pool is leaked on purpose (no destructor races at exit) and re-created in a
// forked child, whose copy of it has no threads.
class CopyPool {
 public:
  static CopyPool& get() {
    static std::mutex mu;
    static CopyPool* pool = nullptr;
    static long pid = -1;
    std::lock_guard<std::mutex> lock(mu);
    const long me = static_cast<long>(getpid());
    if (pool == nullptr || pid != me) {
      pool = new CopyPool();
      pid = me;
    }
    return *pool;
  }
  unsigned threads() const { return workers_ + 1; }

  // fn(part) for every part in [0, parts), on the pool's threads and the caller
  template <typename F>
  void run(unsigned parts, F&& fn) {
    if (parts <= 1 || workers_ == 0) {
      for (unsigned k = 0; k < parts; ++k) fn(k);
      return;
    }
    auto batch = std::make_shared<Batch>();
    batch->parts = parts;
    batch->fn = [&fn](unsigned k) { fn(k); };
    {
      std::lock_guard<std::mutex> lock(mu_);
      queue_.push_back(batch);
    }
    cv_.notify_all();
    work(*batch);
    {
      std::lock_guard<std::mutex> lock(mu_);
      auto it = std::find(queue_.begin(), queue_.end(), batch);
      if (it != queue_.end()) queue_.erase(it);
    }
    std::unique_lock<std::mutex> lock(batch->mu);
    batch->cv.wait(lock, [&] { return batch->done.load() == parts; });
  }

 private:
  struct Batch {
    std::function<void(unsigned)> fn;
    unsigned parts{0};
    std::atomic<unsigned> next{0};
    std::atomic<unsigned> done{0};
    std::mutex mu;
    std::condition_variable cv;
  };
  static void work(Batch& b) {
    for (;;) {
      const unsigned k = b.next.fetch_add(1);
      if (k >= b.parts) return;
      b.fn(k);
      if (b.done.fetch_add(1) + 1 == b.parts) {
        std::lock_guard<std::mutex> lock(b.mu);
        b.cv.notify_all();
      }
    }
  }
  CopyPool() {
    // PYINTERP_B200_COPY_THREADS: staging threads of the process (default: the host's
    // hardware threads, at most 32)
    const char* e = getenv("PYINTERP_B200_COPY_THREADS");
    const int want = e ? atoi(e) : 0;
    const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    const unsigned n = want >= 1 && want <= 256 ? static_cast<unsigned>(want) : std::min(hw, 32u);
    workers_ = n - 1;
    for (unsigned t = 0; t < workers_; ++t) {
      std::thread([this] {
        for (;;) {
          std::shared_ptr<Batch> b;
          {
            std::unique_lock<std::mutex> lock(mu_);
            cv_.wait(lock, [&] { return !queue_.empty(); });
            b = queue_.front();
            if (b->next.load() >= b->parts) {  // nothing left to claim: retire it
              queue_.pop_front();
              continue;
            }
          }
          work(*b);
        }
      }).detach();
    }
  }
  std::mutex mu_;
  std::condition_variable cv_;
  std::deque<std::shared_ptr<Batch>> queue_;
  unsigned workers_{0};
};

// One part of a staging copy with streaming (non-temporal) stores: the
// destination is not read into the cache first, which saves the write-allocate
// third of the copy's DRAM traffic -- the staged gigabytes are never touched by
// the CPU again (the DMA engine reads them).  PYINTERP_B200_COPY=memcpy keeps
// plain memcpy.
void stream_copy(char* dst, const char* src, size_t len) {
  static const bool plain = [] {
    const char* e = getenv("PYINTERP_B200_COPY");
    return e != nullptr && e[0] == 'm';
  }();
  if (plain || len < 4096) {
    std::memcpy(dst, src, len);
    return;
  }
  const size_t head = static_cast<size_t>(-reinterpret_cast<uintptr_t>(dst)) & 63;
  std::memcpy(dst, src, head);
  dst += head;
  src += head;
  len -= head;
  const size_t blocks = len / 64;
  for (size_t b = 0; b < blocks; ++b) {
    const __m128i v0 = _mm_loadu_si128(reinterpret_cast<const __m128i*>(src));
    const __m128i v1 = _mm_loadu_si128(reinterpret_cast<const __m128i*>(src + 16));
    const __m128i v2 = _mm_loadu_si128(reinterpret_cast<const __m128i*>(src + 32));
    const __m128i v3 = _mm_loadu_si128(reinterpret_cast<const __m128i*>(src + 48));
    _mm_stream_si128(reinterpret_cast<__m128i*>(dst), v0);
    _mm_stream_si128(reinterpret_cast<__m128i*>(dst + 16), v1);
    _mm_stream_si128(reinterpret_cast<__m128i*>(dst + 32), v2);
    _mm_stream_si128(reinterpret_cast<__m128i*>(dst + 48), v3);
    src += 64;
    dst += 64;
  }
  _mm_sfence();
  std::memcpy(dst, src, len - blocks * 64);
}

void parallel_memcpy(void* dst, const void* src, size_t bytes) {
  const size_t kMin = 1u << 20;  // bytes per part, at least
  if (bytes < 2 * kMin) {
    stream_copy(static_cast<char*>(dst), static_cast<const char*>(src), bytes);
    return;
  }
  CopyPool& pool = CopyPool::get();
  // a few parts per thread: the threads are shared with other calls and finish unevenly
  const unsigned parts = static_cast<unsigned>(std::min<size_t>(bytes / kMin, 4u * pool.threads()));
  const size_t part = (bytes / parts + 63) & ~static_cast<size_t>(63);
  pool.run(parts, [=](unsigned k) {
    const size_t off = std::min(bytes, k * part);
    const size_t len = std::min(part, bytes - off);
    if (len) stream_copy(static_cast<char*>(dst) + off, static_cast<const char*>(src) + off, len);
  });
}

// ---------------------------------------------------------------------------
// Host-buffer pipeline: the point arrays are cut into chunks that flow
// H2D -> kernel -> D2H through kSlots stream slots, so the two copy engines
// and the SMs overlap.  Caller buffers that are already page-locked are copied
// directly; pageable ones are staged through page-locked buffers.
//
// Re-entrancy (the reference releases the GIL around the drivers and supports
// free-threaded CPython: concurrent calls from several threads are legal): a
// call checks a Workspace out of a per-device pool for its duration -- streams,
// device chunk buffers, staging and the bounds-error key all belong to the
// call.  Nothing is released when the calling thread's device changes.
// ---------------------------------------------------------------------------
constexpr int kSlots = 3;
constexpr long long kChunk = 1LL << 23;     // points per chunk, upper bound
constexpr long long kChunkMin = 1LL << 20;  // lower bound (calls above it are cut in >= 8 chunks)
constexpr int kMaxWorkspacesPerDevice = 4;

struct Slot {
  cudaStream_t stream{nullptr};
  cudaEvent_t done{nullptr};
  char* dev{nullptr};     // device chunk buffers
  char* pinned{nullptr};  // staging (allocated on first pageable use)
  unsigned* queue{nullptr};  // the tolerance kernel's queue of unsettled points
  bool busy{false};
  long long begin{0}, count{0};
};

struct Workspace {
  int device{-1};
  bool in_use{false};
  Slot slot[kSlots];
  cudaEvent_t ready{nullptr};
  size_t dev_bytes{0};
  size_t pinned_bytes{0};
  unsigned long long* oob_key{nullptr};   // device
  unsigned long long* oob_host{nullptr};  // pinned
  // the calling thread's current device must be `device`
  int ensure(size_t need_dev, size_t need_pinned) {
    if (!oob_key) {
      PYI_CUDA(cudaMalloc(reinterpret_cast<void**>(&oob_key), sizeof(unsigned long long)));
      PYI_CUDA(cudaMallocHost(reinterpret_cast<void**>(&oob_host), sizeof(unsigned long long)));
    }
    if (!ready) PYI_CUDA(cudaEventCreateWithFlags(&ready, cudaEventDisableTiming));
    for (auto& s : slot) {
      if (!s.stream) PYI_CUDA(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
      if (!s.done) PYI_CUDA(cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming));
      if (!s.queue) {
        PYI_CUDA(cudaMalloc(reinterpret_cast<void**>(&s.queue), (pyi::kTolQueue + 1) * sizeof(unsigned)));
      }
    }
    if (need_dev > dev_bytes) {
      for (auto& s : slot) {
        cudaFree(s.dev);
        s.dev = nullptr;
      }
      dev_bytes = 0;
      for (auto& s : slot) PYI_CUDA(cudaMalloc(reinterpret_cast<void**>(&s.dev), need_dev));
      dev_bytes = need_dev;
    }
    if (need_pinned > pinned_bytes) {
      for (auto& s : slot) {
        cudaFreeHost(s.pinned);
        s.pinned = nullptr;
      }
      pinned_bytes = 0;
      for (auto& s : slot) PYI_CUDA(cudaMallocHost(reinterpret_cast<void**>(&s.pinned), need_pinned));
      pinned_bytes = need_pinned;
    }
    return PYI_OK;
  }
  // Every exit path of a call ends here: whatever is still in flight on the
  // slot streams is waited for and the slots are marked idle, so a failed call
  // cannot leave stale begin / count behind for the next one.
  void quiesce() {
    for (auto& s : slot) {
      if (s.busy && s.stream) cudaStreamSynchronize(s.stream);
      s.busy = false;
      s.begin = s.count = 0;
    }
    cudaGetLastError();
  }
};

class WorkspacePool {
 public:
  Workspace* acquire(int device) {
    std::unique_lock<std::mutex> lock(mu_);
    for (;;) {
      int have = 0;
      for (auto& w : all_) {
        if (w->device != device) continue;
        ++have;
        if (!w->in_use) {
          w->in_use = true;
          return w.get();
        }
      }
      if (have < kMaxWorkspacesPerDevice) {
        all_.push_back(std::make_unique<Workspace>());
        Workspace* w = all_.back().get();
        w->device = device;
        w->in_use = true;
        return w;
      }
      cv_.wait(lock);
    }
  }
  void release(Workspace* w) {
    {
      std::lock_guard<std::mutex> lock(mu_);
      w->in_use = false;
    }
    cv_.notify_all();
  }

 private:
  std::mutex mu_;
  std::condition_variable cv_;
  std::vector<std::unique_ptr<Workspace>> all_;
};

// never destroyed: no CUDA calls during static destruction
WorkspacePool& pool() {
  static WorkspacePool* p = new WorkspacePool();
  return *p;
}

// A workspace of `device` for the duration of one call (the calling thread's
// current device must be `device`).
struct Lease {
  Workspace* ws;
  explicit Lease(int device) : ws(pool().acquire(device)) {}
  Lease(const Lease&) = delete;
  Lease& operator=(const Lease&) = delete;
  ~Lease() {
    ws->quiesce();
    pool().release(ws);
  }
};

// Binds the calling thread to `device` and restores the previous one.
struct DeviceGuard {
  int prev{-1};
  bool switched{false};
  cudaError_t err{cudaSuccess};
  explicit DeviceGuard(int device) {
    err = cudaGetDevice(&prev);
    if (err == cudaSuccess && prev != device) {
      err = cudaSetDevice(device);
      switched = err == cudaSuccess;
    }
  }
  ~DeviceGuard() {
    if (switched) cudaSetDevice(prev);
  }
};

// One input or output stream of a pipelined call.
struct Column {
  const void* host_in{nullptr};  // input columns
  void* host_out{nullptr};       // output column
  size_t elem{8};
  bool pinned{false};
  size_t dev_off{0};  // offset inside the slot buffers
};

long long chunk_points(long long n) {
  if (n <= kChunkMin) return std::max<long long>(n, 1);
  // 1/8 of the call (measured, round 2: 1/32 chunks lost 20 % of a copy-bound
  // call to per-chunk event waits on three slots)
  long long c = std::min(kChunk, std::max(kChunkMin, (n + 7) / 8));
  return (c + 63) & ~63LL;  // keeps every chunk start 512-byte aligned in the streams
}

// Runs `launch(slot_stream, begin, count, dev_buffer, slot_queue)` over all
// chunks of the n points the columns describe.  With `reset_key` the workspace's bounds-error
// key is reset on the first slot's stream and every other slot waits for it.
template <typename Launch>
int run_pipeline(Workspace& ws, long long n, std::vector<Column>& cols, bool reset_key,
                 Launch&& launch) {
  const long long chunk = chunk_points(n);
  size_t off = 0;
  bool need_staging = false;
  for (auto& c : cols) {
    c.pinned = host_pinned(c.host_in ? c.host_in : c.host_out);
    need_staging |= !c.pinned;
    c.dev_off = off;
    off += (static_cast<size_t>(chunk) * c.elem + 255) & ~static_cast<size_t>(255);
  }
  int st = ws.ensure(off, need_staging ? off : 0);
  if (st != PYI_OK) return st;
  if (reset_key) {
    PYI_CUDA(cudaMemsetAsync(ws.oob_key, 0xff, sizeof(unsigned long long), ws.slot[0].stream));
    PYI_CUDA(cudaEventRecord(ws.ready, ws.slot[0].stream));
    for (int k = 1; k < kSlots; ++k) PYI_CUDA(cudaStreamWaitEvent(ws.slot[k].stream, ws.ready, 0));
  }

  auto drain = [&](Slot& s) -> int {
    if (!s.busy) return PYI_OK;
    PYI_CUDA(cudaEventSynchronize(s.done));
    for (auto& c : cols) {
      if (c.host_out && !c.pinned) {
        parallel_memcpy(static_cast<char*>(c.host_out) + static_cast<size_t>(s.begin) * c.elem,
                        s.pinned + c.dev_off, static_cast<size_t>(s.count) * c.elem);
      }
    }
    s.busy = false;
    return PYI_OK;
  };

  int which = 0;
  for (long long begin = 0; begin < n; begin += chunk, which = (which + 1) % kSlots) {
    Slot& s = ws.slot[which];
    st = drain(s);
    if (st != PYI_OK) return st;
    const long long count = std::min(chunk, n - begin);
    s.begin = begin;
    s.count = count;
    s.busy = true;  // from here on the slot's stream may hold work of this call
    for (auto& c : cols) {
      if (!c.host_in) continue;
      const size_t bytes = static_cast<size_t>(count) * c.elem;
      const char* src = static_cast<const char*>(c.host_in) + static_cast<size_t>(begin) * c.elem;
      if (!c.pinned) {
        parallel_memcpy(s.pinned + c.dev_off, src, bytes);
        src = s.pinned + c.dev_off;
      }
      PYI_CUDA(cudaMemcpyAsync(s.dev + c.dev_off, src, bytes, cudaMemcpyHostToDevice, s.stream));
    }
    cudaError_t e = launch(s.stream, begin, count, s.dev, s.queue);
    if (e != cudaSuccess) return cuda_fail(e, "kernel launch");
    for (auto& c : cols) {
      if (!c.host_out) continue;
      const size_t bytes = static_cast<size_t>(count) * c.elem;
      char* dst = c.pinned ? static_cast<char*>(c.host_out) + static_cast<size_t>(begin) * c.elem
                           : s.pinned + c.dev_off;
      PYI_CUDA(cudaMemcpyAsync(dst, s.dev + c.dev_off, bytes, cudaMemcpyDeviceToHost, s.stream));
    }
    PYI_CUDA(cudaEventRecord(s.done, s.stream));
  }
  for (int k = 0; k < kSlots; ++k) {
    st = drain(ws.slot[(which + k) % kSlots]);
    if (st != PYI_OK) return st;
  }
  return PYI_OK;
}

// Reads the workspace's bounds-error key once every chunk has drained.
int read_key(Workspace& ws, unsigned long long& key) {
  PYI_CUDA(cudaMemcpyAsync(ws.oob_host, ws.oob_key, sizeof(unsigned long long), cudaMemcpyDeviceToHost,
                           ws.slot[0].stream));
  PYI_CUDA(cudaStreamSynchronize(ws.slot[0].stream));
  key = *ws.oob_host;
  return PYI_OK;
}

// pyi_dist_init: how many devices one host-memory call may be sharded over
std::atomic<int> g_dist_devices{1};
constexpr long long kShardMinPoints = 1LL << 21;  // per device

// Shared tail of the compute entry points: decode the atomicMin key.
int report_oob(unsigned long long key, int64_t* oob_index, int* oob_axis) {
  if (key == pyi::kNoOob) return PYI_OK;
  if (oob_index) *oob_index = static_cast<int64_t>(key >> 2);
  if (oob_axis) *oob_axis = static_cast<int>(key & 3);
  return PYI_OOB;
}

}  // namespace

// Device copies of a grid on a device other than the one it was created on
// (single-call multi-GPU sharding, pyi_dist_init).
struct GridReplica {
  void* d_data{nullptr};
  void* d_packed{nullptr};
  void* d_planes{nullptr};
  pyi::DAxis dax[4];
  uint64_t version{0};
};

struct pyi_grid {
  int ndim{0};
  int dtype{0};
  int device{0};
  int64_t shape[4]{1, 1, 1, 1};
  size_t bytes{0};
  size_t packed_bytes{0};
  std::unique_ptr<pyi_axis> axes[4];
  pyi::DAxis dax[4];
  void* d_data{nullptr};
  bool borrowed{false};
  void* d_packed{nullptr};
  // plane-major copy [z][u][x][y] of a 3-D / 4-D grid, built on the first
  // windowed bicubic call (pyi_fastwin_nd.cuh)
  mutable void* d_planes{nullptr};
  mutable bool planes_tried{false};
  mutable std::mutex planes_mu;
  // bumped whenever a device copy on `device` changes; replicas re-sync lazily
  mutable std::atomic<uint64_t> version{1};
  mutable std::mutex rep_mu;
  mutable std::map<int, std::unique_ptr<GridReplica>> replicas;
  // natural-spline constants of a uniformly spaced six-node window on x / y
  // (regular axes with moderate steps), computed on the device at creation
  double u6[2][24]{};
  int u6_ok[2]{0, 0};
  // node-derivative weights of the six-node spline on x / y (pyi_bicubicw.cuh)
  int w_ok{0};
  double w6[2][12]{};
  // cubic coefficients of the separable bicubic stencil (pyi_bicubics.cuh)
  int cw_ok{0};
  double cw[2][8][4]{};
  // The caller's host array when it may change between calls (pyi_grid_track_host): page-locked in
  // place, so that a refresh runs at link speed and a small call can read it over PCIe directly.
  const void* host_src{nullptr};
  const void* host_dev{nullptr};  // device-side alias of host_src, null when it could not be mapped
  bool host_registered{false};    // this handle page-locked it (released in pyi_grid_destroy)
  mutable std::atomic<int> stale{0};
  // the plain device array is current but the packed / plane-major copies are not (a refresh made
  // for a call too small to pay for rebuilding them)
  mutable std::atomic<int> derived_stale{0};
  mutable std::mutex update_mu;
};

extern "C" {
static int grid_planes(const pyi_grid* grid, bool refresh);
}

namespace {

// Window weights of the separable bicubic stencil as cubics in the cell
// coordinate t (pyi_bicubics.cuh): C_j(t) = [j==2] h00 + [j==3] h01 +
// step (h10 W2[j] + h11 W3[j]) with the cubic Hermite basis h00 = 1 - 3t^2 +
// 2t^3, h01 = 3t^2 - 2t^3, h10 = t - 2t^2 + t^3, h11 = t^3 - t^2, and W2 / W3
// the spline derivative weights at the two centre nodes (w6).
void stencil_cubics(const double* w6, double step, double (*cw)[4]) {
  for (int e = 0; e < 8; ++e) {
    for (int c = 0; c < 4; ++c) cw[e][c] = 0.0;
  }
  for (int j = 0; j < 6; ++j) {
    const double a = step * w6[j], b = step * w6[6 + j];
    const double d2 = j == 2 ? 1.0 : 0.0, d3 = j == 3 ? 1.0 : 0.0;
    cw[j + 1][0] = d2;
    cw[j + 1][1] = a;
    cw[j + 1][2] = -3.0 * d2 + 3.0 * d3 - 2.0 * a - b;
    cw[j + 1][3] = 2.0 * d2 - 2.0 * d3 + a + b;
  }
}

// The same table for the fitting models that are LINEAR in the node values and
// cubic on a cell, bivariate/spline.hpp:96-118 (y, then x) being separable for
// them: `linear` (univariate/linear.hpp) and windowed `bilinear`
// (bivariate/bilinear.hpp, the same function), `c_spline` (cspline.hpp:84-139),
// `c_spline_not_a_knot` (cspline_not_a_knot.hpp:114-215) and `c_spline_periodic`
// (cspline_periodic.hpp:129-201).  On uniformly spaced nodes the spline's node
// coefficients c_i (CSplineBase::x_) times step^2 solve, whatever the step,
//   G[i-1] + 4 G[i] + G[i+1] = 3 (y[i+1] - 2 y[i] + y[i-1]),  i = 1..4
// closed by G[0] = G[5] = 0 (natural), by G[0] - 2 G[1] + G[2] = 0 and
// G[3] - 2 G[4] + G[5] = 0 (not-a-knot), or by G[0] = G[5] and the wrap-around
// row G[4] + 4 G[5] + G[1] = 3 ((y[1] - y[0]) - (y[5] - y[4])) (periodic); on
// the centre cell (cspline_base.hpp:32-40,112)
//   s(t) = y2 + t (y3 - y2 - (G3 + 2 G2) / 3) + t^2 G2 + t^3 (G3 - G2) / 3 .
// Solved for the six unit vectors in long double.  False for the other models.
bool model_stencil_cubics(int method, double (*cw)[4]) {
  for (int e = 0; e < 8; ++e) {
    for (int c = 0; c < 4; ++c) cw[e][c] = 0.0;
  }
  if (method == PYI_S_LINEAR || method == PYI_S_BILINEAR) {
    cw[3][0] = 1.0;
    cw[3][1] = -1.0;
    cw[4][1] = 1.0;
    return true;
  }
  if (method != PYI_S_C_SPLINE && method != PYI_S_C_SPLINE_NOT_A_KNOT && method != PYI_S_C_SPLINE_PERIODIC) {
    return false;
  }
  // rows: equations, columns 0..5: G[0..5], columns 6..11: the six unit vectors y = e_j
  long double a[6][12] = {};
  for (int i = 1; i <= 4; ++i) {
    a[i][i - 1] = 1.0L;
    a[i][i] = 4.0L;
    a[i][i + 1] = 1.0L;
    a[i][6 + i - 1] += 3.0L;
    a[i][6 + i] -= 6.0L;
    a[i][6 + i + 1] += 3.0L;
  }
  if (method == PYI_S_C_SPLINE) {
    a[0][0] = 1.0L;
    a[5][5] = 1.0L;
  } else if (method == PYI_S_C_SPLINE_NOT_A_KNOT) {
    a[0][0] = 1.0L, a[0][1] = -2.0L, a[0][2] = 1.0L;
    a[5][3] = 1.0L, a[5][4] = -2.0L, a[5][5] = 1.0L;
  } else {
    a[0][0] = 1.0L, a[0][5] = -1.0L;
    a[5][4] = 1.0L, a[5][5] = 4.0L, a[5][1] = 1.0L;
    a[5][6 + 1] += 3.0L, a[5][6 + 0] -= 3.0L, a[5][6 + 5] -= 3.0L, a[5][6 + 4] += 3.0L;
  }
  for (int c = 0; c < 6; ++c) {  // Gauss-Jordan with partial pivoting
    int piv = c;
    for (int r = c + 1; r < 6; ++r) {
      if (std::fabs(a[r][c]) > std::fabs(a[piv][c])) piv = r;
    }
    if (a[piv][c] == 0.0L) return false;
    for (int k = 0; k < 12; ++k) std::swap(a[c][k], a[piv][k]);
    const long double inv = 1.0L / a[c][c];
    for (int k = 0; k < 12; ++k) a[c][k] *= inv;
    for (int r = 0; r < 6; ++r) {
      if (r == c || a[r][c] == 0.0L) continue;
      const long double f = a[r][c];
      for (int k = 0; k < 12; ++k) a[r][k] -= f * a[c][k];
    }
  }
  for (int j = 0; j < 6; ++j) {
    const long double g2 = a[2][6 + j], g3 = a[3][6 + j];
    const long double d2 = j == 2 ? 1.0L : 0.0L, d3 = j == 3 ? 1.0L : 0.0L;
    cw[j + 1][0] = static_cast<double>(d2);
    cw[j + 1][1] = static_cast<double>(d3 - d2 - (g3 + 2.0L * g2) / 3.0L);
    cw[j + 1][2] = static_cast<double>(g2);
    cw[j + 1][3] = static_cast<double>((g3 - g2) / 3.0L);
  }
  return true;
}

// `polynomial` (univariate/polynomial.hpp:58-102) on six uniformly spaced nodes as six
// quintics in the cell coordinate t: the value is the Newton form
// sum_k work[k] prod_{m<k} (x - xa[m]) of the reference's divided-difference table, which is
// linear in the node values.  The table is the REFERENCE's: its in-place update reads entries
// it has already overwritten (Eigen's SSE2 packet traversal of an aliased expression, restated
// in pyi_univariate.cuh: uni_polynomial), mixing difference orders -- so, unlike the spline
// models, the weights depend on the spacing and are built per axis from its step.
void polynomial_stencil(double step, double (*pw)[6]) {
  constexpr int n = 6;
  const long double h = step;
  long double work[n][n] = {};  // work[k][j]: coefficient of y[j] in work[k]
  work[0][0] = 1.0L;
  for (int k = 1; k < n; ++k) {
    work[k][k] = 1.0L / h;
    work[k][k - 1] = -1.0L / h;
  }
  auto update = [&](long double* dst, const long double* a, const long double* b, int ix) {
    for (int j = 0; j < n; ++j) dst[j] = (a[j] - b[j]) / (ix * h);
  };
  for (int ix = 2; ix < n; ++ix) {
    const int cnt = n - ix;
    int k = 0;
    const int peel = (ix & 1) < cnt ? (ix & 1) : cnt;
    for (; k < peel; ++k) update(work[ix + k], work[ix + k], work[ix - 1 + k], ix);
    for (; k + 2 <= cnt; k += 2) {  // a packet: both lanes read before either is stored
      long double a0[n], a1[n];
      update(a0, work[ix + k], work[ix - 1 + k], ix);
      update(a1, work[ix + k + 1], work[ix + k], ix);
      for (int j = 0; j < n; ++j) {
        work[ix + k][j] = a0[j];
        work[ix + k + 1][j] = a1[j];
      }
    }
    for (; k < cnt; ++k) update(work[ix + k], work[ix + k], work[ix - 1 + k], ix);
  }
  long double poly[n][n] = {};  // poly[j][d]: coefficient of t^d in the weight of y[j]
  long double basis[n] = {1.0L};  // prod_{m<k} (x - xa[m]) with x = xa[2] + t h
  for (int k = 0; k < n; ++k) {
    if (k > 0) {
      const long double c = 2.0L - (k - 1);
      long double next[n];
      for (int d = 0; d < n; ++d) next[d] = ((d > 0 ? basis[d - 1] : 0.0L) + c * basis[d]) * h;
      for (int d = 0; d < n; ++d) basis[d] = next[d];
    }
    for (int j = 0; j < n; ++j) {
      for (int d = 0; d < n; ++d) poly[j][d] += work[k][j] * basis[d];
    }
  }
  for (int j = 0; j < n; ++j) {
    for (int d = 0; d < n; ++d) pw[j][d] = static_cast<double>(poly[j][d]);
  }
}

// What a kernel launch on one device reads of a grid.
struct GridView {
  const void* data{nullptr};
  const void* packed{nullptr};
  const void* planes{nullptr};
  const pyi::DAxis* dax{nullptr};
};

void free_replica(int device, GridReplica& r) {
  DeviceGuard guard(device);
  cudaFree(r.d_data);
  cudaFree(r.d_packed);
  cudaFree(r.d_planes);
  r = GridReplica{};
}

// View of `grid` on `device` (the calling thread's current device).  On the
// grid's own device that is the handle itself; elsewhere a replica is created
// or refreshed with peer copies (NVLink when peer access is available).
// Set for the duration of a small host-memory call on a tracked grid: kernels read the caller's
// page-locked array through its device alias (no packed copy, no plane-major copy: generic kernels).
thread_local const void* tl_inplace = nullptr;

int grid_view(const pyi_grid* grid, int device, GridView& view) {
  if (device == grid->device && tl_inplace != nullptr) {
    view.data = tl_inplace;
    view.packed = nullptr;
    view.planes = nullptr;
    view.dax = grid->dax;
    return PYI_OK;
  }
  if (device == grid->device) {
    view.data = grid->d_data;
    view.packed = grid->d_packed;
    view.planes = grid->d_planes;
    view.dax = grid->dax;
    return PYI_OK;
  }
  std::lock_guard<std::mutex> lock(grid->rep_mu);
  auto& slot = grid->replicas[device];
  if (!slot) slot = std::make_unique<GridReplica>();
  GridReplica& r = *slot;
  const uint64_t version = grid->version.load();
  if (r.version != version) {
    int can = 0;
    if (cudaDeviceCanAccessPeer(&can, device, grid->device) == cudaSuccess && can) {
      const cudaError_t e = cudaDeviceEnablePeerAccess(grid->device, 0);
      if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
    }
    cudaGetLastError();
    auto mirror = [&](void*& dst, const void* src, size_t bytes) -> int {
      if (!src) {
        if (dst) cudaFree(dst);
        dst = nullptr;
        return PYI_OK;
      }
      if (!dst) PYI_CUDA(cudaMalloc(&dst, std::max<size_t>(bytes, 64)));
      PYI_CUDA(cudaMemcpyPeer(dst, device, src, grid->device, bytes));
      return PYI_OK;
    };
    int st = mirror(r.d_data, grid->d_data, grid->bytes);
    if (st == PYI_OK) st = mirror(r.d_packed, grid->d_packed, grid->packed_bytes);
    if (st == PYI_OK) st = mirror(r.d_planes, grid->d_planes, grid->bytes);
    for (int k = 0; st == PYI_OK && k < grid->ndim; ++k) st = device_axis(grid->axes[k].get(), r.dax[k]);
    if (st != PYI_OK) return st;
    PYI_CUDA(cudaDeviceSynchronize());
    r.version = version;
  }
  view.data = r.d_data;
  view.packed = r.d_packed;
  view.planes = r.d_planes;
  view.dax = r.dax;
  return PYI_OK;
}

// Devices a host-memory call of n points is sharded over: the grid's own
// device first, then the others below pyi_dist_init's count; at least
// kShardMinPoints per device.
std::vector<int> shard_devices(const pyi_grid* grid, long long n) {
  std::vector<int> devices{grid->device};
  const int limit = std::min(g_dist_devices.load(), device_count_cached());
  const long long want = std::max<long long>(1, n / kShardMinPoints);
  for (int d = 0; d < limit && static_cast<long long>(devices.size()) < want &&
                  static_cast<int>(devices.size()) < limit; ++d) {
    if (d != grid->device) devices.push_back(d);
  }
  return devices;
}

// A host-memory call: the n points are split into contiguous shards, one per
// device, exactly as the reference's parallel_for splits them over threads
// (parallel_for.hpp:46-53: n / G each, the first n % G shards take one more);
// every shard runs its own chunk pipeline on its own workspace from its own
// host thread.  `launch_on(view, oob_key, stream, global_begin, count, dev,
// columns, slot_queue)` enqueues the kernel of one chunk.  `key` receives the minimum bounds-error
// key over all devices (indexes are global, so the lowest offending point wins
// whatever the sharding).
template <typename LaunchOn>
int run_host_call(const pyi_grid* grid, long long n, const std::vector<Column>& cols, bool want_key,
                  unsigned long long& key, LaunchOn&& launch_on) {
  struct ShardResult {
    int status{PYI_OK};
    std::string error;
    unsigned long long key{pyi::kNoOob};
  };
  const std::vector<int> devices = shard_devices(grid, n);
  const int shards = static_cast<int>(devices.size());
  std::vector<ShardResult> results(shards);
  auto work = [&](int r) {
    ShardResult& out = results[r];
    auto run = [&]() -> int {
      DeviceGuard guard(devices[r]);
      if (guard.err != cudaSuccess) return cuda_fail(guard.err, "cudaSetDevice");
      GridView view;
      int st = grid_view(grid, devices[r], view);
      if (st != PYI_OK) return st;
      const long long each = n / shards, rem = n % shards;
      const long long begin = r * each + std::min<long long>(r, rem);
      const long long count = each + (r < rem ? 1 : 0);
      if (count == 0) return PYI_OK;
      std::vector<Column> mine = cols;
      for (auto& c : mine) {
        const size_t shift = static_cast<size_t>(begin) * c.elem;
        if (c.host_in) c.host_in = static_cast<const char*>(c.host_in) + shift;
        if (c.host_out) c.host_out = static_cast<char*>(c.host_out) + shift;
      }
      Lease lease(devices[r]);
      Workspace& ws = *lease.ws;
      st = run_pipeline(ws, count, mine, true,
                        [&](cudaStream_t s, long long b, long long c, char* dev, unsigned* queue) {
                          return launch_on(view, ws.oob_key, s, begin + b, c, dev, mine, queue);
                        });
      if (st != PYI_OK) return st;
      if (want_key) return read_key(ws, out.key);
      return PYI_OK;
    };
    out.status = run();
    if (out.status != PYI_OK) out.error = g_error;
  };
  if (shards == 1) {
    work(0);
  } else {
    std::vector<std::thread> threads;
    for (int r = 1; r < shards; ++r) threads.emplace_back(work, r);
    work(0);
    for (auto& t : threads) t.join();
  }
  key = pyi::kNoOob;
  for (const auto& r : results) {
    if (r.status != PYI_OK) return fail(r.status, r.error);
    key = std::min(key, r.key);
  }
  return PYI_OK;
}

// The stream-ordered allocator serves the per-launch scratch of device-memory
// calls (bounds-error key, queue of unsettled points).  By default a pool gives
// its memory back to the driver at every synchronisation, which makes the next
// allocation a (process-wide serialising) driver call: keep it cached instead.
void keep_pool_cached(int device) {
  static std::mutex mu;
  static std::vector<char> done;
  std::lock_guard<std::mutex> lock(mu);
  if (device < 0) return;
  if (static_cast<size_t>(device) >= done.size()) done.resize(device + 1, 0);
  if (done[device]) return;
  cudaMemPool_t pool = nullptr;
  if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) {
    unsigned long long threshold = ~0ULL;
    cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &threshold);
  }
  cudaGetLastError();
  done[device] = 1;
}

// Per-call bounds-error key of a device-memory call, from the stream-ordered
// allocator: concurrent calls on different streams do not share it.
struct DeviceKey {
  unsigned long long* ptr{nullptr};
  cudaStream_t stream{nullptr};
  int open(cudaStream_t s) {
    stream = s;
    PYI_CUDA(cudaMallocAsync(reinterpret_cast<void**>(&ptr), sizeof(unsigned long long), s));
    PYI_CUDA(cudaMemsetAsync(ptr, 0xff, sizeof(unsigned long long), s));
    return PYI_OK;
  }
  int read(unsigned long long& key) {
    key = pyi::kNoOob;
    PYI_CUDA(cudaMemcpyAsync(&key, ptr, sizeof(key), cudaMemcpyDeviceToHost, stream));
    PYI_CUDA(cudaStreamSynchronize(stream));
    return PYI_OK;
  }
  ~DeviceKey() {
    if (ptr) cudaFreeAsync(ptr, stream);
  }
};

// A device-memory call works on the caller's buffers: they live on the
// calling thread's current device, which must be the grid's.
int require_grid_device(const pyi_grid* grid) {
  int dev = -1;
  PYI_CUDA(cudaGetDevice(&dev));
  if (dev != grid->device) {
    return fail(PYI_BAD_ARG, "device-memory call: the current device (" + std::to_string(dev) +
                                 ") is not the grid's device (" + std::to_string(grid->device) + ")");
  }
  keep_pool_cached(dev);
  return PYI_OK;
}

// A tracked grid whose host array may have changed since the device copies were made
// (pyi_grid_mark_stale): a small host-memory call reads the array in place -- `touched` bytes of
// scattered sectors against the whole array over the same link -- and leaves the copies stale;
// any other call refreshes them first.
struct InPlace {
  bool on{false};
  ~InPlace() {
    if (on) tl_inplace = nullptr;
  }
};
int rebuild_derived(pyi_grid* grid) {
  if (grid->d_planes) {
    const int st = grid_planes(grid, true);
    if (st != PYI_OK) return st;
  }
  if (grid->d_packed) return pyi_grid_pack(grid);
  return PYI_OK;
}

// `plain_rate`: points per second of the kernels this call would run on the plain array (the generic
// ones; for 2-D windowed calls the fast kernels themselves, which read nothing else).
int sync_grid(const pyi_grid* grid, long long n, long long touched_per_point, int mem, InPlace& inplace,
              double plain_rate) {
  if (grid->host_src == nullptr) return PYI_OK;
  pyi_grid* g = const_cast<pyi_grid*>(grid);
  if (grid->stale.load() != 0) {
    if (mem == PYI_MEM_HOST && grid->host_dev != nullptr && n < kShardMinPoints &&
        static_cast<double>(n) * static_cast<double>(touched_per_point) * 4.0 <= static_cast<double>(grid->bytes)) {
      tl_inplace = grid->host_dev;
      inplace.on = true;
      return PYI_OK;
    }
    std::lock_guard<std::mutex> lock(grid->update_mu);
    if (grid->stale.load() != 0) {
      // the plain array only; the derived copies follow below when this call can use them
      DeviceGuard guard(grid->device);
      if (guard.err != cudaSuccess) return cuda_fail(guard.err, "cudaSetDevice");
      if (grid->borrowed) return fail(PYI_BAD_ARG, "a grid over borrowed device memory tracks no host array");
      PYI_CUDA(cudaMemcpy(g->d_data, grid->host_src, grid->bytes, cudaMemcpyHostToDevice));
      grid->derived_stale.store(1);
      grid->stale.store(0);
      grid->version.fetch_add(1);
    }
  }
  if (grid->derived_stale.load() != 0) {
    // A call that the kernels reading the plain array finish sooner than the packed (2^ndim x the
    // array) and plane-major copies take to rebuild runs on the plain array; the next larger call
    // rebuilds them.
    const double rebuild_s = 50e-6 + static_cast<double>(grid->bytes) *
                                         (static_cast<double>(1 << grid->ndim) / 1.5e12 + 1.0 / 1.0e12);
    if (mem == PYI_MEM_HOST && n < kShardMinPoints && grid->bytes >= (1u << 20) &&
        static_cast<double>(n) < plain_rate * rebuild_s) {
      tl_inplace = grid->d_data;
      inplace.on = true;
      return PYI_OK;
    }
    std::lock_guard<std::mutex> lock(grid->update_mu);
    if (grid->derived_stale.load() != 0) {
      DeviceGuard guard(grid->device);
      if (guard.err != cudaSuccess) return cuda_fail(guard.err, "cudaSetDevice");
      const int st = rebuild_derived(g);
      if (st != PYI_OK) return st;
      grid->derived_stale.store(0);
    }
  }
  return PYI_OK;
}

}  // namespace

extern "C" {

int pyi_abi_version(void) { return PYI_ABI_VERSION; }
int pyi_device_count(void) { return device_count_cached(); }
const char* pyi_last_error(void) { return g_error.c_str(); }
int64_t pyi_launch_count(void) { return pyi::launch_count(); }

int pyi_set_device(int device) {
  int st = require_device();
  if (st != PYI_OK) return st;
  PYI_CUDA(cudaSetDevice(device));
  return PYI_OK;
}

// Single-call multi-GPU sharding (the reference's drop-in
// caller is ONE process whose parallel_for splits a call's points over its
// workers, parallel_for.hpp:30-78): from now on every host-memory
// interpolation call of this process is split into contiguous shards over the
// first `n_devices` CUDA devices.  Returns the count in effect.
int pyi_dist_init(int n_devices) {
  const int have = device_count_cached();
  if (have <= 0) {
    fail(PYI_NO_DEVICE, "no CUDA device: pyinterp_b200 runs on the GPU only (no CPU fallback)");
    return 0;
  }
  const int n = n_devices <= 0 ? have : std::min(n_devices, have);
  g_dist_devices.store(n);
  return n;
}

int pyi_dist_size(void) { return std::min(g_dist_devices.load(), std::max(1, device_count_cached())); }

// ---- axes -----------------------------------------------------------------
pyi_axis* pyi_axis_create_f64(const double* values, int64_t n, double epsilon, int has_period,
                              double period, int* status) {
  std::string err;
  pyi_axis* a = pyi_host::make_axis<double>(values, n, epsilon, has_period != 0, period, err);
  if (!a) g_error = err;
  if (status) *status = a ? PYI_OK : PYI_BAD_ARG;
  return a;
}

pyi_axis* pyi_axis_create_i64(const int64_t* values, int64_t n, int64_t epsilon, int has_period,
                              int64_t period, int* status) {
  std::string err;
  pyi_axis* a = pyi_host::make_axis<int64_t>(values, n, epsilon, has_period != 0, period, err);
  if (!a) g_error = err;
  if (status) *status = a ? PYI_OK : PYI_BAD_ARG;
  return a;
}

pyi_axis* pyi_axis_create_from_desc(const pyi_axis_desc* desc, int* status) {
  if (!desc || desc->size <= 0 ||
      (desc->kind == PYI_AXIS_IRREGULAR && desc->points == nullptr)) {
    g_error = "invalid axis descriptor";
    if (status) *status = PYI_BAD_ARG;
    return nullptr;
  }
  auto a = new pyi_axis();
  a->d = *desc;
  if (desc->kind == PYI_AXIS_IRREGULAR) {
    if (desc->coord_type == PYI_COORD_F64) {
      const double* p = static_cast<const double*>(desc->points);
      a->pf.assign(p, p + desc->size);
    } else {
      const int64_t* p = static_cast<const int64_t*>(desc->points);
      a->pi.assign(p, p + desc->size);
    }
  }
  a->fix_points();
  pyi_host::build_lut(*a);
  if (status) *status = PYI_OK;
  return a;
}

int pyi_axis_flip(pyi_axis* axis) {
  if (!axis) return fail(PYI_BAD_ARG, "null axis");
  {
    std::lock_guard<std::mutex> lock(axis->mu);
    free_axis_device(axis);
  }
  pyi_host::flip_axis(*axis);
  return PYI_OK;
}

int pyi_axis_describe(const pyi_axis* axis, pyi_axis_desc* out) {
  if (!axis || !out) return fail(PYI_BAD_ARG, "null argument");
  *out = axis->d;
  return PYI_OK;
}

void pyi_axis_destroy(pyi_axis* axis) {
  if (axis) {
    free_axis_device(axis);
    delete axis;
  }
}

// ---- axis lookups -----------------------------------------------------------
static int axis_lookup(const pyi_axis* axis_c, int64_t n, const void* coords, int bounded,
                       int64_t* out, int mem, void* stream, bool pairs) {
  int st = require_device();
  if (st != PYI_OK) return st;
  if (!axis_c || n < 0 || (n > 0 && (!coords || !out))) return fail(PYI_BAD_ARG, "null argument");
  pyi_axis* axis = const_cast<pyi_axis*>(axis_c);
  pyi::FindParams p{};
  st = device_axis(axis, p.ax);
  if (st != PYI_OK) return st;
  p.bounded = bounded;
  const size_t out_elem = pairs ? 16 : 8;
  if (mem == PYI_MEM_DEVICE) {
    p.n = n;
    p.coords = coords;
    p.out = reinterpret_cast<long long*>(out);
    cudaError_t e = pairs ? pyi::launch_find_indexes(p, static_cast<cudaStream_t>(stream))
                          : pyi::launch_find_index(p, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return cuda_fail(e, "find_indexes");
    return PYI_OK;
  }
  std::vector<Column> cols(2);
  cols[0].host_in = coords;
  cols[0].elem = 8;
  cols[1].host_out = out;
  cols[1].elem = out_elem;
  int dev = 0;
  PYI_CUDA(cudaGetDevice(&dev));
  Lease lease(dev);
  return run_pipeline(*lease.ws, n, cols, false, [&](cudaStream_t s, long long, long long count, char* d, unsigned*) {
    pyi::FindParams q = p;
    q.n = count;
    q.coords = d + cols[0].dev_off;
    q.out = reinterpret_cast<long long*>(d + cols[1].dev_off);
    return pairs ? pyi::launch_find_indexes(q, s) : pyi::launch_find_index(q, s);
  });
}

int pyi_axis_find_indexes(const pyi_axis* axis, int64_t n, const void* coords,
                          int64_t* out_pairs, int mem, void* stream) {
  return axis_lookup(axis, n, coords, 0, out_pairs, mem, stream, true);
}

int pyi_axis_find_index(const pyi_axis* axis, int64_t n, const void* coords, int bounded,
                        int64_t* out, int mem, void* stream) {
  return axis_lookup(axis, n, coords, bounded, out, mem, stream, false);
}

// ---- grids ------------------------------------------------------------------
pyi_grid* pyi_grid_create(int ndim, const pyi_axis* const* axes, int dtype, const void* data,
                          const int64_t* shape, int mem, unsigned flags, int* status) {
  auto set = [&](int s) {
    if (status) *status = s;
    return static_cast<pyi_grid*>(nullptr);
  };
  if (require_device() != PYI_OK) return set(PYI_NO_DEVICE);
  if (ndim < 1 || ndim > 4 || !axes || !data || !shape || dtype < PYI_I8 || dtype > PYI_F64) {
    g_error = "pyi_grid_create: invalid argument";
    return set(PYI_BAD_ARG);
  }
  auto g = std::make_unique<pyi_grid>();
  g->ndim = ndim;
  g->dtype = dtype;
  size_t count = 1;
  static const char* labels[4] = {"x", "y", "z", "u"};
  for (int k = 0; k < ndim; ++k) {
    if (!axes[k]) {
      g_error = "pyi_grid_create: null axis";
      return set(PYI_BAD_ARG);
    }
    // pybind/grid.hpp:443-481: only x may be periodic, shape must match the
    // axes; src/pybind/grid.cpp:68-92: an int64 (temporal) axis only on z
    if (k > 0 && axes[k]->d.is_periodic) {
      g_error = std::string(labels[k]) + "-axis cannot be a circle";
      return set(PYI_BAD_ARG);
    }
    if (shape[k] != axes[k]->d.size) {
      g_error = std::string(labels[k]) + " axis size (" + std::to_string(axes[k]->d.size) +
                ") doesn't match array shape[" + std::to_string(k) + "] (" +
                std::to_string(shape[k]) + ")";
      return set(PYI_BAD_SHAPE);
    }
    if (axes[k]->d.coord_type == PYI_COORD_I64 && k != 2) {
      g_error = "an int64 (temporal) axis is only supported on the z axis";
      return set(PYI_BAD_ARG);
    }
    g->shape[k] = shape[k];
    count *= static_cast<size_t>(shape[k]);
    g->axes[k] = std::make_unique<pyi_axis>(*axes[k]);
  }
  g->bytes = count * dtype_size(dtype);
  if (cudaGetDevice(&g->device) != cudaSuccess) {
    g_error = "cudaGetDevice failed";
    return set(PYI_CUDA_ERROR);
  }
  for (int k = 0; k < ndim; ++k) {
    int st = device_axis(g->axes[k].get(), g->dax[k]);
    if (st != PYI_OK) return set(st);
  }
  // a borrowed pointer must satisfy the 16-byte vector loads of the windowed
  // kernels; anything else is copied into an aligned allocation
  if (mem == PYI_MEM_DEVICE && (flags & PYI_GRID_BORROW_DEVICE) &&
      (reinterpret_cast<uintptr_t>(data) & 15) == 0) {
    g->d_data = const_cast<void*>(data);
    g->borrowed = true;
  } else {
    cudaError_t e = cudaMalloc(&g->d_data, std::max<size_t>(g->bytes, 16));
    if (e != cudaSuccess) return set(cuda_fail(e, "cudaMalloc(grid)"));
    e = cudaMemcpy(g->d_data, data, g->bytes,
                   mem == PYI_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
      cudaFree(g->d_data);
      return set(cuda_fail(e, "cudaMemcpy(grid)"));
    }
  }
  if (ndim >= 2) {
    double* d_tmp = nullptr;
    if (cudaMalloc(reinterpret_cast<void**>(&d_tmp), 24 * sizeof(double)) == cudaSuccess) {
      for (int k = 0; k < 2; ++k) {
        const pyi::DAxis& a = g->dax[k];
        const double mag = std::fabs(a.step);
        if (a.kind != 0 || a.coord_i64 || !(mag >= 9.313225746154785e-10 && mag <= 1073741824.0)) {
          continue;  // irregular, or outside the spacing range of the fast path (2^-30 .. 2^30)
        }
        if (pyi::launch_spline6_uniform(a.step, d_tmp, nullptr) == cudaSuccess &&
            cudaMemcpy(g->u6[k], d_tmp, 24 * sizeof(double), cudaMemcpyDeviceToHost) == cudaSuccess) {
          g->u6_ok[k] = 1;
        }
      }
      cudaFree(d_tmp);
    }
    cudaGetLastError();
    // derivative weights of the weights kernel: both axes regular (and closing
    // their period when circular), steps in the fast path's range
    double* d_w = nullptr;
    if (cudaMalloc(reinterpret_cast<void**>(&d_w), 12 * sizeof(double)) == cudaSuccess) {
      int have = 0;
      for (int k = 0; k < 2; ++k) {
        const pyi::DAxis& a = g->dax[k];
        const double mag = std::fabs(a.step);
        const bool closes = !a.periodic || std::fabs(static_cast<double>(a.n) * a.step) == a.period;
        // the weights are those of exactly uniform nodes k * step; the
        // reference's nodes start + k * step are rounded to ulp(|coordinate|),
        // a relative derivative error of ~ulp(|coordinate|) / |step|: only
        // axes with max |coordinate| / |step| < 2^16 take the weights kernel
        const double reach = std::max(std::fabs(a.min), std::fabs(a.max)) / mag;
        if (a.coord_i64 || a.n < 6 || a.kind != 0 || !closes || !(reach < 65536.0) ||
            !(mag >= 9.313225746154785e-10 && mag <= 1073741824.0)) {
          break;
        }
        if (pyi::launch_spline6_weights(a.step, d_w, nullptr) != cudaSuccess ||
            cudaMemcpy(g->w6[k], d_w, 12 * sizeof(double), cudaMemcpyDeviceToHost) != cudaSuccess) {
          break;
        }
        ++have;
      }
      g->w_ok = have == 2;
      if (g->w_ok && g->dax[0].step > 0.0 && g->dax[1].step > 0.0) {
        for (int k = 0; k < 2; ++k) stencil_cubics(g->w6[k], g->dax[k].step, g->cw[k]);
        g->cw_ok = 1;
      }
      cudaFree(d_w);
    }
    cudaGetLastError();
  }
  pyi_grid* raw = g.release();
  if (flags & PYI_GRID_PACK_CELLS) {
    int st = pyi_grid_pack(raw);
    if (st != PYI_OK) {
      pyi_grid_destroy(raw);
      return set(st);
    }
  }
  if (status) *status = PYI_OK;
  return raw;
}

// Builds (or refreshes) the plane-major copy of a 3-D / 4-D grid.  It is an
// optimisation: when it does not fit, the windowed kernels keep the C-order
// array.
static int grid_planes(const pyi_grid* grid, bool refresh) {
  if (grid->ndim < 3) return PYI_OK;
  std::lock_guard<std::mutex> lock(grid->planes_mu);
  if (grid->d_planes && !refresh) return PYI_OK;
  if (!grid->d_planes) {
    if (grid->planes_tried) return PYI_OK;
    grid->planes_tried = true;
    size_t free_b = 0, total_b = 0;
    PYI_CUDA(cudaMemGetInfo(&free_b, &total_b));
    if (grid->bytes > free_b / 2) return PYI_OK;
    PYI_CUDA(cudaMalloc(&grid->d_planes, std::max<size_t>(grid->bytes, 64)));
    refresh = true;
  }
  const long long shape[4] = {grid->shape[0], grid->shape[1], grid->shape[2], grid->shape[3]};
  cudaError_t e = pyi::launch_plane_major(grid->d_data, grid->d_planes,
                                          static_cast<int>(dtype_size(grid->dtype)), shape, nullptr);
  if (e != cudaSuccess) return cuda_fail(e, "plane_major");
  PYI_CUDA(cudaStreamSynchronize(nullptr));
  if (refresh) grid->version.fetch_add(1);
  return PYI_OK;
}

int pyi_grid_update(pyi_grid* grid, const void* data, int mem) {
  if (!grid || !data) return fail(PYI_BAD_ARG, "null argument");
  DeviceGuard guard(grid->device);
  if (guard.err != cudaSuccess) return cuda_fail(guard.err, "cudaSetDevice");
  if (grid->borrowed && (reinterpret_cast<uintptr_t>(data) & 15) == 0 && mem == PYI_MEM_DEVICE) {
    grid->d_data = const_cast<void*>(data);
  } else {
    if (grid->borrowed) {  // the new pointer cannot be borrowed: own a copy from now on
      grid->d_data = nullptr;
      grid->borrowed = false;
      PYI_CUDA(cudaMalloc(&grid->d_data, std::max<size_t>(grid->bytes, 16)));
    }
    PYI_CUDA(cudaMemcpy(grid->d_data, data, grid->bytes,
                        mem == PYI_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice));
  }
  grid->version.fetch_add(1);
  grid->derived_stale.store(0);
  if (grid->d_planes) {
    int st = grid_planes(grid, true);
    if (st != PYI_OK) return st;
  }
  if (grid->d_packed) return pyi_grid_pack(grid);
  return PYI_OK;
}

int pyi_grid_pack(pyi_grid* grid) {
  if (!grid) return fail(PYI_BAD_ARG, "null grid");
  if (grid->ndim < 2) return PYI_OK;  // 1-D grids have no cell-packed copy
  size_t cells = 1;
  int shape[4] = {1, 1, 1, 1};
  for (int k = 0; k < grid->ndim; ++k) {
    cells *= static_cast<size_t>(grid->shape[k]);
    shape[k] = static_cast<int>(grid->shape[k]);
  }
  const size_t elem = result_is_f64(grid->dtype) ? 8 : 4;
  const size_t bytes = (cells << grid->ndim) * elem;
  DeviceGuard guard(grid->device);
  if (guard.err != cudaSuccess) return cuda_fail(guard.err, "cudaSetDevice");
  if (!grid->d_packed) {
    // The packed copy is 2^ndim x the grid (in the result type).  It is an
    // optimisation: when it would not leave room for the query streams the
    // kernels keep reading the plain C-order array.
    size_t free_b = 0, total_b = 0;
    PYI_CUDA(cudaMemGetInfo(&free_b, &total_b));
    if (bytes > free_b / 2) return PYI_OK;
    PYI_CUDA(cudaMalloc(&grid->d_packed, std::max<size_t>(bytes, 64)));
  }
  cudaError_t e = pyi::launch_pack_cells(grid->d_data, grid->dtype, grid->ndim, shape,
                                         grid->d_packed, nullptr);
  if (e != cudaSuccess) return cuda_fail(e, "pack_cells");
  PYI_CUDA(cudaStreamSynchronize(nullptr));
  grid->packed_bytes = bytes;
  grid->version.fetch_add(1);
  return PYI_OK;
}

namespace {
// Buffers page-locked by pyi_grid_track_host, counted per base address: several handles may share one
// array, and it stays locked until the last of them is destroyed.
std::mutex g_reg_mu;
std::map<const void*, std::pair<size_t, int>> g_registered;
}  // namespace

int pyi_grid_track_host(pyi_grid* grid, const void* host_data) {
  if (!grid || !host_data) return fail(PYI_BAD_ARG, "null argument");
  if (grid->host_src != nullptr) return fail(PYI_BAD_ARG, "pyi_grid_track_host: the grid already tracks an array");
  DeviceGuard guard(grid->device);
  if (guard.err != cudaSuccess) return cuda_fail(guard.err, "cudaSetDevice");
  grid->host_src = host_data;
  grid->host_dev = nullptr;
  // Page-lock the array where it lies.  An array somebody else page-locked (the caller, an
  // overlapping registration) is still refreshed at link speed but never read in place: its
  // mapping is not this library's to keep alive.  Failure only costs speed.
  std::lock_guard<std::mutex> lock(g_reg_mu);
  auto it = g_registered.find(host_data);
  if (it != g_registered.end() && it->second.first >= grid->bytes) {
    ++it->second.second;
    grid->host_registered = true;
  } else if (it == g_registered.end()) {
    const cudaError_t e = cudaHostRegister(const_cast<void*>(host_data), std::max<size_t>(grid->bytes, 1),
                                           cudaHostRegisterPortable | cudaHostRegisterMapped);
    if (e == cudaSuccess) {
      g_registered[host_data] = {grid->bytes, 1};
      grid->host_registered = true;
    } else {
      cudaGetLastError();
    }
  }
  if (grid->host_registered) {
    void* alias = nullptr;
    if (cudaHostGetDevicePointer(&alias, const_cast<void*>(host_data), 0) == cudaSuccess) {
      grid->host_dev = alias;
    } else {
      cudaGetLastError();
    }
  }
  return PYI_OK;
}

int pyi_grid_mark_stale(pyi_grid* grid) {
  if (!grid) return fail(PYI_BAD_ARG, "null grid");
  if (grid->host_src == nullptr) return fail(PYI_BAD_ARG, "pyi_grid_mark_stale: the grid tracks no host array");
  grid->stale.store(1);
  return PYI_OK;
}

void pyi_grid_destroy(pyi_grid* grid) {
  if (!grid) return;
  {
    DeviceGuard guard(grid->device);
    if (grid->host_registered) {
      std::lock_guard<std::mutex> lock(g_reg_mu);
      auto it = g_registered.find(grid->host_src);
      if (it != g_registered.end() && --it->second.second == 0) {
        cudaHostUnregister(const_cast<void*>(grid->host_src));
        cudaGetLastError();
        g_registered.erase(it);
      }
    }
    if (!grid->borrowed) cudaFree(grid->d_data);
    cudaFree(grid->d_packed);
    cudaFree(grid->d_planes);
  }
  for (auto& kv : grid->replicas) {
    if (kv.second) free_replica(kv.first, *kv.second);
  }
  for (auto& a : grid->axes) {
    if (a) free_axis_device(a.get());
  }
  cudaGetLastError();
  delete grid;
}

// ---- hot path ---------------------------------------------------------------
int pyi_geometric(const pyi_grid* grid, const pyi_geometric_config* cfg, int64_t n,
                  const double* x, const double* y, const void* z, const double* u, void* out,
                  int mem, int64_t* oob_index, int* oob_axis, void* stream) {
  int st = require_device();
  if (st != PYI_OK) return st;
  if (!grid || !cfg || n < 0) return fail(PYI_BAD_ARG, "pyi_geometric: null argument");
  if (n > 0 && (!x || !y || !out || (grid->ndim >= 3 && !z) || (grid->ndim >= 4 && !u))) {
    return fail(PYI_BAD_SHAPE, "pyi_geometric: a coordinate array is missing for this grid");
  }
  if (cfg->method < PYI_BILINEAR || cfg->method > PYI_NEAREST) {
    return fail(PYI_BAD_ARG, "pyi_geometric: unknown method");
  }
  if (cfg->value_mode < PYI_VALUES_DEFAULT || cfg->value_mode > PYI_VALUES_TOLERANCE) {
    return fail(PYI_BAD_ARG, "pyi_geometric: unknown value mode");
  }
  pyi::GeoParams p{};
  p.dtype = grid->dtype;
  p.ndim = grid->ndim;
  p.method = cfg->method;
  p.idw_exp = cfg->idw_exp;
  p.z_method = cfg->z_method;
  p.u_method = cfg->u_method;
  p.bounds_error = cfg->bounds_error;
  p.value_mode = cfg->value_mode;
  if (n == 0) return PYI_OK;
  InPlace inplace;
  // 2^ndim corners, a 32-byte sector each; generic kernel: 5-9 Gpts/s
  st = sync_grid(grid, n, 32LL << grid->ndim, mem, inplace, 5e9);
  if (st != PYI_OK) return st;
  const size_t out_elem = result_is_f64(grid->dtype) ? 8 : 4;
  auto bind = [&](pyi::GeoParams& q, const GridView& v) {
    for (int k = 0; k < grid->ndim; ++k) q.ax[k] = v.dax[k];
    q.data = v.data;
    q.packed = v.packed;
  };

  if (mem == PYI_MEM_DEVICE) {
    st = require_grid_device(grid);
    if (st != PYI_OK) return st;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    GridView view;
    grid_view(grid, grid->device, view);
    bind(p, view);
    p.n = n;
    p.x = x;
    p.y = y;
    p.z = z;
    p.u = u;
    p.out = out;
    DeviceKey key_buf;
    if (cfg->bounds_error) {
      st = key_buf.open(s);
      if (st != PYI_OK) return st;
      p.oob_key = key_buf.ptr;
    }
    cudaError_t e = pyi::launch_geometric(p, s);
    if (e != cudaSuccess) return cuda_fail(e, "geometric kernel");
    if (cfg->bounds_error) {
      unsigned long long key = pyi::kNoOob;
      st = key_buf.read(key);
      if (st != PYI_OK) return st;
      return report_oob(key, oob_index, oob_axis);
    }
    return PYI_OK;
  }

  std::vector<Column> cols;
  auto add_in = [&](const void* ptr) {
    Column c;
    c.host_in = ptr;
    cols.push_back(c);
    return cols.size() - 1;
  };
  const size_t cx = add_in(x), cy = add_in(y);
  const size_t cz = grid->ndim >= 3 ? add_in(z) : 0;
  const size_t cu = grid->ndim >= 4 ? add_in(u) : 0;
  Column co;
  co.host_out = out;
  co.elem = out_elem;
  cols.push_back(co);
  const size_t cout_ = cols.size() - 1;
  unsigned long long key = pyi::kNoOob;
  st = run_host_call(grid, n, cols, cfg->bounds_error != 0, key,
                     [&](const GridView& view, unsigned long long* oob_key, cudaStream_t s,
                         long long begin, long long count, char* dev, const std::vector<Column>& c,
                         unsigned* queue) {
                       pyi::GeoParams q = p;
                       bind(q, view);
                       q.n = count;
                       q.x = reinterpret_cast<const double*>(dev + c[cx].dev_off);
                       q.y = reinterpret_cast<const double*>(dev + c[cy].dev_off);
                       if (grid->ndim >= 3) q.z = dev + c[cz].dev_off;
                       if (grid->ndim >= 4) q.u = reinterpret_cast<const double*>(dev + c[cu].dev_off);
                       q.out = dev + c[cout_].dev_off;
                       q.oob_key = oob_key;
                       q.index_base = begin;
                       q.queue = queue;
                       q.queue_cap = pyi::kTolQueue;
                       return pyi::launch_geometric(q, s);
                     });
  if (st != PYI_OK) return st;
  if (cfg->bounds_error) return report_oob(key, oob_index, oob_axis);
  return PYI_OK;
}

int pyi_windowed2d(const pyi_grid* grid, const pyi_windowed_config* cfg, int64_t n,
                   const double* x, const double* y, void* out, int mem, int64_t* oob_index,
                   int* oob_axis, void* stream) {
  int st = require_device();
  if (st != PYI_OK) return st;
  if (!grid || !cfg || n < 0) return fail(PYI_BAD_ARG, "pyi_windowed2d: null argument");
  if (grid->ndim != 2) return fail(PYI_BAD_ARG, "pyi_windowed2d: bivariate requires 2D grid");
  if (cfg->method != PYI_W_BICUBIC && cfg->method != PYI_W_BILINEAR) {
    return fail(PYI_BAD_ARG, "pyi_windowed2d: unknown method");
  }
  // same kernels as the N-D entry (pyi_windowed_method == the first two pyi_spatial_method)
  pyi_windowed_nd_config c{};
  c.method = cfg->method;
  c.half_x = cfg->half_x;
  c.half_y = cfg->half_y;
  c.boundary = cfg->boundary;
  c.bounds_error = cfg->bounds_error;
  c.num_threads = cfg->num_threads;
  c.value_mode = cfg->value_mode;
  return pyi_windowed(grid, &c, n, x, y, nullptr, nullptr, out, mem, oob_index, oob_axis, stream);
}

int pyi_windowed(const pyi_grid* grid, const pyi_windowed_nd_config* cfg, int64_t n,
                 const double* x, const double* y, const void* z, const double* u, void* out,
                 int mem, int64_t* oob_index, int* oob_axis, void* stream) {
  int st = require_device();
  if (st != PYI_OK) return st;
  if (!grid || !cfg || n < 0) return fail(PYI_BAD_ARG, "pyi_windowed: null argument");
  const int nd = grid->ndim;
  if (cfg->half_x < 1 || (nd >= 2 && cfg->half_y < 1)) {
    return fail(PYI_BAD_ARG, "Half window sizes must be at least 1");  // cache.hpp:258-260
  }
  if (cfg->half_x > 16 || (nd >= 2 && cfg->half_y > 16)) {
    return fail(PYI_BAD_ARG, "pyi_windowed: half window sizes above 16 are not supported");
  }
  if (cfg->method < PYI_S_BICUBIC || cfg->method > PYI_S_POLYNOMIAL) {
    return fail(PYI_BAD_ARG, "Unknown fitting model");  // config/windowed.hpp:54-55
  }
  if (nd == 1 && cfg->method < PYI_S_AKIMA) {
    return fail(PYI_BAD_ARG, "pyi_windowed: a 1-D grid takes a univariate fitting model");
  }
  if (cfg->value_mode < PYI_VALUES_DEFAULT || cfg->value_mode > PYI_VALUES_TOLERANCE) {
    return fail(PYI_BAD_ARG, "pyi_windowed: unknown value mode");
  }
  if (cfg->derivative && nd != 1) {
    return fail(PYI_BAD_ARG, "pyi_windowed: derivatives are defined for 1-D grids only");
  }
  if (n > 0 && (!x || !out || (nd >= 2 && !y) || (nd >= 3 && !z) || (nd >= 4 && !u))) {
    return fail(PYI_BAD_SHAPE, "pyi_windowed: missing array");
  }
  if (n == 0) return PYI_OK;
  pyi::WinNdParams p{};
  p.dtype = grid->dtype;
  p.ndim = nd;
  p.method = cfg->method;
  p.half_x = cfg->half_x;
  p.half_y = nd >= 2 ? cfg->half_y : 1;
  p.boundary = cfg->boundary;
  p.bounds_error = cfg->bounds_error;
  p.z_method = cfg->z_method;
  p.u_method = cfg->u_method;
  p.derivative = cfg->derivative;
  p.value_mode = cfg->value_mode;
  for (int a = 0; a < 2; ++a) {
    p.u6_ok[a] = grid->u6_ok[a];
    std::memcpy(p.u6[a], grid->u6[a], sizeof(p.u6[a]));
  }
  InPlace inplace;
  {
    // sectors a point touches in the C-order array: 2-D window rows are contiguous along y; from
    // 3-D on y is strided and every window value is a sector of its own (the z / u pair shares one
    // about half of the time)
    long long touched;
    if (nd <= 2) {
      touched = 2LL * cfg->half_x * std::max<long long>(64, 2LL * (nd == 2 ? cfg->half_y : 1) * 8 + 32);
    } else {
      touched = 2LL * cfg->half_x * 2LL * cfg->half_y * 48 * (nd >= 4 ? 2 : 1);
    }
    // K6 on the plain array: ~1e9 (1-D), 2e8 (3-D), 1e8 (4-D) points/s; 2-D windowed kernels read
    // the plain array anyway
    const double plain_rate = nd == 2 ? 1e30 : (nd == 1 ? 1e9 : (nd == 3 ? 2e8 : 1e8));
    st = sync_grid(grid, n, touched, mem, inplace, plain_rate);
    if (st != PYI_OK) return st;
  }
  // the plane-major copy serves bicubic in every mode and, unless the caller asks for
  // bit-identical values, the other models' 6 x 6 fast pass (pyi_winmodels.cuh)
  const bool want_planes =
      nd >= 3 && cfg->half_x == 3 && cfg->half_y == 3 &&
      (cfg->method == PYI_S_BICUBIC ||
       (cfg->value_mode != PYI_VALUES_EXACT && grid->cw_ok &&
        (grid->dtype == PYI_F64 || grid->dtype == PYI_F32)));
  if (want_planes && !inplace.on) {
    DeviceGuard guard(grid->device);
    if (guard.err != cudaSuccess) return cuda_fail(guard.err, "cudaSetDevice");
    st = grid_planes(grid, false);
    if (st != PYI_OK) return st;
  }
  p.w_ok = grid->w_ok;
  for (int a = 0; a < 2; ++a) std::memcpy(p.w6[a], grid->w6[a], sizeof(p.w6[a]));
  p.cw_ok = grid->cw_ok;
  std::memcpy(p.cw, grid->cw, sizeof(p.cw));
  if (cfg->method != PYI_S_BICUBIC) {
    // the separable stencil also serves the models that are linear in the node
    // values on axes that qualify for it (pyi_windowed.cu): their own cubics
    const bool local = cfg->method == PYI_S_AKIMA || cfg->method == PYI_S_AKIMA_PERIODIC ||
                       cfg->method == PYI_S_STEFFEN;  // no table: pyi_winmodels.cuh
    p.cw_ok = grid->cw_ok && nd >= 2 && cfg->half_x == 3 && cfg->half_y == 3 &&
              (local || cfg->method == PYI_S_POLYNOMIAL ||
               (model_stencil_cubics(cfg->method, p.cw[0]) && model_stencil_cubics(cfg->method, p.cw[1])));
    if (p.cw_ok && cfg->method == PYI_S_POLYNOMIAL) {
      for (int a = 0; a < 2; ++a) polynomial_stencil(grid->dax[a].step, p.pw[a]);
    }
  }
  p.cells = 1;
  for (int a = 0; a < nd; ++a) p.cells *= grid->shape[a];
  const size_t out_elem = result_is_f64(grid->dtype) ? 8 : 4;
  auto bind = [&](pyi::WinNdParams& q, const GridView& v) {
    for (int k = 0; k < nd; ++k) q.ax[k] = v.dax[k];
    q.data = v.data;
    q.planes = want_planes ? v.planes : nullptr;
  };
  if (mem == PYI_MEM_DEVICE) {
    st = require_grid_device(grid);
    if (st != PYI_OK) return st;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    GridView view;
    grid_view(grid, grid->device, view);
    bind(p, view);
    p.n = n;
    p.x = x;
    p.y = y;
    p.z = z;
    p.u = u;
    p.out = out;
    DeviceKey key_buf;
    if (cfg->bounds_error) {
      st = key_buf.open(s);
      if (st != PYI_OK) return st;
      p.oob_key = key_buf.ptr;
    }
    cudaError_t e = pyi::launch_windowed_nd(p, s);
    if (e != cudaSuccess) return cuda_fail(e, "windowed kernel");
    if (cfg->bounds_error) {
      unsigned long long key = pyi::kNoOob;
      st = key_buf.read(key);
      if (st != PYI_OK) return st;
      return report_oob(key, oob_index, oob_axis);
    }
    return PYI_OK;
  }
  std::vector<Column> cols;
  auto add_in = [&](const void* ptr) {
    Column c;
    c.host_in = ptr;
    cols.push_back(c);
    return cols.size() - 1;
  };
  const size_t cx = add_in(x);
  const size_t cy = nd >= 2 ? add_in(y) : 0;
  const size_t cz = nd >= 3 ? add_in(z) : 0;
  const size_t cu = nd >= 4 ? add_in(u) : 0;
  Column co;
  co.host_out = out;
  co.elem = out_elem;
  cols.push_back(co);
  const size_t cout_ = cols.size() - 1;
  unsigned long long key = pyi::kNoOob;
  st = run_host_call(grid, n, cols, cfg->bounds_error != 0, key,
                     [&](const GridView& view, unsigned long long* oob_key, cudaStream_t s,
                         long long begin, long long count, char* dev, const std::vector<Column>& c,
                         unsigned* queue) {
                       pyi::WinNdParams q = p;
                       bind(q, view);
                       q.n = count;
                       q.x = reinterpret_cast<const double*>(dev + c[cx].dev_off);
                       if (nd >= 2) q.y = reinterpret_cast<const double*>(dev + c[cy].dev_off);
                       if (nd >= 3) q.z = dev + c[cz].dev_off;
                       if (nd >= 4) q.u = reinterpret_cast<const double*>(dev + c[cu].dev_off);
                       q.out = dev + c[cout_].dev_off;
                       q.oob_key = oob_key;
                       q.index_base = begin;
                       q.queue = queue;
                       q.queue_cap = pyi::kTolQueue;
                       return pyi::launch_windowed_nd(q, s);
                     });
  if (st != PYI_OK) return st;
  if (cfg->bounds_error) return report_oob(key, oob_index, oob_axis);
  return PYI_OK;
}

int pyi_datetime_convert(int src, int dst, int64_t n, const int64_t* in, int64_t* out, int mem,
                         int64_t* overflow_index, int* overflow_op, void* stream) {
  int st = require_device();
  if (st != PYI_OK) return st;
  if (src < PYI_RES_YEAR || src > PYI_RES_ATTOSECOND || dst < PYI_RES_YEAR ||
      dst > PYI_RES_ATTOSECOND || n < 0) {
    return fail(PYI_BAD_ARG, "pyi_datetime_convert: unknown resolution");
  }
  if (n > 0 && (!in || !out)) return fail(PYI_BAD_SHAPE, "pyi_datetime_convert: missing array");
  if (n == 0) return PYI_OK;
  auto report = [&](unsigned long long key) {
    if (key == pyi::kNoOob) return static_cast<int>(PYI_OK);
    if (overflow_index) *overflow_index = static_cast<int64_t>(key >> 2);
    if (overflow_op) *overflow_op = static_cast<int>(key & 3);
    // dateutils.hpp:55,99
    return fail(PYI_OVERFLOW, (key & 3) == 1 ? "integer overflow in multiplication"
                                              : "integer overflow in addition");
  };
  if (mem == PYI_MEM_DEVICE) {
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (src == dst) {  // dateutils.hpp:824-826
      if (in != out) {
        PYI_CUDA(cudaMemcpyAsync(out, in, static_cast<size_t>(n) * 8, cudaMemcpyDeviceToDevice, s));
      }
      return PYI_OK;
    }
    DeviceKey key_buf;
    st = key_buf.open(s);
    if (st != PYI_OK) return st;
    cudaError_t e = pyi::launch_datetime_convert(reinterpret_cast<const long long*>(in),
                                                 reinterpret_cast<long long*>(out), n, 0, src, dst,
                                                 key_buf.ptr, s);
    if (e != cudaSuccess) return cuda_fail(e, "datetime kernel");
    unsigned long long key = pyi::kNoOob;
    st = key_buf.read(key);
    if (st != PYI_OK) return st;
    return report(key);
  }
  if (src == dst) {
    if (in != out) std::memcpy(out, in, static_cast<size_t>(n) * 8);
    return PYI_OK;
  }
  std::vector<Column> cols(2);
  cols[0].host_in = in;
  cols[1].host_out = out;
  int dev = 0;
  PYI_CUDA(cudaGetDevice(&dev));
  Lease lease(dev);
  Workspace& ws = *lease.ws;
  st = run_pipeline(ws, n, cols, true, [&](cudaStream_t s, long long begin, long long count, char* d, unsigned*) {
    return pyi::launch_datetime_convert(reinterpret_cast<const long long*>(d + cols[0].dev_off),
                                        reinterpret_cast<long long*>(d + cols[1].dev_off), count,
                                        begin, src, dst, ws.oob_key, s);
  });
  if (st != PYI_OK) return st;
  unsigned long long key = pyi::kNoOob;
  st = read_key(ws, key);
  if (st != PYI_OK) return st;
  return report(key);
}

// fill::gauss_seidel (fill/gauss_seidel.hpp:166-211) and fill::multigrid
// (fill/multi_grid.hpp:276-334) in place.  Host grids are staged through one
// device copy; everything else happens in pyi_fill.cu.
namespace {

int fill_entry(bool multigrid, void* grid, int dtype, int64_t nx, int64_t ny,
               const pyi_fill_config* cfg, int mem, uint64_t* iterations, double* max_residual,
               void* stream) {
  const char* name = multigrid ? "pyi_fill_multigrid" : "pyi_fill_gauss_seidel";
  int st = require_device();
  if (st != PYI_OK) return st;
  if (!cfg || !iterations || !max_residual) {
    return fail(PYI_BAD_ARG, std::string(name) + ": missing argument");
  }
  if (dtype != PYI_F32 && dtype != PYI_F64) {
    return fail(PYI_BAD_ARG, std::string(name) + ": grid must be float32 or float64");
  }
  if (nx < 0 || ny < 0 || (nx > 0 && ny > int64_t{0x7fffffff} / nx)) {
    return fail(PYI_BAD_SHAPE, std::string(name) + ": invalid grid shape (at most 2^31 - 1 cells)");
  }
  if (cfg->first_guess != PYI_FIRST_GUESS_ZERO &&
      cfg->first_guess != PYI_FIRST_GUESS_ZONAL_AVERAGE) {
    // gauss_seidel.hpp:190-193, multi_grid.hpp:296-299
    return fail(PYI_BAD_ARG,
                "Unsupported first guess method: " + std::to_string(cfg->first_guess));
  }
  *iterations = 0;
  *max_residual = 0.0;
  if (nx == 0 || ny == 0) return PYI_OK;
  if (!grid) return fail(PYI_BAD_SHAPE, std::string(name) + ": missing grid");
  unsigned long long its = 0;
  double res = 0.0;
  auto run = [&](void* dev, cudaStream_t s) {
    return multigrid
               ? pyi::fill_multigrid(dev, dtype, nx, ny, cfg->first_guess, cfg->is_periodic != 0,
                                     cfg->max_iterations, cfg->epsilon, cfg->pre_smooth,
                                     cfg->post_smooth, &its, &res, s)
               : pyi::fill_gauss_seidel(dev, dtype, nx, ny, cfg->first_guess,
                                        cfg->is_periodic != 0, cfg->max_iterations, cfg->epsilon,
                                        cfg->relaxation, &its, &res, s);
  };
  if (mem == PYI_MEM_DEVICE) {
    cudaError_t e = run(grid, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return cuda_fail(e, "fill kernels");
    *iterations = its;
    *max_residual = res;
    return PYI_OK;
  }
  const size_t bytes = static_cast<size_t>(nx) * static_cast<size_t>(ny) * dtype_size(dtype);
  int dev = 0;
  PYI_CUDA(cudaGetDevice(&dev));
  Lease lease(dev);
  Workspace& ws = *lease.ws;
  st = ws.ensure(bytes, 0);
  if (st != PYI_OK) return st;
  Slot& slot = ws.slot[0];
  slot.busy = true;
  PYI_CUDA(cudaMemcpyAsync(slot.dev, grid, bytes, cudaMemcpyHostToDevice, slot.stream));
  cudaError_t e = run(slot.dev, slot.stream);
  if (e != cudaSuccess) return cuda_fail(e, "fill kernels");
  // always copied back: with max_iterations == 0 the reference still leaves
  // the first guess in the undefined cells (gauss_seidel.hpp:184-199)
  PYI_CUDA(cudaMemcpyAsync(grid, slot.dev, bytes, cudaMemcpyDeviceToHost, slot.stream));
  PYI_CUDA(cudaStreamSynchronize(slot.stream));
  slot.busy = false;
  *iterations = its;
  *max_residual = res;
  return PYI_OK;
}

}  // namespace

int pyi_fill_gauss_seidel(void* grid, int dtype, int64_t nx, int64_t ny,
                          const pyi_fill_config* cfg, int mem, uint64_t* iterations,
                          double* max_residual, void* stream) {
  return fill_entry(false, grid, dtype, nx, ny, cfg, mem, iterations, max_residual, stream);
}

int pyi_fill_multigrid(void* grid, int dtype, int64_t nx, int64_t ny, const pyi_fill_config* cfg,
                       int mem, uint64_t* iterations, double* max_residual, void* stream) {
  return fill_entry(true, grid, dtype, nx, ny, cfg, mem, iterations, max_residual, stream);
}

// fill::loess (fill/loess.hpp:207-253).  Host grids are staged through two
// device copies; everything else happens in pyi_loess.cu.
int pyi_fill_loess(const void* grid, void* result, int dtype, int64_t nx, int64_t ny,
                   const pyi_loess_config* cfg, int mem, void* stream) {
  int st = require_device();
  if (st != PYI_OK) return st;
  if (!cfg) return fail(PYI_BAD_ARG, "pyi_fill_loess: missing argument");
  if (dtype != PYI_F32 && dtype != PYI_F64) {
    return fail(PYI_BAD_ARG, "pyi_fill_loess: grid must be float32 or float64");
  }
  if (nx < 0 || ny < 0 || (nx > 0 && ny > int64_t{0x7fffffff} / nx)) {
    return fail(PYI_BAD_SHAPE, "pyi_fill_loess: invalid grid shape (at most 2^31 - 1 cells)");
  }
  if (cfg->half_x < 1 || cfg->half_y < 1) {
    return fail(PYI_BAD_ARG, "nx and ny must be at least 1");  // fill/helpers.hpp:117-124
  }
  if (cfg->first_guess != PYI_FIRST_GUESS_ZERO && cfg->first_guess != PYI_FIRST_GUESS_ZONAL_AVERAGE) {
    return fail(PYI_BAD_ARG, "Unsupported first guess method: " + std::to_string(cfg->first_guess));
  }
  if (cfg->value_type < PYI_LOESS_UNDEFINED || cfg->value_type > PYI_LOESS_ALL) {
    return fail(PYI_BAD_ARG, "pyi_fill_loess: unknown value type");
  }
  if (nx == 0 || ny == 0) return PYI_OK;
  if (!grid || !result) return fail(PYI_BAD_SHAPE, "pyi_fill_loess: missing grid");
  auto run = [&](const void* in, void* out, cudaStream_t s) {
    return pyi::fill_loess(in, out, dtype, nx, ny, cfg->half_x, cfg->half_y, cfg->value_type,
                           cfg->first_guess, cfg->is_periodic != 0, cfg->max_iterations, cfg->epsilon, s);
  };
  int dev = 0;
  PYI_CUDA(cudaGetDevice(&dev));
  keep_pool_cached(dev);
  if (mem == PYI_MEM_DEVICE) {
    cudaError_t e = run(grid, result, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return cuda_fail(e, "loess kernels");
    return PYI_OK;
  }
  const size_t bytes = static_cast<size_t>(nx) * static_cast<size_t>(ny) * dtype_size(dtype);
  Lease lease(dev);
  Workspace& ws = *lease.ws;
  st = ws.ensure(2 * ((bytes + 255) & ~static_cast<size_t>(255)), 0);
  if (st != PYI_OK) return st;
  Slot& slot = ws.slot[0];
  slot.busy = true;
  char* d_in = slot.dev;
  char* d_out = slot.dev + ((bytes + 255) & ~static_cast<size_t>(255));
  PYI_CUDA(cudaMemcpyAsync(d_in, grid, bytes, cudaMemcpyHostToDevice, slot.stream));
  cudaError_t e = run(d_in, d_out, slot.stream);
  if (e != cudaSuccess) return cuda_fail(e, "loess kernels");
  PYI_CUDA(cudaMemcpyAsync(result, d_out, bytes, cudaMemcpyDeviceToHost, slot.stream));
  PYI_CUDA(cudaStreamSynchronize(slot.stream));
  slot.busy = false;
  return PYI_OK;
}

// pyi_fill_loess on device pointers with the scratch in a static device array
// (pyi_loess.cu): no allocation of any kind on this path -- no workspace, no
// memory pool, no cudaMallocAsync.  No counterpart in the reference.
int pyi_fill_loess_static(const void* grid, void* result, int dtype, int64_t nx, int64_t ny,
                          const pyi_loess_config* cfg, void* stream) {
  int st = require_device();
  if (st != PYI_OK) return st;
  if (!cfg) return fail(PYI_BAD_ARG, "pyi_fill_loess_static: missing argument");
  if (dtype != PYI_F32 && dtype != PYI_F64) {
    return fail(PYI_BAD_ARG, "pyi_fill_loess_static: grid must be float32 or float64");
  }
  if (nx < 0 || ny < 0 || (nx > 0 && ny > int64_t{0x7fffffff} / nx)) {
    return fail(PYI_BAD_SHAPE, "pyi_fill_loess_static: invalid grid shape (at most 2^31 - 1 cells)");
  }
  const size_t bytes = static_cast<size_t>(nx) * static_cast<size_t>(ny) * dtype_size(dtype);
  if (bytes > PYI_LOESS_STATIC_MAX_BYTES) {
    return fail(PYI_BAD_SHAPE, "pyi_fill_loess_static: grid of " + std::to_string(bytes) +
                                   " bytes exceeds the static scratch of " +
                                   std::to_string(PYI_LOESS_STATIC_MAX_BYTES) +
                                   " bytes (use pyi_fill_loess)");
  }
  if (cfg->half_x < 1 || cfg->half_y < 1) {
    return fail(PYI_BAD_ARG, "nx and ny must be at least 1");  // fill/helpers.hpp:117-124
  }
  if (cfg->first_guess != PYI_FIRST_GUESS_ZERO && cfg->first_guess != PYI_FIRST_GUESS_ZONAL_AVERAGE) {
    return fail(PYI_BAD_ARG, "Unsupported first guess method: " + std::to_string(cfg->first_guess));
  }
  if (cfg->value_type < PYI_LOESS_UNDEFINED || cfg->value_type > PYI_LOESS_ALL) {
    return fail(PYI_BAD_ARG, "pyi_fill_loess_static: unknown value type");
  }
  if (nx == 0 || ny == 0) return PYI_OK;
  if (!grid || !result) return fail(PYI_BAD_SHAPE, "pyi_fill_loess_static: missing grid");
  cudaError_t e = pyi::fill_loess_static(grid, result, dtype, nx, ny, cfg->half_x, cfg->half_y,
                                         cfg->value_type, cfg->first_guess, cfg->is_periodic != 0,
                                         cfg->max_iterations, cfg->epsilon, static_cast<cudaStream_t>(stream));
  if (e != cudaSuccess) return cuda_fail(e, "loess kernels (static scratch)");
  return PYI_OK;
}

// TemporalAxis::coordinate_repr (math/temporal_axis.hpp:80-83,150-156) =
// dateutils::datetime64_to_string / timedelta64_to_string
// (dateutils.hpp:729-812): ISO-8601 text of an int64 in a numpy resolution.
namespace {

struct CivilDate {
  long long year;
  unsigned month, day;
};

// proleptic Gregorian date of a day count since 1970-01-01 (the era / day-of-era
// decomposition the reference uses, dateutils.hpp:563-584)
CivilDate civil_from_days(long long days) {
  days += 719468;
  const long long era = (days >= 0 ? days : days - 146096) / 146097;
  const unsigned doe = static_cast<unsigned>(days - era * 146097);
  const unsigned yoe = (doe - doe / 1460 + doe / 36524 - doe / 146096) / 365;
  const unsigned doy = doe - (365 * yoe + yoe / 4 - yoe / 100);
  const unsigned mp = (5 * doy + 2) / 153;
  const unsigned day = doy - (153 * mp + 2) / 5 + 1;
  const unsigned month = mp < 10 ? mp + 3 : mp - 9;
  // the reference narrows the year to int (dateutils.hpp:580-582)
  const int year = (static_cast<int>(yoe) + static_cast<int>(era) * 400) + (month <= 2 ? 1 : 0);
  return {year, month, day};
}

// floor division / modulo by a positive divisor
void floor_divmod(long long v, long long d, long long& q, long long& r) {
  q = v / d;
  r = v % d;
  if (r < 0) {
    r += d;
    q -= 1;
  }
}

std::string two(unsigned v) {
  char b[16];
  std::snprintf(b, sizeof(b), "%02u", v);
  return b;
}

std::string date_text(const CivilDate& d) {
  return std::to_string(d.year) + "-" + two(d.month) + "-" + two(d.day);
}

std::string temporal_text(bool timedelta, int res, long long value) {
  static const char* plural[] = {"years",        "months",       "weeks",       "days",        "hours",
                                 "minutes",      "seconds",      "milliseconds", "microseconds", "nanoseconds",
                                 "picoseconds",  "femtoseconds", "attoseconds"};
  if (timedelta) return std::to_string(value) + " " + plural[res];
  long long q = 0, r = 0;
  switch (res) {
    case PYI_RES_YEAR:
      return std::to_string(static_cast<int>(1970 + value));
    case PYI_RES_MONTH: {
      floor_divmod(value, 12, q, r);
      return std::to_string(static_cast<int>(1970 + q)) + "-" + two(static_cast<unsigned>(r + 1));
    }
    case PYI_RES_WEEK:
      return date_text(civil_from_days(value * 7));
    case PYI_RES_DAY:
      return date_text(civil_from_days(value));
    case PYI_RES_HOUR: {
      floor_divmod(value, 24, q, r);
      return date_text(civil_from_days(q)) + "T" + two(static_cast<unsigned>(r));
    }
    case PYI_RES_MINUTE: {
      floor_divmod(value, 1440, q, r);
      return date_text(civil_from_days(q)) + "T" + two(static_cast<unsigned>(r / 60)) + ":" +
             two(static_cast<unsigned>(r % 60));
    }
    default: {
      // seconds and below: value = seconds * 10^digits + fraction (floored)
      const int digits = 3 * (res - PYI_RES_SECOND);
      long long magnitude = 1;
      for (int k = 0; k < digits; ++k) magnitude *= 10;
      long long sec = 0, frac = 0, days = 0, in_day = 0;
      floor_divmod(value, magnitude, sec, frac);
      floor_divmod(sec, 86400, days, in_day);
      std::string out = date_text(civil_from_days(days)) + "T" + two(static_cast<unsigned>(in_day / 3600)) +
                        ":" + two(static_cast<unsigned>(in_day % 3600 / 60)) + ":" +
                        two(static_cast<unsigned>(in_day % 60));
      if (digits > 0) {
        char b[32];
        std::snprintf(b, sizeof(b), "%0*lld", digits, frac);
        out += std::string(".") + b;
      }
      return out;
    }
  }
}

}  // namespace

int pyi_model_stencil(int method, double step, double* out36) {
  if (!out36 || !(step > 0.0)) return fail(PYI_BAD_ARG, "pyi_model_stencil: null output or non-positive step");
  for (int k = 0; k < 36; ++k) out36[k] = 0.0;
  if (method == PYI_S_POLYNOMIAL) {
    double pw[6][6];
    polynomial_stencil(step, pw);
    for (int j = 0; j < 6; ++j) {
      for (int d = 0; d < 6; ++d) out36[6 * j + d] = pw[j][d];
    }
    return PYI_OK;
  }
  double cw[8][4];
  if (!model_stencil_cubics(method, cw)) {
    return fail(PYI_BAD_ARG, "pyi_model_stencil: the model is not linear in the node values");
  }
  for (int j = 0; j < 6; ++j) {
    for (int d = 0; d < 4; ++d) out36[6 * j + d] = cw[j + 1][d];
  }
  return PYI_OK;
}

int pyi_format_temporal(int is_timedelta, int resolution, int64_t value, char* buf, size_t len) {
  if (!buf || len == 0 || resolution < PYI_RES_YEAR || resolution > PYI_RES_ATTOSECOND) return -1;
  const std::string s = temporal_text(is_timedelta != 0, resolution, value);
  std::snprintf(buf, len, "%s", s.c_str());
  return static_cast<int>(s.size());
}

// "{coord} is out of bounds for axis {label} [{min}, ..., {max}]"
// (pybind/grid.hpp:362-371).  Doubles are printed like std::format("{}"):
// shortest round-trip digits via std::to_chars.  int64 coordinates are printed
// as integers; the TemporalAxis mirror of the Python layer re-formats them as
// ISO dates like the reference's coordinate_repr (math/temporal_axis.hpp:80-83).
int pyi_format_bounds_error(const pyi_grid* grid, int axis, const void* coord, char* buf,
                            size_t len) {
  if (!grid || axis < 0 || axis >= grid->ndim || !coord || !buf || len == 0) return -1;
  static const char* labels[4] = {"x", "y", "z", "u"};
  const pyi_axis_desc& d = grid->axes[axis]->d;
  auto f64 = [](double v) {
    char tmp[64];
    auto r = std::to_chars(tmp, tmp + sizeof(tmp), v);
    return std::string(tmp, r.ptr);
  };
  std::string s;
  if (d.coord_type == PYI_COORD_I64) {
    s = std::to_string(*static_cast<const int64_t*>(coord)) + " is out of bounds for axis " +
        labels[axis] + " [" + std::to_string(d.min_value.i) + ", ..., " +
        std::to_string(d.max_value.i) + "]";
  } else {
    s = f64(*static_cast<const double*>(coord)) + " is out of bounds for axis " + labels[axis] +
        " [" + f64(d.min_value.f) + ", ..., " + f64(d.max_value.f) + "]";
  }
  std::snprintf(buf, len, "%s", s.c_str());
  return static_cast<int>(s.size());
}

}  // extern "C"
