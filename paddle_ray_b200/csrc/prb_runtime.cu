// Error state, workspace, and the thin CUDA-runtime wrappers the torch-free Python host uses.
#include <mutex>
#include <unordered_map>
#include <vector>

#include "prb_common.cuh"

namespace prb {

static thread_local std::string t_last_error;
std::atomic<int64_t> g_launch_count{0};

void set_error(const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  t_last_error = buf;
}

int32_t require_device() {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n <= 0) {
    cudaGetLastError();
    PRB_FAIL(PRB_EXTERNAL,
             "no usable CUDA device (%s); paddle_ray_b200 has no CPU fallback",
             e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
  }
  return PRB_SUCCESS;
}

int sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

bool pdl_enabled() {
  static const bool v = [] {
    const char* e = getenv("PRB_PDL");
    return !(e != nullptr && atoi(e) == 0);
  }();
  return v;
}

namespace {
struct Slot {
  void* ptr = nullptr;
  size_t bytes = 0;
};
struct WsKey {
  int dev;
  cudaStream_t stream;
  bool operator==(const WsKey& o) const { return dev == o.dev && stream == o.stream; }
};
struct WsKeyHash {
  size_t operator()(const WsKey& k) const { return std::hash<void*>()((void*)k.stream) ^ ((size_t)k.dev * 0x9E3779B97F4A7C15ull); }
};
std::mutex g_ws_mu;
std::unordered_map<WsKey, std::vector<Slot>, WsKeyHash> g_ws;  // (device, stream) -> slots
}  // namespace

int32_t workspace(size_t bytes, void** out, int slot, cudaStream_t stream) {
  int dev = 0;
  PRB_CUDA(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> lk(g_ws_mu);
  auto& slots = g_ws[WsKey{dev, stream}];
  if ((int)slots.size() <= slot) slots.resize(slot + 1);
  Slot& s = slots[slot];
  if (s.bytes < bytes) {
    if (s.ptr) PRB_CUDA(cudaFree(s.ptr));  // implicit device sync: nothing in flight still uses it
    s.ptr = nullptr;
    s.bytes = 0;
    size_t want = bytes < 4096 ? 4096 : bytes + bytes / 4;
    PRB_CUDA(cudaMalloc(&s.ptr, want));
    s.bytes = want;
  }
  *out = s.ptr;
  return PRB_SUCCESS;
}

int32_t ticket_counter(void** out, cudaStream_t stream) {
  int dev = 0;
  PRB_CUDA(cudaGetDevice(&dev));
  static std::mutex mu;
  static std::unordered_map<WsKey, void*, WsKeyHash> counters;
  std::lock_guard<std::mutex> lk(mu);
  const WsKey key{dev, stream};
  auto it = counters.find(key);
  if (it == counters.end()) {
    void* p = nullptr;
    PRB_CUDA(cudaMalloc(&p, 256));
    PRB_CUDA(cudaMemset(p, 0, 256));  // synchronous: complete before the first launch on any stream
    it = counters.emplace(key, p).first;
  }
  *out = it->second;
  return PRB_SUCCESS;
}

// ---------------------------------------------------------------- per-kernel timing
std::atomic<int> g_kernel_timing{0};
namespace {
struct TimedLaunch {
  int kind;
  cudaEvent_t start, stop;
};
std::mutex g_kt_mu;
std::vector<TimedLaunch> g_kt;
}  // namespace

void KernelTimer::begin(int kind, cudaStream_t stream) {
  kind_ = kind;
  stream_ = stream;
  if (cudaEventCreate(&start_) != cudaSuccess) {
    start_ = nullptr;
    return;
  }
  cudaEventRecord(start_, stream);
}
void KernelTimer::end() {
  cudaEvent_t stop = nullptr;
  if (cudaEventCreate(&stop) != cudaSuccess) {
    cudaEventDestroy(start_);
    return;
  }
  cudaEventRecord(stop, stream_);
  std::lock_guard<std::mutex> lk(g_kt_mu);
  g_kt.push_back(TimedLaunch{kind_, start_, stop});
}

__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

__global__ void spin_kernel(unsigned long long ns) {
  const unsigned long long t0 = global_ns();
  while (global_ns() - t0 < ns) {
  }
}

__global__ void flag_or_kernel(bool* dst, const bool* src) {
  if (*src) *dst = true;
}

__global__ void flush_kernel(float4* buf, size_t n) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    buf[i] = make_float4(0.f, 0.f, 0.f, 0.f);
}

}  // namespace prb

using namespace prb;

extern "C" {

const char* prb_last_error_message(void) { return t_last_error.c_str(); }
const char* prb_version(void) { return "paddle_ray_b200 0.1 (sm_100a)"; }

int32_t prb_device_count(int32_t* out) {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess) {
    cudaGetLastError();
    n = 0;
  }
  *out = n;
  return PRB_SUCCESS;
}

int32_t prb_set_device(int32_t device) {
  PRB_TRY(require_device());
  PRB_CUDA(cudaSetDevice(device));
  return PRB_SUCCESS;
}

int32_t prb_device_info(int32_t device, int32_t* sm, int64_t* l2_bytes, int64_t* total_mem, char name[128]) {
  PRB_TRY(require_device());
  cudaDeviceProp p;
  PRB_CUDA(cudaGetDeviceProperties(&p, device));
  if (sm) *sm = p.multiProcessorCount;
  if (l2_bytes) *l2_bytes = p.l2CacheSize;
  if (total_mem) *total_mem = (int64_t)p.totalGlobalMem;
  if (name) snprintf(name, 128, "%s", p.name);
  return PRB_SUCCESS;
}

int32_t prb_malloc(int64_t bytes, void** out) {
  PRB_TRY(require_device());
  PRB_ENFORCE(bytes >= 0, PRB_INVALID_ARGUMENT, "prb_malloc: negative size");
  *out = nullptr;
  if (bytes == 0) bytes = 256;  // keep zero-length tensors addressable
  PRB_CUDA(cudaMalloc(out, (size_t)bytes));
  return PRB_SUCCESS;
}
int32_t prb_free(void* ptr) {
  if (ptr) PRB_CUDA(cudaFree(ptr));
  return PRB_SUCCESS;
}
int32_t prb_host_alloc(int64_t bytes, void** out) {
  PRB_TRY(require_device());
  *out = nullptr;
  PRB_CUDA(cudaMallocHost(out, (size_t)(bytes > 0 ? bytes : 1)));
  return PRB_SUCCESS;
}
int32_t prb_host_free(void* ptr) {
  if (ptr) PRB_CUDA(cudaFreeHost(ptr));
  return PRB_SUCCESS;
}
int32_t prb_memcpy_h2d(void* dst, const void* src, int64_t bytes, void* stream) {
  if (bytes > 0) PRB_CUDA(cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyHostToDevice, (cudaStream_t)stream));
  return PRB_SUCCESS;
}
int32_t prb_memcpy_d2h(void* dst, const void* src, int64_t bytes, void* stream) {
  if (bytes > 0) PRB_CUDA(cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  return PRB_SUCCESS;
}
int32_t prb_memcpy_d2d(void* dst, const void* src, int64_t bytes, void* stream) {
  if (bytes > 0) PRB_CUDA(cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  return PRB_SUCCESS;
}
int32_t prb_memset(void* dst, int32_t value, int64_t bytes, void* stream) {
  if (bytes > 0) PRB_CUDA(cudaMemsetAsync(dst, value, (size_t)bytes, (cudaStream_t)stream));
  return PRB_SUCCESS;
}
int32_t prb_stream_create(int32_t high_priority, void** out) {
  PRB_TRY(require_device());
  cudaStream_t s;
  int lo = 0, hi = 0;
  PRB_CUDA(cudaDeviceGetStreamPriorityRange(&lo, &hi));
  PRB_CUDA(cudaStreamCreateWithPriority(&s, cudaStreamNonBlocking, high_priority ? hi : lo));
  *out = (void*)s;
  return PRB_SUCCESS;
}
int32_t prb_stream_destroy(void* stream) {
  PRB_CUDA(cudaStreamDestroy((cudaStream_t)stream));
  return PRB_SUCCESS;
}
int32_t prb_stream_synchronize(void* stream) {
  PRB_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  return PRB_SUCCESS;
}
int32_t prb_device_synchronize(void) {
  PRB_CUDA(cudaDeviceSynchronize());
  return PRB_SUCCESS;
}
int32_t prb_event_create(void** out) {
  PRB_TRY(require_device());
  cudaEvent_t e;
  PRB_CUDA(cudaEventCreate(&e));
  *out = (void*)e;
  return PRB_SUCCESS;
}
int32_t prb_event_destroy(void* event) {
  PRB_CUDA(cudaEventDestroy((cudaEvent_t)event));
  return PRB_SUCCESS;
}
int32_t prb_event_record(void* event, void* stream) {
  PRB_CUDA(cudaEventRecord((cudaEvent_t)event, (cudaStream_t)stream));
  return PRB_SUCCESS;
}
int32_t prb_event_synchronize(void* event) {
  PRB_CUDA(cudaEventSynchronize((cudaEvent_t)event));
  return PRB_SUCCESS;
}
int32_t prb_event_elapsed_ms(void* start, void* stop, float* out_ms) {
  PRB_CUDA(cudaEventElapsedTime(out_ms, (cudaEvent_t)start, (cudaEvent_t)stop));
  return PRB_SUCCESS;
}
int32_t prb_stream_wait_event(void* stream, void* event) {
  PRB_CUDA(cudaStreamWaitEvent((cudaStream_t)stream, (cudaEvent_t)event, 0));
  return PRB_SUCCESS;
}

int32_t prb_flag_or(void* dst, const void* src, void* stream) {
  PRB_TRY(require_device());
  PRB_ENFORCE(dst && src, PRB_INVALID_ARGUMENT, "flag_or: NULL flag");
  flag_or_kernel<<<1, 1, 0, (cudaStream_t)stream>>>((bool*)dst, (const bool*)src);
  PRB_LAUNCH_CHECK();
  return PRB_SUCCESS;
}

int32_t prb_spin(int64_t microseconds, void* stream) {
  PRB_TRY(require_device());
  spin_kernel<<<1, 1, 0, (cudaStream_t)stream>>>((unsigned long long)(microseconds > 0 ? microseconds : 0) * 1000ull);
  PRB_CUDA(cudaGetLastError());  // bench scaffolding: not counted as a hot-path launch
  return PRB_SUCCESS;
}

int32_t prb_flush_l2(void* stream) {
  PRB_TRY(require_device());
  const size_t bytes = (size_t)256 << 20;  // 256 MiB > 126 MB L2
  void* buf = nullptr;
  PRB_TRY(workspace(bytes, &buf, /*slot=*/7, nullptr));
  flush_kernel<<<sm_count() * 8, 256, 0, (cudaStream_t)stream>>>((float4*)buf, bytes / sizeof(float4));
  PRB_CUDA(cudaGetLastError());  // bench scaffolding: not counted as a hot-path launch
  return PRB_SUCCESS;
}

int64_t prb_launch_count(void) { return g_launch_count.load(); }

int32_t prb_kernel_timing(int32_t enable) {
  {
    std::lock_guard<std::mutex> lk(g_kt_mu);
    if (enable) {
      for (TimedLaunch& t : g_kt) {
        cudaEventDestroy(t.start);
        cudaEventDestroy(t.stop);
      }
      g_kt.clear();
    }
  }
  g_kernel_timing.store(enable ? 1 : 0);
  return PRB_SUCCESS;
}

int32_t prb_kernel_timing_read(int32_t kind, double* total_ms, int64_t* launches) {
  PRB_ENFORCE(kind >= 0 && kind < KK_COUNT && total_ms && launches, PRB_INVALID_ARGUMENT, "kernel_timing_read: bad arguments");
  PRB_CUDA(cudaDeviceSynchronize());
  std::lock_guard<std::mutex> lk(g_kt_mu);
  double ms = 0.0;
  int64_t n = 0;
  for (const TimedLaunch& t : g_kt) {
    if (t.kind != kind) continue;
    float e = 0.f;
    PRB_CUDA(cudaEventElapsedTime(&e, t.start, t.stop));
    ms += e;
    ++n;
  }
  *total_ms = ms;
  *launches = n;
  return PRB_SUCCESS;
}

const char* prb_kernel_kind_name(int32_t kind) {
  static const char* names[KK_COUNT] = {"map_chunks_kernel", "grad_pass_kernel", "grad_pass_kernel<pack>",
                                        "norm_finalize_kernel", "adam_chunks_kernel", "nvl_exchange_kernel"};
  return kind >= 0 && kind < KK_COUNT ? names[kind] : "?";
}

}  // extern "C"
