// Shared device/host helpers for the step-tail kernels (sm_100a only).
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <string>

#include "../../include/paddle_ray_b200.h"

namespace prb {

// ---------------------------------------------------------------- errors
void set_error(const char* fmt, ...);
extern std::atomic<int64_t> g_launch_count;

#define PRB_FAIL(code, ...)         \
  do {                              \
    ::prb::set_error(__VA_ARGS__);  \
    return (code);                  \
  } while (0)

#define PRB_ENFORCE(cond, code, ...) \
  do {                               \
    if (!(cond)) PRB_FAIL(code, __VA_ARGS__); \
  } while (0)

#define PRB_CUDA(expr)                                                                        \
  do {                                                                                        \
    cudaError_t _e = (expr);                                                                  \
    if (_e != cudaSuccess)                                                                    \
      PRB_FAIL(PRB_EXTERNAL, "CUDA error %s at %s:%d: %s", cudaGetErrorName(_e), __FILE__,    \
               __LINE__, cudaGetErrorString(_e));                                             \
  } while (0)

#define PRB_TRY(expr)                     \
  do {                                    \
    int32_t _s = (expr);                  \
    if (_s != PRB_SUCCESS) return _s;     \
  } while (0)

#define PRB_LAUNCH_CHECK()          \
  do {                              \
    ::prb::g_launch_count.fetch_add(1, std::memory_order_relaxed); \
    PRB_CUDA(cudaGetLastError());   \
  } while (0)

inline int dtype_size(int32_t dt) {
  switch (dt) {
    case PRB_BOOL: return 1;
    case PRB_INT32: return 4;
    case PRB_INT64: return 8;
    case PRB_FLOAT32: return 4;
    case PRB_FLOAT64: return 8;
    case PRB_FLOAT16: return 2;
    case PRB_BFLOAT16: return 2;
    default: return 0;
  }
}

inline const char* dtype_name(int32_t dt) {
  switch (dt) {
    case PRB_BOOL: return "bool";
    case PRB_INT32: return "int32";
    case PRB_INT64: return "int64";
    case PRB_FLOAT32: return "float32";
    case PRB_FLOAT64: return "float64";
    case PRB_FLOAT16: return "float16";
    case PRB_BFLOAT16: return "bfloat16";
    default: return "unknown";
  }
}

// ---------------------------------------------------------------- per-kernel timing (measurement aid)
// prb_kernel_timing(1) makes every launch site bracket its kernel with two CUDA events on the launching stream;
// prb_kernel_timing_read sums the elapsed times per kind.  Off (one relaxed load per launch) by default.
enum KernelKind {
  KK_MAP = 0,        // map_chunks_kernel: pack(+scale) without epilogue, unpack, scale, concat, split
  KK_GRAD_PASS,      // grad_pass_kernel modes 0/1: a pass of its own over the gradients
  KK_PACK_EPILOGUE,  // grad_pass_kernel mode 2: pack + scale + inf check + norm partials
  KK_NORM_FINALIZE,  // norm_finalize_kernel
  KK_ADAM,           // adam_chunks_kernel
  KK_EXCHANGE,       // nvl_exchange_kernel / nvls_allreduce_kernel
  KK_COUNT
};
extern std::atomic<int> g_kernel_timing;
struct KernelTimer {
  KernelTimer(int kind, cudaStream_t stream) {
    if (g_kernel_timing.load(std::memory_order_relaxed)) begin(kind, stream);
  }
  ~KernelTimer() {
    if (start_) end();
  }
  void begin(int kind, cudaStream_t stream);
  void end();
  int kind_ = 0;
  cudaStream_t stream_ = nullptr;
  cudaEvent_t start_ = nullptr;
};

int32_t require_device();  // PRB_EXTERNAL if no usable CUDA device (no CPU fallback exists)
int sm_count();            // of the current device (cached)

// ---------------------------------------------------------------- workspace
// Library-owned persistent device scratch, grown on demand (never shrinks); one per (device, stream, slot), so
// two optimizers that step on different streams of one device never share partial sums.
int32_t workspace(size_t bytes, void** out, int slot, cudaStream_t stream);
// One zero-initialised 32-bit device counter per (device, stream): the ticket of the fused beta-pow update.
// Launches on one stream are ordered, so they may share it; launches on different streams get their own.
int32_t ticket_counter(void** out, cudaStream_t stream);

bool pdl_enabled();  // PRB_PDL=0 switches programmatic dependent launch off (A/B aid)

#ifdef __CUDACC__
// ---------------------------------------------------------------- programmatic dependent launch
// The pack and the norm finalize of a step follow each other on one stream.  Launched with
// cudaLaunchAttributeProgrammaticStreamSerialization, the finalize's few CTAs are scheduled while the pack is still
// draining (as soon as every CTA of the pack has executed pdl_trigger or exited); what depends on the pack's results
// comes after pdl_wait, which returns only when that kernel has completed and its writes are visible.  Both
// instructions are no-ops in a normally launched kernel.  (Tried and dropped for the update pass: with 100 000
// small CTAs the wait instruction itself cost 5 % of the kernel, and CTAs parked at the wait held the SMs the
// finalize needed -- BERT-base step 0.67 -> 0.70 ms, profiles/r2_experiments.md.)
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

template <typename... KArgs, typename... Args>
inline cudaError_t launch_kernel(void (*kernel)(KArgs...), int grid, int block, cudaStream_t stream, bool pdl,
                                 Args&&... args) {
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)grid);
  cfg.blockDim = dim3((unsigned)block);
  cfg.stream = stream;
  cfg.attrs = attr;
  cfg.numAttrs = (pdl && pdl_enabled()) ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// ---------------------------------------------------------------- element conversions
// MT = float for every supported T (ref: paddle/phi/common/amp_type_traits.h:24-39).
template <typename T>
struct Conv;
template <>
struct Conv<float> {
  static __device__ __forceinline__ float to_f32(float v) { return v; }
  static __device__ __forceinline__ float from_f32(float v) { return v; }
};
template <>
struct Conv<__nv_bfloat16> {
  static __device__ __forceinline__ float to_f32(__nv_bfloat16 v) { return __bfloat162float(v); }
  static __device__ __forceinline__ __nv_bfloat16 from_f32(float v) { return __float2bfloat16_rn(v); }
};
template <>
struct Conv<__half> {
  static __device__ __forceinline__ float to_f32(__half v) { return __half2float(v); }
  static __device__ __forceinline__ __half from_f32(float v) { return __float2half_rn(v); }
};

// Round an fp32 value through T and back: static_cast<MT>(static_cast<T>(v)).
template <typename T>
__device__ __forceinline__ float round_through(float v) {
  return Conv<T>::to_f32(Conv<T>::from_f32(v));
}
template <>
__device__ __forceinline__ float round_through<float>(float v) {
  return v;
}

// ---------------------------------------------------------------- vector access
// One "lane vector" is kVec = 8 consecutive elements: 256-bit for fp32 (LDG.E.256 / STG.E.256,
// new on sm_100), 128-bit for 16-bit types.  Loads are streaming (L1::no_allocate): every byte of
// this path is touched once per pass.
constexpr int kVec = 8;

template <typename T>
struct alignas(sizeof(T) * kVec) Pack {
  T v[kVec];
};

// Generic over the element size: 4-byte elements move as one 256-bit access, 2-byte elements as
// one 128-bit access (bit patterns are moved untouched, so the same code serves pure copies).
template <typename T>
__device__ __forceinline__ void load_vec(const T* __restrict__ p, Pack<T>& o) {
  static_assert(sizeof(T) == 4 || sizeof(T) == 2, "element size");
  uint32_t* w = reinterpret_cast<uint32_t*>(o.v);
  if constexpr (sizeof(T) == 4) {
    asm volatile("ld.global.L1::no_allocate.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7])
                 : "l"(p));
  } else {
    asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3])
                 : "l"(p));
  }
}
template <typename T>
__device__ __forceinline__ void store_vec(T* __restrict__ p, const Pack<T>& i) {
  static_assert(sizeof(T) == 4 || sizeof(T) == 2, "element size");
  const uint32_t* w = reinterpret_cast<const uint32_t*>(i.v);
  if constexpr (sizeof(T) == 4) {
    asm volatile("st.global.v8.u32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "r"(w[0]), "r"(w[1]), "r"(w[2]),
                 "r"(w[3]), "r"(w[4]), "r"(w[5]), "r"(w[6]), "r"(w[7])
                 : "memory");
  } else {
    asm volatile("st.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3])
                 : "memory");
  }
}

template <typename T>
__host__ __device__ __forceinline__ bool vec_aligned(const void* p) {
  return (reinterpret_cast<uintptr_t>(p) % (sizeof(T) * kVec)) == 0;
}

// ---------------------------------------------------------------- reductions
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Fixed-shape block sum (deterministic): warp shuffles, then warp 0 sums the per-warp values.
template <int kThreads>
__device__ __forceinline__ float block_sum(float v, float* smem /* >= kThreads/32 floats */) {
  v = warp_sum(v);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) smem[warp] = v;
  __syncthreads();
  float r = 0.f;
  if (warp == 0) {
    r = lane < kThreads / 32 ? smem[lane] : 0.f;
    r = warp_sum(r);
  }
  __syncthreads();
  return r;  // valid in warp 0 (all lanes)
}
#endif  // __CUDACC__

// ---------------------------------------------------------------- multi-tensor chunk table
// One entry per (tensor, chunk): 16 bytes, read with a single 128-bit load.
struct Chunk {
  int32_t tensor;  // index into the descriptor table
  int32_t n;       // elements in this chunk (<= chunk_elems)
  int64_t offset;  // element offset inside the tensor (multiple of chunk_elems)
};
static_assert(sizeof(Chunk) == 16, "chunk entry must be one 128-bit word");

}  // namespace prb
