// prb_squared_l2_norm_noalloc: the squared-L2-norm reduction for callers that must not allocate.
//
// Who needs this.  prb_squared_l2_norm keeps its per-chunk partial sums in a workspace that the library
// grows with cudaMalloc on demand and finds its chunk table through a plan cache that also allocates.
// That is fine for a training loop, but some callers cannot tolerate a device allocation at the point
// where they ask for a gradient norm:
//   * code that runs after the process has handed (almost) all device memory to its own pool, where
//     a cudaMalloc may fail or fragment the pool's address space;
//   * latency-critical or "no implicit synchronisation" sections -- cudaMalloc/cudaFree can block on
//     other streams of the device;
//   * health checks and out-of-memory handlers that want a norm precisely when memory is exhausted.
// For them this entry point does the same reduction with its scratch in a fixed array that is part of
// the library image: every GPU that loads the library carries it (kScratchFloats * 4 bytes, ~128 KiB)
// from load until exit, and in exchange no call ever allocates or frees anything.
//
// Limits and sharing rules (also in include/paddle_ray_b200.h):
//   * x may hold at most PRB_NOALLOC_MAX_NUMEL elements; larger inputs are rejected with
//     PRB_INVALID_ARGUMENT before anything is launched.
//   * All calls on a device use the SAME array.  A call never relies on what an earlier call left
//     there: every slot it reads was written earlier in that same call, so nothing has to be cleared.
//     The flip side: two calls that overlap on the GPU (different streams, or different host threads
//     on different streams) overwrite each other's partial sums.  Keep calls on one stream, or order
//     them with events.
//   * The first call on each device looks the array's device address up (cudaGetSymbolAddress, which
//     may also make the runtime load this module).  Make that first call during start-up, outside of
//     any stream capture; later calls only launch two kernels.
#include "prb_common.cuh"

namespace prb {
namespace {

constexpr int kNaThreads = 256;                              // same CTA shape as the default grad pass
constexpr int kNaUnroll = 4;
constexpr int kNaChunk = kNaThreads * kVec * kNaUnroll;      // 8192 elements per CTA
constexpr int kNaMaxChunks = 32768;                          // -> at most 2^28 elements per call
constexpr int kScratchFloats = kNaMaxChunks + 64;            // partial sums + one slot for the total
static_assert((int64_t)kNaMaxChunks * kNaChunk == PRB_NOALLOC_MAX_NUMEL, "header and kernel disagree on the limit");

// The scratch itself.  In host code this name is only a handle for cudaGetSymbolAddress -- it is NOT
// a device pointer and must never be passed to a kernel as one.
static __device__ __align__(256) float g_noalloc_scratch[kScratchFloats];

// Per-device cache of the array's device address (filled on the first call on that device; two threads that race
// on the first call store the same value).
std::atomic<float*> g_scratch_addr[64];

int32_t scratch_for_current_device(float** out) {
  int dev = -1;
  PRB_CUDA(cudaGetDevice(&dev));
  PRB_ENFORCE(dev >= 0 && dev < 64, PRB_OUT_OF_RANGE, "squared_l2_norm_noalloc: device ordinal %d not supported", dev);
  if (g_scratch_addr[dev].load(std::memory_order_acquire) == nullptr) {
    float* p = nullptr;
    // With `dev` current this yields the address of dev's copy of the array.
    PRB_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_noalloc_scratch));
    PRB_ENFORCE(p != nullptr, PRB_EXTERNAL, "squared_l2_norm_noalloc: the scratch array has no device address");
    g_scratch_addr[dev].store(p, std::memory_order_release);
  }
  *out = g_scratch_addr[dev].load(std::memory_order_acquire);
  return PRB_SUCCESS;
}

// Pass 1: one CTA per 8192-element chunk; squares are rounded to T for 16-bit inputs and summed in
// fp32 (same arithmetic as the allocating entry point), block sum -> partials[chunk].
template <typename T>
__global__ void __launch_bounds__(kNaThreads) na_square_sum_kernel(const T* __restrict__ x, int64_t numel,
                                                                   float* __restrict__ partials) {
  __shared__ float red[kNaThreads / 32];
  const int64_t offset = (int64_t)blockIdx.x * kNaChunk;
  const int64_t left = numel - offset;
  const int n = left < kNaChunk ? (int)left : kNaChunk;
  const T* __restrict__ xc = x + offset;
  float acc = 0.0f;
  auto add = [&](T raw) {
    const float v = Conv<T>::to_f32(raw);
    acc = __fadd_rn(acc, round_through<T>(__fmul_rn(v, v)));
  };
  if (vec_aligned<T>(x)) {  // chunk starts are multiples of 8192 elements, so the base decides
    const int nvec = n / kVec;
    Pack<T> v[kNaUnroll];
#pragma unroll
    for (int u = 0; u < kNaUnroll; ++u) {
      const int i = u * kNaThreads + threadIdx.x;
      if (i < nvec) load_vec<T>(xc + i * kVec, v[u]);
    }
#pragma unroll
    for (int u = 0; u < kNaUnroll; ++u) {
      const int i = u * kNaThreads + threadIdx.x;
      if (i < nvec) {
#pragma unroll
        for (int j = 0; j < kVec; ++j) add(v[u].v[j]);
      }
    }
    const int e = nvec * kVec + threadIdx.x;
    if (e < n) add(xc[e]);
  } else {
    for (int e = threadIdx.x; e < n; e += kNaThreads) add(xc[e]);
  }
  const float s = block_sum<kNaThreads>(acc, red);
  if (threadIdx.x == 0) partials[blockIdx.x] = s;
}

// Pass 2: one CTA adds the partial sums in a fixed order and stores the result in x's dtype.
// It names the array directly -- device code may do that.
template <typename T>
__global__ void __launch_bounds__(1024) na_finalize_kernel(int nchunks, T* __restrict__ out) {
  __shared__ float red[32];
  const float* partials = g_noalloc_scratch;
  if (nchunks <= 256) {  // short inputs: one warp, lane-strided, then a fixed butterfly
    if (threadIdx.x < 32) {
      float s = 0.0f;
      for (int c = threadIdx.x; c < nchunks; c += 32) s = __fadd_rn(s, partials[c]);
      s = warp_sum(s);
      if (threadIdx.x == 0) {
        g_noalloc_scratch[kNaMaxChunks] = s;
        out[0] = Conv<T>::from_f32(s);
      }
    }
    return;
  }
  float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
  int c = threadIdx.x;
  for (; c + 3 * 1024 < nchunks; c += 4 * 1024) {
    const float v0 = partials[c], v1 = partials[c + 1024], v2 = partials[c + 2048], v3 = partials[c + 3072];
    s0 = __fadd_rn(s0, v0);
    s1 = __fadd_rn(s1, v1);
    s2 = __fadd_rn(s2, v2);
    s3 = __fadd_rn(s3, v3);
  }
  for (; c < nchunks; c += 1024) s0 = __fadd_rn(s0, partials[c]);
  float s = __fadd_rn(__fadd_rn(s0, s1), __fadd_rn(s2, s3));
  s = block_sum<1024>(s, red);
  if (threadIdx.x == 0) {
    g_noalloc_scratch[kNaMaxChunks] = s;  // the fp32 total stays readable for debugging
    out[0] = Conv<T>::from_f32(s);
  }
}

template <typename T>
int32_t run_noalloc(const PRB_Tensor* x, PRB_Tensor* out, float* scratch, cudaStream_t stream) {
  const int nchunks = (int)((x->numel + kNaChunk - 1) / kNaChunk);
  if (nchunks > 0) {
    na_square_sum_kernel<T><<<nchunks, kNaThreads, 0, stream>>>(reinterpret_cast<const T*>(x->data), x->numel, scratch);
    PRB_LAUNCH_CHECK();
  }
  na_finalize_kernel<T><<<1, 1024, 0, stream>>>(nchunks, reinterpret_cast<T*>(out->data));
  PRB_LAUNCH_CHECK();
  return PRB_SUCCESS;
}

}  // namespace
}  // namespace prb

using namespace prb;

extern "C" int32_t prb_squared_l2_norm_noalloc(const PRB_Tensor* x, PRB_Tensor* out, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  PRB_TRY(require_device());
  PRB_ENFORCE(x && out, PRB_INVALID_ARGUMENT, "squared_l2_norm_noalloc: NULL tensor");
  PRB_ENFORCE(out->numel == 1 && out->dtype == x->dtype && out->device >= 0 && x->device >= 0, PRB_INVALID_ARGUMENT,
              "squared_l2_norm_noalloc: out must be one element of x's dtype, both on the device");
  PRB_ENFORCE(x->numel >= 0 && x->numel <= (int64_t)PRB_NOALLOC_MAX_NUMEL, PRB_INVALID_ARGUMENT,
              "squared_l2_norm_noalloc: x has %lld elements but the fixed scratch covers at most %lld; use "
              "prb_squared_l2_norm (which allocates) or reduce the tensor in pieces",
              (long long)x->numel, (long long)PRB_NOALLOC_MAX_NUMEL);
  PRB_ENFORCE(out->data != nullptr && (x->data != nullptr || x->numel == 0), PRB_INVALID_ARGUMENT,
              "squared_l2_norm_noalloc: NULL data pointer");
  int dev = -1;
  PRB_CUDA(cudaGetDevice(&dev));
  PRB_ENFORCE(x->device == dev && out->device == dev, PRB_INVALID_ARGUMENT,
              "squared_l2_norm_noalloc: tensors live on device %d but device %d is current", x->device, dev);
  float* scratch = nullptr;
  PRB_TRY(scratch_for_current_device(&scratch));
  switch (x->dtype) {
    case PRB_FLOAT32: return run_noalloc<float>(x, out, scratch, stream);
    case PRB_BFLOAT16: return run_noalloc<__nv_bfloat16>(x, out, scratch, stream);
    case PRB_FLOAT16: return run_noalloc<__half>(x, out, scratch, stream);
    default:
      PRB_FAIL(PRB_UNIMPLEMENTED, "squared_l2_norm_noalloc: dtype %s is not built (float32/bfloat16/float16 only)",
               dtype_name(x->dtype));
  }
}
