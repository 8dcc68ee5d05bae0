// One multi-tensor pass over gradients: AMP unscale + inf/nan check + squared-norm partials.
//
//   ref: CheckFiniteAndUnscale            paddle/phi/kernels/gpu/amp_kernel.cu:33-82
//        SquaredL2NormKernel              paddle/phi/kernels/gpu/squared_l2_norm_kernel.cu:23-35
//        ClipGradByGlobalNorm glue        python/paddle/nn/clip.py:643-736
//
// Norm reduction = warp shuffles inside each CTA (fixed shape) + ONE grid-level pass (a single CTA
// that sums the per-chunk partials in chunk order).  No float atomics anywhere, so results are
// bit-identical run to run (ref: test_squared_l2_norm_op.py:98-111).
//
// The per-chunk work (ChunkEpilogue below) is shared by three producers so that they leave
// bit-identical partial sums and flags behind:
//   grad_pass_kernel   mode 0/1   a pass of its own over gradients that already sit somewhere
//   grad_pass_kernel   mode 2     the reducer's pack: gradient -> 1/nranks scale -> bucket slot, check + norm on
//                                 the value just stored (single rank: the bucket IS the reduced gradient)
//   nvl_exchange_kernel           the slice owner of the two-shot allreduce, on the reduced value it is about to
//                                 store into every rank's bucket (prb_nvlink.cu)
// With the last two the gradient is not read again before the update pass.
#pragma once
#include "prb_plan.h"

namespace prb {

struct alignas(16) GradDesc {
  const void* x;
  void* out;             // where the result is stored (mode 1: unscaled value, mode 2: the bucket slot)
  int32_t aligned;
  int32_t partial_base;  // mode 2 / exchange: partials[partial_base + k] belongs to chunk k of this tensor
  int64_t pad1_;
};
static_assert(sizeof(GradDesc) == 32, "GradDesc layout");

// Launch geometry of every kernel that produces norm partials: 256 threads x 4 lane vectors = 8192 elements.
constexpr int kGradThreads = 256;
constexpr int kGradUnroll = 4;
constexpr int kGradChunk = kGradThreads * kVec * kGradUnroll;

inline int grad_chunk_elems() { return kGradChunk; }

struct GradPassOpts {
  int mode = 0;                 // 0: read only; 1: store T(MT(x) * inv) to out; 2: store T(scale * (MT(x) + bias)) to out
  bool norm = false;            // accumulate squares into partials
  bool mp = false;              // squares kept in fp32 (clip.py:195-202) instead of rounded to T
  const float* loss_scale = nullptr;  // nullptr: no unscale and no finite check
  bool* found_inf = nullptr;          // must have been cleared
  float* partials = nullptr;
  float pack_scale = 1.0f, pack_bias = 0.0f;  // mode 2 (ref scale_kernel.cu:35-40, bias_after_scale = false)
};
int32_t launch_grad_pass(int32_t dtype, const DeviceTable& tab, const GradPassOpts& o, cudaStream_t stream);

struct NormGroup {
  const float* partials;       // [nchunks of this group's table]
  const int32_t* first_chunk;  // [ntensors + 1]
  int32_t ntensors;
  int32_t round_mode;          // 0: keep fp32; PRB_BFLOAT16 / PRB_FLOAT16: round sums to that type
  float* tensor_sums;          // scratch [ntensors]
  void* typed_out;             // optional: per-tensor sums stored as the group's dtype (squared_l2_norm op)
};

struct NormFinalizeArgs {
  NormGroup g[3];
  int32_t ngroups;
  float clip_norm;
  int32_t auto_skip_clip;
  int32_t compute_clip;  // 0: only per-tensor sums (squared_l2_norm op)
  float* norm_out;       // optional fp32 device scalar
  float* clip_out;       // fp32 device scalar (1.0f means "leave gradients untouched")
  int32_t grid_hint;     // CTAs worth launching (finalize_grid_hint); the result does not depend on it
  int32_t pdl;           // 1: launch as a programmatic dependent of the kernel before it in the stream
};

// How many CTAs the finalize can keep busy: tensor t goes to CTA t % grid (large) or warp t % (grid * 32) (small).
inline int finalize_grid_hint(const int64_t* numels, int64_t num) {
  int64_t big = 0;
  for (int64_t i = 0; i < num; ++i) big += numels[i] > (int64_t)256 * kGradChunk ? 1 : 0;
  const int64_t g = big > (num + 31) / 32 ? big : (num + 31) / 32;
  return (int)(g < 1 ? 1 : (g > 64 ? 64 : g));
}

int32_t launch_norm_finalize(const NormFinalizeArgs& a, cudaStream_t stream);

// g = T(MT(g) * MT(T(*clip))) in place; the whole launch is a no-op when *clip == 1.0f.
int32_t launch_clip_mul(int32_t dtype, const DeviceTable& tab, const float* clip, cudaStream_t stream);

#ifdef __CUDACC__
// Two 16-bit elements at a time.  For bf16/fp16 gradients the pass is issue-bound, not HBM-bound (ncu: 64 % issue
// slots, 39 % DRAM), so the per-element work is done on packed pairs: one cvt packs two narrowed values, the
// non-finite test is two integer ops on the packed exponents, and the T-rounded square is ONE packed multiply --
// exact, because the product of two 8-bit (11-bit) significands fits fp32, so rounding it once to T is what
// "T(x) * T(x)" means.  Results are bit-identical to the element-wise code.
template <typename T>
struct Pair;
template <>
struct Pair<__nv_bfloat16> {
  static __device__ __forceinline__ void unpack(uint32_t w, float& lo, float& hi) {
    lo = __uint_as_float(w << 16);
    hi = __uint_as_float(w & 0xffff0000u);
  }
  static __device__ __forceinline__ uint32_t pack_rn(float lo, float hi) {
    const __nv_bfloat162 r = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<const uint32_t*>(&r);
  }
  // bit 15 / 31 set iff the exponent field of the low / high half is all ones (inf or nan)
  static __device__ __forceinline__ uint32_t nonfinite(uint32_t w) { return ((w & 0x7F807F80u) + 0x00800080u) & 0x80008000u; }
  static __device__ __forceinline__ uint32_t square(uint32_t w) {
    const __nv_bfloat162 v = *reinterpret_cast<const __nv_bfloat162*>(&w);
    const __nv_bfloat162 r = __hmul2(v, v);
    return *reinterpret_cast<const uint32_t*>(&r);
  }
  static __device__ __forceinline__ uint32_t add_zero(uint32_t w) {
    uint32_t r;
    asm("add.rn.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(w), "r"(0u));
    return r;
  }
};
template <>
struct Pair<__half> {
  static __device__ __forceinline__ void unpack(uint32_t w, float& lo, float& hi) {
    const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&w));
    lo = f.x;
    hi = f.y;
  }
  static __device__ __forceinline__ uint32_t pack_rn(float lo, float hi) {
    const __half2 r = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<const uint32_t*>(&r);
  }
  static __device__ __forceinline__ uint32_t nonfinite(uint32_t w) { return ((w & 0x7C007C00u) + 0x04000400u) & 0x80008000u; }
  static __device__ __forceinline__ uint32_t square(uint32_t w) {
    const __half2 v = *reinterpret_cast<const __half2*>(&w);
    const __half2 r = __hmul2(v, v);
    return *reinterpret_cast<const uint32_t*>(&r);
  }
  static __device__ __forceinline__ uint32_t add_zero(uint32_t w) {
    uint32_t r;
    asm("add.rn.f16x2 %0, %1, %2;" : "=r"(r) : "r"(w), "r"(0u));
    return r;
  }
};

// Scale of the pack (mode 2): out = T(scale * (MT(x) + bias))   (ref scale_kernel.cu:35-40, bias_after_scale = false)
template <typename T>
__device__ __forceinline__ T pack_scale_elem(T raw, float scale, float bias) {
  return Conv<T>::from_f32(__fmul_rn(scale, __fadd_rn(Conv<T>::to_f32(raw), bias)));
}
template <typename T>
__device__ __forceinline__ void pack_scale_vec(Pack<T>& v, float scale, float bias) {
  if constexpr (sizeof(T) == 2) {
    uint32_t* w = reinterpret_cast<uint32_t*>(v.v);
    if (scale == 1.0f && bias == 0.0f) {
      // One rank: T(1.0f * (MT(x) + 0.0f)) only turns -0 into +0 and a NaN into the canonical NaN -- which is exactly
      // what ONE packed add of +0 in T does (x + 0 is exact in any format; subnormals are kept).  1 instruction per
      // pair instead of 7; bit-identical on all 65536 patterns (tests/test_gpu_fused_sync.py).
#pragma unroll
      for (int j = 0; j < kVec / 2; ++j) w[j] = Pair<T>::add_zero(w[j]);
      return;
    }
#pragma unroll
    for (int j = 0; j < kVec / 2; ++j) {
      float lo, hi;
      Pair<T>::unpack(w[j], lo, hi);
      w[j] = Pair<T>::pack_rn(__fmul_rn(scale, __fadd_rn(lo, bias)), __fmul_rn(scale, __fadd_rn(hi, bias)));
    }
  } else {
#pragma unroll
    for (int j = 0; j < kVec; ++j) v.v[j] = pack_scale_elem<T>(v.v[j], scale, bias);
  }
}

// Per-thread state of the check + norm of one chunk.  A thread feeds it its lane vectors in the order
// u = 0 .. kGradUnroll-1 (vector index u * kGradThreads + threadIdx.x), then (one thread in eight) the ragged
// tail element; finish() does the CTA-wide part.  Whoever follows this protocol produces the same bits.
template <typename T, bool kNorm, bool kMp>
struct ChunkEpilogue {
  float acc = 0.0f;
  bool bad = false;
  uint32_t bad_bits = 0;
  bool amp = false;
  float inv = 1.0f;

  __device__ __forceinline__ void begin(const float* __restrict__ loss_scale) {
    amp = loss_scale != nullptr;
    if (amp) inv = (float)(1.0 / (double)__ldg(loss_scale));  // ref amp_kernel.cu:29 (double divide, narrowed)
  }
  // one element: returns the unscaled (narrowed) value
  __device__ __forceinline__ T elem(T raw) {
    float v = Conv<T>::to_f32(raw);
    T narrow = raw;
    if (amp) {
      narrow = Conv<T>::from_f32(__fmul_rn(v, inv));  // val = MT(x) * inv; narrow = T(val)  (ref amp_kernel.cu:66-68)
      v = Conv<T>::to_f32(narrow);
      if (!isfinite(v)) bad = true;                    // finite check on the narrowed value (ref :70-77)
    }
    if (kNorm) {
      // SquareFunctor<T,T>: T(x)*T(x) rounded to T, accumulated in MT=float
      // (ref functor_primitives.h:115-123, reduce_function.h:1030-1048); kMp keeps fp32 squares.
      const float sq = __fmul_rn(v, v);
      acc = __fadd_rn(acc, kMp ? sq : round_through<T>(sq));
    }
    return narrow;
  }
  // one lane vector, in place: v becomes the unscaled value
  __device__ __forceinline__ void vec(Pack<T>& v) {
    if constexpr (sizeof(T) == 2 && !kMp) {
      uint32_t* w = reinterpret_cast<uint32_t*>(v.v);
#pragma unroll
      for (int j = 0; j < kVec / 2; ++j) {
        uint32_t nw = w[j];
        if (amp) {
          float lo, hi;
          Pair<T>::unpack(nw, lo, hi);
          nw = Pair<T>::pack_rn(__fmul_rn(lo, inv), __fmul_rn(hi, inv));  // narrow = T(MT(x) * inv)
          bad_bits |= Pair<T>::nonfinite(nw);                             // finite check on the narrowed value
          w[j] = nw;
        }
        if (kNorm) {
          float lo, hi;
          Pair<T>::unpack(Pair<T>::square(nw), lo, hi);  // T(x)*T(x) rounded to T, both halves at once
          acc = __fadd_rn(__fadd_rn(acc, lo), hi);       // same order as the element-wise code
        }
      }
    } else {
#pragma unroll
      for (int j = 0; j < kVec; ++j) v.v[j] = elem(v.v[j]);
    }
  }
  // CTA-wide: returns (in warp 0) the chunk's sum of squares; *any_bad = some thread saw a non-finite value.
  // Every thread of the CTA must call it.
  __device__ __forceinline__ float finish(float* red /* kGradThreads / 32 floats */, bool* any_bad) {
    if (bad_bits) bad = true;
    *any_bad = amp ? (__syncthreads_or(bad ? 1 : 0) != 0) : false;
    return kNorm ? block_sum<kGradThreads>(acc, red) : 0.0f;
  }
};
#endif  // __CUDACC__

}  // namespace prb
