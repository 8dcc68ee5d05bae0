// check_finite_and_unscale_, squared_l2_norm, ClipGradByGlobalNorm, update_loss_scaling_.
// See prb_grad_pass.cuh for the design and the reference citations.
#include <vector>

#include "prb_fp64.h"
#include "prb_grad_pass.cuh"

namespace prb {
namespace {

// kMode 0: read only.  1: store the unscaled value (check_finite_and_unscale_).  2: the reducer's pack -- store
// T(scale * (x + bias)) into the bucket slot and run the check + norm on the value just stored.
template <typename T, int kMode, bool kNorm, bool kMp>
__global__ void __launch_bounds__(kGradThreads, 2) grad_pass_kernel(const GradDesc* __restrict__ descs,
                                                                    const Chunk* __restrict__ chunks,
                                                                    const float* __restrict__ loss_scale,
                                                                    bool* __restrict__ found_inf,
                                                                    float* __restrict__ partials,
                                                                    const float pack_scale, const float pack_bias) {
  __shared__ float red[kGradThreads / 32];
  pdl_trigger();  // the norm finalize that follows may be scheduled while this grid drains
  const int4 craw = __ldg(reinterpret_cast<const int4*>(chunks) + blockIdx.x);
  const int n = craw.y;
  const int64_t offset = (int64_t)(((uint64_t)(uint32_t)craw.w << 32) | (uint32_t)craw.z);
  const GradDesc* __restrict__ d = descs + craw.x;
  const T* __restrict__ x = reinterpret_cast<const T*>(d->x) + offset;
  T* __restrict__ out = kMode != 0 ? reinterpret_cast<T*>(d->out) + offset : nullptr;
  ChunkEpilogue<T, kNorm, kMp> ep;

  auto one = [&](int e) {  // element-wise path: ragged tail and unaligned tensors
    T r = x[e];
    if (kMode == 2) {
      r = pack_scale_elem<T>(r, pack_scale, pack_bias);
      out[e] = r;
      ep.elem(r);
    } else {
      const T nr = ep.elem(r);
      if (kMode == 1) out[e] = nr;
    }
  };

  if (d->aligned) {
    const int nvec = n / kVec;
    Pack<T> v[kGradUnroll];
#pragma unroll
    for (int u = 0; u < kGradUnroll; ++u) {
      const int i = u * kGradThreads + threadIdx.x;
      if (i < nvec) load_vec<T>(x + i * kVec, v[u]);
    }
    ep.begin(loss_scale);
#pragma unroll
    for (int u = 0; u < kGradUnroll; ++u) {
      const int i = u * kGradThreads + threadIdx.x;
      if (i < nvec) {
        if (kMode == 2) {
          pack_scale_vec<T>(v[u], pack_scale, pack_bias);
          store_vec<T>(out + i * kVec, v[u]);
          ep.vec(v[u]);
        } else {
          ep.vec(v[u]);
          if (kMode == 1) store_vec<T>(out + i * kVec, v[u]);
        }
      }
    }
    const int e = nvec * kVec + threadIdx.x;
    if (e < n) one(e);
  } else {
    ep.begin(loss_scale);
    for (int e = threadIdx.x; e < n; e += kGradThreads) one(e);
  }

  bool any_bad;
  const float s = ep.finish(red, &any_bad);
  // only `true` is ever written, so concurrent stores are benign (ref amp_kernel.cu:79-81)
  if (any_bad && threadIdx.x == 0) *found_inf = true;
  if (kNorm && threadIdx.x == 0)
    partials[kMode == 2 ? d->partial_base + (int)(offset / kGradChunk) : (int)blockIdx.x] = s;
}

__device__ __forceinline__ float round_mode(float v, int mode) {
  if (mode == PRB_BFLOAT16) return round_through<__nv_bfloat16>(v);
  if (mode == PRB_FLOAT16) return round_through<__half>(v);
  return v;
}

// The one grid-level pass.  Phase 1 turns chunk partials into per-tensor sums in a fixed order: a tensor with many
// chunks is summed by a whole 1024-thread CTA (thread-strided with four independent accumulators, then the
// fixed-shape block tree), a small tensor by one warp (lane-strided, then a fixed butterfly).  The tensors are
// spread over a few CTAs -- WHICH CTA sums a tensor changes nothing about HOW it is summed, so the result does not
// depend on the grid (round 1 ran this on one CTA: 52 us for BERT-base's 25 large tensors, one latency chain after
// the other; spread out it is ~10 us).  Phase 2 runs in the CTA that finishes last (ticket): thread 0 adds the
// per-tensor sums in list order per dtype group (add_n), casts, adds the groups, sqrt, clip coefficient
// (ref clip.py:685-713).  No float atomics: bit-identical run to run.
constexpr int kFinThreads = 1024;
constexpr int kFinBigTensor = 256;  // chunks; above this a whole CTA cooperates on the tensor
constexpr int kFinMaxGrid = 64;

__device__ __forceinline__ void store_tensor_sum(const NormGroup& g, int t, float s) {
  g.tensor_sums[t] = s;
  if (g.typed_out != nullptr) {
    if (g.round_mode == PRB_BFLOAT16) reinterpret_cast<__nv_bfloat16*>(g.typed_out)[t] = __float2bfloat16_rn(s);
    else if (g.round_mode == PRB_FLOAT16) reinterpret_cast<__half*>(g.typed_out)[t] = __float2half_rn(s);
    else reinterpret_cast<float*>(g.typed_out)[t] = s;
  }
}

constexpr int kFinMaxTensors = 4096;  // tensors per dtype group staged in shared memory at a time

__global__ void __launch_bounds__(kFinThreads) norm_finalize_kernel(const NormFinalizeArgs a, unsigned int* ticket) {
  __shared__ float red[kFinThreads / 32];
  __shared__ int s_first[kFinMaxTensors + 1];   // chunk ranges of the group's tensors
  __shared__ float s_sum[kFinMaxTensors];       // per-tensor sums of the group (phase 2)
  __shared__ float s_group[3];
  __shared__ unsigned int s_last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = kFinThreads >> 5;
  pdl_wait();  // launched as a programmatic dependent of the kernel that produces the partials: wait for it here
  // ---- phase 1: per-tensor sums.  Tensor t belongs to CTA t % gridDim.x when it is large (the whole CTA sums it)
  // and to warp t % (gridDim.x * 32) of the grid when it is small: no thread ever scans the tensor list (a first
  // version in which every warp walked all tensors to find its share spent 12 M instructions on that walk: 73 us).
  for (int gi = 0; gi < a.ngroups; ++gi) {
    const NormGroup& g = a.g[gi];
    for (int base = 0; base < g.ntensors; base += kFinMaxTensors) {
      const int nt = min(kFinMaxTensors, g.ntensors - base);
      // stage the chunk ranges with one coalesced pass (a serial walk over global memory costs ~0.3 us per tensor)
      for (int i = threadIdx.x; i <= nt; i += kFinThreads) s_first[i] = g.first_chunk[base + i];
      __syncthreads();
      // big tensors: all threads of the owning CTA (uniform control flow inside a CTA)
      for (int t = (int)blockIdx.x; t < nt; t += (int)gridDim.x) {
        const int c0 = s_first[t], c1 = s_first[t + 1];
        if (c1 - c0 <= kFinBigTensor) continue;
        float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
        int c = c0 + threadIdx.x;
        for (; c + 3 * kFinThreads < c1; c += 4 * kFinThreads) {
          const float v0 = g.partials[c], v1 = g.partials[c + kFinThreads], v2 = g.partials[c + 2 * kFinThreads],
                      v3 = g.partials[c + 3 * kFinThreads];
          s0 = __fadd_rn(s0, v0);
          s1 = __fadd_rn(s1, v1);
          s2 = __fadd_rn(s2, v2);
          s3 = __fadd_rn(s3, v3);
        }
        for (; c < c1; c += kFinThreads) s0 = __fadd_rn(s0, g.partials[c]);
        float s = __fadd_rn(__fadd_rn(s0, s1), __fadd_rn(s2, s3));
        s = block_sum<kFinThreads>(s, red);
        if (threadIdx.x == 0) store_tensor_sum(g, base + t, round_mode(s, g.round_mode));  // the op's output is T-typed
      }
      // small tensors: one warp each
      for (int t = (int)blockIdx.x * nwarps + warp; t < nt; t += (int)gridDim.x * nwarps) {
        const int c0 = s_first[t], c1 = s_first[t + 1];
        if (c1 - c0 > kFinBigTensor) continue;
        float s = 0.0f;
        for (int c = c0 + lane; c < c1; c += 32) s = __fadd_rn(s, g.partials[c]);
        s = warp_sum(s);
        if (lane == 0) store_tensor_sum(g, base + t, round_mode(s, g.round_mode));
      }
      __syncthreads();
    }
  }
  // ---- the CTA that finishes last does phase 2 (fence + ticket: it sees every other CTA's sums)
  if (gridDim.x > 1) {
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
      const unsigned int t = atomicAdd(ticket, 1u);
      s_last = (t == gridDim.x - 1) ? 1u : 0u;
      if (s_last) {
        *ticket = 0u;
        __threadfence();
      }
    }
    __syncthreads();
    if (!s_last) return;
  }
  if (!a.compute_clip) return;
  // ---- phase 2: add_n per group in list order
  for (int gi = 0; gi < a.ngroups; ++gi) {
    const NormGroup& g = a.g[gi];
    float group_sum = 0.0f;  // meaningful in thread 0 only
    for (int base = 0; base < g.ntensors; base += kFinMaxTensors) {
      const int nt = min(kFinMaxTensors, g.ntensors - base);
      for (int t = threadIdx.x; t < nt; t += kFinThreads) s_sum[t] = __ldcg(g.tensor_sums + base + t);
      __syncthreads();
      if (threadIdx.x == 0)
        for (int t = 0; t < nt; ++t) group_sum = __fadd_rn(group_sum, s_sum[t]);
      __syncthreads();
    }
    if (threadIdx.x == 0) s_group[gi] = group_sum;
  }
  if (threadIdx.x == 0) {
    float total = 0.0f;
    for (int gi = 0; gi < a.ngroups; ++gi) {
      const NormGroup& g = a.g[gi];
      if (g.ntensors == 0) continue;
      const float s = round_mode(s_group[gi], g.round_mode);  // add_n output is T-typed
      total = __fadd_rn(total, s);                            // astype(fp32) then add_n
    }
    const float norm = __fsqrt_rn(total);
    float clip = 1.0f;
    if (!a.auto_skip_clip) clip = __fdiv_rn(a.clip_norm, fmaxf(norm, a.clip_norm));
    else if (norm > a.clip_norm) clip = __fdiv_rn(a.clip_norm, norm);
    if (a.norm_out) *a.norm_out = norm;
    if (a.clip_out) *a.clip_out = clip;
  }
}

template <typename T>
__global__ void __launch_bounds__(kGradThreads) clip_mul_kernel(const GradDesc* __restrict__ descs,
                                                                const Chunk* __restrict__ chunks,
                                                                const float* __restrict__ clip_ptr) {
  const float craw_clip = __ldg(clip_ptr);
  if (craw_clip == 1.0f) return;  // multiplying by exactly 1 changes nothing: skip the whole pass
  const float c = round_through<T>(craw_clip);  // clip_var.astype(g.dtype)  (ref clip.py:724-728)
  const int4 craw = __ldg(reinterpret_cast<const int4*>(chunks) + blockIdx.x);
  const int n = craw.y;
  const int64_t offset = (int64_t)(((uint64_t)(uint32_t)craw.w << 32) | (uint32_t)craw.z);
  const GradDesc* __restrict__ d = descs + craw.x;
  T* __restrict__ x = reinterpret_cast<T*>(d->out) + offset;
  if (d->aligned) {
    const int nvec = n / kVec;
    for (int i = threadIdx.x; i < nvec; i += kGradThreads) {  // works for any chunk size of the shared table
      Pack<T> v;
      load_vec<T>(x + i * kVec, v);
#pragma unroll
      for (int j = 0; j < kVec; ++j) v.v[j] = Conv<T>::from_f32(__fmul_rn(Conv<T>::to_f32(v.v[j]), c));
      store_vec<T>(x + i * kVec, v);
    }
    const int e = nvec * kVec + threadIdx.x;
    if (e < n) x[e] = Conv<T>::from_f32(__fmul_rn(Conv<T>::to_f32(x[e]), c));
  } else {
    for (int e = threadIdx.x; e < n; e += kGradThreads) x[e] = Conv<T>::from_f32(__fmul_rn(Conv<T>::to_f32(x[e]), c));
  }
}

// ref: FusedFillIf amp_kernel.cu:109-146 (LazyZeros): zero every x when found_inf
template <typename W>
__global__ void __launch_bounds__(kGradThreads) fill_zero_if_kernel(const GradDesc* __restrict__ descs,
                                                                    const Chunk* __restrict__ chunks,
                                                                    const bool* __restrict__ flag) {
  if (!*flag) return;
  const int4 craw = __ldg(reinterpret_cast<const int4*>(chunks) + blockIdx.x);
  const int n = craw.y;
  const int64_t offset = (int64_t)(((uint64_t)(uint32_t)craw.w << 32) | (uint32_t)craw.z);
  W* __restrict__ x = reinterpret_cast<W*>(descs[craw.x].out) + offset;
  for (int e = threadIdx.x; e < n; e += kGradThreads) x[e] = W(0);
}

// ref: GpuUpdateLossScaling amp_kernel.cu:84-107, Update() impl/amp_kernel_impl.h:38-75
__global__ void update_loss_scaling_kernel(const bool* found_inf, const float* prev, const int* good_in,
                                           const int* bad_in, int incr_every_n_steps, int decr_every_n_nan_or_inf,
                                           float incr_ratio, float decr_ratio, float* out, int* good_out, int* bad_out) {
  const float pre = *prev;
  if (*found_inf) {
    *good_out = 0;
    const int b = *bad_in + 1;
    *bad_out = b;
    if (b == decr_every_n_nan_or_inf) {
      const float nl = __fmul_rn(pre, decr_ratio);
      *out = nl < 1.0f ? 1.0f : nl;
      *bad_out = 0;
    }
  } else {
    *bad_out = 0;
    const int g = *good_in + 1;
    *good_out = g;
    if (g == incr_every_n_steps) {
      const float nl = __fmul_rn(pre, incr_ratio);
      *out = isfinite(nl) ? nl : pre;
      *good_out = 0;
    }
  }
}

template <typename T>
int32_t launch_pass_typed(const DeviceTable& tab, const GradPassOpts& o, cudaStream_t stream) {
  if (tab.nchunks == 0) return PRB_SUCCESS;
  const GradDesc* d = reinterpret_cast<const GradDesc*>(tab.descs);
  KernelTimer kt(o.mode == 2 ? KK_PACK_EPILOGUE : KK_GRAD_PASS, stream);
#define PRB_GP(M, N, P)                                                                                              \
  grad_pass_kernel<T, M, N, P><<<tab.nchunks, kGradThreads, 0, stream>>>(d, tab.chunks, o.loss_scale, o.found_inf, \
                                                                         o.partials, o.pack_scale, o.pack_bias)
#define PRB_GPM(M)                          \
  do {                                      \
    if (!o.norm) PRB_GP(M, false, false);   \
    else if (!o.mp) PRB_GP(M, true, false); \
    else PRB_GP(M, true, true);             \
  } while (0)
  if (o.mode == 0) PRB_GPM(0);
  else if (o.mode == 1) PRB_GPM(1);
  else PRB_GPM(2);
#undef PRB_GPM
#undef PRB_GP
  PRB_LAUNCH_CHECK();
  return PRB_SUCCESS;
}

}  // namespace

int32_t launch_grad_pass(int32_t dtype, const DeviceTable& tab, const GradPassOpts& o, cudaStream_t stream) {
  PRB_ENFORCE(o.loss_scale == nullptr || o.found_inf != nullptr, PRB_INVALID_ARGUMENT, "grad pass: unscale needs a found_inf flag");
  PRB_ENFORCE(!o.norm || o.partials != nullptr, PRB_INVALID_ARGUMENT, "grad pass: norm needs a partials buffer");
  switch (dtype) {
    case PRB_FLOAT32: return launch_pass_typed<float>(tab, o, stream);
    case PRB_BFLOAT16: return launch_pass_typed<__nv_bfloat16>(tab, o, stream);
    case PRB_FLOAT16: return launch_pass_typed<__half>(tab, o, stream);
    default: PRB_FAIL(PRB_UNIMPLEMENTED, "dtype %s is not built (float32/bfloat16/float16 only)", dtype_name(dtype));
  }
}

int32_t launch_norm_finalize(const NormFinalizeArgs& a, cudaStream_t stream) {
  int grid = a.grid_hint < 1 ? 1 : (a.grid_hint > kFinMaxGrid ? kFinMaxGrid : a.grid_hint);
  void* tk = nullptr;
  if (grid > 1) PRB_TRY(ticket_counter(&tk, stream));
  KernelTimer kt(KK_NORM_FINALIZE, stream);
  PRB_CUDA(launch_kernel(norm_finalize_kernel, grid, kFinThreads, stream, a.pdl != 0, a, reinterpret_cast<unsigned int*>(tk)));
  PRB_LAUNCH_CHECK();
  return PRB_SUCCESS;
}

int32_t launch_clip_mul(int32_t dtype, const DeviceTable& tab, const float* clip, cudaStream_t stream) {
  if (tab.nchunks == 0) return PRB_SUCCESS;
  const GradDesc* d = reinterpret_cast<const GradDesc*>(tab.descs);
  switch (dtype) {
    case PRB_FLOAT32: clip_mul_kernel<float><<<tab.nchunks, kGradThreads, 0, stream>>>(d, tab.chunks, clip); break;
    case PRB_BFLOAT16: clip_mul_kernel<__nv_bfloat16><<<tab.nchunks, kGradThreads, 0, stream>>>(d, tab.chunks, clip); break;
    case PRB_FLOAT16: clip_mul_kernel<__half><<<tab.nchunks, kGradThreads, 0, stream>>>(d, tab.chunks, clip); break;
    default: PRB_FAIL(PRB_UNIMPLEMENTED, "clip: dtype %s is not built", dtype_name(dtype));
  }
  PRB_LAUNCH_CHECK();
  return PRB_SUCCESS;
}

namespace {

// Build the GradDesc table of the tensors of `list` whose dtype is `dtype` (all when dtype < 0).
int32_t grad_table(const PRB_Tensor* xs, const PRB_Tensor* outs, int64_t num, int32_t dtype, cudaStream_t stream,
                   DeviceTable* tab, std::vector<int64_t>* index_of = nullptr) {
  std::vector<GradDesc> descs;
  std::vector<int64_t> numels;
  for (int64_t i = 0; i < num; ++i) {
    if (dtype >= 0 && xs[i].dtype != dtype) continue;
    GradDesc d{};
    d.x = xs[i].data;
    d.out = outs ? outs[i].data : xs[i].data;
    const uintptr_t a = (uintptr_t)dtype_size(xs[i].dtype) * kVec;
    d.aligned = (((uintptr_t)d.x % a) == 0 && ((uintptr_t)d.out % a) == 0) ? 1 : 0;
    descs.push_back(d);
    numels.push_back(xs[i].numel);
    if (index_of) index_of->push_back(i);
  }
  return get_table(descs.data(), sizeof(GradDesc), (int)descs.size(), numels.data(), grad_chunk_elems(), stream, tab);
}

}  // namespace
}  // namespace prb

using namespace prb;

extern "C" {

int32_t prb_check_finite_and_unscale(const PRB_Tensor* xs, int64_t num, const PRB_Tensor* scale, PRB_Tensor* outs,
                                     PRB_Tensor* found_infinite, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  PRB_TRY(require_device());
  PRB_ENFORCE(scale && found_infinite && num >= 0 && (xs || num == 0), PRB_INVALID_ARGUMENT,
              "check_finite_and_unscale: NULL argument");
  const bool f64 = num > 0 && xs[0].dtype == PRB_FLOAT64;  // T = MT = double: the scale is a double tensor too
  PRB_ENFORCE(scale->numel == 1 && scale->dtype == (f64 ? PRB_FLOAT64 : PRB_FLOAT32) && scale->device >= 0,
              PRB_INVALID_ARGUMENT, "check_finite_and_unscale: Scale must be one %s on the device",
              f64 ? "float64 (float64 gradients)" : "float32");
  PRB_ENFORCE(found_infinite->numel == 1 && found_infinite->dtype == PRB_BOOL && found_infinite->device >= 0,
              PRB_INVALID_ARGUMENT, "check_finite_and_unscale: FoundInfinite must be one bool on the device");
  // InverseAndMemset: *found_inf = false (ref amp_kernel.cu:27-31); the inverse is taken in-kernel.
  PRB_CUDA(cudaMemsetAsync(found_infinite->data, 0, 1, stream));
  if (num == 0) return PRB_SUCCESS;  // ref amp_kernel.cu:283-284
  const int32_t dt = xs[0].dtype;
  for (int64_t i = 0; i < num; ++i) {
    PRB_ENFORCE(xs[i].dtype == dt, PRB_INVALID_ARGUMENT, "check_finite_and_unscale: xs must share one dtype");
    PRB_ENFORCE(xs[i].device >= 0 || xs[i].numel == 0, PRB_INVALID_ARGUMENT,
                "check_finite_and_unscale: device tensors required");
    if (outs) PRB_ENFORCE(outs[i].numel == xs[i].numel && outs[i].dtype == dt, PRB_INVALID_ARGUMENT,
                          "check_finite_and_unscale: outs[%lld] does not match xs[%lld]", (long long)i, (long long)i);
  }
  if (f64) {  // ref amp_kernel.cu:352-357: one launch per tensor, not on the north-star path
    for (int64_t i = 0; i < num; ++i)
      PRB_TRY(launch_unscale_f64(reinterpret_cast<const double*>(xs[i].data),
                                 reinterpret_cast<double*>((outs ? outs[i] : xs[i]).data), xs[i].numel,
                                 reinterpret_cast<const double*>(scale->data),
                                 reinterpret_cast<bool*>(found_infinite->data), stream));
    return PRB_SUCCESS;
  }
  DeviceTable tab;
  PRB_TRY(grad_table(xs, outs, num, -1, stream, &tab));
  GradPassOpts o;
  o.mode = 1;
  o.loss_scale = reinterpret_cast<const float*>(scale->data);
  o.found_inf = reinterpret_cast<bool*>(found_infinite->data);
  return launch_grad_pass(dt, tab, o, stream);
}

int32_t prb_squared_l2_norm(const PRB_Tensor* x, PRB_Tensor* out, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  PRB_TRY(require_device());
  PRB_ENFORCE(x && out, PRB_INVALID_ARGUMENT, "squared_l2_norm: NULL tensor");
  PRB_ENFORCE(out->numel == 1 && out->dtype == x->dtype && out->device >= 0 && x->device >= 0, PRB_INVALID_ARGUMENT,
              "squared_l2_norm: out must be one element of x's dtype on the device");
  if (x->dtype == PRB_FLOAT64)  // ref squared_l2_norm_kernel.cu:38-43
    return launch_squared_l2_norm_f64(reinterpret_cast<const double*>(x->data), x->numel,
                                      reinterpret_cast<double*>(out->data), stream);
  DeviceTable tab;
  PRB_TRY(grad_table(x, nullptr, 1, -1, stream, &tab));
  void* ws = nullptr;
  const size_t npart = (size_t)tab.nchunks + 1;
  PRB_TRY(workspace((npart + 1) * sizeof(float), &ws, /*slot=*/1, stream));
  float* partials = reinterpret_cast<float*>(ws);
  GradPassOpts o;
  o.norm = true;
  o.partials = partials;
  PRB_TRY(launch_grad_pass(x->dtype, tab, o, stream));
  NormFinalizeArgs a{};
  a.ngroups = 1;
  a.g[0].partials = partials;
  a.g[0].first_chunk = tab.first_chunk;
  a.g[0].ntensors = 1;
  a.g[0].round_mode = x->dtype == PRB_FLOAT32 ? 0 : x->dtype;
  a.g[0].tensor_sums = partials + npart;
  a.g[0].typed_out = out->data;
  a.compute_clip = 0;
  return launch_norm_finalize(a, stream);
}

int32_t prb_clip_grad_by_global_norm(PRB_Tensor* grads, int64_t num, float clip_norm, int32_t auto_skip_clip,
                                     int32_t use_mp_type, PRB_Tensor* global_norm_out, PRB_Tensor* clip_out,
                                     void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  PRB_TRY(require_device());
  PRB_ENFORCE(num >= 0 && (grads || num == 0), PRB_INVALID_ARGUMENT, "clip: NULL argument");
  if (num == 0) return PRB_SUCCESS;  // "all parameters have been filtered out" (ref clip.py:676-682)
  const int32_t classes[3] = {PRB_FLOAT16, PRB_BFLOAT16, PRB_FLOAT32};
  for (int64_t i = 0; i < num; ++i) {
    PRB_ENFORCE(grads[i].dtype == PRB_FLOAT32 || grads[i].dtype == PRB_BFLOAT16 || grads[i].dtype == PRB_FLOAT16,
                PRB_UNIMPLEMENTED, "clip: dtype %s is not built", dtype_name(grads[i].dtype));
    PRB_ENFORCE(grads[i].device >= 0 || grads[i].numel == 0, PRB_INVALID_ARGUMENT, "clip: device tensors required");
  }
  DeviceTable tabs[3];
  size_t total_chunks = 0, total_tensors = 0;
  for (int c = 0; c < 3; ++c) {
    PRB_TRY(grad_table(grads, nullptr, num, classes[c], stream, &tabs[c]));
    total_chunks += tabs[c].nchunks;
    total_tensors += tabs[c].ndescs;
  }
  void* ws = nullptr;
  PRB_TRY(workspace((total_chunks + total_tensors + 2) * sizeof(float), &ws, /*slot=*/1, stream));
  float* base = reinterpret_cast<float*>(ws);
  float* clip_dev = clip_out ? reinterpret_cast<float*>(clip_out->data) : base + total_chunks + total_tensors;
  float* norm_dev = global_norm_out ? reinterpret_cast<float*>(global_norm_out->data) : nullptr;
  NormFinalizeArgs a{};
  float* part = base;
  float* sums = base + total_chunks;
  for (int c = 0; c < 3; ++c) {
    if (tabs[c].ndescs == 0) continue;
    GradPassOpts o;
    o.norm = true;
    o.mp = use_mp_type != 0;
    o.partials = part;
    PRB_TRY(launch_grad_pass(classes[c], tabs[c], o, stream));
    NormGroup& g = a.g[a.ngroups++];
    g.partials = part;
    g.first_chunk = tabs[c].first_chunk;
    g.ntensors = tabs[c].ndescs;
    // fp16 and bf16 squared norms share one add_n list in the reference (clip.py:664-669); they
    // never coexist in practice, and here each keeps its own rounding.
    g.round_mode = (classes[c] == PRB_FLOAT32 || use_mp_type) ? 0 : classes[c];
    g.tensor_sums = sums;
    g.typed_out = nullptr;
    part += tabs[c].nchunks;
    sums += tabs[c].ndescs;
  }
  a.clip_norm = clip_norm;
  a.auto_skip_clip = auto_skip_clip;
  a.compute_clip = 1;
  {
    std::vector<int64_t> ns((size_t)num);
    for (int64_t i = 0; i < num; ++i) ns[(size_t)i] = grads[i].numel;
    a.grid_hint = finalize_grid_hint(ns.data(), num);
  }
  a.norm_out = norm_dev;
  a.clip_out = clip_dev;
  PRB_TRY(launch_norm_finalize(a, stream));
  for (int c = 0; c < 3; ++c)
    if (tabs[c].ndescs) PRB_TRY(launch_clip_mul(classes[c], tabs[c], clip_dev, stream));
  return PRB_SUCCESS;
}

int32_t prb_multiply_by_clip(PRB_Tensor* xs, int64_t num, const PRB_Tensor* clip, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  PRB_TRY(require_device());
  PRB_ENFORCE(num >= 0 && (xs || num == 0) && clip, PRB_INVALID_ARGUMENT, "multiply_by_clip: NULL argument");
  PRB_ENFORCE(clip->numel == 1 && clip->dtype == PRB_FLOAT32 && clip->device >= 0, PRB_INVALID_ARGUMENT,
              "multiply_by_clip: the factor must be one float32 on the device");
  const int32_t classes[3] = {PRB_FLOAT16, PRB_BFLOAT16, PRB_FLOAT32};
  for (int64_t i = 0; i < num; ++i) {
    PRB_ENFORCE(xs[i].dtype == PRB_FLOAT32 || xs[i].dtype == PRB_BFLOAT16 || xs[i].dtype == PRB_FLOAT16, PRB_UNIMPLEMENTED,
                "multiply_by_clip: dtype %s is not built", dtype_name(xs[i].dtype));
    PRB_ENFORCE(xs[i].device >= 0 || xs[i].numel == 0, PRB_INVALID_ARGUMENT, "multiply_by_clip: device tensors required");
  }
  for (int c = 0; c < 3; ++c) {
    DeviceTable tab;
    PRB_TRY(grad_table(xs, nullptr, num, classes[c], stream, &tab));
    if (tab.ndescs) PRB_TRY(launch_clip_mul(classes[c], tab, reinterpret_cast<const float*>(clip->data), stream));
  }
  return PRB_SUCCESS;
}

int32_t prb_update_loss_scaling(const PRB_Tensor* xs, int64_t num, const PRB_Tensor* found_infinite,
                                const PRB_Tensor* prev_loss_scaling, const PRB_Tensor* in_good_steps,
                                const PRB_Tensor* in_bad_steps, int32_t incr_every_n_steps,
                                int32_t decr_every_n_nan_or_inf, float incr_ratio, float decr_ratio,
                                int32_t stop_update, PRB_Tensor* outs, PRB_Tensor* loss_scaling,
                                PRB_Tensor* out_good_steps, PRB_Tensor* out_bad_steps, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  PRB_TRY(require_device());
  PRB_ENFORCE(found_infinite && prev_loss_scaling && in_good_steps && in_bad_steps && loss_scaling && out_good_steps &&
                  out_bad_steps,
              PRB_INVALID_ARGUMENT, "update_loss_scaling: NULL argument");
  PRB_ENFORCE(found_infinite->numel == 1, PRB_INVALID_ARGUMENT, "FoundInfinite must has only one element.");
  PRB_ENFORCE(found_infinite->device >= 0 && prev_loss_scaling->device >= 0 && in_good_steps->device >= 0 &&
                  in_bad_steps->device >= 0,
              PRB_INVALID_ARGUMENT, "update_loss_scaling: all state must live on the device");
  const bool* flag = reinterpret_cast<const bool*>(found_infinite->data);
  if (num > 0) {
    // LazyZeros: outs = 0 when found_inf (ref impl/amp_kernel_impl.h:127-136)
    for (int c = 0; c < 2; ++c) {
      const int es = c == 0 ? 4 : 2;
      std::vector<GradDesc> descs;
      std::vector<int64_t> numels;
      for (int64_t i = 0; i < num; ++i) {
        if (dtype_size(xs[i].dtype) != es) continue;
        GradDesc d{};
        d.x = xs[i].data;
        d.out = outs ? outs[i].data : xs[i].data;
        descs.push_back(d);
        numels.push_back(xs[i].numel);
      }
      if (descs.empty()) continue;
      DeviceTable tab;
      PRB_TRY(get_table(descs.data(), sizeof(GradDesc), (int)descs.size(), numels.data(), grad_chunk_elems(), stream, &tab));
      if (tab.nchunks == 0) continue;
      const GradDesc* d = reinterpret_cast<const GradDesc*>(tab.descs);
      if (es == 4) fill_zero_if_kernel<uint32_t><<<tab.nchunks, kGradThreads, 0, stream>>>(d, tab.chunks, flag);
      else fill_zero_if_kernel<uint16_t><<<tab.nchunks, kGradThreads, 0, stream>>>(d, tab.chunks, flag);
      PRB_LAUNCH_CHECK();
    }
  }
  if (stop_update) return PRB_SUCCESS;
  update_loss_scaling_kernel<<<1, 1, 0, stream>>>(
      flag, reinterpret_cast<const float*>(prev_loss_scaling->data), reinterpret_cast<const int*>(in_good_steps->data),
      reinterpret_cast<const int*>(in_bad_steps->data), incr_every_n_steps, decr_every_n_nan_or_inf, incr_ratio,
      decr_ratio, reinterpret_cast<float*>(loss_scaling->data), reinterpret_cast<int*>(out_good_steps->data),
      reinterpret_cast<int*>(out_bad_steps->data));
  PRB_LAUNCH_CHECK();
  return PRB_SUCCESS;
}

}  // extern "C"
