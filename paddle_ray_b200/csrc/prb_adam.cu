// adamw_ / adam_ / merged_adam_ / fused_adam_ : one multi-tensor kernel over a flat chunk table.
//
// Arithmetic follows the reference's GPU kernels operation for operation
//   ref: paddle/phi/kernels/gpu/adamw_kernel.cu:32-79, adam_kernel.cu:31-72,
//        fused_adam_kernel.cu:198-227
// with every floating-point operation individually rounded (__fmul_rn/__fadd_rn/__fdiv_rn/
// __fsqrt_rn never contract into FMA), which makes the result independent of compiler
// contraction choices and bit-identical to the CPU oracle (oracle/dp_tail_oracle.c).
//
// HBM traffic per element: fp32 r16+w12 = 28 B; bf16 param + fp32 master r14+w14 = 28 B
// (30 B with fp32 main-grad).  Each CTA owns one 1024-element chunk: 128 threads x one lane vector of 8
// elements per array, 256-bit (fp32) / 128-bit (16-bit) streaming loads, all loads issued before the
// first use, >= 8 CTAs per SM (profiles/r1_adam_variants.md: 6891 vs 6084 GB/s for fat CTAs).
#include <map>
#include <vector>

#include <type_traits>

#include "prb_adam.cuh"
#include "prb_fp64.h"

namespace prb {

namespace {

struct Coef {
  float beta1, beta2, c1, c2, eps, decay, bc2, step, inv, clip, b1p, b2p;
  bool use_decay, amp, clipon;
};

// AMP unscale, in place in the reference: out = T(MT(x) * inv)   (ref amp_kernel.cu:66-68).
// This is the value a reference user sees in p.grad after the step.
template <typename TG>
__device__ __forceinline__ float unscale_grad(float g, const Coef& k) {
  return k.amp ? round_through<TG>(__fmul_rn(g, k.inv)) : g;
}
// Global-norm clip: new_grad = g * clip.astype(g.dtype), computed in T, a NEW tensor in the
// reference (ref python/paddle/nn/clip.py:715-732) -- it feeds the optimizer but never replaces p.grad.
template <typename TG>
__device__ __forceinline__ float clip_grad(float g, const Coef& k) {
  return k.clipon ? round_through<TG>(__fmul_rn(g, k.clip)) : g;
}

__device__ __forceinline__ void adam_elem(float& p, float g, float& m1, float& m2, const Coef& k) {
  if (k.use_decay) p = __fmul_rn(p, k.decay);                                    // p *= 1 - lr*coeff
  adam_moments_and_step(p, g, m1, m2, k.beta1, k.beta2, k.c1, k.c2, k.bc2, k.eps, k.step);
}


__device__ __forceinline__ float ld_relaxed_gpu(const float* p) {
  float v;
  asm volatile("ld.relaxed.gpu.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}

template <typename TG>
__device__ __forceinline__ Coef make_coef(const AdamDesc* __restrict__ d, const AdamArgs& a) {
  Coef k;
  const float lr = __fmul_rn(__ldg(d->lr), d->lr_ratio);  // ref adamw_kernel.cu:50
  // The accumulators may be overwritten by the last CTA of THIS launch (advance_pows_once), so they are read with
  // coherent relaxed loads, never through the read-only path.
  const float b1p = ld_relaxed_gpu(d->b1pow), b2p = ld_relaxed_gpu(d->b2pow);
  k.b1p = b1p;
  k.b2p = b2p;
  k.beta1 = a.beta1;
  k.beta2 = a.beta2;
  k.eps = a.epsilon;
  k.c1 = __fsub_rn(1.0f, a.beta1);
  k.c2 = __fsub_rn(1.0f, a.beta2);
  k.use_decay = d->use_decay != 0;
  k.decay = __fsub_rn(1.0f, __fmul_rn(lr, d->coeff));
  k.bc2 = __fsqrt_rn(__fsub_rn(1.0f, b2p));
  k.step = -__fdiv_rn(lr, __fsub_rn(1.0f, b1p));
  k.amp = a.loss_scale != nullptr;
  k.inv = k.amp ? (float)(1.0 / (double)__ldg(a.loss_scale)) : 1.0f;  // ref amp_kernel.cu:29
  k.clipon = a.clip != nullptr;
  k.clip = k.clipon ? round_through<TG>(__ldg(a.clip)) : 1.0f;
  return k;
}

// Shared beta-pow accumulators advanced inside the update kernel (ref UpdateAdamWBetaPow
// adamw_kernel.cu:130-139), which removes the separate 1-thread launch.  The coefficient thread of each CTA loads
// the old values with coherent gpu-scope loads (make_coef) and then takes a ticket whose INCREMENT is computed
// from the loaded bits (their sign bits, always 0 for a product of betas): a true data dependency, so the atomic
// cannot be issued before both loads have returned -- at no cost, the atomic's latency still hides under the CTA's
// data loads.  (A fence in its place waits for the thread's data loads too and cost 20 % of the pass: r2
// measurement, BERT-base 0.46 -> 0.60 ms.)  The CTA that draws the last ticket therefore knows that every CTA
// holds the old values in registers, fences, and overwrites them.
__device__ __forceinline__ void advance_pows_once(const AdamArgs& a, float b1p, float b2p) {
  if (a.pow_ticket == nullptr) return;
  const unsigned int inc = 1u + ((__float_as_uint(b1p) | __float_as_uint(b2p)) >> 31);
  const unsigned int t = atomicAdd(a.pow_ticket, inc);
  if (t == gridDim.x - 1) {
    __threadfence();
    *a.pow_ticket = 0u;
    *a.pow_b1 = __fmul_rn(a.beta1, b1p);
    *a.pow_b2 = __fmul_rn(a.beta2, b2p);
  }
}

// What lands in grad_out: the unscaled gradient (what p.grad holds in the reference after a step), or, in a
// sharded step, the updated parameter on its way to the peers.
template <typename T, typename TG>
__device__ __forceinline__ void store_grad_or_param(TG* dst, float g, float p, int param_to_grad) {
  if constexpr (std::is_same<T, TG>::value) {
    *dst = param_to_grad ? Conv<T>::from_f32(p) : Conv<TG>::from_f32(g);
  } else {
    *dst = Conv<TG>::from_f32(g);
  }
}

template <typename T, typename TG, bool kMaster, int kAdamThreads, int kAdamUnroll, int kMinBlocks>
__global__ void __launch_bounds__(kAdamThreads, kMinBlocks) adam_chunks_kernel(const AdamArgs a) {
  // Per-tensor coefficients are fetched and derived by ONE thread (while the data loads of the
  // whole CTA are in flight) and broadcast through shared memory.
  __shared__ Coef sk;
  const bool skipped = a.skip != nullptr && *a.skip;
  if (skipped && !a.zero_grad) return;

  const int4 craw = __ldg(reinterpret_cast<const int4*>(a.chunks) + blockIdx.x);
  const int tensor = craw.x;
  const int n = craw.y;
  const int64_t offset = (int64_t)(((uint64_t)(uint32_t)craw.w << 32) | (uint32_t)craw.z);
  const AdamDesc* __restrict__ d = a.descs + tensor;

  T* __restrict__ p_ptr = reinterpret_cast<T*>(d->param) + offset;
  const TG* __restrict__ g_ptr = reinterpret_cast<const TG*>(d->grad) + offset;
  float* __restrict__ m1_ptr = d->mom1 + offset;
  float* __restrict__ m2_ptr = d->mom2 + offset;
  float* __restrict__ mp_ptr = kMaster ? d->master + offset : nullptr;
  TG* __restrict__ go_ptr = d->grad_out ? reinterpret_cast<TG*>(d->grad_out) + offset : nullptr;
  const bool aligned = d->aligned != 0;

  if (skipped) {
    // found_inf: the update is skipped, but the folded-in clear_grad still zeroes the gradient
    TG* __restrict__ gz = const_cast<TG*>(g_ptr);
    for (int e = threadIdx.x; e < n; e += kAdamThreads) gz[e] = Conv<TG>::from_f32(0.0f);
    return;
  }

  if (aligned) {
    const int nvec = n / kVec;
    Pack<T> pv[kAdamUnroll];
    Pack<float> mpv[kAdamUnroll];
    Pack<TG> gv[kAdamUnroll];
    Pack<float> m1v[kAdamUnroll], m2v[kAdamUnroll];
    // ---- load phase: every load of this thread is in flight before the first use
#pragma unroll
    for (int u = 0; u < kAdamUnroll; ++u) {
      const int v = u * kAdamThreads + threadIdx.x;
      if (v < nvec) {
        const int e = v * kVec;
        if (kMaster) load_vec<float>(mp_ptr + e, mpv[u]);
        else load_vec<T>(p_ptr + e, pv[u]);
        load_vec<TG>(g_ptr + e, gv[u]);
        load_vec<float>(m1_ptr + e, m1v[u]);
        load_vec<float>(m2_ptr + e, m2v[u]);
      }
    }
    if (threadIdx.x == 0) {
      sk = make_coef<TG>(d, a);
      advance_pows_once(a, sk.b1p, sk.b2p);
    }
    __syncthreads();
    const Coef k = sk;
    // ---- compute + store
#pragma unroll
    for (int u = 0; u < kAdamUnroll; ++u) {
      const int v = u * kAdamThreads + threadIdx.x;
      if (v < nvec) {
        const int e = v * kVec;
        Pack<T> po;
        Pack<TG> go;
#pragma unroll
        for (int j = 0; j < kVec; ++j) {
          float p = kMaster ? mpv[u].v[j] : Conv<T>::to_f32(pv[u].v[j]);
          const float gu = unscale_grad<TG>(Conv<TG>::to_f32(gv[u].v[j]), k);
          adam_elem(p, clip_grad<TG>(gu, k), m1v[u].v[j], m2v[u].v[j], k);
          po.v[j] = Conv<T>::from_f32(p);
          if (kMaster) mpv[u].v[j] = p;
          go.v[j] = a.zero_grad ? Conv<TG>::from_f32(0.0f) : Conv<TG>::from_f32(gu);
        }
        store_vec<float>(m1_ptr + e, m1v[u]);
        store_vec<float>(m2_ptr + e, m2v[u]);
        store_vec<T>(p_ptr + e, po);
        if (kMaster) store_vec<float>(mp_ptr + e, mpv[u]);
        if (a.zero_grad) store_vec<TG>(const_cast<TG*>(g_ptr) + e, go);
        else if (go_ptr) {
          if constexpr (std::is_same<T, TG>::value) {
            if (a.param_to_grad) store_vec<T>(reinterpret_cast<T*>(go_ptr) + e, po);
            else store_vec<TG>(go_ptr + e, go);
          } else {
            store_vec<TG>(go_ptr + e, go);
          }
        }
      }
    }
    // ---- ragged tail of the tensor's last chunk (< kVec elements)
    const int e = nvec * kVec + threadIdx.x;
    if (e < n) {
      float p = kMaster ? mp_ptr[e] : Conv<T>::to_f32(p_ptr[e]);
      const float g = unscale_grad<TG>(Conv<TG>::to_f32(g_ptr[e]), k);
      float m1 = m1_ptr[e], m2 = m2_ptr[e];
      adam_elem(p, clip_grad<TG>(g, k), m1, m2, k);
      m1_ptr[e] = m1;
      m2_ptr[e] = m2;
      p_ptr[e] = Conv<T>::from_f32(p);
      if (kMaster) mp_ptr[e] = p;
      if (a.zero_grad) const_cast<TG*>(g_ptr)[e] = Conv<TG>::from_f32(0.0f);
      else if (go_ptr) store_grad_or_param<T, TG>(go_ptr + e, g, p, a.param_to_grad);
    }
  } else {
    // Unaligned tensor (caller passed a pointer that is not 32-/16-byte aligned): coalesced
    // element-wise accesses, same arithmetic.
    if (threadIdx.x == 0) {
      sk = make_coef<TG>(d, a);
      advance_pows_once(a, sk.b1p, sk.b2p);
    }
    __syncthreads();
    const Coef k = sk;
    for (int e = threadIdx.x; e < n; e += kAdamThreads) {
      float p = kMaster ? mp_ptr[e] : Conv<T>::to_f32(p_ptr[e]);
      const float g = unscale_grad<TG>(Conv<TG>::to_f32(g_ptr[e]), k);
      float m1 = m1_ptr[e], m2 = m2_ptr[e];
      adam_elem(p, clip_grad<TG>(g, k), m1, m2, k);
      m1_ptr[e] = m1;
      m2_ptr[e] = m2;
      p_ptr[e] = Conv<T>::from_f32(p);
      if (kMaster) mp_ptr[e] = p;
      if (a.zero_grad) const_cast<TG*>(g_ptr)[e] = Conv<TG>::from_f32(0.0f);
      else if (go_ptr) store_grad_or_param<T, TG>(go_ptr + e, g, p, a.param_to_grad);
    }
  }

}

struct PowPair {
  const float* b1_in;
  float* b1_out;
  const float* b2_in;
  float* b2_out;
};

// ref: UpdateAdamWBetaPow adamw_kernel.cu:130-139, UpdateBetaPowGroup fused_adam_kernel.cu:230-238
__global__ void beta_pow_kernel(const PowPair* pairs, int n, float beta1, float beta2, const bool* skip) {
  if (skip != nullptr && *skip) return;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    const PowPair pp = pairs[i];
    *pp.b1_out = __fmul_rn(beta1, *pp.b1_in);
    *pp.b2_out = __fmul_rn(beta2, *pp.b2_in);
  }
}

// Launch geometry: 128 threads x 1 lane vector, >= 8 CTAs/SM -- the best of the six shapes measured in round 1
// (profiles/r1_adam_variants.md); the other five were dropped from the binary.
constexpr int kAdamThreads = 128, kAdamUnroll = 1, kAdamMinBlocks = 8;

template <typename T, typename TG>
int32_t launch_typed(bool has_master, const AdamArgs& a, cudaStream_t stream) {
  if (a.nchunks == 0) return PRB_SUCCESS;
  KernelTimer kt(KK_ADAM, stream);
  if (has_master)
    adam_chunks_kernel<T, TG, true, kAdamThreads, kAdamUnroll, kAdamMinBlocks><<<a.nchunks, kAdamThreads, 0, stream>>>(a);
  else
    adam_chunks_kernel<T, TG, false, kAdamThreads, kAdamUnroll, kAdamMinBlocks><<<a.nchunks, kAdamThreads, 0, stream>>>(a);
  PRB_LAUNCH_CHECK();
  return PRB_SUCCESS;
}

}  // namespace

int adam_chunk_elems() { return kAdamThreads * kVec * kAdamUnroll; }

int32_t launch_adam(int32_t dtype_param, int32_t dtype_grad, bool has_master, const AdamArgs& a, cudaStream_t stream) {
  if (dtype_param == PRB_FLOAT32 && dtype_grad == PRB_FLOAT32) return launch_typed<float, float>(has_master, a, stream);
  if (dtype_param == PRB_BFLOAT16 && dtype_grad == PRB_BFLOAT16)
    return launch_typed<__nv_bfloat16, __nv_bfloat16>(has_master, a, stream);
  if (dtype_param == PRB_BFLOAT16 && dtype_grad == PRB_FLOAT32)
    return launch_typed<__nv_bfloat16, float>(has_master, a, stream);
  if (dtype_param == PRB_FLOAT16 && dtype_grad == PRB_FLOAT16) return launch_typed<__half, __half>(has_master, a, stream);
  if (dtype_param == PRB_FLOAT16 && dtype_grad == PRB_FLOAT32) return launch_typed<__half, float>(has_master, a, stream);
  PRB_FAIL(PRB_UNIMPLEMENTED, "adam: unsupported dtype combination param=%s grad=%s (fp64 is not built)",
           dtype_name(dtype_param), dtype_name(dtype_grad));
}

// ------------------------------------------------------------------------------------------
// Host side shared by the four entry points
// ------------------------------------------------------------------------------------------
namespace {

struct ListArgs {
  const PRB_Tensor* param;
  const PRB_Tensor* grad;
  const PRB_Tensor* lr;        // one tensor (lr_stride 0) or a list (lr_stride 1)
  int lr_stride;
  const PRB_Tensor* mom1;
  const PRB_Tensor* mom2;
  const PRB_Tensor* b1pow;     // list
  const PRB_Tensor* b2pow;     // list
  bool pow_from_first;         // fused_adam: tensor 0's beta-pows for every tensor
  const PRB_Tensor* master;    // list or nullptr
  int64_t num;
  const PRB_Tensor* skip;      // optional
  float beta1, beta2, epsilon;
  double beta1_d, beta2_d, epsilon_d;  // the attributes as given (float64 kernels use them unrounded)
  const double* lr_ratio_d;    // float64 path: per tensor or nullptr
  const double* coeff_d;
  const float* lr_ratio;       // per tensor or nullptr (1.0)
  const float* coeff;          // per tensor or nullptr (0)
  int use_decay;               // decay line present
  bool multi_precision, use_global_beta_pow;
  PRB_Tensor* param_out;
  PRB_Tensor* mom1_out;
  PRB_Tensor* mom2_out;
  PRB_Tensor* b1pow_out;
  PRB_Tensor* b2pow_out;
  PRB_Tensor* master_out;
  const char* op;
};

int32_t copy_if_different(const PRB_Tensor& src, const PRB_Tensor* dst, cudaStream_t stream, const char* what,
                          const char* op) {
  // ref: CopyTensorIfDifferent fused_adam_kernel.cu:240-252 -- outputs normally alias inputs.
  if (dst == nullptr || dst->data == src.data || src.numel == 0) return PRB_SUCCESS;
  PRB_ENFORCE(dst->numel == src.numel && dst->dtype == src.dtype, PRB_INVALID_ARGUMENT,
              "%s: %s_out must match %s in numel and dtype", op, what, what);
  const size_t bytes = (size_t)src.numel * dtype_size(src.dtype);
  const cudaMemcpyKind kind = src.device < 0 ? (dst->device < 0 ? cudaMemcpyHostToHost : cudaMemcpyHostToDevice)
                                             : (dst->device < 0 ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice);
  PRB_CUDA(cudaMemcpyAsync(dst->data, src.data, bytes, kind, stream));
  return PRB_SUCCESS;
}

// float64 parameters (T = MT = double): moments, learning rate and beta-pows are double tensors too
// (ref adamw_kernel.cu:147-343 with MPDType = double).  One launch per tensor: not on the north-star path.
int32_t run_adam_list_f64(const ListArgs& A, cudaStream_t stream) {
  const int64_t n = A.num;
  PRB_ENFORCE(!A.multi_precision, PRB_INVALID_ARGUMENT, "%s: multi_precision has no meaning for float64 parameters", A.op);
  const bool pows_on_host = A.b1pow[0].device < 0;
  const bool* skip_dev = nullptr;
  if (A.skip != nullptr) {
    PRB_ENFORCE(A.skip->numel == 1, PRB_INVALID_ARGUMENT, "Input(SkipUpdate) size must be 1, but get %lld",
                (long long)A.skip->numel);
    if (A.skip->device < 0) {
      if (*reinterpret_cast<const bool*>(A.skip->data)) return PRB_SUCCESS;
    } else if (pows_on_host && !A.use_global_beta_pow) {
      bool h = false;  // host-resident accumulators must not advance on a skipped step (ref adamw_kernel.cu:177-187)
      PRB_CUDA(cudaMemcpyAsync(&h, A.skip->data, 1, cudaMemcpyDeviceToHost, stream));
      PRB_CUDA(cudaStreamSynchronize(stream));
      if (h) return PRB_SUCCESS;
    } else {
      skip_dev = reinterpret_cast<const bool*>(A.skip->data);
    }
  }
  for (int64_t i = 0; i < n; ++i) {
    const PRB_Tensor &p = A.param[i], &g = A.grad[i], &m1 = A.mom1[i], &m2 = A.mom2[i];
    const int64_t pi = A.pow_from_first ? 0 : i;
    const PRB_Tensor &lr = A.lr[i * A.lr_stride], &b1 = A.b1pow[pi], &b2 = A.b2pow[pi];
    PRB_ENFORCE(p.dtype == PRB_FLOAT64 && g.dtype == PRB_FLOAT64 && m1.dtype == PRB_FLOAT64 && m2.dtype == PRB_FLOAT64 &&
                    lr.dtype == PRB_FLOAT64 && b1.dtype == PRB_FLOAT64 && b2.dtype == PRB_FLOAT64,
                PRB_INVALID_ARGUMENT,
                "%s: float64 parameters need float64 grad, moments, learning_rate and beta-pows (MPType = double)", A.op);
    PRB_ENFORCE(g.numel == p.numel && m1.numel == p.numel && m2.numel == p.numel, PRB_INVALID_ARGUMENT,
                "%s: tensor %lld: param/grad/moment sizes differ", A.op, (long long)i);
    PRB_ENFORCE(p.device >= 0 && g.device >= 0 && m1.device >= 0 && m2.device >= 0, PRB_INVALID_ARGUMENT,
                "%s: param/grad/moments must be device tensors (there is no CPU path)", A.op);
    PRB_ENFORCE(lr.numel == 1 && b1.numel == 1 && b2.numel == 1, PRB_INVALID_ARGUMENT,
                "%s: learning_rate and beta-pows must hold one element", A.op);
    PRB_ENFORCE((b1.device < 0) == pows_on_host && (b2.device < 0) == pows_on_host, PRB_INVALID_ARGUMENT,
                "All Beta1Pow/Beta2Pow must be in the same place.");
    PRB_TRY(copy_if_different(p, A.param_out ? &A.param_out[i] : nullptr, stream, "param", A.op));
    PRB_TRY(copy_if_different(m1, A.mom1_out ? &A.mom1_out[i] : nullptr, stream, "moment1", A.op));
    PRB_TRY(copy_if_different(m2, A.mom2_out ? &A.mom2_out[i] : nullptr, stream, "moment2", A.op));
    Adam64 a{};
    a.param = reinterpret_cast<double*>((A.param_out ? A.param_out[i] : p).data);
    a.grad = reinterpret_cast<const double*>(g.data);
    a.mom1 = reinterpret_cast<double*>((A.mom1_out ? A.mom1_out[i] : m1).data);
    a.mom2 = reinterpret_cast<double*>((A.mom2_out ? A.mom2_out[i] : m2).data);
    a.n = p.numel;
    if (lr.device < 0) a.lr_host = *reinterpret_cast<const double*>(lr.data);
    else a.lr_dev = reinterpret_cast<const double*>(lr.data);
    if (pows_on_host) {
      a.b1pow_host = *reinterpret_cast<const double*>(b1.data);
      a.b2pow_host = *reinterpret_cast<const double*>(b2.data);
    } else {
      a.b1pow_dev = reinterpret_cast<const double*>(b1.data);
      a.b2pow_dev = reinterpret_cast<const double*>(b2.data);
    }
    a.lr_ratio = A.lr_ratio_d ? A.lr_ratio_d[i] : 1.0;
    a.beta1 = A.beta1_d;
    a.beta2 = A.beta2_d;
    a.epsilon = A.epsilon_d;
    a.coeff = A.coeff_d ? A.coeff_d[i] : 0.0;
    a.use_decay = A.use_decay;
    a.skip = skip_dev;
    PRB_TRY(launch_adam_f64(a, stream));
  }
  if (A.use_global_beta_pow) return PRB_SUCCESS;
  for (int64_t i = 0; i < n; ++i) {  // ref adamw_kernel.cu:283-289 (host) / :332-341 (device)
    const PRB_Tensor* o1 = A.b1pow_out ? &A.b1pow_out[i] : &A.b1pow[i];
    const PRB_Tensor* o2 = A.b2pow_out ? &A.b2pow_out[i] : &A.b2pow[i];
    bool seen = false;  // an accumulator shared by several parameters advances once per step
    for (int64_t j = 0; j < i && !seen; ++j) seen = (A.b1pow_out ? A.b1pow_out[j].data : A.b1pow[j].data) == o1->data;
    if (seen) continue;
    if (pows_on_host) {
      *reinterpret_cast<double*>(o1->data) = A.beta1_d * *reinterpret_cast<const double*>(A.b1pow[i].data);
      *reinterpret_cast<double*>(o2->data) = A.beta2_d * *reinterpret_cast<const double*>(A.b2pow[i].data);
    } else {
      PRB_TRY(launch_beta_pow_f64(reinterpret_cast<const double*>(A.b1pow[i].data), reinterpret_cast<double*>(o1->data),
                                  reinterpret_cast<const double*>(A.b2pow[i].data), reinterpret_cast<double*>(o2->data),
                                  A.beta1_d, A.beta2_d, skip_dev, stream));
    }
  }
  return PRB_SUCCESS;
}

int32_t run_adam_list(const ListArgs& A, cudaStream_t stream) {
  PRB_TRY(require_device());
  const int64_t n = A.num;
  PRB_ENFORCE(n >= 0, PRB_INVALID_ARGUMENT, "%s: negative tensor count", A.op);
  if (n == 0) return PRB_SUCCESS;
  PRB_ENFORCE(A.param && A.grad && A.lr && A.mom1 && A.mom2 && A.b1pow && A.b2pow, PRB_INVALID_ARGUMENT,
              "%s: a required input is NULL", A.op);
  PRB_ENFORCE(!A.multi_precision || A.master != nullptr, PRB_INVALID_ARGUMENT,
              "%s: multi_precision requires master_param", A.op);
  if (A.param[0].dtype == PRB_FLOAT64) return run_adam_list_f64(A, stream);

  // ---- skip_update (ref adamw_kernel.cu:177-201, fused_adam_kernel.cu:325-343)
  const bool* skip_dev = nullptr;
  const bool pows_on_host = A.b1pow[0].device < 0;
  if (A.skip != nullptr) {
    PRB_ENFORCE(A.skip->numel == 1, PRB_INVALID_ARGUMENT, "Input(SkipUpdate) size must be 1, but get %lld",
                (long long)A.skip->numel);
    if (A.skip->device < 0) {
      if (*reinterpret_cast<const bool*>(A.skip->data)) return PRB_SUCCESS;  // copy-through == no-op in place
    } else if (pows_on_host && !A.use_global_beta_pow) {
      // Host-resident accumulators must not advance on a skipped step: read the flag back, exactly
      // like the reference's TensorToVector.  Keep beta-pows on the device to stay sync-free.
      bool h = false;
      PRB_CUDA(cudaMemcpyAsync(&h, A.skip->data, 1, cudaMemcpyDeviceToHost, stream));
      PRB_CUDA(cudaStreamSynchronize(stream));
      if (h) return PRB_SUCCESS;
    } else {
      skip_dev = reinterpret_cast<const bool*>(A.skip->data);
    }
  }

  const int32_t dt_p = A.param[0].dtype, dt_g = A.grad[0].dtype;
  const bool has_master = A.multi_precision;

  // ---- validate + host scalars
  int n_host_scalars = 0;
  for (int64_t i = 0; i < n; ++i) {
    const PRB_Tensor &p = A.param[i], &g = A.grad[i], &m1 = A.mom1[i], &m2 = A.mom2[i];
    PRB_ENFORCE(p.dtype == dt_p && g.dtype == dt_g, PRB_INVALID_ARGUMENT, "%s: tensor %lld has a different dtype", A.op,
                (long long)i);
    PRB_ENFORCE(g.numel == p.numel && m1.numel == p.numel && m2.numel == p.numel, PRB_INVALID_ARGUMENT,
                "%s: tensor %lld: param/grad/moment sizes differ", A.op, (long long)i);
    PRB_ENFORCE(m1.dtype == PRB_FLOAT32 && m2.dtype == PRB_FLOAT32, PRB_INVALID_ARGUMENT,
                "%s: moments must be float32 (MPType)", A.op);
    PRB_ENFORCE(p.device >= 0 && g.device >= 0 && m1.device >= 0 && m2.device >= 0, PRB_INVALID_ARGUMENT,
                "%s: param/grad/moments must be device tensors (there is no CPU path)", A.op);
    if (has_master)
      PRB_ENFORCE(A.master[i].numel == p.numel && A.master[i].dtype == PRB_FLOAT32 && A.master[i].device >= 0,
                  PRB_INVALID_ARGUMENT, "%s: master_param %lld must be a float32 device tensor of param's size", A.op,
                  (long long)i);
    const int64_t pi = A.pow_from_first ? 0 : i;
    PRB_ENFORCE(A.b1pow[pi].numel == 1, PRB_INVALID_ARGUMENT,
                "beta1 pow output size should be 1, but received value is:%lld.", (long long)A.b1pow[pi].numel);
    PRB_ENFORCE(A.b2pow[pi].numel == 1, PRB_INVALID_ARGUMENT,
                "beta2 pow output size should be 1, but received value is:%lld.", (long long)A.b2pow[pi].numel);
    PRB_ENFORCE((A.b1pow[i].device < 0) == pows_on_host && (A.b2pow[i].device < 0) == pows_on_host,
                PRB_INVALID_ARGUMENT, "All Beta1Pow/Beta2Pow must be in the same place.");
    const PRB_Tensor& lr = A.lr[i * A.lr_stride];
    PRB_ENFORCE(lr.numel == 1 && lr.dtype == PRB_FLOAT32, PRB_INVALID_ARGUMENT,
                "%s: learning_rate must be one float32", A.op);
    if (lr.device < 0) n_host_scalars++;
  }
  if (pows_on_host) n_host_scalars += 2 * (int)n;

  // ---- outputs that do not alias their inputs
  for (int64_t i = 0; i < n; ++i) {
    PRB_TRY(copy_if_different(A.param[i], A.param_out ? &A.param_out[i] : nullptr, stream, "param", A.op));
    PRB_TRY(copy_if_different(A.mom1[i], A.mom1_out ? &A.mom1_out[i] : nullptr, stream, "moment1", A.op));
    PRB_TRY(copy_if_different(A.mom2[i], A.mom2_out ? &A.mom2_out[i] : nullptr, stream, "moment2", A.op));
    if (has_master)
      PRB_TRY(copy_if_different(A.master[i], A.master_out ? &A.master_out[i] : nullptr, stream, "master_param", A.op));
  }

  // ---- host-resident scalars (lr / beta-pows given as CPU tensors): staged into device scratch
  float* scal_dev = nullptr;
  std::vector<float> scal_host;
  if (n_host_scalars > 0) {
    void* ws = nullptr;
    PRB_TRY(workspace((size_t)(3 * n) * sizeof(float), &ws, /*slot=*/0, stream));
    scal_dev = reinterpret_cast<float*>(ws);
    scal_host.assign((size_t)3 * n, 0.0f);
  }

  // ---- descriptors
  std::vector<AdamDesc> descs((size_t)n);
  std::vector<int64_t> numels((size_t)n);
  for (int64_t i = 0; i < n; ++i) {
    AdamDesc& d = descs[(size_t)i];
    memset(&d, 0, sizeof(d));
    const PRB_Tensor* po = A.param_out ? &A.param_out[i] : &A.param[i];
    const PRB_Tensor* m1o = A.mom1_out ? &A.mom1_out[i] : &A.mom1[i];
    const PRB_Tensor* m2o = A.mom2_out ? &A.mom2_out[i] : &A.mom2[i];
    d.param = po->data;
    d.grad = A.grad[i].data;
    d.mom1 = reinterpret_cast<float*>(m1o->data);
    d.mom2 = reinterpret_cast<float*>(m2o->data);
    d.master = has_master ? reinterpret_cast<float*>((A.master_out ? A.master_out[i] : A.master[i]).data) : nullptr;
    d.grad_out = nullptr;
    const PRB_Tensor& lr = A.lr[i * A.lr_stride];
    if (lr.device < 0) {
      scal_host[(size_t)i] = *reinterpret_cast<const float*>(lr.data);
      d.lr = scal_dev + i;
    } else {
      d.lr = reinterpret_cast<const float*>(lr.data);
    }
    const int64_t pi = A.pow_from_first ? 0 : i;
    if (pows_on_host) {
      scal_host[(size_t)(n + i)] = *reinterpret_cast<const float*>(A.b1pow[pi].data);
      scal_host[(size_t)(2 * n + i)] = *reinterpret_cast<const float*>(A.b2pow[pi].data);
      d.b1pow = scal_dev + n + i;
      d.b2pow = scal_dev + 2 * n + i;
    } else {
      d.b1pow = reinterpret_cast<const float*>(A.b1pow[pi].data);
      d.b2pow = reinterpret_cast<const float*>(A.b2pow[pi].data);
    }
    d.lr_ratio = A.lr_ratio ? A.lr_ratio[i] : 1.0f;
    d.coeff = A.coeff ? A.coeff[i] : 0.0f;
    d.use_decay = A.use_decay;
    const bool al = (dt_p == PRB_FLOAT32 ? vec_aligned<float>(d.param) : vec_aligned<uint16_t>(d.param)) &&
                    (dt_g == PRB_FLOAT32 ? vec_aligned<float>(d.grad) : vec_aligned<uint16_t>(d.grad)) &&
                    vec_aligned<float>(d.mom1) && vec_aligned<float>(d.mom2) &&
                    (!has_master || vec_aligned<float>(d.master));
    d.aligned = al ? 1 : 0;
    numels[(size_t)i] = A.param[i].numel;
  }
  if (n_host_scalars > 0)
    PRB_CUDA(cudaMemcpyAsync(scal_dev, scal_host.data(), scal_host.size() * sizeof(float), cudaMemcpyHostToDevice,
                             stream));  // pageable source: staged before return

  DeviceTable tab;
  PRB_TRY(get_table(descs.data(), sizeof(AdamDesc), (int)n, numels.data(), adam_chunk_elems(), stream, &tab));

  AdamArgs a;
  memset(&a, 0, sizeof(a));
  a.descs = reinterpret_cast<const AdamDesc*>(tab.descs);
  a.chunks = tab.chunks;
  a.nchunks = tab.nchunks;
  a.beta1 = A.beta1;
  a.beta2 = A.beta2;
  a.epsilon = A.epsilon;
  a.skip = skip_dev;
  PRB_TRY(launch_adam(dt_p, dt_g, has_master, a, stream));

  // ---- beta-pow accumulators (ref adamw_kernel.cu:283-289 host, :332-341 device)
  if (!A.use_global_beta_pow) {
    if (pows_on_host) {
      for (int64_t i = 0; i < n; ++i) {
        const PRB_Tensor* o1 = A.b1pow_out ? &A.b1pow_out[i] : &A.b1pow[i];
        const PRB_Tensor* o2 = A.b2pow_out ? &A.b2pow_out[i] : &A.b2pow[i];
        bool seen = false;  // an accumulator shared by several parameters advances once per step
        for (int64_t j = 0; j < i && !seen; ++j)
          seen = (A.b1pow_out ? A.b1pow_out[j].data : A.b1pow[j].data) == o1->data;
        if (seen) continue;
        *reinterpret_cast<float*>(o1->data) = A.beta1 * *reinterpret_cast<const float*>(A.b1pow[i].data);
        *reinterpret_cast<float*>(o2->data) = A.beta2 * *reinterpret_cast<const float*>(A.b2pow[i].data);
      }
    } else {
      PRB_TRY(beta_pow_update_device(A.b1pow, A.b2pow, A.b1pow_out ? A.b1pow_out : nullptr,
                                     A.b2pow_out ? A.b2pow_out : nullptr, n, A.beta1, A.beta2, skip_dev, stream));
    }
  }
  return PRB_SUCCESS;
}

}  // namespace

int32_t beta_pow_update_device(const PRB_Tensor* b1_in, const PRB_Tensor* b2_in, const PRB_Tensor* b1_out,
                               const PRB_Tensor* b2_out, int64_t n, float beta1, float beta2, const bool* skip,
                               cudaStream_t stream) {
  std::vector<PowPair> pairs;
  std::map<const void*, bool> seen;
  for (int64_t i = 0; i < n; ++i) {
    PowPair pp;
    pp.b1_in = reinterpret_cast<const float*>(b1_in[i].data);
    pp.b2_in = reinterpret_cast<const float*>(b2_in[i].data);
    pp.b1_out = reinterpret_cast<float*>((b1_out ? b1_out[i] : b1_in[i]).data);
    pp.b2_out = reinterpret_cast<float*>((b2_out ? b2_out[i] : b2_in[i]).data);
    if (seen.count(pp.b1_out)) continue;  // an accumulator shared by several parameters advances once
    seen[pp.b1_out] = true;
    pairs.push_back(pp);
  }
  std::vector<int64_t> ones(pairs.size(), 1);
  DeviceTable pt;
  PRB_TRY(get_table(pairs.data(), sizeof(PowPair), (int)pairs.size(), ones.data(), 1, stream, &pt));
  const int np = (int)pairs.size();
  beta_pow_kernel<<<(np + 127) / 128, 128, 0, stream>>>(reinterpret_cast<const PowPair*>(pt.descs), np, beta1, beta2,
                                                        skip);
  PRB_LAUNCH_CHECK();
  return PRB_SUCCESS;
}

}  // namespace prb

using namespace prb;

extern "C" {

int32_t prb_adamw(const PRB_Tensor* param, const PRB_Tensor* grad, const PRB_Tensor* learning_rate,
                  const PRB_Tensor* moment1, const PRB_Tensor* moment2, const PRB_Tensor* beta1_pow,
                  const PRB_Tensor* beta2_pow, const PRB_Tensor* master_param, const PRB_Tensor* skip_update,
                  double beta1, double beta2, double epsilon, float lr_ratio, float coeff, int32_t with_decay,
                  int32_t lazy_mode, int64_t min_row_size_to_use_multithread, int32_t multi_precision,
                  int32_t use_global_beta_pow, PRB_Tensor* param_out, PRB_Tensor* moment1_out, PRB_Tensor* moment2_out,
                  PRB_Tensor* beta1_pow_out, PRB_Tensor* beta2_pow_out, PRB_Tensor* master_param_out, void* stream) {
  (void)lazy_mode;
  (void)min_row_size_to_use_multithread;
  const float coeff_eff = with_decay ? coeff : 0.0f;  // ref adamw_kernel.cu:203-206
  const double coeff_d = coeff_eff, ratio_d = lr_ratio;
  ListArgs A{};
  A.lr_ratio_d = &ratio_d; A.coeff_d = &coeff_d;
  A.param = param; A.grad = grad; A.lr = learning_rate; A.lr_stride = 0; A.mom1 = moment1; A.mom2 = moment2;
  A.b1pow = beta1_pow; A.b2pow = beta2_pow; A.pow_from_first = false; A.master = multi_precision ? master_param : nullptr;
  A.num = 1; A.skip = skip_update; A.beta1 = (float)beta1; A.beta2 = (float)beta2; A.epsilon = (float)epsilon;
  A.beta1_d = beta1; A.beta2_d = beta2; A.epsilon_d = epsilon;
  A.lr_ratio = &lr_ratio; A.coeff = &coeff_eff; A.use_decay = 1; A.multi_precision = multi_precision != 0;
  A.use_global_beta_pow = use_global_beta_pow != 0; A.param_out = param_out; A.mom1_out = moment1_out;
  A.mom2_out = moment2_out; A.b1pow_out = beta1_pow_out; A.b2pow_out = beta2_pow_out;
  A.master_out = multi_precision ? master_param_out : nullptr; A.op = "adamw";
  return run_adam_list(A, (cudaStream_t)stream);
}

int32_t prb_adam(const PRB_Tensor* param, const PRB_Tensor* grad, const PRB_Tensor* learning_rate,
                 const PRB_Tensor* moment1, const PRB_Tensor* moment2, const PRB_Tensor* beta1_pow,
                 const PRB_Tensor* beta2_pow, const PRB_Tensor* master_param, const PRB_Tensor* skip_update,
                 double beta1, double beta2, double epsilon, int32_t lazy_mode,
                 int64_t min_row_size_to_use_multithread, int32_t multi_precision, int32_t use_global_beta_pow,
                 PRB_Tensor* param_out, PRB_Tensor* moment1_out, PRB_Tensor* moment2_out, PRB_Tensor* beta1_pow_out,
                 PRB_Tensor* beta2_pow_out, PRB_Tensor* master_param_out, void* stream) {
  (void)lazy_mode;
  (void)min_row_size_to_use_multithread;
  ListArgs A{};
  A.param = param; A.grad = grad; A.lr = learning_rate; A.lr_stride = 0; A.mom1 = moment1; A.mom2 = moment2;
  A.b1pow = beta1_pow; A.b2pow = beta2_pow; A.pow_from_first = false; A.master = multi_precision ? master_param : nullptr;
  A.num = 1; A.skip = skip_update; A.beta1 = (float)beta1; A.beta2 = (float)beta2; A.epsilon = (float)epsilon;
  A.beta1_d = beta1; A.beta2_d = beta2; A.epsilon_d = epsilon;
  A.lr_ratio = nullptr; A.coeff = nullptr; A.use_decay = 0; A.multi_precision = multi_precision != 0;
  A.use_global_beta_pow = use_global_beta_pow != 0; A.param_out = param_out; A.mom1_out = moment1_out;
  A.mom2_out = moment2_out; A.b1pow_out = beta1_pow_out; A.b2pow_out = beta2_pow_out;
  A.master_out = multi_precision ? master_param_out : nullptr; A.op = "adam";
  return run_adam_list(A, (cudaStream_t)stream);
}

int32_t prb_merged_adam(const PRB_Tensor* param, const PRB_Tensor* grad, const PRB_Tensor* learning_rate,
                        const PRB_Tensor* moment1, const PRB_Tensor* moment2, const PRB_Tensor* beta1_pow,
                        const PRB_Tensor* beta2_pow, const PRB_Tensor* master_param, int64_t num, double beta1,
                        double beta2, double epsilon, int32_t multi_precision, int32_t use_global_beta_pow,
                        PRB_Tensor* param_out, PRB_Tensor* moment1_out, PRB_Tensor* moment2_out,
                        PRB_Tensor* beta1_pow_out, PRB_Tensor* beta2_pow_out, PRB_Tensor* master_param_out,
                        void* stream) {
  ListArgs A{};
  A.param = param; A.grad = grad; A.lr = learning_rate; A.lr_stride = 1; A.mom1 = moment1; A.mom2 = moment2;
  A.b1pow = beta1_pow; A.b2pow = beta2_pow; A.pow_from_first = false; A.master = multi_precision ? master_param : nullptr;
  A.num = num; A.skip = nullptr; A.beta1 = (float)beta1; A.beta2 = (float)beta2; A.epsilon = (float)epsilon;
  A.beta1_d = beta1; A.beta2_d = beta2; A.epsilon_d = epsilon;
  A.lr_ratio = nullptr; A.coeff = nullptr; A.use_decay = 0; A.multi_precision = multi_precision != 0;
  A.use_global_beta_pow = use_global_beta_pow != 0; A.param_out = param_out; A.mom1_out = moment1_out;
  A.mom2_out = moment2_out; A.b1pow_out = beta1_pow_out; A.b2pow_out = beta2_pow_out;
  A.master_out = multi_precision ? master_param_out : nullptr; A.op = "merged_adam";
  return run_adam_list(A, (cudaStream_t)stream);
}

int32_t prb_fused_adam(const PRB_Tensor* params, const PRB_Tensor* grads, const PRB_Tensor* learning_rate,
                       const PRB_Tensor* moments1, const PRB_Tensor* moments2, const PRB_Tensor* beta1_pows,
                       const PRB_Tensor* beta2_pows, const PRB_Tensor* master_params, int64_t num,
                       const PRB_Tensor* skip_update, double beta1, double beta2, double epsilon, int32_t chunk_size,
                       float weight_decay, int32_t use_adamw, int32_t multi_precision, int32_t use_global_beta_pow,
                       PRB_Tensor* params_out, PRB_Tensor* moments1_out, PRB_Tensor* moments2_out,
                       PRB_Tensor* beta1_pows_out, PRB_Tensor* beta2_pows_out, PRB_Tensor* master_params_out,
                       void* stream) {
  (void)chunk_size;  // partitioning is internal; it never changes per-element results
  std::vector<float> coeffs((size_t)(num > 0 ? num : 0), weight_decay);
  ListArgs A{};
  A.param = params; A.grad = grads; A.lr = learning_rate; A.lr_stride = 0; A.mom1 = moments1; A.mom2 = moments2;
  A.b1pow = beta1_pows; A.b2pow = beta2_pows; A.pow_from_first = true;  // ref fused_adam_kernel.cu:292-293,368-369
  A.master = multi_precision ? master_params : nullptr; A.num = num; A.skip = skip_update; A.beta1 = (float)beta1;
  A.beta2 = (float)beta2; A.epsilon = (float)epsilon; A.lr_ratio = nullptr; A.coeff = coeffs.data();
  A.beta1_d = beta1; A.beta2_d = beta2; A.epsilon_d = epsilon;
  std::vector<double> coeffs_d(coeffs.begin(), coeffs.end());
  A.coeff_d = coeffs_d.data();
  A.use_decay = use_adamw ? 1 : 0; A.multi_precision = multi_precision != 0;
  A.use_global_beta_pow = use_global_beta_pow != 0; A.param_out = params_out; A.mom1_out = moments1_out;
  A.mom2_out = moments2_out; A.b1pow_out = beta1_pows_out; A.b2pow_out = beta2_pows_out;
  A.master_out = multi_precision ? master_params_out : nullptr; A.op = "fused_adam";
  return run_adam_list(A, (cudaStream_t)stream);
}

}  // extern "C"
