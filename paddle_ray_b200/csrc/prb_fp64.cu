// float64 operator kernels (see prb_fp64.h).
#include "prb_fp64.h"

namespace prb {
namespace {

constexpr int k64Threads = 256;

int grid_for(int64_t n) {
  const int64_t want = (n + k64Threads - 1) / k64Threads;
  const int64_t cap = (int64_t)sm_count() * 8;
  return (int)(want < 1 ? 1 : (want > cap ? cap : want));
}

// ref adamw_kernel.cu:32-79 / adam_kernel.cu:31-72 with T = MT = double
__global__ void __launch_bounds__(k64Threads) adam_f64_kernel(const Adam64 a) {
  if (a.skip != nullptr && *a.skip) return;
  const double lr = __dmul_rn(a.lr_dev ? *a.lr_dev : a.lr_host, a.lr_ratio);
  const double b1p = a.b1pow_dev ? *a.b1pow_dev : a.b1pow_host;
  const double b2p = a.b2pow_dev ? *a.b2pow_dev : a.b2pow_host;
  const double decay = __dsub_rn(1.0, __dmul_rn(lr, a.coeff));
  const double c1 = __dsub_rn(1.0, a.beta1), c2 = __dsub_rn(1.0, a.beta2);
  const double bc2 = __dsqrt_rn(__dsub_rn(1.0, b2p));
  const double step = -__ddiv_rn(lr, __dsub_rn(1.0, b1p));
  const int64_t stride = (int64_t)gridDim.x * k64Threads;
  for (int64_t i = (int64_t)blockIdx.x * k64Threads + threadIdx.x; i < a.n; i += stride) {
    double p = a.param[i], m1 = a.mom1[i], m2 = a.mom2[i];
    const double g = a.grad[i];
    if (a.use_decay) p = __dmul_rn(p, decay);
    m1 = __dadd_rn(__dmul_rn(a.beta1, m1), __dmul_rn(c1, g));
    m2 = __dadd_rn(__dmul_rn(a.beta2, m2), __dmul_rn(__dmul_rn(c2, g), g));
    const double denom = __dadd_rn(__ddiv_rn(__dsqrt_rn(m2), bc2), a.epsilon);
    p = __dadd_rn(p, __dmul_rn(__ddiv_rn(m1, denom), step));
    a.mom1[i] = m1;
    a.mom2[i] = m2;
    a.param[i] = p;
  }
}

__global__ void beta_pow_f64_kernel(const double* b1_in, double* b1_out, const double* b2_in, double* b2_out,
                                    double beta1, double beta2, const bool* skip) {
  if (skip != nullptr && *skip) return;
  *b1_out = __dmul_rn(beta1, *b1_in);
  *b2_out = __dmul_rn(beta2, *b2_in);
}

// ref scale_kernel.cu:35-40 with MT = double
__global__ void __launch_bounds__(k64Threads) scale_f64_kernel(const double* __restrict__ x, double* __restrict__ out,
                                                               int64_t n, double scale, double bias, bool after) {
  const int64_t stride = (int64_t)gridDim.x * k64Threads;
  for (int64_t i = (int64_t)blockIdx.x * k64Threads + threadIdx.x; i < n; i += stride)
    out[i] = after ? __dadd_rn(__dmul_rn(scale, x[i]), bias) : __dmul_rn(scale, __dadd_rn(x[i], bias));
}

// ref amp_kernel.cu:27-82 with T = MT = double
__global__ void __launch_bounds__(k64Threads) unscale_f64_kernel(const double* __restrict__ x, double* __restrict__ out,
                                                                 int64_t n, const double* __restrict__ scale,
                                                                 bool* __restrict__ found_inf) {
  const double inv = __ddiv_rn(1.0, *scale);
  bool bad = false;
  const int64_t stride = (int64_t)gridDim.x * k64Threads;
  for (int64_t i = (int64_t)blockIdx.x * k64Threads + threadIdx.x; i < n; i += stride) {
    const double v = __dmul_rn(x[i], inv);
    out[i] = v;
    if (!isfinite(v)) bad = true;
  }
  if (__syncthreads_or(bad ? 1 : 0) && threadIdx.x == 0) *found_inf = true;
}

__device__ __forceinline__ double block_sum_f64(double v, double* smem /* k64Threads */) {
  smem[threadIdx.x] = v;
  __syncthreads();
  for (int s = k64Threads / 2; s > 0; s >>= 1) {
    if ((int)threadIdx.x < s) smem[threadIdx.x] = __dadd_rn(smem[threadIdx.x], smem[threadIdx.x + s]);
    __syncthreads();
  }
  return smem[0];
}

constexpr int kNorm64Blocks = 1024;  // fixed grid: the summation order never depends on the device

__global__ void __launch_bounds__(k64Threads) sqnorm_f64_partial_kernel(const double* __restrict__ x, int64_t n,
                                                                        double* __restrict__ partials) {
  __shared__ double sm[k64Threads];
  double acc = 0.0;
  const int64_t stride = (int64_t)gridDim.x * k64Threads;
  for (int64_t i = (int64_t)blockIdx.x * k64Threads + threadIdx.x; i < n; i += stride) acc = __dadd_rn(acc, __dmul_rn(x[i], x[i]));
  const double s = block_sum_f64(acc, sm);
  if (threadIdx.x == 0) partials[blockIdx.x] = s;
}

__global__ void __launch_bounds__(k64Threads) sqnorm_f64_final_kernel(const double* __restrict__ partials, int nparts,
                                                                      double* __restrict__ out) {
  __shared__ double sm[k64Threads];
  double acc = 0.0;
  for (int i = threadIdx.x; i < nparts; i += k64Threads) acc = __dadd_rn(acc, partials[i]);
  const double s = block_sum_f64(acc, sm);
  if (threadIdx.x == 0) *out = s;
}

}  // namespace

int32_t launch_adam_f64(const Adam64& a, cudaStream_t stream) {
  if (a.n <= 0) return PRB_SUCCESS;
  adam_f64_kernel<<<grid_for(a.n), k64Threads, 0, stream>>>(a);
  PRB_LAUNCH_CHECK();
  return PRB_SUCCESS;
}

int32_t launch_beta_pow_f64(const double* b1_in, double* b1_out, const double* b2_in, double* b2_out, double beta1,
                            double beta2, const bool* skip, cudaStream_t stream) {
  beta_pow_f64_kernel<<<1, 1, 0, stream>>>(b1_in, b1_out, b2_in, b2_out, beta1, beta2, skip);
  PRB_LAUNCH_CHECK();
  return PRB_SUCCESS;
}

int32_t launch_scale_f64(const double* x, double* out, int64_t n, double scale, double bias, bool bias_after_scale,
                         cudaStream_t stream) {
  if (n <= 0) return PRB_SUCCESS;
  scale_f64_kernel<<<grid_for(n), k64Threads, 0, stream>>>(x, out, n, scale, bias, bias_after_scale);
  PRB_LAUNCH_CHECK();
  return PRB_SUCCESS;
}

int32_t launch_unscale_f64(const double* x, double* out, int64_t n, const double* scale, bool* found_inf,
                           cudaStream_t stream) {
  if (n <= 0) return PRB_SUCCESS;
  unscale_f64_kernel<<<grid_for(n), k64Threads, 0, stream>>>(x, out, n, scale, found_inf);
  PRB_LAUNCH_CHECK();
  return PRB_SUCCESS;
}

int32_t launch_squared_l2_norm_f64(const double* x, int64_t n, double* out, cudaStream_t stream) {
  void* ws = nullptr;
  PRB_TRY(workspace(kNorm64Blocks * sizeof(double), &ws, /*slot=*/2, stream));
  double* partials = reinterpret_cast<double*>(ws);
  sqnorm_f64_partial_kernel<<<kNorm64Blocks, k64Threads, 0, stream>>>(x, n, partials);
  PRB_LAUNCH_CHECK();
  sqnorm_f64_final_kernel<<<1, k64Threads, 0, stream>>>(partials, kNorm64Blocks, out);
  PRB_LAUNCH_CHECK();
  return PRB_SUCCESS;
}

}  // namespace prb
