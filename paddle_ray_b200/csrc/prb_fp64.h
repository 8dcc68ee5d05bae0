// float64 flavour of the operator entry points.  The reference registers `double` for adamw / adam / merged_adam /
// fused_adam / scale / check_finite_and_unscale / squared_l2_norm (ref adamw_kernel.cu:347-352,
// amp_kernel.cu:352-368, scale_kernel.cu:70-75, squared_l2_norm_kernel.cu:38-43); there T = MT = double, so the
// moments, the learning rate, the beta-pow accumulators and the loss scale are double tensors too
// (ref amp_type_traits.h:24-39).  Not on the north-star path (fp32 / bf16): plain grid-stride kernels, one launch
// per tensor, same expression trees with every operation individually rounded.
#pragma once
#include "prb_common.cuh"

namespace prb {

struct Adam64 {
  double* param;
  const double* grad;
  double* mom1;
  double* mom2;
  int64_t n;
  const double* lr_dev;     // device scalar, or nullptr and lr_host
  double lr_host;
  const double* b1pow_dev;  // device scalars, or nullptr and the host values
  const double* b2pow_dev;
  double b1pow_host, b2pow_host;
  double lr_ratio, beta1, beta2, epsilon, coeff;
  int use_decay;
  const bool* skip;         // optional device flag
};
int32_t launch_adam_f64(const Adam64& a, cudaStream_t stream);
// *b1_out = beta1 * *b1_in etc. on the device unless *skip
int32_t launch_beta_pow_f64(const double* b1_in, double* b1_out, const double* b2_in, double* b2_out, double beta1,
                            double beta2, const bool* skip, cudaStream_t stream);
int32_t launch_scale_f64(const double* x, double* out, int64_t n, double scale, double bias, bool bias_after_scale,
                         cudaStream_t stream);
// out = x * (1.0 / *scale); *found_inf |= !isfinite(out)   (found_inf cleared by the caller)
int32_t launch_unscale_f64(const double* x, double* out, int64_t n, const double* scale, bool* found_inf,
                           cudaStream_t stream);
// *out = sum of x*x in double: fixed grid, fixed-shape trees -> bit-identical run to run
int32_t launch_squared_l2_norm_f64(const double* x, int64_t n, double* out, cudaStream_t stream);

}  // namespace prb
