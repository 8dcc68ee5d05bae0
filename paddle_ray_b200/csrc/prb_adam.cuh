// Multi-tensor Adam/AdamW over a chunk table: the one device function every optimizer entry point
// (adamw_, adam_, merged_adam_, fused_adam_, dp_step_tail) shares, so they are bit-identical to one
// another (ref parity: test_merged_adam_op.py:204-217, test_fused_adam_kernel.cc:453-459).
#pragma once
#include "prb_plan.h"

namespace prb {

// Per-tensor descriptor (device-resident, 16-byte aligned; read once per CTA through the
// read-only path).
struct alignas(16) AdamDesc {
  void* param;          // T*
  const void* grad;     // TG* (T or float)
  float* mom1;
  float* mom2;
  float* master;        // fp32 master weights or nullptr
  void* grad_out;       // TG* where the transformed gradient is written back (tail only) or nullptr
  const float* lr;      // device scalar
  const float* b1pow;   // device scalar
  const float* b2pow;   // device scalar
  float lr_ratio;
  float coeff;          // already 0 when with_decay == false (ref adamw_kernel.cu:203-206)
  int32_t use_decay;    // 1: AdamW decay line present (ref adamw_kernel.cu:61); 0: plain Adam
  int32_t aligned;      // 1: every pointer allows kVec-wide vector access
  int64_t pad_;
};
static_assert(sizeof(AdamDesc) == 96, "AdamDesc layout");

// Launch-wide arguments.
struct AdamArgs {
  const AdamDesc* descs;
  const Chunk* chunks;
  int32_t nchunks;
  float beta1, beta2, epsilon;
  const bool* skip;          // device flag: when set, the launch is a no-op (found_inf / skip_update)
  const float* loss_scale;   // AMP: g = T(MT(g) * inv), inv = float(1.0 / double(*loss_scale)); nullptr = off
  const float* clip;         // global-norm clip: g = T(MT(g) * MT(T(*clip))); nullptr = off
  int32_t zero_grad;         // 1: store zeros to grad after consuming it (clear_grad folded in)
  int32_t param_to_grad;     // 1: grad_out receives the updated PARAMETER (same dtype) instead of the unscaled
                             //    gradient: a sharded step ships it to the peers through the bucket arena
  // Shared beta-pow accumulators advanced by the CTA that draws the LAST ticket, which removes the separate
  // 1-thread launch: every CTA has read the old value (and fenced) before it takes its ticket.
  float* pow_b1;             // nullptr = off
  float* pow_b2;
  unsigned int* pow_ticket;  // zero-initialised device counter of the launching stream, reset by the last CTA
};

#ifdef __CUDACC__
// The Adam expression every optimizer kernel of the library shares, each operation individually rounded
// (ref adam_kernel.cu:31-72, adamw_kernel.cu:32-79, selected_rows/gpu/adam_kernel.cu:41-93):
//   m1 = b1*m1 + (1-b1)*g;  m2 = b2*m2 + (1-b2)*g*g;  p += (m1 / (sqrt(m2)/sqrt(1-b2^t) + eps)) * -(lr/(1-b1^t))
// c1 = 1-b1, c2 = 1-b2, bc2 = sqrt(1-b2^t), step = -(lr/(1-b1^t)).
__device__ __forceinline__ void adam_moments_and_step(float& p, float g, float& m1, float& m2, float beta1, float beta2,
                                                      float c1, float c2, float bc2, float eps, float step) {
  m1 = __fadd_rn(__fmul_rn(beta1, m1), __fmul_rn(c1, g));
  m2 = __fadd_rn(__fmul_rn(beta2, m2), __fmul_rn(__fmul_rn(c2, g), g));
  const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(m2), bc2), eps);
  p = __fadd_rn(p, __fmul_rn(__fdiv_rn(m1, denom), step));
}
#endif

// Elements per CTA of the active launch geometry (threads * kVec * lane-vectors per thread).
int adam_chunk_elems();

int32_t launch_adam(int32_t dtype_param, int32_t dtype_grad, bool has_master, const AdamArgs& args,
                    cudaStream_t stream);

// Device-resident beta-pow accumulators: out = beta * in, once per distinct accumulator, unless
// *skip (ref UpdateAdamWBetaPow adamw_kernel.cu:130-139, UpdateBetaPowGroup fused_adam_kernel.cu:230-238).
int32_t beta_pow_update_device(const PRB_Tensor* b1_in, const PRB_Tensor* b2_in, const PRB_Tensor* b1_out,
                               const PRB_Tensor* b2_out, int64_t n, float beta1, float beta2, const bool* skip,
                               cudaStream_t stream);

}  // namespace prb
