// One multi-tensor pass over gradients: AMP unscale + inf/nan check + squared-norm partials.
//
//   ref: CheckFiniteAndUnscale            paddle/phi/kernels/gpu/amp_kernel.cu:33-82
//        SquaredL2NormKernel              paddle/phi/kernels/gpu/squared_l2_norm_kernel.cu:23-35
//        ClipGradByGlobalNorm glue        python/paddle/nn/clip.py:643-736
//
// Norm reduction = warp shuffles inside each CTA (fixed shape) + ONE grid-level pass (a single CTA
// that sums the per-chunk partials in chunk order).  No float atomics anywhere, so results are
// bit-identical run to run (ref: test_squared_l2_norm_op.py:98-111).
#pragma once
#include "prb_plan.h"

namespace prb {

struct alignas(16) GradDesc {
  const void* x;
  void* out;        // where the unscaled value is stored (kWrite) -- normally == x
  int32_t aligned;
  int32_t pad0_;
  int64_t pad1_;
};
static_assert(sizeof(GradDesc) == 32, "GradDesc layout");

int grad_chunk_elems();      // elements per CTA of the active grad-pass launch geometry

// write: store T(MT(x) * inv) to out; norm: accumulate squares into partials[chunk];
// mp: squares kept in fp32 (clip.py:195-202) instead of rounded to T.
// loss_scale == nullptr: no unscale and no finite check.  found_inf must have been cleared.
int32_t launch_grad_pass(int32_t dtype, const DeviceTable& tab, bool write, bool norm, bool mp,
                         const float* loss_scale, bool* found_inf, float* partials, cudaStream_t stream);

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
};

int32_t launch_norm_finalize(const NormFinalizeArgs& a, cudaStream_t stream);

// g = T(MT(g) * MT(T(*clip))) in place; the whole launch is a no-op when *clip == 1.0f.
int32_t launch_clip_mul(int32_t dtype, const DeviceTable& tab, const float* clip, cudaStream_t stream);

}  // namespace prb
