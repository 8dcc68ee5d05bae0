// prb_dp_step_tail: AMP unscale + inf/nan check + global-norm clip + Adam/AdamW in two passes.
//
// What the reference does per step, per gradient element:
//   check_finite_and_unscale (r+w) -> squared_l2_norm (r) -> multiply by clip (r+w, new tensor)
//   -> host branch on found_inf (D2H sync) -> adamw_ per tensor (r) -> clear_grad (w)
// Here:
//   pass 1  grad_pass_kernel   reads each gradient once: unscale in registers, finite check, squared
//                              norm partials; one single-CTA pass turns partials into the clip factor
//   pass 2  adam_chunks_kernel reads gradient + param/master + moments once, applies unscale and clip
//                              in registers (with the reference's intermediate roundings to T),
//                              predicated on found_inf on the device, optionally writes the final
//                              gradient back or zeroes it.
// Every parameter/moment byte crosses HBM once.  The global norm is a grid-wide dependency, so the check + norm
// cannot ride in the update pass; it rides in whatever kernel touches the gradient BEFORE the update instead: the
// reducer's pack (one rank) or the exchange kernel's slice owner (several ranks) leave found_inf and the norm
// partials behind (cfg->pass1_done), and pass 1 disappears.  Only a caller without a reducer in front pays the
// second read of the gradient (ncu: it comes from HBM, not from L2 -- the update pass streams 10x the L2 size).
#include <vector>

#include "prb_adam.cuh"
#include "prb_grad_pass.cuh"

using namespace prb;

extern "C" int32_t prb_dp_step_tail(const PRB_TailConfig* cfg, const PRB_Tensor* params, PRB_Tensor* grads,
                                    const PRB_Tensor* learning_rate, const PRB_Tensor* moments1,
                                    const PRB_Tensor* moments2, PRB_Tensor* beta1_pow, PRB_Tensor* beta2_pow,
                                    const PRB_Tensor* master_params, int64_t num, const float* lr_ratio,
                                    const uint8_t* with_decay, const PRB_Tensor* loss_scale, PRB_Tensor* found_inf,
                                    PRB_Tensor* global_norm_out, PRB_Tensor* clip_out, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  PRB_TRY(require_device());
  PRB_ENFORCE(cfg && params && grads && num > 0, PRB_INVALID_ARGUMENT, "dp_step_tail: a required argument is NULL");
  PRB_ENFORCE(cfg->stage >= 0 && cfg->stage <= 2, PRB_INVALID_ARGUMENT, "dp_step_tail: stage must be 0, 1 or 2");
  // A sharded step runs the tail as two calls: stage 1 (check + norm + clip factor over ALL gradients, no update)
  // and stage 2 (the update over the sub-tensors this rank owns, clip factor and skip flag taken as inputs).
  const bool check_only = cfg->stage == 1, update_only = cfg->stage == 2;
  PRB_ENFORCE(check_only || (learning_rate && moments1 && moments2 && beta1_pow && beta2_pow), PRB_INVALID_ARGUMENT,
              "dp_step_tail: a required argument is NULL");
  PRB_ENFORCE(check_only || !cfg->multi_precision || master_params, PRB_INVALID_ARGUMENT,
              "dp_step_tail: multi_precision requires master_params");
  if (!check_only) {
    PRB_ENFORCE(learning_rate->numel == 1 && learning_rate->dtype == PRB_FLOAT32 && learning_rate->device >= 0,
                PRB_INVALID_ARGUMENT, "dp_step_tail: learning_rate must be one float32 on the device");
    PRB_ENFORCE(beta1_pow->numel == 1 && beta2_pow->numel == 1 && beta1_pow->device >= 0 && beta2_pow->device >= 0,
                PRB_INVALID_ARGUMENT, "dp_step_tail: beta-pow accumulators must be single device floats");
  }
  if (update_only && cfg->use_clip)
    PRB_ENFORCE(clip_out && clip_out->numel == 1 && clip_out->dtype == PRB_FLOAT32 && clip_out->device >= 0,
                PRB_INVALID_ARGUMENT, "dp_step_tail: stage 2 with use_clip reads the clip factor from clip_out");
  if (cfg->param_to_grad_slot)
    PRB_ENFORCE(params[0].dtype == grads[0].dtype && !cfg->zero_grads && !cfg->write_back_grads, PRB_INVALID_ARGUMENT,
                "dp_step_tail: param_to_grad_slot needs gradients of the parameters' dtype and excludes zero_grads / "
                "write_back_grads");
  if (cfg->use_amp) {
    PRB_ENFORCE(loss_scale && loss_scale->numel == 1 && loss_scale->dtype == PRB_FLOAT32 && loss_scale->device >= 0,
                PRB_INVALID_ARGUMENT, "dp_step_tail: use_amp needs a float32 device loss_scale");
    PRB_ENFORCE(found_inf && found_inf->numel == 1 && found_inf->dtype == PRB_BOOL && found_inf->device >= 0,
                PRB_INVALID_ARGUMENT, "dp_step_tail: use_amp needs a bool device found_inf");
  }
  const int32_t dt_p = params[0].dtype, dt_g = grads[0].dtype;
  const bool master = cfg->multi_precision != 0 && !check_only;

  std::vector<AdamDesc> descs((size_t)num);
  std::vector<GradDesc> gdescs((size_t)num);
  std::vector<int64_t> numels((size_t)num);
  for (int64_t i = 0; i < num; ++i) {
    const PRB_Tensor &p = params[i], &g = grads[i];
    PRB_ENFORCE(p.dtype == dt_p && g.dtype == dt_g, PRB_INVALID_ARGUMENT,
                "dp_step_tail: tensor %lld: all params (and all grads) must share one dtype per call", (long long)i);
    PRB_ENFORCE(g.numel == p.numel && p.device >= 0 && g.device >= 0, PRB_INVALID_ARGUMENT,
                "dp_step_tail: tensor %lld: parameter and gradient must be device tensors of one size", (long long)i);
    numels[(size_t)i] = p.numel;
    GradDesc& gd = gdescs[(size_t)i];
    memset(&gd, 0, sizeof(gd));
    gd.x = g.data;
    gd.out = g.data;
    gd.aligned = ((uintptr_t)g.data % ((uintptr_t)dtype_size(dt_g) * kVec)) == 0 ? 1 : 0;
    if (check_only) continue;
    const PRB_Tensor &m1 = moments1[i], &m2 = moments2[i];
    PRB_ENFORCE(m1.numel == p.numel && m2.numel == p.numel && m1.dtype == PRB_FLOAT32 && m2.dtype == PRB_FLOAT32,
                PRB_INVALID_ARGUMENT, "dp_step_tail: tensor %lld: sizes differ or moments are not float32", (long long)i);
    PRB_ENFORCE(m1.device >= 0 && m2.device >= 0, PRB_INVALID_ARGUMENT, "dp_step_tail: device tensors required");
    AdamDesc& d = descs[(size_t)i];
    memset(&d, 0, sizeof(d));
    d.param = p.data;
    d.grad = g.data;
    d.mom1 = reinterpret_cast<float*>(m1.data);
    d.mom2 = reinterpret_cast<float*>(m2.data);
    d.master = master ? reinterpret_cast<float*>(master_params[i].data) : nullptr;
    d.grad_out = ((cfg->write_back_grads && !cfg->zero_grads) || cfg->param_to_grad_slot) ? g.data : nullptr;
    d.lr = reinterpret_cast<const float*>(learning_rate->data);
    d.b1pow = reinterpret_cast<const float*>(beta1_pow->data);
    d.b2pow = reinterpret_cast<const float*>(beta2_pow->data);
    d.lr_ratio = lr_ratio ? lr_ratio[i] : 1.0f;
    const bool decay = with_decay ? with_decay[i] != 0 : true;
    d.coeff = (cfg->use_adamw && decay) ? cfg->coeff : 0.0f;  // with_decay=false -> coeff 0 (ref adamw_kernel.cu:203-206)
    d.use_decay = cfg->use_adamw ? 1 : 0;
    const uintptr_t ap = (uintptr_t)dtype_size(dt_p) * kVec, ag = (uintptr_t)dtype_size(dt_g) * kVec;
    const bool al = ((uintptr_t)d.param % ap) == 0 && ((uintptr_t)d.grad % ag) == 0 && vec_aligned<float>(d.mom1) &&
                    vec_aligned<float>(d.mom2) && (!master || vec_aligned<float>(d.master));
    d.aligned = al ? 1 : 0;
  }

  const float* ls = cfg->use_amp ? reinterpret_cast<const float*>(loss_scale->data) : nullptr;
  // use_amp: cleared here, set by the check.  Otherwise a non-NULL flag is an input-only skip flag.
  bool* found = found_inf ? reinterpret_cast<bool*>(found_inf->data) : nullptr;
  if (found_inf)
    PRB_ENFORCE(found_inf->numel == 1 && found_inf->dtype == PRB_BOOL && found_inf->device >= 0, PRB_INVALID_ARGUMENT,
                "dp_step_tail: found_inf must be one bool on the device");
  const float* clip_dev = nullptr;

  // pass1_done: the producer of the gradients (the reducer's pack at one rank, the exchange kernel at several; see
  // prb_reducer_set_grad_epilogue) already cleared found_inf, ran the check and left the norm partials in
  // cfg->pass1_partials, so the gradient is not read a second time here.
  const bool pass1_done = cfg->pass1_done != 0;
  if (pass1_done)
    PRB_ENFORCE(!cfg->use_clip || cfg->pass1_partials != nullptr, PRB_INVALID_ARGUMENT,
                "dp_step_tail: pass1_done with use_clip needs pass1_partials");
  if (update_only) {
    if (cfg->use_clip) clip_dev = reinterpret_cast<const float*>(clip_out->data);
  } else if (cfg->use_amp && !pass1_done) {
    PRB_CUDA(cudaMemsetAsync(found, 0, 1, stream));
  }
  if (!update_only && (cfg->use_amp || cfg->use_clip)) {
    DeviceTable gt;
    PRB_TRY(get_table(gdescs.data(), sizeof(GradDesc), (int)num, numels.data(), grad_chunk_elems(), stream, &gt));
    float* partials = nullptr;
    float* scratch_clip = nullptr;
    if (cfg->use_clip) {
      if (pass1_done) {
        partials = cfg->pass1_partials;  // sized by prb_tail_partials_numel: [nchunks | num tensor sums | 2]
      } else {
        void* ws = nullptr;
        PRB_TRY(workspace(((size_t)gt.nchunks + (size_t)num + 2) * sizeof(float), &ws, /*slot=*/1, stream));
        partials = reinterpret_cast<float*>(ws);
      }
      scratch_clip = partials + gt.nchunks + num;
    }
    const bool mp = cfg->use_mp_type != 0;
    if (!pass1_done) {
      GradPassOpts o;
      o.norm = cfg->use_clip != 0;
      o.mp = mp;
      o.loss_scale = ls;
      o.found_inf = cfg->use_amp ? found : nullptr;
      o.partials = partials;
      PRB_TRY(launch_grad_pass(dt_g, gt, o, stream));
    }
    if (cfg->use_clip) {
      NormFinalizeArgs a{};
      a.ngroups = 1;
      a.g[0].partials = partials;
      a.g[0].first_chunk = gt.first_chunk;
      a.g[0].ntensors = (int32_t)num;
      a.g[0].round_mode = (dt_g == PRB_FLOAT32 || mp) ? 0 : dt_g;
      a.g[0].tensor_sums = partials + gt.nchunks;
      a.g[0].typed_out = nullptr;
      a.clip_norm = cfg->clip_norm;
      a.auto_skip_clip = cfg->auto_skip_clip;
      a.compute_clip = 1;
      a.grid_hint = finalize_grid_hint(numels.data(), num);
      a.pdl = 1;  // its predecessor in the stream is the producer of the partials (pack / grad pass) or an event wait
      a.norm_out = global_norm_out ? reinterpret_cast<float*>(global_norm_out->data) : nullptr;
      a.clip_out = clip_out ? reinterpret_cast<float*>(clip_out->data) : scratch_clip;
      PRB_TRY(launch_norm_finalize(a, stream));
      clip_dev = a.clip_out;
    }
  }

  if (check_only) return PRB_SUCCESS;

  DeviceTable tab;
  PRB_TRY(get_table(descs.data(), sizeof(AdamDesc), (int)num, numels.data(), adam_chunk_elems(), stream, &tab));
  AdamArgs a;
  memset(&a, 0, sizeof(a));
  a.descs = reinterpret_cast<const AdamDesc*>(tab.descs);
  a.chunks = tab.chunks;
  a.nchunks = tab.nchunks;
  a.beta1 = (float)cfg->beta1;
  a.beta2 = (float)cfg->beta2;
  a.epsilon = (float)cfg->epsilon;
  a.skip = found;
  a.loss_scale = ls;
  a.clip = clip_dev;
  a.zero_grad = cfg->zero_grads;
  a.param_to_grad = cfg->param_to_grad_slot;
  // beta-pow accumulators advance inside the update kernel (last-CTA ticket) unless found_inf
  void* tk = nullptr;
  PRB_TRY(ticket_counter(&tk, stream));
  if (!cfg->keep_beta_pow) {
    a.pow_b1 = reinterpret_cast<float*>(beta1_pow->data);
    a.pow_b2 = reinterpret_cast<float*>(beta2_pow->data);
    a.pow_ticket = reinterpret_cast<unsigned int*>(tk);
  }
  return launch_adam(dt_p, dt_g, master, a, stream);
}
