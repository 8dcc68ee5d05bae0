/*
 * paddle_ray_b200.h -- C ABI of the B200-native data-parallel step tail.
 *
 * Every entry point replaces one operator / binding of the reference on the path
 *   EagerReducer bucket sync -> check_finite_and_unscale -> squared_l2_norm clip -> adam(w)
 * and keeps that operator's argument order and meaning.  Citations ("ref:") are relative
 * to the reference tree (Rayman96/Paddle_ray).
 *
 * Conventions
 *  - Tensors are passed as PRB_Tensor {data, numel, dtype, device}; lists as pointer+count;
 *    optional tensors as NULL.  dtype codes are phi::DataType's
 *    (ref: paddle/phi/common/data_type.h:37-70).  device = -1 means host memory,
 *    >= 0 a CUDA device ordinal.
 *  - `Scalar` attributes are double, `float` attributes float, bool attributes int32_t.
 *  - In-place outputs alias inputs exactly as the YAML `inplace:` maps of the reference say;
 *    a distinct output buffer is honoured by a device copy first.
 *  - All work is enqueued on `stream` (a cudaStream_t passed as void*); nothing synchronises
 *    the device unless documented.  The caller owns every buffer.
 *  - Every function returns a PRB_Status; prb_last_error_message() returns the thread-local
 *    text of the last failure.  Nothing throws across the ABI.
 *  - There is no CPU fallback: on a machine without a usable CUDA device every compute entry
 *    point fails with PRB_EXTERNAL.
 */
#ifndef PADDLE_RAY_B200_H_
#define PADDLE_RAY_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ref: paddle/phi/backends/device_ext.h:55-62 (C_Status) + PADDLE_ENFORCE categories */
typedef enum PRB_Status {
  PRB_SUCCESS = 0,
  PRB_INVALID_ARGUMENT = 1,    /* errors::InvalidArgument */
  PRB_PRECONDITION_NOT_MET = 2, /* errors::PreconditionNotMet */
  PRB_OUT_OF_RANGE = 3,        /* errors::OutOfRange */
  PRB_UNIMPLEMENTED = 4,       /* errors::Unimplemented */
  PRB_EXTERNAL = 5             /* errors::External: CUDA / NCCL failure */
} PRB_Status;

/* ref: paddle/phi/common/data_type.h:37-70 */
typedef enum PRB_DataType {
  PRB_BOOL = 1,
  PRB_INT32 = 7,
  PRB_INT64 = 9,
  PRB_FLOAT32 = 10,
  PRB_FLOAT64 = 11,
  PRB_FLOAT16 = 15,
  PRB_BFLOAT16 = 16
} PRB_DataType;

typedef struct PRB_Tensor {
  void* data;
  int64_t numel;
  int32_t dtype;  /* PRB_DataType */
  int32_t device; /* -1 = host, >= 0 = CUDA ordinal */
} PRB_Tensor;

const char* prb_last_error_message(void);
const char* prb_version(void);

/* ------------------------------------------------------------------------------------------
 * Optimizer kernels
 * ---------------------------------------------------------------------------------------- */

/* adamw_  ref: paddle/phi/kernels/adamw_kernel.h:22-49, gpu/adamw_kernel.cu:141-343,
 *              paddle/phi/api/yaml/legacy_ops.yaml:68-77.
 * beta1_pow/beta2_pow may live on the host (value captured at call time, updated on the host;
 * ref :234-289) or on the device (read in-kernel, updated by a device kernel; ref :290-341).
 * skip_update on the device is consumed on the device (no D2H sync) unless the beta-pows are on
 * the host, in which case the flag is read back like the reference does (ref :177-187). */
int32_t prb_adamw(const PRB_Tensor* param, const PRB_Tensor* grad, const PRB_Tensor* learning_rate,
                  const PRB_Tensor* moment1, const PRB_Tensor* moment2, const PRB_Tensor* beta1_pow,
                  const PRB_Tensor* beta2_pow, const PRB_Tensor* master_param /* optional */,
                  const PRB_Tensor* skip_update /* optional */, double beta1, double beta2, double epsilon,
                  float lr_ratio, float coeff, int32_t with_decay, int32_t lazy_mode,
                  int64_t min_row_size_to_use_multithread, int32_t multi_precision,
                  int32_t use_global_beta_pow, PRB_Tensor* param_out, PRB_Tensor* moment1_out,
                  PRB_Tensor* moment2_out, PRB_Tensor* beta1_pow_out, PRB_Tensor* beta2_pow_out,
                  PRB_Tensor* master_param_out /* optional */, void* stream);

/* adam_  ref: paddle/phi/kernels/adam_kernel.h:22-45, gpu/adam_kernel.cu:128-266,
 *             legacy_ops.yaml:45-55 */
int32_t prb_adam(const PRB_Tensor* param, const PRB_Tensor* grad, const PRB_Tensor* learning_rate,
                 const PRB_Tensor* moment1, const PRB_Tensor* moment2, const PRB_Tensor* beta1_pow,
                 const PRB_Tensor* beta2_pow, const PRB_Tensor* master_param /* optional */,
                 const PRB_Tensor* skip_update /* optional */, double beta1, double beta2, double epsilon,
                 int32_t lazy_mode, int64_t min_row_size_to_use_multithread, int32_t multi_precision,
                 int32_t use_global_beta_pow, PRB_Tensor* param_out, PRB_Tensor* moment1_out,
                 PRB_Tensor* moment2_out, PRB_Tensor* beta1_pow_out, PRB_Tensor* beta2_pow_out,
                 PRB_Tensor* master_param_out /* optional */, void* stream);

/* merged_adam_  ref: paddle/phi/kernels/adam_kernel.h:47-68, gpu/adam_kernel.cu:268-366,
 *                    legacy_ops.yaml:1090-1099.  One launch for the whole list (the reference
 *                    loops over per-tensor launches); bit-equal to per-tensor prb_adam. */
int32_t prb_merged_adam(const PRB_Tensor* param, const PRB_Tensor* grad, const PRB_Tensor* learning_rate,
                        const PRB_Tensor* moment1, const PRB_Tensor* moment2, const PRB_Tensor* beta1_pow,
                        const PRB_Tensor* beta2_pow, const PRB_Tensor* master_param /* optional list */,
                        int64_t num, double beta1, double beta2, double epsilon, int32_t multi_precision,
                        int32_t use_global_beta_pow, PRB_Tensor* param_out, PRB_Tensor* moment1_out,
                        PRB_Tensor* moment2_out, PRB_Tensor* beta1_pow_out, PRB_Tensor* beta2_pow_out,
                        PRB_Tensor* master_param_out /* optional list */, void* stream);

/* fused_adam_  ref: paddle/phi/kernels/fused_adam_kernel.h:22-47, gpu/fused_adam_kernel.cu:263-486,
 *                   legacy_ops.yaml:718-727.  One learning rate; tensor 0's beta-pows are applied
 *                   to every tensor (ref :292-293,368-369); zero-length tensors are legal;
 *                   chunk_size is accepted for signature parity and does not change results. */
int32_t prb_fused_adam(const PRB_Tensor* params, const PRB_Tensor* grads, const PRB_Tensor* learning_rate,
                       const PRB_Tensor* moments1, const PRB_Tensor* moments2, const PRB_Tensor* beta1_pows,
                       const PRB_Tensor* beta2_pows, const PRB_Tensor* master_params /* optional list */,
                       int64_t num, const PRB_Tensor* skip_update /* optional */, double beta1, double beta2,
                       double epsilon, int32_t chunk_size, float weight_decay, int32_t use_adamw,
                       int32_t multi_precision, int32_t use_global_beta_pow, PRB_Tensor* params_out,
                       PRB_Tensor* moments1_out, PRB_Tensor* moments2_out, PRB_Tensor* beta1_pows_out,
                       PRB_Tensor* beta2_pows_out, PRB_Tensor* master_params_out /* optional list */,
                       void* stream);

/* A SelectedRows gradient (ref: paddle/phi/core/selected_rows.h): `nrows` row indices (a HOST array, as the reference
 * keeps them; any order, duplicates allowed) into a [height, width] parameter and the dense [nrows, width] values. */
typedef struct PRB_SelectedRows {
  const int64_t* rows;
  int64_t nrows;
  int64_t height;
  PRB_Tensor value;
} PRB_SelectedRows;

/* adam_ with a sparse gradient  ref: paddle/phi/kernels/selected_rows/gpu/adam_kernel.cu:95-260
 *   (adam_dense_param_sparse_grad, legacy_ops.yaml:45-55).  Duplicate rows are merged first (scatter::MergeAdd);
 *   lazy_mode updates only the rows that are present, otherwise every element is updated with g = 0 where its row is
 *   absent.  In place (outputs must alias their inputs); the learning rate is a device scalar, the beta-pow
 *   accumulators live both on the device or both on the host (the reference's default). */
int32_t prb_adam_dense_param_sparse_grad(
    const PRB_Tensor* param, const PRB_SelectedRows* grad, const PRB_Tensor* learning_rate, const PRB_Tensor* moment1,
    const PRB_Tensor* moment2, const PRB_Tensor* beta1_pow, const PRB_Tensor* beta2_pow,
    const PRB_Tensor* master_param /* optional */, const PRB_Tensor* skip_update /* optional */, double beta1, double beta2,
    double epsilon, int32_t lazy_mode, int64_t min_row_size_to_use_multithread, int32_t multi_precision,
    int32_t use_global_beta_pow, PRB_Tensor* param_out, PRB_Tensor* moment1_out, PRB_Tensor* moment2_out,
    PRB_Tensor* beta1_pow_out, PRB_Tensor* beta2_pow_out, PRB_Tensor* master_param_out /* optional */, void* stream);

/* merge_selected_rows  ref: legacy_ops.yaml:1082-1088, paddle/phi/kernels/selected_rows/merge_selected_rows_kernel.h,
 *   scatter::MergeAdd (selected_rows_functor.cu:340-420): unique rows in ascending order (written to the HOST array
 *   out_rows, capacity x->nrows; their number to *out_nrows), the value rows of duplicates added in input order in
 *   fp32 and rounded once (written to out_value, a device tensor with room for x->nrows x width elements).
 *   ClipGradByGlobalNorm takes the norm of a sparse gradient from this (clip.py:655-664). */
int32_t prb_merge_selected_rows(const PRB_SelectedRows* x, int64_t* out_rows, int64_t* out_nrows, PRB_Tensor* out_value,
                                void* stream);

/* ------------------------------------------------------------------------------------------
 * AMP + global-norm clip
 * ---------------------------------------------------------------------------------------- */

/* check_finite_and_unscale_  ref: paddle/phi/kernels/amp_kernel.h:24-29,
 *                            gpu/amp_kernel.cu:27-82,267-348, legacy_ops.yaml:292-302.
 * found_infinite: one bool on the device. */
int32_t prb_check_finite_and_unscale(const PRB_Tensor* xs, int64_t num, const PRB_Tensor* scale,
                                     PRB_Tensor* outs, PRB_Tensor* found_infinite, void* stream);

/* update_loss_scaling_  ref: paddle/phi/kernels/amp_kernel.h:31-46, impl/amp_kernel_impl.h:38-180,
 *                       gpu/amp_kernel.cu:84-146, legacy_ops.yaml:1708-1717.  All state on device. */
int32_t prb_update_loss_scaling(const PRB_Tensor* xs, int64_t num, const PRB_Tensor* found_infinite,
                                const PRB_Tensor* prev_loss_scaling, const PRB_Tensor* in_good_steps,
                                const PRB_Tensor* in_bad_steps, int32_t incr_every_n_steps,
                                int32_t decr_every_n_nan_or_inf, float incr_ratio, float decr_ratio,
                                int32_t stop_update, PRB_Tensor* outs, PRB_Tensor* loss_scaling,
                                PRB_Tensor* out_good_steps, PRB_Tensor* out_bad_steps, void* stream);

/* squared_l2_norm  ref: paddle/phi/kernels/squared_l2_norm_kernel.h:21-24,
 *                  gpu/squared_l2_norm_kernel.cu:23-35, legacy_ops.yaml:1504-1511.
 * out: one element of x's dtype on the device.  Deterministic run to run
 * (ref: test_squared_l2_norm_op.py:98-111).  Allocates per-call scratch from the library's
 * workspace (grown with cudaMalloc when needed). */
int32_t prb_squared_l2_norm(const PRB_Tensor* x, PRB_Tensor* out, void* stream);

/* Same result as prb_squared_l2_norm for callers that cannot allocate device memory when they need a
 * norm: code running inside a process whose memory pool already owns the whole GPU, sections that must
 * not block on cudaMalloc/cudaFree, out-of-memory handlers.  This entry point has NO counterpart in the
 * reference (its reduce kernels always allocate a temporary); it exists for those callers only.
 *  - It never calls cudaMalloc/cudaFree: the partial sums live in a fixed array inside the library image
 *    (~128 KiB of device memory on every GPU that loads the library, held until the process exits).
 *  - x may have at most PRB_NOALLOC_MAX_NUMEL elements; more is rejected with PRB_INVALID_ARGUMENT.
 *  - x and out must live on the CURRENT device.  The first call on each device resolves the array's
 *    address (cudaGetSymbolAddress); make that call at start-up, before any stream capture.
 *  - Every call on a device shares the array.  A call overwrites what it later reads, so nothing needs
 *    clearing between calls, but two calls that overlap on the GPU (different streams) corrupt each
 *    other: keep them on one stream or order them with events. */
#define PRB_NOALLOC_MAX_NUMEL 268435456LL /* 32768 chunks x 8192 elements */
int32_t prb_squared_l2_norm_noalloc(const PRB_Tensor* x, PRB_Tensor* out, void* stream);

/* Multi-tensor form of the per-gradient loop in ClipGradByGlobalNorm._dygraph_clip
 * (ref: python/paddle/nn/clip.py:643-736): per-tensor squared norms (typed like the op),
 * their grouped sum, sqrt, clip = max_norm / max(norm, max_norm) (or the auto_skip_clip rule),
 * then g *= T(clip) in place.  global_norm_out / clip_out: optional fp32 device scalars.
 * use_mp_type mirrors _clip_by_global_norm_using_mp_type (ref: clip.py:179-202). */
int32_t prb_clip_grad_by_global_norm(PRB_Tensor* grads, int64_t num, float clip_norm, int32_t auto_skip_clip,
                                     int32_t use_mp_type, PRB_Tensor* global_norm_out /* optional */,
                                     PRB_Tensor* clip_out /* optional */, void* stream);
/* The multiply of _dygraph_clip alone (ref clip.py:715-727: paddle.multiply(g, clip_var.astype(g.dtype))), in place on a
 * list: x = T(MT(x) * MT(T(*clip))); a no-op when *clip == 1.  Used for the VALUES of sparse (SelectedRows) gradients,
 * whose norm is taken from their merged rows (prb_merge_selected_rows) but which are scaled as they are. */
int32_t prb_multiply_by_clip(PRB_Tensor* xs, int64_t num, const PRB_Tensor* clip, void* stream);

/* ------------------------------------------------------------------------------------------
 * Bucket operators
 * ---------------------------------------------------------------------------------------- */

/* scale  ref: paddle/phi/kernels/scale_kernel.h:23-29, gpu/scale_kernel.cu:24-67, ops.yaml:1251-1262 */
int32_t prb_scale(const PRB_Tensor* x, double scale, float bias, int32_t bias_after_scale, PRB_Tensor* out,
                  void* stream);

/* Dense 1-D concat / split of same-dtype tensors (what EagerGroup::ConcatTensors / SplitTensors
 * do to a bucket).  ref: paddle/fluid/distributed/collective/reducer.cc:358-459,
 * paddle/phi/kernels/funcs/concat_and_split_functor.cu:553-632,758-872.  Bit-exact byte moves. */
int32_t prb_concat(const PRB_Tensor* ins, int64_t num, PRB_Tensor* out, void* stream);
int32_t prb_split(const PRB_Tensor* in, PRB_Tensor* outs, int64_t num, void* stream);

/* out = x * y with y one float32 element on the device: the broadcast multiply of AmpScaler.scale(var)
 * (ref: python/paddle/amp/grad_scaler.py:152-200, `var * self._scale`), without reading the loss scale back. */
int32_t prb_multiply_by_scalar_tensor(const PRB_Tensor* x, const PRB_Tensor* y, PRB_Tensor* out, void* stream);

/* cast between float32 / bfloat16 / float16 (init-time helper: master weights are the fp32 cast of
 * 16-bit parameters, ref: python/paddle/optimizer/optimizer.py:640-666). */
int32_t prb_cast(const PRB_Tensor* x, PRB_Tensor* out, void* stream);

/* ------------------------------------------------------------------------------------------
 * Reducer (EagerReducer) and communicator (ProcessGroupNCCL, allreduce only)
 * ---------------------------------------------------------------------------------------- */

/* eager_assign_group_by_size  ref: reducer.cc:45-156, paddle/fluid/pybind/distributed_py.cc:64-77.
 * Output: groups as a flat index list + per-group sizes (caller provides num-sized arrays). */
int32_t prb_assign_group_by_size(const int64_t* numels, const int32_t* dtypes, const uint8_t* is_sparse,
                                 int64_t num, const uint64_t* group_size_limits, int64_t num_limits,
                                 const int64_t* tensor_indices, int64_t num_indices, int64_t* out_flat,
                                 int64_t* out_group_sizes, int64_t* out_num_groups);

typedef struct PRB_Comm PRB_Comm;
typedef struct PRB_Reducer PRB_Reducer;

#define PRB_UNIQUE_ID_BYTES 128
/* ref: process_group_nccl.cc:559-571 (rank 0 creates the ncclUniqueId, peers fetch it from the store) */
int32_t prb_comm_get_unique_id(uint8_t id[PRB_UNIQUE_ID_BYTES]);
/* ref: process_group_nccl.cc:688-737 (communicator + dedicated comm stream per device) */
int32_t prb_comm_init_rank(const uint8_t id[PRB_UNIQUE_ID_BYTES], int32_t nranks, int32_t rank, int32_t device,
                           PRB_Comm** out);
/* A communicator of one rank that needs no NCCL (nranks == 1). */
int32_t prb_comm_init_single(int32_t device, PRB_Comm** out);
int32_t prb_comm_destroy(PRB_Comm* comm);
int32_t prb_comm_nranks(const PRB_Comm* comm);
int32_t prb_comm_rank(const PRB_Comm* comm);
void* prb_comm_stream(const PRB_Comm* comm);
/* ref: ProcessGroupNCCL::AllReduce process_group_nccl.cc:841-865 (in place when out == in).
 * op: 0 = sum, 1 = max.  Enqueued on `stream` (pass prb_comm_stream(comm) for the side stream). */
int32_t prb_comm_all_reduce(PRB_Comm* comm, const PRB_Tensor* in, PRB_Tensor* out, int32_t op, void* stream);
/* ref: ProcessGroupNCCL::AllGather (the sparse-gradient path, reducer.cc:1165-1215): out holds nranks x in.numel
 * elements, rank r's contribution at offset r * in.numel. */
int32_t prb_comm_all_gather(PRB_Comm* comm, const PRB_Tensor* in, PRB_Tensor* out, void* stream);
/* ref: ProcessGroupNCCL::Broadcast process_group_nccl.cc:867-... (init-time parameter sync) */
int32_t prb_comm_broadcast(PRB_Comm* comm, PRB_Tensor* buf, int32_t root, void* stream);

/* EagerReducer(tensors, group_indices, is_sparse, process_group, group_size_limits, find_unused)
 * ref: reducer.h:88-142, reducer.cc:461-618, distributed_py.cc:1285-1315.
 * `grads[i]` is the gradient buffer of parameter i (dense only).  Buckets live in one persistent
 * arena owned by the reducer (no per-step allocation, unlike ref reducer.cc:359-360); segment
 * starts inside a bucket are padded to 256 B (internal layout, never visible to the caller).
 * prb_reducer_bucket_view exposes each parameter's slot so the optimizer can consume the reduced
 * gradients straight from the arena.
 * comm == NULL builds a host-only reducer: the hook/group state machine without any device work
 * (used to validate ordering and error behaviour on machines without a GPU). */
int32_t prb_reducer_create(const PRB_Tensor* grads, int64_t num, const int64_t* group_flat,
                           const int64_t* group_sizes, int64_t num_groups, PRB_Comm* comm,
                           int32_t find_unused_parameters, PRB_Reducer** out);
int32_t prb_reducer_destroy(PRB_Reducer* r);
/* ref: EagerReducer::PrepareForBackward reducer.cc:669-728 */
int32_t prb_reducer_prepare_for_backward(PRB_Reducer* r);
/* The result of the reference's TraverseBackwardGraph (reducer.cc:620-667), which belongs to the
 * host autograd engine: parameters that will receive no gradient in this backward.  Call after
 * prepare_for_backward (requires find_unused_parameters).  Their bucket slots are zero-filled and
 * they are marked ready together with the first real hook (ref :761-771,833-855).  Unlike the
 * reference (ProcessUnusedDenseVars :941-1011: int32 bitmap allreduce + host sync + copy), the
 * reduced slot is simply unpacked into the gradient buffer like any other -- no host sync. */
int32_t prb_reducer_set_unused_vars(PRB_Reducer* r, const int64_t* var_indices, int64_t num);
/* ref: EagerReducer::AddDistHook / MarkVarReady / MarkGroupReady reducer.cc:730-930.
 * `stream` is the calc stream the gradient was produced on.  When a group (and every group
 * before it) is complete: pack+scale on `stream`, event fence, allreduce on the comm stream
 * (ref: FusedAllReduceSchedule :1037-1070).  When the last group is launched the reducer
 * finalises (ref: FinalizeBackward :1013-1035): `stream` waits for the comm stream, and, unless
 * unpack is disabled, reduced buckets are scattered back into the gradient buffers. */
int32_t prb_reducer_mark_var_ready(PRB_Reducer* r, int64_t var_index, void* stream);
/* Batch form (one ABI crossing for many hooks; the reference's hooks are C++ callbacks). */
int32_t prb_reducer_mark_vars_ready(PRB_Reducer* r, const int64_t* var_indices, int64_t num, void* stream);
/* Gradients produced IN the bucket arena (what MarkVarReady achieves by aliasing the gradient into the group's
 * dense tensor, ref reducer.cc:774-901, without the concat that follows there): parameter i's gradient is the
 * slot prb_reducer_bucket_view(r, i) and the producer writes it there, so there is nothing to pack or to split
 * back.  With one rank a bucket launch does no device work at all; with several the 1/nranks scale (ref
 * reducer.cc:1049-1050) is applied to each rank's value inside the exchange kernel (same rounding as scale_),
 * or in place before ncclAllReduce.  A reducer created with grads[i].data == NULL starts in this mode. */
int32_t prb_reducer_set_grads_in_arena(PRB_Reducer* r, int32_t enabled);

/* Producer-side pass 1.  The consumer of the reduced gradients (prb_dp_step_tail) needs, before it can update,
 * (a) found_inf of check_finite_and_unscale (ref amp_kernel.cu:33-82) and (b) the squared-norm partial of every
 * 8192-element chunk (ref squared_l2_norm_kernel.cu:23-35, clip.py:643-736).  Both are functions of the REDUCED
 * gradient, and some kernel of the reducer is the last one to hold that value in registers: the pack (one rank)
 * or the slice owner of the peer-memory exchange (several ranks).  With an epilogue set, that kernel clears and
 * sets `found_inf` and stores the partials, and the tail is called with pass1_done = 1 -- the gradient is not
 * read a second time.  Results are bit-identical to the tail's own pass 1.
 *  partial_base[i]: where parameter i's first chunk sits in the consumer's list (sum of ceil(numel/8192) of the
 *  tensors before it in the list handed to prb_dp_step_tail).  *partials_out: the reducer-owned fp32 buffer to pass
 *  as cfg->pass1_partials.  *active: 1 if this reducer will run the epilogue (one rank with gradients outside the
 *  arena, or several ranks on the library's own exchange kernel); 0 = the consumer keeps its own pass 1.
 * e == NULL switches it off.  Not to be called between prepare_for_backward and the end of that backward. */
typedef struct PRB_GradEpilogue {
  const PRB_Tensor* loss_scale; /* fp32 device scalar; NULL = no unscale / finite check */
  PRB_Tensor* found_inf;        /* bool device scalar, required with loss_scale */
  int32_t want_norm;            /* 1: squared-norm partials */
  int32_t use_mp_type;          /* squares kept in fp32 for 16-bit gradients (ref clip.py:179-202) */
  const int32_t* partial_base;  /* [num parameters] host, required with want_norm */
} PRB_GradEpilogue;
int32_t prb_reducer_set_grad_epilogue(PRB_Reducer* r, const PRB_GradEpilogue* e, PRB_Tensor* partials_out,
                                      int32_t* active);
/* Elements per chunk of the norm partials (8192). */
int32_t prb_grad_chunk_elems(void);

/* 1 -> split reduced buckets back into the grads at finalize (reference behaviour, default);
 * 0 -> leave them in the arena for prb_dp_step_tail to consume directly. */
int32_t prb_reducer_set_unpack(PRB_Reducer* r, int32_t enabled);
/* Pipelined consumption.  With deferred waiting enabled, finalising no longer makes the calc stream wait
 * for every bucket (ref: FinalizeBackward reducer.cc:1013-1035 waits for all tasks); instead the consumer
 * calls prb_reducer_wait_group(r, k, stream) right before it touches the k-th LAUNCHED group's slots, so the
 * optimizer update of early buckets overlaps the allreduce of later ones. */
int32_t prb_reducer_set_deferred_wait(PRB_Reducer* r, int32_t enabled);
int32_t prb_reducer_wait_group(PRB_Reducer* r, int64_t group_index, void* stream);
int32_t prb_reducer_num_groups(const PRB_Reducer* r);
/* Index (in launch order, i.e. as passed to prb_reducer_create) of the group that holds parameter i. */
int32_t prb_reducer_group_of(const PRB_Reducer* r, int64_t var_index, int64_t* out_group);
/* Which allreduce runs the buckets: 0 = none (one rank / host-only), 1 = NCCL, 2 = the library's two-shot kernel
 * over NVLink peer memory (CUDA IPC; the default), 3 = the library's kernel over NVLink multicast (NVLS: the switch
 * adds and replicates).  PRB_ALLREDUCE = nvls | nvlink | nccl picks the first level to try; each level falls back
 * to the next, collectively, when it cannot be set up (single node, one GPU per process required for 2 and 3).
 * A wait inside those kernels that runs into its time limit (PRB_NVL_TIMEOUT_MS, default 600 000: a peer is gone)
 * does not fall through to the reduction: the rank leaves the data untouched, tells its peers (they fail at once),
 * and marks the exchange dead.  prb_reducer_comm_error reports that state without synchronising (a mapped host
 * word); prb_reducer_prepare_for_backward and prb_reducer_wait_group check it and fail with PRB_EXTERNAL. */
int32_t prb_reducer_allreduce_backend(const PRB_Reducer* r);
/* Only the allreduce of one group's bucket region, in place, on `stream` (measurement aid: lets a benchmark time
 * the exchange step of ref reducer.cc:1058 by itself with either backend). */
int32_t prb_reducer_allreduce_group(PRB_Reducer* r, int64_t group_index, void* stream);
int32_t prb_reducer_comm_error(PRB_Reducer* r, int32_t* out);
/* Grid of the library's exchange kernel (0 = the default, 64 CTAs: as fast as one CTA per SM on this NVLink-bound
 * kernel, and it leaves the backward pass room -- exposed communication 9 us instead of 77 us per step at 8 ranks).
 * A job whose backward needs still more of the GPU can go lower (NCCL runs 16-32 CTAs). */
int32_t prb_reducer_set_exchange_ctas(PRB_Reducer* r, int32_t max_ctas);
/* 1 when every group of the current backward has been launched and finalised. */
int32_t prb_reducer_is_finalized(const PRB_Reducer* r);
/* Device address / element offset of parameter i's slot inside the bucket arena. */
int32_t prb_reducer_bucket_view(const PRB_Reducer* r, int64_t var_index, PRB_Tensor* out);
/* Number of groups launched so far in this backward, and their order (for tests). */
int32_t prb_reducer_launched_groups(const PRB_Reducer* r, int64_t* out_order, int64_t capacity, int64_t* out_num);

/* Sharded step (ZeRO-style), built on the reducer's own NVLink exchange.  ref: the reference's optimizer-state
 * sharding, python/paddle/distributed/fleet/meta_parallel/sharding/group_sharded_optimizer_stage2.py:54-612 and
 * group_sharded_stage2.py (gradients reduced to the rank that owns them, owner-only update, parameters broadcast);
 * here the owner of an element is decided by position in the bucket arena, not by parameter:
 *   prb_reducer_set_sharded      the exchange becomes ONE reduce-scatter over all buckets, launched when the last
 *                                bucket is packed; rank r owns the r-th n-th of the 8192-element chunk list (groups in
 *                                launch order) and alone holds the reduced gradient there; inf flag and norm
 *                                partials of prb_reducer_set_grad_epilogue still reach every rank.  *active = 0 when
 *                                the reducer cannot do it (one rank, or no peer-memory exchange).
 *   prb_reducer_shard_ranges     the (variable, element offset, length) ranges a rank owns, in chunk order.
 *   prb_reducer_allgather_params after the owners updated their ranges with PRB_TailConfig.param_to_grad_slot: every
 *                                rank pulls the ranges it does not own from the owners' arenas into params[v]
 *                                (one tensor per variable).  skip (optional bool device flag, equal on all ranks):
 *                                nothing is copied when set. */
int32_t prb_reducer_set_sharded(PRB_Reducer* r, int32_t enabled, int32_t* active);
int32_t prb_reducer_shard_ranges(const PRB_Reducer* r, int32_t rank, int32_t nranks /* 0: the communicator's */,
                                 int64_t* out_var, int64_t* out_offset, int64_t* out_length, int64_t capacity,
                                 int64_t* out_num);
/* Ownership layout, chosen before prb_reducer_set_sharded and before anybody asks for shard_ranges:
 *   per_bucket = 0 (default)  rank r owns the r-th n-th of the chunk list of ALL buckets; ONE reduce-scatter, launched
 *                             after the last bucket is packed (fewest launches and barriers: best when no backward
 *                             pass is left to hide under);
 *   per_bucket = 1            rank r owns the r-th n-th of EACH bucket; one reduce-scatter per bucket, launched when
 *                             that bucket is packed, on the comm streams like the allreduce (ref: the per-parameter
 *                             reduce hooks of group_sharded_stage2.py:308-420, which also overlap the backward). */
int32_t prb_reducer_set_shard_layout(PRB_Reducer* r, int32_t per_bucket);
int32_t prb_reducer_allgather_params(PRB_Reducer* r, const PRB_Tensor* params, int64_t num, const PRB_Tensor* skip,
                                     void* stream);

/* ------------------------------------------------------------------------------------------
 * Fused step tail: unscale + inf/nan check + global-norm clip + Adam/AdamW in two passes
 * (norm pass, update pass) over the reducer's bucket arena or over caller gradients.
 * Composition of prb_check_finite_and_unscale, prb_clip_grad_by_global_norm and
 * prb_merged_adam / prb_adamw with identical per-element arithmetic
 * (ref: python/paddle/amp/grad_scaler.py:253-399,742-790; python/paddle/nn/clip.py:643-736;
 *  python/paddle/optimizer/optimizer.py:983-1010; python/paddle/optimizer/adamw.py:390-470).
 * ---------------------------------------------------------------------------------------- */
typedef struct PRB_TailConfig {
  double beta1, beta2, epsilon;
  float coeff;               /* weight decay coefficient (AdamW) */
  int32_t use_adamw;         /* 1: adamw (decay line present), 0: adam */
  int32_t multi_precision;   /* master params present */
  int32_t use_amp;           /* 1: unscale by 1/loss_scale and check finiteness */
  int32_t use_clip;          /* 1: ClipGradByGlobalNorm */
  float clip_norm;
  int32_t auto_skip_clip;
  int32_t use_mp_type;       /* clip norm in fp32 for 16-bit grads (ref: clip.py:179-202) */
  int32_t write_back_grads;  /* 1: store the unscaled (not clipped) gradient back: what p.grad holds in the reference */
  int32_t zero_grads;        /* 1: clear_grad(set_to_zero=True) folded in (ref: optimizer.py:1333-1381) */
  int32_t keep_beta_pow;     /* 1: do not advance the beta-pow accumulators in this call (a step split into
                                several calls, one per bucket, advances them in its last call only) */
  int32_t pass1_done;        /* 1: the producer already cleared/set found_inf and stored the norm partials
                                (prb_reducer_set_grad_epilogue); the gradient is read once, by the update pass */
  float* pass1_partials;     /* device: *partials_out of prb_reducer_set_grad_epilogue (required with use_clip) */
  int32_t stage;             /* 0: the whole tail.  A sharded step (prb_reducer_set_sharded) splits it in two calls:
                                1 = check + norm + clip factor only, over ALL gradients (moments, masters, learning
                                rate and beta-pows may be NULL; clip_out / global_norm_out are written);
                                2 = update only, over the sub-tensors this rank owns: the clip factor is READ from
                                clip_out (when use_clip), found_inf is read as the skip flag, use_amp still unscales */
  int32_t param_to_grad_slot; /* 1: the updated parameter (dtype of the gradient) is also stored over its gradient:
                                the bucket arena then carries it to the peers (prb_reducer_allgather_params) */
} PRB_TailConfig;

/* Per-tensor attributes follow adamw_: lr_ratio[i], with_decay[i] (NULL = 1.0 / all decayed).
 * found_inf (bool, device): with use_amp it is cleared, then set by the inf/nan check (output); with
 * use_amp == 0 a non-NULL found_inf is an input-only skip flag (the caller unscaled earlier).
 * global_norm/clip (fp32, device) are optional outputs.  beta-pows live on the device and are
 * advanced on the device unless found_inf. */
int32_t prb_dp_step_tail(const PRB_TailConfig* cfg, const PRB_Tensor* params, PRB_Tensor* grads,
                         const PRB_Tensor* learning_rate, const PRB_Tensor* moments1,
                         const PRB_Tensor* moments2, PRB_Tensor* beta1_pow, PRB_Tensor* beta2_pow,
                         const PRB_Tensor* master_params /* optional list */, int64_t num,
                         const float* lr_ratio /* optional [num] host */,
                         const uint8_t* with_decay /* optional [num] host */,
                         const PRB_Tensor* loss_scale /* fp32 device scalar, required if use_amp */,
                         PRB_Tensor* found_inf, PRB_Tensor* global_norm_out, PRB_Tensor* clip_out,
                         void* stream);

/* ------------------------------------------------------------------------------------------
 * Runtime helpers for the torch-free Python host (thin wrappers over the CUDA runtime).
 * ---------------------------------------------------------------------------------------- */
int32_t prb_device_count(int32_t* out);
int32_t prb_set_device(int32_t device);
int32_t prb_device_info(int32_t device, int32_t* sm_count, int64_t* l2_bytes, int64_t* total_mem, char name[128]);
int32_t prb_malloc(int64_t bytes, void** out);
int32_t prb_free(void* ptr);
int32_t prb_host_alloc(int64_t bytes, void** out); /* pinned */
int32_t prb_host_free(void* ptr);
int32_t prb_memcpy_h2d(void* dst, const void* src, int64_t bytes, void* stream);
int32_t prb_memcpy_d2h(void* dst, const void* src, int64_t bytes, void* stream);
int32_t prb_memcpy_d2d(void* dst, const void* src, int64_t bytes, void* stream);
int32_t prb_memset(void* dst, int32_t value, int64_t bytes, void* stream);
int32_t prb_stream_create(int32_t high_priority, void** out);
int32_t prb_stream_destroy(void* stream);
int32_t prb_stream_synchronize(void* stream);
int32_t prb_device_synchronize(void);
int32_t prb_event_create(void** out);
int32_t prb_event_destroy(void* event);
int32_t prb_event_record(void* event, void* stream);
int32_t prb_event_synchronize(void* event);
int32_t prb_event_elapsed_ms(void* start, void* stop, float* out_ms);
int32_t prb_stream_wait_event(void* stream, void* event);
/* *dst = *dst || *src for two device bools (the bitwise_or the reference's AmpScaler applies to its
 * per-dtype found_inf flags, ref: python/paddle/amp/grad_scaler.py:373-396). */
int32_t prb_flag_or(void* dst, const void* src, void* stream);
/* Busy kernel of roughly `microseconds` on `stream` (synthetic backward producer for overlap runs). */
int32_t prb_spin(int64_t microseconds, void* stream);
/* Writes `bytes` of scratch to evict L2 between timed iterations. */
int32_t prb_flush_l2(void* stream);
/* Number of kernels this library has launched since load (bench.py's gpu_launches). */
int64_t prb_launch_count(void);
/* Per-kernel timing for roofline reports: while enabled, every launch of a hot-path kernel is bracketed by two CUDA
 * events on the stream it is launched on.  prb_kernel_timing(1) clears and starts, (0) stops;
 * prb_kernel_timing_read synchronises the device and returns the summed event time and the launch count of one
 * kernel kind (0 .. 5, named by prb_kernel_kind_name; out-of-range kinds fail). */
int32_t prb_kernel_timing(int32_t enable);
int32_t prb_kernel_timing_read(int32_t kind, double* total_ms, int64_t* launches);
const char* prb_kernel_kind_name(int32_t kind);

#ifdef __cplusplus
}
#endif
#endif /* PADDLE_RAY_B200_H_ */
