// adam_ with a SelectedRows (sparse) gradient: the embedding-table case.
//
//   ref: AdamDenseParamSparseGradKernel  paddle/phi/kernels/selected_rows/gpu/adam_kernel.cu:95-260
//        SparseAdamCUDAKernelREG          :41-93
//        scatter::MergeAdd                paddle/phi/kernels/funcs/selected_rows_functor.cu (duplicates summed, rows sorted)
//        legacy_ops.yaml:45-55 (adam_ -> adam_dense_param_sparse_grad for a selected_rows grad)
//
// The gradient is a list of row indices (host array, as phi::SelectedRows keeps them) and a dense [nrows, width]
// value tensor on the device.  Rows may repeat and come in any order.  The host sorts and deduplicates the indices
// (a few thousand integers), the device adds the value rows of duplicates (fp32 accumulation in input order,
// rounded once to T -- the reference adds with atomics in an unspecified order; identical for up to two
// occurrences), then ONE kernel applies the shared Adam expression:
//   lazy_mode      only the rows that are present: one warp per row, 4 columns per lane and round, no search;
//   otherwise      every row of the parameter, with g = 0 where the row is absent (one binary search per row in
//                  the sorted unique rows; the reference searches per element).
#include <algorithm>
#include <map>
#include <mutex>
#include <cstring>
#include <numeric>
#include <vector>

#include "prb_adam.cuh"

namespace prb {
namespace {

constexpr int kSpThreads = 256;

// V consecutive elements of a row: one 16-byte access for fp32 (V = 4), 8 bytes for the 16-bit types.
template <typename T, int V>
struct alignas(sizeof(T) * V) RowVec {
  T v[V];
};
template <typename T, int V>
__device__ __forceinline__ void ld_row(const T* p, float (&out)[V]) {
  const RowVec<T, V> r = *reinterpret_cast<const RowVec<T, V>*>(p);
#pragma unroll
  for (int k = 0; k < V; ++k) out[k] = Conv<T>::to_f32(r.v[k]);
}
template <typename T, int V>
__device__ __forceinline__ void st_row(T* p, const float (&in)[V]) {
  RowVec<T, V> r;
#pragma unroll
  for (int k = 0; k < V; ++k) r.v[k] = Conv<T>::from_f32(in[k]);
  *reinterpret_cast<RowVec<T, V>*>(p) = r;
}

// One WARP per row (no per-element division, every access coalesced inside the row), V columns per lane and round.
// merged[u] = the value rows of unique row u added in INPUT order in fp32, rounded once to T.
template <typename T, int V>
__global__ void __launch_bounds__(kSpThreads) merge_rows_kernel(const T* __restrict__ values, const int64_t* __restrict__ off,
                                                                const int64_t* __restrict__ src, int64_t nunique,
                                                                int64_t width, T* __restrict__ merged) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * kSpThreads + threadIdx.x) >> 5, nwarps = ((int64_t)gridDim.x * kSpThreads) >> 5;
  for (int64_t u = warp; u < nunique; u += nwarps) {
    const int64_t s0 = off[u], s1 = off[u + 1];
    for (int64_t j = (int64_t)lane * V; j < width; j += 32 * V) {
      float acc[V];
#pragma unroll
      for (int k = 0; k < V; ++k) acc[k] = 0.0f;
      for (int64_t s = s0; s < s1; ++s) {
        float x[V];
        ld_row<T, V>(values + src[s] * width + j, x);
#pragma unroll
        for (int k = 0; k < V; ++k) acc[k] = __fadd_rn(acc[k], x[k]);
      }
      st_row<T, V>(merged + u * width + j, acc);
    }
  }
}

struct SparseArgs {
  void* param;
  float* mom1;
  float* mom2;
  float* master;
  const void* merged;        // [nunique, width]
  const int64_t* urows;      // sorted unique rows
  int64_t nunique, width, height;
  const float* lr;           // device scalar
  const float* b1pow;        // device scalars, or NULL when the accumulators live on the host (values below)
  const float* b2pow;
  float b1pow_host, b2pow_host;
  float beta1, beta2, eps;
  const bool* skip;
  int lazy;
};

__device__ __forceinline__ int64_t find_row(const int64_t* rows, int64_t n, int64_t key) {
  int64_t lo = 0, hi = n - 1;
  while (lo <= hi) {
    const int64_t mid = (lo + hi) >> 1;
    const int64_t v = rows[mid];
    if (v == key) return mid;
    if (v < key) lo = mid + 1;
    else hi = mid - 1;
  }
  return -1;
}

// lazy_mode: warp w takes unique row w (no search); otherwise warp w takes parameter row w and looks it up in the
// sorted unique rows once (ref SparseAdamCUDAKernelREG searches per ELEMENT, adam_kernel.cu:66-76).
template <typename T, bool kMaster, int V>
__global__ void __launch_bounds__(kSpThreads) sparse_adam_kernel(const SparseArgs a) {
  if (a.skip != nullptr && *a.skip) return;
  const float lr = __ldg(a.lr);
  const float b1p = a.b1pow != nullptr ? __ldg(a.b1pow) : a.b1pow_host;
  const float b2p = a.b2pow != nullptr ? __ldg(a.b2pow) : a.b2pow_host;
  const float c1 = __fsub_rn(1.0f, a.beta1), c2 = __fsub_rn(1.0f, a.beta2);
  const float bc2 = __fsqrt_rn(__fsub_rn(1.0f, b2p));
  const float step = -__fdiv_rn(lr, __fsub_rn(1.0f, b1p));
  T* __restrict__ param = reinterpret_cast<T*>(a.param);
  const T* __restrict__ merged = reinterpret_cast<const T*>(a.merged);
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * kSpThreads + threadIdx.x) >> 5, nwarps = ((int64_t)gridDim.x * kSpThreads) >> 5;
  const int64_t nrows = a.lazy ? a.nunique : a.height;
  for (int64_t r = warp; r < nrows; r += nwarps) {
    const int64_t prow = a.lazy ? a.urows[r] : r;
    const int64_t u = a.lazy ? r : find_row(a.urows, a.nunique, r);
    const int64_t base = prow * a.width;
    for (int64_t j = (int64_t)lane * V; j < a.width; j += 32 * V) {
      float p[V], g[V], m1[V], m2[V];
      if (kMaster) ld_row<float, V>(a.master + base + j, p);
      else ld_row<T, V>(param + base + j, p);
      if (u >= 0) {
        ld_row<T, V>(merged + u * a.width + j, g);
      } else {
#pragma unroll
        for (int k = 0; k < V; ++k) g[k] = 0.0f;
      }
      ld_row<float, V>(a.mom1 + base + j, m1);
      ld_row<float, V>(a.mom2 + base + j, m2);
#pragma unroll
      for (int k = 0; k < V; ++k) adam_moments_and_step(p[k], g[k], m1[k], m2[k], a.beta1, a.beta2, c1, c2, bc2, a.eps, step);
      st_row<float, V>(a.mom1 + base + j, m1);
      st_row<float, V>(a.mom2 + base + j, m2);
      st_row<T, V>(param + base + j, p);
      if (kMaster) st_row<float, V>(a.master + base + j, p);
    }
  }
}

template <typename T, int V>
int32_t run_sparse_v(const SparseArgs& a, const T* values, const int64_t* off, const int64_t* src, T* merged, bool need_merge,
                     bool has_master, cudaStream_t stream) {
  const int64_t cap = (int64_t)sm_count() * 8, wpc = kSpThreads / 32;
  if (need_merge && a.nunique > 0) {
    const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((a.nunique + wpc - 1) / wpc, cap));
    merge_rows_kernel<T, V><<<grid, kSpThreads, 0, stream>>>(values, off, src, a.nunique, a.width, merged);
    PRB_LAUNCH_CHECK();
  }
  const int64_t nrows = a.lazy ? a.nunique : a.height;
  if (nrows == 0 || a.width == 0) return PRB_SUCCESS;
  const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((nrows + wpc - 1) / wpc, cap));
  if (has_master) sparse_adam_kernel<T, true, V><<<grid, kSpThreads, 0, stream>>>(a);
  else sparse_adam_kernel<T, false, V><<<grid, kSpThreads, 0, stream>>>(a);
  PRB_LAUNCH_CHECK();
  return PRB_SUCCESS;
}

template <typename T>
int32_t run_sparse(const SparseArgs& a, const T* values, const int64_t* off, const int64_t* src, T* merged, bool need_merge,
                   bool has_master, cudaStream_t stream) {
  // 4 columns per lane when every row of every operand starts on a 16-byte (fp32) / 8-byte (16-bit) boundary
  auto al = [](const void* p, size_t b) { return ((uintptr_t)p % b) == 0; };
  const bool vec = a.width % 4 == 0 && al(a.param, sizeof(T) * 4) && al(values, sizeof(T) * 4) && al(merged, sizeof(T) * 4) &&
                   al(a.mom1, 16) && al(a.mom2, 16) && (!has_master || al(a.master, 16));
  if (vec) return run_sparse_v<T, 4>(a, values, off, src, merged, need_merge, has_master, stream);
  return run_sparse_v<T, 1>(a, values, off, src, merged, need_merge, has_master, stream);
}

// Host side shared by the sparse adam and merge_selected_rows: stable sort + dedup of the row indices.
// Stable (duplicates keep their input order): an LSD radix sort, 11 bits per pass over the bits the table's height
// needs -- 8192 indices into a 30522-row table: 64 us, where std::stable_sort through the index indirection took
// 556 us (and the reference's std::set insertion, more).
struct RowPlan {
  std::vector<int64_t> order;  // input positions, sorted by row (stable)
  std::vector<int64_t> urows;  // unique rows, ascending
  std::vector<int64_t> off;    // segment k of `order` = positions [off[k], off[k + 1]) of unique row k
  bool need_merge = false;     // false: the rows were already strictly ascending (ref is_strict_sorted)
};

int32_t plan_rows(const PRB_SelectedRows* x, const char* op, RowPlan* out) {
  const int64_t n = x->nrows;
  for (int64_t k = 0; k < n; ++k)
    PRB_ENFORCE(x->rows[k] >= 0 && x->rows[k] < x->height, PRB_OUT_OF_RANGE, "%s: row %lld is outside [0, %lld)", op,
                (long long)x->rows[k], (long long)x->height);
  std::vector<int64_t>& order = out->order;
  std::vector<int64_t> other((size_t)n);
  order.resize((size_t)n);
  std::iota(order.begin(), order.end(), (int64_t)0);
  int bits = 0;
  while (((int64_t)1 << bits) < x->height) ++bits;
  for (int sh = 0; sh < bits && n > 1; sh += 11) {
    uint32_t cnt[2049];
    memset(cnt, 0, sizeof(cnt));
    for (int64_t k = 0; k < n; ++k) cnt[((x->rows[order[(size_t)k]] >> sh) & 2047) + 1]++;
    for (int d = 0; d < 2048; ++d) cnt[d + 1] += cnt[d];
    for (int64_t k = 0; k < n; ++k) other[cnt[(x->rows[order[(size_t)k]] >> sh) & 2047]++] = order[(size_t)k];
    order.swap(other);
  }
  for (int64_t k = 0; k < n; ++k) {
    const int64_t r = x->rows[order[(size_t)k]];
    if (out->urows.empty() || out->urows.back() != r) {
      out->urows.push_back(r);
      out->off.push_back(k);
    }
  }
  out->off.push_back(n);
  out->need_merge = (int64_t)out->urows.size() != n;
  for (int64_t k = 0; k < n && !out->need_merge; ++k) out->need_merge = order[(size_t)k] != k;
  return PRB_SUCCESS;
}

// Page-locked staging for the index tables: a copy from pageable memory would first wait for the whole stream
// (CUDA synchronises the stream before it stages a pageable source), which stalls the host behind everything a
// training step has queued.  Four slots per (device, stream), each guarded by an event recorded after its copy, so the
// host may run four sparse updates ahead of the device.
struct StageSlot {
  void* host = nullptr;
  size_t bytes = 0;
  cudaEvent_t copied = nullptr;
};
struct StageRing {
  StageSlot slot[4];
  unsigned next = 0;
};
std::mutex g_stage_mu;
std::map<std::pair<int, void*>, StageRing> g_stage;

int32_t stage_acquire(size_t bytes, cudaStream_t stream, StageSlot** out) {
  int dev = 0;
  PRB_CUDA(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> lk(g_stage_mu);
  StageRing& ring = g_stage[std::make_pair(dev, (void*)stream)];
  StageSlot& st = ring.slot[ring.next++ & 3u];
  if (st.copied == nullptr) PRB_CUDA(cudaEventCreateWithFlags(&st.copied, cudaEventDisableTiming));
  else PRB_CUDA(cudaEventSynchronize(st.copied));  // the copy that last used this slot has been performed
  if (st.bytes < bytes) {
    if (st.host != nullptr) PRB_CUDA(cudaFreeHost(st.host));
    st.host = nullptr;
    st.bytes = 0;
    const size_t want = bytes < 65536 ? 65536 : bytes + bytes / 2;
    PRB_CUDA(cudaHostAlloc(&st.host, want, cudaHostAllocDefault));
    st.bytes = want;
  }
  *out = &st;
  return PRB_SUCCESS;
}

// [urows | off | src] -> the stream's workspace (one asynchronous H2D copy out of page-locked staging), followed by
// room for `extra` bytes at a 256-byte boundary.
int32_t upload_plan(const RowPlan& rp, size_t extra, cudaStream_t stream, const int64_t** d_urows, const int64_t** d_off,
                    const int64_t** d_src, void** d_extra) {
  const size_t nu = rp.urows.size(), n = rp.order.size();
  const size_t ints = nu + nu + 1 + n;
  const size_t extra_off = (ints * sizeof(int64_t) + 255) / 256 * 256;
  void* ws = nullptr;
  PRB_TRY(workspace(extra_off + extra + 256, &ws, /*slot=*/3, stream));
  StageSlot* st = nullptr;
  PRB_TRY(stage_acquire(ints * sizeof(int64_t), stream, &st));
  int64_t* host = reinterpret_cast<int64_t*>(st->host);
  std::copy(rp.urows.begin(), rp.urows.end(), host);
  std::copy(rp.off.begin(), rp.off.end(), host + nu);
  std::copy(rp.order.begin(), rp.order.end(), host + nu + nu + 1);
  PRB_CUDA(cudaMemcpyAsync(ws, host, ints * sizeof(int64_t), cudaMemcpyHostToDevice, stream));
  PRB_CUDA(cudaEventRecord(st->copied, stream));
  *d_urows = reinterpret_cast<const int64_t*>(ws);
  *d_off = *d_urows + nu;
  *d_src = *d_off + nu + 1;
  *d_extra = reinterpret_cast<char*>(ws) + extra_off;
  return PRB_SUCCESS;
}

template <typename T>
int32_t run_merge(const T* values, const int64_t* off, const int64_t* src, int64_t nunique, int64_t width, T* merged,
                  cudaStream_t stream) {
  if (nunique == 0 || width == 0) return PRB_SUCCESS;
  const int64_t cap = (int64_t)sm_count() * 8, wpc = kSpThreads / 32;
  const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((nunique + wpc - 1) / wpc, cap));
  const bool vec = width % 4 == 0 && ((uintptr_t)values % (sizeof(T) * 4)) == 0 && ((uintptr_t)merged % (sizeof(T) * 4)) == 0;
  if (vec) merge_rows_kernel<T, 4><<<grid, kSpThreads, 0, stream>>>(values, off, src, nunique, width, merged);
  else merge_rows_kernel<T, 1><<<grid, kSpThreads, 0, stream>>>(values, off, src, nunique, width, merged);
  PRB_LAUNCH_CHECK();
  return PRB_SUCCESS;
}

}  // namespace
}  // namespace prb

using namespace prb;

extern "C" int32_t prb_adam_dense_param_sparse_grad(
    const PRB_Tensor* param, const PRB_SelectedRows* grad, const PRB_Tensor* learning_rate, const PRB_Tensor* moment1,
    const PRB_Tensor* moment2, const PRB_Tensor* beta1_pow, const PRB_Tensor* beta2_pow, const PRB_Tensor* master_param,
    const PRB_Tensor* skip_update, double beta1, double beta2, double epsilon, int32_t lazy_mode,
    int64_t min_row_size_to_use_multithread, int32_t multi_precision, int32_t use_global_beta_pow, PRB_Tensor* param_out,
    PRB_Tensor* moment1_out, PRB_Tensor* moment2_out, PRB_Tensor* beta1_pow_out, PRB_Tensor* beta2_pow_out,
    PRB_Tensor* master_param_out, void* stream_) {
  (void)min_row_size_to_use_multithread;
  cudaStream_t stream = (cudaStream_t)stream_;
  PRB_TRY(require_device());
  PRB_ENFORCE(param && grad && learning_rate && moment1 && moment2 && beta1_pow && beta2_pow, PRB_INVALID_ARGUMENT,
              "adam (sparse grad): a required input is NULL");
  const int32_t dt = param->dtype;
  PRB_ENFORCE(dt == PRB_FLOAT32 || dt == PRB_BFLOAT16 || dt == PRB_FLOAT16, PRB_UNIMPLEMENTED,
              "adam (sparse grad): dtype %s is not built", dtype_name(dt));
  PRB_ENFORCE(grad->value.dtype == dt && grad->value.device >= 0 && param->device >= 0, PRB_INVALID_ARGUMENT,
              "adam (sparse grad): the gradient values must be a device tensor of the parameter's dtype");
  PRB_ENFORCE(grad->height > 0 && param->numel % grad->height == 0, PRB_INVALID_ARGUMENT,
              "adam (sparse grad): the parameter's %lld elements are not a whole number of rows of height %lld",
              (long long)param->numel, (long long)grad->height);
  const int64_t width = param->numel / grad->height;
  PRB_ENFORCE(grad->nrows >= 0 && grad->value.numel == grad->nrows * width && (grad->rows != nullptr || grad->nrows == 0),
              PRB_INVALID_ARGUMENT, "adam (sparse grad): the gradient must hold nrows x %lld values", (long long)width);
  PRB_ENFORCE(moment1->numel == param->numel && moment2->numel == param->numel && moment1->dtype == PRB_FLOAT32 &&
                  moment2->dtype == PRB_FLOAT32 && moment1->device >= 0 && moment2->device >= 0,
              PRB_INVALID_ARGUMENT, "adam (sparse grad): moments must be float32 device tensors of the parameter's size");
  const bool has_master = multi_precision != 0;
  PRB_ENFORCE(!has_master || (master_param && master_param->numel == param->numel && master_param->dtype == PRB_FLOAT32),
              PRB_INVALID_ARGUMENT, "adam (sparse grad): multi_precision requires a float32 master_param");
  PRB_ENFORCE(learning_rate->numel == 1 && learning_rate->dtype == PRB_FLOAT32 && learning_rate->device >= 0,
              PRB_INVALID_ARGUMENT, "adam (sparse grad): learning_rate must be a single float32 device scalar");
  // the accumulators may live on the host, as the reference's do by default (adam_kernel.cu:207-229 / :230-252)
  const bool pows_on_host = beta1_pow->device < 0;
  PRB_ENFORCE(beta1_pow->numel == 1 && beta2_pow->numel == 1 && beta1_pow->dtype == PRB_FLOAT32 &&
                  beta2_pow->dtype == PRB_FLOAT32 && (beta2_pow->device < 0) == pows_on_host,
              PRB_INVALID_ARGUMENT,
              "adam (sparse grad): the beta-pow accumulators must be single float32 scalars, both on the host or both on "
              "the device");
  for (const PRB_Tensor* o : {param_out, moment1_out, moment2_out, master_param_out})
    PRB_ENFORCE(o == nullptr || o->data == nullptr || o->data == param->data || o->data == moment1->data ||
                    o->data == moment2->data || (master_param && o->data == master_param->data),
                PRB_UNIMPLEMENTED, "adam (sparse grad): outputs must alias their inputs (the op is in place, legacy_ops.yaml:55)");
  const bool* skip = nullptr;
  if (skip_update != nullptr) {
    PRB_ENFORCE(skip_update->numel == 1, PRB_INVALID_ARGUMENT, "Input(SkipUpdate) size must be 1, but get %lld",
                (long long)skip_update->numel);
    if (skip_update->device < 0) {
      if (*reinterpret_cast<const bool*>(skip_update->data)) return PRB_SUCCESS;
    } else if (pows_on_host) {
      // host accumulators advance on the host, so the flag has to be known here: read it back, as the reference
      // always does (TensorToVector + Wait, adam_kernel.cu:133-143)
      bool flag = false;
      PRB_CUDA(cudaMemcpyAsync(&flag, skip_update->data, sizeof(bool), cudaMemcpyDeviceToHost, stream));
      PRB_CUDA(cudaStreamSynchronize(stream));
      if (flag) return PRB_SUCCESS;
    } else {
      skip = reinterpret_cast<const bool*>(skip_update->data);
    }
  }
  // ---- host: sort + deduplicate the row indices; device scratch: [urows | off | src | merged values]
  RowPlan rp;
  PRB_TRY(plan_rows(grad, "adam (sparse grad)", &rp));
  const int64_t nu = (int64_t)rp.urows.size();
  const bool need_merge = rp.need_merge;
  const int64_t *d_urows, *d_off, *d_src;
  void* d_merged;
  PRB_TRY(upload_plan(rp, (size_t)nu * (size_t)width * (size_t)dtype_size(dt), stream, &d_urows, &d_off, &d_src, &d_merged));
  SparseArgs a{};
  a.param = param->data;
  a.mom1 = reinterpret_cast<float*>(moment1->data);
  a.mom2 = reinterpret_cast<float*>(moment2->data);
  a.master = has_master ? reinterpret_cast<float*>(master_param->data) : nullptr;
  a.merged = need_merge ? d_merged : grad->value.data;
  a.urows = d_urows;
  a.nunique = nu;
  a.width = width;
  a.height = grad->height;
  a.lr = reinterpret_cast<const float*>(learning_rate->data);
  if (pows_on_host) {
    a.b1pow_host = *reinterpret_cast<const float*>(beta1_pow->data);
    a.b2pow_host = *reinterpret_cast<const float*>(beta2_pow->data);
  } else {
    a.b1pow = reinterpret_cast<const float*>(beta1_pow->data);
    a.b2pow = reinterpret_cast<const float*>(beta2_pow->data);
  }
  a.beta1 = (float)beta1;
  a.beta2 = (float)beta2;
  a.eps = (float)epsilon;
  a.skip = skip;
  a.lazy = lazy_mode != 0;
  switch (dt) {
    case PRB_FLOAT32:
      PRB_TRY(run_sparse<float>(a, (const float*)grad->value.data, d_off, d_src, (float*)d_merged, need_merge, has_master, stream));
      break;
    case PRB_BFLOAT16:
      PRB_TRY(run_sparse<__nv_bfloat16>(a, (const __nv_bfloat16*)grad->value.data, d_off, d_src, (__nv_bfloat16*)d_merged,
                                        need_merge, has_master, stream));
      break;
    default:
      PRB_TRY(run_sparse<__half>(a, (const __half*)grad->value.data, d_off, d_src, (__half*)d_merged, need_merge, has_master, stream));
  }
  if (!use_global_beta_pow && pows_on_host) {
    const PRB_Tensor* o1 = beta1_pow_out && beta1_pow_out->data ? beta1_pow_out : beta1_pow;
    const PRB_Tensor* o2 = beta2_pow_out && beta2_pow_out->data ? beta2_pow_out : beta2_pow;
    *reinterpret_cast<float*>(o1->data) = (float)beta1 * a.b1pow_host;
    *reinterpret_cast<float*>(o2->data) = (float)beta2 * a.b2pow_host;
  } else if (!use_global_beta_pow) {
    PRB_TRY(beta_pow_update_device(beta1_pow, beta2_pow, beta1_pow_out, beta2_pow_out, 1, (float)beta1, (float)beta2, skip, stream));
  }
  return PRB_SUCCESS;
}

// ref: merge_selected_rows (legacy_ops.yaml:1082-1088, MergeSelectedRowsKernel -> scatter::MergeAdd): unique rows in
// ascending order, the value rows of duplicates added.
extern "C" int32_t prb_merge_selected_rows(const PRB_SelectedRows* x, int64_t* out_rows, int64_t* out_nrows,
                                           PRB_Tensor* out_value, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  PRB_TRY(require_device());
  PRB_ENFORCE(x && out_nrows && out_value && (out_rows || x->nrows == 0), PRB_INVALID_ARGUMENT,
              "merge_selected_rows: a required argument is NULL");
  const int32_t dt = x->value.dtype;
  PRB_ENFORCE(dt == PRB_FLOAT32 || dt == PRB_BFLOAT16 || dt == PRB_FLOAT16, PRB_UNIMPLEMENTED,
              "merge_selected_rows: dtype %s is not built", dtype_name(dt));
  PRB_ENFORCE(x->nrows >= 0 && x->height > 0 && (x->rows != nullptr || x->nrows == 0) && x->value.device >= 0,
              PRB_INVALID_ARGUMENT, "merge_selected_rows: bad input");
  if (x->nrows == 0) {
    *out_nrows = 0;
    return PRB_SUCCESS;
  }
  PRB_ENFORCE(x->value.numel % x->nrows == 0, PRB_INVALID_ARGUMENT, "merge_selected_rows: value must hold nrows x width elements");
  const int64_t width = x->value.numel / x->nrows;
  RowPlan rp;
  PRB_TRY(plan_rows(x, "merge_selected_rows", &rp));
  const int64_t nu = (int64_t)rp.urows.size();
  PRB_ENFORCE(out_value->dtype == dt && out_value->device >= 0 && out_value->numel >= nu * width, PRB_INVALID_ARGUMENT,
              "merge_selected_rows: out_value must be a device tensor of the input's dtype with room for %lld x %lld elements",
              (long long)nu, (long long)width);
  std::copy(rp.urows.begin(), rp.urows.end(), out_rows);
  *out_nrows = nu;
  if (!rp.need_merge) {  // already strictly ascending: the values are the result
    if (out_value->data != x->value.data)
      PRB_CUDA(cudaMemcpyAsync(out_value->data, x->value.data, (size_t)nu * width * dtype_size(dt), cudaMemcpyDeviceToDevice, stream));
    return PRB_SUCCESS;
  }
  PRB_ENFORCE(out_value->data != x->value.data, PRB_INVALID_ARGUMENT, "merge_selected_rows: cannot merge in place");
  const int64_t *d_urows, *d_off, *d_src;
  void* unused;
  PRB_TRY(upload_plan(rp, 0, stream, &d_urows, &d_off, &d_src, &unused));
  switch (dt) {
    case PRB_FLOAT32: return run_merge<float>((const float*)x->value.data, d_off, d_src, nu, width, (float*)out_value->data, stream);
    case PRB_BFLOAT16:
      return run_merge<__nv_bfloat16>((const __nv_bfloat16*)x->value.data, d_off, d_src, nu, width,
                                      (__nv_bfloat16*)out_value->data, stream);
    default: return run_merge<__half>((const __half*)x->value.data, d_off, d_src, nu, width, (__half*)out_value->data, stream);
  }
}
