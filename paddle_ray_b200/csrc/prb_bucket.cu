// Bucket operators: scale, dense concat (pack) and split (unpack) as ONE multi-tensor map kernel.
//
//   ref: EagerGroup::ConcatTensors / SplitTensors  paddle/fluid/distributed/collective/reducer.cc:358-459
//        ConcatFunctor / SplitFunctor (GPU)         paddle/phi/kernels/funcs/concat_and_split_functor.cu:553-872
//        ScaleKernel / ScaleFunctor                 paddle/phi/kernels/gpu/scale_kernel.cu:24-67
//
// The reference moves every gradient byte three extra times per step (concat, scale, split; the
// split with element-granular accesses and a per-thread linear segment search).  Here pack and the
// 1/nranks scale are one pass (read s + write s bytes per element), and unpack is one pass; each
// CTA owns one 8192-element chunk of one tensor, so no thread ever searches for its segment.
#include <vector>

#include "prb_bucket.cuh"
#include "prb_fp64.h"

namespace prb {
namespace {

// Launch geometry: 256 threads x 4 lane vectors = 8192 elements per CTA (the six shapes measured in round 1 were
// within 1 % of each other, profiles/r1_experiments.md; one is kept).
constexpr int kMapThreads = 256, kMapUnroll = 4, kMapMinBlocks = 2;

struct CopyOp {
  template <typename T>
  __device__ __forceinline__ T operator()(T v) const {
    return v;
  }
};

// out = T(scale * (MT(x) + bias)) or T(scale * MT(x) + bias)   (ref scale_kernel.cu:35-40)
struct ScaleOp {
  float scale, bias;
  bool bias_after_scale;
  template <typename T>
  __device__ __forceinline__ T operator()(T v) const {
    const float x = Conv<T>::to_f32(v);
    const float r = bias_after_scale ? __fadd_rn(__fmul_rn(scale, x), bias) : __fmul_rn(scale, __fadd_rn(x, bias));
    return Conv<T>::from_f32(r);
  }
};

template <typename T, typename Op>
__global__ void __launch_bounds__(kMapThreads, kMapMinBlocks) map_chunks_kernel(const MapDesc* __restrict__ descs,
                                                                                const Chunk* __restrict__ chunks,
                                                                                const Op op) {
  pdl_trigger();  // a dependent launch behind this one may start filling SMs as the grid drains
  const int4 craw = __ldg(reinterpret_cast<const int4*>(chunks) + blockIdx.x);
  const int n = craw.y;
  const int64_t offset = (int64_t)(((uint64_t)(uint32_t)craw.w << 32) | (uint32_t)craw.z);
  const MapDesc* __restrict__ d = descs + craw.x;
  const T* __restrict__ src = reinterpret_cast<const T*>(d->src) + offset;
  T* __restrict__ dst = reinterpret_cast<T*>(d->dst) + offset;
  if (d->aligned) {
    const int nvec = n / kVec;
    Pack<T> v[kMapUnroll];
#pragma unroll
    for (int u = 0; u < kMapUnroll; ++u) {
      const int i = u * kMapThreads + threadIdx.x;
      if (i < nvec) load_vec<T>(src + i * kVec, v[u]);
    }
#pragma unroll
    for (int u = 0; u < kMapUnroll; ++u) {
      const int i = u * kMapThreads + threadIdx.x;
      if (i < nvec) {
#pragma unroll
        for (int j = 0; j < kVec; ++j) v[u].v[j] = op(v[u].v[j]);
        store_vec<T>(dst + i * kVec, v[u]);
      }
    }
    const int e = nvec * kVec + threadIdx.x;
    if (e < n) dst[e] = op(src[e]);
  } else {
    // Segment whose source or destination is only element-aligned (dense reference packing):
    // coalesced element-wise accesses.
    for (int e = threadIdx.x; e < n; e += kMapThreads) dst[e] = op(src[e]);
  }
}

template <typename T, typename Op>
int32_t launch_map_typed(const DeviceTable& tab, const Op& op, cudaStream_t stream) {
  if (tab.nchunks == 0) return PRB_SUCCESS;
  const MapDesc* d = reinterpret_cast<const MapDesc*>(tab.descs);
  KernelTimer kt(KK_MAP, stream);
  map_chunks_kernel<T, Op><<<tab.nchunks, kMapThreads, 0, stream>>>(d, tab.chunks, op);
  PRB_LAUNCH_CHECK();
  return PRB_SUCCESS;
}

}  // namespace

int map_chunk_elems() { return kMapThreads * kVec * kMapUnroll; }

bool map_aligned(const void* src, const void* dst, int elem_size) {
  const uintptr_t a = (uintptr_t)elem_size * kVec;
  return ((uintptr_t)src % a) == 0 && ((uintptr_t)dst % a) == 0;
}

int32_t launch_copy(int elem_size, const DeviceTable& tab, cudaStream_t stream) {
  if (elem_size == 4) return launch_map_typed<uint32_t>(tab, CopyOp{}, stream);
  if (elem_size == 2) return launch_map_typed<uint16_t>(tab, CopyOp{}, stream);
  PRB_FAIL(PRB_UNIMPLEMENTED, "copy: unsupported element size %d", elem_size);
}

int32_t launch_scale(int32_t dtype, const DeviceTable& tab, double scale, float bias, bool bias_after_scale,
                     cudaStream_t stream) {
  ScaleOp op{(float)scale, bias, bias_after_scale};  // Scalar::to<MT>() (ref scale_kernel.cu:64-66)
  switch (dtype) {
    case PRB_FLOAT32: return launch_map_typed<float>(tab, op, stream);
    case PRB_BFLOAT16: return launch_map_typed<__nv_bfloat16>(tab, op, stream);
    case PRB_FLOAT16: return launch_map_typed<__half>(tab, op, stream);
    default: PRB_FAIL(PRB_UNIMPLEMENTED, "scale: dtype %s is not built (float32/bfloat16/float16 only)", dtype_name(dtype));
  }
}

}  // namespace prb

using namespace prb;

extern "C" {

int32_t prb_scale(const PRB_Tensor* x, double scale, float bias, int32_t bias_after_scale, PRB_Tensor* out,
                  void* stream) {
  PRB_TRY(require_device());
  PRB_ENFORCE(x && out, PRB_INVALID_ARGUMENT, "scale: NULL tensor");
  PRB_ENFORCE(x->numel == out->numel && x->dtype == out->dtype, PRB_INVALID_ARGUMENT, "scale: x/out mismatch");
  PRB_ENFORCE(x->device >= 0 && out->device >= 0, PRB_INVALID_ARGUMENT, "scale: device tensors required");
  if (x->numel <= 0) return PRB_SUCCESS;  // ref scale_kernel.cu:59-61
  if (x->dtype == PRB_FLOAT64)              // MT = double (ref scale_kernel.cu:70-75)
    return launch_scale_f64(reinterpret_cast<const double*>(x->data), reinterpret_cast<double*>(out->data), x->numel, scale,
                            (double)bias, bias_after_scale != 0, (cudaStream_t)stream);
  MapDesc d{};
  d.src = x->data;
  d.dst = out->data;
  d.aligned = map_aligned(d.src, d.dst, dtype_size(x->dtype)) ? 1 : 0;
  DeviceTable tab;
  PRB_TRY(get_table(&d, sizeof(d), 1, &x->numel, map_chunk_elems(), (cudaStream_t)stream, &tab));
  return launch_scale(x->dtype, tab, scale, bias, bias_after_scale != 0, (cudaStream_t)stream);
}

static int32_t concat_split(const PRB_Tensor* flat, const PRB_Tensor* parts, int64_t num, bool is_concat,
                            cudaStream_t stream) {
  PRB_TRY(require_device());
  const char* op = is_concat ? "concat" : "split";
  PRB_ENFORCE(flat && (parts || num == 0) && num >= 0, PRB_INVALID_ARGUMENT, "%s: NULL argument", op);
  PRB_ENFORCE(flat->device >= 0, PRB_INVALID_ARGUMENT, "%s: device tensors required", op);
  int es = dtype_size(flat->dtype);
  PRB_ENFORCE(es == 2 || es == 4 || es == 8, PRB_UNIMPLEMENTED, "%s: unsupported dtype %s", op, dtype_name(flat->dtype));
  const int64_t words = es == 8 ? 2 : 1;  // 8-byte elements move as pairs of 32-bit words
  if (es == 8) es = 4;
  std::vector<MapDesc> descs;
  std::vector<int64_t> numels;
  int64_t off = 0;
  for (int64_t i = 0; i < num; ++i) {
    PRB_ENFORCE(parts[i].dtype == flat->dtype, PRB_PRECONDITION_NOT_MET, "%s: tensor %lld has unexpected dtype.", op,
                (long long)i);  // ref reducer.cc:607-612
    PRB_ENFORCE(parts[i].numel >= 0 && (parts[i].device >= 0 || parts[i].numel == 0), PRB_INVALID_ARGUMENT,
                "%s: tensor %lld must be a device tensor", op, (long long)i);
    if (parts[i].numel > 0 && parts[i].data != nullptr) {  // a NULL output is skipped (ref concat_and_split_functor.cc:115)
      MapDesc d{};
      char* slot = reinterpret_cast<char*>(flat->data) + off * es * words;
      d.src = is_concat ? parts[i].data : (const void*)slot;
      d.dst = is_concat ? (void*)slot : parts[i].data;
      d.aligned = map_aligned(d.src, d.dst, es) ? 1 : 0;
      descs.push_back(d);
      numels.push_back(parts[i].numel * words);
    }
    off += parts[i].numel;
  }
  PRB_ENFORCE(off == flat->numel, PRB_INVALID_ARGUMENT, "%s: sizes of the parts (%lld) do not add up to %lld", op,
              (long long)off, (long long)flat->numel);
  if (descs.empty()) return PRB_SUCCESS;
  DeviceTable tab;
  PRB_TRY(get_table(descs.data(), sizeof(MapDesc), (int)descs.size(), numels.data(), map_chunk_elems(), stream, &tab));
  return launch_copy(es, tab, stream);
}

int32_t prb_concat(const PRB_Tensor* ins, int64_t num, PRB_Tensor* out, void* stream) {
  return concat_split(out, ins, num, true, (cudaStream_t)stream);
}

int32_t prb_split(const PRB_Tensor* in, PRB_Tensor* outs, int64_t num, void* stream) {
  return concat_split(in, outs, num, false, (cudaStream_t)stream);
}

}  // extern "C"

// ------------------------------------------------------------------------------------------
// cast (init-time helper of the host shim: master weights = fp32 copy of 16-bit params,
// ref python/paddle/optimizer/optimizer.py:640-666; not on the per-step path)
// ------------------------------------------------------------------------------------------
namespace prb {
namespace {
template <typename TI, typename TO>
__global__ void cast_kernel(const TI* __restrict__ in, TO* __restrict__ out, int64_t n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    out[i] = Conv<TO>::from_f32(Conv<TI>::to_f32(in[i]));
}
template <typename TI>
int32_t cast_from(const void* in, void* out, int32_t dt_out, int64_t n, cudaStream_t s) {
  const int blocks = (int)((n + 255) / 256 < 148 * 16 ? (n + 255) / 256 : 148 * 16);
  switch (dt_out) {
    case PRB_FLOAT32: cast_kernel<TI, float><<<blocks, 256, 0, s>>>((const TI*)in, (float*)out, n); break;
    case PRB_BFLOAT16: cast_kernel<TI, __nv_bfloat16><<<blocks, 256, 0, s>>>((const TI*)in, (__nv_bfloat16*)out, n); break;
    case PRB_FLOAT16: cast_kernel<TI, __half><<<blocks, 256, 0, s>>>((const TI*)in, (__half*)out, n); break;
    default: PRB_FAIL(PRB_UNIMPLEMENTED, "cast: dtype %s is not built", dtype_name(dt_out));
  }
  PRB_LAUNCH_CHECK();
  return PRB_SUCCESS;
}
}  // namespace
}  // namespace prb

namespace prb {
namespace {
// out = T(MT(x) * *y): elementwise multiply by a one-element float32 device tensor (the broadcast multiply behind
// AmpScaler.scale, ref python/paddle/amp/grad_scaler.py:152-200 `var * self._scale`)
template <typename T>
__global__ void mul_scalar_tensor_kernel(const T* __restrict__ x, const float* __restrict__ y, T* __restrict__ out, int64_t n) {
  const float s = __ldg(y);
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    out[i] = Conv<T>::from_f32(__fmul_rn(Conv<T>::to_f32(x[i]), s));
}
}  // namespace
}  // namespace prb

extern "C" int32_t prb_multiply_by_scalar_tensor(const PRB_Tensor* x, const PRB_Tensor* y, PRB_Tensor* out, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  PRB_TRY(prb::require_device());
  PRB_ENFORCE(x && y && out && x->numel == out->numel && x->dtype == out->dtype, PRB_INVALID_ARGUMENT,
              "multiply: x/out mismatch");
  PRB_ENFORCE(y->numel == 1 && y->dtype == PRB_FLOAT32 && y->device >= 0 && x->device >= 0 && out->device >= 0,
              PRB_INVALID_ARGUMENT, "multiply: y must be one float32 on the device, x and out device tensors");
  if (x->numel == 0) return PRB_SUCCESS;
  const int64_t n = x->numel;
  const int blocks = (int)((n + 255) / 256 < 148 * 16 ? (n + 255) / 256 : 148 * 16);
  const float* yp = reinterpret_cast<const float*>(y->data);
  switch (x->dtype) {
    case PRB_FLOAT32: prb::mul_scalar_tensor_kernel<float><<<blocks, 256, 0, stream>>>((const float*)x->data, yp, (float*)out->data, n); break;
    case PRB_BFLOAT16: prb::mul_scalar_tensor_kernel<__nv_bfloat16><<<blocks, 256, 0, stream>>>((const __nv_bfloat16*)x->data, yp, (__nv_bfloat16*)out->data, n); break;
    case PRB_FLOAT16: prb::mul_scalar_tensor_kernel<__half><<<blocks, 256, 0, stream>>>((const __half*)x->data, yp, (__half*)out->data, n); break;
    default: PRB_FAIL(PRB_UNIMPLEMENTED, "multiply: dtype %s is not built", prb::dtype_name(x->dtype));
  }
  PRB_LAUNCH_CHECK();
  return PRB_SUCCESS;
}

extern "C" int32_t prb_cast(const PRB_Tensor* x, PRB_Tensor* out, void* stream) {
  PRB_TRY(prb::require_device());
  PRB_ENFORCE(x && out && x->numel == out->numel, PRB_INVALID_ARGUMENT, "cast: x/out mismatch");
  PRB_ENFORCE(x->device >= 0 && out->device >= 0, PRB_INVALID_ARGUMENT, "cast: device tensors required");
  if (x->numel == 0) return PRB_SUCCESS;
  switch (x->dtype) {
    case PRB_FLOAT32: return prb::cast_from<float>(x->data, out->data, out->dtype, x->numel, (cudaStream_t)stream);
    case PRB_BFLOAT16: return prb::cast_from<__nv_bfloat16>(x->data, out->data, out->dtype, x->numel, (cudaStream_t)stream);
    case PRB_FLOAT16: return prb::cast_from<__half>(x->data, out->data, out->dtype, x->numel, (cudaStream_t)stream);
    default: PRB_FAIL(PRB_UNIMPLEMENTED, "cast: dtype %s is not built", prb::dtype_name(x->dtype));
  }
}
