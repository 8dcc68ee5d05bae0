// NVLink peer-memory characterisation on two GPUs of one process: how fast can SM-issued remote LOADS (pull) and
// remote STORES (push) go, as a function of the vector width, the loads in flight per thread and the grid?
// Both GPUs run the same kernel at the same time (both link directions loaded, as in the exchange kernels).
//
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/probe/p2p_probe tools/probe/p2p_probe.cu
//   tools/probe/p2p_probe [MiB]
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x)                                                                          \
  do {                                                                                 \
    cudaError_t e_ = (x);                                                              \
    if (e_ != cudaSuccess) {                                                           \
      printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
      exit(1);                                                                         \
    }                                                                                  \
  } while (0)

struct alignas(32) V32 {
  unsigned int w[8];
};
struct alignas(16) V16 {
  unsigned int w[4];
};

__device__ __forceinline__ void ld(const V32* p, V32& v) {
  asm volatile("ld.global.L1::no_allocate.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(v.w[0]), "=r"(v.w[1]), "=r"(v.w[2]), "=r"(v.w[3]), "=r"(v.w[4]), "=r"(v.w[5]), "=r"(v.w[6]), "=r"(v.w[7])
               : "l"(p));
}
__device__ __forceinline__ void st(V32* p, const V32& v) {
  asm volatile("st.global.v8.u32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "r"(v.w[0]), "r"(v.w[1]), "r"(v.w[2]),
               "r"(v.w[3]), "r"(v.w[4]), "r"(v.w[5]), "r"(v.w[6]), "r"(v.w[7])
               : "memory");
}
__device__ __forceinline__ void ld(const V16* p, V16& v) {
  asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.w[0]), "=r"(v.w[1]), "=r"(v.w[2]), "=r"(v.w[3]) : "l"(p));
}
__device__ __forceinline__ void st(V16* p, const V16& v) {
  asm volatile("st.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.w[0]), "r"(v.w[1]), "r"(v.w[2]), "r"(v.w[3]) : "memory");
}

// dst[i] = src[i]; U vectors in flight per thread.  Which side is remote is the caller's choice of pointers.
template <typename V, int U>
__global__ void __launch_bounds__(256) copy_kernel(const V* __restrict__ src, V* __restrict__ dst, long long n) {
  const long long stride = (long long)gridDim.x * 256;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += stride * U) {
    V v[U];
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (i + u * stride < n) ld(src + i + u * stride, v[u]);
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (i + u * stride < n) st(dst + i + u * stride, v[u]);
  }
}

// allreduce-like mix: read remote AND local, write remote AND local (what the two-rank exchange kernel moves)
template <typename V, int U>
__global__ void __launch_bounds__(256) mix_kernel(const V* __restrict__ rsrc, const V* __restrict__ lsrc, V* __restrict__ rdst,
                                                  V* __restrict__ ldst, long long n) {
  const long long stride = (long long)gridDim.x * 256;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += stride * U) {
    V a[U], b[U];
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (i + u * stride < n) {
        ld(rsrc + i + u * stride, a[u]);
        ld(lsrc + i + u * stride, b[u]);
      }
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (i + u * stride < n) {
        a[u].w[0] ^= b[u].w[0];
        st(rdst + i + u * stride, a[u]);
        st(ldst + i + u * stride, a[u]);
      }
  }
}

template <typename V, int U>
void run(const char* what, int mode, int ctas, char* buf[2], char* buf2[2], size_t bytes, cudaStream_t s[2], cudaEvent_t e0[2],
         cudaEvent_t e1[2]) {
  const long long n = (long long)(bytes / sizeof(V));
  const int iters = 10;
  float best = 0;
  for (int rep = 0; rep < 2; ++rep) {
    for (int d = 0; d < 2; ++d) {
      CK(cudaSetDevice(d));
      CK(cudaEventRecord(e0[d], s[d]));
      for (int it = 0; it < iters; ++it) {
        const V* remote = reinterpret_cast<const V*>(buf[1 - d]);
        V* local = reinterpret_cast<V*>(buf2[d]);
        if (mode == 0) copy_kernel<V, U><<<ctas, 256, 0, s[d]>>>(remote, local, n);                                      // pull
        else if (mode == 1) copy_kernel<V, U><<<ctas, 256, 0, s[d]>>>(reinterpret_cast<const V*>(buf[d]), reinterpret_cast<V*>(buf2[1 - d]), n);  // push
        else mix_kernel<V, U><<<ctas, 256, 0, s[d]>>>(remote, reinterpret_cast<const V*>(buf[d]), reinterpret_cast<V*>(buf2[1 - d]), local, n / 2);
      }
      CK(cudaEventRecord(e1[d], s[d]));
    }
    float ms[2];
    for (int d = 0; d < 2; ++d) {
      CK(cudaSetDevice(d));
      CK(cudaEventSynchronize(e1[d]));
      CK(cudaEventElapsedTime(&ms[d], e0[d], e1[d]));
    }
    best = (ms[0] > ms[1] ? ms[0] : ms[1]) / iters;
  }
  const double moved = mode == 2 ? (double)bytes / 2 : (double)bytes;  // bytes per link direction per kernel
  printf("%-5s V%-2d U%-2d ctas %4d  %8.1f us  %7.1f GB/s per direction%s\n", what, (int)sizeof(V), U, ctas, best * 1e3,
         moved / (best * 1e-3) / 1e9, mode == 2 ? " of reads + the same of writes" : "");
  fflush(stdout);
}

int main(int argc, char** argv) {
  const size_t bytes = (size_t)(argc > 1 ? atoi(argv[1]) : 256) << 20;
  int ndev = 0;
  CK(cudaGetDeviceCount(&ndev));
  if (ndev < 2) {
    printf("needs 2 GPUs\n");
    return 0;
  }
  char *buf[2], *buf2[2];
  cudaStream_t s[2];
  cudaEvent_t e0[2], e1[2];
  for (int d = 0; d < 2; ++d) {
    CK(cudaSetDevice(d));
    CK(cudaDeviceEnablePeerAccess(1 - d, 0));
    CK(cudaMalloc(&buf[d], bytes));
    CK(cudaMalloc(&buf2[d], bytes));
    CK(cudaMemset(buf[d], d + 1, bytes));
    CK(cudaStreamCreateWithFlags(&s[d], cudaStreamNonBlocking));
    CK(cudaEventCreate(&e0[d]));
    CK(cudaEventCreate(&e1[d]));
  }
  for (int d = 0; d < 2; ++d) {
    CK(cudaSetDevice(d));
    CK(cudaDeviceSynchronize());
  }
  printf("%zu MiB per kernel, both GPUs at once\n", bytes >> 20);
  const int grids[] = {32, 64, 148, 296, 592};
  for (int mode = 0; mode < 3; ++mode) {
    const char* what = mode == 0 ? "pull" : (mode == 1 ? "push" : "mix");
    for (int g : grids) {
      run<V16, 4>(what, mode, g, buf, buf2, bytes, s, e0, e1);
      run<V16, 16>(what, mode, g, buf, buf2, bytes, s, e0, e1);
      run<V32, 4>(what, mode, g, buf, buf2, bytes, s, e0, e1);
      run<V32, 8>(what, mode, g, buf, buf2, bytes, s, e0, e1);
      if (g <= 148) run<V32, 16>(what, mode, g, buf, buf2, bytes, s, e0, e1);
    }
  }
  // copy engine for reference
  for (int d = 0; d < 2; ++d) {
    CK(cudaSetDevice(d));
    CK(cudaEventRecord(e0[d], s[d]));
    for (int it = 0; it < 10; ++it) CK(cudaMemcpyPeerAsync(buf2[d], d, buf[1 - d], 1 - d, bytes, s[d]));
    CK(cudaEventRecord(e1[d], s[d]));
  }
  for (int d = 0; d < 2; ++d) {
    float ms;
    CK(cudaSetDevice(d));
    CK(cudaEventSynchronize(e1[d]));
    CK(cudaEventElapsedTime(&ms, e0[d], e1[d]));
    printf("cudaMemcpyPeerAsync into GPU %d: %.1f GB/s\n", d, (double)bytes / (ms / 10 * 1e-3) / 1e9);
  }
  return 0;
}
