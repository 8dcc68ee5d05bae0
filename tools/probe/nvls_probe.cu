// Probe: is NVLink multicast (NVLS) usable from user code on this box?  Single process, all visible GPUs.
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
#define CK(x) do { CUresult r_ = (x); if (r_ != CUDA_SUCCESS) { const char* s_; cuGetErrorString(r_, &s_); printf("FAIL %s -> %s\n", #x, s_); return 1; } } while (0)

__global__ void fill(float* p, float v, size_t n) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}
__global__ void reduce_bcast(float* mc, size_t n4) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i < n4) {
    float a, b, c, d;
    asm volatile("multimem.ld_reduce.global.add.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(a), "=f"(b), "=f"(c), "=f"(d) : "l"(mc + 4 * i) : "memory");
    asm volatile("multimem.st.global.v4.f32 [%0], {%1,%2,%3,%4};" :: "l"(mc + 4 * i), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
  }
}

int main() {
  CK(cuInit(0));
  int ndev = 0;
  cudaGetDeviceCount(&ndev);
  printf("devices %d\n", ndev);
  for (int d = 0; d < ndev; ++d) {
    int mc = 0;
    CK(cuDeviceGetAttribute(&mc, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, d));
    printf("device %d multicast_supported %d\n", d, mc);
    if (!mc) return 2;
  }
  const size_t want = 64ull << 20;
  CUmulticastObjectProp mp = {};
  mp.numDevices = ndev;
  mp.size = want;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  size_t gran = 0;
  CK(cuMulticastGetGranularity(&gran, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED));
  printf("multicast granularity %zu\n", gran);
  mp.size = (want + gran - 1) / gran * gran;
  CUmemGenericAllocationHandle mch;
  CK(cuMulticastCreate(&mch, &mp));
  int fd = -1;
  CK(cuMemExportToShareableHandle(&fd, mch, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  printf("exported fd %d\n", fd);
  for (int d = 0; d < ndev; ++d) CK(cuMulticastAddDevice(mch, d));
  std::vector<CUdeviceptr> uc(ndev), mcva(ndev);
  for (int d = 0; d < ndev; ++d) {
    cudaSetDevice(d);
    cudaFree(0);
    CUmemAllocationProp ap = {};
    ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    ap.location.id = d;
    ap.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t g2 = 0;
    CK(cuMemGetAllocationGranularity(&g2, &ap, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
    size_t sz = (mp.size + g2 - 1) / g2 * g2;
    CUmemGenericAllocationHandle h;
    CK(cuMemCreate(&h, sz, &ap, 0));
    CK(cuMulticastBindMem(mch, 0, h, 0, mp.size, 0));
    CUmemAccessDesc ad = {};
    ad.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    ad.location.id = d;
    ad.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    CK(cuMemAddressReserve(&uc[d], sz, g2, 0, 0));
    CK(cuMemMap(uc[d], sz, 0, h, 0));
    CK(cuMemSetAccess(uc[d], sz, &ad, 1));
    CK(cuMemAddressReserve(&mcva[d], mp.size, gran, 0, 0));
    CK(cuMemMap(mcva[d], mp.size, 0, mch, 0));
    CK(cuMemSetAccess(mcva[d], mp.size, &ad, 1));
    size_t n = mp.size / 4;
    fill<<<(unsigned)((n + 255) / 256), 256>>>((float*)uc[d], (float)(d + 1), n);
    cudaDeviceSynchronize();
  }
  cudaSetDevice(0);
  size_t n4 = mp.size / 16;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  reduce_bcast<<<(unsigned)((n4 + 255) / 256), 256>>>((float*)mcva[0], n4);
  cudaError_t ce = cudaDeviceSynchronize();
  printf("kernel: %s\n", cudaGetErrorString(ce));
  cudaEventRecord(e0);
  for (int i = 0; i < 10; ++i) reduce_bcast<<<(unsigned)((n4 + 255) / 256), 256>>>((float*)mcva[0], n4);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  printf("one GPU reducing+broadcasting ALL %zu MiB: %.1f us per pass\n", mp.size >> 20, ms / 10 * 1e3);
  float want_sum = ndev * (ndev + 1) / 2.0f;
  for (int d = 0; d < ndev; ++d) {
    cudaSetDevice(d);
    float h[4];
    cudaMemcpy(h, (void*)uc[d], 16, cudaMemcpyDeviceToHost);
    printf("device %d first element %g (first pass expected %g)\n", d, h[0], want_sum);
  }
  return 0;
}
