// In-place bucket allreduce through NVLink multicast ("NVLS": the NVSwitch does the arithmetic).
//
// The reducer's bucket arena is allocated with the virtual-memory API and bound, on every rank, to ONE multicast
// object that spans all GPUs of the job.  Through the multicast address
//   multimem.ld_reduce.add  returns the SUM over all ranks' copies of a vector -- the switch adds while it gathers,
//   multimem.st             writes a vector into all ranks' copies at once      -- the switch replicates.
// One kernel per rank and bucket: ready barrier -> for its slice (1/n of the bucket): v = ld_reduce; st(v) ->
// done barrier.  Per rank a bucket of S bytes costs about S out + S in on NVLink instead of 2(n-1)/n * S each way
// for the peer-memory two-shot kernel (prb_nvlink.cu) -- 1.75x less at n = 8 -- and the adds cost no SM time.
// Every element is reduced by exactly one rank and then broadcast, so results are bit-equal across ranks; for two
// ranks they equal a + b exactly; 16-bit types accumulate in fp32 inside the switch (.acc::f32).
//
// The multicast handle travels from rank 0 to its peers as a POSIX file descriptor over a unix-domain socket
// (SCM_RIGHTS), like NCCL does; driver entry points are resolved at run time (cudaGetDriverEntryPoint), so the
// library still loads where libcuda is absent.  Any failure on any rank makes ALL ranks fall back (collective vote).
#include <cuda.h>
#include <sys/socket.h>
#include <sys/un.h>
#include <unistd.h>

#include <cerrno>
#include <cstring>

#include "prb_nvl_common.cuh"
#include "prb_nvls.h"

namespace prb {
namespace {

// ---------------------------------------------------------------- driver API, resolved at run time
struct Drv {
  CUresult (*MulticastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*) = nullptr;
  CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice) = nullptr;
  CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t,
                               unsigned long long) = nullptr;
  CUresult (*MulticastGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags) = nullptr;
  CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long) = nullptr;
  CUresult (*MemRelease)(CUmemGenericAllocationHandle) = nullptr;
  CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
  CUresult (*MemAddressFree)(CUdeviceptr, size_t) = nullptr;
  CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
  CUresult (*MemUnmap)(CUdeviceptr, size_t) = nullptr;
  CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t) = nullptr;
  CUresult (*MemExportToShareableHandle)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType,
                                         unsigned long long) = nullptr;
  CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType) = nullptr;
  CUresult (*MemGetAllocationGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags) = nullptr;
  CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice) = nullptr;
  bool ok = false;
};

Drv& drv() {
  static Drv d = [] {
    Drv x;
    bool ok = true;
    auto get = [&](const char* name, void** fn) {
      cudaDriverEntryPointQueryResult q;
      if (cudaGetDriverEntryPoint(name, fn, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess ||
          *fn == nullptr)
        ok = false;
    };
    get("cuMulticastCreate", (void**)&x.MulticastCreate);
    get("cuMulticastAddDevice", (void**)&x.MulticastAddDevice);
    get("cuMulticastBindMem", (void**)&x.MulticastBindMem);
    get("cuMulticastGetGranularity", (void**)&x.MulticastGetGranularity);
    get("cuMemCreate", (void**)&x.MemCreate);
    get("cuMemRelease", (void**)&x.MemRelease);
    get("cuMemAddressReserve", (void**)&x.MemAddressReserve);
    get("cuMemAddressFree", (void**)&x.MemAddressFree);
    get("cuMemMap", (void**)&x.MemMap);
    get("cuMemUnmap", (void**)&x.MemUnmap);
    get("cuMemSetAccess", (void**)&x.MemSetAccess);
    get("cuMemExportToShareableHandle", (void**)&x.MemExportToShareableHandle);
    get("cuMemImportFromShareableHandle", (void**)&x.MemImportFromShareableHandle);
    get("cuMemGetAllocationGranularity", (void**)&x.MemGetAllocationGranularity);
    get("cuDeviceGetAttribute", (void**)&x.DeviceGetAttribute);
    cudaGetLastError();
    x.ok = ok;
    return x;
  }();
  return d;
}

// ---------------------------------------------------------------- file descriptor passing (rank 0 -> peers)
void socket_path(sockaddr_un* addr, int pid0, int serial) {
  memset(addr, 0, sizeof(*addr));
  addr->sun_family = AF_UNIX;
  // abstract namespace (leading NUL): nothing to unlink, vanishes with the process
  snprintf(addr->sun_path + 1, sizeof(addr->sun_path) - 1, "prb_nvls_%d_%d", pid0, serial);
}

bool send_fd(int sock, int fd) {
  char byte = 'x';
  iovec iov{&byte, 1};
  char ctrl[CMSG_SPACE(sizeof(int))];
  memset(ctrl, 0, sizeof(ctrl));
  msghdr msg{};
  msg.msg_iov = &iov;
  msg.msg_iovlen = 1;
  msg.msg_control = ctrl;
  msg.msg_controllen = sizeof(ctrl);
  cmsghdr* c = CMSG_FIRSTHDR(&msg);
  c->cmsg_level = SOL_SOCKET;
  c->cmsg_type = SCM_RIGHTS;
  c->cmsg_len = CMSG_LEN(sizeof(int));
  memcpy(CMSG_DATA(c), &fd, sizeof(int));
  return sendmsg(sock, &msg, 0) == 1;
}

int recv_fd(int sock) {
  char byte = 0;
  iovec iov{&byte, 1};
  char ctrl[CMSG_SPACE(sizeof(int))];
  memset(ctrl, 0, sizeof(ctrl));
  msghdr msg{};
  msg.msg_iov = &iov;
  msg.msg_iovlen = 1;
  msg.msg_control = ctrl;
  msg.msg_controllen = sizeof(ctrl);
  if (recvmsg(sock, &msg, 0) != 1) return -1;
  cmsghdr* c = CMSG_FIRSTHDR(&msg);
  if (c == nullptr || c->cmsg_type != SCM_RIGHTS) return -1;
  int fd = -1;
  memcpy(&fd, CMSG_DATA(c), sizeof(int));
  return fd;
}

// rank 0: serve `fd` to nranks-1 peers; returns false on any error (with a 20 s limit per accept)
bool serve_fd(int listen_sock, int fd, int npeers) {
  timeval tv{20, 0};
  setsockopt(listen_sock, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof(tv));
  bool ok = true;
  for (int i = 0; i < npeers; ++i) {
    const int s = accept(listen_sock, nullptr, nullptr);
    if (s < 0) return false;
    ok = send_fd(s, fd) && ok;
    close(s);
  }
  return ok;
}

int fetch_fd(int pid0, int serial) {
  sockaddr_un addr;
  socket_path(&addr, pid0, serial);
  for (int attempt = 0; attempt < 2000; ++attempt) {  // up to ~20 s
    const int s = socket(AF_UNIX, SOCK_STREAM, 0);
    if (s < 0) return -1;
    if (connect(s, (sockaddr*)&addr, sizeof(sa_family_t) + 1 + strlen(addr.sun_path + 1)) == 0) {
      timeval tv{20, 0};
      setsockopt(s, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof(tv));
      const int fd = recv_fd(s);
      close(s);
      return fd;
    }
    close(s);
    usleep(10000);
  }
  return -1;
}

size_t round_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// ---------------------------------------------------------------- kernel
struct NvlsArgs {
  Bar bar;                   // barriers (prb_nvl_common.cuh)
  char* mc;                  // multicast address of the arena
  size_t region_off;
  long long nvec;            // 16-byte vectors in the bucket
};

struct V16 {
  uint32_t a, b, c, d;
};
template <typename T>
__device__ __forceinline__ V16 mc_load_reduce(const char* p) {
  V16 v;
  if constexpr (sizeof(T) == 4) {
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.a), "=r"(v.b), "=r"(v.c), "=r"(v.d) : "l"(p) : "memory");
  } else if constexpr (std::is_same<T, __nv_bfloat16>::value) {
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.a), "=r"(v.b), "=r"(v.c), "=r"(v.d) : "l"(p) : "memory");
  } else {
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.a), "=r"(v.b), "=r"(v.c), "=r"(v.d) : "l"(p) : "memory");
  }
  return v;
}
__device__ __forceinline__ void mc_store(char* p, const V16& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.a), "r"(v.b), "r"(v.c),
               "r"(v.d)
               : "memory");
}
constexpr int kNvlsUnroll = 8;  // 16-byte vectors in flight per thread: the round trip through the switch is long

template <typename T>
__global__ void __launch_bounds__(kNvlThreads, 2) nvls_allreduce_kernel(const NvlsArgs a) {
  const int slot = (int)(a.bar.seq & (kSlots - 1));
  // ---- ready barrier: every rank's packed bucket is complete
  if (!ready_barrier(a.bar, slot)) return;
  // ---- my slice: the switch reduces on the way in and replicates on the way out
  const long long v0 = a.nvec * a.bar.rank / a.bar.nranks, v1 = a.nvec * (a.bar.rank + 1) / a.bar.nranks;
  const long long stride = (long long)gridDim.x * kNvlThreads;
  char* base = a.mc + a.region_off;
  for (long long v = v0 + (long long)blockIdx.x * kNvlThreads + threadIdx.x; v < v1; v += stride * kNvlsUnroll) {
    V16 r[kNvlsUnroll];
#pragma unroll
    for (int u = 0; u < kNvlsUnroll; ++u)
      if (v + u * stride < v1) r[u] = mc_load_reduce<T>(base + (size_t)(v + u * stride) * 16);
#pragma unroll
    for (int u = 0; u < kNvlsUnroll; ++u)
      if (v + u * stride < v1) mc_store(base + (size_t)(v + u * stride) * 16, r[u]);
  }
  // ---- done barrier
  done_barrier(a.bar, slot);
}

}  // namespace

struct NvlsContext {
  NvlsArgs args;
  SignalSet sigs;
  CUmemGenericAllocationHandle mch = 0, mem = 0;
  CUdeviceptr uc = 0, mc = 0;
  bool uc_mapped = false, mc_mapped = false, have_mch = false, have_mem = false;
  size_t size = 0;
  unsigned long long seq = 0;
  int ctas = 0;
};

namespace {
int g_serial = 0;  // setup calls are collective and ordered, so the count is the same on every rank

int32_t vote(const NvlCollectives& nc, int ok, cudaStream_t s, int* agreed) { return nc.all_min_int(nc.comm, ok, agreed, s); }
}  // namespace

void nvls_destroy(NvlsContext* c) {
  if (!c) return;
  Drv& d = drv();
  if (c->mc_mapped) d.MemUnmap(c->mc, c->size);
  if (c->mc) d.MemAddressFree(c->mc, c->size);
  if (c->uc_mapped) d.MemUnmap(c->uc, c->size);
  if (c->uc) d.MemAddressFree(c->uc, c->size);
  if (c->have_mem) d.MemRelease(c->mem);
  if (c->have_mch) d.MemRelease(c->mch);
  signals_destroy(&c->sigs);
  delete c;
}

int32_t nvls_setup(const NvlCollectives& nc, int nranks, int rank, int device, size_t bytes, cudaStream_t stream,
                   NvlsContext** out, void** arena_out) {
  *out = nullptr;
  *arena_out = nullptr;
  if (nranks < 2 || nranks > kMaxRanks) return PRB_SUCCESS;
  const int serial = g_serial++;
  Drv& d = drv();
  int agreed = 0;
  // ---- 0: is multicast there at all?
  int ok = d.ok ? 1 : 0;
  if (ok) {
    int sup = 0;
    if (d.DeviceGetAttribute(&sup, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, device) != CUDA_SUCCESS || !sup) ok = 0;
  }
  PRB_TRY(vote(nc, ok, stream, &agreed));
  if (!agreed) return PRB_SUCCESS;

  NvlsContext* c = new NvlsContext();
  memset(&c->args, 0, sizeof(c->args));
  // ---- 1: sizes
  CUmulticastObjectProp mp{};
  mp.numDevices = (unsigned)nranks;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  mp.size = bytes;
  CUmemAllocationProp ap{};
  ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  ap.location.id = device;
  ap.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  size_t g_mc = 0, g_mem = 0;
  if (d.MulticastGetGranularity(&g_mc, &mp, CU_MULTICAST_GRANULARITY_MINIMUM) != CUDA_SUCCESS) ok = 0;
  if (d.MemGetAllocationGranularity(&g_mem, &ap, CU_MEM_ALLOC_GRANULARITY_MINIMUM) != CUDA_SUCCESS) ok = 0;
  if (ok) {
    const size_t g = g_mc > g_mem ? g_mc : g_mem;  // both are powers of two
    c->size = round_up(bytes < 1 ? 1 : bytes, g);
    mp.size = c->size;
  }
  // ---- 2: everybody learns rank 0's pid (names the rendezvous socket)
  int pid = (int)getpid(), pids[kMaxRanks] = {0};
  {
    void *ds = nullptr, *dr = nullptr;
    PRB_CUDA(cudaMalloc(&ds, sizeof(int)));
    PRB_CUDA(cudaMalloc(&dr, sizeof(int) * nranks));
    PRB_CUDA(cudaMemcpyAsync(ds, &pid, sizeof(int), cudaMemcpyHostToDevice, stream));
    PRB_TRY(nc.all_gather_bytes(nc.comm, ds, dr, sizeof(int), stream));
    PRB_CUDA(cudaMemcpyAsync(pids, dr, sizeof(int) * nranks, cudaMemcpyDeviceToHost, stream));
    PRB_CUDA(cudaStreamSynchronize(stream));
    cudaFree(ds);
    cudaFree(dr);
  }
  // ---- 3: rank 0 creates the multicast object and hands its file descriptor to the peers
  int fd = -1, lsock = -1;
  if (rank == 0 && ok) {
    if (d.MulticastCreate(&c->mch, &mp) == CUDA_SUCCESS) c->have_mch = true;
    else ok = 0;
    if (ok && d.MemExportToShareableHandle(&fd, c->mch, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0) != CUDA_SUCCESS) ok = 0;
    if (ok) {
      sockaddr_un addr;
      socket_path(&addr, pid, serial);
      lsock = socket(AF_UNIX, SOCK_STREAM, 0);
      if (lsock < 0 || bind(lsock, (sockaddr*)&addr, sizeof(sa_family_t) + 1 + strlen(addr.sun_path + 1)) != 0 ||
          listen(lsock, nranks) != 0)
        ok = 0;
    }
  }
  PRB_TRY(vote(nc, ok, stream, &agreed));
  if (agreed) {
    if (rank == 0) {
      if (!serve_fd(lsock, fd, nranks - 1)) ok = 0;
    } else {
      fd = fetch_fd(pids[0], serial);
      if (fd < 0) ok = 0;
      if (ok && d.MemImportFromShareableHandle(&c->mch, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR) ==
                    CUDA_SUCCESS)
        c->have_mch = true;
      else
        ok = 0;
    }
  }
  if (lsock >= 0) close(lsock);
  if (fd >= 0) close(fd);
  if (agreed) PRB_TRY(vote(nc, ok, stream, &agreed));
  // ---- 4: every device joins (all must have joined before memory is bound)
  if (agreed) {
    if (d.MulticastAddDevice(c->mch, device) != CUDA_SUCCESS) ok = 0;
    PRB_TRY(vote(nc, ok, stream, &agreed));
  }
  // ---- 5: this rank's arena: physical memory, unicast mapping, zero fill, bind into the multicast object
  CUmemAccessDesc ad{};
  ad.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  ad.location.id = device;
  ad.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  if (agreed) {
    if (d.MemCreate(&c->mem, c->size, &ap, 0) == CUDA_SUCCESS) c->have_mem = true;
    else ok = 0;
    if (ok && d.MemAddressReserve(&c->uc, c->size, g_mc > g_mem ? g_mc : g_mem, 0, 0) != CUDA_SUCCESS) ok = 0;
    if (ok && d.MemMap(c->uc, c->size, 0, c->mem, 0) == CUDA_SUCCESS) c->uc_mapped = true;
    else ok = 0;
    if (ok && d.MemSetAccess(c->uc, c->size, &ad, 1) != CUDA_SUCCESS) ok = 0;
    if (ok && cudaMemset((void*)c->uc, 0, c->size) != cudaSuccess) ok = 0;
    if (ok && d.MulticastBindMem(c->mch, 0, c->mem, 0, c->size, 0) != CUDA_SUCCESS) ok = 0;
    cudaGetLastError();
    PRB_TRY(vote(nc, ok, stream, &agreed));
  }
  // ---- 6: multicast mapping
  if (agreed) {
    if (d.MemAddressReserve(&c->mc, c->size, g_mc > g_mem ? g_mc : g_mem, 0, 0) != CUDA_SUCCESS) ok = 0;
    if (ok && d.MemMap(c->mc, c->size, 0, c->mch, 0) == CUDA_SUCCESS) c->mc_mapped = true;
    else ok = 0;
    if (ok && d.MemSetAccess(c->mc, c->size, &ad, 1) != CUDA_SUCCESS) ok = 0;
    PRB_TRY(vote(nc, ok, stream, &agreed));
  }
  // ---- 7: signal blocks for the barriers
  if (agreed) {
    int sok = 0;
    signals_setup(nc, nranks, rank, stream, &c->sigs, &sok);
    ok = ok && sok;
    PRB_TRY(vote(nc, ok, stream, &agreed));
  }
  if (!agreed) {
    nvls_destroy(c);
    return PRB_SUCCESS;  // the caller falls back to peer memory / NCCL
  }
  c->args.mc = reinterpret_cast<char*>(c->mc);
  fill_bar(c->sigs, nranks, rank, &c->args.bar);
  int per_sm = 1;
  if (const char* e = getenv("PRB_NVL_CTAS_PER_SM")) per_sm = atoi(e) > 0 ? atoi(e) : per_sm;
  c->ctas = sm_count() * per_sm;
  *out = c;
  *arena_out = reinterpret_cast<void*>(c->uc);
  return PRB_SUCCESS;
}

int nvls_next_slot(const NvlsContext* c) { return c ? (int)((c->seq + 1) & (kSlots - 1)) : 0; }

int32_t nvls_allreduce(NvlsContext* c, size_t region_off_bytes, int64_t numel, int32_t dtype, cudaStream_t stream) {
  PRB_ENFORCE(c != nullptr, PRB_INVALID_ARGUMENT, "nvls allreduce: no context");
  const size_t bytes = (size_t)numel * dtype_size(dtype);
  PRB_ENFORCE(bytes % 16 == 0 && region_off_bytes % 16 == 0, PRB_INVALID_ARGUMENT,
              "nvls allreduce: the bucket must be a whole number of 16-byte vectors at a 16-byte offset");
  NvlsArgs a = c->args;
  a.region_off = region_off_bytes;
  a.nvec = (long long)(bytes / 16);
  a.bar.seq = ++c->seq;
  const long long per_rank = (a.nvec + a.bar.nranks - 1) / a.bar.nranks;
  int ctas = (int)((per_rank + kNvlThreads - 1) / kNvlThreads);
  if (ctas > c->ctas) ctas = c->ctas;
  if (ctas < 1) ctas = 1;
  KernelTimer kt(KK_EXCHANGE, stream);
  switch (dtype) {
    case PRB_FLOAT32: PRB_CUDA(nvl_launch(nvls_allreduce_kernel<float>, ctas, kNvlThreads, stream, a)); break;
    case PRB_BFLOAT16: PRB_CUDA(nvl_launch(nvls_allreduce_kernel<__nv_bfloat16>, ctas, kNvlThreads, stream, a)); break;
    case PRB_FLOAT16: PRB_CUDA(nvl_launch(nvls_allreduce_kernel<__half>, ctas, kNvlThreads, stream, a)); break;
    default: PRB_FAIL(PRB_UNIMPLEMENTED, "nvls allreduce: dtype %s is not built", dtype_name(dtype));
  }
  PRB_LAUNCH_CHECK();
  return PRB_SUCCESS;
}

int nvls_error_flag(const NvlsContext* c) {
  return (c != nullptr && c->sigs.host_error != nullptr) ? (int)*reinterpret_cast<volatile unsigned int*>(c->sigs.host_error) : 0;
}

}  // namespace prb
