// Cross-GPU signalling shared by the NVLink peer-memory allreduce (prb_nvlink.cu) and the NVLS multicast
// allreduce (prb_nvls.cu): per-rank signal blocks exported through CUDA IPC, sequence-number barriers.
#pragma once
#include "prb_nvlink.h"

namespace prb {

constexpr int kMaxRanks = 8;
constexpr int kNvlThreads = 256;
constexpr int kNvlMinCtasPerSm = 4;  // register cap (<= 64): a CTA must find room next to the optimizer's CTAs
// CTAs launched per SM.  The kernel is NVLink-bound and its CTAs mostly wait; 1 per SM (256 threads, ~12 % of an
// SM's thread slots) leaves the SM to the optimizer kernels that run next to it: ResNet-50 N=2 step 0.30 ms vs
// 0.35 ms with 2 or 4 per SM, same stand-alone bandwidth (profiles/r1_multi_gpu.md).
constexpr int kNvlCtasPerSm = 1;
constexpr unsigned long long kWaitLimitNs = 4000000000ull;  // 4 s

// Two allreduces may be in flight at once (the reducer alternates two comm streams so that one bucket's
// barriers hide under the next bucket's transfers); call number `seq` uses slot seq & 1, and calls that share
// a slot are ordered by their stream, so the sequence numbers seen in a slot only grow.
constexpr int kSlots = 2;
struct Signals {                       // lives in each rank's device memory, written by its peers
  unsigned long long ready[kSlots][kMaxRanks];
  unsigned long long done[kSlots][kMaxRanks];
  unsigned int cta_counter[kSlots];    // local: CTAs of this rank that finished their stores
  unsigned int error;                  // local: a wait ran into the time limit
};

__device__ __forceinline__ unsigned long long now_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ bool wait_at_least(const unsigned long long* p, unsigned long long v) {
  const unsigned long long t0 = now_ns();
  while (ld_acquire_sys(p) < v) {
    if (now_ns() - t0 > kWaitLimitNs) return false;
  }
  return true;
}

// Done barrier of an allreduce kernel (Args has sig[], nranks, rank, seq, parallel_release).  Every thread fences
// its own peer stores, the rank's last CTA publishes "all my stores are out" to every peer, and a rank's kernel
// ends only when every peer has published.  parallel_release: one publishing thread per peer, so the release
// fences run side by side instead of one after the other (PRB_NVL_PARALLEL_RELEASE=0 restores the loop).
template <typename Args>
__device__ __forceinline__ void done_barrier(const Args& a, Signals* mine, int slot) {
  __threadfence_system();
  __syncthreads();
  if (a.parallel_release) {
    __shared__ unsigned int s_last;
    if (threadIdx.x == 0) {
      const unsigned int t = atomicAdd(&mine->cta_counter[slot], 1u);
      s_last = (t == gridDim.x - 1) ? 1u : 0u;
      if (s_last) mine->cta_counter[slot] = 0u;
    }
    __syncthreads();
    if (threadIdx.x < a.nranks && s_last) st_release_sys(&a.sig[threadIdx.x]->done[slot][a.rank], a.seq);
  } else if (threadIdx.x == 0) {
    const unsigned int t = atomicAdd(&mine->cta_counter[slot], 1u);
    if (t == gridDim.x - 1) {
      mine->cta_counter[slot] = 0u;
      __threadfence_system();
      for (int q = 0; q < a.nranks; ++q) st_release_sys(&a.sig[q]->done[slot][a.rank], a.seq);
    }
  }
  if (threadIdx.x < a.nranks && !wait_at_least(&mine->done[slot][threadIdx.x], a.seq)) mine->error = 1;
}

// Host side: allocate this rank's signal block, exchange IPC handles through the communicator and map the peers'.
// Collective.  *ok is the rank-local result; the caller agrees on it across ranks.
struct SignalSet {
  Signals* sig[kMaxRanks] = {nullptr};
  Signals* mine = nullptr;
  void* opened[kMaxRanks] = {nullptr};
  int nopened = 0;
};
int32_t signals_setup(const NvlCollectives& nc, int nranks, int rank, cudaStream_t stream, SignalSet* out, int* ok);
void signals_destroy(SignalSet* s);

}  // namespace prb
