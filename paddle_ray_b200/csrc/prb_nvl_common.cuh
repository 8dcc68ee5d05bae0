// Cross-GPU signalling shared by the NVLink peer-memory exchange (prb_nvlink.cu) and the NVLS multicast
// allreduce (prb_nvls.cu): per-rank signal blocks exported through CUDA IPC, sequence-number barriers.
//
// Round-2 design (round 1 had every CTA poll its peers with ld.acquire.sys and every thread issue a
// sequentially-consistent system fence, which held the calc stream's kernels back by ~100 us per step at 8 ranks):
//  * ONE warp of ONE CTA talks to the peers.  Ready barrier: warp 0 of CTA 0 publishes "my bucket is packed" to
//    every peer, polls the peers' flags with relaxed system-scope loads, issues ONE acquire fence and releases the
//    other CTAs through a gpu-scope flag in local memory.  Done barrier: each CTA issues one acq_rel system fence
//    after its stores (the CTA barrier makes it cumulative over all its threads) and takes a ticket; the CTA that
//    draws the last ticket publishes "my stores are out" and alone waits for the peers' publications.
//  * The kernels are launched into the REMOTE memory-synchronisation domain (cudaLaunchAttributeMemSyncDomain), so
//    their fences do not wait for -- and are not waited on by -- the optimizer's HBM traffic on the calc stream.
//  * A wait that runs into its time limit (PRB_NVL_TIMEOUT_MS, default 10 minutes: a peer is gone, not late) does
//    NOT fall through to the reduction: the rank marks the context dead, tells its peers (an abort value in their
//    flags, so they fail at once instead of waiting out their own limit), leaves the data untouched and raises a
//    host-visible error flag that the reducer checks on its next host call.
#pragma once
#include <cstring>

#include "prb_nvlink.h"

namespace prb {

constexpr int kMaxRanks = 8;
constexpr int kNvlThreads = 256;
constexpr unsigned long long kAbort = ~0ull;  // larger than any sequence number

// Two exchanges may be in flight at once (the reducer alternates two comm streams so that one bucket's
// barriers hide under the next bucket's transfers); call number `seq` uses slot seq & 1, and calls that share
// a slot are ordered by their stream, so the sequence numbers seen in a slot only grow.
constexpr int kSlots = 2;
struct Signals {                       // lives in each rank's device memory
  unsigned long long ready[kSlots][kMaxRanks];  // written by peers
  unsigned long long done[kSlots][kMaxRanks];   // written by peers
  unsigned long long go[kSlots];       // local: CTA 0 releases the other CTAs of this rank's kernel
  unsigned int cta_counter[kSlots];    // local: CTAs of this rank that finished their stores
  unsigned int error;                  // local: this context is dead (a wait timed out or a peer aborted)
};

// What a kernel needs to run the two barriers.
struct Bar {
  Signals* sig[kMaxRanks];             // every rank's signal block as mapped into this process
  int nranks, rank;
  unsigned long long seq;
  unsigned long long timeout_ns;
  unsigned int* host_error;            // mapped pinned host word (the reducer polls it without a sync) or nullptr
};

#ifdef __CUDACC__
__device__ __forceinline__ unsigned long long now_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ unsigned long long ld_relaxed_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_gpu(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_gpu(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void fence_acq_rel_sys() { asm volatile("fence.acq_rel.sys;" ::: "memory"); }

// Relaxed polling (no L1 invalidation per iteration); the caller fences once after all its flags have arrived.
// false: the peer aborted or the time limit passed.
__device__ __forceinline__ bool wait_at_least(const unsigned long long* p, unsigned long long v,
                                              unsigned long long timeout_ns) {
  const unsigned long long t0 = now_ns();
  unsigned long long x;
  unsigned int spins = 0;
  while ((x = ld_relaxed_sys(p)) < v) {
    if ((++spins & 63u) == 0 && now_ns() - t0 > timeout_ns) return false;
  }
  return x != kAbort;
}

// This rank gives up: mark the context dead (device and host flags) and put the abort value into every flag a peer
// might be waiting on, for both slots, so that the peers fail at once.
__device__ __forceinline__ void raise_comm_error(const Bar& b, Signals* mine) {
  mine->error = 1u;
  if (b.host_error != nullptr) {
    *reinterpret_cast<volatile unsigned int*>(b.host_error) = 1u;
    __threadfence_system();
  }
  for (int q = 0; q < b.nranks; ++q) {
    if (q == b.rank) continue;
    for (int s = 0; s < kSlots; ++s) {
      st_release_sys(&b.sig[q]->ready[s][b.rank], kAbort);
      st_release_sys(&b.sig[q]->done[s][b.rank], kAbort);
    }
  }
}

// Ready barrier.  Returns (uniformly in the CTA) false when the exchange must not run: the kernel then returns
// without touching data, counters or flags.  Every thread of every CTA must call it.
__device__ __forceinline__ bool ready_barrier(const Bar& b, int slot) {
  Signals* mine = b.sig[b.rank];
  __shared__ int s_go;
  if (blockIdx.x == 0 && threadIdx.x < 32) {
    const int q = threadIdx.x;
    bool ok = true;
    if (q < b.nranks) {
      if (*reinterpret_cast<volatile unsigned int*>(&mine->error) != 0u) {
        ok = false;  // an earlier exchange of this context failed: fail fast, stay dead
      } else {
        st_release_sys(&b.sig[q]->ready[slot][b.rank], b.seq);  // my packed bucket is complete (kernel order)
        ok = wait_at_least(&mine->ready[slot][q], b.seq, b.timeout_ns);
      }
    }
    ok = __all_sync(0xffffffffu, ok);
    if (q == 0) {
      if (ok) {
        fence_acq_rel_sys();  // the one acquire for all polled flags
        st_release_gpu(&mine->go[slot], b.seq);
      } else {
        raise_comm_error(b, mine);
        st_release_gpu(&mine->go[slot], kAbort);
      }
    }
  }
  if (threadIdx.x == 0) {
    unsigned long long v;
    while ((v = ld_acquire_gpu(&mine->go[slot])) < b.seq) __nanosleep(20);
    s_go = v != kAbort;
  }
  __syncthreads();
  return s_go != 0;
}

// Done barrier.  Every thread of every CTA calls it after its last peer store.  Returns true in the ONE CTA that
// drew the last ticket, after every peer has published that its stores into this rank's memory are out (or after
// the failure was raised): that CTA may then read what the peers wrote.
__device__ __forceinline__ bool done_barrier(const Bar& b, int slot) {
  Signals* mine = b.sig[b.rank];
  __shared__ unsigned int s_last;
  __syncthreads();  // all stores of this CTA are issued ...
  if (threadIdx.x == 0) {
    fence_acq_rel_sys();  // ... and performed system-wide before the ticket (cumulative over the CTA barrier)
    const unsigned int t = atomicAdd(&mine->cta_counter[slot], 1u);
    s_last = (t == gridDim.x - 1) ? 1u : 0u;
    if (s_last) {
      mine->cta_counter[slot] = 0u;
      fence_acq_rel_sys();  // acquire side of the ticket chain: every CTA's fence happened before this point
    }
  }
  __syncthreads();
  if (!s_last) return false;
  if (threadIdx.x < 32) {
    const int q = threadIdx.x;
    bool ok = true;
    if (q < b.nranks) {
      st_release_sys(&b.sig[q]->done[slot][b.rank], b.seq);
      ok = wait_at_least(&mine->done[slot][q], b.seq, b.timeout_ns);
    }
    ok = __all_sync(0xffffffffu, ok);
    if (q == 0) {
      if (ok) fence_acq_rel_sys();
      else raise_comm_error(b, mine);
    }
  }
  __syncthreads();
  return true;
}
#endif  // __CUDACC__

// Launch into the remote memory-synchronisation domain unless PRB_NVL_DOMAIN=default.
bool nvl_remote_domain();
template <typename Kernel, typename Args>
inline cudaError_t nvl_launch(Kernel kernel, int ctas, int threads, cudaStream_t stream, const Args& a) {
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeMemSyncDomain;
  attr[0].val.memSyncDomain = nvl_remote_domain() ? cudaLaunchMemSyncDomainRemote : cudaLaunchMemSyncDomainDefault;
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3((unsigned)ctas);
  cfg.blockDim = dim3((unsigned)threads);
  cfg.stream = stream;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, a);
}
unsigned long long nvl_timeout_ns();  // PRB_NVL_TIMEOUT_MS (default 600 000)

// Host side: allocate this rank's signal block, exchange IPC handles through the communicator and map the peers'.
// Collective: a local failure is folded into *ok and the rank keeps taking part; the caller agrees on *ok across ranks.
struct SignalSet {
  Signals* sig[kMaxRanks] = {nullptr};
  Signals* mine = nullptr;
  void* opened[kMaxRanks] = {nullptr};
  int nopened = 0;
  unsigned int* host_error = nullptr;      // pinned, mapped
  unsigned int* host_error_dev = nullptr;  // its device address
};
void signals_setup(const NvlCollectives& nc, int nranks, int rank, cudaStream_t stream, SignalSet* out, int* ok);
void signals_destroy(SignalSet* s);
void fill_bar(const SignalSet& s, int nranks, int rank, Bar* b);

}  // namespace prb
