// In-place bucket allreduce over NVLink peer memory (single node, one process per GPU).
//
// The reference hands every bucket to ncclAllReduce (ref: process_group_nccl.cc:841-865).  On an
// NVSwitch box every GPU can load and store every peer's memory at NVLink speed, so the reducer's
// persistent bucket arena (one per rank, exported through CUDA IPC once) allows a two-shot allreduce
// written as one kernel per rank:
//
//   ready barrier   every rank tells its peers "my packed bucket is complete" (the kernel runs after the
//                   pack in stream order) and waits for theirs;
//   reduce-scatter  rank r owns slice r of the bucket: it loads that slice from ALL ranks' arenas
//                   (256-/128-bit peer loads), adds the values in rank order 0..n-1 in fp32 -- a fixed
//                   order, so the result does not depend on who computes it -- rounds once to T ...
//   all-gather      ... and stores the reduced vector straight into every rank's arena (peer stores);
//   done barrier    the last CTA of each rank publishes "all my stores are out"; a rank's kernel ends
//                   only after every peer has published, so the next kernel in the stream (the optimizer)
//                   sees the complete reduced bucket.
//
// Per rank and bucket of S bytes: (n-1)/n * S over NVLink in each direction, no staging copies, no
// extra HBM pass.  Barriers are 64-bit sequence numbers written with st.release.sys into the peer's
// signal block and polled with ld.acquire.sys; every wait has a time limit (a missing peer turns into
// an error flag, never into a hung GPU).  Results are bit-equal across ranks by construction and equal
// to NCCL's for two ranks (one addition); for more ranks the fp32 accumulation with one final rounding
// is at least as accurate as pairwise sums in T.
//
// The default for multi-rank reducers (PRB_ALLREDUCE=nccl forces NCCL).  Needs CUDA IPC between the ranks;
// falls back to NCCL if any rank cannot set it up -- the decision is taken collectively.
#include "prb_nvl_common.cuh"

namespace prb {
namespace {

struct NvlArgs {
  char* arena[kMaxRanks];              // every rank's bucket arena as mapped into this process
  Signals* sig[kMaxRanks];             // every rank's signal block
  int nranks, rank;
  int parallel_release;        // see done_barrier
  size_t region_off;                   // byte offset of the bucket inside the arena
  long long nvec;                      // 8-element vectors in the bucket
  unsigned long long seq;
};

// Peer loads/stores: plain (coherent at system scope after the barriers), 8 elements per access.
template <typename T>
__device__ __forceinline__ void peer_load(const T* p, Pack<T>& o) {
  uint32_t* w = reinterpret_cast<uint32_t*>(o.v);
  if constexpr (sizeof(T) == 4) {
    asm volatile("ld.global.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7])
                 : "l"(p));
  } else {
    asm volatile("ld.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]) : "l"(p));
  }
}

template <typename T, int kRanks>
__global__ void __launch_bounds__(kNvlThreads, kNvlMinCtasPerSm) nvl_allreduce_kernel(const NvlArgs a) {
  Signals* mine = a.sig[a.rank];
  const int slot = (int)(a.seq & (kSlots - 1));
  // ---- ready barrier
  if (blockIdx.x == 0 && threadIdx.x < a.nranks) st_release_sys(&a.sig[threadIdx.x]->ready[slot][a.rank], a.seq);
  if (threadIdx.x < a.nranks && !wait_at_least(&mine->ready[slot][threadIdx.x], a.seq)) mine->error = 1;
  __syncthreads();

  // ---- reduce-scatter + all-gather of my slice
  const long long v0 = a.nvec * a.rank / a.nranks, v1 = a.nvec * (a.rank + 1) / a.nranks;
  const long long stride = (long long)gridDim.x * kNvlThreads;
  for (long long v = v0 + (long long)blockIdx.x * kNvlThreads + threadIdx.x; v < v1; v += stride) {
    const size_t off = a.region_off + (size_t)v * kVec * sizeof(T);
    // kRanks is the communicator size rounded up to 2, 4 or 8; surplus slots are skipped at run time.
    // Peers are loaded four at a time (all four loads in flight together) and added in rank order.
    constexpr int kBatch = kRanks > 4 ? 4 : kRanks;
    float acc[kVec];
#pragma unroll
    for (int b = 0; b < kRanks; b += kBatch) {
      Pack<T> in[kBatch];
#pragma unroll
      for (int q = 0; q < kBatch; ++q)
        if (b + q < a.nranks) peer_load<T>(reinterpret_cast<const T*>(a.arena[b + q] + off), in[q]);
#pragma unroll
      for (int q = 0; q < kBatch; ++q) {
        if (b + q < a.nranks) {
#pragma unroll
          for (int j = 0; j < kVec; ++j) {
            const float x = Conv<T>::to_f32(in[q].v[j]);
            acc[j] = (b + q == 0) ? x : __fadd_rn(acc[j], x);  // rank order: the same on every rank
          }
        }
      }
    }
    Pack<T> out;
#pragma unroll
    for (int j = 0; j < kVec; ++j) out.v[j] = Conv<T>::from_f32(acc[j]);
#pragma unroll
    for (int q = 0; q < kRanks; ++q)
      if (q < a.nranks) store_vec<T>(reinterpret_cast<T*>(a.arena[q] + off), out);
  }

  // ---- done barrier: my stores are out (last CTA publishes), then wait for every peer's
  done_barrier(a, mine, slot);
}

}  // namespace

struct NvlContext {
  NvlArgs args;
  void* opened[2 * kMaxRanks];
  int nopened = 0;
  Signals* my_signals = nullptr;
  unsigned long long seq = 0;
  int ctas = 0;
};

int32_t nvl_setup(const NvlCollectives& nc, int nranks, int rank, void* arena, size_t arena_bytes, cudaStream_t stream,
                  NvlContext** out) {
  *out = nullptr;
  if (nranks < 2 || nranks > kMaxRanks) return PRB_SUCCESS;  // nothing to set up: caller keeps NCCL
  (void)arena_bytes;
  NvlContext* c = new NvlContext();
  memset(&c->args, 0, sizeof(c->args));
  int ok = 1;
  cudaIpcMemHandle_t mine[2];
  if (cudaMalloc((void**)&c->my_signals, sizeof(Signals)) != cudaSuccess) ok = 0;
  if (ok && cudaMemset(c->my_signals, 0, sizeof(Signals)) != cudaSuccess) ok = 0;
  if (ok && cudaIpcGetMemHandle(&mine[0], arena) != cudaSuccess) ok = 0;
  if (ok && cudaIpcGetMemHandle(&mine[1], c->my_signals) != cudaSuccess) ok = 0;
  cudaGetLastError();
  // exchange {ok, handles} through the communicator (collective: every rank takes part even when it failed)
  struct Packet {
    int ok, pad;
    cudaIpcMemHandle_t h[2];
  } pk, all[kMaxRanks];
  memset(&pk, 0, sizeof(pk));
  pk.ok = ok;
  if (ok) {
    pk.h[0] = mine[0];
    pk.h[1] = mine[1];
  }
  void *dsend = nullptr, *drecv = nullptr;
  PRB_CUDA(cudaMalloc(&dsend, sizeof(Packet)));
  PRB_CUDA(cudaMalloc(&drecv, sizeof(Packet) * nranks));
  PRB_CUDA(cudaMemcpyAsync(dsend, &pk, sizeof(Packet), cudaMemcpyHostToDevice, stream));
  PRB_TRY(nc.all_gather_bytes(nc.comm, dsend, drecv, sizeof(Packet), stream));
  PRB_CUDA(cudaMemcpyAsync(all, drecv, sizeof(Packet) * nranks, cudaMemcpyDeviceToHost, stream));
  PRB_CUDA(cudaStreamSynchronize(stream));
  cudaFree(dsend);
  cudaFree(drecv);
  for (int q = 0; q < nranks; ++q) ok = ok && all[q].ok;
  if (ok) {
    for (int q = 0; q < nranks && ok; ++q) {
      if (q == rank) {
        c->args.arena[q] = reinterpret_cast<char*>(arena);
        c->args.sig[q] = c->my_signals;
        continue;
      }
      void *pa = nullptr, *ps = nullptr;
      if (cudaIpcOpenMemHandle(&pa, all[q].h[0], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) ok = 0;
      if (ok) c->opened[c->nopened++] = pa;
      if (ok && cudaIpcOpenMemHandle(&ps, all[q].h[1], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) ok = 0;
      if (ok) c->opened[c->nopened++] = ps;
      c->args.arena[q] = reinterpret_cast<char*>(pa);
      c->args.sig[q] = reinterpret_cast<Signals*>(ps);
    }
    cudaGetLastError();
  }
  // second agreement round: did everybody manage to map everybody?
  int agreed = 0;
  PRB_TRY(nc.all_min_int(nc.comm, ok, &agreed, stream));
  if (!agreed) {
    nvl_destroy(c);
    return PRB_SUCCESS;  // *out stays NULL: the caller keeps using NCCL
  }
  c->args.nranks = nranks;
  c->args.rank = rank;
  c->args.parallel_release = 1;
  if (const char* e = getenv("PRB_NVL_PARALLEL_RELEASE")) c->args.parallel_release = atoi(e) != 0;
  int per_sm = kNvlCtasPerSm;
  if (const char* e = getenv("PRB_NVL_CTAS_PER_SM")) per_sm = atoi(e) > 0 ? atoi(e) : per_sm;
  c->ctas = sm_count() * per_sm;
  *out = c;
  return PRB_SUCCESS;
}

int32_t signals_setup(const NvlCollectives& nc, int nranks, int rank, cudaStream_t stream, SignalSet* out, int* ok_out) {
  int ok = 1;
  cudaIpcMemHandle_t mine;
  memset(&mine, 0, sizeof(mine));
  if (cudaMalloc((void**)&out->mine, sizeof(Signals)) != cudaSuccess) ok = 0;
  if (ok && cudaMemset(out->mine, 0, sizeof(Signals)) != cudaSuccess) ok = 0;
  if (ok && cudaIpcGetMemHandle(&mine, out->mine) != cudaSuccess) ok = 0;
  cudaGetLastError();
  struct Packet {
    int ok, pad;
    cudaIpcMemHandle_t h;
  } pk, all[kMaxRanks];
  memset(&pk, 0, sizeof(pk));
  pk.ok = ok;
  pk.h = mine;
  void *dsend = nullptr, *drecv = nullptr;
  PRB_CUDA(cudaMalloc(&dsend, sizeof(Packet)));
  PRB_CUDA(cudaMalloc(&drecv, sizeof(Packet) * nranks));
  PRB_CUDA(cudaMemcpyAsync(dsend, &pk, sizeof(Packet), cudaMemcpyHostToDevice, stream));
  PRB_TRY(nc.all_gather_bytes(nc.comm, dsend, drecv, sizeof(Packet), stream));
  PRB_CUDA(cudaMemcpyAsync(all, drecv, sizeof(Packet) * nranks, cudaMemcpyDeviceToHost, stream));
  PRB_CUDA(cudaStreamSynchronize(stream));
  cudaFree(dsend);
  cudaFree(drecv);
  for (int q = 0; q < nranks; ++q) ok = ok && all[q].ok;
  for (int q = 0; q < nranks && ok; ++q) {
    if (q == rank) {
      out->sig[q] = out->mine;
      continue;
    }
    void* ps = nullptr;
    if (cudaIpcOpenMemHandle(&ps, all[q].h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) ok = 0;
    if (ok) out->opened[out->nopened++] = ps;
    out->sig[q] = reinterpret_cast<Signals*>(ps);
  }
  cudaGetLastError();
  *ok_out = ok;
  return PRB_SUCCESS;
}

void signals_destroy(SignalSet* s) {
  for (int i = 0; i < s->nopened; ++i) cudaIpcCloseMemHandle(s->opened[i]);
  s->nopened = 0;
  if (s->mine) cudaFree(s->mine);
  s->mine = nullptr;
}

int nvl_next_slot(const NvlContext* c) { return c ? (int)((c->seq + 1) & (kSlots - 1)) : 0; }

int32_t nvl_allreduce(NvlContext* c, size_t region_off_bytes, int64_t numel, int32_t dtype, cudaStream_t stream) {
  PRB_ENFORCE(c != nullptr, PRB_INVALID_ARGUMENT, "nvlink allreduce: no context");
  PRB_ENFORCE(numel % kVec == 0 && region_off_bytes % 32 == 0, PRB_INVALID_ARGUMENT,
              "nvlink allreduce: the bucket must be a whole number of 8-element vectors at a 32-byte offset");
  NvlArgs a = c->args;
  a.region_off = region_off_bytes;
  a.nvec = numel / kVec;
  a.seq = ++c->seq;
  const long long per_rank = (a.nvec + a.nranks - 1) / a.nranks;
  int ctas = (int)((per_rank + kNvlThreads - 1) / kNvlThreads);
  if (ctas > c->ctas) ctas = c->ctas;  // a few CTAs per SM, each looping over its share of the slice
  if (ctas < 1) ctas = 1;
#define PRB_NVL(T_)                                                                              \
  do {                                                                                           \
    if (a.nranks <= 2) nvl_allreduce_kernel<T_, 2><<<ctas, kNvlThreads, 0, stream>>>(a);         \
    else if (a.nranks <= 4) nvl_allreduce_kernel<T_, 4><<<ctas, kNvlThreads, 0, stream>>>(a);    \
    else nvl_allreduce_kernel<T_, 8><<<ctas, kNvlThreads, 0, stream>>>(a);                       \
  } while (0)
  switch (dtype) {
    case PRB_FLOAT32: PRB_NVL(float); break;
    case PRB_BFLOAT16: PRB_NVL(__nv_bfloat16); break;
    case PRB_FLOAT16: PRB_NVL(__half); break;
    default: PRB_FAIL(PRB_UNIMPLEMENTED, "nvlink allreduce: dtype %s is not built", dtype_name(dtype));
  }
#undef PRB_NVL
  PRB_LAUNCH_CHECK();
  return PRB_SUCCESS;
}

int32_t nvl_error_flag(NvlContext* c, int* out) {
  unsigned int e = 0;
  PRB_CUDA(cudaMemcpy(&e, &c->my_signals->error, sizeof(e), cudaMemcpyDeviceToHost));
  *out = (int)e;
  return PRB_SUCCESS;
}

void nvl_destroy(NvlContext* c) {
  if (!c) return;
  for (int i = 0; i < c->nopened; ++i) cudaIpcCloseMemHandle(c->opened[i]);
  if (c->my_signals) cudaFree(c->my_signals);
  delete c;
}

}  // namespace prb
