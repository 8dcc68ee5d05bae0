// In-place bucket exchange (allreduce) over NVLink peer memory (single node, one process per GPU).
//
// The reference hands every bucket to ncclAllReduce (ref: process_group_nccl.cc:841-865).  On an
// NVSwitch box every GPU can load and store every peer's memory at NVLink speed, so the reducer's
// persistent bucket arena (one per rank, exported through CUDA IPC once) allows a two-shot allreduce
// written as one kernel per rank:
//
//   ready barrier   every rank tells its peers "my packed bucket is complete" (the kernel runs after the
//                   pack in stream order) and waits for theirs (one warp does the talking, prb_nvl_common.cuh);
//   reduce-scatter  rank r owns a slice of the bucket: it loads that slice from ALL ranks' arenas
//                   (256-/128-bit loads, ~8-16 of them in flight per thread), adds the values in rank order
//                   0..n-1 in fp32 -- a fixed order, so the result does not depend on who computes it --
//                   rounds once to T ...
//   all-gather      ... and stores the reduced vector straight into every rank's arena (peer stores);
//   done barrier    the last CTA of each rank publishes "all my stores are out"; a rank's kernel ends
//                   only after every peer has published, so the next kernel in the stream (the optimizer)
//                   sees the complete reduced bucket.
//
// Two things ride on the reduced value while the slice owner holds it in registers:
//   * the 1/nranks scale of gradients that were produced in the arena (each rank's value goes through
//     T(scale * (x + 0)) before the add: ScaleFunctor's rounding, ref scale_kernel.cu:24-41), so that mode needs
//     no pass of its own for the scale;
//   * producer-side pass 1 (chunk mode): check_finite_and_unscale's inf/nan test and the squared-norm partial of
//     every 8192-element chunk (ChunkEpilogue, prb_grad_pass.cuh), stored into EVERY rank's sidecar with the same
//     peer stores -- the gradient is not read again before the update pass, and the bits equal those of a
//     separate pass over the reduced bucket.
//
// Per rank and bucket of S bytes: (n-1)/n * S over NVLink in each direction, no staging copies, no
// extra HBM pass.  Results are bit-equal across ranks by construction and equal to NCCL's for two ranks (one
// addition); for more ranks the fp32 accumulation with one final rounding is at least as accurate as pairwise
// sums in T.
//
// The default for multi-rank reducers (PRB_ALLREDUCE=nccl forces NCCL).  Needs CUDA IPC between the ranks;
// falls back to NCCL if any rank cannot set it up -- the decision is taken collectively.
#include "prb_grad_pass.cuh"
#include "prb_nvl_common.cuh"

namespace prb {
namespace {

static_assert(kNvlThreads == kGradThreads, "the exchange kernel produces norm partials with the grad-pass CTA shape");

struct NvlArgs {
  Bar bar;
  char* arena[kMaxRanks];  // every rank's bucket arena as mapped into this process
  // flat mode
  size_t region_off;       // byte offset of the bucket inside the arena
  long long nvec;          // 8-element vectors in the bucket
  // chunk mode
  const GradDesc* descs;
  const Chunk* chunks;
  int nchunks;
  // scale fold
  float scale, bias;
  // epilogue
  const float* loss_scale;
  bool* found_local;
  size_t partials_off, flags_off;
  // sharded step
  int local_only;          // reduce-scatter: the reduced value stays with its owner
  const bool* skip;        // gather: nothing was updated (found_inf)
  int nseg;                // gather: ownership segments of the chunk list
  int seg_first[kMaxShardSegments + 1];
};

// Loads in flight per thread.  NVLink latency is ~2 us: a thread that waits for ONE remote vector at a time
// moves ~230 GB/s per GPU at two ranks (round-2 measurement), so the number of lane vectors a thread handles per
// round grows as the rank count shrinks: ~8 loads of 32 bytes (fp32) or ~16 loads of 16 bytes (16-bit types).
template <typename T, int kRanks>
struct Flight {
  static constexpr int kTarget = sizeof(T) == 4 ? 8 : 16;
  static constexpr int kVecs = (kTarget / kRanks) < 1 ? 1 : ((kTarget / kRanks) > kGradUnroll ? kGradUnroll : (kTarget / kRanks));
};

// Loads of the vector at byte offset `off` from every rank's arena (issued, not yet used).
// kRanks is the communicator size rounded up to 2, 4 or 8; surplus slots are skipped at run time.
template <typename T, int kRanks>
__device__ __forceinline__ void load_all(const NvlArgs& a, size_t off, Pack<T> (&in)[kRanks]) {
#pragma unroll
  for (int q = 0; q < kRanks; ++q)
    if (q < a.bar.nranks) load_vec<T>(reinterpret_cast<const T*>(a.arena[q] + off), in[q]);
}

// Sum over ranks in rank order (a fixed order: the same on every rank), rounded once to T.
template <typename T, int kRanks, bool kFold>
__device__ __forceinline__ void combine(const NvlArgs& a, Pack<T> (&in)[kRanks], Pack<T>& out) {
  float acc[kVec];
#pragma unroll
  for (int q = 0; q < kRanks; ++q) {
    if (q < a.bar.nranks) {
      if (kFold) pack_scale_vec<T>(in[q], a.scale, a.bias);
#pragma unroll
      for (int j = 0; j < kVec; ++j) {
        const float x = Conv<T>::to_f32(in[q].v[j]);
        acc[j] = (q == 0) ? x : __fadd_rn(acc[j], x);
      }
    }
  }
#pragma unroll
  for (int j = 0; j < kVec; ++j) out.v[j] = Conv<T>::from_f32(acc[j]);
}

// One element straight from memory (never from a stale L1 line): the ragged tail path.
template <typename T>
__device__ __forceinline__ T load_elem_cv(const char* p) {
  if constexpr (sizeof(T) == 4) {
    uint32_t w = *reinterpret_cast<const volatile uint32_t*>(p);
    return *reinterpret_cast<T*>(&w);
  } else {
    uint16_t w = *reinterpret_cast<const volatile uint16_t*>(p);
    return *reinterpret_cast<T*>(&w);
  }
}

template <typename T, int kRanks>
__device__ __forceinline__ void store_all(const NvlArgs& a, size_t off, const Pack<T>& v) {
  if (a.local_only) {
    store_vec<T>(reinterpret_cast<T*>(a.arena[a.bar.rank] + off), v);
    return;
  }
#pragma unroll
  for (int q = 0; q < kRanks; ++q)
    if (q < a.bar.nranks) store_vec<T>(reinterpret_cast<T*>(a.arena[q] + off), v);
}

// ---- flat mode: rank r owns the r-th n-th of the region
template <typename T, int kRanks, bool kFold>
__global__ void __launch_bounds__(kNvlThreads, 2) nvl_exchange_kernel(const NvlArgs a) {
  const int slot = (int)(a.bar.seq & (kSlots - 1));
  if (!ready_barrier(a.bar, slot)) return;
  const long long v0 = a.nvec * a.bar.rank / a.bar.nranks, v1 = a.nvec * (a.bar.rank + 1) / a.bar.nranks;
  const long long stride = (long long)gridDim.x * kNvlThreads;
  constexpr int kV = Flight<T, kRanks>::kVecs;
  for (long long v = v0 + (long long)blockIdx.x * kNvlThreads + threadIdx.x; v < v1; v += stride * kV) {
    Pack<T> in[kV][kRanks];
#pragma unroll
    for (int u = 0; u < kV; ++u)
      if (v + u * stride < v1) load_all<T, kRanks>(a, a.region_off + (size_t)(v + u * stride) * kVec * sizeof(T), in[u]);
#pragma unroll
    for (int u = 0; u < kV; ++u) {
      if (v + u * stride < v1) {
        Pack<T> out;
        combine<T, kRanks, kFold>(a, in[u], out);
        store_all<T, kRanks>(a, a.region_off + (size_t)(v + u * stride) * kVec * sizeof(T), out);
      }
    }
  }
  done_barrier(a.bar, slot);
}

// ---- chunk mode: rank r owns the r-th n-th of the (tensor, 8192-element chunk) list; one CTA per chunk at a
// time, with exactly the thread -> element assignment of grad_pass_kernel, so the partials are the same bits
template <typename T, int kRanks, bool kFold, bool kNorm, bool kMp>
__global__ void __launch_bounds__(kNvlThreads, 2) nvl_exchange_chunks_kernel(const NvlArgs a) {
  __shared__ float red[kNvlThreads / 32];
  const int slot = (int)(a.bar.seq & (kSlots - 1));
  if (!ready_barrier(a.bar, slot)) return;
  const int n = a.bar.nranks, r = a.bar.rank;
  const int c0 = (int)((long long)a.nchunks * r / n), c1 = (int)((long long)a.nchunks * (r + 1) / n);
  const char* my_arena = a.arena[r];
  for (int c = c0 + (int)blockIdx.x; c < c1; c += (int)gridDim.x) {
    const int4 craw = __ldg(reinterpret_cast<const int4*>(a.chunks) + c);
    const int cn = craw.y;
    const int64_t offset = (int64_t)(((uint64_t)(uint32_t)craw.w << 32) | (uint32_t)craw.z);
    const GradDesc* __restrict__ d = a.descs + craw.x;
    const size_t off = (size_t)(reinterpret_cast<const char*>(d->x) - my_arena) + (size_t)offset * sizeof(T);
    ChunkEpilogue<T, kNorm, kMp> ep;
    ep.begin(a.loss_scale);
    const int nvec = cn / kVec;
    constexpr int kV = Flight<T, kRanks>::kVecs;
#pragma unroll 1
    for (int u0 = 0; u0 < kGradUnroll; u0 += kV) {
      Pack<T> in[kV][kRanks];
#pragma unroll
      for (int u = 0; u < kV; ++u) {
        const int i = (u0 + u) * kNvlThreads + threadIdx.x;
        if (i < nvec) load_all<T, kRanks>(a, off + (size_t)i * kVec * sizeof(T), in[u]);
      }
#pragma unroll
      for (int u = 0; u < kV; ++u) {  // reduce, store and feed the epilogue in the order u = 0, 1, 2, 3
        const int i = (u0 + u) * kNvlThreads + threadIdx.x;
        if (i < nvec) {
          Pack<T> v;
          combine<T, kRanks, kFold>(a, in[u], v);
          store_all<T, kRanks>(a, off + (size_t)i * kVec * sizeof(T), v);
          ep.vec(v);
        }
      }
    }
    const int e = nvec * kVec + threadIdx.x;
    if (e < cn) {  // ragged tail of the tensor's last chunk (the slot's zero padding is never touched)
      const size_t o = off + (size_t)e * sizeof(T);
      float acc = 0.0f;
      for (int q = 0; q < n; ++q) {
        T x = load_elem_cv<T>(a.arena[q] + o);
        if (kFold) x = pack_scale_elem<T>(x, a.scale, a.bias);
        acc = (q == 0) ? Conv<T>::to_f32(x) : __fadd_rn(acc, Conv<T>::to_f32(x));
      }
      const T out = Conv<T>::from_f32(acc);
      if (a.local_only) *reinterpret_cast<T*>(a.arena[r] + o) = out;
      else
        for (int q = 0; q < n; ++q) *reinterpret_cast<T*>(a.arena[q] + o) = out;
      ep.elem(out);
    }
    bool any_bad;
    const float s = ep.finish(red, &any_bad);
    if (threadIdx.x == 0) {
      if (kNorm) {
        const size_t po = a.partials_off + ((size_t)d->partial_base + (size_t)(offset / kGradChunk)) * sizeof(float);
        for (int q = 0; q < n; ++q) *reinterpret_cast<float*>(a.arena[q] + po) = s;
      }
      if (any_bad)  // only "set" is ever written, so concurrent stores are benign (ref amp_kernel.cu:79-81)
        for (int q = 0; q < n; ++q) *reinterpret_cast<unsigned int*>(a.arena[q] + a.flags_off) = 1u;
    }
  }
  const bool last = done_barrier(a.bar, slot);
  // every peer's stores are out: the flag in MY sidecar is the OR over all slices of all ranks so far in this step
  if (last && threadIdx.x == 0 && a.found_local != nullptr &&
      *reinterpret_cast<const volatile unsigned int*>(my_arena + a.flags_off) != 0u)
    *a.found_local = true;
}

// ---- sharded step, second half: pull the chunks the peers own from their arenas into my parameters.
// CTA b serves peer (rank + 1 + b mod (n-1)) mod n, so every rank reads from all its peers at once and no rank's
// links serve more than their share.  E = element size: the copy does not care about the type.
template <typename T>
__global__ void __launch_bounds__(kNvlThreads, 2) nvl_gather_chunks_kernel(const NvlArgs a) {
  const int slot = (int)(a.bar.seq & (kSlots - 1));
  if (!ready_barrier(a.bar, slot)) return;  // every owner's update pass has finished (kernel order + release)
  const int n = a.bar.nranks, r = a.bar.rank;
  const bool skipped = a.skip != nullptr && *a.skip;
  if (!skipped) {
    const int peers = n - 1;
    const int pi = (int)blockIdx.x % peers;
    const int q = (r + 1 + pi) % n;
    const int first = (int)blockIdx.x / peers, step = ((int)gridDim.x - pi + peers - 1) / peers;
    const char* my_arena = a.arena[r];
    // peer q owns the q-th n-th of every segment; the CTAs that serve q walk its chunks as one list
    int seg = 0, seg_base = 0;  // chunks of q in the segments before `seg`
    int s0 = a.seg_first[0] + (int)((long long)(a.seg_first[1] - a.seg_first[0]) * q / n);
    int s1 = a.seg_first[0] + (int)((long long)(a.seg_first[1] - a.seg_first[0]) * (q + 1) / n);
    for (int fi = first;; fi += step) {
      while (seg < a.nseg && fi - seg_base >= s1 - s0) {
        seg_base += s1 - s0;
        ++seg;
        if (seg < a.nseg) {
          const int f = a.seg_first[seg], len = a.seg_first[seg + 1] - f;
          s0 = f + (int)((long long)len * q / n);
          s1 = f + (int)((long long)len * (q + 1) / n);
        }
      }
      if (seg >= a.nseg) break;
      const int c = s0 + (fi - seg_base);
      const int4 craw = __ldg(reinterpret_cast<const int4*>(a.chunks) + c);
      const int cn = craw.y;
      const int64_t offset = (int64_t)(((uint64_t)(uint32_t)craw.w << 32) | (uint32_t)craw.z);
      const GradDesc* __restrict__ d = a.descs + craw.x;
      const size_t off = (size_t)(reinterpret_cast<const char*>(d->x) - my_arena) + (size_t)offset * sizeof(T);
      const char* src = a.arena[q] + off;
      T* dst = reinterpret_cast<T*>(d->out) + offset;
      if (d->aligned) {
        const int nvec = cn / kVec;
        Pack<T> v[kGradUnroll];
#pragma unroll
        for (int u = 0; u < kGradUnroll; ++u) {
          const int i = u * kNvlThreads + threadIdx.x;
          if (i < nvec) load_vec<T>(reinterpret_cast<const T*>(src) + (size_t)i * kVec, v[u]);
        }
#pragma unroll
        for (int u = 0; u < kGradUnroll; ++u) {
          const int i = u * kNvlThreads + threadIdx.x;
          if (i < nvec) store_vec<T>(dst + (size_t)i * kVec, v[u]);
        }
        const int e = nvec * kVec + threadIdx.x;
        if (e < cn) dst[e] = load_elem_cv<T>(src + (size_t)e * sizeof(T));
      } else {  // a parameter that is not 32-/16-byte aligned: element-wise
        for (int e = threadIdx.x; e < cn; e += kNvlThreads) dst[e] = load_elem_cv<T>(src + (size_t)e * sizeof(T));
      }
    }
  }
  done_barrier(a.bar, slot);  // no peer overwrites its slots (the next pack) before everybody has read them
}

}  // namespace

struct NvlContext {
  NvlArgs args;
  SignalSet sigs;
  void* opened[kMaxRanks];
  int nopened = 0;
  unsigned long long seq = 0;
  int ctas = 0;
};

bool nvl_remote_domain() {
  static const bool v = [] {
    const char* e = getenv("PRB_NVL_DOMAIN");
    return !(e != nullptr && strcmp(e, "default") == 0);
  }();
  return v;
}

unsigned long long nvl_timeout_ns() {
  static const unsigned long long v = [] {
    const char* e = getenv("PRB_NVL_TIMEOUT_MS");
    const long long ms = e ? atoll(e) : 600000ll;
    return (unsigned long long)(ms > 0 ? ms : 600000ll) * 1000000ull;
  }();
  return v;
}

// Exchange one fixed-size packet per rank through the communicator.  Takes part in the collective whatever the
// local state; returns false if anything local or the collective itself failed.
static bool gather_packets(const NvlCollectives& nc, const void* mine, void* all, size_t bytes, int nranks, cudaStream_t stream) {
  void *dsend = nullptr, *drecv = nullptr;
  bool ok = cudaMalloc(&dsend, bytes) == cudaSuccess && cudaMalloc(&drecv, bytes * nranks) == cudaSuccess;
  if (ok) ok = cudaMemcpyAsync(dsend, mine, bytes, cudaMemcpyHostToDevice, stream) == cudaSuccess;
  // a rank that could not even allocate still has to enter the collective: it sends from a spare buffer
  static char* spare = nullptr;
  if (!ok && spare == nullptr) cudaMalloc((void**)&spare, 4096 * (size_t)(kMaxRanks + 1));
  const bool coll_ok = nc.all_gather_bytes(nc.comm, ok ? dsend : spare, ok ? drecv : spare + 4096, bytes, stream) == PRB_SUCCESS;
  if (ok && coll_ok) ok = cudaMemcpyAsync(all, drecv, bytes * nranks, cudaMemcpyDeviceToHost, stream) == cudaSuccess;
  if (cudaStreamSynchronize(stream) != cudaSuccess) ok = false;
  if (dsend) cudaFree(dsend);
  if (drecv) cudaFree(drecv);
  cudaGetLastError();
  return ok && coll_ok;
}

void signals_setup(const NvlCollectives& nc, int nranks, int rank, cudaStream_t stream, SignalSet* out, int* ok_out) {
  int ok = 1;
  cudaIpcMemHandle_t mine;
  memset(&mine, 0, sizeof(mine));
  if (cudaMalloc((void**)&out->mine, sizeof(Signals)) != cudaSuccess) ok = 0;
  if (ok && cudaMemset(out->mine, 0, sizeof(Signals)) != cudaSuccess) ok = 0;
  if (ok && cudaIpcGetMemHandle(&mine, out->mine) != cudaSuccess) ok = 0;
  if (ok && cudaHostAlloc((void**)&out->host_error, sizeof(unsigned int), cudaHostAllocMapped) == cudaSuccess) {
    *out->host_error = 0u;
    if (cudaHostGetDevicePointer((void**)&out->host_error_dev, out->host_error, 0) != cudaSuccess) out->host_error_dev = nullptr;
  }
  cudaGetLastError();
  struct Packet {
    int ok, pad;
    cudaIpcMemHandle_t h;
  } pk, all[kMaxRanks];
  memset(&pk, 0, sizeof(pk));
  memset(all, 0, sizeof(all));
  pk.ok = ok;
  pk.h = mine;
  if (!gather_packets(nc, &pk, all, sizeof(Packet), nranks, stream)) ok = 0;
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
}

void signals_destroy(SignalSet* s) {
  for (int i = 0; i < s->nopened; ++i) cudaIpcCloseMemHandle(s->opened[i]);
  s->nopened = 0;
  if (s->mine) cudaFree(s->mine);
  s->mine = nullptr;
  if (s->host_error) cudaFreeHost(s->host_error);
  s->host_error = nullptr;
}

void fill_bar(const SignalSet& s, int nranks, int rank, Bar* b) {
  for (int q = 0; q < kMaxRanks; ++q) b->sig[q] = s.sig[q];
  b->nranks = nranks;
  b->rank = rank;
  b->seq = 0;
  b->timeout_ns = nvl_timeout_ns();
  b->host_error = s.host_error_dev;
}

// Grid of the exchange kernels.  The kernel is NVLink-bound: 64 CTAs move as many bytes per second as one CTA on
// each of the 148 SMs (8 ranks, 102 MB: 366 vs 374 us), but a grid that sits on EVERY SM slows the kernels of the
// calc stream it is supposed to hide under -- with a 730 us backward at 8 ranks the exposed communication was 77 us
// with 148 CTAs and 9 us with 64, 32 or 16 (profiles/r2_exchange_sweep_n8.log; NCCL, 580 us of exchanges: 12 us).
// PRB_NVL_CTAS overrides; prb_reducer_set_exchange_ctas changes it at run time.
// The two kernels of a sharded step run when the backward is over (the reduce-scatter starts after the LAST bucket,
// the gather after the update): nothing on the calc stream to slow down, and remote loads need bytes in flight --
// two ranks, 16-byte vectors, 4 per thread: 449 GB/s per direction with 64 CTAs, 614 with 148, 619 with 296
// (tools/probe/p2p_probe.cu, profiles/r2_p2p_probe.log).  PRB_NVL_SHARD_CTAS overrides.
static int shard_ctas() {
  int n = 2 * sm_count();
  if (const char* e = getenv("PRB_NVL_SHARD_CTAS")) n = atoi(e) > 0 ? atoi(e) : n;
  return n;
}

static int default_ctas() {
  int n = sm_count() < 64 ? sm_count() : 64;
  if (const char* e = getenv("PRB_NVL_CTAS")) n = atoi(e) > 0 ? atoi(e) : n;
  return n;
}

int32_t nvl_setup(const NvlCollectives& nc, int nranks, int rank, void* arena, size_t arena_bytes, cudaStream_t stream,
                  NvlContext** out) {
  *out = nullptr;
  if (nranks < 2 || nranks > kMaxRanks) return PRB_SUCCESS;  // nothing to set up: caller keeps NCCL
  (void)arena_bytes;
  NvlContext* c = new NvlContext();
  memset(&c->args, 0, sizeof(c->args));
  // Every step below is collective: a local failure is folded into `ok` and the rank keeps taking part in every
  // exchange and in the final vote, so that no peer is left waiting in a collective this rank has abandoned.
  int ok = 1;
  signals_setup(nc, nranks, rank, stream, &c->sigs, &ok);
  cudaIpcMemHandle_t mine;
  memset(&mine, 0, sizeof(mine));
  if (ok && cudaIpcGetMemHandle(&mine, arena) != cudaSuccess) ok = 0;
  cudaGetLastError();
  struct Packet {
    int ok, pad;
    cudaIpcMemHandle_t h;
  } pk, all[kMaxRanks];
  memset(&pk, 0, sizeof(pk));
  memset(all, 0, sizeof(all));
  pk.ok = ok;
  pk.h = mine;
  if (!gather_packets(nc, &pk, all, sizeof(Packet), nranks, stream)) ok = 0;
  for (int q = 0; q < nranks; ++q) ok = ok && all[q].ok;
  for (int q = 0; q < nranks && ok; ++q) {
    if (q == rank) {
      c->args.arena[q] = reinterpret_cast<char*>(arena);
      continue;
    }
    void* pa = nullptr;
    if (cudaIpcOpenMemHandle(&pa, all[q].h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) ok = 0;
    if (ok) c->opened[c->nopened++] = pa;
    c->args.arena[q] = reinterpret_cast<char*>(pa);
  }
  cudaGetLastError();
  // final agreement: did everybody manage to map everybody?
  int agreed = 0;
  const int32_t vote = nc.all_min_int(nc.comm, ok, &agreed, stream);
  if (vote != PRB_SUCCESS || !agreed) {
    nvl_destroy(c);
    return vote;  // *out stays NULL: the caller keeps using NCCL (or reports the broken communicator)
  }
  fill_bar(c->sigs, nranks, rank, &c->args.bar);
  c->ctas = default_ctas();
  *out = c;
  return PRB_SUCCESS;
}

int nvl_next_slot(const NvlContext* c) { return c ? (int)((c->seq + 1) & (kSlots - 1)) : 0; }

void nvl_set_max_ctas(NvlContext* c, int max_ctas) {
  if (c) c->ctas = max_ctas > 0 ? max_ctas : default_ctas();
}

namespace {

template <typename T, int kRanks>
cudaError_t launch_exchange(const NvlArgs& a, const NvlWork& w, int ctas, cudaStream_t stream) {
  if (w.table == nullptr) {
    if (w.fold_scale) return nvl_launch(nvl_exchange_kernel<T, kRanks, true>, ctas, kNvlThreads, stream, a);
    return nvl_launch(nvl_exchange_kernel<T, kRanks, false>, ctas, kNvlThreads, stream, a);
  }
#define PRB_NVC(F)                                                                                                 \
  do {                                                                                                             \
    if (!w.norm) return nvl_launch(nvl_exchange_chunks_kernel<T, kRanks, F, false, false>, ctas, kNvlThreads, stream, a); \
    if (!w.mp) return nvl_launch(nvl_exchange_chunks_kernel<T, kRanks, F, true, false>, ctas, kNvlThreads, stream, a);    \
    return nvl_launch(nvl_exchange_chunks_kernel<T, kRanks, F, true, true>, ctas, kNvlThreads, stream, a);         \
  } while (0)
  if (w.fold_scale) PRB_NVC(true);
  PRB_NVC(false);
#undef PRB_NVC
}

template <typename T>
cudaError_t launch_exchange_ranks(const NvlArgs& a, const NvlWork& w, int ctas, cudaStream_t stream) {
  if (a.bar.nranks <= 2) return launch_exchange<T, 2>(a, w, ctas, stream);
  if (a.bar.nranks <= 4) return launch_exchange<T, 4>(a, w, ctas, stream);
  return launch_exchange<T, 8>(a, w, ctas, stream);
}

}  // namespace

int32_t nvl_exchange(NvlContext* c, const NvlWork& w, cudaStream_t stream) {
  PRB_ENFORCE(c != nullptr, PRB_INVALID_ARGUMENT, "nvlink exchange: no context");
  PRB_ENFORCE(w.numel % kVec == 0 && w.region_off % 32 == 0, PRB_INVALID_ARGUMENT,
              "nvlink exchange: the bucket must be a whole number of 8-element vectors at a 32-byte offset");
  NvlArgs a = c->args;
  a.region_off = w.region_off;
  a.nvec = w.numel / kVec;
  a.bar.seq = ++c->seq;
  a.scale = w.scale;
  a.bias = 0.0f;
  a.local_only = w.scatter_only ? 1 : 0;
  int ctas;
  if (w.table != nullptr) {
    PRB_ENFORCE(w.loss_scale == nullptr || w.found_local != nullptr, PRB_INVALID_ARGUMENT,
                "nvlink exchange: the finite check needs a found_inf flag");
    a.descs = reinterpret_cast<const GradDesc*>(w.table->descs);
    a.chunks = w.table->chunks;
    a.nchunks = w.table->nchunks;
    a.loss_scale = w.loss_scale;
    a.found_local = w.found_local;
    a.partials_off = w.partials_off;
    a.flags_off = w.flags_off;
    ctas = (a.nchunks + a.bar.nranks - 1) / a.bar.nranks;  // one CTA per chunk of my share, capped below
  } else {
    const long long per_rank = (a.nvec + a.bar.nranks - 1) / a.bar.nranks;
    ctas = (int)((per_rank + kNvlThreads - 1) / kNvlThreads);
  }
  const int cap = (w.scatter_only && !w.overlap_grid) ? shard_ctas() : c->ctas;
  if (ctas > cap) ctas = cap;  // each CTA loops over its share of the slice
  if (ctas < 1) ctas = 1;
  KernelTimer kt(KK_EXCHANGE, stream);
  cudaError_t e;
  switch (w.dtype) {
    case PRB_FLOAT32: e = launch_exchange_ranks<float>(a, w, ctas, stream); break;
    case PRB_BFLOAT16: e = launch_exchange_ranks<__nv_bfloat16>(a, w, ctas, stream); break;
    case PRB_FLOAT16: e = launch_exchange_ranks<__half>(a, w, ctas, stream); break;
    default: PRB_FAIL(PRB_UNIMPLEMENTED, "nvlink exchange: dtype %s is not built", dtype_name(w.dtype));
  }
  PRB_CUDA(e);
  PRB_LAUNCH_CHECK();
  return PRB_SUCCESS;
}

int32_t nvl_gather_params(NvlContext* c, const NvlGather& w, cudaStream_t stream) {
  PRB_ENFORCE(c != nullptr && w.table != nullptr, PRB_INVALID_ARGUMENT, "nvlink gather: no context or no table");
  NvlArgs a = c->args;
  a.bar.seq = ++c->seq;
  a.descs = reinterpret_cast<const GradDesc*>(w.table->descs);
  a.chunks = w.table->chunks;
  a.nchunks = w.table->nchunks;
  a.skip = w.skip;
  PRB_ENFORCE(w.nseg >= 1 && w.nseg <= kMaxShardSegments, PRB_INVALID_ARGUMENT, "nvlink gather: bad segment count %d", w.nseg);
  a.nseg = w.nseg;
  for (int s = 0; s <= w.nseg; ++s) a.seg_first[s] = w.seg_first[s];
  const int peers = a.bar.nranks - 1;
  int ctas = (a.nchunks + a.bar.nranks - 1) / a.bar.nranks * peers;  // one CTA per chunk I pull, capped below
  if (ctas > shard_ctas()) ctas = shard_ctas() / peers * peers;
  if (ctas < peers) ctas = peers;
  KernelTimer kt(KK_EXCHANGE, stream);
  cudaError_t e;
  switch (dtype_size(w.dtype)) {
    case 4: e = nvl_launch(nvl_gather_chunks_kernel<float>, ctas, kNvlThreads, stream, a); break;
    case 2: e = nvl_launch(nvl_gather_chunks_kernel<__nv_bfloat16>, ctas, kNvlThreads, stream, a); break;
    default: PRB_FAIL(PRB_UNIMPLEMENTED, "nvlink gather: dtype %s is not built", dtype_name(w.dtype));
  }
  PRB_CUDA(e);
  PRB_LAUNCH_CHECK();
  return PRB_SUCCESS;
}

int nvl_error_flag(const NvlContext* c) {
  return (c != nullptr && c->sigs.host_error != nullptr) ? (int)*reinterpret_cast<volatile unsigned int*>(c->sigs.host_error) : 0;
}

void nvl_destroy(NvlContext* c) {
  if (!c) return;
  for (int i = 0; i < c->nopened; ++i) cudaIpcCloseMemHandle(c->opened[i]);
  signals_destroy(&c->sigs);
  delete c;
}

}  // namespace prb
