// Bucket assignment, the reducer state machine and the NCCL communicator.
//
//   ref: Eager_AssignGroupBySize   paddle/fluid/distributed/collective/reducer.cc:45-156
//        EagerReducer              reducer.h:88-142, reducer.cc:461-1070
//        ProcessGroupNCCL          process_group_nccl.cc:559-571,658-667,688-789,841-865
//
// B200-first differences from the reference (none changes a result):
//  * one persistent bucket arena, allocated once (the reference allocates a fresh bucket per group
//    per step, reducer.cc:359-360), segment starts padded to 256 B so pack/unpack are pure
//    256-bit vector passes;
//  * pack and the 1/nranks scale are ONE kernel (the reference runs concat, then scale_);
//  * NCCL is loaded at run time (dlopen), so the library also loads where NCCL is absent and a
//    single-rank communicator needs no NCCL at all.
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <map>
#include <mutex>
#include <vector>

#include "prb_bucket.cuh"
#include "prb_grad_pass.cuh"
#include "prb_nvlink.h"
#include "prb_nvls.h"

namespace prb {
namespace {

// ---------------------------------------------------------------- NCCL dynload
struct NcclApi {
  void* handle = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  ncclResult_t (*GetVersion)(int*) = nullptr;
  // optional (NCCL >= 2.19): user-buffer registration lets NVLink/NVLS collectives skip staging copies
  ncclResult_t (*CommRegister)(const ncclComm_t, void*, size_t, void**) = nullptr;
  ncclResult_t (*CommDeregister)(const ncclComm_t, void*) = nullptr;
};
NcclApi g_nccl;
std::mutex g_nccl_mu;

int32_t load_nccl() {
  std::lock_guard<std::mutex> lk(g_nccl_mu);
  if (g_nccl.handle) return PRB_SUCCESS;
  const char* env = getenv("PRB_NCCL_LIB");
  const char* names[] = {env, "libnccl.so.2", "libnccl.so"};
  void* h = nullptr;
  for (const char* n : names) {
    if (n == nullptr || *n == 0) continue;
    h = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    if (h) break;
  }
  PRB_ENFORCE(h != nullptr, PRB_EXTERNAL, "cannot load NCCL (libnccl.so.2): %s", dlerror());
#define PRB_SYM(field, name)                                                   \
  g_nccl.field = reinterpret_cast<decltype(g_nccl.field)>(dlsym(h, name));     \
  PRB_ENFORCE(g_nccl.field != nullptr, PRB_EXTERNAL, "NCCL symbol %s not found", name)
  PRB_SYM(GetUniqueId, "ncclGetUniqueId");
  PRB_SYM(CommInitRank, "ncclCommInitRank");
  PRB_SYM(CommDestroy, "ncclCommDestroy");
  PRB_SYM(AllReduce, "ncclAllReduce");
  PRB_SYM(Broadcast, "ncclBroadcast");
  PRB_SYM(AllGather, "ncclAllGather");
  PRB_SYM(GetErrorString, "ncclGetErrorString");
  PRB_SYM(GetVersion, "ncclGetVersion");
#undef PRB_SYM
  g_nccl.CommRegister = reinterpret_cast<decltype(g_nccl.CommRegister)>(dlsym(h, "ncclCommRegister"));
  g_nccl.CommDeregister = reinterpret_cast<decltype(g_nccl.CommDeregister)>(dlsym(h, "ncclCommDeregister"));
  g_nccl.handle = h;
  return PRB_SUCCESS;
}

#define PRB_NCCL(expr)                                                                               \
  do {                                                                                               \
    ncclResult_t _r = (expr);                                                                        \
    if (_r != ncclSuccess)                                                                           \
      PRB_FAIL(PRB_EXTERNAL, "NCCL error at %s:%d: %s", __FILE__, __LINE__, g_nccl.GetErrorString(_r)); \
  } while (0)

int32_t nccl_dtype(int32_t dt, ncclDataType_t* out) {
  switch (dt) {  // ref paddle/phi/core/utils/data_type.h:224-234
    case PRB_FLOAT32: *out = ncclFloat; return PRB_SUCCESS;
    case PRB_FLOAT64: *out = ncclDouble; return PRB_SUCCESS;
    case PRB_FLOAT16: *out = ncclHalf; return PRB_SUCCESS;
    case PRB_BFLOAT16: *out = ncclBfloat16; return PRB_SUCCESS;
    case PRB_INT32: *out = ncclInt32; return PRB_SUCCESS;
    case PRB_INT64: *out = ncclInt64; return PRB_SUCCESS;
    default: PRB_FAIL(PRB_INVALID_ARGUMENT, "no NCCL type for dtype %s", dtype_name(dt));
  }
}

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// the two small collectives prb_nvlink.cu needs during set-up
int32_t nvl_all_gather_bytes(void* comm, const void* send, void* recv, size_t bytes, cudaStream_t s) {
  PRB_NCCL(g_nccl.AllGather(send, recv, bytes, ncclChar, (ncclComm_t)comm, s));
  return PRB_SUCCESS;
}
int32_t nvl_all_min_int(void* comm, int value, int* out, cudaStream_t s) {
  int* d = nullptr;
  PRB_CUDA(cudaMalloc((void**)&d, sizeof(int)));
  PRB_CUDA(cudaMemcpyAsync(d, &value, sizeof(int), cudaMemcpyHostToDevice, s));
  ncclResult_t rc = g_nccl.AllReduce(d, d, 1, ncclInt32, ncclMin, (ncclComm_t)comm, s);
  if (rc == ncclSuccess) {
    cudaMemcpyAsync(out, d, sizeof(int), cudaMemcpyDeviceToHost, s);
    cudaStreamSynchronize(s);
  }
  cudaFree(d);
  PRB_ENFORCE(rc == ncclSuccess, PRB_EXTERNAL, "NCCL error: %s", g_nccl.GetErrorString(rc));
  return PRB_SUCCESS;
}

}  // namespace
}  // namespace prb

using namespace prb;

struct PRB_Comm {
  int nranks = 1, rank = 0, device = 0;
  ncclComm_t comm = nullptr;
  cudaStream_t stream = nullptr;  // dedicated high-priority comm stream (ref process_group_nccl.cc:688-737)
  cudaStream_t stream2 = nullptr; // second comm stream: the NVLink allreduce alternates between the two
};

namespace {
struct Group {
  std::vector<int64_t> vars;
  std::vector<int64_t> slot_off;  // element offset of each var inside the group's region (256-B padded)
  int64_t length = 0;             // padded element count handed to the allreduce
  int32_t dtype = 0;
  size_t arena_off = 0;           // byte offset of the region inside the arena
  int64_t pending = 0;
  cudaEvent_t ready = nullptr;    // recorded on the calc stream after pack
  cudaEvent_t done = nullptr;     // recorded on the comm stream after allreduce (+unpack)
  bool has_unused = false;
  // plans cached per group (rebuilt only when the set of unused parameters changes)
  DeviceTable pack_tab, unpack_tab;
  bool pack_valid = false, unpack_valid = false, pack_had_unused = false;
  DeviceTable packep_tab;         // pack with the check + norm epilogue (GradDesc table, grad_pass_kernel mode 2)
  bool packep_valid = false, packep_had_unused = false;
  DeviceTable exch_tab;           // chunk table of the group's slots for the exchange kernel's epilogue
  bool exch_valid = false;
  int64_t first_chunk = 0;        // index of the group's first 8192-element chunk in the reducer's own numbering
};

// Producer-side pass 1 (prb_reducer_set_grad_epilogue): what the kernel that writes a bucket leaves behind.
struct Epilogue {
  bool on = false;
  const float* loss_scale = nullptr;  // device; nullptr = no unscale / finite check
  bool* found_inf = nullptr;          // device (this rank's flag)
  bool norm = false, mp = false;
  std::vector<int32_t> partial_base;  // per var: first partial in the consumer's layout
  bool cleared = false;               // found_inf was cleared in this backward
  bool sidecar_cleared = false;       // ... and so was the flag the peers write (several ranks)
};
}  // namespace

struct PRB_Reducer {
  std::vector<PRB_Tensor> grads;
  std::vector<Group> groups;
  std::vector<std::pair<int64_t, int64_t>> locator;  // var -> (group, slot)   (ref reducer.h:82-86)
  PRB_Comm* comm = nullptr;
  char* arena = nullptr;
  size_t arena_bytes = 0;
  // Sidecar at the end of the arena (so that peers can reach it like the buckets): norm partials of every
  // 8192-element chunk of every gradient, the per-tensor sums / clip scratch of the consumer, and flags.
  size_t sidecar_off = 0;
  int64_t sidecar_floats = 0;
  int64_t total_chunks = 0;
  Epilogue ep;
  bool grads_in_arena = false;  // gradients are produced in their bucket slots: no pack, no unpack
  // Sharded step (prb_reducer_set_sharded): ONE reduce-scatter over all groups when the last one is ready instead of
  // an allreduce per group; rank r owns the r-th n-th of the chunk list of all groups in launch order.
  bool sharded = false;
  // Ownership layout of a sharded step (prb_reducer_set_shard_layout, fixed before the optimizer builds its views):
  //   false  rank r owns the r-th n-th of the chunk list of ALL groups: one reduce-scatter after the last pack;
  //   true   rank r owns the r-th n-th of EACH group's chunk list: one reduce-scatter per bucket, launched as soon as
  //          the bucket is packed, so that all but the last hide under the rest of the backward pass.
  bool shard_per_bucket = false;
  DeviceTable shard_tab;          // GradDesc chunk table of every slot, groups in launch order
  bool shard_valid = false;
  DeviceTable gather_tab;         // the same chunks with out = the caller's parameters (prb_reducer_allgather_params)
  std::vector<void*> gather_key;
  void* arena_reg = nullptr;  // ncclCommRegister handle of the arena (NULL when unavailable)
  NvlContext* nvl = nullptr;    // allreduce over NVLink peer memory (prb_nvlink.cu)
  NvlsContext* nvls = nullptr;  // allreduce through NVLink multicast (prb_nvls.cu); owns the arena when set
  size_t next_group = 0;
  std::vector<uint8_t> ready;
  std::vector<uint8_t> unused;
  std::vector<int64_t> unused_list;
  bool has_marked_unused = false;
  bool need_finalize = false;
  bool grad_need_hooks = false;
  bool finalized = false;
  bool unpack = true;
  bool deferred_wait = false;  // the consumer waits per group (prb_reducer_wait_group)
  bool find_unused = false;
  bool dry = false;  // host-only: state machine without device work (comm == NULL)
  // Batch mode (prb_reducer_mark_vars_ready with one rank): groups that complete inside one call are
  // packed by ONE launch instead of one launch per group -- with nothing to communicate there is no
  // pipelining to preserve.
  bool batching = false;
  std::vector<size_t> batch;
  std::vector<size_t> batch_key;
  // Tables of a batch.  With several ranks a batch is issued as two waves (a small first one, the rest) so that
  // the first wave's exchange runs under the second wave's pack.
  struct BatchWave {
    std::vector<size_t> members;
    DeviceTable pack_tab, unpack_tab, packep_tab, exch_tab;
    bool valid = false, packep_valid = false, exch_valid = false;
  };
  BatchWave waves[2];
  std::vector<int64_t> launched;
};

namespace {

int32_t group_tables(PRB_Reducer* r, Group& g, bool for_unpack, cudaStream_t stream, DeviceTable* tab) {
  std::vector<MapDesc> descs;
  std::vector<int64_t> numels;
  const int es = dtype_size(g.dtype);
  for (size_t s = 0; s < g.vars.size(); ++s) {
    const int64_t v = g.vars[s];
    if (!for_unpack && r->unused[v]) continue;  // slot is zero-filled instead (ref reducer.cc:833-855)
    MapDesc d{};
    char* slot = r->arena + g.arena_off + (size_t)g.slot_off[s] * es;
    d.src = for_unpack ? (const void*)slot : (const void*)r->grads[v].data;
    d.dst = for_unpack ? r->grads[v].data : (void*)slot;
    d.aligned = map_aligned(d.src, d.dst, es) ? 1 : 0;
    descs.push_back(d);
    numels.push_back(r->grads[v].numel);
  }
  return get_table(descs.data(), sizeof(MapDesc), (int)descs.size(), numels.data(), map_chunk_elems(), stream, tab,
                   /*pin=*/true);
}

float* sidecar(PRB_Reducer* r) { return reinterpret_cast<float*>(r->arena + r->sidecar_off); }

// GradDesc of var v for the pack with epilogue: gradient -> slot, partials at the consumer's position.
GradDesc packep_desc(PRB_Reducer* r, const Group& g, size_t s) {
  const int64_t v = g.vars[s];
  const int es = dtype_size(g.dtype);
  GradDesc d{};
  d.x = r->grads[v].data;
  d.out = r->arena + g.arena_off + (size_t)g.slot_off[s] * es;
  d.aligned = map_aligned(d.x, d.out, es) ? 1 : 0;
  d.partial_base = r->ep.norm ? r->ep.partial_base[(size_t)v] : 0;
  return d;
}

int32_t group_packep_table(PRB_Reducer* r, Group& g, cudaStream_t stream, DeviceTable* tab) {
  std::vector<GradDesc> descs;
  std::vector<int64_t> numels;
  for (size_t s = 0; s < g.vars.size(); ++s) {
    if (r->unused[g.vars[s]]) continue;
    descs.push_back(packep_desc(r, g, s));
    numels.push_back(r->grads[g.vars[s]].numel);
  }
  return get_table(descs.data(), sizeof(GradDesc), (int)descs.size(), numels.data(), grad_chunk_elems(), stream, tab,
                   /*pin=*/true);
}

GradPassOpts packep_opts(PRB_Reducer* r) {
  GradPassOpts o;
  o.mode = 2;
  o.norm = r->ep.norm;
  o.mp = r->ep.mp;
  o.loss_scale = r->ep.loss_scale;
  o.found_inf = r->ep.found_inf;
  o.partials = sidecar(r);
  o.pack_scale = (float)(1.0 / (double)r->comm->nranks);  // Scalar::to<MT>() (ref scale_kernel.cu:64-66)
  o.pack_bias = 0.0f;
  return o;
}

// Is the epilogue carried by the pack kernel in this configuration?  (One rank, gradients outside the arena.
// With several ranks the check must see the REDUCED value: it rides in the exchange kernel instead.)
bool epilogue_in_pack(const PRB_Reducer* r) { return r->ep.on && r->comm->nranks == 1 && !r->grads_in_arena; }
// Several ranks: only the library's own peer-memory exchange kernel can carry the epilogue and fold the scale.
bool epilogue_in_exchange(const PRB_Reducer* r) { return r->ep.on && r->comm->nranks > 1 && r->nvl != nullptr; }
bool epilogue_active(const PRB_Reducer* r) { return epilogue_in_pack(r) || epilogue_in_exchange(r); }
bool exchange_folds_scale(const PRB_Reducer* r) { return r->comm->nranks > 1 && r->nvl != nullptr; }

// Once per backward, before the first bucket is written: clear this rank's found_inf.
int32_t epilogue_begin(PRB_Reducer* r, cudaStream_t calc) {
  if (!epilogue_active(r) || r->ep.cleared) return PRB_SUCCESS;
  if (r->ep.loss_scale != nullptr) PRB_CUDA(cudaMemsetAsync(r->ep.found_inf, 0, 1, calc));
  r->ep.cleared = true;
  return PRB_SUCCESS;
}

// A parameter without a gradient in this backward: zero slot (ref reducer.cc:833-855) and, with the epilogue,
// zero norm partials (nothing will visit its chunks on this rank when the pack carries the epilogue).
int32_t zero_unused_slots(PRB_Reducer* r, Group& g, cudaStream_t calc) {
  const int es = dtype_size(g.dtype);
  char* region = r->arena + g.arena_off;
  for (size_t s = 0; s < g.vars.size(); ++s) {
    const int64_t v = g.vars[s];
    if (!r->unused[v]) continue;
    PRB_CUDA(cudaMemsetAsync(region + (size_t)g.slot_off[s] * es, 0, (size_t)r->grads[v].numel * es, calc));
    if (epilogue_in_pack(r) && r->ep.norm) {
      const int64_t nch = (r->grads[v].numel + grad_chunk_elems() - 1) / grad_chunk_elems();
      PRB_CUDA(cudaMemsetAsync(sidecar(r) + r->ep.partial_base[(size_t)v], 0, (size_t)nch * sizeof(float), calc));
    }
  }
  return PRB_SUCCESS;
}

// GradDesc of var v for the exchange kernel: x = the slot in THIS rank's arena (peers are reached by offset).
GradDesc exch_desc(PRB_Reducer* r, const Group& g, size_t s) {
  GradDesc d{};
  d.x = r->arena + g.arena_off + (size_t)g.slot_off[s] * dtype_size(g.dtype);
  d.out = nullptr;
  d.aligned = 1;
  d.partial_base = r->ep.norm ? r->ep.partial_base[(size_t)g.vars[s]] : 0;
  return d;
}

int32_t exchange_table(PRB_Reducer* r, const std::vector<size_t>& gis, cudaStream_t stream, DeviceTable* tab) {
  std::vector<GradDesc> descs;
  std::vector<int64_t> numels;
  for (size_t gi : gis) {
    const Group& g = r->groups[gi];
    for (size_t s = 0; s < g.vars.size(); ++s) {
      descs.push_back(exch_desc(r, g, s));
      numels.push_back(r->grads[(size_t)g.vars[s]].numel);
    }
  }
  return get_table(descs.data(), sizeof(GradDesc), (int)descs.size(), numels.data(), grad_chunk_elems(), stream, tab,
                   /*pin=*/true);
}

bool batch_exchange_enabled() {
  static const bool v = [] {
    const char* e = getenv("PRB_BATCH_EXCHANGE");
    return !(e != nullptr && atoi(e) == 0);
  }();
  return v;
}

size_t sidecar_flags_off(const PRB_Reducer* r) {
  return r->sidecar_off + align_up((size_t)r->sidecar_floats * sizeof(float), 256);
}

// The library's own exchange of the contiguous arena region of groups [first, last] on comm stream `cs`.
int32_t nvl_exchange_groups(PRB_Reducer* r, size_t first, size_t last, const DeviceTable* table, cudaStream_t cs,
                            bool scatter_only = false) {
  const Group &g0 = r->groups[first], &g1 = r->groups[last];
  const int es = dtype_size(g0.dtype);
  NvlWork w;
  w.region_off = g0.arena_off;
  w.numel = (int64_t)((g1.arena_off + (size_t)g1.length * es - g0.arena_off) / es);
  w.dtype = g0.dtype;
  w.fold_scale = r->grads_in_arena;
  w.scale = (float)(1.0 / (double)r->comm->nranks);
  if (scatter_only) {
    // reduce-scatter of one bucket while the backward pass is still running: the owner keeps the reduced values;
    // the grid stays at the allreduce's cap so that the calc stream's kernels are not slowed down
    w.table = table;
    w.scatter_only = true;
    w.overlap_grid = true;
  }
  if (table != nullptr && epilogue_in_exchange(r)) {
    w.table = table;
    w.loss_scale = r->ep.loss_scale;
    w.found_local = r->ep.found_inf;
    w.norm = r->ep.norm;
    w.mp = r->ep.mp;
    w.partials_off = r->sidecar_off;
    w.flags_off = sidecar_flags_off(r);
  }
  return nvl_exchange(r->nvl, w, cs);
}

// Once per backward with the epilogue in the exchange: clear the sidecar flag the peers write into.
int32_t epilogue_begin_exchange(PRB_Reducer* r, cudaStream_t calc) {
  if (!epilogue_in_exchange(r) || r->ep.loss_scale == nullptr || r->ep.sidecar_cleared) return PRB_SUCCESS;
  PRB_CUDA(cudaMemsetAsync(r->arena + sidecar_flags_off(r), 0, 256, calc));
  r->ep.sidecar_cleared = true;
  return PRB_SUCCESS;
}

// A dead exchange context (a peer never arrived): raise on the host instead of carrying on with unreduced gradients.
int32_t check_comm_alive(const PRB_Reducer* r) {
  const int dead = r->nvls != nullptr ? nvls_error_flag(r->nvls) : (r->nvl != nullptr ? nvl_error_flag(r->nvl) : 0);
  PRB_ENFORCE(!dead, PRB_EXTERNAL,
              "bucket exchange failed: a peer rank did not arrive within PRB_NVL_TIMEOUT_MS or aborted; the gradients of "
              "that step were NOT reduced and this communicator is dead (restart the job)");
  return PRB_SUCCESS;
}

bool sharded_active(const PRB_Reducer* r) { return r->sharded && r->comm->nranks > 1 && r->nvl != nullptr; }

std::vector<size_t> all_groups(const PRB_Reducer* r) {
  std::vector<size_t> gis(r->groups.size());
  for (size_t i = 0; i < gis.size(); ++i) gis[i] = i;
  return gis;
}

// Sharded step, first half: every bucket is packed; one reduce-scatter over all of them on the comm stream.  The
// slice owner keeps the reduced gradient and leaves the inf flag and the norm partials on EVERY rank.
int32_t launch_sharded_reduce_scatter(PRB_Reducer* r, cudaStream_t calc) {
  PRB_Comm* c = r->comm;
  if (!r->shard_valid) {
    PRB_TRY(exchange_table(r, all_groups(r), calc, &r->shard_tab));
    r->shard_valid = true;
  }
  cudaStream_t cs = nvl_next_slot(r->nvl) == 1 ? c->stream2 : c->stream;
  Group& g0 = r->groups.front();
  Group& g1 = r->groups.back();
  PRB_CUDA(cudaEventRecord(g0.ready, calc));
  PRB_CUDA(cudaStreamWaitEvent(cs, g0.ready, 0));
  const int es = dtype_size(g0.dtype);
  NvlWork w;
  w.region_off = g0.arena_off;
  w.numel = (int64_t)((g1.arena_off + (size_t)g1.length * es - g0.arena_off) / es);
  w.dtype = g0.dtype;
  w.fold_scale = r->grads_in_arena;
  w.scale = (float)(1.0 / (double)c->nranks);
  w.table = &r->shard_tab;
  w.scatter_only = true;
  if (r->ep.on) {
    w.loss_scale = r->ep.loss_scale;
    w.found_local = r->ep.found_inf;
    w.norm = r->ep.norm;
    w.mp = r->ep.mp;
    w.partials_off = r->sidecar_off;
    w.flags_off = sidecar_flags_off(r);
  }
  PRB_TRY(nvl_exchange(r->nvl, w, cs));
  for (Group& g : r->groups) PRB_CUDA(cudaEventRecord(g.done, cs));
  return PRB_SUCCESS;
}

// ref: FusedAllReduceSchedule reducer.cc:1037-1070 ("concat > div_nranks > allreduce > split")
int32_t schedule_group(PRB_Reducer* r, size_t gi, cudaStream_t calc) {
  Group& g = r->groups[gi];
  if (r->dry) {
    r->launched.push_back((int64_t)gi);
    return PRB_SUCCESS;
  }
  PRB_Comm* c = r->comm;
  const int es = dtype_size(g.dtype);
  // Batch mode: groups that complete inside ONE mark_vars_ready call.  One rank: a single pack launch for all of
  // them.  Several ranks: only when the consumer has a global dependency anyway (epilogue set: it cannot start
  // before the last bucket is reduced) -- then one pack and ONE exchange (one barrier pair) serve all of them;
  // without such a dependency buckets stay separate so that the optimizer can consume them one by one.
  // (With find_unused_parameters the ranks may disagree about which groups hold unused parameters, and every rank
  // must issue the same sequence of exchanges: no batching then.)
  const bool shard_unit = sharded_active(r) && !r->shard_per_bucket;   // one reduce-scatter for all buckets
  const bool shard_each = sharded_active(r) && r->shard_per_bucket;    // one per bucket, like the allreduce
  const bool batch_here = !shard_each &&
                          ((c->nranks == 1 && !r->grads_in_arena) ||
                           (c->nranks > 1 && epilogue_in_exchange(r) && batch_exchange_enabled() && !r->find_unused) ||
                           (shard_unit && !r->grads_in_arena));
  if (r->batching && batch_here && !g.has_unused && !g.pack_had_unused && !g.packep_had_unused &&
      (r->batch.empty() || r->groups[r->batch.front()].dtype == g.dtype)) {
    r->batch.push_back(gi);
    r->launched.push_back((int64_t)gi);
    return PRB_SUCCESS;
  }
  char* region = r->arena + g.arena_off;
  PRB_TRY(epilogue_begin(r, calc));
  PRB_TRY(epilogue_begin_exchange(r, calc));
  if (g.has_unused) PRB_TRY(zero_unused_slots(r, g, calc));
  if (r->grads_in_arena) {
    // the producer wrote the slots: nothing to pack; with several ranks the 1/nranks scale runs in place
    // (ref scale_ reducer.cc:1049-1050) unless the exchange kernel folds it in
    if (c->nranks > 1 && !exchange_folds_scale(r)) {
      if (!g.pack_valid) {
        MapDesc d{};
        d.src = region;
        d.dst = region;
        d.aligned = 1;
        PRB_TRY(get_table(&d, sizeof(d), 1, &g.length, map_chunk_elems(), calc, &g.pack_tab, /*pin=*/true));
        g.pack_valid = true;
      }
      PRB_TRY(launch_scale(g.dtype, g.pack_tab, 1.0 / (double)c->nranks, 0.0f, false, calc));
    }
  } else if (epilogue_in_pack(r)) {
    // pack + scale + unscale/inf check + norm partials in one pass, calc stream
    if (!g.packep_valid || g.has_unused || g.packep_had_unused) {
      PRB_TRY(group_packep_table(r, g, calc, &g.packep_tab));
      g.packep_valid = true;
      g.packep_had_unused = g.has_unused;
    }
    PRB_TRY(launch_grad_pass(g.dtype, g.packep_tab, packep_opts(r), calc));
  } else {
    // pack + scale_(bucket, 1.0/nranks, 0.0, false) in one pass, calc stream
    if (!g.pack_valid || g.has_unused || g.pack_had_unused) {
      PRB_TRY(group_tables(r, g, false, calc, &g.pack_tab));
      g.pack_valid = true;
      g.pack_had_unused = g.has_unused;
    }
    PRB_TRY(launch_scale(g.dtype, g.pack_tab, 1.0 / (double)c->nranks, 0.0f, false, calc));
  }
  if (shard_unit) {
    // nothing to exchange yet: the reduce-scatter covers all groups and starts when the last one is packed
    r->launched.push_back((int64_t)gi);
    if (gi + 1 == r->groups.size()) PRB_TRY(launch_sharded_reduce_scatter(r, calc));
    return PRB_SUCCESS;
  }
  if (c->nranks > 1) {
    // SyncDefaultStream: comm stream waits for the calc stream (ref process_group_nccl.cc:658-667)
    cudaStream_t cs = c->stream;
    // our own kernels alternate between two comm streams: one bucket's barriers hide under the next one's transfers
    if (r->nvls != nullptr ? nvls_next_slot(r->nvls) == 1 : (r->nvl != nullptr && nvl_next_slot(r->nvl) == 1)) cs = c->stream2;
    PRB_CUDA(cudaEventRecord(g.ready, calc));
    PRB_CUDA(cudaStreamWaitEvent(cs, g.ready, 0));
    if (r->nvls != nullptr) {
      PRB_TRY(nvls_allreduce(r->nvls, g.arena_off, g.length, g.dtype, cs));
    } else if (r->nvl != nullptr) {
      if ((epilogue_in_exchange(r) || shard_each) && !g.exch_valid) {
        PRB_TRY(exchange_table(r, {gi}, cs, &g.exch_tab));
        g.exch_valid = true;
      }
      PRB_TRY(nvl_exchange_groups(r, gi, gi, (epilogue_in_exchange(r) || shard_each) ? &g.exch_tab : nullptr, cs, shard_each));
    } else {
      ncclDataType_t nt;
      PRB_TRY(nccl_dtype(g.dtype, &nt));
      PRB_NCCL(g_nccl.AllReduce(region, region, (size_t)g.length, nt, ncclSum, c->comm, cs));
    }
    if (r->unpack && !r->grads_in_arena && !shard_each) {  // the reference also splits on the comm stream (reducer.cc:1062-1069)
      if (!g.unpack_valid) {
        PRB_TRY(group_tables(r, g, true, cs, &g.unpack_tab));
        g.unpack_valid = true;
      }
      PRB_TRY(launch_copy(es, g.unpack_tab, cs));
    }
    PRB_CUDA(cudaEventRecord(g.done, cs));
  } else if (r->unpack && !r->grads_in_arena) {
    if (!g.unpack_valid) {
      PRB_TRY(group_tables(r, g, true, calc, &g.unpack_tab));
      g.unpack_valid = true;
    }
    PRB_TRY(launch_copy(es, g.unpack_tab, calc));
  }
  r->launched.push_back((int64_t)gi);
  return PRB_SUCCESS;
}

int batch_waves() {
  static const int v = [] {
    const char* e = getenv("PRB_BATCH_WAVES");
    return (e != nullptr && atoi(e) == 1) ? 1 : 2;
  }();
  return v;
}

int32_t flush_wave(PRB_Reducer* r, PRB_Reducer::BatchWave& w, cudaStream_t calc) {
  if (w.members.empty()) return PRB_SUCCESS;
  PRB_Comm* c = r->comm;
  const bool multi = c->nranks > 1;
  const bool with_ep = epilogue_in_pack(r);
  const bool need_pack = !r->grads_in_arena;
  if (need_pack && (with_ep ? !w.packep_valid : !w.valid)) {
    std::vector<MapDesc> pack, unpack;
    std::vector<GradDesc> packep;
    std::vector<int64_t> numels;
    for (size_t gi : w.members) {
      Group& g = r->groups[gi];
      const int es = dtype_size(g.dtype);
      for (size_t s = 0; s < g.vars.size(); ++s) {
        const int64_t v = g.vars[s];
        char* slot = r->arena + g.arena_off + (size_t)g.slot_off[s] * es;
        MapDesc d{};
        d.src = r->grads[v].data;
        d.dst = slot;
        d.aligned = map_aligned(d.src, d.dst, es) ? 1 : 0;
        pack.push_back(d);
        MapDesc u{};
        u.src = slot;
        u.dst = r->grads[v].data;
        u.aligned = d.aligned;
        unpack.push_back(u);
        if (with_ep) packep.push_back(packep_desc(r, g, s));
        numels.push_back(r->grads[v].numel);
      }
    }
    if (with_ep) {
      PRB_TRY(get_table(packep.data(), sizeof(GradDesc), (int)packep.size(), numels.data(), grad_chunk_elems(), calc,
                        &w.packep_tab, /*pin=*/true));
      w.packep_valid = true;
    } else {
      PRB_TRY(get_table(pack.data(), sizeof(MapDesc), (int)pack.size(), numels.data(), map_chunk_elems(), calc,
                        &w.pack_tab, /*pin=*/true));
      w.valid = true;
    }
    PRB_TRY(get_table(unpack.data(), sizeof(MapDesc), (int)unpack.size(), numels.data(), map_chunk_elems(), calc,
                      &w.unpack_tab, /*pin=*/true));
  }
  if (multi && epilogue_in_exchange(r) && !w.exch_valid && !sharded_active(r)) {
    PRB_TRY(exchange_table(r, w.members, calc, &w.exch_tab));
    w.exch_valid = true;
  }
  const size_t first = w.members.front(), last = w.members.back();
  const int32_t dt = r->groups[first].dtype;
  if (need_pack) {
    if (with_ep) PRB_TRY(launch_grad_pass(dt, w.packep_tab, packep_opts(r), calc));
    else PRB_TRY(launch_scale(dt, w.pack_tab, 1.0 / (double)c->nranks, 0.0f, false, calc));
  }
  if (!multi) {
    if (r->unpack && need_pack) PRB_TRY(launch_copy(dtype_size(dt), w.unpack_tab, calc));
    return PRB_SUCCESS;
  }
  if (sharded_active(r)) return PRB_SUCCESS;  // flush_batch launches the one reduce-scatter after the last pack
  // several ranks: ONE exchange for the whole wave (the groups' regions are adjacent in the arena)
  cudaStream_t cs = nvl_next_slot(r->nvl) == 1 ? c->stream2 : c->stream;
  Group& g0 = r->groups[first];
  PRB_CUDA(cudaEventRecord(g0.ready, calc));
  PRB_CUDA(cudaStreamWaitEvent(cs, g0.ready, 0));
  PRB_TRY(nvl_exchange_groups(r, first, last, &w.exch_tab, cs));
  if (r->unpack && need_pack) PRB_TRY(launch_copy(dtype_size(dt), w.unpack_tab, cs));
  for (size_t gi : w.members) PRB_CUDA(cudaEventRecord(r->groups[gi].done, cs));
  return PRB_SUCCESS;
}

int32_t flush_batch(PRB_Reducer* r, cudaStream_t calc) {
  if (r->batch.empty()) return PRB_SUCCESS;
  if (r->batch_key != r->batch) {
    // split: one rank -> one wave.  Several ranks -> a first wave of ~1/4 of the bytes (its exchange starts early and
    // runs under the pack of the rest), then everything else; the exchange, not the pack, bounds the batch.
    for (auto& w : r->waves) {
      for (const DeviceTable* t : {&w.pack_tab, &w.unpack_tab, &w.packep_tab, &w.exch_tab}) release_table(*t);
      w = PRB_Reducer::BatchWave();
    }
    size_t split = r->batch.size();
    if (r->comm->nranks > 1 && batch_waves() > 1 && r->batch.size() >= 2 && !sharded_active(r)) {
      size_t total = 0, run = 0;
      for (size_t gi : r->batch) total += (size_t)r->groups[gi].length * dtype_size(r->groups[gi].dtype);
      for (split = 0; split + 1 < r->batch.size(); ++split) {
        run += (size_t)r->groups[r->batch[split]].length * dtype_size(r->groups[r->batch[split]].dtype);
        if (run * 4 >= total) {
          ++split;
          break;
        }
      }
    }
    r->waves[0].members.assign(r->batch.begin(), r->batch.begin() + (long)split);
    r->waves[1].members.assign(r->batch.begin() + (long)split, r->batch.end());
    r->batch_key = r->batch;
  }
  r->batch.clear();
  PRB_TRY(epilogue_begin(r, calc));
  PRB_TRY(epilogue_begin_exchange(r, calc));
  const bool holds_last = !r->waves[0].members.empty() &&
                          (r->waves[1].members.empty() ? r->waves[0].members.back() : r->waves[1].members.back()) + 1 ==
                              r->groups.size();
  PRB_TRY(flush_wave(r, r->waves[0], calc));
  PRB_TRY(flush_wave(r, r->waves[1], calc));
  if (sharded_active(r) && !r->shard_per_bucket && holds_last) PRB_TRY(launch_sharded_reduce_scatter(r, calc));
  return PRB_SUCCESS;
}

// ref: FinalizeBackward reducer.cc:1013-1035 -- task->Synchronize(): calc stream waits on comm events
int32_t finalize(PRB_Reducer* r, cudaStream_t calc) {
  r->need_finalize = false;
  r->grad_need_hooks = false;
  if (!r->dry && r->comm->nranks > 1 && !r->deferred_wait)
    for (Group& g : r->groups) PRB_CUDA(cudaStreamWaitEvent(calc, g.done, 0));
  r->finalized = true;
  return PRB_SUCCESS;
}

int32_t mark_group_ready(PRB_Reducer* r, size_t gi, cudaStream_t calc) {
  // ref: MarkGroupReady reducer.cc:903-930
  PRB_ENFORCE(gi >= r->next_group, PRB_PRECONDITION_NOT_MET,
              "The index of the incoming group must be greater than or equal to the previously synchronized group "
              "index, expect it to greater than or equal to %zu, but got %zu.",
              r->next_group, gi);
  if (gi > r->next_group) return PRB_SUCCESS;  // launched later, strictly in order
  for (; r->next_group < r->groups.size() && r->groups[r->next_group].pending == 0; ++r->next_group)
    PRB_TRY(schedule_group(r, r->next_group, calc));
  return PRB_SUCCESS;
}

int32_t mark_var_ready_impl(PRB_Reducer* r, int64_t v, bool is_used, cudaStream_t calc) {
  // ref: MarkVarReady reducer.cc:774-901
  PRB_ENFORCE(!r->ready[v], PRB_PRECONDITION_NOT_MET,
              "Error happened, when parameter[%lld] has been ready before. Please set find_unused_parameters=True "
              "to traverse backward graph in each step to prepare reduce in advance.",
              (long long)v);
  r->ready[v] = 1;
  r->need_finalize = true;
  (void)is_used;
  const size_t gi = (size_t)r->locator[v].first;
  if (--r->groups[gi].pending == 0) PRB_TRY(mark_group_ready(r, gi, calc));
  if (r->next_group == r->groups.size()) PRB_TRY(finalize(r, calc));
  return PRB_SUCCESS;
}

}  // namespace

extern "C" {

int32_t prb_assign_group_by_size(const int64_t* numels, const int32_t* dtypes, const uint8_t* is_sparse, int64_t num,
                                 const uint64_t* limits, int64_t num_limits, const int64_t* tensor_indices,
                                 int64_t num_indices, int64_t* out_flat, int64_t* out_group_sizes,
                                 int64_t* out_num_groups) {
  PRB_ENFORCE(num >= 0 && num_limits >= 1 && limits != nullptr, PRB_INVALID_ARGUMENT,
              "assign_group_by_size: need at least one group size limit");
  if (num_indices > 0) {  // check_perm (ref reducer.cc:57-75)
    std::vector<uint8_t> cnt((size_t)num_indices, 0);
    for (int64_t i = 0; i < num_indices; ++i) {
      const int64_t x = tensor_indices[i];
      PRB_ENFORCE(x >= 0 && x < num_indices && !cnt[(size_t)x], PRB_PRECONDITION_NOT_MET,
                  "tensor_indices must be a permutation from 0 to %lld", (long long)num_indices);
      cnt[(size_t)x] = 1;
    }
  }
  std::vector<std::vector<int64_t>> res;
  std::map<int32_t, size_t> limit_index;                               // dtype -> index into limits
  std::map<int32_t, std::pair<std::vector<int64_t>, uint64_t>> next;   // dtype -> (open group, bytes)
  for (int64_t i = 0; i < num; ++i) {
    const int64_t real = num_indices > 0 ? tensor_indices[i] : i;
    if (is_sparse[real]) {  // a sparse var is a group of its own
      res.push_back({real});
      continue;
    }
    const int32_t dt = dtypes[i];
    const int es = dtype_size(dt);
    PRB_ENFORCE(es > 0, PRB_INVALID_ARGUMENT, "assign_group_by_size: unknown dtype code %d", dt);
    auto& grp = next[dt];
    grp.first.push_back(real);
    grp.second += (uint64_t)es * (uint64_t)numels[i];
    if (!limit_index.count(dt)) limit_index[dt] = 0;
    size_t& li = limit_index[dt];
    if (grp.second >= limits[li]) {  // ">=" after adding: a bucket may overshoot by one tensor
      res.emplace_back(std::move(grp.first));
      grp = std::pair<std::vector<int64_t>, uint64_t>();
      li = std::min(li + 1, (size_t)num_limits - 1);
    }
  }
  for (auto& e : next)
    if (!e.second.first.empty()) res.emplace_back(std::move(e.second.first));
  if (num_indices == 0)
    std::sort(res.begin(), res.end(),
              [](const std::vector<int64_t>& x, const std::vector<int64_t>& y) { return x.front() < y.front(); });
  int64_t k = 0;
  for (size_t g = 0; g < res.size(); ++g) {
    out_group_sizes[g] = (int64_t)res[g].size();
    for (int64_t v : res[g]) out_flat[k++] = v;
  }
  *out_num_groups = (int64_t)res.size();
  return PRB_SUCCESS;
}

// ---------------------------------------------------------------- communicator
int32_t prb_comm_get_unique_id(uint8_t id[PRB_UNIQUE_ID_BYTES]) {
  PRB_TRY(load_nccl());
  static_assert(sizeof(ncclUniqueId) == PRB_UNIQUE_ID_BYTES, "ncclUniqueId size");
  ncclUniqueId uid;
  PRB_NCCL(g_nccl.GetUniqueId(&uid));
  memcpy(id, &uid, sizeof(uid));
  return PRB_SUCCESS;
}

int32_t prb_comm_init_rank(const uint8_t id[PRB_UNIQUE_ID_BYTES], int32_t nranks, int32_t rank, int32_t device,
                           PRB_Comm** out) {
  PRB_TRY(require_device());
  PRB_ENFORCE(nranks >= 1 && rank >= 0 && rank < nranks, PRB_INVALID_ARGUMENT, "comm_init_rank: bad rank %d of %d", rank,
              nranks);
  PRB_TRY(load_nccl());
  PRB_CUDA(cudaSetDevice(device));
  PRB_Comm* c = new PRB_Comm();
  c->nranks = nranks;
  c->rank = rank;
  c->device = device;
  ncclUniqueId uid;
  memcpy(&uid, id, sizeof(uid));
  ncclResult_t rc = g_nccl.CommInitRank(&c->comm, nranks, uid, rank);
  if (rc != ncclSuccess) {
    delete c;
    PRB_FAIL(PRB_EXTERNAL, "ncclCommInitRank failed: %s", g_nccl.GetErrorString(rc));
  }
  int lo = 0, hi = 0;
  PRB_CUDA(cudaDeviceGetStreamPriorityRange(&lo, &hi));
  PRB_CUDA(cudaStreamCreateWithPriority(&c->stream, cudaStreamNonBlocking, hi));
  PRB_CUDA(cudaStreamCreateWithPriority(&c->stream2, cudaStreamNonBlocking, hi));
  *out = c;
  return PRB_SUCCESS;
}

int32_t prb_comm_init_single(int32_t device, PRB_Comm** out) {
  PRB_TRY(require_device());
  PRB_CUDA(cudaSetDevice(device));
  PRB_Comm* c = new PRB_Comm();
  c->device = device;
  int lo = 0, hi = 0;
  PRB_CUDA(cudaDeviceGetStreamPriorityRange(&lo, &hi));
  PRB_CUDA(cudaStreamCreateWithPriority(&c->stream, cudaStreamNonBlocking, hi));
  *out = c;
  return PRB_SUCCESS;
}

int32_t prb_comm_destroy(PRB_Comm* c) {
  if (!c) return PRB_SUCCESS;
  if (c->comm) g_nccl.CommDestroy(c->comm);
  if (c->stream) cudaStreamDestroy(c->stream);
  if (c->stream2) cudaStreamDestroy(c->stream2);
  delete c;
  return PRB_SUCCESS;
}

int32_t prb_comm_nranks(const PRB_Comm* c) { return c ? c->nranks : 0; }
int32_t prb_comm_rank(const PRB_Comm* c) { return c ? c->rank : -1; }
void* prb_comm_stream(const PRB_Comm* c) { return c ? (void*)c->stream : nullptr; }

int32_t prb_comm_all_reduce(PRB_Comm* c, const PRB_Tensor* in, PRB_Tensor* out, int32_t op, void* stream) {
  PRB_ENFORCE(c && in && out && in->numel == out->numel && in->dtype == out->dtype, PRB_INVALID_ARGUMENT,
              "all_reduce: bad arguments");
  PRB_ENFORCE(op == 0 || op == 1, PRB_INVALID_ARGUMENT, "all_reduce: op must be 0 (sum) or 1 (max)");
  if (c->nranks == 1) {
    if (out->data != in->data)
      PRB_CUDA(cudaMemcpyAsync(out->data, in->data, (size_t)in->numel * dtype_size(in->dtype), cudaMemcpyDeviceToDevice,
                               (cudaStream_t)stream));
    return PRB_SUCCESS;
  }
  ncclDataType_t nt;
  PRB_TRY(nccl_dtype(in->dtype, &nt));
  PRB_NCCL(g_nccl.AllReduce(in->data, out->data, (size_t)in->numel, nt, op == 0 ? ncclSum : ncclMax, c->comm,
                            (cudaStream_t)stream));
  return PRB_SUCCESS;
}

int32_t prb_comm_all_gather(PRB_Comm* c, const PRB_Tensor* in, PRB_Tensor* out, void* stream) {
  PRB_ENFORCE(c && in && out && in->dtype == out->dtype && out->numel == in->numel * c->nranks, PRB_INVALID_ARGUMENT,
              "all_gather: out must hold nranks x in.numel elements of in's dtype");
  const size_t bytes = (size_t)in->numel * dtype_size(in->dtype);
  if (c->nranks == 1) {
    if (out->data != in->data && bytes > 0)
      PRB_CUDA(cudaMemcpyAsync(out->data, in->data, bytes, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    return PRB_SUCCESS;
  }
  PRB_NCCL(g_nccl.AllGather(in->data, out->data, bytes, ncclChar, c->comm, (cudaStream_t)stream));
  return PRB_SUCCESS;
}

int32_t prb_comm_broadcast(PRB_Comm* c, PRB_Tensor* buf, int32_t root, void* stream) {
  PRB_ENFORCE(c && buf && root >= 0 && root < c->nranks, PRB_INVALID_ARGUMENT, "broadcast: bad arguments");
  if (c->nranks == 1) return PRB_SUCCESS;
  ncclDataType_t nt;
  PRB_TRY(nccl_dtype(buf->dtype, &nt));
  PRB_NCCL(g_nccl.Broadcast(buf->data, buf->data, (size_t)buf->numel, nt, root, c->comm, (cudaStream_t)stream));
  return PRB_SUCCESS;
}

// ---------------------------------------------------------------- reducer
int32_t prb_reducer_create(const PRB_Tensor* grads, int64_t num, const int64_t* group_flat, const int64_t* group_sizes,
                           int64_t num_groups, PRB_Comm* comm, int32_t find_unused_parameters, PRB_Reducer** out) {
  if (comm != nullptr) PRB_TRY(require_device());
  PRB_ENFORCE(grads && group_flat && group_sizes && out && num > 0 && num_groups > 0, PRB_INVALID_ARGUMENT,
              "reducer_create: bad arguments");
  PRB_Reducer* r = new PRB_Reducer();
  r->dry = comm == nullptr;
  r->grads.assign(grads, grads + num);
  r->comm = comm;
  r->find_unused = find_unused_parameters != 0;
  r->locator.assign((size_t)num, {-1, -1});
  r->ready.assign((size_t)num, 0);
  r->unused.assign((size_t)num, 0);
  size_t arena = 0;
  int64_t k = 0;
  int32_t status = PRB_SUCCESS;
  for (int64_t gi = 0; gi < num_groups && status == PRB_SUCCESS; ++gi) {
    Group g;
    if (group_sizes[gi] <= 0) {  // ref reducer.cc:536-540
      set_error("The number of group[%lld]'s elements is 0.", (long long)gi);
      status = PRB_PRECONDITION_NOT_MET;
      break;
    }
    int64_t off = 0;
    for (int64_t s = 0; s < group_sizes[gi]; ++s) {
      const int64_t v = group_flat[k++];
      if (v < 0 || v >= num || r->locator[(size_t)v].first >= 0) {
        set_error("reducer_create: group %lld holds an invalid or repeated tensor index %lld", (long long)gi, (long long)v);
        status = PRB_INVALID_ARGUMENT;
        break;
      }
      const PRB_Tensor& t = r->grads[(size_t)v];
      if (t.numel <= 0) {  // ref reducer.cc:589-593
        set_error("The number of tensor %lld's elements is 0.", (long long)v);
        status = PRB_PRECONDITION_NOT_MET;
        break;
      }
      if (s == 0) g.dtype = t.dtype;
      if (t.dtype != g.dtype) {  // ref reducer.cc:607-612
        set_error("Tensor %lld has unexpected dtype.", (long long)v);
        status = PRB_PRECONDITION_NOT_MET;
        break;
      }
      const int es = dtype_size(t.dtype);
      if (t.data == nullptr && !r->dry) r->grads_in_arena = true;  // no buffer given: the gradient lives in its slot
      if (!(es == 2 || es == 4) || (t.device < 0 && !r->dry)) {
        set_error("reducer_create: tensor %lld must be a float32/bfloat16/float16 device tensor", (long long)v);
        status = PRB_UNIMPLEMENTED;
        break;
      }
      r->locator[(size_t)v] = {gi, s};
      g.vars.push_back(v);
      g.slot_off.push_back(off);
      off += (int64_t)(align_up((size_t)t.numel * es, 256) / es);
    }
    if (status != PRB_SUCCESS) break;
    g.length = off;
    g.arena_off = arena;
    arena += align_up((size_t)off * dtype_size(g.dtype), 256);
    g.first_chunk = r->total_chunks;
    for (int64_t v : g.vars) r->total_chunks += (r->grads[(size_t)v].numel + grad_chunk_elems() - 1) / grad_chunk_elems();
    r->groups.push_back(std::move(g));
  }
  if (status == PRB_SUCCESS)
    for (int64_t v = 0; v < num; ++v)
      if (r->locator[(size_t)v].first < 0) {
        set_error("reducer_create: tensor %lld belongs to no group", (long long)v);
        status = PRB_INVALID_ARGUMENT;
        break;
      }
  if (status != PRB_SUCCESS) {
    delete r;
    return status;
  }
  // sidecar: [total_chunks partials | num per-tensor sums | 8 spare floats (clip scratch)] + 256 B of flags
  r->sidecar_off = arena;
  r->sidecar_floats = r->total_chunks + num + 8;
  arena += align_up((size_t)r->sidecar_floats * sizeof(float), 256) + 256;
  r->arena_bytes = arena;
  if (r->dry) {
    r->finalized = true;
    *out = r;
    return PRB_SUCCESS;
  }
  // Which allreduce runs the buckets of a multi-rank job (PRB_ALLREDUCE = nvls | nvlink | nccl):
  //   nvls              NVLink multicast: the arena is multicast-bound memory and the NVSwitch does the adds
  //                     (wins for buckets >= ~100 MiB at 8 ranks, loses below 8 ranks: profiles/r1_multi_gpu.md);
  //   nvlink (default)  two-shot kernel over CUDA-IPC peer mappings of a cudaMalloc'ed arena;
  //   nccl              ncclAllReduce.
  // Each level falls back to the next when the ranks agree (collectively) that it cannot be set up.
  const char* ar = getenv("PRB_ALLREDUCE");
  const int mode = ar == nullptr ? 1 : (strcmp(ar, "nvls") == 0 ? 2 : (strcmp(ar, "nvlink") == 0 ? 1 : 0));
  NvlCollectives nc{comm->nranks > 1 ? (void*)comm->comm : nullptr, nvl_all_gather_bytes, nvl_all_min_int};
  if (comm->nranks > 1 && mode >= 2) {
    void* a = nullptr;
    int32_t st = nvls_setup(nc, comm->nranks, comm->rank, comm->device, arena, comm->stream, &r->nvls, &a);
    if (st != PRB_SUCCESS) {
      prb_reducer_destroy(r);
      return st;
    }
    if (r->nvls != nullptr) r->arena = reinterpret_cast<char*>(a);  // zero-filled by nvls_setup
  }
  cudaError_t e = cudaSuccess;
  if (r->arena == nullptr) {
    e = cudaMalloc((void**)&r->arena, arena);
    if (e == cudaSuccess) e = cudaMemset(r->arena, 0, arena);  // padding gaps stay zero forever
  }
  for (Group& g : r->groups) {
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&g.ready, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&g.done, cudaEventDisableTiming);
  }
  if (e != cudaSuccess) {
    prb_reducer_destroy(r);
    PRB_FAIL(PRB_EXTERNAL, "reducer_create: CUDA error %s", cudaGetErrorString(e));
  }
  // Opt-in (PRB_NCCL_REGISTER=1): register the persistent arena with the communicator so NCCL can reduce
  // straight out of it.  Measured on 8 x B200 with NCCL 2.27.3 it is a large LOSS (8 MiB allreduce 2.7 ms
  // registered vs 72 us unregistered, profiles/r1_multi_gpu.md), so it stays off by default.
  if (comm->nranks > 1 && g_nccl.CommRegister != nullptr && getenv("PRB_NCCL_REGISTER") != nullptr) {
    if (g_nccl.CommRegister(comm->comm, r->arena, arena, &r->arena_reg) != ncclSuccess) r->arena_reg = nullptr;
  }
  if (comm->nranks > 1 && r->nvls == nullptr && mode >= 1) {
    int32_t st = nvl_setup(nc, comm->nranks, comm->rank, r->arena, arena, comm->stream, &r->nvl);
    if (st != PRB_SUCCESS) {
      prb_reducer_destroy(r);
      return st;
    }
  }
  r->finalized = true;  // nothing pending before the first prepare_for_backward
  *out = r;
  return PRB_SUCCESS;
}

int32_t prb_reducer_destroy(PRB_Reducer* r) {
  if (!r) return PRB_SUCCESS;
  if (!r->dry) cudaDeviceSynchronize();
  for (Group& g : r->groups)
    for (const DeviceTable* t : {&g.pack_tab, &g.unpack_tab, &g.packep_tab, &g.exch_tab}) release_table(*t);
  for (auto& w : r->waves)
    for (const DeviceTable* t : {&w.pack_tab, &w.unpack_tab, &w.packep_tab, &w.exch_tab}) release_table(*t);
  release_table(r->shard_tab);
  release_table(r->gather_tab);
  if (r->arena_reg && g_nccl.CommDeregister) g_nccl.CommDeregister(r->comm->comm, r->arena_reg);
  if (r->nvl) nvl_destroy(r->nvl);
  for (Group& g : r->groups) {
    if (g.ready) cudaEventDestroy(g.ready);
    if (g.done) cudaEventDestroy(g.done);
  }
  if (r->nvls) {
    nvls_destroy(r->nvls);  // unmaps and releases the arena it allocated
    r->arena = nullptr;
  }
  if (r->arena) cudaFree(r->arena);
  delete r;
  return PRB_SUCCESS;
}

int32_t prb_reducer_prepare_for_backward(PRB_Reducer* r) {
  PRB_ENFORCE(r, PRB_INVALID_ARGUMENT, "reducer: NULL handle");
  if (!r->dry) PRB_TRY(check_comm_alive(r));
  // ref: PrepareForBackward reducer.cc:669-728
  PRB_ENFORCE(!r->need_finalize, PRB_PRECONDITION_NOT_MET,
              "A serious error has occurred here. Please set find_unused_parameters=True to traverse backward graph "
              "in each step to prepare reduce in advance. (the previous backward did not mark every parameter ready)");
  r->grad_need_hooks = true;
  r->next_group = 0;
  for (Group& g : r->groups) {
    g.pending = (int64_t)g.vars.size();
    g.has_unused = false;
  }
  std::fill(r->ready.begin(), r->ready.end(), 0);
  std::fill(r->unused.begin(), r->unused.end(), 0);
  r->unused_list.clear();
  r->has_marked_unused = false;
  r->finalized = false;
  r->launched.clear();
  r->ep.cleared = false;
  r->ep.sidecar_cleared = false;
  return PRB_SUCCESS;
}

int32_t prb_reducer_set_unused_vars(PRB_Reducer* r, const int64_t* var_indices, int64_t num) {
  PRB_ENFORCE(r && (var_indices || num == 0), PRB_INVALID_ARGUMENT, "reducer: bad arguments");
  PRB_ENFORCE(r->find_unused || num == 0, PRB_PRECONDITION_NOT_MET,
              "unused parameters were reported but the reducer was built with find_unused_parameters=False");
  for (int64_t i = 0; i < num; ++i) {
    const int64_t v = var_indices[i];
    PRB_ENFORCE(v >= 0 && v < (int64_t)r->grads.size(), PRB_OUT_OF_RANGE, "Out of bounds variable index %lld", (long long)v);
    if (!r->unused[(size_t)v]) {
      r->unused[(size_t)v] = 1;
      r->unused_list.push_back(v);
      r->groups[(size_t)r->locator[(size_t)v].first].has_unused = true;
    }
  }
  return PRB_SUCCESS;
}

int32_t prb_reducer_mark_var_ready(PRB_Reducer* r, int64_t v, void* stream) {
  PRB_ENFORCE(r, PRB_INVALID_ARGUMENT, "reducer: NULL handle");
  // ref: AddDistHook reducer.cc:730-772
  PRB_ENFORCE(v >= 0 && v < (int64_t)r->locator.size(), PRB_OUT_OF_RANGE,
              "Out of bounds variable index. it must be less than %zu, but it is %lld", r->locator.size(), (long long)v);
  if (!r->grad_need_hooks) return PRB_SUCCESS;  // no_sync / already finalized: nothing to synchronise
  if (!r->has_marked_unused) {
    r->has_marked_unused = true;
    for (int64_t u : r->unused_list) PRB_TRY(mark_var_ready_impl(r, u, false, (cudaStream_t)stream));
  }
  return mark_var_ready_impl(r, v, true, (cudaStream_t)stream);
}

int32_t prb_reducer_mark_vars_ready(PRB_Reducer* r, const int64_t* var_indices, int64_t num, void* stream) {
  PRB_ENFORCE(r && (var_indices || num == 0), PRB_INVALID_ARGUMENT, "reducer: bad arguments");
  r->batching = !r->dry;
  int32_t st = PRB_SUCCESS;
  for (int64_t i = 0; i < num && st == PRB_SUCCESS; ++i) st = prb_reducer_mark_var_ready(r, var_indices[i], stream);
  r->batching = false;
  if (st == PRB_SUCCESS && !r->dry) st = flush_batch(r, (cudaStream_t)stream);
  r->batch.clear();
  return st;
}

int32_t prb_reducer_set_grads_in_arena(PRB_Reducer* r, int32_t enabled) {
  PRB_ENFORCE(r, PRB_INVALID_ARGUMENT, "reducer: NULL handle");
  PRB_ENFORCE(r->finalized, PRB_PRECONDITION_NOT_MET, "reducer: cannot change the gradient layout in the middle of a backward");
  if (!enabled)
    for (const PRB_Tensor& t : r->grads)
      PRB_ENFORCE(t.data != nullptr || r->dry, PRB_INVALID_ARGUMENT,
                  "reducer: this reducer was created without gradient buffers; its gradients can only live in the arena");
  r->grads_in_arena = enabled != 0;
  for (Group& g : r->groups) g.pack_valid = false;  // the table slot is shared by the pack and the in-place scale
  return PRB_SUCCESS;
}

int32_t prb_reducer_set_grad_epilogue(PRB_Reducer* r, const PRB_GradEpilogue* e, PRB_Tensor* partials_out,
                                      int32_t* active) {
  PRB_ENFORCE(r && !r->dry, PRB_INVALID_ARGUMENT, "reducer: bad handle");
  PRB_ENFORCE(r->finalized, PRB_PRECONDITION_NOT_MET, "reducer: cannot change the epilogue in the middle of a backward");
  Epilogue ep;
  if (e != nullptr) {
    if (e->loss_scale != nullptr) {
      PRB_ENFORCE(e->loss_scale->numel == 1 && e->loss_scale->dtype == PRB_FLOAT32 && e->loss_scale->device >= 0,
                  PRB_INVALID_ARGUMENT, "grad epilogue: loss_scale must be one float32 on the device");
      PRB_ENFORCE(e->found_inf && e->found_inf->numel == 1 && e->found_inf->dtype == PRB_BOOL && e->found_inf->device >= 0,
                  PRB_INVALID_ARGUMENT, "grad epilogue: loss_scale needs a bool device found_inf");
      ep.loss_scale = reinterpret_cast<const float*>(e->loss_scale->data);
      ep.found_inf = reinterpret_cast<bool*>(e->found_inf->data);
    }
    ep.norm = e->want_norm != 0;
    ep.mp = e->use_mp_type != 0;
    if (ep.norm) {
      PRB_ENFORCE(e->partial_base != nullptr, PRB_INVALID_ARGUMENT, "grad epilogue: want_norm needs partial_base");
      ep.partial_base.assign(e->partial_base, e->partial_base + r->grads.size());
      for (size_t v = 0; v < r->grads.size(); ++v) {
        const int64_t nch = (r->grads[v].numel + grad_chunk_elems() - 1) / grad_chunk_elems();
        PRB_ENFORCE(ep.partial_base[v] >= 0 && ep.partial_base[v] + nch <= r->total_chunks, PRB_OUT_OF_RANGE,
                    "grad epilogue: partial_base[%zu] = %d does not fit the %lld chunks of this reducer", v,
                    ep.partial_base[v], (long long)r->total_chunks);
      }
    }
    ep.on = ep.norm || ep.loss_scale != nullptr;
  }
  r->ep = ep;
  r->shard_valid = false;
  for (Group& g : r->groups) g.packep_valid = g.exch_valid = false;
  for (auto& w : r->waves) w.packep_valid = w.exch_valid = false;
  if (partials_out != nullptr) {
    partials_out->data = sidecar(r);
    partials_out->numel = r->sidecar_floats;
    partials_out->dtype = PRB_FLOAT32;
    partials_out->device = r->comm->device;
  }
  if (active != nullptr) *active = epilogue_active(r) ? 1 : 0;
  return PRB_SUCCESS;
}

int32_t prb_grad_chunk_elems(void) { return grad_chunk_elems(); }

int32_t prb_reducer_set_unpack(PRB_Reducer* r, int32_t enabled) {
  PRB_ENFORCE(r, PRB_INVALID_ARGUMENT, "reducer: NULL handle");
  r->unpack = enabled != 0;
  return PRB_SUCCESS;
}

int32_t prb_reducer_is_finalized(const PRB_Reducer* r) { return r && r->finalized ? 1 : 0; }

int32_t prb_reducer_allreduce_backend(const PRB_Reducer* r) {
  if (r == nullptr || r->dry || r->comm->nranks == 1) return 0;
  return r->nvls != nullptr ? 3 : (r->nvl != nullptr ? 2 : 1);
}

int32_t prb_reducer_allreduce_group(PRB_Reducer* r, int64_t gi, void* stream) {
  PRB_ENFORCE(r && !r->dry, PRB_INVALID_ARGUMENT, "reducer: bad handle");
  PRB_ENFORCE(gi >= 0 && gi < (int64_t)r->groups.size(), PRB_OUT_OF_RANGE, "Out of bounds group index %lld", (long long)gi);
  Group& g = r->groups[(size_t)gi];
  PRB_Comm* c = r->comm;
  if (c->nranks == 1) return PRB_SUCCESS;
  if (r->nvls != nullptr) return nvls_allreduce(r->nvls, g.arena_off, g.length, g.dtype, (cudaStream_t)stream);
  if (r->nvl != nullptr) {
    // stand-alone calls all go to the caller's one stream, which orders them whatever slot they use
    NvlWork w;
    w.region_off = g.arena_off;
    w.numel = g.length;
    w.dtype = g.dtype;
    return nvl_exchange(r->nvl, w, (cudaStream_t)stream);
  }
  ncclDataType_t nt;
  PRB_TRY(nccl_dtype(g.dtype, &nt));
  char* region = r->arena + g.arena_off;
  PRB_NCCL(g_nccl.AllReduce(region, region, (size_t)g.length, nt, ncclSum, c->comm, (cudaStream_t)stream));
  return PRB_SUCCESS;
}

int32_t prb_reducer_comm_error(PRB_Reducer* r, int32_t* out) {
  PRB_ENFORCE(r && out, PRB_INVALID_ARGUMENT, "reducer: bad arguments");
  *out = r->nvls != nullptr ? nvls_error_flag(r->nvls) : (r->nvl != nullptr ? nvl_error_flag(r->nvl) : 0);
  return PRB_SUCCESS;
}

int32_t prb_reducer_set_exchange_ctas(PRB_Reducer* r, int32_t max_ctas) {
  PRB_ENFORCE(r, PRB_INVALID_ARGUMENT, "reducer: NULL handle");
  if (r->nvl != nullptr) nvl_set_max_ctas(r->nvl, max_ctas);
  return PRB_SUCCESS;
}

int32_t prb_reducer_set_deferred_wait(PRB_Reducer* r, int32_t enabled) {
  PRB_ENFORCE(r, PRB_INVALID_ARGUMENT, "reducer: NULL handle");
  r->deferred_wait = enabled != 0;
  return PRB_SUCCESS;
}

int32_t prb_reducer_wait_group(PRB_Reducer* r, int64_t gi, void* stream) {
  PRB_ENFORCE(r, PRB_INVALID_ARGUMENT, "reducer: NULL handle");
  PRB_ENFORCE(gi >= 0 && gi < (int64_t)r->groups.size(), PRB_OUT_OF_RANGE, "Out of bounds group index %lld", (long long)gi);
  PRB_ENFORCE((size_t)gi < r->next_group, PRB_PRECONDITION_NOT_MET,
              "group %lld has not been launched in this backward yet", (long long)gi);
  if (!r->dry) PRB_TRY(check_comm_alive(r));
  if (!r->dry && r->comm->nranks > 1) PRB_CUDA(cudaStreamWaitEvent((cudaStream_t)stream, r->groups[(size_t)gi].done, 0));
  return PRB_SUCCESS;
}

int32_t prb_reducer_num_groups(const PRB_Reducer* r) { return r ? (int32_t)r->groups.size() : 0; }

int32_t prb_reducer_group_of(const PRB_Reducer* r, int64_t v, int64_t* out_group) {
  PRB_ENFORCE(r && out_group, PRB_INVALID_ARGUMENT, "reducer: bad arguments");
  PRB_ENFORCE(v >= 0 && v < (int64_t)r->locator.size(), PRB_OUT_OF_RANGE, "Out of bounds variable index %lld", (long long)v);
  *out_group = r->locator[(size_t)v].first;
  return PRB_SUCCESS;
}

int32_t prb_reducer_bucket_view(const PRB_Reducer* r, int64_t v, PRB_Tensor* out) {
  PRB_ENFORCE(r && out, PRB_INVALID_ARGUMENT, "reducer: bad arguments");
  PRB_ENFORCE(v >= 0 && v < (int64_t)r->locator.size(), PRB_OUT_OF_RANGE, "Out of bounds variable index %lld", (long long)v);
  const Group& g = r->groups[(size_t)r->locator[(size_t)v].first];
  const size_t s = (size_t)r->locator[(size_t)v].second;
  out->data = r->arena + g.arena_off + (size_t)g.slot_off[s] * dtype_size(g.dtype);
  out->numel = r->grads[(size_t)v].numel;
  out->dtype = g.dtype;
  out->device = r->grads[(size_t)v].device;
  return PRB_SUCCESS;
}

int32_t prb_reducer_launched_groups(const PRB_Reducer* r, int64_t* out_order, int64_t capacity, int64_t* out_num) {
  PRB_ENFORCE(r && out_num, PRB_INVALID_ARGUMENT, "reducer: bad arguments");
  *out_num = (int64_t)r->launched.size();
  for (int64_t i = 0; i < capacity && i < (int64_t)r->launched.size(); ++i) out_order[i] = r->launched[(size_t)i];
  return PRB_SUCCESS;
}

// ---------------------------------------------------------------- sharded step
int32_t prb_reducer_set_sharded(PRB_Reducer* r, int32_t enabled, int32_t* active) {
  PRB_ENFORCE(r && !r->dry, PRB_INVALID_ARGUMENT, "reducer: bad handle");
  PRB_ENFORCE(r->finalized, PRB_PRECONDITION_NOT_MET, "reducer: cannot change the exchange in the middle of a backward");
  if (enabled) {
    PRB_ENFORCE(!r->find_unused, PRB_UNIMPLEMENTED,
                "sharded step: find_unused_parameters is not supported (every rank must pack every bucket)");
    for (const Group& g : r->groups)
      PRB_ENFORCE(g.dtype == r->groups.front().dtype, PRB_UNIMPLEMENTED,
                  "sharded step: all gradients must share one dtype (one reduce-scatter covers every bucket)");
  }
  r->sharded = enabled != 0;
  r->shard_valid = false;
  r->batch_key.clear();  // the wave split differs between the two modes
  if (active != nullptr) *active = sharded_active(r) ? 1 : 0;
  return PRB_SUCCESS;
}

int32_t prb_reducer_set_shard_layout(PRB_Reducer* r, int32_t per_bucket) {
  PRB_ENFORCE(r, PRB_INVALID_ARGUMENT, "reducer: NULL handle");
  PRB_ENFORCE(r->finalized && !r->sharded, PRB_PRECONDITION_NOT_MET,
              "reducer: the ownership layout is fixed before the sharded step is switched on");
  PRB_ENFORCE(!per_bucket || (int)r->groups.size() <= kMaxShardSegments, PRB_UNIMPLEMENTED,
              "sharded step: a per-bucket layout supports at most %d buckets, this reducer has %zu", kMaxShardSegments,
              r->groups.size());
  r->shard_per_bucket = per_bucket != 0;
  r->gather_key.clear();
  return PRB_SUCCESS;
}

int32_t prb_reducer_shard_ranges(const PRB_Reducer* r, int32_t rank, int32_t nranks, int64_t* out_var, int64_t* out_offset,
                                 int64_t* out_length, int64_t capacity, int64_t* out_num) {
  PRB_ENFORCE(r && out_num, PRB_INVALID_ARGUMENT, "reducer: bad arguments");
  const int n = nranks > 0 ? nranks : (r->dry ? 1 : r->comm->nranks);
  PRB_ENFORCE(rank >= 0 && rank < n, PRB_OUT_OF_RANGE, "shard_ranges: rank %d of %d", rank, n);
  // The chunk list of the exchange kernels: groups in launch order, slots in order, 8192-element chunks.  One
  // ownership segment = the whole list, or one per group (prb_reducer_set_shard_layout); rank r owns the r-th n-th
  // of every segment: chunks [f + len * r / n, f + len * (r + 1) / n) -- the formula the kernels use.
  const int64_t ce = grad_chunk_elems();
  int64_t k = 0, c = 0;
  auto emit = [&](const Group& g, int64_t c0, int64_t c1, int64_t& cc) {
    for (int64_t v : g.vars) {
      const int64_t numel = r->grads[(size_t)v].numel, nch = (numel + ce - 1) / ce;
      const int64_t lo = std::max(cc, c0), hi = std::min(cc + nch, c1);
      if (lo < hi) {
        if (k < capacity) {
          out_var[k] = v;
          out_offset[k] = (lo - cc) * ce;
          out_length[k] = std::min(numel, (hi - cc) * ce) - (lo - cc) * ce;
        }
        ++k;
      }
      cc += nch;
    }
  };
  for (size_t gi = 0; gi < r->groups.size(); ++gi) {
    const Group& g = r->groups[gi];
    int64_t f = 0, len = r->total_chunks;
    if (r->shard_per_bucket) {
      f = g.first_chunk;
      len = (gi + 1 < r->groups.size() ? r->groups[gi + 1].first_chunk : r->total_chunks) - f;
    }
    emit(g, f + len * rank / n, f + len * (rank + 1) / n, c);
  }
  *out_num = k;
  return PRB_SUCCESS;
}

int32_t prb_reducer_allgather_params(PRB_Reducer* r, const PRB_Tensor* params, int64_t num, const PRB_Tensor* skip,
                                     void* stream) {
  PRB_ENFORCE(r && !r->dry && params, PRB_INVALID_ARGUMENT, "reducer: bad arguments");
  PRB_ENFORCE(num == (int64_t)r->grads.size(), PRB_INVALID_ARGUMENT,
              "allgather_params: expected one parameter per gradient (%zu), got %lld", r->grads.size(), (long long)num);
  if (r->comm->nranks == 1) return PRB_SUCCESS;
  PRB_ENFORCE(r->nvl != nullptr, PRB_UNIMPLEMENTED, "allgather_params: needs the NVLink peer-memory exchange (PRB_ALLREDUCE=nvlink)");
  PRB_TRY(check_comm_alive(r));
  if (skip != nullptr)
    PRB_ENFORCE(skip->numel == 1 && skip->dtype == PRB_BOOL && skip->device >= 0, PRB_INVALID_ARGUMENT,
                "allgather_params: skip must be one bool on the device");
  std::vector<void*> key((size_t)num);
  for (int64_t v = 0; v < num; ++v) key[(size_t)v] = params[v].data;
  if (key != r->gather_key) {
    std::vector<GradDesc> descs;
    std::vector<int64_t> numels;
    for (const Group& g : r->groups) {
      const int es = dtype_size(g.dtype);
      for (size_t s = 0; s < g.vars.size(); ++s) {
        const PRB_Tensor& p = params[(size_t)g.vars[s]];
        PRB_ENFORCE(p.dtype == g.dtype && p.numel == r->grads[(size_t)g.vars[s]].numel && p.device >= 0,
                    PRB_INVALID_ARGUMENT, "allgather_params: parameter %lld must have its gradient's dtype and size",
                    (long long)g.vars[s]);
        GradDesc d{};
        d.x = r->arena + g.arena_off + (size_t)g.slot_off[s] * es;
        d.out = p.data;
        d.aligned = ((uintptr_t)p.data % ((uintptr_t)es * kVec)) == 0 ? 1 : 0;
        descs.push_back(d);
        numels.push_back(p.numel);
      }
    }
    release_table(r->gather_tab);
    PRB_TRY(get_table(descs.data(), sizeof(GradDesc), (int)descs.size(), numels.data(), grad_chunk_elems(),
                      (cudaStream_t)stream, &r->gather_tab, /*pin=*/true));
    r->gather_key = key;
  }
  NvlGather w;
  w.table = &r->gather_tab;
  w.dtype = r->groups.front().dtype;
  if (r->shard_per_bucket) {
    w.nseg = (int)r->groups.size();
    for (size_t gi = 0; gi < r->groups.size(); ++gi) w.seg_first[gi] = (int)r->groups[gi].first_chunk;
    w.seg_first[w.nseg] = (int)r->total_chunks;
  } else {
    w.nseg = 1;
    w.seg_first[0] = 0;
    w.seg_first[1] = (int)r->total_chunks;
  }
  w.skip = skip ? reinterpret_cast<const bool*>(skip->data) : nullptr;
  return nvl_gather_params(r->nvl, w, (cudaStream_t)stream);
}

}  // extern "C"
