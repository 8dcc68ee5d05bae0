// Two-shot bucket exchange over NVLink peer memory for the reducer's bucket arena (see prb_nvlink.cu).
#pragma once
#include "prb_plan.h"

namespace prb {

struct NvlContext;

// The two tiny collectives the set-up needs, provided by the owner of the NCCL communicator.
struct NvlCollectives {
  void* comm;
  int32_t (*all_gather_bytes)(void* comm, const void* send_dev, void* recv_dev, size_t bytes_per_rank, cudaStream_t s);
  int32_t (*all_min_int)(void* comm, int value, int* out, cudaStream_t s);
};

// One exchange: the in-place sum over ranks of arena[region_off .. region_off + numel * sizeof(T)) on every rank.
struct NvlWork {
  size_t region_off = 0;
  int64_t numel = 0;
  int32_t dtype = 0;
  // gradients were produced in the arena unscaled: every rank's value goes through T(scale * (x + 0)) before it is
  // added (the reducer's scale_(1/nranks), ref reducer.cc:1049-1050, same rounding as ScaleFunctor)
  bool fold_scale = false;
  float scale = 1.0f;
  // Chunk mode: walk a GradDesc chunk table of the region's tensors (x = the tensor's slot in THIS rank's arena)
  // instead of the flat region, so that the slice owner can leave the inf/nan flag and the squared-norm partial
  // of every 8192-element chunk of the REDUCED gradient in every rank's sidecar (producer-side pass 1).
  const DeviceTable* table = nullptr;
  const float* loss_scale = nullptr;  // device; nullptr = no unscale / finite check
  bool* found_local = nullptr;        // this rank's found_inf (set from the sidecar flag when the exchange ends)
  bool norm = false, mp = false;
  size_t partials_off = 0, flags_off = 0;  // byte offsets of the sidecar inside the arena
  // Reduce-scatter only: the slice owner keeps the reduced values (its own arena only; flags and partials still go
  // to every rank).  The first half of a sharded step; nvl_gather_params is the second.
  bool scatter_only = false;
  bool overlap_grid = false;  // scatter_only while a backward is running: keep the allreduce's grid cap
};

constexpr int kMaxShardSegments = 64;

// Second half of a sharded step: every rank PULLS the chunks its peers own (the same n-th of the chunk table as in
// the reduce-scatter) from the owner's arena -- where the owner's update pass left the new parameter values in
// place of the gradients -- into desc.out, its own copy of that parameter.  Nothing moves when *skip is set
// (found_inf: no rank updated anything, the arena still holds gradients).
struct NvlGather {
  const DeviceTable* table = nullptr;  // GradDesc: x = the tensor's slot in THIS rank's arena, out = the parameter
  int32_t dtype = 0;
  const bool* skip = nullptr;          // device flag, equal on all ranks
  // Ownership: rank q owns the q-th n-th of every segment [seg_first[s], seg_first[s + 1]) of the chunk list
  // (one segment = the whole table, or one segment per bucket: prb_reducer_set_shard_layout)
  int nseg = 1;
  int seg_first[kMaxShardSegments + 1] = {0};
};

// Collective over all ranks.  On success *out is a context, or NULL when the ranks agreed that peer
// mapping is not available (the caller then keeps NCCL).  A local failure never makes a rank leave early.
int32_t nvl_setup(const NvlCollectives& nc, int nranks, int rank, void* arena, size_t arena_bytes, cudaStream_t stream,
                  NvlContext** out);
// Up to two calls may be in flight at once, PROVIDED call number k runs on stream (k & 1) of a fixed pair of
// streams: nvl_next_slot tells which of the two the next call must use.
int nvl_next_slot(const NvlContext* c);
int32_t nvl_exchange(NvlContext* c, const NvlWork& w, cudaStream_t stream);
int32_t nvl_gather_params(NvlContext* c, const NvlGather& w, cudaStream_t stream);
// Cap on the grid of the exchange kernels (0 = one CTA per SM).
void nvl_set_max_ctas(NvlContext* c, int max_ctas);
// 1 if an exchange of this context ran into its time limit or a peer aborted (read from a mapped host word: no sync).
int nvl_error_flag(const NvlContext* c);
void nvl_destroy(NvlContext* c);

}  // namespace prb
