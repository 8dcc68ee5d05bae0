// Two-shot allreduce over NVLink peer memory for the reducer's bucket arena (see prb_nvlink.cu).
#pragma once
#include "prb_common.cuh"

namespace prb {

struct NvlContext;

// The two tiny collectives the set-up needs, provided by the owner of the NCCL communicator.
struct NvlCollectives {
  void* comm;
  int32_t (*all_gather_bytes)(void* comm, const void* send_dev, void* recv_dev, size_t bytes_per_rank, cudaStream_t s);
  int32_t (*all_min_int)(void* comm, int value, int* out, cudaStream_t s);
};

// Collective over all ranks.  On success *out is a context, or NULL when the ranks agreed that peer
// mapping is not available (the caller then keeps NCCL).
int32_t nvl_setup(const NvlCollectives& nc, int nranks, int rank, void* arena, size_t arena_bytes, cudaStream_t stream,
                  NvlContext** out);
// Up to two calls may be in flight at once, PROVIDED call number k runs on stream (k & 1) of a fixed pair of
// streams: nvl_next_slot tells which of the two the next call must use.
int nvl_next_slot(const NvlContext* c);
// In-place sum over ranks of arena[region_off .. region_off + numel * sizeof(T)) on every rank.
int32_t nvl_allreduce(NvlContext* c, size_t region_off_bytes, int64_t numel, int32_t dtype, cudaStream_t stream);
// 1 if any wait inside a kernel of this context hit its time limit (a peer never arrived).
int32_t nvl_error_flag(NvlContext* c, int* out);
void nvl_destroy(NvlContext* c);

}  // namespace prb
