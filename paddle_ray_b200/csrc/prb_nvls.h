// Bucket allreduce through NVLink multicast (NVLS): the NVSwitch adds the ranks' copies of a vector while it
// is being loaded (multimem.ld_reduce) and replicates a store to every rank (multimem.st).  See prb_nvls.cu.
#pragma once
#include "prb_nvlink.h"

namespace prb {

struct NvlsContext;

// Collective over all ranks.  Allocates the bucket arena (`bytes`, zero-filled) as multicast-bound memory and
// returns it in *arena_out.  On success *out is a context; when the ranks agree that multicast is not
// available *out and *arena_out stay NULL and the caller allocates the arena itself.
int32_t nvls_setup(const NvlCollectives& nc, int nranks, int rank, int device, size_t bytes, cudaStream_t stream,
                   NvlsContext** out, void** arena_out);
// In-place sum over ranks of arena[region_off .. region_off + numel * sizeof(T)) on every rank.
// Same two-stream / two-slot rule as nvl_allreduce.
int nvls_next_slot(const NvlsContext* c);
int32_t nvls_allreduce(NvlsContext* c, size_t region_off_bytes, int64_t numel, int32_t dtype, cudaStream_t stream);
int nvls_error_flag(const NvlsContext* c);
void nvls_destroy(NvlsContext* c);

}  // namespace prb
