// Multi-tensor plan cache: descriptor table + chunk table resident on the device.
//
// A "plan" is what one multi-tensor launch walks: `ndescs` fixed-size descriptors (pointers and
// per-tensor attributes) and one 16-byte Chunk per (tensor, chunk_elems-sized piece).  Plans are
// cached per device, keyed by a hash of their host image, so a training loop that calls the same
// operator on the same buffers every step uploads its tables once (the reference re-uploads its
// pointer tables on every call: paddle/phi/kernels/gpu/amp_kernel.cu:287-335,
// paddle/phi/kernels/funcs/multi_tensor_apply.h:110-175).
#pragma once
#include "prb_common.cuh"

namespace prb {

struct DeviceTable {
  const void* descs = nullptr;    // device
  const Chunk* chunks = nullptr;  // device
  int32_t nchunks = 0;
  int32_t ndescs = 0;
  int64_t total_elems = 0;
  const int32_t* first_chunk = nullptr;  // device, [ndescs+1]: chunk range of each tensor
};

// host_descs: ndescs * desc_bytes bytes.  numels[i]: elements of tensor i (0 allowed -> no chunks).
// pin: the caller keeps the returned DeviceTable beyond this call (the reducer caches its pack/unpack plans),
// so the entry is exempt from LRU eviction for the life of the process.
int32_t get_table(const void* host_descs, size_t desc_bytes, int ndescs, const int64_t* numels, int chunk_elems,
                  cudaStream_t stream, DeviceTable* out, bool pin = false);

// The owner of a pinned plan no longer needs it: it becomes an ordinary LRU entry.
void release_table(const DeviceTable& t);

}  // namespace prb
