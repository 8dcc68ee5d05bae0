// Multi-tensor map (copy / scale) used by scale, concat, split and the reducer's pack/unpack.
#pragma once
#include "prb_plan.h"

namespace prb {

struct alignas(16) MapDesc {
  const void* src;
  void* dst;
  int32_t aligned;  // 1: src and dst both allow kVec-wide vector access
  int32_t pad0_;
  int64_t pad1_;
};
static_assert(sizeof(MapDesc) == 32, "MapDesc layout");

// Elements per CTA of the active launch geometry.
int map_chunk_elems();

bool map_aligned(const void* src, const void* dst, int elem_size);
// Pure bit copy of 2- or 4-byte elements.
int32_t launch_copy(int elem_size, const DeviceTable& tab, cudaStream_t stream);
// out = T(scale * (MT(x) + bias))  /  T(scale * MT(x) + bias)
int32_t launch_scale(int32_t dtype, const DeviceTable& tab, double scale, float bias, bool bias_after_scale,
                     cudaStream_t stream);

}  // namespace prb
