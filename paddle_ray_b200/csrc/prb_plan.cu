#include <cstring>
#include <list>
#include <mutex>
#include <unordered_map>
#include <vector>

#include "prb_plan.h"

namespace prb {
namespace {

struct Entry {
  uint64_t key;
  int device;
  std::vector<uint8_t> image;  // host image, compared on hit to rule out hash collisions
  void* dev_blob = nullptr;
  DeviceTable table;
  bool pinned = false;
};

std::mutex g_mu;
std::list<Entry> g_lru;  // front = most recent
constexpr size_t kMaxPlans = 64;
constexpr size_t kGrace = 16;
std::list<void*> g_parked;  // evicted device tables awaiting their free

uint64_t fnv1a(const uint8_t* p, size_t n, uint64_t h = 1469598103934665603ull) {
  for (size_t i = 0; i < n; ++i) {
    h ^= p[i];
    h *= 1099511628211ull;
  }
  return h;
}

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

}  // namespace

int32_t get_table(const void* host_descs, size_t desc_bytes, int ndescs, const int64_t* numels, int chunk_elems,
                  cudaStream_t stream, DeviceTable* out, bool pin) {
  int dev = 0;
  PRB_CUDA(cudaGetDevice(&dev));
  // host image = [chunk_elems][desc_bytes][descs...][numels...]
  std::vector<uint8_t> image(16 + (size_t)ndescs * desc_bytes + (size_t)ndescs * 8);
  int64_t hdr[2] = {chunk_elems, (int64_t)desc_bytes};
  memcpy(image.data(), hdr, 16);
  memcpy(image.data() + 16, host_descs, (size_t)ndescs * desc_bytes);
  memcpy(image.data() + 16 + (size_t)ndescs * desc_bytes, numels, (size_t)ndescs * 8);
  const uint64_t key = fnv1a(image.data(), image.size()) ^ ((uint64_t)dev << 56);

  std::lock_guard<std::mutex> lk(g_mu);
  for (auto it = g_lru.begin(); it != g_lru.end(); ++it) {
    if (it->key == key && it->device == dev && it->image == image) {
      g_lru.splice(g_lru.begin(), g_lru, it);
      g_lru.front().pinned |= pin;
      *out = g_lru.front().table;
      return PRB_SUCCESS;
    }
  }

  // build chunk table
  std::vector<Chunk> chunks;
  std::vector<int32_t> first(ndescs + 1, 0);
  int64_t total = 0;
  for (int t = 0; t < ndescs; ++t) {
    first[t] = (int32_t)chunks.size();
    const int64_t n = numels[t];
    total += n;
    for (int64_t off = 0; off < n; off += chunk_elems) {
      Chunk c;
      c.tensor = t;
      c.n = (int32_t)((n - off) < chunk_elems ? (n - off) : chunk_elems);
      c.offset = off;
      chunks.push_back(c);
    }
    PRB_ENFORCE(chunks.size() < (size_t)INT32_MAX, PRB_OUT_OF_RANGE, "too many chunks in one multi-tensor plan");
  }
  first[ndescs] = (int32_t)chunks.size();

  const size_t descs_off = 0;
  const size_t chunks_off = align_up((size_t)ndescs * desc_bytes, 256);
  const size_t first_off = chunks_off + align_up(chunks.size() * sizeof(Chunk), 256);
  const size_t blob_bytes = first_off + align_up(first.size() * 4, 256);
  std::vector<uint8_t> blob(blob_bytes, 0);
  memcpy(blob.data() + descs_off, host_descs, (size_t)ndescs * desc_bytes);
  if (!chunks.empty()) memcpy(blob.data() + chunks_off, chunks.data(), chunks.size() * sizeof(Chunk));
  memcpy(blob.data() + first_off, first.data(), first.size() * 4);

  // The LRU bounds the UNPINNED plans only: pinned ones (a reducer's pack / unpack / exchange tables) live as long as
  // their owner says.  (Counting them used to let a few reducers fill the cache, after which every step of an
  // unrelated optimizer rebuilt -- and synchronously uploaded -- its tables: nested benchmark configs ran 2x slower.)
  size_t unpinned = 0;
  for (const Entry& x : g_lru) unpinned += x.pinned ? 0 : 1;
  if (unpinned >= kMaxPlans) {
    for (auto it = g_lru.end(); it != g_lru.begin();) {  // least recently used entry that nobody holds on to
      --it;
      if (it->pinned) continue;
      // The evicted table is not freed at once: another host thread may have fetched it a moment ago and be about
      // to launch with it.  It is parked and freed kGrace evictions later (cudaFree synchronises the device, so no
      // kernel still reads it then).
      g_parked.push_back(it->dev_blob);
      g_lru.erase(it);
      if (g_parked.size() > kGrace) {
        cudaFree(g_parked.front());
        g_parked.pop_front();
      }
      break;
    }
  }
  Entry e;
  e.key = key;
  e.device = dev;
  e.pinned = pin;
  e.image.swap(image);
  PRB_CUDA(cudaMalloc(&e.dev_blob, blob_bytes));
  // A plan is built once and may then be used from any stream (calc or comm), so the upload must be
  // COMPLETE before this call returns: enqueue on the caller's stream, then wait for that stream.
  // (A plain cudaMemcpy from pageable memory may return while its DMA is still in flight on the legacy
  // stream, which non-blocking streams do not wait for.)  Plan creation is a warm-up event, never per step.
  PRB_CUDA(cudaMemcpyAsync(e.dev_blob, blob.data(), blob_bytes, cudaMemcpyHostToDevice, stream));
  PRB_CUDA(cudaStreamSynchronize(stream));
  e.table.descs = (const uint8_t*)e.dev_blob + descs_off;
  e.table.chunks = (const Chunk*)((const uint8_t*)e.dev_blob + chunks_off);
  e.table.first_chunk = (const int32_t*)((const uint8_t*)e.dev_blob + first_off);
  e.table.nchunks = (int32_t)chunks.size();
  e.table.ndescs = ndescs;
  e.table.total_elems = total;
  g_lru.push_front(std::move(e));
  *out = g_lru.front().table;
  return PRB_SUCCESS;
}

void release_table(const DeviceTable& t) {
  if (t.chunks == nullptr && t.descs == nullptr) return;
  std::lock_guard<std::mutex> lk(g_mu);
  for (Entry& e : g_lru)
    if (e.table.descs == t.descs && e.table.chunks == t.chunks) e.pinned = false;  // evictable from now on
}

}  // namespace prb
