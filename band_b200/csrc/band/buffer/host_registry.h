// band/buffer/host_registry.h -- opt-in registration cache for caller-owned (pageable) host frames.
//
// The reference API takes arbitrary user memory (/root/reference/band/c/c_api_buffer.cc:40-53,
// band/buffer/buffer.cc:24-78). The row-staged kernel can read host frames IN PLACE over PCIe (only the rows the
// resize samples cross the bus), but only when the pages are locked and mapped into the device address space.
// Frames from BandHostAlloc are; an unmodified caller's malloc'ed camera ring is not and falls to the HBM staging
// pipeline. With the cache on (BandHostRegistrationCacheConfigure / BAND_REGISTER_HOST=1) a frame's address range
// is cudaHostRegister'ed on first sight and found again on every later call (camera rings reuse a handful of
// buffers), bounded by a byte budget with least-recently-used eviction, and everything is unregistered when the last
// processor that used the cache goes away.
#ifndef BAND_BUFFER_HOST_REGISTRY_H_
#define BAND_BUFFER_HOST_REGISTRY_H_

#include <cstddef>
#include <cstdint>
#include <map>
#include <mutex>
#include <vector>

namespace band {
namespace cuda {

class HostRegistry {
 public:
  struct Entry {
    uintptr_t lo = 0, hi = 0;  // page-aligned [lo, hi)
    int busy = 0;              // passes in flight that read this range
    uint64_t last_use = 0;
  };
  static HostRegistry& Get();

  // process-wide switch and budget; the first call to Enabled() seeds them from BAND_REGISTER_HOST / BAND_REGISTER_HOST_MB
  void Configure(bool enable, size_t max_bytes);
  bool Enabled();

  // Makes [p, p + bytes) readable by kernels in place. On success the covering entries are appended to `held` and stay
  // registered until Release(held). false: not enabled, the platform cannot use registered host pointers directly, the
  // budget is exhausted by ranges in use, or cudaHostRegister refused -- the caller stages the frame instead.
  // existing_only: succeed only when the whole range is already registered here (a pure lookup, nothing is pinned).
  bool Acquire(const void* p, size_t bytes, std::vector<Entry*>* held, bool existing_only = false);
  void Release(std::vector<Entry*>* held);

  // processors that were created while the cache was on; the last one out unregisters everything
  void AddUser();
  void RemoveUser();

  size_t registered_bytes();
  size_t num_entries();
  uint64_t hits() const { return hits_; }
  uint64_t registrations() const { return registrations_; }

 private:
  HostRegistry() = default;
  bool EvictFor(size_t bytes);  // lock held
  void ClearIdle();             // lock held

  std::mutex mu_;
  std::map<uintptr_t, Entry*> entries_;  // by lo; ranges never overlap
  size_t total_ = 0;
  size_t budget_ = size_t(4) << 30;
  int enabled_ = -1;  // -1: not yet read from the environment
  int usable_ = -1;   // device can dereference registered host pointers as they are
  int users_ = 0;
  uint64_t clock_ = 0, hits_ = 0, registrations_ = 0;
};

}  // namespace cuda
}  // namespace band
#endif  // BAND_BUFFER_HOST_REGISTRY_H_
