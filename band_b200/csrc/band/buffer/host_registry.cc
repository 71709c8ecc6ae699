// band/buffer/host_registry.cc -- see host_registry.h.
#include "band/buffer/host_registry.h"

#include <cstdlib>

#include "bandcu.h"

namespace band {
namespace cuda {
namespace {
constexpr uintptr_t kPage = 4096;
}

HostRegistry& HostRegistry::Get() {
  static HostRegistry* r = new HostRegistry();  // never destroyed: registrations die with the process
  return *r;
}

void HostRegistry::Configure(bool enable, size_t max_bytes) {
  std::lock_guard<std::mutex> l(mu_);
  enabled_ = enable ? 1 : 0;
  if (max_bytes) budget_ = max_bytes;
  if (!enable) ClearIdle();
}

bool HostRegistry::Enabled() {
  std::lock_guard<std::mutex> l(mu_);
  if (enabled_ < 0) {
    const char* e = getenv("BAND_REGISTER_HOST");
    enabled_ = (e && e[0] == '1') ? 1 : 0;
    if (const char* mb = getenv("BAND_REGISTER_HOST_MB"))
      if (atoll(mb) > 0) budget_ = static_cast<size_t>(atoll(mb)) << 20;
  }
  return enabled_ == 1;
}

bool HostRegistry::EvictFor(size_t bytes) {
  while (total_ + bytes > budget_) {
    Entry* victim = nullptr;
    for (auto& kv : entries_)
      if (kv.second->busy == 0 && (!victim || kv.second->last_use < victim->last_use)) victim = kv.second;
    if (!victim) return false;
    bandcu_host_unregister(reinterpret_cast<void*>(victim->lo));
    total_ -= victim->hi - victim->lo;
    entries_.erase(victim->lo);
    delete victim;
  }
  return true;
}

void HostRegistry::ClearIdle() {
  for (auto it = entries_.begin(); it != entries_.end();) {
    Entry* e = it->second;
    if (e->busy == 0) {
      bandcu_host_unregister(reinterpret_cast<void*>(e->lo));
      total_ -= e->hi - e->lo;
      it = entries_.erase(it);
      delete e;
    } else {
      ++it;
    }
  }
}

bool HostRegistry::Acquire(const void* p, size_t bytes, std::vector<Entry*>* held, bool existing_only) {
  if (!p || bytes == 0 || !held) return false;
  std::lock_guard<std::mutex> l(mu_);
  if (enabled_ != 1) return false;
  if (usable_ < 0) usable_ = bandcu_registered_host_pointers_usable() == 1 ? 1 : 0;
  if (!usable_) return false;
  const uintptr_t a = reinterpret_cast<uintptr_t>(p) & ~(kPage - 1);
  const uintptr_t b = (reinterpret_cast<uintptr_t>(p) + bytes + kPage - 1) & ~(kPage - 1);
  const size_t first_held = held->size();
  auto undo = [&]() {
    for (size_t i = first_held; i < held->size(); ++i) --(*held)[i]->busy;
    held->resize(first_held);
    return false;
  };
  ++clock_;
  uintptr_t cur = a;
  bool registered_new = false;
  while (cur < b) {
    // entry covering `cur`, if any
    auto it = entries_.upper_bound(cur);
    Entry* cover = nullptr;
    if (it != entries_.begin()) {
      auto prev = std::prev(it);
      if (prev->second->hi > cur) cover = prev->second;
    }
    if (cover) {
      ++cover->busy;
      cover->last_use = clock_;
      held->push_back(cover);
      cur = cover->hi;
      continue;
    }
    if (existing_only) return undo();
    // gap [cur, next entry or b): pin it first, so eviction below can never pick what this call already holds
    const uintptr_t gap_hi = (it != entries_.end() && it->first < b) ? it->first : b;
    const size_t gap = gap_hi - cur;
    if (!EvictFor(gap)) return undo();
    if (bandcu_host_register(reinterpret_cast<void*>(cur), gap) != BANDCU_OK) return undo();
    Entry* e = new Entry();
    e->lo = cur;
    e->hi = gap_hi;
    e->busy = 1;
    e->last_use = clock_;
    entries_[cur] = e;
    total_ += gap;
    ++registrations_;
    registered_new = true;
    held->push_back(e);
    cur = gap_hi;
  }
  if (!registered_new) ++hits_;
  return true;
}

void HostRegistry::Release(std::vector<Entry*>* held) {
  if (!held || held->empty()) return;
  std::lock_guard<std::mutex> l(mu_);
  for (Entry* e : *held) --e->busy;
  held->clear();
  if (enabled_ == 0 || users_ == 0) ClearIdle();
}

void HostRegistry::AddUser() {
  std::lock_guard<std::mutex> l(mu_);
  ++users_;
}

void HostRegistry::RemoveUser() {
  std::lock_guard<std::mutex> l(mu_);
  if (--users_ <= 0) {
    users_ = 0;
    ClearIdle();
  }
}

size_t HostRegistry::registered_bytes() {
  std::lock_guard<std::mutex> l(mu_);
  return total_;
}
size_t HostRegistry::num_entries() {
  std::lock_guard<std::mutex> l(mu_);
  return entries_.size();
}

}  // namespace cuda
}  // namespace band
