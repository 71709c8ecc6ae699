// band/c/c_api_sharded.cc -- ADDITIVE C entry points of include/band_c_api.h: the single-call multi-GPU batch
// (band::ShardedBufferProcessor; sits beside BandImageProcessorProcess, /root/reference/band/c/c_api_buffer.cc:198-214)
// and the host registration cache (band::cuda::HostRegistry).
#include <string>
#include <vector>

#include "band/buffer/host_registry.h"
#include "band/buffer/sharded_processor.h"
#include "band/c/c_api_internal.h"
#include "band/logger.h"
#include "bandcu.h"

using band::Buffer;
using band::LogSeverity;

struct BandShardedImageProcessor {
  std::unique_ptr<band::ShardedBufferProcessor> impl;
};

namespace {
BandStatus ShardedBatch(BandShardedImageProcessor* sharded, BandBuffer* const* buffers, BandTensor* const* tensors,
                        size_t count, bool async) {
  if (!sharded || !buffers || !tensors) {
    BAND_LOG(LogSeverity::kError, "BandImageProcessorProcessBatchSharded: null argument");
    return kBandErr;
  }
  std::vector<const Buffer*> in(count);
  std::vector<Buffer*> out(count);
  for (size_t i = 0; i < count; ++i) {
    if (!buffers[i] || !buffers[i]->impl || !tensors[i]) {
      BAND_LOG(LogSeverity::kError, "BandImageProcessorProcessBatchSharded: entry %d is null", static_cast<int>(i));
      return kBandErr;
    }
    in[i] = buffers[i]->impl.get();
    out[i] = tensors[i]->View();
    if (!out[i]) return kBandErr;
  }
  absl::Status status = async ? sharded->impl->ProcessBatchAsync(in.data(), out.data(), count)
                              : sharded->impl->ProcessBatch(in.data(), out.data(), count);
  if (!status.ok()) BandSetLastError(status.ToString().c_str());
  return status.ok() ? kBandOk : kBandErr;
}
}  // namespace

extern "C" {

BandShardedImageProcessor* BandImageProcessorBuilderBuildSharded(BandImageProcessorBuilder* builder, const int* devices,
                                                                 size_t num_devices) {
  if (!builder) {
    BAND_LOG(LogSeverity::kError, "BandImageProcessorBuilder is null");
    return nullptr;
  }
  std::vector<int> devs;
  if (devices && num_devices) {
    devs.assign(devices, devices + num_devices);
  } else {
    int n = 0;
    if (bandcu_device_count(&n) != BANDCU_OK) n = 0;
    for (int d = 0; d < n; ++d) devs.push_back(d);
  }
  absl::StatusOr<std::unique_ptr<band::ShardedBufferProcessor>> built =
      band::ShardedBufferProcessor::Create(*builder->impl, devs);
  if (!built.ok()) {
    BandSetLastError(built.status().ToString().c_str());
    return nullptr;
  }
  BandShardedImageProcessor* s = new BandShardedImageProcessor();
  s->impl = std::move(built).value();
  return s;
}

void BandShardedImageProcessorDelete(BandShardedImageProcessor* sharded) { delete sharded; }

size_t BandShardedImageProcessorGetNumShards(BandShardedImageProcessor* sharded) {
  return sharded ? sharded->impl->NumShards() : 0;
}

int BandShardedImageProcessorGetDevice(BandShardedImageProcessor* sharded, size_t shard) {
  return (sharded && shard < sharded->impl->NumShards()) ? sharded->impl->DeviceOf(shard) : -1;
}

void BandShardedImageProcessorGetSlice(BandShardedImageProcessor* sharded, size_t count, size_t shard, size_t* begin,
                                       size_t* end) {
  size_t b = 0, e = 0;
  if (sharded) band::ShardedBufferProcessor::SliceOf(count, sharded->impl->NumShards(), shard, &b, &e);
  if (begin) *begin = b;
  if (end) *end = e;
}

void BandShardSlice(size_t count, size_t shards, size_t shard, size_t* begin, size_t* end) {
  size_t b = 0, e = 0;
  if (shards > 0 && shard < shards) band::ShardedBufferProcessor::SliceOf(count, shards, shard, &b, &e);
  if (begin) *begin = b;
  if (end) *end = e;
}

BandStatus BandImageProcessorProcessBatchSharded(BandShardedImageProcessor* sharded, BandBuffer* const* buffers,
                                                 BandTensor* const* target_tensors, size_t count) {
  return ShardedBatch(sharded, buffers, target_tensors, count, false);
}

BandStatus BandImageProcessorProcessBatchShardedAsync(BandShardedImageProcessor* sharded, BandBuffer* const* buffers,
                                                      BandTensor* const* target_tensors, size_t count) {
  return ShardedBatch(sharded, buffers, target_tensors, count, true);
}

BandStatus BandShardedImageProcessorSynchronize(BandShardedImageProcessor* sharded) {
  if (!sharded) return kBandErr;
  absl::Status status = sharded->impl->Synchronize();
  if (!status.ok()) BandSetLastError(status.ToString().c_str());
  return status.ok() ? kBandOk : kBandErr;
}

void* BandShardedImageProcessorGetStream(BandShardedImageProcessor* sharded, size_t shard) {
  return sharded ? sharded->impl->GetComputeStream(shard) : nullptr;
}

int BandShardedImageProcessorIsFused(BandShardedImageProcessor* sharded, size_t shard) {
  return sharded && sharded->impl->IsSingleFastLaunch(shard) ? 1 : 0;
}

BandStatus BandHostRegistrationCacheConfigure(int enable, size_t max_bytes) {
  band::cuda::HostRegistry::Get().Enabled();  // seed from the environment first, then override
  band::cuda::HostRegistry::Get().Configure(enable != 0, max_bytes);
  return kBandOk;
}

void BandHostRegistrationCacheStats(size_t* registered_bytes, size_t* entries, unsigned long long* hits,
                                    unsigned long long* registrations) {
  band::cuda::HostRegistry& r = band::cuda::HostRegistry::Get();
  if (registered_bytes) *registered_bytes = r.registered_bytes();
  if (entries) *entries = r.num_entries();
  if (hits) *hits = r.hits();
  if (registrations) *registrations = r.registrations();
}

}  // extern "C"
