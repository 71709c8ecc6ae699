// band/buffer/sharded_processor.cc -- see sharded_processor.h.
#include "band/buffer/sharded_processor.h"

#include "absl/strings/str_format.h"
#include "band/buffer/buffer.h"
#include "band/logger.h"
#include "bandcu.h"

namespace band {

void ShardedBufferProcessor::SliceOf(size_t n, size_t shards, size_t shard, size_t* begin, size_t* end) {
  const size_t per = shards ? (n + shards - 1) / shards : n;
  const size_t b = std::min(n, shard * per);
  *begin = b;
  *end = std::min(n, b + per);
}

absl::StatusOr<std::unique_ptr<ShardedBufferProcessor>> ShardedBufferProcessor::Create(
    IBufferProcessorBuilder& builder, const std::vector<int>& devices) {
  if (devices.empty()) return absl::InvalidArgumentError("ShardedBufferProcessor: no devices");
  int count = 0;
  if (bandcu_device_count(&count) != BANDCU_OK || count <= 0)
    return absl::InternalError(absl::StrFormat("ShardedBufferProcessor: no CUDA device (%s)", bandcu_last_error_string()));
  for (int d : devices)
    if (d < 0 || d >= count)
      return absl::InvalidArgumentError(absl::StrFormat("ShardedBufferProcessor: device %d of %d", d, count));
  std::unique_ptr<ShardedBufferProcessor> sharded(new ShardedBufferProcessor());
  for (int d : devices) {
    absl::StatusOr<std::unique_ptr<BufferProcessor>> built = builder.Build();
    if (!built.ok()) return built.status();
    std::unique_ptr<Shard> s(new Shard());
    s->device = d;
    s->processor = std::move(built).value();
    sharded->shards_.push_back(std::move(s));
  }
  for (auto& s : sharded->shards_) s->worker = std::thread(WorkerLoop, s.get());
  return sharded;
}

ShardedBufferProcessor::~ShardedBufferProcessor() {
  for (auto& s : shards_) {
    Task t;
    t.kind = Task::kStop;
    Enqueue(s.get(), std::move(t));
  }
  for (auto& s : shards_)
    if (s->worker.joinable()) s->worker.join();
}

// One host thread per shard: binds its device once (cudaSetDevice is per thread), then runs its queue in order.
// The processor -- plans, streams, staging memory -- is created, used and destroyed on this thread's device.
void ShardedBufferProcessor::WorkerLoop(Shard* s) {
  const bool bound = bandcu_set_device(s->device) == BANDCU_OK;
  for (;;) {
    Task t;
    {
      std::unique_lock<std::mutex> l(s->mu);
      s->cv.wait(l, [&] { return !s->queue.empty(); });
      t = std::move(s->queue.front());
      s->queue.pop_front();
    }
    absl::Status st = absl::OkStatus();
    if (!bound) {
      st = absl::InternalError(absl::StrFormat("cudaSetDevice(%d) failed: %s", s->device, bandcu_last_error_string()));
    } else if (t.kind == Task::kRun) {
      st = s->processor->ProcessBatchAsync(t.in.data(), t.out.data(), t.in.size());
    } else if (t.kind == Task::kSync) {
      st = s->processor->Synchronize();
    } else {
      (void)s->processor->Synchronize();
      s->processor.reset();  // frees device memory while this thread's device is current
    }
    {
      std::lock_guard<std::mutex> l(s->mu);
      if (!st.ok() && s->status.ok()) s->status = st;
      s->completed = t.ticket;
    }
    s->done_cv.notify_all();
    if (t.kind == Task::kStop) return;
  }
}

uint64_t ShardedBufferProcessor::Enqueue(Shard* s, Task&& t) {
  uint64_t ticket;
  {
    std::lock_guard<std::mutex> l(s->mu);
    ticket = t.ticket = ++s->issued;
    s->queue.push_back(std::move(t));
  }
  s->cv.notify_one();
  return ticket;
}

absl::Status ShardedBufferProcessor::Wait(Shard* s, uint64_t ticket) {
  std::unique_lock<std::mutex> l(s->mu);
  s->done_cv.wait(l, [&] { return s->completed >= ticket; });
  absl::Status st = s->status;
  s->status = absl::OkStatus();
  return st;
}

absl::Status ShardedBufferProcessor::ProcessBatchAsync(const Buffer* const* inputs, Buffer* const* outputs, size_t n) {
  if (n == 0) return absl::OkStatus();
  if (inputs == nullptr || outputs == nullptr) return absl::InvalidArgumentError("ProcessBatchSharded: null buffer list");
  for (size_t i = 0; i < n; ++i)
    if (inputs[i] == nullptr || outputs[i] == nullptr)
      return absl::InvalidArgumentError(absl::StrFormat("ProcessBatchSharded: entry %d is null", static_cast<int>(i)));
  // Device-resident frames and outputs are used where they are: a slice that lives on another GPU than its shard
  // would fault inside the kernel (and take the CUDA context with it). The ends of every slice are checked here -- two
  // driver queries per shard, not per frame.
  for (size_t k = 0; k < shards_.size(); ++k) {
    size_t b, e;
    SliceOf(n, shards_.size(), k, &b, &e);
    if (b >= e) continue;
    for (size_t i : {b, e - 1}) {
      for (const Buffer* buf : {inputs[i], static_cast<const Buffer*>(outputs[i])}) {
        if (buf->GetMemorySpace() != MemorySpace::kDevice || buf->GetNumPlanes() == 0) continue;
        const int where = bandcu_pointer_device((*buf)[0].data);
        if (where >= 0 && where != shards_[k]->device)
          return absl::InvalidArgumentError(absl::StrFormat(
              "ProcessBatchSharded: entry %d lives on device %d but belongs to shard %d on device %d (slices are "
              "contiguous: see SliceOf / DeviceOf)", static_cast<int>(i), where, static_cast<int>(k), shards_[k]->device));
      }
    }
  }
  for (size_t k = 0; k < shards_.size(); ++k) {
    size_t b, e;
    SliceOf(n, shards_.size(), k, &b, &e);
    if (b >= e) continue;
    Task t;
    t.kind = Task::kRun;
    t.in.assign(inputs + b, inputs + e);
    t.out.assign(outputs + b, outputs + e);
    Enqueue(shards_[k].get(), std::move(t));
  }
  return absl::OkStatus();
}

absl::Status ShardedBufferProcessor::Synchronize() {
  std::vector<uint64_t> tickets(shards_.size());
  for (size_t k = 0; k < shards_.size(); ++k) {
    Task t;
    t.kind = Task::kSync;
    tickets[k] = Enqueue(shards_[k].get(), std::move(t));
  }
  absl::Status first = absl::OkStatus();
  for (size_t k = 0; k < shards_.size(); ++k) {
    absl::Status st = Wait(shards_[k].get(), tickets[k]);
    if (!st.ok()) {
      BAND_LOG(LogSeverity::kError, "ShardedBufferProcessor: shard %d (device %d) failed %s", static_cast<int>(k),
               shards_[k]->device, st.ToString().c_str());
      if (first.ok()) first = st;
    }
  }
  return first;
}

absl::Status ShardedBufferProcessor::ProcessBatch(const Buffer* const* inputs, Buffer* const* outputs, size_t n) {
  RETURN_IF_ERROR(ProcessBatchAsync(inputs, outputs, n));
  return Synchronize();
}

void* ShardedBufferProcessor::GetComputeStream(size_t shard) {
  if (shard >= shards_.size()) return nullptr;
  std::lock_guard<std::mutex> l(shards_[shard]->mu);  // the worker is idle whenever the caller has synchronised
  return shards_[shard]->processor ? shards_[shard]->processor->GetComputeStream() : nullptr;
}

bool ShardedBufferProcessor::IsSingleFastLaunch(size_t shard) {
  if (shard >= shards_.size()) return false;
  std::lock_guard<std::mutex> l(shards_[shard]->mu);
  return shards_[shard]->processor && shards_[shard]->processor->IsSingleFastLaunch();
}

}  // namespace band
