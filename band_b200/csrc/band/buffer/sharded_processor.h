// band/buffer/sharded_processor.h -- one batch of frames across the GPUs of one box, through ONE call.
//
// The reference processes a frame on the caller's thread and leaves batching as a TODO
// (/root/reference/band/buffer/buffer_processor.h:54-58 "a processor can have multiple inputs and outputs"); its
// only entry is Process(input, output) (buffer_processor.cc:22-43, C: c_api_buffer.cc:198-214). Frames are
// independent, so a batch shards across devices with no exchange step: frame i of n goes to shard
// i / ceil(n / shards) -- contiguous slices -- and every shard is an ordinary BufferProcessor with its own compiled
// plans and streams on its own device, driven by its own host thread. There is no collective and no peer traffic.
//
// Memory: host frames/tensors (pinned or pageable) work on any shard. DEVICE-resident frames and tensors must live
// on the device of the shard that owns their slice (SliceOf / DeviceOf tell which); nothing is migrated.
#ifndef BAND_BUFFER_SHARDED_PROCESSOR_H_
#define BAND_BUFFER_SHARDED_PROCESSOR_H_

#include <condition_variable>
#include <deque>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>

#include "absl/status/status.h"
#include "absl/status/statusor.h"
#include "band/buffer/buffer_processor.h"

namespace band {

class ShardedBufferProcessor {
 public:
  // Builds one processor per listed device from the same builder (Build() clones the operators each time,
  // image_processor.cc:25-67). Devices may repeat (two shards on one GPU) but normally list each GPU once.
  static absl::StatusOr<std::unique_ptr<ShardedBufferProcessor>> Create(IBufferProcessorBuilder& builder,
                                                                        const std::vector<int>& devices);
  ~ShardedBufferProcessor();

  // [begin, end) of a shard's contiguous slice of an n-frame batch.
  static void SliceOf(size_t n, size_t shards, size_t shard, size_t* begin, size_t* end);
  size_t NumShards() const { return shards_.size(); }
  int DeviceOf(size_t shard) const { return shards_[shard]->device; }

  // n frames of one layout -> n outputs; returns when every shard's results are in place.
  absl::Status ProcessBatch(const Buffer* const* inputs, Buffer* const* outputs, size_t n);
  // Enqueue only (every shard's worker calls ProcessBatchAsync on its slice); complete after Synchronize(). Buffers
  // and outputs must stay alive until then. Several calls may be queued back to back (e.g. one per target shape).
  absl::Status ProcessBatchAsync(const Buffer* const* inputs, Buffer* const* outputs, size_t n);
  absl::Status Synchronize();

  // for harness timing: the stream a shard's fused launches go to (nullptr before its first batch)
  void* GetComputeStream(size_t shard);
  bool IsSingleFastLaunch(size_t shard);

 private:
  struct Task {
    enum Kind { kRun, kSync, kStop } kind = kRun;
    std::vector<const Buffer*> in;
    std::vector<Buffer*> out;
    uint64_t ticket = 0;
  };
  struct Shard {
    int device = 0;
    std::unique_ptr<BufferProcessor> processor;
    std::thread worker;
    std::mutex mu;
    std::condition_variable cv, done_cv;
    std::deque<Task> queue;
    uint64_t issued = 0, completed = 0;  // tickets
    absl::Status status = absl::OkStatus();  // first failure since the last Synchronize
  };
  ShardedBufferProcessor() = default;
  static void WorkerLoop(Shard* s);
  uint64_t Enqueue(Shard* s, Task&& t);
  absl::Status Wait(Shard* s, uint64_t ticket);

  std::vector<std::unique_ptr<Shard>> shards_;
};

}  // namespace band
#endif  // BAND_BUFFER_SHARDED_PROCESSOR_H_
