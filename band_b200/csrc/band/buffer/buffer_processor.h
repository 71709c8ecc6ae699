// band/buffer/buffer_processor.h -- BufferProcessor / IBufferProcessorBuilder
// with the reference's surface (/root/reference/band/buffer/buffer_processor.h:29-72).
// Process(input, output) keeps its meaning -- run the operator chain, the last
// operator writes into `output` -- but the chain is compiled once per
// (input layout, output layout) into a cuda::FusedPlan and replayed as fused
// launches. Additive: ProcessBatch for n independent frames in one call.
#ifndef BAND_BUFFER_BUFFER_PROCESSOR_H_
#define BAND_BUFFER_BUFFER_PROCESSOR_H_

#include <memory>
#include <vector>

#include "absl/status/status.h"
#include "absl/status/statusor.h"
#include "band/buffer/operator.h"

namespace band {

class Buffer;
class IBufferOperator;
class BufferProcessor;
namespace cuda {
class FusedPlan;
}

class IBufferProcessorBuilder {
 public:
  IBufferProcessorBuilder() = default;
  virtual ~IBufferProcessorBuilder() = default;
  // Build a processor from the operations added to this builder.
  virtual absl::StatusOr<std::unique_ptr<BufferProcessor>> Build() = 0;

  // e.g. builder.AddOperation(std::make_unique<buffer::Resize>(224, 224));
  IBufferProcessorBuilder& AddOperation(std::unique_ptr<IBufferOperator> operation);

 protected:
  std::unique_ptr<BufferProcessor> CreateProcessor(std::vector<IBufferOperator*> operations);

  IBufferProcessorBuilder(const IBufferProcessorBuilder&) = delete;
  IBufferProcessorBuilder& operator=(const IBufferProcessorBuilder&) = delete;

  std::vector<std::unique_ptr<IBufferOperator>> operations_;
};

// Additive: one batch recorded into a CUDA graph (BufferProcessor::RecordBatch). The graph's kernels read plan-owned
// device memory (Normalize table, scratch images, pointers resolved at record time), so the recording keeps its plan
// alive -- the processor may since have compiled plans for other layouts, or be gone. The recorded frames and
// outputs must stay where they are and outlive the recording.
struct RecordedBatch {
  ~RecordedBatch();
  std::shared_ptr<cuda::FusedPlan> plan;
  void* graph_exec = nullptr;
};

// Not re-entrant (like the reference: Process mutates operator state); use one
// processor per thread and per GPU.
class BufferProcessor {
 public:
  virtual ~BufferProcessor();
  virtual absl::Status Process(const Buffer& input, Buffer& output);
  // Additive: n frames sharing one layout, each written to its own output.
  // Frames may be host or device memory; returns after all results are in place.
  virtual absl::Status ProcessBatch(const Buffer* const* inputs, Buffer* const* outputs, size_t n);
  // Additive: enqueue only; results are complete after Synchronize(). The buffers must stay alive until then.
  virtual absl::Status ProcessBatchAsync(const Buffer* const* inputs, Buffer* const* outputs, size_t n);
  absl::Status Synchronize();
  // Additive: record one batch into a CUDA graph (cuda::FusedPlan::Record), replayed with ReplayBatch. Frames and
  // outputs must be device-resident (or pinned with the zero-copy conditions met) and stay in place.
  absl::StatusOr<std::unique_ptr<RecordedBatch>> RecordBatch(const Buffer* const* inputs, Buffer* const* outputs,
                                                             size_t n);
  static absl::Status ReplayBatch(RecordedBatch& recorded, bool synchronize = true);
  // CUDA stream the fused launches go to (nullptr before the first Process*); for event timing.
  void* GetComputeStream() const;
  // Additive introspection for tests / benchmarks.
  size_t GetNumLaunchSegments() const;
  bool IsSingleFastLaunch() const;

 protected:
  friend class IBufferProcessorBuilder;
  BufferProcessor(std::vector<IBufferOperator*> operations);
  BufferProcessor(const BufferProcessor&) = delete;
  BufferProcessor& operator=(const BufferProcessor&) = delete;

  // Runs the operator protocol (ValidateInput / ValidateOrCreateOutput) down the
  // chain with shape-only intermediates and compiles the plan if the layouts changed.
  absl::Status Prepare(const Buffer& input, Buffer& output);

  std::vector<IBufferOperator*> operations_;
  // compiled plans, most recently used first: a processor fed alternating layouts (two camera formats, a frame and a
  // thumbnail) re-lowers nothing. plan_ is the one the current call runs.
  static constexpr size_t kPlanCacheSize = 4;
  std::vector<std::shared_ptr<cuda::FusedPlan>> plans_;
  std::shared_ptr<cuda::FusedPlan> plan_;
  // The previous call, for the repeat fast path: a camera loop (or a benchmark) hands over the SAME Buffer objects call after
  // call. While Buffer::Epoch() stands still those pointers name the same views over the same memory, so validation, plan
  // look-up and pointer resolution are skipped and the plan replays its launches (the per-frame host work otherwise bounds
  // batches of very small frames at ~14 M frames/s and costs a single-frame call ~4 us).
  bool SameAsLastCall(const Buffer* const* inputs, Buffer* const* outputs, size_t n) const;
  void RememberCall(const Buffer* const* inputs, Buffer* const* outputs, size_t n);
  std::vector<const Buffer*> last_in_;
  std::vector<Buffer*> last_out_;
  uint64_t last_epoch_ = 0;
  const cuda::FusedPlan* last_plan_ = nullptr;
};
}  // namespace band

#endif  // BAND_BUFFER_BUFFER_PROCESSOR_H_
