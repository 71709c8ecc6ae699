// band/buffer/buffer_processor.cc -- see buffer_processor.h
// (/root/reference/band/buffer/buffer_processor.cc:22-64).
#include "band/buffer/buffer_processor.h"

#include <algorithm>
#include <cstring>

#include "band/buffer/buffer.h"
#include "band/buffer/fused_plan.h"
#include "band/logger.h"
#include "bandcu.h"

namespace band {

BufferProcessor::BufferProcessor(std::vector<IBufferOperator*> operations) : operations_(std::move(operations)) {
  // intermediates between operators are never materialised: shapes only
  for (IBufferOperator* op : operations_) op->virtual_outputs_ = true;
}

RecordedBatch::~RecordedBatch() {
  if (graph_exec) {
    if (plan) (void)plan->Synchronize();  // a replay may still be running on the plan's stream
    bandcu_graph_destroy(graph_exec);
  }
}

BufferProcessor::~BufferProcessor() {
  plan_.reset();
  plans_.clear();
  for (IBufferOperator* op : operations_) delete op;
}

absl::Status BufferProcessor::Prepare(const Buffer& input, Buffer& output) {
  if (operations_.empty()) return absl::InternalError("IProcessor: no operations are specified.");
  // the last operation writes the caller's buffer (buffer_processor.cc:28)
  operations_.back()->SetOutput(&output);

  const Buffer* next_input = &input;
  for (size_t i = 0; i < operations_.size(); ++i) {
    if (next_input == nullptr) return absl::InternalError("an in-place operator has no output to pass on");
    IBufferOperator* op = operations_[i];
    absl::Status status = op->ValidateInput(*next_input);
    if (status.ok()) status = op->ValidateOrCreateOutput(*next_input);
    if (!status.ok()) {
      BAND_LOG(LogSeverity::kError, "BufferProcessor::Process failed %s", status.ToString().c_str());
      return status;
    }
    next_input = op->GetOutput();
  }

  const cuda::ImageLayout in = cuda::ImageLayout::Of(input);
  const cuda::ImageLayout out = cuda::ImageLayout::Of(output);
  for (size_t i = 0; i < plans_.size(); ++i) {
    const std::shared_ptr<cuda::FusedPlan>& p = plans_[i];
    if (p->input_layout() == in && (!p->writes_output() || p->output_layout() == out)) {
      plan_ = p;
      if (i != 0) std::rotate(plans_.begin(), plans_.begin() + i, plans_.begin() + i + 1);
      return absl::OkStatus();
    }
  }

  cuda::ChainLowering lowering(input);
  next_input = &input;
  for (IBufferOperator* op : operations_) {
    absl::Status status = lowering.LowerNested(*op, *next_input);
    if (!status.ok()) {
      BAND_LOG(LogSeverity::kError, "BufferProcessor::Process failed %s", status.ToString().c_str());
      return status;
    }
    next_input = op->GetOutput();
  }
  absl::StatusOr<std::unique_ptr<cuda::FusedPlan>> plan = lowering.Finish(&output);
  if (!plan.ok()) {
    BAND_LOG(LogSeverity::kError, "BufferProcessor::Process failed %s", plan.status().ToString().c_str());
    return plan.status();
  }
  plan_ = std::move(plan).value();
  if (plans_.size() >= kPlanCacheSize) {
    // the evicted plan may still have an asynchronous pass in flight: its destructor drains its streams
    plans_.pop_back();
  }
  plans_.insert(plans_.begin(), plan_);
  return absl::OkStatus();
}

bool BufferProcessor::SameAsLastCall(const Buffer* const* inputs, Buffer* const* outputs, size_t n) const {
  return plan_ && last_plan_ == plan_.get() && n == last_in_.size() && n > 0 && Buffer::Epoch() == last_epoch_ &&
         plan_->CanRerun() && memcmp(inputs, last_in_.data(), n * sizeof(*inputs)) == 0 &&
         memcmp(outputs, last_out_.data(), n * sizeof(*outputs)) == 0;
}

void BufferProcessor::RememberCall(const Buffer* const* inputs, Buffer* const* outputs, size_t n) {
  if (!plan_ || !plan_->CanRerun()) {
    last_plan_ = nullptr;
    return;
  }
  last_in_.assign(inputs, inputs + n);
  last_out_.assign(outputs, outputs + n);
  last_plan_ = plan_.get();
  last_epoch_ = Buffer::Epoch();  // read AFTER the run: the first call of a layout creates the shape-only intermediates
}

absl::Status BufferProcessor::Process(const Buffer& input, Buffer& output) {
  const Buffer* in = &input;
  Buffer* out = &output;
  if (SameAsLastCall(&in, &out, 1)) return plan_->Rerun(/*synchronize=*/true);
  last_plan_ = nullptr;
  RETURN_IF_ERROR(Prepare(input, output));
  absl::Status status = plan_->Run(&in, &out, 1);
  if (!status.ok()) {
    BAND_LOG(LogSeverity::kError, "BufferProcessor::Process failed %s", status.ToString().c_str());
  } else {
    RememberCall(&in, &out, 1);
  }
  return status;
}

absl::Status BufferProcessor::ProcessBatch(const Buffer* const* inputs, Buffer* const* outputs, size_t n) {
  if (n == 0) return absl::OkStatus();
  if (inputs == nullptr || outputs == nullptr || inputs[0] == nullptr || outputs[0] == nullptr)
    return absl::InvalidArgumentError("ProcessBatch: null buffer list");
  if (SameAsLastCall(inputs, outputs, n)) return plan_->Rerun(/*synchronize=*/true);
  last_plan_ = nullptr;
  RETURN_IF_ERROR(Prepare(*inputs[0], *outputs[0]));
  absl::Status status = plan_->Run(inputs, outputs, static_cast<int>(n));
  if (!status.ok()) {
    BAND_LOG(LogSeverity::kError, "BufferProcessor::ProcessBatch failed %s", status.ToString().c_str());
  } else {
    RememberCall(inputs, outputs, n);
  }
  return status;
}

absl::Status BufferProcessor::ProcessBatchAsync(const Buffer* const* inputs, Buffer* const* outputs, size_t n) {
  if (n == 0) return absl::OkStatus();
  if (inputs == nullptr || outputs == nullptr || inputs[0] == nullptr || outputs[0] == nullptr)
    return absl::InvalidArgumentError("ProcessBatchAsync: null buffer list");
  if (SameAsLastCall(inputs, outputs, n)) return plan_->Rerun(/*synchronize=*/false);
  last_plan_ = nullptr;
  RETURN_IF_ERROR(Prepare(*inputs[0], *outputs[0]));
  absl::Status status = plan_->Run(inputs, outputs, static_cast<int>(n), /*synchronize=*/false);
  if (!status.ok()) {
    BAND_LOG(LogSeverity::kError, "BufferProcessor::ProcessBatchAsync failed %s", status.ToString().c_str());
  } else {
    RememberCall(inputs, outputs, n);
  }
  return status;
}

absl::StatusOr<std::unique_ptr<RecordedBatch>> BufferProcessor::RecordBatch(const Buffer* const* inputs,
                                                                            Buffer* const* outputs, size_t n) {
  if (n == 0 || inputs == nullptr || outputs == nullptr || inputs[0] == nullptr || outputs[0] == nullptr)
    return absl::InvalidArgumentError("RecordBatch: null buffer list");
  for (size_t f = 0; f < n; ++f) {
    const bool in_ok = inputs[f] && (inputs[f]->GetMemorySpace() == MemorySpace::kDevice || inputs[f]->IsPinnedHost());
    const bool out_ok = outputs[f] && (outputs[f]->GetMemorySpace() == MemorySpace::kDevice || outputs[f]->IsPinnedHost());
    if (!in_ok || !out_ok)
      return absl::InvalidArgumentError("RecordBatch: frames and outputs must be device-resident or pinned host memory");
  }
  RETURN_IF_ERROR(Prepare(*inputs[0], *outputs[0]));
  std::unique_ptr<RecordedBatch> recorded(new RecordedBatch());
  recorded->plan = plan_;
  absl::Status status = plan_->Record(inputs, outputs, static_cast<int>(n), &recorded->graph_exec);
  if (!status.ok()) {
    BAND_LOG(LogSeverity::kError, "BufferProcessor::RecordBatch failed %s", status.ToString().c_str());
    return status;
  }
  return recorded;
}

absl::Status BufferProcessor::ReplayBatch(RecordedBatch& recorded, bool synchronize) {
  if (!recorded.plan || recorded.graph_exec == nullptr) return absl::InvalidArgumentError("ReplayBatch: nothing recorded");
  void* stream = recorded.plan->ComputeStream();
  if (bandcu_graph_launch(recorded.graph_exec, stream) != BANDCU_OK) return absl::InternalError(bandcu_last_error_string());
  if (synchronize && bandcu_stream_sync(stream) != BANDCU_OK) return absl::InternalError(bandcu_last_error_string());
  return absl::OkStatus();
}

absl::Status BufferProcessor::Synchronize() {
  // every cached plan: asynchronous calls on alternating layouts leave work on more than one plan's streams
  absl::Status first = absl::OkStatus();
  for (const std::shared_ptr<cuda::FusedPlan>& p : plans_) {
    absl::Status s = p->Synchronize();
    if (first.ok() && !s.ok()) first = s;
  }
  return first;
}
void* BufferProcessor::GetComputeStream() const { return plan_ ? plan_->ComputeStream() : nullptr; }

size_t BufferProcessor::GetNumLaunchSegments() const { return plan_ ? plan_->NumSegments() : 0; }
bool BufferProcessor::IsSingleFastLaunch() const { return plan_ && plan_->IsSingleFastLaunch(); }

IBufferProcessorBuilder& IBufferProcessorBuilder::AddOperation(std::unique_ptr<IBufferOperator> operation) {
  operations_.emplace_back(std::move(operation));
  return *this;
}

std::unique_ptr<BufferProcessor> IBufferProcessorBuilder::CreateProcessor(std::vector<IBufferOperator*> operations) {
  return std::unique_ptr<BufferProcessor>(new BufferProcessor(std::move(operations)));
}

}  // namespace band
