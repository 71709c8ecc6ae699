// band/buffer/operator.cc -- the operator protocol
// (/root/reference/band/buffer/operator.cc:52-97).
#include "band/buffer/operator.h"

#include "band/buffer/fused_plan.h"
#include "band/logger.h"

namespace band {

IBufferOperator::IBufferOperator(const IBufferOperator&) {}
IBufferOperator::IBufferOperator(IBufferOperator&&) {}
IBufferOperator& IBufferOperator::operator=(const IBufferOperator&) {
  owned_output_.reset();
  single_plan_.reset();
  output_ = nullptr;
  output_assigned_ = false;
  return *this;
}

// operator.cc:52-57
absl::Status IBufferOperator::Process(const Buffer& input) {
  RETURN_IF_ERROR(ValidateInput(input));
  RETURN_IF_ERROR(ValidateOrCreateOutput(input));
  RETURN_IF_ERROR(ProcessImpl(input));
  return absl::OkStatus();
}

// operator.cc:59-66
void IBufferOperator::SetOutput(Buffer* output) {
  if (output) {
    output_assigned_ = true;
    owned_output_.reset();
    output_ = output;
  }
}

Buffer* IBufferOperator::GetOutput() { return output_; }
const Buffer* IBufferOperator::GetOutput() const { return output_; }

void IBufferOperator::OwnOutput(Buffer* created) {
  owned_output_.reset(created);
  output_ = created;
}

absl::Status IBufferOperator::ValidateInput(const Buffer&) const { return absl::OkStatus(); }

// operator.cc:75-97: an assigned output must validate; a cached or missing one
// is (re)created and validated again.
absl::Status IBufferOperator::ValidateOrCreateOutput(const Buffer& input) {
  absl::Status status = output_ == nullptr ? absl::InternalError("Null output buffer") : ValidateOutput(input);
  if (status.ok()) return status;
  if (output_assigned_) return status;
  RETURN_IF_ERROR(CreateOutput(input));
  status = ValidateOutput(input);
  if (!status.ok()) {
    BAND_LOG(LogSeverity::kError, "Failed to create valid output buffer: %s", status.message().c_str());
  }
  return status;
}

// A stand-alone operator is a one-stage chain: compile it and run it on the GPU.
absl::Status IBufferOperator::ProcessImpl(const Buffer& input) {
  return cuda::RunSingleOperator(*this, input);
}

}  // namespace band
