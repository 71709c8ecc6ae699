// band/buffer/image_processor.cc -- ImageProcessorBuilder::Build
// (/root/reference/band/buffer/image_processor.cc:25-67): operator classes may
// appear once each, every operator is cloned into the processor, and an empty
// builder yields the AutoConvert default. The chain is kept in the order given
// (the reference never reorders either, :28-29); fusion happens at plan level.
#include "band/buffer/image_processor.h"

#include <typeinfo>

#include "absl/strings/str_format.h"
#include "band/buffer/image_operator.h"

namespace band {

absl::StatusOr<std::unique_ptr<BufferProcessor>> ImageProcessorBuilder::Build() {
  for (const auto& operation : operations_) {
    if (operation == nullptr) return absl::InvalidArgumentError("operation is nullptr.");
  }
  // Same pair scan as the reference (j starts at 1, so a duplicate of the FIRST
  // operator is still caught from the other side).
  for (size_t i = 0; i < operations_.size(); ++i) {
    for (size_t j = 1; j < operations_.size(); ++j) {
      if (i == j) continue;
      const IBufferOperator& lhs = *operations_[i];
      const IBufferOperator& rhs = *operations_[j];
      if (typeid(lhs) == typeid(rhs)) {
        return absl::InvalidArgumentError(absl::StrFormat("operation %s is duplicated.", typeid(lhs).name()));
      }
    }
  }
  std::vector<IBufferOperator*> cloned;
  for (const auto& operation : operations_) {
    const IBufferOperator::Type type = operation->GetOpType();
    if (type != IBufferOperator::Type::kImage && type != IBufferOperator::Type::kCommon) {
      for (IBufferOperator* op : cloned) delete op;
      return absl::InvalidArgumentError(
          absl::StrFormat("operation type %d is not supported.", static_cast<int>(type)));
    }
    cloned.push_back(operation->Clone());
  }
  if (cloned.empty()) cloned.push_back(new buffer::AutoConvert());
  return CreateProcessor(std::move(cloned));
}
}  // namespace band
