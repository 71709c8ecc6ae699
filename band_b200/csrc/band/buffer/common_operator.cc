// band/buffer/common_operator.cc -- see common_operator.h.
#include "band/buffer/common_operator.h"

#include "absl/strings/str_format.h"
#include "band/buffer/fused_plan.h"
#include "band/logger.h"

namespace band {
namespace buffer {

IBufferOperator* Normalize::Clone() const { return new Normalize(*this); }
IBufferOperator* DataTypeConvert::Clone() const { return new DataTypeConvert(*this); }

IBufferOperator::Type Normalize::GetOpType() const { return Type::kCommon; }

// common_operator.cc:71-78: an in-place Normalize refuses an output.
void Normalize::SetOutput(Buffer* output) {
  if (inplace_) {
    BAND_LOG(LogSeverity::kError, "Normalize: setting output buffer is not allowed for inplace");
    return;
  }
  IBufferOperator::SetOutput(output);
}

// common_operator.cc:111-126: packed single-plane buffers only.
absl::Status Normalize::ValidateInput(const Buffer& input) const {
  if (Buffer::IsYUV(input.GetBufferFormat())) {
    return absl::InvalidArgumentError(absl::StrFormat("input buffer format %d is not supported.",
                                                      static_cast<int>(input.GetBufferFormat())));
  }
  if (input.GetNumPlanes() != 1) {
    return absl::InvalidArgumentError(absl::StrFormat(
        "input buffer should have only one plane, but it has %d planes.", static_cast<int>(input.GetNumPlanes())));
  }
  return absl::OkStatus();
}

// common_operator.cc:128-151: same format and dimensions (the element type may differ).
absl::Status Normalize::ValidateOutput(const Buffer& input) const {
  if (inplace_) return absl::OkStatus();
  if (input.GetBufferFormat() != output_->GetBufferFormat()) {
    return absl::InvalidArgumentError(absl::StrFormat(
        "input buffer should have the same format as output buffer, but input format is %s and output format is %s.",
        ToString(input.GetBufferFormat()), ToString(output_->GetBufferFormat())));
  }
  if (input.GetDimension() != output_->GetDimension()) {
    return absl::InvalidArgumentError(absl::StrFormat(
        "input buffer should have the same dimension as output buffer, but input dimension is %s and output "
        "dimension is %s.",
        ToString(input.GetDimension()), ToString(output_->GetDimension())));
  }
  return absl::OkStatus();
}

// common_operator.cc:153-163: same shape AND element type as the input.
absl::Status Normalize::CreateOutput(const Buffer& input) {
  if (inplace_) return absl::OkStatus();
  if (input.GetDimension().size() < 2) return absl::InvalidArgumentError("Normalize: input buffer has no 2-D shape");
  Buffer* created = Buffer::CreateEmpty(input.GetDimension()[0], input.GetDimension()[1], input.GetBufferFormat(),
                                        input.GetDataType(), input.GetOrientation(), OutputSpace());
  if (created == nullptr) return absl::InternalError("Normalize: failed to create the output buffer");
  OwnOutput(created);
  return absl::OkStatus();
}

absl::Status Normalize::Lower(cuda::ChainLowering& lowering, const Buffer& input) const {
  // common_operator.cc:80-109: inputs may be u8/i8/i16/i32/f32/f64
  switch (input.GetDataType()) {
    case DataType::kUInt8: case DataType::kInt8: case DataType::kInt16: case DataType::kInt32:
    case DataType::kFloat32: case DataType::kFloat64:
      break;
    default:
      return absl::InvalidArgumentError(
          absl::StrFormat("data type %d is not supported.", static_cast<int>(input.GetDataType())));
  }
  const DataType out = inplace_ ? input.GetDataType() : GetOutput()->GetDataType();
  return lowering.Normalize(mean_, std_, out, inplace_);
}

// common_operator.cc:169-175
absl::Status DataTypeConvert::Lower(cuda::ChainLowering& lowering, const Buffer& input) const {
  if (input.GetDataType() == GetOutput()->GetDataType()) return lowering.PassThrough();
  return Normalize::Lower(lowering, input);
}

}  // namespace buffer
}  // namespace band
