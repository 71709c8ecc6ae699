// band/buffer/common_operator.h -- Normalize / DataTypeConvert with the
// reference's constructors and validation
// (/root/reference/band/buffer/common_operator.h:27-55, common_operator.cc:67-175).
// The arithmetic runs on the GPU: fused into the store stage of the pixel
// kernel when it follows image operators, an element-wise launch otherwise.
#ifndef BAND_BUFFER_COMMON_OPERATOR_H_
#define BAND_BUFFER_COMMON_OPERATOR_H_

#include "absl/status/status.h"
#include "band/buffer/buffer.h"
#include "band/buffer/operator.h"

namespace band {
namespace buffer {

// out[i] = static_cast<Out>(in[i] - mean) / std over W*H*channels elements.
class Normalize : public IBufferOperator {
 public:
  Normalize(float mean, float std, bool inplace) : mean_(mean), std_(std), inplace_(inplace) {}
  Type GetOpType() const override;

  void SetOutput(Buffer* output) override;
  absl::Status ValidateInput(const Buffer& input) const override;
  absl::Status ValidateOutput(const Buffer& input) const override;
  absl::Status CreateOutput(const Buffer& input) override;

 protected:
  absl::Status Lower(cuda::ChainLowering& lowering, const Buffer& input) const override;
  IBufferOperator* Clone() const override;

  float mean_, std_;
  bool inplace_;
};

// Normalize(0, 1): a cast between element types; skipped when they are equal.
class DataTypeConvert : public Normalize {
 public:
  DataTypeConvert() : Normalize(0.f, 1.f, false) {}

 protected:
  absl::Status Lower(cuda::ChainLowering& lowering, const Buffer& input) const override;
  IBufferOperator* Clone() const override;
};

}  // namespace buffer
}  // namespace band

#endif  // BAND_BUFFER_COMMON_OPERATOR_H_
