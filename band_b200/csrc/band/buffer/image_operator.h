// band/buffer/image_operator.h -- the image operator classes with the
// reference's constructors, validation rules and status codes
// (/root/reference/band/buffer/image_operator.h:28-159, image_operator.cc).
// None of them touches pixels: each describes its stage to the plan compiler
// (Lower), which fuses the chain into CUDA launches (fused_plan.h).
#ifndef BAND_BUFFER_IMAGE_OPERATOR_H_
#define BAND_BUFFER_IMAGE_OPERATOR_H_

#include <memory>
#include <vector>

#include "absl/status/status.h"
#include "band/buffer/buffer.h"
#include "band/buffer/common_operator.h"
#include "band/buffer/operator.h"

namespace band {
namespace buffer {

// Inclusive box (x0,y0)-(x1,y1); output is (x1-x0+1) x (y1-y0+1).
class Crop : public IBufferOperator {
 public:
  Crop(int x0, int y0, int x1, int y1) : x0_(x0), y0_(y0), x1_(x1), y1_(y1) {}
  Type GetOpType() const override;
  absl::Status ValidateInput(const Buffer& input) const override;
  absl::Status ValidateOutput(const Buffer& input) const override;
  absl::Status CreateOutput(const Buffer& input) override;

 protected:
  absl::Status Lower(cuda::ChainLowering& lowering, const Buffer& input) const override;
  IBufferOperator* Clone() const override;

 private:
  int x0_, y0_, x1_, y1_;
};

// Bilinear resize. -1 = take that dimension from an assigned output.
class Resize : public IBufferOperator {
 public:
  Resize(int width = -1, int height = -1) : dims_({width, height}) {}
  Resize(const std::vector<int>& dims) : dims_(dims) {}
  Type GetOpType() const override;
  absl::Status ValidateInput(const Buffer& input) const override;
  absl::Status ValidateOutput(const Buffer& input) const override;
  absl::Status CreateOutput(const Buffer& input) override;

 protected:
  absl::Status Lower(cuda::ChainLowering& lowering, const Buffer& input) const override;
  IBufferOperator* Clone() const override;

 private:
  bool IsAuto(size_t dim) const { return dims_[dim] == -1; }
  std::vector<int> dims_;
};

// Counter-clockwise rotation by a multiple of 90 degrees (any integer is reduced mod 360).
class Rotate : public IBufferOperator {
 public:
  Rotate(int angle_deg) : angle_deg_(angle_deg % 360) {
    if (angle_deg_ < 0) angle_deg_ += 360;
  }
  Type GetOpType() const override;
  absl::Status ValidateInput(const Buffer& input) const override;
  absl::Status ValidateOutput(const Buffer& input) const override;
  absl::Status CreateOutput(const Buffer& input) override;

 protected:
  absl::Status Lower(cuda::ChainLowering& lowering, const Buffer& input) const override;
  IBufferOperator* Clone() const override;

 private:
  int angle_deg_;
};

class Flip : public IBufferOperator {
 public:
  Flip(bool horizontal, bool vertical) : horizontal_(horizontal), vertical_(vertical) {}
  Type GetOpType() const override;
  absl::Status ValidateInput(const Buffer& input) const override;
  absl::Status ValidateOutput(const Buffer& input) const override;
  absl::Status CreateOutput(const Buffer& input) override;

 protected:
  absl::Status Lower(cuda::ChainLowering& lowering, const Buffer& input) const override;
  IBufferOperator* Clone() const override;

 private:
  bool horizontal_, vertical_;
};

// ADDITIVE. The reference carries Buffer::GetOrientation() (EXIF 1..8, band/common.h:146-161) from the
// camera to the processor and never applies it. AutoOrient turns the frame upright: the orientation becomes the
// equivalent Flip / Rotate pair (2: mirror, 3: 180, 4: vertical mirror, 5: transpose, 6: quarter-turn clockwise,
// 7: transverse, 8: quarter-turn counter-clockwise), which the plan compiler folds into the same index map as any
// other Crop / Rotate / Flip -- no extra pass. Output: same format, dimensions swapped for 5..8, kTopLeft.
class AutoOrient : public IBufferOperator {
 public:
  AutoOrient() = default;
  Type GetOpType() const override;
  absl::Status ValidateInput(const Buffer& input) const override;
  absl::Status ValidateOutput(const Buffer& input) const override;
  absl::Status CreateOutput(const Buffer& input) override;
  // orientation -> (mirror horizontally first, then rotate counter-clockwise by angle)
  static void Decompose(BufferOrientation orientation, bool* mirror, int* angle_deg);

 protected:
  absl::Status Lower(cuda::ChainLowering& lowering, const Buffer& input) const override;
  IBufferOperator* Clone() const override;
};

class ColorSpaceConvert : public IBufferOperator {
 public:
  ColorSpaceConvert() : output_format_(BufferFormat::kRaw), is_format_specified_(false) {}
  ColorSpaceConvert(BufferFormat buffer_format) : output_format_(buffer_format), is_format_specified_(true) {}
  Type GetOpType() const override;
  absl::Status ValidateOutput(const Buffer& input) const override;
  absl::Status CreateOutput(const Buffer& input) override;

 protected:
  absl::Status Lower(cuda::ChainLowering& lowering, const Buffer& input) const override;
  IBufferOperator* Clone() const override;

 private:
  BufferFormat output_format_;
  bool is_format_specified_;
};

// The default processor: ColorSpaceConvert -> Resize -> DataTypeConvert, each
// only if the assigned output differs from the input in that respect.
class AutoConvert : public IBufferOperator {
 public:
  AutoConvert() = default;
  AutoConvert(const AutoConvert& rhs);
  AutoConvert& operator=(const AutoConvert& rhs);
  ~AutoConvert() override;

  Type GetOpType() const override;
  absl::Status ValidateInput(const Buffer& input) const override;
  absl::Status ValidateOutput(const Buffer& input) const override;
  absl::Status CreateOutput(const Buffer& input) override;
  void SetOutput(Buffer* output) override;

 protected:
  absl::Status Lower(cuda::ChainLowering& lowering, const Buffer& input) const override;
  IBufferOperator* Clone() const override;

 private:
  bool RequiresColorSpaceConvert(const Buffer& input) const;
  bool RequiresResize(const Buffer& input) const;
  bool RequiresDataTypeConvert(const Buffer& input) const;

  std::unique_ptr<ColorSpaceConvert> color_space_convert_;
  std::unique_ptr<Resize> resize_;
  std::unique_ptr<DataTypeConvert> data_type_convert_;
};

}  // namespace buffer
}  // namespace band

#endif  // BAND_BUFFER_IMAGE_OPERATOR_H_
