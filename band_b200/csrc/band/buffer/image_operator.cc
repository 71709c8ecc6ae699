// band/buffer/image_operator.cc -- see image_operator.h. Validation follows
// /root/reference/band/buffer/image_operator.cc (cited per method); pixel work
// is described to cuda::ChainLowering instead of calling libyuv.
#include "band/buffer/image_operator.h"

#include "absl/strings/str_format.h"
#include "band/buffer/fused_plan.h"
#include "band/logger.h"

namespace band {
namespace buffer {
namespace {

bool IsRaw(const Buffer& b) { return b.GetBufferFormat() == BufferFormat::kRaw; }

absl::Status RawNotSupported(const char* op) {
  return absl::InvalidArgumentError(absl::StrFormat("%s: Raw buffer format type is not supported.", op));
}

// An intermediate shaped like `input` but w x h (and possibly another format / element type).
Buffer* EmptyLike(const Buffer& input, size_t w, size_t h, BufferFormat format, DataType dtype, MemorySpace space) {
  return Buffer::CreateEmpty(w, h, format, dtype, input.GetOrientation(), space);
}

bool QuarterTurn(int angle_deg) { return (angle_deg / 90) % 2 == 1; }

// An index operator (or a YUV-domain Resize) whose cached output carries another layout than this call's input. The
// reference accepts every pair inside {RGB, RGBA} and inside the four 4:2:0 layouts (Buffer::IsBufferFormatCompatible,
// buffer.cc:367-388). What it then does is only well defined for the two PLANAR layouts: CropYv / RotateYv / FlipYv /
// ResizeYv resolve the U and V planes of both buffers by name (libyuv_image_operator.cc GetYuvDataFromBuffer), so a
// processor that first saw YV21 and now sees YV12 produces the same pixels as a fresh one -- a camera that swaps its
// planar layout mid-stream keeps working. The other pairs are refused here: semi-planar pairs copy the interleaved
// plane raw and relabel it (U and V change places) or, for Rotate, write it one byte off (:816-887); planar <->
// semi-planar and RGBA -> RGB move rows of one geometry into buffers of another, past their end.
bool PlanarPair(BufferFormat a, BufferFormat b) {
  const bool pa = a == BufferFormat::kYV12 || a == BufferFormat::kYV21;
  const bool pb = b == BufferFormat::kYV12 || b == BufferFormat::kYV21;
  return pa && pb;
}
bool LayoutsAgree(const Buffer& input, const Buffer& output) {
  return input.GetBufferFormat() == output.GetBufferFormat() || PlanarPair(input.GetBufferFormat(), output.GetBufferFormat());
}

}  // namespace

IBufferOperator::Type Crop::GetOpType() const { return Type::kImage; }
IBufferOperator::Type Resize::GetOpType() const { return Type::kImage; }
IBufferOperator::Type Rotate::GetOpType() const { return Type::kImage; }
IBufferOperator::Type Flip::GetOpType() const { return Type::kImage; }
IBufferOperator::Type AutoOrient::GetOpType() const { return Type::kImage; }
IBufferOperator::Type ColorSpaceConvert::GetOpType() const { return Type::kImage; }
// image_operator.cc:72: value-initialised Type, i.e. kImage
IBufferOperator::Type AutoConvert::GetOpType() const { return Type(); }

IBufferOperator* Crop::Clone() const { return new Crop(*this); }
IBufferOperator* Resize::Clone() const { return new Resize(*this); }
IBufferOperator* Rotate::Clone() const { return new Rotate(*this); }
IBufferOperator* Flip::Clone() const { return new Flip(*this); }
IBufferOperator* AutoOrient::Clone() const { return new AutoOrient(*this); }
IBufferOperator* ColorSpaceConvert::Clone() const { return new ColorSpaceConvert(*this); }
IBufferOperator* AutoConvert::Clone() const { return new AutoConvert(*this); }

// ---- Crop (image_operator.cc:74-157) ----------------------------------------
absl::Status Crop::ValidateInput(const Buffer& input) const {
  if (IsRaw(input)) return RawNotSupported("Crop");
  if (x0_ < 0 || y0_ < 0 || x1_ < 0 || y1_ < 0) {
    return absl::InvalidArgumentError(absl::StrFormat(
        "Crop: negative crop region is not allowed. (%d, %d) -> (%d, %d)", x0_, y0_, x1_, y1_));
  }
  if (x0_ >= x1_ || y0_ >= y1_) {
    return absl::InvalidArgumentError(absl::StrFormat(
        "Crop: invalid crop region is not allowed. (%d, %d) -> (%d, %d)", x0_, y0_, x1_, y1_));
  }
  if (static_cast<size_t>(x1_) >= input.GetDimension()[0] || static_cast<size_t>(y1_) >= input.GetDimension()[1]) {
    return absl::InvalidArgumentError(absl::StrFormat(
        "Crop: crop region is out of bounds. (%d, %d) -> (%d, %d)", x0_, y0_, x1_, y1_));
  }
  if (input.GetDataType() != DataType::kUInt8) {
    return absl::InvalidArgumentError("Crop: only UInt8 data type is supported.");
  }
  return absl::OkStatus();
}

absl::Status Crop::ValidateOutput(const Buffer& input) const {
  if (!input.IsBufferFormatCompatible(*output_)) {
    return absl::InvalidArgumentError(absl::StrFormat("Crop: output buffer format type is not compatible. %s vs %s",
                                                      ToString(input.GetBufferFormat()),
                                                      ToString(output_->GetBufferFormat())));
  }
  if (input.GetDataType() != output_->GetDataType()) {
    return absl::InvalidArgumentError(absl::StrFormat("Crop: output buffer data type is not compatible. %s vs %s",
                                                      ToString(input.GetDataType()),
                                                      ToString(output_->GetDataType())));
  }
  const std::vector<size_t> want = Buffer::GetCropDimension(x0_, x1_, y0_, y1_);
  if (want[0] != output_->GetDimension()[0] || want[1] != output_->GetDimension()[1]) {
    return absl::InvalidArgumentError(absl::StrFormat(
        "Crop: output buffer dimension is not compatible. %d x %d vs %d x %d", static_cast<int>(want[0]),
        static_cast<int>(want[1]), static_cast<int>(output_->GetDimension()[0]),
        static_cast<int>(output_->GetDimension()[1])));
  }
  return absl::OkStatus();
}

absl::Status Crop::CreateOutput(const Buffer& input) {
  const std::vector<size_t> dims = Buffer::GetCropDimension(x0_, x1_, y0_, y1_);
  Buffer* created = EmptyLike(input, dims[0], dims[1], input.GetBufferFormat(), input.GetDataType(), OutputSpace());
  if (created == nullptr) return absl::InternalError("Crop: failed to create the output buffer");
  OwnOutput(created);
  return absl::OkStatus();
}

absl::Status Crop::Lower(cuda::ChainLowering& lowering, const Buffer& input) const {
  if (!LayoutsAgree(input, *GetOutput()))
    return absl::InvalidArgumentError("Crop: input and output buffer formats must match");
  return lowering.Crop(x0_, y0_, x1_, y1_);
}

// ---- Resize (image_operator.cc:159-239) --------------------------------------
absl::Status Resize::ValidateInput(const Buffer& input) const {
  if (IsRaw(input)) return RawNotSupported("Resize");
  if (dims_.size() < 2) return absl::InvalidArgumentError("Resize: invalid dimension size.");
  return absl::OkStatus();
}

absl::Status Resize::ValidateOutput(const Buffer& input) const {
  if (input.GetDimension() == output_->GetDimension()) {
    return absl::InvalidArgumentError("Resize: input and output buffer dimensions are the same.");
  }
  if (!input.IsBufferFormatCompatible(*output_)) {
    return absl::InvalidArgumentError("Resize: input and output buffer formats are not compatible.");
  }
  const BufferFormat in = input.GetBufferFormat(), out = output_->GetBufferFormat();
  if (in == BufferFormat::kRGBA) {
    if (out != BufferFormat::kRGB && out != BufferFormat::kRGBA)
      return absl::InvalidArgumentError("Resize: output buffer format type is not compatible.");
  } else if (in == BufferFormat::kGrayScale || in == BufferFormat::kRGB || Buffer::IsYUV(in)) {
    if (in != out) return absl::InvalidArgumentError("Resize: output buffer format type is not compatible.");
  } else {
    return absl::InternalError(absl::StrFormat("Unsupported buffer format: %s.", ToString(in)));
  }
  for (size_t i = 0; i < dims_.size() && i < output_->GetDimension().size(); ++i) {
    if (!IsAuto(i) && static_cast<size_t>(dims_[i]) != output_->GetDimension()[i]) {
      return absl::InvalidArgumentError(absl::StrFormat("Resize: output buffer dimension is not compatible. %d != %d",
                                                        dims_[i], static_cast<int>(output_->GetDimension()[i])));
    }
  }
  return absl::OkStatus();
}

absl::Status Resize::CreateOutput(const Buffer& input) {
  if (IsAuto(0) || IsAuto(1)) {
    return absl::InvalidArgumentError("Resize: cannot create output buffer with auto dimension.");
  }
  if (dims_[0] <= 0 || dims_[1] <= 0) return absl::InvalidArgumentError("Resize: non-positive output dimension.");
  Buffer* created = EmptyLike(input, dims_[0], dims_[1], input.GetBufferFormat(), input.GetDataType(), OutputSpace());
  if (created == nullptr) return absl::InternalError("Resize: failed to create the output buffer");
  OwnOutput(created);
  return absl::OkStatus();
}

absl::Status Resize::Lower(cuda::ChainLowering& lowering, const Buffer& input) const {
  // Deliberate deviation (memory safety): the reference lets RGBA resize into an
  // RGB buffer and writes 4 bytes per pixel into it (libyuv_image_operator.cc:1213-1229).
  if (!LayoutsAgree(input, *GetOutput()))
    return absl::InvalidArgumentError("Resize: input and output buffer formats must match");
  return lowering.Scale(static_cast<int>(GetOutput()->GetDimension()[0]),
                        static_cast<int>(GetOutput()->GetDimension()[1]));
}

// ---- Rotate (image_operator.cc:241-297) --------------------------------------
absl::Status Rotate::ValidateInput(const Buffer& input) const {
  if (IsRaw(input)) return RawNotSupported("Rotate");
  return absl::OkStatus();
}

absl::Status Rotate::ValidateOutput(const Buffer& input) const {
  if (!input.IsBufferFormatCompatible(*output_)) {
    return absl::InvalidArgumentError("Rotate: output buffer format type is not compatible.");
  }
  if (angle_deg_ > 360 || angle_deg_ < 0 || angle_deg_ % 90 != 0) {
    return absl::InvalidArgumentError(absl::StrFormat(
        "Rotation angle (%d) must be between 0 and 360, in multiples of 90 degrees.", angle_deg_));
  }
  const std::vector<size_t>& in = input.GetDimension();
  const std::vector<size_t>& out = output_->GetDimension();
  const bool fits = QuarterTurn(angle_deg_) ? (in[0] == out[1] && in[1] == out[0]) : (in == out);
  if (!fits) return absl::InvalidArgumentError("Output buffer has invalid dimensions for rotation.");
  return absl::OkStatus();
}

absl::Status Rotate::CreateOutput(const Buffer& input) {
  const bool turn = QuarterTurn(angle_deg_);
  const size_t w = input.GetDimension()[turn ? 1 : 0], h = input.GetDimension()[turn ? 0 : 1];
  Buffer* created = EmptyLike(input, w, h, input.GetBufferFormat(), input.GetDataType(), OutputSpace());
  if (created == nullptr) return absl::InternalError("Rotate: failed to create the output buffer");
  OwnOutput(created);
  return absl::OkStatus();
}

absl::Status Rotate::Lower(cuda::ChainLowering& lowering, const Buffer& input) const {
  if (!LayoutsAgree(input, *GetOutput()))
    return absl::InvalidArgumentError("Rotate: input and output buffer formats must match");
  return lowering.Rotate(angle_deg_);
}

// ---- Flip (image_operator.cc:299-357) ----------------------------------------
absl::Status Flip::ValidateInput(const Buffer& input) const {
  if (!horizontal_ && !vertical_) {
    return absl::InvalidArgumentError("Flip: either horizontal or vertical flip must be enabled.");
  }
  if (IsRaw(input)) return RawNotSupported("Flip");
  return absl::OkStatus();
}

absl::Status Flip::ValidateOutput(const Buffer& input) const {
  if (!input.IsBufferFormatCompatible(*output_)) {
    return absl::InvalidArgumentError("Flip: output buffer format type is not compatible.");
  }
  if (input.GetDimension() != output_->GetDimension()) {
    return absl::InvalidArgumentError("Flip: input and output buffer dimensions must be same.");
  }
  return absl::OkStatus();
}

absl::Status Flip::CreateOutput(const Buffer& input) {
  // (the reference also allocates an intermediate for h+v; one index map needs none)
  Buffer* created = EmptyLike(input, input.GetDimension()[0], input.GetDimension()[1], input.GetBufferFormat(),
                              input.GetDataType(), OutputSpace());
  if (created == nullptr) return absl::InternalError("Flip: failed to create the output buffer");
  OwnOutput(created);
  return absl::OkStatus();
}

absl::Status Flip::Lower(cuda::ChainLowering& lowering, const Buffer& input) const {
  if (!LayoutsAgree(input, *GetOutput()))
    return absl::InvalidArgumentError("Flip: input and output buffer formats must match");
  return lowering.Flip(horizontal_, vertical_);
}

// ---- AutoOrient (additive; applies band/common.h:146-161 BufferOrientation) ----
void AutoOrient::Decompose(BufferOrientation orientation, bool* mirror, int* angle_deg) {
  // Rotate() is counter-clockwise (libyuv_image_operator.cc:697-708). Derivations: transpose out[r][c] = in[c][r] is a
  // horizontal mirror followed by Rotate(90); transverse out[r][c] = in[H-1-c][W-1-r] a mirror followed by Rotate(270).
  switch (orientation) {
    case BufferOrientation::kTopRight: *mirror = true; *angle_deg = 0; break;
    case BufferOrientation::kBottomRight: *mirror = false; *angle_deg = 180; break;
    case BufferOrientation::kBottomLeft: *mirror = true; *angle_deg = 180; break;   // == vertical mirror
    case BufferOrientation::kLeftTop: *mirror = true; *angle_deg = 90; break;       // transpose
    case BufferOrientation::kRightTop: *mirror = false; *angle_deg = 270; break;    // quarter-turn clockwise
    case BufferOrientation::kRightBottom: *mirror = true; *angle_deg = 270; break;  // transverse
    case BufferOrientation::kLeftBottom: *mirror = false; *angle_deg = 90; break;   // quarter-turn counter-clockwise
    default: *mirror = false; *angle_deg = 0; break;
  }
}

absl::Status AutoOrient::ValidateInput(const Buffer& input) const {
  if (IsRaw(input)) return RawNotSupported("AutoOrient");
  const size_t o = static_cast<size_t>(input.GetOrientation());
  if (o < 1 || o > 8) return absl::InvalidArgumentError(absl::StrFormat("AutoOrient: invalid orientation %d", static_cast<int>(o)));
  return absl::OkStatus();
}

absl::Status AutoOrient::ValidateOutput(const Buffer& input) const {
  if (!input.IsBufferFormatCompatible(*output_))
    return absl::InvalidArgumentError("AutoOrient: output buffer format type is not compatible.");
  bool mirror;
  int angle;
  Decompose(input.GetOrientation(), &mirror, &angle);
  const std::vector<size_t>& in = input.GetDimension();
  const std::vector<size_t>& out = output_->GetDimension();
  const bool fits = QuarterTurn(angle) ? (in[0] == out[1] && in[1] == out[0]) : (in == out);
  if (!fits) return absl::InvalidArgumentError("AutoOrient: output buffer has invalid dimensions.");
  return absl::OkStatus();
}

absl::Status AutoOrient::CreateOutput(const Buffer& input) {
  bool mirror;
  int angle;
  Decompose(input.GetOrientation(), &mirror, &angle);
  const bool turn = QuarterTurn(angle);
  const size_t w = input.GetDimension()[turn ? 1 : 0], h = input.GetDimension()[turn ? 0 : 1];
  Buffer* created = Buffer::CreateEmpty(w, h, input.GetBufferFormat(), input.GetDataType(), BufferOrientation::kTopLeft,
                                        OutputSpace());
  if (created == nullptr) return absl::InternalError("AutoOrient: failed to create the output buffer");
  OwnOutput(created);
  return absl::OkStatus();
}

absl::Status AutoOrient::Lower(cuda::ChainLowering& lowering, const Buffer& input) const {
  if (!LayoutsAgree(input, *GetOutput()))
    return absl::InvalidArgumentError("AutoOrient: input and output buffer formats must match");
  bool mirror;
  int angle;
  Decompose(input.GetOrientation(), &mirror, &angle);
  if (mirror) RETURN_IF_ERROR(lowering.Flip(true, false));
  if (angle != 0) RETURN_IF_ERROR(lowering.Rotate(angle));
  return absl::OkStatus();
}

// ---- ColorSpaceConvert (image_operator.cc:359-419) ---------------------------
absl::Status ColorSpaceConvert::ValidateOutput(const Buffer& input) const {
  const BufferFormat in = input.GetBufferFormat(), out = output_->GetBufferFormat();
  if (in == out) {
    return absl::InvalidArgumentError("ColorSpaceConvert: input and output buffer formats must be different.");
  }
  if (input.GetDataType() != output_->GetDataType()) {
    return absl::InvalidArgumentError("Data types must be the same.");
  }
  if (in == BufferFormat::kGrayScale) {
    return absl::InvalidArgumentError("Grayscale format does not convert to other formats.");
  }
  if (in == BufferFormat::kRGB && out == BufferFormat::kRGBA) {
    return absl::InvalidArgumentError("RGB format does not convert to RGBA");
  }
  if (in == BufferFormat::kRGB || in == BufferFormat::kRGBA || Buffer::IsYUV(in)) return absl::OkStatus();
  return absl::InternalError(absl::StrFormat("Unsupported buffer format: %s.", ToString(in)));
}

absl::Status ColorSpaceConvert::CreateOutput(const Buffer& input) {
  if (!is_format_specified_) return absl::InvalidArgumentError("Convert: output buffer format is not set.");
  const DataType dtype = Buffer::IsYUV(output_format_) ? DataType::kUInt8 : input.GetDataType();
  Buffer* created =
      EmptyLike(input, input.GetDimension()[0], input.GetDimension()[1], output_format_, dtype, OutputSpace());
  if (created == nullptr) return absl::InternalError("ColorSpaceConvert: failed to create the output buffer");
  OwnOutput(created);
  return absl::OkStatus();
}

absl::Status ColorSpaceConvert::Lower(cuda::ChainLowering& lowering, const Buffer& input) const {
  // libyuv_image_operator.cc:1709-1727: dimensions must agree
  if (input.GetDimension() != GetOutput()->GetDimension())
    return absl::InvalidArgumentError("ColorSpaceConvert: input and output buffer dimensions must be the same.");
  return lowering.ConvertColor(GetOutput()->GetBufferFormat());
}

// ---- AutoConvert (image_operator.cc:421-606) ---------------------------------
// Copies carry no state: the nested operators are rebuilt against the next
// input (the reference aliases them between copies).
AutoConvert::AutoConvert(const AutoConvert& rhs) : IBufferOperator(rhs) {}
AutoConvert& AutoConvert::operator=(const AutoConvert& rhs) {
  IBufferOperator::operator=(rhs);
  color_space_convert_.reset();
  resize_.reset();
  data_type_convert_.reset();
  return *this;
}
AutoConvert::~AutoConvert() = default;

// image_operator.cc:580-587: does NOT mark the output as assigned.
void AutoConvert::SetOutput(Buffer* output) {
  if (output) {
    owned_output_.reset();
    output_ = output;
  }
}

bool AutoConvert::RequiresColorSpaceConvert(const Buffer& input) const {
  return input.GetBufferFormat() != output_->GetBufferFormat();
}
bool AutoConvert::RequiresResize(const Buffer& input) const { return input.GetDimension() != output_->GetDimension(); }
bool AutoConvert::RequiresDataTypeConvert(const Buffer& input) const {
  return input.GetDataType() != output_->GetDataType();
}

absl::Status AutoConvert::ValidateInput(const Buffer& input) const {
  if (Buffer::IsYUV(input.GetBufferFormat())) {
    return absl::InvalidArgumentError(absl::StrFormat("input buffer format %d is not supported.",
                                                      static_cast<int>(input.GetBufferFormat())));
  }
  return absl::OkStatus();
}

// image_operator.cc:494-515: the nested operators needed for this input exist and have outputs.
absl::Status AutoConvert::ValidateOutput(const Buffer& input) const {
  if (output_ == nullptr) return absl::InternalError("Null output buffer");
  if (RequiresColorSpaceConvert(input) && (!color_space_convert_ || !color_space_convert_->GetOutput())) {
    return absl::InvalidArgumentError("color_space_convert_ output is nullptr.");
  }
  if (RequiresResize(input) && (!resize_ || !resize_->GetOutput())) {
    return absl::InvalidArgumentError("resize_ output is nullptr.");
  }
  if (RequiresDataTypeConvert(input) && (!data_type_convert_ || !data_type_convert_->GetOutput())) {
    return absl::InvalidArgumentError("data_type_convert_ output is nullptr.");
  }
  return absl::OkStatus();
}

// image_operator.cc:517-578: build the nested chain for this input, hand the
// assigned output to its last stage, then validate stage by stage.
absl::Status AutoConvert::CreateOutput(const Buffer& input) {
  if (output_ == nullptr) return absl::InternalError("Null output buffer");
  IBufferOperator* last = nullptr;
  const Buffer* current = &input;
  if (RequiresColorSpaceConvert(*current)) {
    color_space_convert_ = std::make_unique<ColorSpaceConvert>(output_->GetBufferFormat());
    InheritOutputSpace(*this, *color_space_convert_);
    RETURN_IF_ERROR(color_space_convert_->CreateOutput(*current));
    current = color_space_convert_->GetOutput();
    last = color_space_convert_.get();
  }
  if (RequiresResize(*current)) {
    resize_ = std::make_unique<Resize>(static_cast<int>(output_->GetDimension()[0]),
                                       static_cast<int>(output_->GetDimension()[1]));
    InheritOutputSpace(*this, *resize_);
    RETURN_IF_ERROR(resize_->CreateOutput(*current));
    current = resize_->GetOutput();
    last = resize_.get();
  }
  if (RequiresDataTypeConvert(*current)) {
    data_type_convert_ = std::make_unique<DataTypeConvert>();
    InheritOutputSpace(*this, *data_type_convert_);
    last = data_type_convert_.get();
  }
  if (last) last->SetOutput(output_);

  current = &input;
  if (RequiresColorSpaceConvert(*current)) {
    if (!color_space_convert_->GetOutput()) return absl::InvalidArgumentError("color_space_convert_ output is nullptr.");
    RETURN_IF_ERROR(color_space_convert_->ValidateOutput(*current));
    current = color_space_convert_->GetOutput();
  }
  if (RequiresResize(*current)) {
    if (!resize_->GetOutput()) return absl::InvalidArgumentError("resize_ output is nullptr.");
    RETURN_IF_ERROR(resize_->ValidateOutput(*current));
    current = resize_->GetOutput();
  }
  if (RequiresDataTypeConvert(*current)) {
    RETURN_IF_ERROR(data_type_convert_->ValidateOutput(*current));
  }
  return absl::OkStatus();
}

// image_operator.cc:465-492. With nothing to do the input is copied to the output.
absl::Status AutoConvert::Lower(cuda::ChainLowering& lowering, const Buffer& input) const {
  const Buffer* current = &input;
  if (RequiresColorSpaceConvert(*current)) {
    RETURN_IF_ERROR(lowering.LowerNested(*color_space_convert_, *current));
    current = color_space_convert_->GetOutput();
  }
  if (RequiresResize(*current)) {
    RETURN_IF_ERROR(lowering.LowerNested(*resize_, *current));
    current = resize_->GetOutput();
  }
  if (RequiresDataTypeConvert(*current)) {
    RETURN_IF_ERROR(lowering.LowerNested(*data_type_convert_, *current));
  }
  return absl::OkStatus();
}

}  // namespace buffer
}  // namespace band
