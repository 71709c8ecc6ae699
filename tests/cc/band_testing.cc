// tests/cc/band_testing.cc -- TEST SUPPORT ONLY (not part of the product libraries).
//
// The reference's own tests drive the C++ classes directly
// (/root/reference/band/test/image_operation_test.cc: `Crop op(...);
// op.Process(*buffer); op.GetOutput()`), which the C API cannot express (it only
// writes into tensors). This file exposes exactly those C++ call sequences as
// plain C entry points so pytest can run the same scenarios against
// libband_c.so and compare with the oracle. It contains no pixel arithmetic.
#include <cstring>
#include <memory>
#include <vector>

#include "band/buffer/buffer.h"
#include "band/buffer/common_operator.h"
#include "band/buffer/image_operator.h"
#include "band/buffer/image_processor.h"

using namespace band;
using namespace band::buffer;

namespace {

struct OpDesc {  // same field meaning as oracle/band_chain.h orc_op_desc
  int kind;      // BandImageProcessorBuilderField order, 7 = AutoConvert
  int i[4];
  float f[2];
  int inplace;
};

std::unique_ptr<IBufferOperator> MakeOp(const OpDesc& d) {
  switch (d.kind) {
    case 0: return std::make_unique<Crop>(d.i[0], d.i[1], d.i[2], d.i[3]);
    case 1: return std::make_unique<Resize>(d.i[0], d.i[1]);
    case 2: return std::make_unique<Rotate>(d.i[0]);
    case 3: return std::make_unique<Flip>(d.i[0] != 0, d.i[1] != 0);
    case 4:
      if (d.i[0] < 0) return std::make_unique<ColorSpaceConvert>();
      return std::make_unique<ColorSpaceConvert>(static_cast<BufferFormat>(d.i[0]));
    case 5: return std::make_unique<Normalize>(d.f[0], d.f[1], d.inplace != 0);
    case 6: return std::make_unique<DataTypeConvert>();
    case 7: return std::make_unique<AutoConvert>();
    default: return nullptr;
  }
}

struct OutInfo {
  int width, height, format, dtype;
  size_t bytes;
};

// Copies a host buffer's contents plane by plane into `out` in CreateFromRaw order.
size_t Export(const Buffer& b, unsigned char* out, size_t cap, OutInfo* info) {
  info->width = static_cast<int>(b.GetDimension()[0]);
  info->height = b.GetDimension().size() > 1 ? static_cast<int>(b.GetDimension()[1]) : 1;
  info->format = static_cast<int>(b.GetBufferFormat());
  info->dtype = static_cast<int>(b.GetDataType());
  size_t bytes;
  if (Buffer::IsYUV(b.GetBufferFormat())) bytes = Buffer::GetBufferByteSize(b.GetDimension(), b.GetBufferFormat());
  else bytes = b.GetNumElements() * b.GetNumPixelElements() * GetDataTypeBytes(b.GetDataType());
  info->bytes = bytes;
  if (bytes <= cap && b.GetMemorySpace() == MemorySpace::kHost) {
    // owning buffers from CreateEmpty are one contiguous allocation starting at plane 0
    memcpy(out, b[0].data, bytes);
  }
  return bytes;
}

}  // namespace

extern "C" {

// IBufferOperator::Process stand-alone. If assign_format >= 0 an output buffer of
// (assign_w x assign_h, assign_format, assign_dtype) is created and SetOutput() first.
// Returns the absl status code; on OK the operator's output is exported.
__attribute__((visibility("default"))) int bandt_run_op(const OpDesc* desc, const unsigned char* frame, int w, int h,
                                                        int fmt, int in_dtype, int assign_format, int assign_dtype,
                                                        int assign_w, int assign_h, unsigned char* out,
                                                        size_t out_cap, OutInfo* info) {
  std::unique_ptr<IBufferOperator> op = MakeOp(*desc);
  if (!op) return -1;
  std::unique_ptr<Buffer> input(
      Buffer::CreateFromRaw(frame, w, h, static_cast<BufferFormat>(fmt), static_cast<DataType>(in_dtype)));
  if (!input) return -2;
  std::unique_ptr<Buffer> assigned;
  if (assign_format >= 0) {
    assigned.reset(Buffer::CreateEmpty(assign_w, assign_h, static_cast<BufferFormat>(assign_format),
                                       static_cast<DataType>(assign_dtype)));
    if (!assigned) return -3;
    op->SetOutput(assigned.get());
  }
  absl::Status st = op->Process(*input);
  if (!st.ok()) return static_cast<int>(st.code());
  if (op->GetOutput() && out) Export(*op->GetOutput(), out, out_cap, info);
  return 0;
}

// ImageProcessorBuilder -> Build -> Process(input, output) with an image-buffer
// output (any format), as C++ users of the reference do. build_status receives
// the Build() status code.
__attribute__((visibility("default"))) int bandt_run_chain(const OpDesc* descs, int n, const unsigned char* frame,
                                                           int w, int h, int fmt, int out_format, int out_dtype,
                                                           int out_w, int out_h, unsigned char* out, size_t out_cap,
                                                           OutInfo* info, int* build_status) {
  ImageProcessorBuilder builder;
  for (int k = 0; k < n; ++k) builder.AddOperation(MakeOp(descs[k]));
  absl::StatusOr<std::unique_ptr<BufferProcessor>> built = builder.Build();
  if (build_status) *build_status = static_cast<int>(built.status().code());
  if (!built.ok()) return static_cast<int>(built.status().code());
  std::unique_ptr<Buffer> input(Buffer::CreateFromRaw(frame, w, h, static_cast<BufferFormat>(fmt)));
  std::unique_ptr<Buffer> output(
      Buffer::CreateEmpty(out_w, out_h, static_cast<BufferFormat>(out_format), static_cast<DataType>(out_dtype)));
  if (!input || !output) return -2;
  absl::Status st = built.value()->Process(*input, *output);
  if (!st.ok()) return static_cast<int>(st.code());
  if (out) Export(*output, out, out_cap, info);
  return 0;
}

// Build() status only (duplicate detection etc.), no device needed.
__attribute__((visibility("default"))) int bandt_build_status(const OpDesc* descs, int n) {
  ImageProcessorBuilder builder;
  for (int k = 0; k < n; ++k) builder.AddOperation(MakeOp(descs[k]));
  return static_cast<int>(builder.Build().status().code());
}

}  // extern "C"
