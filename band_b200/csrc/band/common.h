// band/common.h -- the thin slice of the reference's band/common.h the
// preprocessing path needs: enums with the reference's numeric values
// (/root/reference/band/common.h:71-161), ToString, GetDataTypeBytes
// (band/common.cc:338-367), Quantization (common.h:246-262), RETURN_IF_ERROR
// (common.h:388-394). Engine/scheduler types are out of scope.
#ifndef BAND_COMMON_H_
#define BAND_COMMON_H_

#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>

namespace band {
typedef int CallbackId;

enum class LogSeverity : size_t { kInternal = 0, kInfo, kWarning, kError };

enum class DataType : size_t {
  kNoType = 0,
  kFloat32,
  kInt32,
  kUInt8,
  kInt64,
  kString,
  kBool,
  kInt16,
  kComplex64,
  kInt8,
  kFloat16,
  kFloat64,
};

size_t GetDataTypeBytes(DataType type);

enum class BufferFormat : size_t {
  kGrayScale = 0,
  kRGB,
  kRGBA,
  kYV12,
  kYV21,
  kNV21,
  kNV12,
  kRaw
};

enum class BufferOrientation : size_t {
  kTopLeft = 1,
  kTopRight,
  kBottomRight,
  kBottomLeft,
  kLeftTop,
  kRightTop,
  kRightBottom,
  kLeftBottom,
};

enum class QuantizationType : size_t { kNoQuantization = 0, kAffineQuantization };

// Additive (not in the reference): where a Buffer's / Tensor's bytes live.
enum class MemorySpace : size_t {
  kHost = 0,    // ordinary host memory (the reference's only case)
  kDevice = 1,  // CUDA device memory: the operator chain reads/writes it in place
  kVirtual = 2  // shape-only intermediate of a fused plan (never materialised)
};

const char* ToString(LogSeverity v);
const char* ToString(DataType v);
const char* ToString(BufferFormat v);
const char* ToString(BufferOrientation v);
const char* ToString(QuantizationType v);
const char* ToString(MemorySpace v);
std::string ToString(const std::vector<size_t>& dims);

struct AffineQuantizationParams {
  std::vector<float> scale;
  std::vector<int32_t> zero_point;
  int32_t quantized_dimension;
};

class Quantization {
 public:
  Quantization(QuantizationType type, void* params) : type_(type), params_(params) {}
  QuantizationType GetType() { return type_; }
  void* GetParams() { return params_; }
  void SetParams(void* params) { params_ = params; }

 private:
  QuantizationType type_;
  void* params_;
};
}  // namespace band

#define RETURN_IF_ERROR(expr) \
  {                           \
    auto status = (expr);     \
    if (!status.ok()) {       \
      return status;          \
    }                         \
  }

#endif  // BAND_COMMON_H_
