#include "band/common.h"

#include <cstdio>

#include "band/logger.h"

namespace band {

size_t GetDataTypeBytes(DataType type) {
  switch (type) {
    case DataType::kNoType: return 0;
    case DataType::kFloat32: return sizeof(float);
    case DataType::kInt32: return sizeof(int32_t);
    case DataType::kUInt8: return sizeof(uint8_t);
    case DataType::kInt8: return sizeof(int8_t);
    case DataType::kInt16: return sizeof(int16_t);
    case DataType::kInt64: return sizeof(int64_t);
    case DataType::kString: return sizeof(char);
    case DataType::kBool: return sizeof(bool);
    case DataType::kComplex64: return sizeof(double);
    case DataType::kFloat16: return sizeof(float) / 2;
    case DataType::kFloat64: return sizeof(double);
    default: break;
  }
  return 0;
}

const char* ToString(LogSeverity v) {
  switch (v) {
    case LogSeverity::kInternal: return "INTERNAL";
    case LogSeverity::kInfo: return "INFO";
    case LogSeverity::kWarning: return "WARNING";
    case LogSeverity::kError: return "ERROR";
  }
  return "Unknown LogSeverity";
}
const char* ToString(DataType v) {
  switch (v) {
    case DataType::kNoType: return "NoType";
    case DataType::kFloat32: return "Float32";
    case DataType::kInt32: return "Int32";
    case DataType::kUInt8: return "UInt8";
    case DataType::kInt64: return "Int64";
    case DataType::kString: return "String";
    case DataType::kBool: return "Bool";
    case DataType::kInt16: return "Int16";
    case DataType::kComplex64: return "Complex64";
    case DataType::kInt8: return "Int8";
    case DataType::kFloat16: return "Float16";
    case DataType::kFloat64: return "Float64";
  }
  return "Unknown data type";
}
const char* ToString(BufferFormat v) {
  switch (v) {
    case BufferFormat::kGrayScale: return "GrayScale";
    case BufferFormat::kRGB: return "RGB";
    case BufferFormat::kRGBA: return "RGBA";
    case BufferFormat::kYV12: return "YV12";
    case BufferFormat::kYV21: return "YV21";
    case BufferFormat::kNV21: return "NV21";
    case BufferFormat::kNV12: return "NV12";
    case BufferFormat::kRaw: return "Raw";
  }
  return "Unknown format type";
}
const char* ToString(BufferOrientation v) {
  switch (v) {
    case BufferOrientation::kTopLeft: return "TopLeft";
    case BufferOrientation::kTopRight: return "TopRight";
    case BufferOrientation::kBottomRight: return "BottomRight";
    case BufferOrientation::kBottomLeft: return "BottomLeft";
    case BufferOrientation::kLeftTop: return "LeftTop";
    case BufferOrientation::kRightTop: return "RightTop";
    case BufferOrientation::kRightBottom: return "RightBottom";
    case BufferOrientation::kLeftBottom: return "LeftBottom";
  }
  return "Unknown buffer orientation";
}
const char* ToString(QuantizationType v) {
  switch (v) {
    case QuantizationType::kNoQuantization: return "NoQuantization";
    case QuantizationType::kAffineQuantization: return "AffineQuantization";
  }
  return "Unknown quantization type";
}
const char* ToString(MemorySpace v) {
  switch (v) {
    case MemorySpace::kHost: return "Host";
    case MemorySpace::kDevice: return "Device";
    case MemorySpace::kVirtual: return "Virtual";
  }
  return "Unknown memory space";
}
std::string ToString(const std::vector<size_t>& dims) {
  std::string s;
  for (size_t i = 0; i < dims.size(); ++i) {
    if (i) s += " x ";
    s += std::to_string(dims[i]);
  }
  return s;
}

Logger& Logger::Get() {
  static Logger* logger = new Logger();
  return *logger;
}
void Logger::SetVerbosity(LogSeverity severity) {
  std::lock_guard<std::mutex> l(mu_);
  verbosity_ = severity;
}
CallbackId Logger::SetReporter(std::function<void(LogSeverity, const char*)> reporter) {
  std::lock_guard<std::mutex> l(mu_);
  CallbackId id = next_callback_id_++;
  reporters_[id] = std::move(reporter);
  return id;
}
absl::Status Logger::RemoveReporter(CallbackId callback_id) {
  std::lock_guard<std::mutex> l(mu_);
  if (reporters_.erase(callback_id) == 0) return absl::NotFoundError("The given callback id does not exist.");
  return absl::OkStatus();
}
std::pair<LogSeverity, std::string> Logger::GetLastLog() const {
  std::lock_guard<std::mutex> l(mu_);
  return last_message_;
}
void Logger::Log(LogSeverity severity, const char* format, ...) {
  {
    std::lock_guard<std::mutex> l(mu_);
    if (static_cast<size_t>(severity) < static_cast<size_t>(verbosity_)) return;
  }
  char buf[1024];
  va_list args;
  va_start(args, format);
  vsnprintf(buf, sizeof(buf), format, args);
  va_end(args);
  std::lock_guard<std::mutex> l(mu_);
  last_message_ = {severity, buf};
  for (auto& r : reporters_) r.second(severity, buf);
  fprintf(stderr, "%s: %s\n", ToString(severity), buf);
}

}  // namespace band
