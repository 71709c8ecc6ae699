// band/buffer/buffer.cc -- see buffer.h. Layout rules follow
// /root/reference/band/buffer/buffer.cc (cited per function).
#include "band/buffer/buffer.h"

#include <cstring>

#include "absl/strings/str_format.h"
#include "band/logger.h"
#include "bandcu.h"

namespace band {
namespace {
// Storage for an owning buffer, released by the space it came from.
std::shared_ptr<unsigned char> Allocate(size_t bytes, MemorySpace* space) {
  if (*space == MemorySpace::kDevice) {
    void* p = nullptr;
    if (bandcu_malloc(&p, bytes) != BANDCU_OK) return nullptr;
    return std::shared_ptr<unsigned char>(static_cast<unsigned char*>(p), [](unsigned char* q) { bandcu_free(q); });
  }
  // host: pinned when the CUDA runtime is usable, so uploads/downloads can stream
  void* p = nullptr;
  if (bandcu_host_alloc(&p, bytes) == BANDCU_OK && p != nullptr) {
    return std::shared_ptr<unsigned char>(static_cast<unsigned char*>(p),
                                          [](unsigned char* q) { bandcu_host_free(q); });
  }
  return std::shared_ptr<unsigned char>(new unsigned char[bytes], [](unsigned char* q) { delete[] q; });
}
}  // namespace

namespace {
std::atomic<uint64_t> g_buffer_epoch{1};
}
uint64_t Buffer::Epoch() { return g_buffer_epoch.load(std::memory_order_relaxed); }
void Buffer::SetOrientation(BufferOrientation orientation) {
  orientation_ = orientation;
  g_buffer_epoch.fetch_add(1, std::memory_order_relaxed);
}

Buffer::~Buffer() { g_buffer_epoch.fetch_add(1, std::memory_order_relaxed); }

Buffer::Buffer(std::vector<size_t> dimension, std::vector<DataPlane> data_planes, BufferFormat buffer_format,
               DataType data_type, BufferOrientation orientation, MemorySpace space)
    : dimension_(std::move(dimension)),
      data_planes_(std::move(data_planes)),
      buffer_format_(buffer_format),
      data_type_(data_type),
      orientation_(orientation),
      space_(space) {
  g_buffer_epoch.fetch_add(1, std::memory_order_relaxed);
}

// buffer.cc:16-22
Buffer* Buffer::CreateFromPlanes(const std::vector<DataPlane>& data_planes, const std::vector<size_t>& dims,
                                 BufferFormat buffer_format, BufferOrientation orientation, MemorySpace space) {
  return new Buffer(dims, data_planes, buffer_format, DataType::kUInt8, orientation, space);
}

// buffer.cc:24-78: one contiguous frame -> planes.
Buffer* Buffer::CreateFromRaw(const unsigned char* data, size_t width, size_t height, BufferFormat buffer_format,
                              DataType data_type, BufferOrientation orientation, bool owns_data,
                              MemorySpace space) {
  Buffer* b = nullptr;
  if (buffer_format <= BufferFormat::kRGBA) {
    const size_t ch = GetNumPixelElements(buffer_format);
    b = new Buffer({width, height}, {DataPlane{data, width * ch, ch}}, buffer_format, data_type, orientation, space);
  } else {
    const size_t uvw = (width + 1) / 2, uvh = (height + 1) / 2;
    const size_t nv_stride = uvw * 2;  // == width when even (buffer.cc:38,46)
    const unsigned char* c0 = data + width * height;
    switch (buffer_format) {
      case BufferFormat::kNV21:  // VU interleaved: v first
        b = CreateFromYUVPlanes(data, c0 + 1, c0, width, height, width, nv_stride, 2, buffer_format, orientation,
                                false, space);
        break;
      case BufferFormat::kNV12:  // UV interleaved: u first
        b = CreateFromYUVPlanes(data, c0, c0 + 1, width, height, width, nv_stride, 2, buffer_format, orientation,
                                false, space);
        break;
      case BufferFormat::kYV21:  // planar, u then v
        b = CreateFromYUVPlanes(data, c0, c0 + uvw * uvh, width, height, width, uvw, 1, buffer_format, orientation,
                                false, space);
        break;
      case BufferFormat::kYV12:  // planar, v then u
        b = CreateFromYUVPlanes(data, c0 + uvw * uvh, c0, width, height, width, uvw, 1, buffer_format, orientation,
                                false, space);
        break;
      default:
        BAND_LOG(LogSeverity::kError, "Unsupported format type : %s", ToString(buffer_format));
        return nullptr;
    }
  }
  if (b && owns_data && data) {
    unsigned char* p = const_cast<unsigned char*>(data);
    if (space == MemorySpace::kDevice) b->storage_.reset(p, [](unsigned char* q) { bandcu_free(q); });
    else b->storage_.reset(p, [](unsigned char* q) { delete[] q; });  // the reference's contract: new[]'d by the caller
  }
  return b;
}

// buffer.cc:80-104: plane order stored is {Y,V,U} for NV21/YV12 and {Y,U,V} for NV12/YV21.
Buffer* Buffer::CreateFromYUVPlanes(const unsigned char* y_data, const unsigned char* u_data,
                                    const unsigned char* v_data, size_t width, size_t height, size_t row_stride_y,
                                    size_t row_stride_uv, size_t pixel_stride_uv, BufferFormat buffer_format,
                                    BufferOrientation orientation, bool owns_data, MemorySpace space) {
  std::vector<DataPlane> planes;
  if (buffer_format == BufferFormat::kNV21 || buffer_format == BufferFormat::kYV12) {
    planes = {{y_data, row_stride_y, 1}, {v_data, row_stride_uv, pixel_stride_uv}, {u_data, row_stride_uv, pixel_stride_uv}};
  } else if (buffer_format == BufferFormat::kNV12 || buffer_format == BufferFormat::kYV21) {
    planes = {{y_data, row_stride_y, 1}, {u_data, row_stride_uv, pixel_stride_uv}, {v_data, row_stride_uv, pixel_stride_uv}};
  } else {
    BAND_LOG(LogSeverity::kError, "Unsupported YUV format type : %s", ToString(buffer_format));
    return nullptr;
  }
  Buffer* b = new Buffer({width, height}, planes, buffer_format, DataType::kUInt8, orientation, space);
  if (owns_data && y_data) {
    unsigned char* p = const_cast<unsigned char*>(y_data);
    if (space == MemorySpace::kDevice) b->storage_.reset(p, [](unsigned char* q) { bandcu_free(q); });
    else b->storage_.reset(p, [](unsigned char* q) { delete[] q; });
  }
  return b;
}

// buffer.cc:106-152: view over a tensor. Dims equal to 1 are dropped; [H,W,3] is
// an RGB image {W,H} whose row stride is W*3 (the reference does not scale it by
// the element size; the GPU path treats tensors as contiguous NHWC anyway);
// anything else is a single-row kRaw plane.
Buffer* Buffer::CreateFromTensor(const interface::ITensor* tensor) {
  if (tensor == nullptr) {
    BAND_LOG(LogSeverity::kError, "Given tensor is null");
    return nullptr;
  }
  if (tensor->GetNumDims() == 0) {
    BAND_LOG(LogSeverity::kError, "Given tensor has no dimension");
    return nullptr;
  }
  std::vector<size_t> dims;
  for (size_t i = 0; i < tensor->GetNumDims(); ++i) {
    const int d = tensor->GetDims()[i];
    if (d == 1) continue;
    if (d <= 0) {
      BAND_LOG(LogSeverity::kError, "Given tensor has invalid dimension : %d", d);
      return nullptr;
    }
    dims.push_back(static_cast<size_t>(d));
  }
  const unsigned char* data = reinterpret_cast<const unsigned char*>(tensor->GetData());
  const size_t esz = GetDataTypeBytes(tensor->GetType());
  const MemorySpace space = tensor->GetMemorySpace();
  if (dims.size() == 3 && dims[2] == 3) {
    const size_t w = dims[1], h = dims[0];
    return new Buffer({w, h}, {DataPlane{data, w * 3, 3 * esz}}, BufferFormat::kRGB, tensor->GetType(),
                      BufferOrientation::kTopLeft, space);
  }
  return new Buffer(dims, {DataPlane{data, tensor->GetBytes(), esz}}, BufferFormat::kRaw, tensor->GetType(),
                    BufferOrientation::kTopLeft, space);
}

// buffer.cc:154-197. Deliberate deviation (memory safety):
// the reference sizes RGB/RGBA storage as W*H*channels BYTES whatever the
// element type; here the element size is honoured so a float intermediate fits.
Buffer* Buffer::CreateEmpty(size_t width, size_t height, BufferFormat buffer_format, DataType data_type,
                            BufferOrientation orientation, MemorySpace space) {
  size_t total = 0;
  switch (buffer_format) {
    case BufferFormat::kRGB:
    case BufferFormat::kRGBA: {
      const size_t esz = GetDataTypeBytes(data_type);
      total = width * height * GetNumPixelElements(buffer_format) * (esz ? esz : 1);
      break;
    }
    case BufferFormat::kGrayScale:
      total = width * height;
      data_type = DataType::kUInt8;
      break;
    case BufferFormat::kNV21:
    case BufferFormat::kNV12:
    case BufferFormat::kYV21:
    case BufferFormat::kYV12:
      total = GetBufferByteSize({width, height}, buffer_format);
      data_type = DataType::kUInt8;
      break;
    case BufferFormat::kRaw:
      BAND_LOG(LogSeverity::kError, "Raw format type requires external input to create empty buffer");
      return nullptr;
    default:
      BAND_LOG(LogSeverity::kError, "Unsupported format type : %s", ToString(buffer_format));
      return nullptr;
  }
  if (total == 0) return nullptr;
  if (space == MemorySpace::kVirtual) {
    // shape only: plane strides are real, data pointers are null
    return CreateFromRaw(nullptr, width, height, buffer_format, data_type, orientation, false, space);
  }
  std::shared_ptr<unsigned char> mem = Allocate(total, &space);
  if (!mem) {
    BAND_LOG(LogSeverity::kError, "Failed to allocate %zu bytes in %s memory", total, ToString(space));
    return nullptr;
  }
  Buffer* b = CreateFromRaw(mem.get(), width, height, buffer_format, data_type, orientation, false, space);
  if (b) b->storage_ = std::move(mem);
  return b;
}

size_t Buffer::GetNumPixelElements(BufferFormat buffer_format) {
  switch (buffer_format) {
    case BufferFormat::kGrayScale: return 1;
    case BufferFormat::kRGB: return 3;
    case BufferFormat::kRGBA: return 4;
    default: return 0;  // buffer.cc:216-231: YUV / Raw have no per-pixel element count
  }
}

std::vector<size_t> Buffer::GetUvDims(const std::vector<size_t>& dims, BufferFormat buffer_format) {
  if (dims.size() != 2 || dims[0] == 0 || dims[1] == 0) {
    BAND_LOG(LogSeverity::kError, "Given dims is not valid for UV plane : %s", ToString(dims).c_str());
    return {};
  }
  if (!IsYUV(buffer_format)) {
    BAND_LOG(LogSeverity::kError, "Given format type is not YUV format : %s", ToString(buffer_format));
    return {};
  }
  return {(dims[0] + 1) / 2, (dims[1] + 1) / 2};
}

size_t Buffer::GetBufferByteSize(const std::vector<size_t>& dims, BufferFormat buffer_format) {
  if (IsYUV(buffer_format)) {
    std::vector<size_t> uv = GetUvDims(dims, buffer_format);
    if (uv.empty()) return 0;
    return GetSize(dims) + GetSize(uv) * 2;
  }
  return GetSize(dims) * GetNumPixelElements(buffer_format);
}

// inclusive box (buffer.cc:268-271)
std::vector<size_t> Buffer::GetCropDimension(size_t x0, size_t x1, size_t y0, size_t y1) {
  return {x1 - x0 + 1, y1 - y0 + 1};
}

size_t Buffer::GetSize(const std::vector<size_t>& dims) {
  size_t size = 1;
  for (size_t d : dims) {
    if (d == 0) return 0;
    size *= d;
  }
  return size;
}

const std::vector<size_t>& Buffer::GetDimension() const { return dimension_; }
const Buffer::DataPlane& Buffer::operator[](size_t index) const { return data_planes_[index]; }
Buffer::DataPlane& Buffer::operator[](size_t index) {
  g_buffer_epoch.fetch_add(1, std::memory_order_relaxed);  // the caller may re-point the plane
  return data_planes_[index];
}
size_t Buffer::GetNumPlanes() const { return data_planes_.size(); }
size_t Buffer::GetNumElements() const { return GetSize(dimension_); }
size_t Buffer::GetPixelBytes() const {
  if (buffer_format_ == BufferFormat::kRaw) return data_planes_[0].pixel_stride_bytes;
  return GetNumPixelElements(buffer_format_) * GetDataTypeBytes(data_type_);
}
size_t Buffer::GetNumPixelElements() const { return GetNumPixelElements(buffer_format_); }
size_t Buffer::GetBytes() const { return GetPixelBytes() * GetNumElements(); }
DataType Buffer::GetDataType() const { return data_type_; }
BufferFormat Buffer::GetBufferFormat() const { return buffer_format_; }
BufferOrientation Buffer::GetOrientation() const { return orientation_; }

bool Buffer::IsPinnedHost() const {
  if (pinned_cache_ < 0) {
    int pinned = (space_ == MemorySpace::kHost && !data_planes_.empty()) ? 1 : 0;
    for (size_t k = 0; pinned && k < data_planes_.size(); ++k)
      if (data_planes_[k].data == nullptr || bandcu_pointer_kind(data_planes_[k].data) != 1) pinned = 0;
    pinned_cache_ = pinned;
  }
  return pinned_cache_ == 1;
}

bool Buffer::IsYUV(BufferFormat f) {
  return f == BufferFormat::kNV21 || f == BufferFormat::kNV12 || f == BufferFormat::kYV21 || f == BufferFormat::kYV12;
}

// buffer.cc:367-388
bool Buffer::IsBufferFormatCompatible(const Buffer& rhs) const {
  if (data_type_ != rhs.data_type_) return false;
  if (buffer_format_ == BufferFormat::kRGB || buffer_format_ == BufferFormat::kRGBA)
    return rhs.buffer_format_ == BufferFormat::kRGB || rhs.buffer_format_ == BufferFormat::kRGBA;
  if (IsYUV(buffer_format_)) return IsYUV(rhs.buffer_format_);
  return buffer_format_ == rhs.buffer_format_;
}

// buffer.cc:331-358: same dtype/format/dims, single plane, not YUV. Crosses
// memory spaces (host<->device) through the CUDA shim.
absl::Status Buffer::CopyFrom(const Buffer& rhs) {
  if (data_type_ != rhs.data_type_)
    return absl::InvalidArgumentError(absl::StrFormat("Buffer::CopyFrom: data type is not compatible : %s %s",
                                                      ToString(data_type_), ToString(rhs.data_type_)));
  if (buffer_format_ != rhs.buffer_format_)
    return absl::InvalidArgumentError(absl::StrFormat("Buffer::CopyFrom: buffer format is not compatible : %s %s",
                                                      ToString(buffer_format_), ToString(rhs.buffer_format_)));
  if (dimension_ != rhs.dimension_)
    return absl::InvalidArgumentError(absl::StrFormat("Buffer::CopyFrom: dimension is not compatible : %s %s",
                                                      ToString(dimension_), ToString(rhs.dimension_)));
  if (IsYUV(buffer_format_)) return absl::InvalidArgumentError("Buffer::CopyFrom: YUV format is not supported");
  if (space_ == MemorySpace::kVirtual || rhs.space_ == MemorySpace::kVirtual)
    return absl::InternalError("Buffer::CopyFrom: virtual buffers have no storage");
  unsigned char* dst = data_planes_[0].GetMutableData();
  const unsigned char* src = rhs.data_planes_[0].data;
  const size_t bytes = GetBytes();
  if (space_ == MemorySpace::kHost && rhs.space_ == MemorySpace::kHost) {
    memcpy(dst, src, bytes);
    return absl::OkStatus();
  }
  int e;
  if (space_ == MemorySpace::kDevice && rhs.space_ == MemorySpace::kDevice) e = bandcu_memcpy_d2d(dst, src, bytes, nullptr);
  else if (space_ == MemorySpace::kDevice) e = bandcu_memcpy_h2d(dst, src, bytes, nullptr);
  else e = bandcu_memcpy_d2h(dst, src, bytes, nullptr);
  if (e == BANDCU_OK) e = bandcu_stream_sync(nullptr);
  if (e != BANDCU_OK) return absl::InternalError(bandcu_last_error_string());
  return absl::OkStatus();
}

std::ostream& operator<<(std::ostream& os, const Buffer& buffer) {
  return os << "Buffer: " << ToString(buffer.GetBufferFormat()) << " " << ToString(buffer.GetDataType()) << " "
            << ToString(buffer.GetOrientation()) << " " << ToString(buffer.GetDimension()) << " "
            << buffer.GetNumPlanes() << " planes in " << ToString(buffer.GetMemorySpace()) << " memory";
}
}  // namespace band
