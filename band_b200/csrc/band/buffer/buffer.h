// band/buffer/buffer.h -- band::Buffer with the reference's public surface
// (/root/reference/band/buffer/buffer.h:15-95) and one additive notion: WHERE
// the planes live (MemorySpace). A Buffer is a non-owning multi-plane view
// (1 plane for Gray/RGB/RGBA/Raw, 3 planes Y + two chroma for the 4:2:0
// formats); only CreateEmpty buffers own storage. Additions over the reference:
//   * MemorySpace tag (host / CUDA device / virtual = shape only);
//   * CreateEmpty(..., space): device- or pinned-host-backed intermediates;
//   * storage is reference counted and freed by the space it came from.
#ifndef BAND_BUFFER_BUFFER_H_
#define BAND_BUFFER_BUFFER_H_

#include <atomic>
#include <cstdint>
#include <memory>
#include <ostream>
#include <vector>

#include "absl/status/status.h"
#include "band/common.h"
#include "band/interface/tensor.h"

namespace band {

class Buffer {
 public:
  struct DataPlane {
    unsigned char* GetMutableData() { return const_cast<unsigned char*>(data); }
    const unsigned char* data;
    // bytes between two consecutive rows / two consecutive pixels of this plane
    size_t row_stride_bytes = 1;
    size_t pixel_stride_bytes = 1;
  };
  ~Buffer();

  static Buffer* CreateFromPlanes(const std::vector<DataPlane>& data_planes, const std::vector<size_t>& dims,
                                  BufferFormat buffer_format,
                                  BufferOrientation orientation = BufferOrientation::kTopLeft,
                                  MemorySpace space = MemorySpace::kHost);

  static Buffer* CreateFromRaw(const unsigned char* data, size_t width, size_t height, BufferFormat buffer_format,
                               DataType data_type = DataType::kUInt8,
                               BufferOrientation orientation = BufferOrientation::kTopLeft, bool owns_data = false,
                               MemorySpace space = MemorySpace::kHost);

  static Buffer* CreateFromYUVPlanes(const unsigned char* y_data, const unsigned char* u_data,
                                     const unsigned char* v_data, size_t width, size_t height, size_t row_stride_y,
                                     size_t row_stride_uv, size_t pixel_stride_uv, BufferFormat buffer_format,
                                     BufferOrientation orientation = BufferOrientation::kTopLeft,
                                     bool owns_data = false, MemorySpace space = MemorySpace::kHost);

  static Buffer* CreateFromTensor(const interface::ITensor* tensor);

  // space: kHost -> pinned host memory when a CUDA device is usable (so the
  // processor can stream it), plain heap otherwise; kDevice -> HBM on the
  // current device; kVirtual -> no storage (shape propagation inside a fused plan).
  static Buffer* CreateEmpty(size_t width, size_t height, BufferFormat buffer_format, DataType data_type,
                             BufferOrientation orientation = BufferOrientation::kTopLeft,
                             MemorySpace space = MemorySpace::kHost);

  const std::vector<size_t>& GetDimension() const;
  const DataPlane& operator[](size_t index) const;
  DataPlane& operator[](size_t index);
  size_t GetNumPlanes() const;
  size_t GetNumElements() const;
  size_t GetPixelBytes() const;
  size_t GetNumPixelElements() const;
  size_t GetBytes() const;
  DataType GetDataType() const;
  BufferFormat GetBufferFormat() const;
  BufferOrientation GetOrientation() const;
  // additive: the camera layer knows the sensor orientation only after the view exists (BandBufferSetOrientation)
  void SetOrientation(BufferOrientation orientation);
  // additive: every plane of this HOST buffer lies in pinned (page-locked, device-mapped) memory. Asked of the CUDA
  // runtime once per buffer and remembered: batch calls hand the same views over and over.
  bool IsPinnedHost() const;
  MemorySpace GetMemorySpace() const { return space_; }  // additive
  // additive: a process-wide counter that moves whenever ANY Buffer is created, destroyed or may have been changed in
  // place (non-const plane access, SetOrientation). While it stands still, a list of Buffer pointers names the same
  // views over the same memory as before, so a processor may replay its last batch without re-validating it.
  static uint64_t Epoch();

  absl::Status CopyFrom(const Buffer& rhs);

  bool IsBufferFormatCompatible(const Buffer& rhs) const;

  static std::vector<size_t> GetUvDims(const std::vector<size_t>& dims, BufferFormat buffer_format);
  static size_t GetBufferByteSize(const std::vector<size_t>& dims, BufferFormat buffer_format);
  static bool IsYUV(BufferFormat buffer_format);
  static std::vector<size_t> GetCropDimension(size_t x0, size_t x1, size_t y0, size_t y1);
  static size_t GetNumPixelElements(BufferFormat buffer_format);

 private:
  Buffer(std::vector<size_t> dimension, std::vector<DataPlane> data_planes, BufferFormat buffer_format,
         DataType data_type, BufferOrientation orientation, MemorySpace space);
  static size_t GetSize(const std::vector<size_t>& dims);

  const std::vector<size_t> dimension_;
  std::vector<DataPlane> data_planes_;
  BufferFormat buffer_format_;
  DataType data_type_ = DataType::kUInt8;
  BufferOrientation orientation_;
  MemorySpace space_ = MemorySpace::kHost;
  mutable int pinned_cache_ = -1;  // IsPinnedHost(): -1 unknown, 0 no, 1 yes (a view's planes never change)
  // owning buffers keep their allocation alive here; the deleter knows the space
  std::shared_ptr<unsigned char> storage_;
};

std::ostream& operator<<(std::ostream& os, const Buffer& buffer);

}  // namespace band

#endif  // BAND_BUFFER_BUFFER_H_
