// band/interface/tensor.cc -- ITensor helpers
// (/root/reference/band/interface/tensor.cc:25-69); CopyDataFrom additionally
// crosses memory spaces through the CUDA shim.
#include "band/interface/tensor.h"

#include <cstring>

#include "bandcu.h"

namespace band {
namespace interface {

bool ITensor::operator==(const ITensor& rhs) const {
  return GetType() == rhs.GetType() && GetDimsVector() == rhs.GetDimsVector();
}
bool ITensor::operator!=(const ITensor& rhs) const { return !(*this == rhs); }

size_t ITensor::GetBytes() const { return GetDataTypeBytes(GetType()) * GetNumElements(); }

size_t ITensor::GetNumElements() const {
  size_t n = 1;
  for (size_t i = 0; i < GetNumDims(); ++i) n *= static_cast<size_t>(GetDims()[i]);
  return n;
}

std::vector<int> ITensor::GetDimsVector() const { return std::vector<int>(GetDims(), GetDims() + GetNumDims()); }

absl::Status ITensor::CopyDataFrom(const ITensor& rhs) {
  if (*this != rhs) return absl::InternalError("");
  const bool dst_dev = GetMemorySpace() == MemorySpace::kDevice, src_dev = rhs.GetMemorySpace() == MemorySpace::kDevice;
  if (!dst_dev && !src_dev) {
    memcpy(GetData(), rhs.GetData(), GetBytes());
    return absl::OkStatus();
  }
  int e = dst_dev && src_dev ? bandcu_memcpy_d2d(GetData(), rhs.GetData(), GetBytes(), nullptr)
          : dst_dev          ? bandcu_memcpy_h2d(GetData(), rhs.GetData(), GetBytes(), nullptr)
                             : bandcu_memcpy_d2h(GetData(), rhs.GetData(), GetBytes(), nullptr);
  if (e == BANDCU_OK) e = bandcu_stream_sync(nullptr);
  if (e != BANDCU_OK) return absl::InternalError(bandcu_last_error_string());
  return absl::OkStatus();
}

absl::Status ITensor::CopyDataFrom(const ITensor* rhs) {
  if (!rhs) return absl::InternalError("Tried to copy null tensor");
  return CopyDataFrom(*rhs);
}
}  // namespace interface
}  // namespace band
