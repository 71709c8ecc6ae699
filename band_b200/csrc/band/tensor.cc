// band/tensor.cc -- band::Tensor (/root/reference/band/tensor.cc:8-31) with an
// optional device backing (additive). Host tensors are PINNED when the CUDA
// runtime is usable so results stream back at full PCIe rate.
#include "band/tensor.h"

#include <cstdlib>
#include <cstring>

#include "band/logger.h"
#include "bandcu.h"

namespace band {
void Tensor::Allocate(size_t bytes) {
  data_ = nullptr;
  pinned_ = false;
  if (bytes == 0) return;
  void* p = nullptr;
  if (space_ == MemorySpace::kDevice) {
    if (bandcu_malloc(&p, bytes) != BANDCU_OK) {
      BAND_LOG(LogSeverity::kError, "Tensor: device allocation of %zu bytes failed: %s", bytes,
               bandcu_last_error_string());
      return;
    }
    data_ = static_cast<char*>(p);
    return;
  }
  if (bandcu_host_alloc(&p, bytes) == BANDCU_OK && p != nullptr) {
    data_ = static_cast<char*>(p);
    pinned_ = true;
    return;
  }
  data_ = new char[bytes];
}

Tensor::Tensor(interface::ITensor* tensor_view, bool copy_data)
    : type_(tensor_view->GetType()),
      quantization_({QuantizationType::kNoQuantization, nullptr}),
      dims_(tensor_view->GetDims(), tensor_view->GetDims() + tensor_view->GetNumDims()),
      name_(tensor_view->GetName()),
      space_(MemorySpace::kHost) {
  Allocate(tensor_view->GetBytes());
  absl::Status status = SetQuantization(tensor_view->GetQuantization());
  if (!status.ok()) BAND_LOG(LogSeverity::kError, "Failed to set quantization: %s", status.message().c_str());
  if (copy_data && data_ && tensor_view->GetMemorySpace() == MemorySpace::kHost)
    memcpy(data_, tensor_view->GetData(), tensor_view->GetBytes());
}

Tensor::Tensor(DataType type, const std::vector<int>& dims, const std::string& name, MemorySpace space)
    : type_(type), quantization_({QuantizationType::kNoQuantization, nullptr}), dims_(dims), name_(name), space_(space) {
  Allocate(GetBytes());
}

Tensor::Tensor(const interface::ITensor* like, MemorySpace space, char* external)
    : type_(like->GetType()),
      quantization_({QuantizationType::kNoQuantization, nullptr}),
      dims_(like->GetDims(), like->GetDims() + like->GetNumDims()),
      data_(external),
      name_(like->GetName()),
      space_(space),
      owns_(false) {
  absl::Status status = SetQuantization(like->GetQuantization());
  if (!status.ok()) BAND_LOG(LogSeverity::kError, "Failed to set quantization: %s", status.message().c_str());
}

Tensor::~Tensor() {
  if (data_ && owns_) {
    if (space_ == MemorySpace::kDevice) bandcu_free(data_);
    else if (pinned_) bandcu_host_free(data_);
    else delete[] data_;
  }
  if (quantization_.GetParams() != nullptr)
    delete reinterpret_cast<AffineQuantizationParams*>(quantization_.GetParams());
}

DataType Tensor::GetType() const { return type_; }
void Tensor::SetType(DataType type) { type_ = type; }
const char* Tensor::GetData() const { return data_; }
char* Tensor::GetData() { return data_; }
const int* Tensor::GetDims() const { return dims_.data(); }
size_t Tensor::GetNumDims() const { return dims_.size(); }
void Tensor::SetDims(const std::vector<int>& dims) { dims_ = dims; }
const char* Tensor::GetName() const { return name_.c_str(); }
Quantization Tensor::GetQuantization() const { return quantization_; }

// tensor.cc:55-90: affine parameters are deep-copied (the reference mallocs a
// struct holding std::vectors; here it is a properly constructed object).
absl::Status Tensor::SetQuantization(Quantization quantization) {
  if (quantization.GetType() == QuantizationType::kAffineQuantization) {
    const auto* in = reinterpret_cast<const AffineQuantizationParams*>(quantization.GetParams());
    if (in == nullptr) return absl::InternalError("Failed to allocate memory for quantization params");
    auto* copy = new AffineQuantizationParams(*in);
    if (quantization_.GetParams() != nullptr)
      delete reinterpret_cast<AffineQuantizationParams*>(quantization_.GetParams());
    quantization_ = Quantization(QuantizationType::kAffineQuantization, copy);
  }
  return absl::OkStatus();
}
}  // namespace band
