// band/c/c_api_tensor.cc -- BandTensor getters (/root/reference/band/c/c_api.cc:262-343)
// plus the additive stand-alone BandTensorCreate.
#include <vector>

#include "band/c/c_api_internal.h"
#include "band/logger.h"

using band::LogSeverity;

namespace {
bool Null(BandTensor* tensor) {
  if (tensor) return false;
  BAND_LOG(LogSeverity::kError, "BandTensor is null");
  return true;
}
}  // namespace

extern "C" {

void BandTensorDelete(BandTensor* tensor) {
  if (Null(tensor)) return;
  delete tensor;
}
BandDataType BandTensorGetType(BandTensor* tensor) {
  return Null(tensor) ? kBandNumDataType : static_cast<BandDataType>(tensor->impl->GetType());
}
void* BandTensorGetData(BandTensor* tensor) { return Null(tensor) ? nullptr : tensor->impl->GetData(); }
size_t BandTensorGetNumDims(BandTensor* tensor) { return Null(tensor) ? 0 : tensor->impl->GetNumDims(); }
const int* BandTensorGetDims(BandTensor* tensor) { return Null(tensor) ? nullptr : tensor->impl->GetDims(); }
size_t BandTensorGetBytes(BandTensor* tensor) { return Null(tensor) ? 0 : tensor->impl->GetBytes(); }
const char* BandTensorGetName(BandTensor* tensor) { return Null(tensor) ? nullptr : tensor->impl->GetName(); }
BandQuantizationType BandTensorGetQuantizationType(BandTensor* tensor) {
  return Null(tensor) ? kBandNumQuantizationType
                      : static_cast<BandQuantizationType>(tensor->impl->GetQuantization().GetType());
}
void* BandTensorGetQuantizationParams(BandTensor* tensor) {
  return Null(tensor) ? nullptr : tensor->impl->GetQuantization().GetParams();
}

// ADDITIVE
BandTensor* BandTensorCreate(BandDataType type, size_t num_dims, const int* dims, BandMemorySpace space) {
  if (!dims || num_dims == 0) {
    BAND_LOG(LogSeverity::kError, "BandTensorCreate: dims is null or empty");
    return nullptr;
  }
  if (band::GetDataTypeBytes(static_cast<band::DataType>(type)) == 0) {
    BAND_LOG(LogSeverity::kError, "BandTensorCreate: unsupported data type %d", static_cast<int>(type));
    return nullptr;
  }
  std::vector<int> d(dims, dims + num_dims);
  for (int v : d) {
    if (v <= 0) {
      BAND_LOG(LogSeverity::kError, "BandTensorCreate: non-positive dimension %d", v);
      return nullptr;
    }
  }
  auto tensor = std::make_unique<band::Tensor>(
      static_cast<band::DataType>(type), d, "",
      space == kBandDeviceMemory ? band::MemorySpace::kDevice : band::MemorySpace::kHost);
  if (!tensor->IsValid()) return nullptr;
  return new BandTensor(std::move(tensor));
}

BandMemorySpace BandTensorGetMemorySpace(BandTensor* tensor) {
  if (Null(tensor)) return kBandHostMemory;
  return tensor->impl->GetMemorySpace() == band::MemorySpace::kDevice ? kBandDeviceMemory : kBandHostMemory;
}

}  // extern "C"
