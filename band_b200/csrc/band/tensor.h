// band/tensor.h -- band::Tensor, the handoff target of the preprocessing path
// (/root/reference/band/tensor.h:13-36, tensor.cc:8-31): an owning NHWC
// buffer. Additive: a second constructor that builds a tensor from
// (type, dims, name) without an engine, and an optional DEVICE backing so the
// fused kernels write model input straight into HBM.
#ifndef BAND_TENSOR_H_
#define BAND_TENSOR_H_

#include <string>
#include <vector>

#include "band/interface/tensor.h"

namespace band {
class Tensor : public interface::ITensor {
 public:
  explicit Tensor(interface::ITensor* tensor_view, bool copy_data = false);
  // additive
  Tensor(DataType type, const std::vector<int>& dims, const std::string& name = "",
         MemorySpace space = MemorySpace::kHost);
  // additive: a tensor shaped like `like` over EXTERNAL memory of the given space (not owned, never freed here);
  // the slots of a TensorRingBuffer are such views into the ring's slab. external == nullptr => zero-byte tensor.
  Tensor(const interface::ITensor* like, MemorySpace space, char* external);
  ~Tensor();

  DataType GetType() const override;
  void SetType(DataType type) override;
  const char* GetData() const override;
  char* GetData() override;
  const int* GetDims() const override;
  size_t GetNumDims() const override;
  void SetDims(const std::vector<int>& dims) override;
  const char* GetName() const override;
  Quantization GetQuantization() const override;
  absl::Status SetQuantization(Quantization quantization) override;
  MemorySpace GetMemorySpace() const override { return space_; }
  bool IsValid() const { return data_ != nullptr || GetBytes() == 0; }

 private:
  Tensor(const Tensor&) = delete;
  Tensor& operator=(const Tensor&) = delete;
  void Allocate(size_t bytes);

  DataType type_;
  Quantization quantization_;
  std::vector<int> dims_;
  char* data_ = nullptr;
  std::string name_;
  MemorySpace space_ = MemorySpace::kHost;
  bool pinned_ = false;
  bool owns_ = true;
};
}  // namespace band
#endif  // BAND_TENSOR_H_
