// band/tensor_ring_buffer.h -- TensorRingBuffer, the engine's hand-off buffer between the preprocessing path and
// model execution (/root/reference/band/tensor_ring_buffer.h:33-66; used by Engine::RequestAsync,
// /root/reference/band/engine.cc:506-520, and TryCopyInputTensors, :1247-1319).
//
// Same public interface and status behaviour as the reference. What is new:
//  * the slots can live in HBM (MemorySpace::kDevice): all `size x tensors` slot tensors are carved out of ONE
//    device slab (256-byte aligned sub-allocations), so a ring of 128 x [1,224,224,3] f32 slots is one 77 MB
//    allocation instead of 128 cudaMallocs;
//  * GetSlot() lends a slot tensor to a producer. The image processor then writes model input straight into the
//    ring (BandImageProcessorProcess(proc, buffer, slot)), which removes both host memcpys the reference does
//    between the processor output and the model input (user tensor -> ring -> backend tensor).
// Put/Get still copy (H2D / D2H / D2D as the memory spaces require) for callers that hold their own tensors.
#ifndef BAND_TENSOR_RING_BUFFER_H_
#define BAND_TENSOR_RING_BUFFER_H_

#include <map>
#include <memory>
#include <mutex>
#include <vector>

#include "absl/status/status.h"
#include "band/interface/tensor.h"

namespace band {

class Tensor;

class TensorRingBuffer {
 public:
  TensorRingBuffer(std::vector<std::shared_ptr<interface::ITensor>> tensors, std::vector<int> tensor_indices,
                   int size = 128, MemorySpace space = MemorySpace::kHost);
  ~TensorRingBuffer();

  const int GetTensorsLength() const;
  int Alloc();
  bool IsTensorIndexValid(int tensor_index) const;
  bool IsHandleValid(int handle) const;
  absl::Status GetTensorFromHandle(interface::ITensor* dst, int tensor_index, int handle) const;
  absl::Status PutTensorToHandle(const interface::ITensor* src, int tensor_index, int handle);
  absl::Status GetTensorsFromHandle(std::vector<interface::ITensor*>& dst_tensors, int handle) const;
  absl::Status PutTensorsToHandle(const std::vector<interface::ITensor*>& src_tensors, int handle);

  // additive
  MemorySpace GetMemorySpace() const { return space_; }
  bool IsValid() const { return valid_; }
  // The slot tensor itself (owned by the ring; valid until the handle is overwritten `size` allocations later).
  // nullptr for an invalid tensor index or handle.
  interface::ITensor* GetSlot(int tensor_index, int handle) const;

 private:
  int GetIndex(int handle) const;
  absl::Status Locate(const char* who, int tensor_index, int handle, bool whole_slot,
                      interface::ITensor** slot) const;
  absl::Status CopyTensors(const std::vector<interface::ITensor*>& src_tensors,
                           std::vector<interface::ITensor*>& dst_tensors) const;
  absl::Status CopyTensor(const interface::ITensor* src, interface::ITensor* dst) const;

  mutable std::mutex head_mtx_;
  int head_ = 0;
  const int size_;
  const MemorySpace space_;
  bool valid_ = false;
  char* slab_ = nullptr;  // one allocation for every slot (device memory or pinned host memory)
  bool slab_pinned_ = false;
  std::vector<interface::ITensor*>* tensors_;
  // Model's tensor index to ring buffer's index
  std::map<int, int> tensor_to_buffer_;
};
}  // namespace band

#endif  // BAND_TENSOR_RING_BUFFER_H_
