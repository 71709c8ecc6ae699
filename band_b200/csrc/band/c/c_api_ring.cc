// band/c/c_api_ring.cc -- ADDITIVE C entry points over band::TensorRingBuffer (band/tensor_ring_buffer.h): the
// hand-off step right after the preprocessing path (reference call sites
// /root/reference/band/engine.cc:273-276,506-520,608).
#include <map>
#include <memory>
#include <mutex>
#include <utility>
#include <vector>

#include "band/c/c_api_internal.h"
#include "band/logger.h"
#include "band/tensor_ring_buffer.h"

using band::LogSeverity;

struct BandTensorRingBuffer {
  std::unique_ptr<band::TensorRingBuffer> impl;
  int size = 0;
  // C handles of lent slots, keyed by (ring index, tensor index); each wraps a non-owning view of the slot memory
  std::mutex mtx;
  std::map<std::pair<int, int>, std::unique_ptr<BandTensor>> lent;
};

extern "C" {

BandTensorRingBuffer* BandTensorRingBufferCreate(BandTensor* const* model_tensors, const int* tensor_indices,
                                                 size_t num_tensors, int size, BandMemorySpace space) {
  if (!model_tensors || !tensor_indices || num_tensors == 0 || size <= 0) {
    BAND_LOG(LogSeverity::kError, "BandTensorRingBufferCreate: null tensor list or non-positive size");
    return nullptr;
  }
  std::vector<std::shared_ptr<band::interface::ITensor>> shapes;
  std::vector<int> indices(tensor_indices, tensor_indices + num_tensors);
  for (size_t i = 0; i < num_tensors; ++i) {
    if (!model_tensors[i]) {
      BAND_LOG(LogSeverity::kError, "BandTensorRingBufferCreate: model tensor %zu is null", i);
      return nullptr;
    }
    // shape-only stand-in: the ring reads type, dims, name and quantization, never the data
    shapes.push_back(std::make_shared<band::Tensor>(model_tensors[i]->impl.get(), band::MemorySpace::kHost, nullptr));
  }
  auto ring = std::make_unique<BandTensorRingBuffer>();
  ring->size = size;
  ring->impl = std::make_unique<band::TensorRingBuffer>(
      shapes, indices, size, space == kBandDeviceMemory ? band::MemorySpace::kDevice : band::MemorySpace::kHost);
  if (!ring->impl->IsValid()) return nullptr;
  return ring.release();
}

void BandTensorRingBufferDelete(BandTensorRingBuffer* ring) { delete ring; }

int BandTensorRingBufferAlloc(BandTensorRingBuffer* ring) {
  if (!ring) {
    BAND_LOG(LogSeverity::kError, "BandTensorRingBuffer is null");
    return -1;
  }
  return ring->impl->Alloc();
}

int BandTensorRingBufferIsHandleValid(BandTensorRingBuffer* ring, int handle) {
  return ring && ring->impl->IsHandleValid(handle) ? 1 : 0;
}

BandTensor* BandTensorRingBufferGetSlot(BandTensorRingBuffer* ring, int tensor_index, int handle) {
  if (!ring) {
    BAND_LOG(LogSeverity::kError, "BandTensorRingBuffer is null");
    return nullptr;
  }
  band::interface::ITensor* slot = ring->impl->GetSlot(tensor_index, handle);
  if (!slot) {
    BAND_LOG(LogSeverity::kError, "BandTensorRingBufferGetSlot: invalid tensor index %d or handle %d", tensor_index,
             handle);
    return nullptr;
  }
  std::lock_guard<std::mutex> lock(ring->mtx);
  std::unique_ptr<BandTensor>& h = ring->lent[{handle % ring->size, tensor_index}];
  if (!h) h = std::make_unique<BandTensor>(std::make_unique<band::Tensor>(slot, slot->GetMemorySpace(), slot->GetData()));
  return h.get();
}

BandStatus BandTensorRingBufferPutTensor(BandTensorRingBuffer* ring, BandTensor* src, int tensor_index, int handle) {
  if (!ring || !src) {
    BAND_LOG(LogSeverity::kError, "BandTensorRingBufferPutTensor: null argument");
    return kBandErr;
  }
  absl::Status st = ring->impl->PutTensorToHandle(src->impl.get(), tensor_index, handle);
  if (!st.ok()) BAND_LOG(LogSeverity::kError, "%s", st.message().c_str());
  return st.ok() ? kBandOk : kBandErr;
}

BandStatus BandTensorRingBufferGetTensor(BandTensorRingBuffer* ring, BandTensor* dst, int tensor_index, int handle) {
  if (!ring || !dst) {
    BAND_LOG(LogSeverity::kError, "BandTensorRingBufferGetTensor: null argument");
    return kBandErr;
  }
  absl::Status st = ring->impl->GetTensorFromHandle(dst->impl.get(), tensor_index, handle);
  if (!st.ok()) BAND_LOG(LogSeverity::kError, "%s", st.message().c_str());
  return st.ok() ? kBandOk : kBandErr;
}

}  // extern "C"
