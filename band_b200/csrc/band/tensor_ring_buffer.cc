// band/tensor_ring_buffer.cc -- see tensor_ring_buffer.h. Follows
// /root/reference/band/tensor_ring_buffer.cc:27-161 for handle arithmetic, validity rules and error texts.
#include "band/tensor_ring_buffer.h"

#include <cassert>
#include <cstdlib>

#include "absl/strings/str_format.h"
#include "band/logger.h"
#include "band/tensor.h"
#include "bandcu.h"

namespace band {
namespace {
constexpr size_t kSlotAlign = 256;  // every slot tensor starts on a 256-byte boundary (TMA / 128-bit stores)
size_t AlignUp(size_t v) { return (v + kSlotAlign - 1) / kSlotAlign * kSlotAlign; }
}  // namespace

TensorRingBuffer::TensorRingBuffer(std::vector<std::shared_ptr<interface::ITensor>> tensors,
                                   std::vector<int> tensor_indices, int size, MemorySpace space)
    : size_(size), space_(space), tensors_(new std::vector<interface::ITensor*>[size]) {
  assert(size_ > 0);
  size_t per_slot = 0;
  for (const auto& t : tensors) per_slot += AlignUp(t->GetBytes());
  const size_t total = per_slot * static_cast<size_t>(size_);
  valid_ = true;
  if (total > 0) {
    void* p = nullptr;
    if (space_ == MemorySpace::kDevice) {
      if (bandcu_malloc(&p, total) != BANDCU_OK) p = nullptr;
    } else if (bandcu_host_alloc(&p, total) == BANDCU_OK && p != nullptr) {
      slab_pinned_ = true;
    } else {
      p = malloc(total);
    }
    if (p == nullptr) {
      BAND_LOG(LogSeverity::kError, "TensorRingBuffer: allocation of %zu bytes in %s memory failed", total,
               ToString(space_));
      valid_ = false;
    }
    slab_ = static_cast<char*>(p);
  }
  size_t off = 0;
  for (int i = 0; i < size_; i++) {
    tensors_[i].resize(tensors.size());
    for (size_t j = 0; j < tensors_[i].size(); j++) {
      tensors_[i][j] = new Tensor(tensors[j].get(), space_, slab_ ? slab_ + off : nullptr);
      off += AlignUp(tensors[j]->GetBytes());
    }
  }
  for (size_t i = 0; i < tensor_indices.size(); i++) tensor_to_buffer_[tensor_indices[i]] = static_cast<int>(i);
}

TensorRingBuffer::~TensorRingBuffer() {
  for (int i = 0; i < size_; i++)
    for (size_t j = 0; j < tensors_[i].size(); j++) delete tensors_[i][j];
  delete[] tensors_;
  if (slab_) {
    if (space_ == MemorySpace::kDevice) bandcu_free(slab_);
    else if (slab_pinned_) bandcu_host_free(slab_);
    else free(slab_);
  }
}

const int TensorRingBuffer::GetTensorsLength() const { return static_cast<int>(tensors_[0].size()); }

int TensorRingBuffer::Alloc() {
  std::lock_guard<std::mutex> lock(head_mtx_);
  return head_++;
}

bool TensorRingBuffer::IsTensorIndexValid(int tensor_index) const {
  return tensor_to_buffer_.find(tensor_index) != tensor_to_buffer_.end();
}

bool TensorRingBuffer::IsHandleValid(int handle) const {
  return (handle >= 0) && (head_ - size_ <= handle) && (handle < head_);
}

// One place for the reference's two validity rules and their messages (tensor_ring_buffer.cc:71-129): the caller
// holds head_mtx_. `tensor_index` < 0 means "whole slot" (no per-tensor lookup).
absl::Status TensorRingBuffer::Locate(const char* who, int tensor_index, int handle, bool whole_slot,
                                      interface::ITensor** slot) const {
  if (!whole_slot && !IsTensorIndexValid(tensor_index))
    return absl::InternalError(absl::StrFormat("%s: Invalid tensor index: %d.", who, tensor_index));
  if (!IsHandleValid(handle))
    return absl::InternalError(absl::StrFormat("%s: Invalid memory handle: %d head: %d.", who, handle, head_));
  if (slot) *slot = whole_slot ? nullptr : tensors_[GetIndex(handle)][tensor_to_buffer_.at(tensor_index)];
  return absl::OkStatus();
}

interface::ITensor* TensorRingBuffer::GetSlot(int tensor_index, int handle) const {
  if (!valid_) return nullptr;
  std::lock_guard<std::mutex> lock(head_mtx_);
  interface::ITensor* slot = nullptr;
  return Locate("GetSlot", tensor_index, handle, false, &slot).ok() ? slot : nullptr;
}

absl::Status TensorRingBuffer::GetTensorFromHandle(interface::ITensor* dst, int tensor_index, int handle) const {
  std::lock_guard<std::mutex> lock(head_mtx_);
  interface::ITensor* slot = nullptr;
  absl::Status st = Locate("GetTensorFromHandle", tensor_index, handle, false, &slot);
  return st.ok() ? CopyTensor(slot, dst) : st;
}

absl::Status TensorRingBuffer::PutTensorToHandle(const interface::ITensor* src, int tensor_index, int handle) {
  std::lock_guard<std::mutex> lock(head_mtx_);
  interface::ITensor* slot = nullptr;
  absl::Status st = Locate("PutTensorToHandle", tensor_index, handle, false, &slot);
  return st.ok() ? CopyTensor(src, slot) : st;
}

absl::Status TensorRingBuffer::GetTensorsFromHandle(std::vector<interface::ITensor*>& dst_tensors, int handle) const {
  std::lock_guard<std::mutex> lock(head_mtx_);
  absl::Status st = Locate("GetTensorsFromHandle", -1, handle, true, nullptr);
  return st.ok() ? CopyTensors(tensors_[GetIndex(handle)], dst_tensors) : st;
}

absl::Status TensorRingBuffer::PutTensorsToHandle(const std::vector<interface::ITensor*>& src_tensors, int handle) {
  std::lock_guard<std::mutex> lock(head_mtx_);
  absl::Status st = Locate("PutTensorsToHandle", -1, handle, true, nullptr);
  return st.ok() ? CopyTensors(src_tensors, tensors_[GetIndex(handle)]) : st;
}

absl::Status TensorRingBuffer::CopyTensors(const std::vector<interface::ITensor*>& src_tensors,
                                           std::vector<interface::ITensor*>& dst_tensors) const {
  const size_t tensors_length = static_cast<size_t>(GetTensorsLength());
  if (src_tensors.size() != tensors_length || dst_tensors.size() != tensors_length)
    return absl::InternalError(absl::StrFormat("Invalid tensor length. src tensors: %d dst tensors: %d expected: %d",
                                               static_cast<int>(src_tensors.size()),
                                               static_cast<int>(dst_tensors.size()), static_cast<int>(tensors_length)));
  for (size_t i = 0; i < tensors_length; i++)
    if (!CopyTensor(src_tensors[i], dst_tensors[i]).ok()) return absl::InternalError("Failed to copy tensors.");
  return absl::OkStatus();
}

absl::Status TensorRingBuffer::CopyTensor(const interface::ITensor* src, interface::ITensor* dst) const {
  if (!valid_ || dst == nullptr || !dst->CopyDataFrom(src).ok())
    return absl::InternalError(absl::StrFormat("Tensor data copy failure. src name : %s, dst name : %s",
                                               src ? src->GetName() : "null", dst ? dst->GetName() : "null"));
  return absl::OkStatus();
}

int TensorRingBuffer::GetIndex(int handle) const { return handle % size_; }
}  // namespace band
