// band/c/c_api_internal.h -- what the opaque C handles hold
// (/root/reference/band/c/c_api_internal.h:30-57, buffer/tensor/processor part).
#ifndef BAND_C_C_API_INTERNAL_H_
#define BAND_C_C_API_INTERNAL_H_

#include <memory>

#include "band/buffer/buffer.h"
#include "band/buffer/image_processor.h"
#include "band/tensor.h"
#include "band_c_api.h"

struct BandTensor {
  explicit BandTensor(std::unique_ptr<band::Tensor> tensor) : impl(std::move(tensor)) {}
  std::unique_ptr<band::Tensor> impl;
  // Buffer view over the tensor memory (Buffer::CreateFromTensor), built on first use: a C-API tensor's
  // type and dims never change after creation, so batch calls do not rebuild it per frame.
  std::unique_ptr<band::Buffer> view;
  band::Buffer* View() {
    if (!view) view.reset(band::Buffer::CreateFromTensor(impl.get()));
    return view.get();
  }
};

struct BandBuffer {
  BandBuffer() : impl(nullptr) {}
  // replaced by every BandBufferSetFrom* call
  std::shared_ptr<band::Buffer> impl;
};

struct BandImageProcessorBuilder {
  BandImageProcessorBuilder() : impl(std::make_unique<band::ImageProcessorBuilder>()) {}
  std::unique_ptr<band::ImageProcessorBuilder> impl;
};

struct BandImageProcessor {
  explicit BandImageProcessor(std::unique_ptr<band::BufferProcessor> processor) : impl(std::move(processor)) {}
  std::unique_ptr<band::BufferProcessor> impl;
};

// text BandGetLastError() returns on the calling thread (c_api_buffer.cc)
void BandSetLastError(const char* message);

#endif  // BAND_C_C_API_INTERNAL_H_
