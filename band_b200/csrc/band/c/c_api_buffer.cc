// band/c/c_api_buffer.cc -- the buffer / image-processor part of Band's C API
// (/root/reference/band/c/c_api_buffer.cc:30-218) over the CUDA-backed
// processor, plus the additive device / batch entry points of band_c_api.h.
#include <cstdarg>
#include <vector>

#include "band/buffer/fused_plan.h"
#include "band/buffer/image_operator.h"
#include "band/c/c_api_internal.h"
#include "band/logger.h"
#include "bandcu.h"

using band::Buffer;
using band::BufferFormat;
using band::LogSeverity;
using band::MemorySpace;

namespace {

BandStatus SetRaw(BandBuffer* buffer, const void* data, size_t width, size_t height, BandBufferFormat format,
                  MemorySpace space) {
  if (!buffer || !data) {
    BAND_LOG(LogSeverity::kError, "BandBuffer (%p) or data (%p) is null", static_cast<void*>(buffer), data);
    return kBandErr;
  }
  buffer->impl = std::shared_ptr<Buffer>(
      Buffer::CreateFromRaw(static_cast<const unsigned char*>(data), width, height, BufferFormat(format),
                            band::DataType::kUInt8, band::BufferOrientation::kTopLeft, false, space));
  return buffer->impl ? kBandOk : kBandErr;
}

BandStatus SetYuv(BandBuffer* buffer, const void* y, const void* u, const void* v, size_t width, size_t height,
                  size_t row_stride_y, size_t row_stride_uv, size_t pixel_stride_uv, BandBufferFormat format,
                  MemorySpace space) {
  if (!buffer || !y || !u || !v) {
    BAND_LOG(LogSeverity::kError, "BandBuffer (%p) or y_data (%p) or u_data (%p) or v_data (%p) is null",
             static_cast<void*>(buffer), y, u, v);
    return kBandErr;
  }
  buffer->impl = std::shared_ptr<Buffer>(Buffer::CreateFromYUVPlanes(
      static_cast<const unsigned char*>(y), static_cast<const unsigned char*>(u),
      static_cast<const unsigned char*>(v), width, height, row_stride_y, row_stride_uv, pixel_stride_uv,
      BufferFormat(format), band::BufferOrientation::kTopLeft, false, space));
  return buffer->impl ? kBandOk : kBandErr;
}

thread_local std::string g_last_error;

}  // namespace

void BandSetLastError(const char* message) { g_last_error = message ? message : ""; }

extern "C" {

BandBuffer* BandBufferCreate() { return new BandBuffer(); }

void BandBufferDelete(BandBuffer* buffer) {
  if (!buffer) BAND_LOG(LogSeverity::kError, "BandBuffer is null");
  delete buffer;
}

BandStatus BandBufferSetFromRawData(BandBuffer* buffer, const void* data, size_t width, size_t height,
                                    BandBufferFormat format) {
  return SetRaw(buffer, data, width, height, format, MemorySpace::kHost);
}

BandStatus BandBufferSetFromYUVData(BandBuffer* buffer, const void* y_data, const void* u_data, const void* v_data,
                                    size_t width, size_t height, size_t row_stride_y, size_t row_stride_uv,
                                    size_t pixel_stride_uv, BandBufferFormat buffer_format) {
  return SetYuv(buffer, y_data, u_data, v_data, width, height, row_stride_y, row_stride_uv, pixel_stride_uv,
                buffer_format, MemorySpace::kHost);
}

BandStatus BandBufferSetFromDeviceRawData(BandBuffer* buffer, const void* device_data, size_t width, size_t height,
                                          BandBufferFormat format) {
  return SetRaw(buffer, device_data, width, height, format, MemorySpace::kDevice);
}

BandStatus BandBufferSetFromDeviceYUVData(BandBuffer* buffer, const void* y_data, const void* u_data,
                                          const void* v_data, size_t width, size_t height, size_t row_stride_y,
                                          size_t row_stride_uv, size_t pixel_stride_uv,
                                          BandBufferFormat buffer_format) {
  return SetYuv(buffer, y_data, u_data, v_data, width, height, row_stride_y, row_stride_uv, pixel_stride_uv,
                buffer_format, MemorySpace::kDevice);
}

BandImageProcessorBuilder* BandImageProcessorBuilderCreate() { return new BandImageProcessorBuilder(); }

void BandImageProcessorBuilderDelete(BandImageProcessorBuilder* builder) {
  if (!builder) {
    BAND_LOG(LogSeverity::kError, "BandImageProcessorBuilder is null");
    return;
  }
  delete builder;
}

BandImageProcessor* BandImageProcessorBuilderBuild(BandImageProcessorBuilder* builder) {
  if (!builder) {
    BAND_LOG(LogSeverity::kError, "BandImageProcessorBuilder is null");
    return nullptr;
  }
  absl::StatusOr<std::unique_ptr<band::BufferProcessor>> built = builder->impl->Build();
  if (!built.ok()) {
    g_last_error = built.status().ToString();
    return nullptr;
  }
  return new BandImageProcessor(std::move(built).value());
}

// varargs: ints for the image operators, doubles for Normalize (c_api_buffer.cc:108-196)
BandStatus BandAddOperator(BandImageProcessorBuilder* builder, BandImageProcessorBuilderField field, int count, ...) {
  if (!builder) {
    BAND_LOG(LogSeverity::kError, "BandImageProcessorBuilder is null");
    return kBandErr;
  }
  static const int kArity[] = {4, 2, 1, 2, 1, 2, 0, 0};
  if (field < BAND_CROP || field > BAND_AUTO_ORIENT || count != kArity[field]) return kBandErr;
  va_list vl;
  va_start(vl, count);
  std::unique_ptr<band::IBufferOperator> op;
  switch (field) {
    case BAND_CROP: {
      const int x0 = va_arg(vl, int), y0 = va_arg(vl, int), x1 = va_arg(vl, int), y1 = va_arg(vl, int);
      op = std::make_unique<band::buffer::Crop>(x0, y0, x1, y1);
      break;
    }
    case BAND_RESIZE: {
      const int w = va_arg(vl, int), h = va_arg(vl, int);
      op = std::make_unique<band::buffer::Resize>(w, h);
      break;
    }
    case BAND_ROTATE:
      op = std::make_unique<band::buffer::Rotate>(va_arg(vl, int));
      break;
    case BAND_FLIP: {
      const bool horizontal = va_arg(vl, int) != 0, vertical = va_arg(vl, int) != 0;
      op = std::make_unique<band::buffer::Flip>(horizontal, vertical);
      break;
    }
    case BAND_COLOR_SPACE_CONVERT:
      op = std::make_unique<band::buffer::ColorSpaceConvert>(BufferFormat(va_arg(vl, int)));
      break;
    case BAND_NORMALIZE: {
      const float mean = static_cast<float>(va_arg(vl, double)), std = static_cast<float>(va_arg(vl, double));
      op = std::make_unique<band::buffer::Normalize>(mean, std, false);
      break;
    }
    case BAND_DATA_TYPE_CONVERT:
      op = std::make_unique<band::buffer::DataTypeConvert>();
      break;
    case BAND_AUTO_ORIENT:
      op = std::make_unique<band::buffer::AutoOrient>();
      break;
  }
  va_end(vl);
  builder->impl->AddOperation(std::move(op));
  return kBandOk;
}

BandStatus BandBufferSetOrientation(BandBuffer* buffer, int exif_orientation) {
  if (!buffer || !buffer->impl) {
    BAND_LOG(LogSeverity::kError, "BandBufferSetOrientation: buffer is null or has no data");
    return kBandErr;
  }
  if (exif_orientation < 1 || exif_orientation > 8) {
    BAND_LOG(LogSeverity::kError, "BandBufferSetOrientation: orientation %d is not in 1..8", exif_orientation);
    return kBandErr;
  }
  buffer->impl->SetOrientation(static_cast<band::BufferOrientation>(exif_orientation));
  return kBandOk;
}

BandStatus BandImageProcessorProcess(BandImageProcessor* image_processor, BandBuffer* buffer,
                                     BandTensor* target_tensor) {
  if (!image_processor || !buffer || !target_tensor) {
    BAND_LOG(LogSeverity::kError, "BandImageProcessor (%p) or BandBuffer (%p) or BandTensor (%p) is null",
             static_cast<void*>(image_processor), static_cast<void*>(buffer), static_cast<void*>(target_tensor));
    return kBandErr;
  }
  if (!buffer->impl) {
    BAND_LOG(LogSeverity::kError, "BandBuffer has no data");
    return kBandErr;
  }
  // the cached view of the tensor (its type and dims never change): the reference builds a temporary Buffer per call
  // (c_api_buffer.cc:209-212); reusing one keeps the processor's repeat fast path alive
  Buffer* tensor_buffer = target_tensor->View();
  if (!tensor_buffer) return kBandErr;
  absl::Status status = image_processor->impl->Process(*buffer->impl, *tensor_buffer);
  if (!status.ok()) g_last_error = status.ToString();
  return status.ok() ? kBandOk : kBandErr;
}

static BandStatus ProcessBatchImpl(BandImageProcessor* image_processor, BandBuffer* const* buffers,
                                   BandTensor* const* target_tensors, size_t count, bool async) {
  if (!image_processor || !buffers || !target_tensors) {
    BAND_LOG(LogSeverity::kError, "BandImageProcessorProcessBatch: null argument");
    return kBandErr;
  }
  // per-thread scratch: a harness calls this in a loop with thousands of frames
  thread_local std::vector<const Buffer*> in;
  thread_local std::vector<Buffer*> out;
  in.resize(count);
  out.resize(count);
  for (size_t i = 0; i < count; ++i) {
    if (!buffers[i] || !buffers[i]->impl || !target_tensors[i]) {
      BAND_LOG(LogSeverity::kError, "BandImageProcessorProcessBatch: entry %d is null", static_cast<int>(i));
      return kBandErr;
    }
    in[i] = buffers[i]->impl.get();
    out[i] = target_tensors[i]->View();
    if (!out[i]) return kBandErr;
  }
  absl::Status status = async ? image_processor->impl->ProcessBatchAsync(in.data(), out.data(), count)
                              : image_processor->impl->ProcessBatch(in.data(), out.data(), count);
  if (!status.ok()) g_last_error = status.ToString();
  return status.ok() ? kBandOk : kBandErr;
}

BandStatus BandImageProcessorProcessBatch(BandImageProcessor* image_processor, BandBuffer* const* buffers,
                                          BandTensor* const* target_tensors, size_t count) {
  return ProcessBatchImpl(image_processor, buffers, target_tensors, count, false);
}

BandStatus BandImageProcessorProcessBatchAsync(BandImageProcessor* image_processor, BandBuffer* const* buffers,
                                               BandTensor* const* target_tensors, size_t count) {
  return ProcessBatchImpl(image_processor, buffers, target_tensors, count, true);
}

struct BandRecordedBatch {
  // graph + the plan whose device memory the graph reads (kept alive here, whatever the processor does afterwards)
  std::unique_ptr<band::RecordedBatch> impl;
};

BandRecordedBatch* BandImageProcessorRecordBatch(BandImageProcessor* image_processor, BandBuffer* const* buffers,
                                                 BandTensor* const* target_tensors, size_t count) {
  if (!image_processor || !buffers || !target_tensors || count == 0) {
    BAND_LOG(LogSeverity::kError, "BandImageProcessorRecordBatch: null argument");
    return nullptr;
  }
  std::vector<const Buffer*> in(count);
  std::vector<Buffer*> out(count);
  for (size_t i = 0; i < count; ++i) {
    if (!buffers[i] || !buffers[i]->impl || !target_tensors[i]) {
      BAND_LOG(LogSeverity::kError, "BandImageProcessorRecordBatch: entry %d is null", static_cast<int>(i));
      return nullptr;
    }
    in[i] = buffers[i]->impl.get();
    out[i] = target_tensors[i]->View();
    if (!out[i]) return nullptr;
  }
  absl::StatusOr<std::unique_ptr<band::RecordedBatch>> recorded =
      image_processor->impl->RecordBatch(in.data(), out.data(), count);
  if (!recorded.ok()) {
    g_last_error = recorded.status().ToString();
    return nullptr;
  }
  BandRecordedBatch* r = new BandRecordedBatch();
  r->impl = std::move(recorded).value();
  return r;
}

BandStatus BandRecordedBatchReplay(BandRecordedBatch* recorded, int synchronize) {
  if (!recorded) {
    BAND_LOG(LogSeverity::kError, "BandRecordedBatch is null");
    return kBandErr;
  }
  absl::Status status = band::BufferProcessor::ReplayBatch(*recorded->impl, synchronize != 0);
  if (!status.ok()) g_last_error = status.ToString();
  return status.ok() ? kBandOk : kBandErr;
}

void BandRecordedBatchDelete(BandRecordedBatch* recorded) {
  delete recorded;  // drains a replay still in flight, then frees the graph and (if last owner) the plan
}

BandStatus BandImageProcessorProcessBuffers(BandImageProcessor* image_processor, BandBuffer* const* inputs,
                                            BandBuffer* const* outputs, size_t count, int async) {
  if (!image_processor || !inputs || !outputs) return kBandErr;
  thread_local std::vector<const Buffer*> in;
  thread_local std::vector<Buffer*> out;
  in.resize(count);
  out.resize(count);
  for (size_t i = 0; i < count; ++i) {
    if (!inputs[i] || !inputs[i]->impl || !outputs[i] || !outputs[i]->impl) return kBandErr;
    in[i] = inputs[i]->impl.get();
    out[i] = outputs[i]->impl.get();
  }
  absl::Status status = async ? image_processor->impl->ProcessBatchAsync(in.data(), out.data(), count)
                              : image_processor->impl->ProcessBatch(in.data(), out.data(), count);
  if (!status.ok()) g_last_error = status.ToString();
  return status.ok() ? kBandOk : kBandErr;
}

BandStatus BandImageProcessorSynchronize(BandImageProcessor* image_processor) {
  if (!image_processor) return kBandErr;
  absl::Status status = image_processor->impl->Synchronize();
  if (!status.ok()) g_last_error = status.ToString();
  return status.ok() ? kBandOk : kBandErr;
}

void* BandImageProcessorGetStream(BandImageProcessor* image_processor) {
  return image_processor ? image_processor->impl->GetComputeStream() : nullptr;
}

int BandImageProcessorIsFused(BandImageProcessor* image_processor) {
  return image_processor && image_processor->impl->IsSingleFastLaunch() ? 1 : 0;
}

void BandImageProcessorDelete(BandImageProcessor* processor) { delete processor; }

int BandCudaGetDeviceCount() {
  int n = 0;
  return bandcu_device_count(&n) == BANDCU_OK ? n : 0;
}
BandStatus BandCudaSetDevice(int device) { return band::cuda::SetDevice(device).ok() ? kBandOk : kBandErr; }
void* BandHostAlloc(size_t bytes) {
  void* p = nullptr;
  return bandcu_host_alloc(&p, bytes) == BANDCU_OK ? p : nullptr;
}
void BandHostFree(void* ptr) { bandcu_host_free(ptr); }
void* BandDeviceAlloc(size_t bytes) {
  void* p = nullptr;
  return bandcu_malloc(&p, bytes) == BANDCU_OK ? p : nullptr;
}
void BandDeviceFree(void* ptr) { bandcu_free(ptr); }
BandStatus BandCopyToDevice(void* device_dst, const void* host_src, size_t bytes) {
  if (bandcu_memcpy_h2d(device_dst, host_src, bytes, nullptr) != BANDCU_OK) return kBandErr;
  return bandcu_stream_sync(nullptr) == BANDCU_OK ? kBandOk : kBandErr;
}
BandStatus BandCopyToHost(void* host_dst, const void* device_src, size_t bytes) {
  if (bandcu_memcpy_d2h(host_dst, device_src, bytes, nullptr) != BANDCU_OK) return kBandErr;
  return bandcu_stream_sync(nullptr) == BANDCU_OK ? kBandOk : kBandErr;
}
long long BandCudaLaunchCount() { return bandcu_launch_count(); }
const char* BandGetLastError() { return g_last_error.c_str(); }

}  // extern "C"
