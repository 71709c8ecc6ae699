/*
 * include/band_c_api.h -- the C boundary of libband_c.so for the preprocessing path.
 *
 * Part 1 keeps the reference's exported names, signatures, enum values and
 * status conventions so existing dlopen() users bind unchanged:
 *   /root/reference/band/c/c_api_buffer.h:55-121  (BandBuffer*, BandImageProcessor*, BandAddOperator, PFN_* typedefs)
 *   /root/reference/band/c/c_api.h:82-92          (BandTensor* getters, BandTensorDelete)
 *   /root/reference/band/c/c_api_type.h:43,88-133,178-186 (BandStatus, BandDataType, BandBufferFormat,
 *                                                  BandBufferOrientation, BandImageProcessorBuilderField)
 * Part 2 is additive (marked ADDITIVE): device-resident buffers / tensors, batch
 * processing, device selection and pinned host memory. The reference can only
 * obtain a BandTensor from an engine + model (c_api.cc:428-442); the engine is
 * out of scope here, so BandTensorCreate fills that gap.
 *
 * Everything is plain C: opaque handles, pointers and sizes. There is no CPU
 * fallback behind these calls: without a CUDA device Process* returns kBandErr.
 */
#ifndef BAND_INCLUDE_BAND_C_API_H_
#define BAND_INCLUDE_BAND_C_API_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BAND_CAPI_EXPORT __attribute__((visibility("default")))

/* ---- c_api_type.h ---------------------------------------------------------- */
typedef enum BandStatus { kBandOk = 0, kBandErr, kBandDelegateErr } BandStatus;

typedef enum {
  kBandNoType = 0,
  kBandFloat32,
  kBandInt32,
  kBandUInt8,
  kBandInt64,
  kBandString,
  kBandBool,
  kBandInt16,
  kBandComplex64,
  kBandInt8,
  kBandFloat16,
  kBandFloat64,
  kBandNumDataType,
} BandDataType;

typedef enum {
  kBandGrayScale = 0,
  kBandRGB,
  kBandRGBA,
  kBandYV12,
  kBandYV21,
  kBandNV21,
  kBandNV12,
  kBandRaw, /* from tensor; element type follows BandDataType */
  kBandNumBufferFormat,
} BandBufferFormat;

typedef enum {
  kBandTopLeft = 1,
  kBandTopRight,
  kBandBottomRight,
  kBandBottomLeft,
  kBandLeftTop,
  kBandRightTop,
  kBandRightBottom,
  kBandLeftBottom,
  kBandNumBufferOrientation,
} BandBufferOrientation;

typedef enum BandQuantizationType {
  kBandNoQuantization = 0,
  kBandAffineQuantization,
  kBandNumQuantizationType,
} BandQuantizationType;

typedef enum BandImageProcessorBuilderField {
  BAND_CROP = 0,            /* 4 int: x0, y0, x1, y1 (inclusive box) */
  BAND_RESIZE,              /* 2 int: width, height */
  BAND_ROTATE,              /* 1 int: counter-clockwise degrees, multiple of 90 */
  BAND_FLIP,                /* 2 int (bool): horizontal, vertical */
  BAND_COLOR_SPACE_CONVERT, /* 1 int: BandBufferFormat */
  BAND_NORMALIZE,           /* 2 double: mean, std */
  BAND_DATA_TYPE_CONVERT,   /* 0 */
  BAND_AUTO_ORIENT          /* 0; ADDITIVE: apply the buffer's EXIF orientation (BandBufferSetOrientation) */
} BandImageProcessorBuilderField;

/* ---- handles ---------------------------------------------------------------- */
typedef struct BandTensor BandTensor;
/* A view of user-owned memory: Band never takes ownership and the memory must
 * outlive the BandBuffer (c_api_buffer.h:27-30). */
typedef struct BandBuffer BandBuffer;
typedef struct BandImageProcessor BandImageProcessor;
typedef struct BandImageProcessorBuilder BandImageProcessorBuilder;

/* ---- c_api.h:82-92 ------------------------------------------------------------ */
BAND_CAPI_EXPORT extern void BandTensorDelete(BandTensor* tensor);
BAND_CAPI_EXPORT extern BandDataType BandTensorGetType(BandTensor* tensor);
BAND_CAPI_EXPORT extern void* BandTensorGetData(BandTensor* tensor);
BAND_CAPI_EXPORT extern size_t BandTensorGetNumDims(BandTensor* tensor);
BAND_CAPI_EXPORT extern const int* BandTensorGetDims(BandTensor* tensor);
BAND_CAPI_EXPORT extern size_t BandTensorGetBytes(BandTensor* tensor);
BAND_CAPI_EXPORT extern const char* BandTensorGetName(BandTensor* tensor);
BAND_CAPI_EXPORT extern BandQuantizationType BandTensorGetQuantizationType(BandTensor* tensor);
BAND_CAPI_EXPORT extern void* BandTensorGetQuantizationParams(BandTensor* tensor);

/* ---- c_api_buffer.h:55-102 ------------------------------------------------------ */
BAND_CAPI_EXPORT extern BandBuffer* BandBufferCreate();
BAND_CAPI_EXPORT extern void BandBufferDelete(BandBuffer* buffer);
/* One contiguous frame: RGB / RGBA / GRAY packed, NV21 / NV12 semi-planar, YV12 / YV21 planar. */
BAND_CAPI_EXPORT extern BandStatus BandBufferSetFromRawData(BandBuffer* buffer, const void* data, size_t width,
                                                            size_t height, BandBufferFormat format);
BAND_CAPI_EXPORT extern BandStatus BandBufferSetFromYUVData(BandBuffer* buffer, const void* y_data,
                                                            const void* u_data, const void* v_data, size_t width,
                                                            size_t height, size_t row_stride_y, size_t row_stride_uv,
                                                            size_t pixel_stride_uv, BandBufferFormat buffer_format);
BAND_CAPI_EXPORT extern BandImageProcessorBuilder* BandImageProcessorBuilderCreate();
BAND_CAPI_EXPORT extern void BandImageProcessorBuilderDelete(BandImageProcessorBuilder* builder);
BAND_CAPI_EXPORT extern BandImageProcessor* BandImageProcessorBuilderBuild(BandImageProcessorBuilder* builder);
/* Operators apply in the order added; kBandErr if `count` does not fit `field`. */
BAND_CAPI_EXPORT extern BandStatus BandAddOperator(BandImageProcessorBuilder* b, BandImageProcessorBuilderField field,
                                                   int count, ...);
BAND_CAPI_EXPORT extern BandStatus BandImageProcessorProcess(BandImageProcessor* image_processor, BandBuffer* buffer,
                                                             BandTensor* target_tensor);
BAND_CAPI_EXPORT extern void BandImageProcessorDelete(BandImageProcessor* processor);

/* ---- ADDITIVE: device memory, batches, devices ---------------------------------- */
typedef enum BandMemorySpace { kBandHostMemory = 0, kBandDeviceMemory = 1 } BandMemorySpace;

/* A stand-alone NHWC tensor (no engine): host tensors are pinned, device tensors live in HBM. */
BAND_CAPI_EXPORT extern BandTensor* BandTensorCreate(BandDataType type, size_t num_dims, const int* dims,
                                                     BandMemorySpace space);
BAND_CAPI_EXPORT extern BandMemorySpace BandTensorGetMemorySpace(BandTensor* tensor);
/* Same as the SetFrom* calls above, for frames already resident in device memory. */
BAND_CAPI_EXPORT extern BandStatus BandBufferSetFromDeviceRawData(BandBuffer* buffer, const void* device_data,
                                                                  size_t width, size_t height,
                                                                  BandBufferFormat format);
BAND_CAPI_EXPORT extern BandStatus BandBufferSetFromDeviceYUVData(BandBuffer* buffer, const void* y_data,
                                                                  const void* u_data, const void* v_data,
                                                                  size_t width, size_t height, size_t row_stride_y,
                                                                  size_t row_stride_uv, size_t pixel_stride_uv,
                                                                  BandBufferFormat buffer_format);
/* EXIF orientation 1..8 of the frame (band::BufferOrientation, band/common.h:146-161; default 1 = upright). The
 * reference carries it and never applies it; here the BAND_AUTO_ORIENT operator does. Call after SetFrom*. */
BAND_CAPI_EXPORT extern BandStatus BandBufferSetOrientation(BandBuffer* buffer, int exif_orientation);
/* n independent frames of one layout -> n tensors of one shape, as fused batch launches. */
BAND_CAPI_EXPORT extern BandStatus BandImageProcessorProcessBatch(BandImageProcessor* image_processor,
                                                                  BandBuffer* const* buffers,
                                                                  BandTensor* const* target_tensors, size_t count);
/* Enqueue only: results are complete after BandImageProcessorSynchronize. Buffers and tensors must stay alive
 * (and host memory untouched) until then. */
BAND_CAPI_EXPORT extern BandStatus BandImageProcessorProcessBatchAsync(BandImageProcessor* image_processor,
                                                                       BandBuffer* const* buffers,
                                                                       BandTensor* const* target_tensors,
                                                                       size_t count);
BAND_CAPI_EXPORT extern BandStatus BandImageProcessorSynchronize(BandImageProcessor* image_processor);
/* n frames -> n IMAGE buffers (any BandBufferFormat, e.g. gray or NV12 outputs that a tensor cannot describe):
 * the C++ BufferProcessor::Process(input, output) of the reference, which its C API does not expose. Enqueue only
 * when `async` is non-zero. */
BAND_CAPI_EXPORT extern BandStatus BandImageProcessorProcessBuffers(BandImageProcessor* image_processor,
                                                                    BandBuffer* const* inputs,
                                                                    BandBuffer* const* outputs, size_t count,
                                                                    int async);
/* Record one batch into a CUDA graph and replay it: for launch-bound use (one camera frame per call into a fixed
 * tensor, BASELINE's single-frame latency configuration) a replay is ONE graph launch -- no validation, no plan lookup,
 * no pointer resolution. Frames and tensors must be device-resident or pinned host memory, stay where they are while
 * the recording is used, and (count > 1) be laid out at a uniform stride. The processor must outlive the recording. */
typedef struct BandRecordedBatch BandRecordedBatch;
BAND_CAPI_EXPORT extern BandRecordedBatch* BandImageProcessorRecordBatch(BandImageProcessor* image_processor,
                                                                         BandBuffer* const* buffers,
                                                                         BandTensor* const* target_tensors,
                                                                         size_t count);
BAND_CAPI_EXPORT extern BandStatus BandRecordedBatchReplay(BandRecordedBatch* recorded, int synchronize);
BAND_CAPI_EXPORT extern void BandRecordedBatchDelete(BandRecordedBatch* recorded);
/* cudaStream_t the processor launches on (NULL before its first Process*); lets a harness bracket the
 * launches with CUDA events. */
BAND_CAPI_EXPORT extern void* BandImageProcessorGetStream(BandImageProcessor* image_processor);
/* 1 when the last Process/ProcessBatch ran as a single staged fast-kernel launch per batch. */
BAND_CAPI_EXPORT extern int BandImageProcessorIsFused(BandImageProcessor* image_processor);
BAND_CAPI_EXPORT extern int BandCudaGetDeviceCount();
BAND_CAPI_EXPORT extern BandStatus BandCudaSetDevice(int device);
/* Pinned host memory for frames handed to BandBufferSetFromRawData (uploads then run at full PCIe rate). */
BAND_CAPI_EXPORT extern void* BandHostAlloc(size_t bytes);
BAND_CAPI_EXPORT extern void BandHostFree(void* ptr);
BAND_CAPI_EXPORT extern void* BandDeviceAlloc(size_t bytes);
BAND_CAPI_EXPORT extern void BandDeviceFree(void* ptr);
BAND_CAPI_EXPORT extern BandStatus BandCopyToDevice(void* device_dst, const void* host_src, size_t bytes);
BAND_CAPI_EXPORT extern BandStatus BandCopyToHost(void* host_dst, const void* device_src, size_t bytes);
/* ---- ADDITIVE: the engine's tensor ring with device-resident slots ----
 * band::TensorRingBuffer (/root/reference/band/tensor_ring_buffer.h:33-66) is what Engine::RequestAsync copies
 * model inputs into (/root/reference/band/engine.cc:506-520). Here its slots can live in HBM (one slab) and a slot
 * can be lent out, so BandImageProcessorProcess[Batch] writes model input straight into the ring:
 *     int h = BandTensorRingBufferAlloc(ring);
 *     BandImageProcessorProcess(proc, frame, BandTensorRingBufferGetSlot(ring, tensor_index, h));
 * `model_tensors` give the shape/type of each slot tensor (they are not retained); `tensor_indices[i]` is the
 * model's index of tensor i. Slot tensors are owned by the ring: never BandTensorDelete them. */
typedef struct BandTensorRingBuffer BandTensorRingBuffer;
BAND_CAPI_EXPORT extern BandTensorRingBuffer* BandTensorRingBufferCreate(BandTensor* const* model_tensors,
                                                                         const int* tensor_indices,
                                                                         size_t num_tensors, int size,
                                                                         BandMemorySpace space);
BAND_CAPI_EXPORT extern void BandTensorRingBufferDelete(BandTensorRingBuffer* ring);
BAND_CAPI_EXPORT extern int BandTensorRingBufferAlloc(BandTensorRingBuffer* ring);
BAND_CAPI_EXPORT extern int BandTensorRingBufferIsHandleValid(BandTensorRingBuffer* ring, int handle);
BAND_CAPI_EXPORT extern BandTensor* BandTensorRingBufferGetSlot(BandTensorRingBuffer* ring, int tensor_index,
                                                                int handle);
BAND_CAPI_EXPORT extern BandStatus BandTensorRingBufferPutTensor(BandTensorRingBuffer* ring, BandTensor* src,
                                                                 int tensor_index, int handle);
BAND_CAPI_EXPORT extern BandStatus BandTensorRingBufferGetTensor(BandTensorRingBuffer* ring, BandTensor* dst,
                                                                 int tensor_index, int handle);
/* ---- ADDITIVE: one batch across the GPUs of the box through ONE call ----
 * The reference processes one frame per call on the caller's thread (c_api_buffer.cc:198-214) and leaves batching as
 * a TODO (band/buffer/buffer_processor.h:54-58). Frames are independent, so a batch shards with no exchange step:
 * frame i of `count` goes to shard i / ceil(count / shards) (contiguous slices); every shard is an ordinary image
 * processor with its own plans and streams on its own device, driven by its own host thread. No collective.
 *   devices == NULL: one shard per visible GPU. Host frames / tensors (pinned or pageable) work on any shard;
 *   DEVICE-resident ones must live on the device of the shard that owns their slice (GetSlice / GetDevice). */
typedef struct BandShardedImageProcessor BandShardedImageProcessor;
BAND_CAPI_EXPORT extern BandShardedImageProcessor* BandImageProcessorBuilderBuildSharded(
    BandImageProcessorBuilder* builder, const int* devices, size_t num_devices);
BAND_CAPI_EXPORT extern void BandShardedImageProcessorDelete(BandShardedImageProcessor* sharded);
BAND_CAPI_EXPORT extern size_t BandShardedImageProcessorGetNumShards(BandShardedImageProcessor* sharded);
BAND_CAPI_EXPORT extern int BandShardedImageProcessorGetDevice(BandShardedImageProcessor* sharded, size_t shard);
BAND_CAPI_EXPORT extern void BandShardedImageProcessorGetSlice(BandShardedImageProcessor* sharded, size_t count,
                                                               size_t shard, size_t* begin, size_t* end);
/* The slicing rule on its own (no processor, no device): [begin, end) of `shard` among `shards` for `count` frames. */
BAND_CAPI_EXPORT extern void BandShardSlice(size_t count, size_t shards, size_t shard, size_t* begin, size_t* end);
/* Returns when every tensor is complete. */
BAND_CAPI_EXPORT extern BandStatus BandImageProcessorProcessBatchSharded(BandShardedImageProcessor* sharded,
                                                                         BandBuffer* const* buffers,
                                                                         BandTensor* const* target_tensors,
                                                                         size_t count);
/* Enqueue only; several batches may be queued (e.g. one per target shape). Complete after ...Synchronize. */
BAND_CAPI_EXPORT extern BandStatus BandImageProcessorProcessBatchShardedAsync(BandShardedImageProcessor* sharded,
                                                                              BandBuffer* const* buffers,
                                                                              BandTensor* const* target_tensors,
                                                                              size_t count);
BAND_CAPI_EXPORT extern BandStatus BandShardedImageProcessorSynchronize(BandShardedImageProcessor* sharded);
BAND_CAPI_EXPORT extern void* BandShardedImageProcessorGetStream(BandShardedImageProcessor* sharded, size_t shard);
BAND_CAPI_EXPORT extern int BandShardedImageProcessorIsFused(BandShardedImageProcessor* sharded, size_t shard);

/* ---- ADDITIVE: registration cache for caller-owned (pageable) host frames ----
 * BandBufferSetFromRawData receives whatever memory the application has (c_api_buffer.cc:40-53). With the cache on,
 * a frame's address range is page-locked and mapped on first sight (cudaHostRegister) and found again on every later
 * call, so a camera ring of malloc'ed buffers reaches the same in-place PCIe read path as BandHostAlloc memory.
 * Bounded by `max_bytes` (0 keeps the current budget, default 4 GiB) with least-recently-used eviction; everything is
 * unregistered when the last processor created while the cache was on is deleted. Processors created BEFORE the
 * call keep their setting. Also switched on by BAND_REGISTER_HOST=1 (budget BAND_REGISTER_HOST_MB). The registered
 * memory must stay mapped while registered: use it for long-lived frame buffers, not for transient allocations. */
BAND_CAPI_EXPORT extern BandStatus BandHostRegistrationCacheConfigure(int enable, size_t max_bytes);
BAND_CAPI_EXPORT extern void BandHostRegistrationCacheStats(size_t* registered_bytes, size_t* entries,
                                                            unsigned long long* hits,
                                                            unsigned long long* registrations);
/* Kernel launches issued so far by this process (for harness accounting). */
BAND_CAPI_EXPORT extern long long BandCudaLaunchCount();
BAND_CAPI_EXPORT extern const char* BandGetLastError();

/* ---- dlopen() typedefs (c_api_buffer.h:103-121, c_api.h:153-161) ----------------- */
typedef void (*PFN_BandTensorDelete)(BandTensor*);
typedef BandDataType (*PFN_BandTensorGetType)(BandTensor*);
typedef void* (*PFN_BandTensorGetData)(BandTensor*);
typedef size_t (*PFN_BandTensorGetNumDims)(BandTensor*);
typedef const int* (*PFN_BandTensorGetDims)(BandTensor*);
typedef size_t (*PFN_BandTensorGetBytes)(BandTensor*);
typedef const char* (*PFN_BandTensorGetName)(BandTensor*);
typedef BandQuantizationType (*PFN_BandTensorGetQuantizationType)(BandTensor*);
typedef void* (*PFN_BandTensorGetQuantizationParams)(BandTensor*);
typedef BandBuffer* (*PFN_BandBufferCreate)();
typedef void (*PFN_BandBufferDelete)(BandBuffer*);
typedef BandStatus (*PFN_BandBufferSetFromRawData)(BandBuffer*, const void*, size_t, size_t, BandBufferFormat);
typedef BandStatus (*PFN_BandBufferSetFromYUVData)(BandBuffer*, const void*, const void*, const void*, size_t, size_t,
                                                   size_t, size_t, size_t, BandBufferFormat);
typedef BandImageProcessorBuilder* (*PFN_BandImageProcessorBuilderCreate)();
typedef void (*PFN_BandImageProcessorBuilderDelete)(BandImageProcessorBuilder*);
typedef BandImageProcessor* (*PFN_BandImageProcessorBuilderBuild)(BandImageProcessorBuilder*);
typedef BandStatus (*PFN_BandAddOperator)(BandImageProcessorBuilder*, BandImageProcessorBuilderField, int, ...);
typedef BandStatus (*PFN_BandImageProcessorProcess)(BandImageProcessor*, BandBuffer*, BandTensor*);
typedef void (*PFN_BandImageProcessorDelete)(BandImageProcessor*);
typedef BandStatus (*PFN_BandImageProcessorProcessBatch)(BandImageProcessor*, BandBuffer* const*, BandTensor* const*,
                                                         size_t);

#ifdef __cplusplus
} /* extern "C" */
#endif
#endif /* BAND_INCLUDE_BAND_C_API_H_ */
