/*
 * oracle/band_chain.h -- TEST INFRASTRUCTURE ONLY.
 *
 * C restatement of the reference's band/buffer operator-chain semantics
 * (Buffer, IBufferOperator protocol, the operator classes, BufferProcessor,
 * ImageProcessorBuilder) on top of the libyuv restatement in yuvref.h.
 * Every stage materialises its output exactly like the reference does; this is
 * the checker for the fused CUDA plans and the timed `cpu_baseline`.
 * Follows: band/buffer/buffer.cc, operator.cc, image_operator.cc,
 * common_operator.cc, buffer_processor.cc, image_processor.cc,
 * libyuv_image_operator.cc (cited per function in band_chain.c).
 */
#ifndef BAND_ORACLE_BAND_CHAIN_H_
#define BAND_ORACLE_BAND_CHAIN_H_

#include <stddef.h>
#include <stdint.h>

#include "yuvref.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct orc_plane {
  uint8_t* data;
  size_t row_stride_bytes;
  size_t pixel_stride_bytes;
} orc_plane;

/* band::Buffer (band/buffer/buffer.h:15-95) */
typedef struct orc_buffer {
  int format;      /* ORC_FMT_* */
  int dtype;       /* ORC_DT_*  */
  int orientation; /* carried, never applied */
  size_t width, height; /* dimension_ = {W, H} (for kRaw: dims as given) */
  size_t raw_ndims;     /* kRaw only */
  size_t raw_dims[8];
  int nplanes;
  orc_plane planes[3];
  int owns_data;
} orc_buffer;

orc_buffer* orc_buffer_from_raw(const uint8_t* data, size_t width, size_t height, int format, int dtype);
orc_buffer* orc_buffer_from_yuv_planes(const uint8_t* y, const uint8_t* u, const uint8_t* v, size_t width,
                                       size_t height, size_t row_stride_y, size_t row_stride_uv,
                                       size_t pixel_stride_uv, int format);
orc_buffer* orc_buffer_from_tensor(void* data, int dtype, const int* dims, int ndims);
orc_buffer* orc_buffer_empty(size_t width, size_t height, int format, int dtype);
void orc_buffer_free(orc_buffer* b);
size_t orc_buffer_byte_size(size_t width, size_t height, int format);
size_t orc_dtype_bytes(int dtype);

enum {
  ORC_OP_CROP = 0, /* values follow BandImageProcessorBuilderField (c_api_type.h:178-186) */
  ORC_OP_RESIZE,
  ORC_OP_ROTATE,
  ORC_OP_FLIP,
  ORC_OP_COLOR_SPACE_CONVERT,
  ORC_OP_NORMALIZE,
  ORC_OP_DATA_TYPE_CONVERT,
  ORC_OP_AUTO_CONVERT
};

typedef struct orc_op orc_op;
/* i0..i3: Crop x0,y0,x1,y1 | Resize w,h | Rotate angle | Flip horizontal,vertical |
 * ColorSpaceConvert format (or -1 = unspecified). f0,f1: Normalize mean,std.
 * inplace: Normalize only. */
orc_op* orc_op_create(int kind, int i0, int i1, int i2, int i3, float f0, float f1, int inplace);
void orc_op_free(orc_op* op);
int orc_op_process(orc_op* op, const orc_buffer* input); /* IBufferOperator::Process */
void orc_op_set_output(orc_op* op, orc_buffer* output);
orc_buffer* orc_op_output(orc_op* op);

typedef struct orc_processor orc_processor;
/* ImageProcessorBuilder::Build over (kind,i[4],f[2]) descriptors; *status gets
 * the absl code (duplicate op type => INVALID_ARGUMENT, NULL returned). */
typedef struct orc_op_desc {
  int kind;
  int i[4];
  float f[2];
  int inplace;
} orc_op_desc;
orc_processor* orc_processor_build(const orc_op_desc* descs, int n, int* status);
int orc_processor_process(orc_processor* p, const orc_buffer* input, orc_buffer* output);
void orc_processor_free(orc_processor* p);

/* Convenience for the timed CPU baseline: run `n` frames laid out back to back
 * (frame_bytes apart) through a fresh processor per call, writing outputs
 * out_bytes apart. Returns the first non-OK status. */
int orc_run_batch(const orc_op_desc* descs, int nops, const uint8_t* frames, size_t frame_bytes, size_t width,
                  size_t height, int format, uint8_t* outs, size_t out_bytes, int out_dtype, const int* out_dims,
                  int out_ndims, int n);

#ifdef __cplusplus
}
#endif
#endif
