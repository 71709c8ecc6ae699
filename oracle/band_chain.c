/*
 * oracle/band_chain.c -- TEST INFRASTRUCTURE ONLY (see band_chain.h).
 * Part 1: Buffer + LibyuvImageOperator glue.
 */
#include "band_chain.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define RETURN_IF_ERROR(expr)   \
  do {                          \
    int _st = (expr);           \
    if (_st != ORC_OK) return _st; \
  } while (0)

/* band/common.cc:338-367 GetDataTypeBytes */
size_t orc_dtype_bytes(int dtype) {
  switch (dtype) {
    case ORC_DT_F32: return 4;
    case ORC_DT_I32: return 4;
    case ORC_DT_U8: return 1;
    case ORC_DT_I8: return 1;
    case ORC_DT_I16: return 2;
    case ORC_DT_I64: return 8;
    case ORC_DT_F64: return 8;
    default: return 0;
  }
}

static int is_yuv(int f) { return f == ORC_FMT_NV21 || f == ORC_FMT_NV12 || f == ORC_FMT_YV21 || f == ORC_FMT_YV12; }
/* buffer.cc:216-231 GetNumPixelElements(format): 0 for YUV/Raw */
static size_t num_pixel_elements(int f) {
  switch (f) {
    case ORC_FMT_GRAY: return 1;
    case ORC_FMT_RGB: return 3;
    case ORC_FMT_RGBA: return 4;
    default: return 0;
  }
}

/* buffer.cc:250-266 GetBufferByteSize */
size_t orc_buffer_byte_size(size_t w, size_t h, int format) {
  if (is_yuv(format)) return w * h + ((w + 1) / 2) * ((h + 1) / 2) * 2;
  return w * h * num_pixel_elements(format);
}

static orc_buffer* new_buffer(size_t w, size_t h, int format, int dtype) {
  orc_buffer* b = (orc_buffer*)calloc(1, sizeof(orc_buffer));
  b->format = format;
  b->dtype = dtype;
  b->orientation = 1; /* kTopLeft */
  b->width = w;
  b->height = h;
  return b;
}

/* buffer.cc:80-104 CreateFromYUVPlanes: plane order NV21/YV12 = {Y,V,U}, NV12/YV21 = {Y,U,V} */
orc_buffer* orc_buffer_from_yuv_planes(const uint8_t* y, const uint8_t* u, const uint8_t* v, size_t width,
                                       size_t height, size_t row_stride_y, size_t row_stride_uv,
                                       size_t pixel_stride_uv, int format) {
  if (!is_yuv(format)) return NULL;
  orc_buffer* b = new_buffer(width, height, format, ORC_DT_U8);
  b->nplanes = 3;
  b->planes[0] = (orc_plane){(uint8_t*)y, row_stride_y, 1};
  if (format == ORC_FMT_NV21 || format == ORC_FMT_YV12) {
    b->planes[1] = (orc_plane){(uint8_t*)v, row_stride_uv, pixel_stride_uv};
    b->planes[2] = (orc_plane){(uint8_t*)u, row_stride_uv, pixel_stride_uv};
  } else {
    b->planes[1] = (orc_plane){(uint8_t*)u, row_stride_uv, pixel_stride_uv};
    b->planes[2] = (orc_plane){(uint8_t*)v, row_stride_uv, pixel_stride_uv};
  }
  return b;
}

/* buffer.cc:24-78 CreateFromRaw */
orc_buffer* orc_buffer_from_raw(const uint8_t* data, size_t width, size_t height, int format, int dtype) {
  if (format <= ORC_FMT_RGBA) {
    orc_buffer* b = new_buffer(width, height, format, dtype);
    b->nplanes = 1;
    b->planes[0] = (orc_plane){(uint8_t*)data, width * num_pixel_elements(format), num_pixel_elements(format)};
    return b;
  }
  size_t uvw = (width + 1) / 2, uvh = (height + 1) / 2;
  switch (format) {
    case ORC_FMT_NV21: {
      size_t rs = (width % 2 == 1) ? (width + 1) / 2 * 2 : width;
      return orc_buffer_from_yuv_planes(data, data + width * height + 1, data + width * height, width, height,
                                        width, rs, 2, format);
    }
    case ORC_FMT_NV12: {
      size_t rs = (width % 2 == 1) ? (width + 1) / 2 * 2 : width;
      return orc_buffer_from_yuv_planes(data, data + width * height, data + width * height + 1, width, height,
                                        width, rs, 2, format);
    }
    case ORC_FMT_YV21:
      return orc_buffer_from_yuv_planes(data, data + width * height, data + width * height + uvw * uvh, width,
                                        height, width, uvw, 1, format);
    case ORC_FMT_YV12:
      return orc_buffer_from_yuv_planes(data, data + width * height + uvw * uvh, data + width * height, width,
                                        height, width, uvw, 1, format);
    default:
      return NULL;
  }
}

/* buffer.cc:106-152 CreateFromTensor */
orc_buffer* orc_buffer_from_tensor(void* data, int dtype, const int* dims_in, int ndims) {
  if (ndims == 0) return NULL;
  size_t dims[8];
  size_t nd = 0;
  size_t bytes = orc_dtype_bytes(dtype);
  for (int i = 0; i < ndims; ++i) {
    bytes *= (size_t)dims_in[i];
    if (dims_in[i] == 1) continue;
    if (dims_in[i] <= 0 || nd >= 8) return NULL;
    dims[nd++] = (size_t)dims_in[i];
  }
  if (nd == 3 && dims[2] == 3) {
    orc_buffer* b = new_buffer(dims[1], dims[0], ORC_FMT_RGB, dtype);
    b->nplanes = 1;
    b->planes[0] = (orc_plane){(uint8_t*)data, dims[1] * 3, 3 * orc_dtype_bytes(dtype)};
    return b;
  }
  orc_buffer* b = new_buffer(nd > 0 ? dims[0] : 0, nd > 1 ? dims[1] : 0, ORC_FMT_RAW, dtype);
  b->raw_ndims = nd;
  memcpy(b->raw_dims, dims, sizeof(size_t) * nd);
  b->nplanes = 1;
  b->planes[0] = (orc_plane){(uint8_t*)data, bytes, orc_dtype_bytes(dtype)};
  return b;
}

/* buffer.cc:154-197 CreateEmpty. DEVIATION (memory safety only): the
 * reference allocates W*H*channels BYTES for RGB/RGBA whatever the dtype
 * (buffer.cc:157-164); the oracle multiplies by the dtype size so that a float
 * intermediate cannot overrun. No u8 result changes. */
orc_buffer* orc_buffer_empty(size_t width, size_t height, int format, int dtype) {
  size_t total;
  if (format == ORC_FMT_RGB || format == ORC_FMT_RGBA) {
    size_t eb = orc_dtype_bytes(dtype);
    total = width * height * num_pixel_elements(format) * (eb ? eb : 1);
  } else if (format == ORC_FMT_GRAY) {
    total = width * height;
    dtype = ORC_DT_U8;
  } else if (is_yuv(format)) {
    total = width * height + ((width + 1) / 2) * ((height + 1) / 2) * 2;
    dtype = ORC_DT_U8;
  } else {
    return NULL;
  }
  uint8_t* mem = (uint8_t*)malloc(total + 64); /* +64: libyuv-style row over-read slack */
  memset(mem, 0, total + 64);
  orc_buffer* b = orc_buffer_from_raw(mem, width, height, format, dtype);
  if (!b) {
    free(mem);
    return NULL;
  }
  b->owns_data = 1;
  return b;
}

void orc_buffer_free(orc_buffer* b) {
  if (!b) return;
  if (b->owns_data) free(b->planes[0].data);
  free(b);
}

static int same_dims(const orc_buffer* a, const orc_buffer* b) {
  if (a->format == ORC_FMT_RAW || b->format == ORC_FMT_RAW) {
    if (a->format != b->format || a->raw_ndims != b->raw_ndims) return 0;
    return memcmp(a->raw_dims, b->raw_dims, sizeof(size_t) * a->raw_ndims) == 0;
  }
  return a->width == b->width && a->height == b->height;
}
/* buffer.cc:367-388 IsBufferFormatCompatible */
static int format_compatible(const orc_buffer* a, const orc_buffer* b) {
  if (a->dtype != b->dtype) return 0;
  switch (a->format) {
    case ORC_FMT_RGB:
    case ORC_FMT_RGBA:
      return b->format == ORC_FMT_RGB || b->format == ORC_FMT_RGBA;
    case ORC_FMT_NV21:
    case ORC_FMT_NV12:
    case ORC_FMT_YV21:
    case ORC_FMT_YV12:
      return is_yuv(b->format);
    default:
      return a->format == b->format;
  }
}
static size_t buffer_num_elements(const orc_buffer* b) {
  if (b->format == ORC_FMT_RAW) {
    size_t n = 1;
    for (size_t i = 0; i < b->raw_ndims; ++i) n *= b->raw_dims[i];
    return n;
  }
  return b->width * b->height;
}
/* buffer.cc:302-309 GetPixelBytes */
static size_t buffer_pixel_bytes(const orc_buffer* b) {
  if (b->format == ORC_FMT_RAW) return b->planes[0].pixel_stride_bytes;
  return num_pixel_elements(b->format) * orc_dtype_bytes(b->dtype);
}

/* libyuv_image_operator.cc:51-88 GetYuvDataFromBuffer */
typedef struct {
  size_t y_row_stride, uv_row_stride, uv_pixel_stride;
  uint8_t *y, *u, *v;
} yuv_data;
static int get_yuv(const orc_buffer* b, yuv_data* d) {
  if (!is_yuv(b->format) || b->nplanes != 3) return ORC_INVALID_ARGUMENT;
  d->y = b->planes[0].data;
  if (b->format == ORC_FMT_NV21 || b->format == ORC_FMT_YV12) {
    d->v = b->planes[1].data;
    d->u = b->planes[2].data;
  } else {
    d->u = b->planes[1].data;
    d->v = b->planes[2].data;
  }
  d->y_row_stride = b->planes[0].row_stride_bytes;
  d->uv_row_stride = b->planes[1].row_stride_bytes;
  d->uv_pixel_stride = b->planes[1].pixel_stride_bytes;
  return ORC_OK;
}
/* :933-946 GetUvRawBuffer */
static uint8_t* uv_raw(const orc_buffer* b) {
  yuv_data d;
  if ((b->format != ORC_FMT_NV12 && b->format != ORC_FMT_NV21) || get_yuv(b, &d)) return NULL;
  return b->format == ORC_FMT_NV12 ? d.u : d.v;
}

#define W(b) ((int)(b)->width)
#define H(b) ((int)(b)->height)
#define P0(b) ((b)->planes[0].data)
#define S0(b) ((int)(b)->planes[0].row_stride_bytes)

/* ---- ColorSpaceConvert (:92-373, :523-694, :1709-1727) ------------------ */
static int convert_from_yv(const orc_buffer* in, orc_buffer* out);

static int convert_from_nv(const orc_buffer* in, orc_buffer* out, int nv21) {
  yuv_data s, d;
  RETURN_IF_ERROR(get_yuv(in, &s));
  const uint8_t* chroma = nv21 ? s.v : s.u; /* NV21ToRAW gets v_buffer (:190), NV12ToRAW u_buffer (:102) */
  int ret = 0;
  switch (out->format) {
    case ORC_FMT_RGB:
      ret = nv21 ? orc_NV21ToRAW(s.y, (int)s.y_row_stride, chroma, (int)s.uv_row_stride, P0(out), S0(out), W(in), H(in))
                 : orc_NV12ToRAW(s.y, (int)s.y_row_stride, chroma, (int)s.uv_row_stride, P0(out), S0(out), W(in), H(in));
      return ret ? ORC_UNKNOWN : ORC_OK;
    case ORC_FMT_RGBA:
      ret = nv21 ? orc_NV21ToABGR(s.y, (int)s.y_row_stride, chroma, (int)s.uv_row_stride, P0(out), S0(out), W(in), H(in))
                 : orc_NV12ToABGR(s.y, (int)s.y_row_stride, chroma, (int)s.uv_row_stride, P0(out), S0(out), W(in), H(in));
      return ret ? ORC_UNKNOWN : ORC_OK;
    case ORC_FMT_YV12:
    case ORC_FMT_YV21:
      RETURN_IF_ERROR(get_yuv(out, &d));
      ret = nv21 ? orc_NV21ToI420(s.y, (int)s.y_row_stride, chroma, (int)s.uv_row_stride, d.y, (int)d.y_row_stride,
                                  d.u, (int)d.uv_row_stride, d.v, (int)d.uv_row_stride, W(out), H(out))
                 : orc_NV12ToI420(s.y, (int)s.y_row_stride, chroma, (int)s.uv_row_stride, d.y, (int)d.y_row_stride,
                                  d.u, (int)d.uv_row_stride, d.v, (int)d.uv_row_stride, W(out), H(out));
      return ret ? ORC_UNKNOWN : ORC_OK;
    case ORC_FMT_NV21:
    case ORC_FMT_NV12: {
      if ((out->format == ORC_FMT_NV21) == nv21) return ORC_INTERNAL; /* same format: default branch */
      RETURN_IF_ERROR(get_yuv(out, &d));
      orc_CopyPlane(s.y, (int)s.y_row_stride, d.y, (int)d.y_row_stride, W(in), H(in));
      /* :155-160 / :243-248 SwapUVPlane into the other order's first chroma pointer */
      orc_SwapUVPlane(chroma, (int)s.uv_row_stride, nv21 ? d.u : d.v, (int)d.uv_row_stride, (W(out) + 1) / 2,
                      (H(out) + 1) / 2);
      return ORC_OK;
    }
    case ORC_FMT_GRAY:
      orc_CopyPlane(s.y, (int)s.y_row_stride, P0(out), S0(out), W(out), H(out));
      return ORC_OK;
    default:
      return ORC_INTERNAL;
  }
}

static int convert_from_yv(const orc_buffer* in, orc_buffer* out) {
  yuv_data s, d;
  RETURN_IF_ERROR(get_yuv(in, &s));
  int ret;
  switch (out->format) {
    case ORC_FMT_RGB:
      ret = orc_I420ToRAW(s.y, (int)s.y_row_stride, s.u, (int)s.uv_row_stride, s.v, (int)s.uv_row_stride, P0(out),
                          S0(out), W(in), H(in));
      return ret ? ORC_UNKNOWN : ORC_OK;
    case ORC_FMT_RGBA:
      ret = orc_I420ToABGR(s.y, (int)s.y_row_stride, s.u, (int)s.uv_row_stride, s.v, (int)s.uv_row_stride, P0(out),
                           S0(out), W(in), H(in));
      return ret ? ORC_UNKNOWN : ORC_OK;
    case ORC_FMT_NV12:
      RETURN_IF_ERROR(get_yuv(out, &d));
      ret = orc_I420ToNV12(s.y, (int)s.y_row_stride, s.u, (int)s.uv_row_stride, s.v, (int)s.uv_row_stride, d.y,
                           (int)d.y_row_stride, d.u, (int)d.uv_row_stride, W(out), H(out));
      return ret ? ORC_UNKNOWN : ORC_OK;
    case ORC_FMT_NV21:
      RETURN_IF_ERROR(get_yuv(out, &d));
      ret = orc_I420ToNV21(s.y, (int)s.y_row_stride, s.u, (int)s.uv_row_stride, s.v, (int)s.uv_row_stride, d.y,
                           (int)d.y_row_stride, d.v, (int)d.uv_row_stride, W(out), H(out));
      return ret ? ORC_UNKNOWN : ORC_OK;
    case ORC_FMT_GRAY:
      orc_CopyPlane(s.y, (int)s.y_row_stride, P0(out), S0(out), W(out), H(out));
      return ORC_OK;
    case ORC_FMT_YV12:
    case ORC_FMT_YV21: {
      RETURN_IF_ERROR(get_yuv(out, &d));
      int uw = (W(out) + 1) / 2, uh = (H(out) + 1) / 2;
      orc_CopyPlane(s.y, (int)s.y_row_stride, d.y, (int)d.y_row_stride, W(in), H(in));
      orc_CopyPlane(s.u, (int)s.uv_row_stride, d.u, (int)d.uv_row_stride, uw, uh);
      orc_CopyPlane(s.v, (int)s.uv_row_stride, d.v, (int)d.uv_row_stride, uw, uh);
      return ORC_OK;
    }
    default:
      return ORC_INTERNAL;
  }
}

static int convert_from_rgb(const orc_buffer* in, orc_buffer* out) {
  if (out->format == ORC_FMT_GRAY) {
    return orc_RAWToJ400(P0(in), S0(in), P0(out), S0(out), W(in), H(in)) ? ORC_INTERNAL : ORC_OK;
  } else if (is_yuv(out->format)) {
    yuv_data d;
    orc_buffer* tmp = NULL;
    if (out->format == ORC_FMT_NV12 || out->format == ORC_FMT_NV21) {
      tmp = orc_buffer_empty(in->width, in->height, ORC_FMT_YV21, ORC_DT_U8); /* :548-563 */
      RETURN_IF_ERROR(get_yuv(tmp, &d));
    } else {
      RETURN_IF_ERROR(get_yuv(out, &d));
    }
    int ret = orc_RAWToI420(P0(in), S0(in), d.y, (int)d.y_row_stride, d.u, (int)d.uv_row_stride, d.v,
                            (int)d.uv_row_stride, W(in), H(in));
    int st = ret ? ORC_INTERNAL : ORC_OK;
    if (st == ORC_OK && tmp) st = convert_from_yv(tmp, out);
    orc_buffer_free(tmp);
    return st;
  }
  return ORC_INTERNAL;
}

static int convert_from_rgba(const orc_buffer* in, orc_buffer* out) {
  yuv_data d;
  switch (out->format) {
    case ORC_FMT_GRAY: {
      int row = W(in) * 4;
      uint8_t* argb = (uint8_t*)malloc((size_t)row * in->height);
      int st = orc_ABGRToARGB(P0(in), S0(in), argb, row, W(in), H(in)) ? ORC_INTERNAL : ORC_OK;
      if (st == ORC_OK) st = orc_ARGBToJ400(argb, row, P0(out), S0(out), W(in), H(in)) ? ORC_UNKNOWN : ORC_OK;
      free(argb);
      return st;
    }
    case ORC_FMT_NV12:
      RETURN_IF_ERROR(get_yuv(out, &d));
      return orc_ABGRToNV12(P0(in), S0(in), d.y, (int)d.y_row_stride, d.u, (int)d.uv_row_stride, W(in), H(in))
                 ? ORC_UNKNOWN : ORC_OK;
    case ORC_FMT_NV21:
      RETURN_IF_ERROR(get_yuv(out, &d));
      return orc_ABGRToNV21(P0(in), S0(in), d.y, (int)d.y_row_stride, d.v, (int)d.uv_row_stride, W(in), H(in))
                 ? ORC_UNKNOWN : ORC_OK;
    case ORC_FMT_YV12:
    case ORC_FMT_YV21:
      RETURN_IF_ERROR(get_yuv(out, &d));
      return orc_ABGRToI420(P0(in), S0(in), d.y, (int)d.y_row_stride, d.u, (int)d.uv_row_stride, d.v,
                            (int)d.uv_row_stride, W(in), H(in)) ? ORC_UNKNOWN : ORC_OK;
    case ORC_FMT_RGB: /* :675-687 ARGBToRGB24 on RGBA bytes = drop the 4th byte */
      return orc_ARGBToRGB24(P0(in), S0(in), P0(out), S0(out), W(in), H(in)) ? ORC_UNKNOWN : ORC_OK;
    default:
      return ORC_INTERNAL;
  }
}

static int yuvop_convert(const orc_buffer* in, orc_buffer* out) {
  switch (in->format) {
    case ORC_FMT_NV12: return convert_from_nv(in, out, 0);
    case ORC_FMT_NV21: return convert_from_nv(in, out, 1);
    case ORC_FMT_YV12:
    case ORC_FMT_YV21: return convert_from_yv(in, out);
    case ORC_FMT_RGB: return convert_from_rgb(in, out);
    case ORC_FMT_RGBA: return convert_from_rgba(in, out);
    default: return ORC_INTERNAL;
  }
}

/* ---- Resize (:376-435, :1146-1229, :1367-1380, :1627-1646) -------------- */
static int yuvop_resize(const orc_buffer* in, orc_buffer* out) {
  yuv_data s, d;
  switch (in->format) {
    case ORC_FMT_YV12:
    case ORC_FMT_YV21:
      RETURN_IF_ERROR(get_yuv(in, &s));
      RETURN_IF_ERROR(get_yuv(out, &d));
      return orc_I420Scale(s.y, (int)s.y_row_stride, s.u, (int)s.uv_row_stride, s.v, (int)s.uv_row_stride, W(in),
                           H(in), d.y, (int)d.y_row_stride, d.u, (int)d.uv_row_stride, d.v, (int)d.uv_row_stride,
                           W(out), H(out), ORC_FILTER_BILINEAR) ? ORC_UNKNOWN : ORC_OK;
    case ORC_FMT_NV12:
    case ORC_FMT_NV21: {
      RETURN_IF_ERROR(get_yuv(in, &s));
      RETURN_IF_ERROR(get_yuv(out, &d));
      const uint8_t* src_uv = s.u;
      uint8_t* dst_uv = d.u;
      if (in->format == ORC_FMT_NV21) { /* :417-420 */
        src_uv = s.v;
        dst_uv = d.v;
      }
      return orc_NV12Scale(s.y, (int)s.y_row_stride, src_uv, (int)s.uv_row_stride, W(in), H(in), d.y,
                           (int)d.y_row_stride, dst_uv, (int)d.uv_row_stride, W(out), H(out), ORC_FILTER_BILINEAR)
                 ? ORC_UNKNOWN : ORC_OK;
    }
    case ORC_FMT_RGB: { /* :1146-1187: RGB24ToARGB -> ARGBScale -> ARGBToRGB24 */
      int arow = W(in) * 4, rrow = W(out) * 4;
      uint8_t* argb = (uint8_t*)malloc((size_t)arow * in->height + 64);
      uint8_t* rargb = (uint8_t*)malloc((size_t)rrow * out->height + 64);
      int st = orc_RGB24ToARGB(P0(in), S0(in), argb, arow, W(in), H(in)) ? ORC_UNKNOWN : ORC_OK;
      if (st == ORC_OK)
        st = orc_ARGBScale(argb, arow, W(in), H(in), rargb, rrow, W(out), H(out), ORC_FILTER_BILINEAR) ? ORC_UNKNOWN : ORC_OK;
      if (st == ORC_OK) st = orc_ARGBToRGB24(rargb, rrow, P0(out), S0(out), W(out), H(out)) ? ORC_UNKNOWN : ORC_OK;
      free(argb);
      free(rargb);
      return st;
    }
    case ORC_FMT_RGBA:
      /* DEVIATION (memory safety): the reference lets an RGBA
       * input resize into an RGB output and then writes 4 B/px into it. */
      if (out->format != ORC_FMT_RGBA) return ORC_INVALID_ARGUMENT;
      return orc_ARGBScale(P0(in), S0(in), W(in), H(in), P0(out), S0(out), W(out), H(out), ORC_FILTER_BILINEAR)
                 ? ORC_UNKNOWN : ORC_OK;
    case ORC_FMT_GRAY:
      orc_ScalePlane(P0(in), S0(in), W(in), H(in), P0(out), S0(out), W(out), H(out), ORC_FILTER_BILINEAR);
      return ORC_OK;
    default:
      return ORC_INTERNAL;
  }
}

/* ---- Rotate (:697-887, :1648-1667): angle is COUNTER-clockwise ---------- */
static int libyuv_mode(int angle_deg) {
  switch (angle_deg) {
    case 90: return ORC_ROTATE_270;
    case 270: return ORC_ROTATE_90;
    case 180: return ORC_ROTATE_180;
    default: return ORC_ROTATE_0;
  }
}
static int yuvop_rotate(const orc_buffer* in, int angle_deg, orc_buffer* out) {
  yuv_data s, d;
  int mode = libyuv_mode(angle_deg % 360);
  switch (in->format) {
    case ORC_FMT_GRAY:
      return orc_RotatePlane(P0(in), S0(in), P0(out), S0(out), W(in), H(in), mode) ? ORC_UNKNOWN : ORC_OK;
    case ORC_FMT_RGBA:
      return orc_ARGBRotate(P0(in), S0(in), P0(out), S0(out), W(in), H(in), mode) ? ORC_UNKNOWN : ORC_OK;
    case ORC_FMT_RGB: { /* :730-765 via ARGB temporaries */
      int arow = W(in) * 4, rrow = W(out) * 4;
      size_t sz = (size_t)arow * in->height;
      uint8_t* argb = (uint8_t*)malloc(sz);
      uint8_t* rot = (uint8_t*)malloc(sz);
      int st = orc_RGB24ToARGB(P0(in), S0(in), argb, arow, W(in), H(in)) ? ORC_UNKNOWN : ORC_OK;
      if (st == ORC_OK) st = orc_ARGBRotate(argb, arow, rot, rrow, W(in), H(in), mode) ? ORC_UNKNOWN : ORC_OK;
      if (st == ORC_OK) st = orc_ARGBToRGB24(rot, rrow, P0(out), S0(out), W(out), H(out)) ? ORC_UNKNOWN : ORC_OK;
      free(argb);
      free(rot);
      return st;
    }
    case ORC_FMT_YV12:
    case ORC_FMT_YV21:
      RETURN_IF_ERROR(get_yuv(in, &s));
      RETURN_IF_ERROR(get_yuv(out, &d));
      return orc_I420Rotate(s.y, (int)s.y_row_stride, s.u, (int)s.uv_row_stride, s.v, (int)s.uv_row_stride, d.y,
                            (int)d.y_row_stride, d.u, (int)d.uv_row_stride, d.v, (int)d.uv_row_stride, W(in), H(in),
                            mode) ? ORC_UNKNOWN : ORC_OK;
    case ORC_FMT_NV12:
    case ORC_FMT_NV21: { /* :816-887 NV12ToI420Rotate + MergeUVPlane */
      RETURN_IF_ERROR(get_yuv(in, &s));
      RETURN_IF_ERROR(get_yuv(out, &d));
      orc_buffer* tmp = orc_buffer_empty(out->width, out->height, ORC_FMT_YV21, ORC_DT_U8);
      yuv_data t;
      RETURN_IF_ERROR(get_yuv(tmp, &t));
      const uint8_t* chroma = in->format == ORC_FMT_NV12 ? s.u : s.v;
      int st = orc_NV12ToI420Rotate(s.y, (int)s.y_row_stride, chroma, (int)s.uv_row_stride, d.y, (int)d.y_row_stride,
                                    t.u, (int)t.uv_row_stride, t.v, (int)t.uv_row_stride, W(in), H(in), mode)
                   ? ORC_UNKNOWN : ORC_OK;
      if (st == ORC_OK) {
        uint8_t* out_chroma = in->format == ORC_FMT_NV12 ? d.u : d.v;
        orc_MergeUVPlane(t.u, (int)t.uv_row_stride, t.v, (int)t.uv_row_stride, out_chroma, (int)d.uv_row_stride,
                         (W(out) + 1) / 2, (H(out) + 1) / 2);
      }
      orc_buffer_free(tmp);
      return st;
    }
    default:
      return ORC_INTERNAL;
  }
}

/* ---- Crop (:910-1106, :1383-1410, :1609-1625). The Crop operator's output
 * always equals the crop dims, so only the copy branches are reachable. ---- */
static int yuvop_crop(const orc_buffer* in, int x0, int y0, int x1, int y1, orc_buffer* out) {
  int cw = x1 - x0 + 1, ch = y1 - y0 + 1;
  if ((size_t)cw != out->width || (size_t)ch != out->height) return ORC_INTERNAL; /* crop+resize: unreachable */
  yuv_data s, d;
  switch (in->format) {
    case ORC_FMT_RGBA:
    case ORC_FMT_RGB:
    case ORC_FMT_GRAY: {
      size_t px = buffer_pixel_bytes(in);
      int off = S0(in) * y0 + x0 * (int)px;
      orc_CopyPlane(P0(in) + off, S0(in), P0(out), S0(out), cw * (int)px, ch);
      return ORC_OK;
    }
    case ORC_FMT_NV12:
    case ORC_FMT_NV21: {
      RETURN_IF_ERROR(get_yuv(in, &s));
      RETURN_IF_ERROR(get_yuv(out, &d));
      orc_CopyPlane(s.y + (int)s.y_row_stride * y0 + x0, (int)s.y_row_stride, d.y, (int)d.y_row_stride, cw, ch);
      int coff = (int)s.uv_row_stride * (y0 / 2) + (int)s.uv_pixel_stride * (x0 / 2);
      const uint8_t* ic = uv_raw(in);
      uint8_t* oc = uv_raw(out);
      if (!ic || !oc) return ORC_INTERNAL;
      orc_CopyPlane(ic + coff, (int)s.uv_row_stride, oc, (int)d.uv_row_stride, (cw + 1) / 2 * 2, (ch + 1) / 2);
      return ORC_OK;
    }
    case ORC_FMT_YV12:
    case ORC_FMT_YV21: {
      RETURN_IF_ERROR(get_yuv(in, &s));
      RETURN_IF_ERROR(get_yuv(out, &d));
      orc_CopyPlane(s.y + (int)s.y_row_stride * y0 + x0, (int)s.y_row_stride, d.y, (int)d.y_row_stride, cw, ch);
      int coff = (int)s.uv_row_stride * (y0 / 2) + (int)s.uv_pixel_stride * (x0 / 2);
      orc_CopyPlane(s.u + coff, (int)s.uv_row_stride, d.u, (int)d.uv_row_stride, (cw + 1) / 2, (ch + 1) / 2);
      orc_CopyPlane(s.v + coff, (int)s.uv_row_stride, d.v, (int)d.uv_row_stride, (cw + 1) / 2, (ch + 1) / 2);
      return ORC_OK;
    }
    default:
      return ORC_INTERNAL;
  }
}

/* ---- Flip (:890-907, :1108-1144, :1189-1363, :1669-1707) ---------------- */
static int yuvop_flip_h(const orc_buffer* in, orc_buffer* out) {
  yuv_data s, d;
  switch (in->format) {
    case ORC_FMT_RGBA:
      return orc_ARGBMirror(P0(in), S0(in), P0(out), S0(out), W(out), H(out)) ? ORC_UNKNOWN : ORC_OK;
    case ORC_FMT_RGB:
      return orc_RGB24Mirror(P0(in), S0(in), P0(out), S0(out), W(in), H(in)) ? ORC_UNKNOWN : ORC_OK;
    case ORC_FMT_GRAY:
      orc_MirrorPlane(P0(in), S0(in), P0(out), S0(out), W(out), H(out));
      return ORC_OK;
    case ORC_FMT_YV12:
    case ORC_FMT_YV21:
      RETURN_IF_ERROR(get_yuv(in, &s));
      RETURN_IF_ERROR(get_yuv(out, &d));
      return orc_I420Mirror(s.y, (int)s.y_row_stride, s.u, (int)s.uv_row_stride, s.v, (int)s.uv_row_stride, d.y,
                            (int)d.y_row_stride, d.u, (int)d.uv_row_stride, d.v, (int)d.uv_row_stride, W(in), H(in))
                 ? ORC_UNKNOWN : ORC_OK;
    case ORC_FMT_NV12:
    case ORC_FMT_NV21: {
      RETURN_IF_ERROR(get_yuv(in, &s));
      RETURN_IF_ERROR(get_yuv(out, &d));
      const uint8_t* ic = uv_raw(in);
      uint8_t* oc = uv_raw(out);
      if (!ic || !oc) return ORC_INTERNAL;
      return orc_NV12Mirror(s.y, (int)s.y_row_stride, ic, (int)s.uv_row_stride, d.y, (int)d.y_row_stride, oc,
                            (int)d.uv_row_stride, W(in), H(in)) ? ORC_UNKNOWN : ORC_OK;
    }
    default:
      return ORC_INTERNAL;
  }
}
static int yuvop_flip_v(const orc_buffer* in, orc_buffer* out) {
  yuv_data s, d;
  switch (in->format) {
    case ORC_FMT_RGBA:
    case ORC_FMT_RGB:
    case ORC_FMT_GRAY:
      orc_CopyPlane(P0(in), S0(in), P0(out), S0(out), W(out) * (int)buffer_pixel_bytes(in), -H(out));
      return ORC_OK;
    case ORC_FMT_NV12:
    case ORC_FMT_NV21: {
      RETURN_IF_ERROR(get_yuv(in, &s));
      RETURN_IF_ERROR(get_yuv(out, &d));
      orc_CopyPlane(s.y, (int)s.y_row_stride, d.y, (int)d.y_row_stride, W(in), -H(out));
      const uint8_t* ic = uv_raw(in);
      uint8_t* oc = uv_raw(out);
      if (!ic || !oc) return ORC_INTERNAL;
      orc_CopyPlane(ic, (int)s.uv_row_stride, oc, (int)d.uv_row_stride, (W(in) + 1) / 2 * 2, -((H(in) + 1) / 2));
      return ORC_OK;
    }
    case ORC_FMT_YV12:
    case ORC_FMT_YV21:
      RETURN_IF_ERROR(get_yuv(in, &s));
      RETURN_IF_ERROR(get_yuv(out, &d));
      return orc_I420Copy(s.y, (int)s.y_row_stride, s.u, (int)s.uv_row_stride, s.v, (int)s.uv_row_stride, d.y,
                          (int)d.y_row_stride, d.u, (int)d.uv_row_stride, d.v, (int)d.uv_row_stride, W(in), -H(in))
                 ? ORC_UNKNOWN : ORC_OK;
    default:
      return ORC_INTERNAL;
  }
}

/* =========================================================================
 * Part 2: operators (band/buffer/operator.cc, image_operator.cc,
 * common_operator.cc) and the processor (buffer_processor.cc,
 * image_processor.cc).
 * ========================================================================= */
struct orc_op {
  int kind;
  int i[4];
  float f[2];
  int inplace;
  int format_specified; /* ColorSpaceConvert */
  orc_buffer* output;
  int output_owned;
  int output_assigned;
  orc_buffer* flip_intermediate;
  /* AutoConvert sub-operators (image_operator.h:137-159) */
  orc_op *sub_csc, *sub_resize, *sub_dtc;
};

orc_op* orc_op_create(int kind, int i0, int i1, int i2, int i3, float f0, float f1, int inplace) {
  orc_op* op = (orc_op*)calloc(1, sizeof(orc_op));
  op->kind = kind;
  op->i[0] = i0; op->i[1] = i1; op->i[2] = i2; op->i[3] = i3;
  op->f[0] = f0; op->f[1] = f1;
  op->inplace = inplace;
  if (kind == ORC_OP_ROTATE) { /* image_operator.h:66-71 */
    int a = i0 % 360;
    if (a < 0) a += 360;
    op->i[0] = a;
  }
  if (kind == ORC_OP_COLOR_SPACE_CONVERT) {
    op->format_specified = (i0 >= 0);
    if (!op->format_specified) op->i[0] = ORC_FMT_RAW;
  }
  if (kind == ORC_OP_DATA_TYPE_CONVERT) { /* common_operator.h:50: Normalize(0,1,false) */
    op->f[0] = 0.f;
    op->f[1] = 1.f;
    op->inplace = 0;
  }
  return op;
}
static void op_drop_output(orc_op* op) {
  if (op->output && op->output_owned) orc_buffer_free(op->output);
  op->output = NULL;
  op->output_owned = 0;
}
void orc_op_free(orc_op* op) {
  if (!op) return;
  op_drop_output(op);
  orc_buffer_free(op->flip_intermediate);
  orc_op_free(op->sub_csc);
  orc_op_free(op->sub_resize);
  orc_op_free(op->sub_dtc);
  free(op);
}
orc_buffer* orc_op_output(orc_op* op) { return op->output; }

/* operator.cc:59-66 SetOutput; Normalize (common_operator.cc:71-78) refuses when
 * inplace; AutoConvert (image_operator.cc:588-595) stores without "assigned". */
void orc_op_set_output(orc_op* op, orc_buffer* output) {
  if (!output) return;
  if ((op->kind == ORC_OP_NORMALIZE || op->kind == ORC_OP_DATA_TYPE_CONVERT) && op->inplace) return;
  op_drop_output(op);
  op->output = output;
  op->output_owned = 0;
  if (op->kind != ORC_OP_AUTO_CONVERT) op->output_assigned = 1;
}
static void op_own_output(orc_op* op, orc_buffer* b) {
  op_drop_output(op);
  op->output = b;
  op->output_owned = 1;
}

/* ---- Normalize arithmetic (common_operator.cc:24-40):
 *   out[i] = static_cast<Out>((in[i] - mean)) / std
 * i.e. the cast binds to the difference, the division runs in float and the
 * quotient is converted to Out on assignment. Linear index, strides ignored. */
#define NORM_LOOP(IT, OT)                                                      \
  do {                                                                         \
    const IT* ip = (const IT*)in->planes[0].data;                              \
    OT* opx = (OT*)out->planes[0].data;                                        \
    for (size_t k = 0; k < n; ++k) opx[k] = (OT)((OT)(ip[k] - mean) / std);    \
  } while (0)
#define NORM_FROM(IT)                                            \
  do {                                                           \
    switch (out->dtype) {                                        \
      case ORC_DT_U8: NORM_LOOP(IT, uint8_t); break;             \
      case ORC_DT_I8: NORM_LOOP(IT, int8_t); break;              \
      case ORC_DT_I16: NORM_LOOP(IT, int16_t); break;            \
      case ORC_DT_I32: NORM_LOOP(IT, int32_t); break;            \
      case ORC_DT_F32: NORM_LOOP(IT, float); break;              \
      default: break;                                            \
    }                                                            \
  } while (0)
static int normalize_impl(const orc_buffer* in, orc_buffer* out, float mean, float std) {
  size_t n = buffer_num_elements(in) * num_pixel_elements(in->format);
  switch (in->dtype) {
    case ORC_DT_U8: NORM_FROM(uint8_t); break;
    case ORC_DT_I8: NORM_FROM(int8_t); break;
    case ORC_DT_I16: NORM_FROM(int16_t); break;
    case ORC_DT_I32: NORM_FROM(int32_t); break;
    case ORC_DT_F32: NORM_FROM(float); break;
    case ORC_DT_F64: NORM_FROM(double); break;
    default: return ORC_INVALID_ARGUMENT;
  }
  return ORC_OK;
}

static int auto_requires_csc(const orc_op* op, const orc_buffer* in) { return in->format != op->output->format; }
static int auto_requires_resize(const orc_op* op, const orc_buffer* in) { return !same_dims(in, op->output); }
static int auto_requires_dtc(const orc_op* op, const orc_buffer* in) { return in->dtype != op->output->dtype; }

static int op_validate_input(const orc_op* op, const orc_buffer* in) {
  switch (op->kind) {
    case ORC_OP_CROP: { /* image_operator.cc:85-115 */
      int x0 = op->i[0], y0 = op->i[1], x1 = op->i[2], y1 = op->i[3];
      if (in->format == ORC_FMT_RAW) return ORC_INVALID_ARGUMENT;
      if (x0 < 0 || y0 < 0 || x1 < 0 || y1 < 0) return ORC_INVALID_ARGUMENT;
      if (x0 >= x1 || y0 >= y1) return ORC_INVALID_ARGUMENT;
      if ((size_t)x1 >= in->width || (size_t)y1 >= in->height) return ORC_INVALID_ARGUMENT;
      if (in->dtype != ORC_DT_U8) return ORC_INVALID_ARGUMENT;
      return ORC_OK;
    }
    case ORC_OP_RESIZE: /* :166-177 */
    case ORC_OP_ROTATE: /* :249-255 */
      return in->format == ORC_FMT_RAW ? ORC_INVALID_ARGUMENT : ORC_OK;
    case ORC_OP_FLIP: /* :311-323 */
      if (!op->i[0] && !op->i[1]) return ORC_INVALID_ARGUMENT;
      return in->format == ORC_FMT_RAW ? ORC_INVALID_ARGUMENT : ORC_OK;
    case ORC_OP_NORMALIZE:
    case ORC_OP_DATA_TYPE_CONVERT: /* common_operator.cc:111-126 */
      if (is_yuv(in->format)) return ORC_INVALID_ARGUMENT;
      if (in->nplanes != 1) return ORC_INVALID_ARGUMENT;
      return ORC_OK;
    case ORC_OP_AUTO_CONVERT: /* image_operator.cc:493-501 */
      return is_yuv(in->format) ? ORC_INVALID_ARGUMENT : ORC_OK;
    default:
      return ORC_OK;
  }
}

static int op_validate_output(const orc_op* op, const orc_buffer* in) {
  const orc_buffer* out = op->output;
  switch (op->kind) {
    case ORC_OP_CROP: /* :117-144 */
      if (!format_compatible(in, out)) return ORC_INVALID_ARGUMENT;
      if (in->dtype != out->dtype) return ORC_INVALID_ARGUMENT;
      if ((size_t)(op->i[2] - op->i[0] + 1) != out->width || (size_t)(op->i[3] - op->i[1] + 1) != out->height)
        return ORC_INVALID_ARGUMENT;
      return ORC_OK;
    case ORC_OP_RESIZE: /* :179-223 */
      if (same_dims(in, out)) return ORC_INVALID_ARGUMENT;
      if (!format_compatible(in, out)) return ORC_INVALID_ARGUMENT;
      switch (in->format) {
        case ORC_FMT_GRAY: case ORC_FMT_RGB: case ORC_FMT_NV12: case ORC_FMT_NV21: case ORC_FMT_YV12: case ORC_FMT_YV21:
          if (in->format != out->format) return ORC_INVALID_ARGUMENT;
          break;
        case ORC_FMT_RGBA:
          if (out->format != ORC_FMT_RGB && out->format != ORC_FMT_RGBA) return ORC_INVALID_ARGUMENT;
          break;
        default:
          return ORC_INTERNAL;
      }
      if (op->i[0] != -1 && (size_t)op->i[0] != out->width) return ORC_INVALID_ARGUMENT;
      if (op->i[1] != -1 && (size_t)op->i[1] != out->height) return ORC_INVALID_ARGUMENT;
      return ORC_OK;
    case ORC_OP_ROTATE: { /* :257-283 */
      if (!format_compatible(in, out)) return ORC_INVALID_ARGUMENT;
      int a = op->i[0];
      int change = (a / 90) % 2 == 1;
      int rotated = in->width == out->height && in->height == out->width;
      int equal = same_dims(in, out);
      if (a > 360 || a < 0 || a % 90 != 0) return ORC_INVALID_ARGUMENT;
      if ((change && !rotated) || (!change && !equal)) return ORC_INVALID_ARGUMENT;
      return ORC_OK;
    }
    case ORC_OP_FLIP: /* :325-340 */
      if (!format_compatible(in, out)) return ORC_INVALID_ARGUMENT;
      if (in->width != out->width || in->height != out->height) return ORC_INVALID_ARGUMENT;
      return ORC_OK;
    case ORC_OP_COLOR_SPACE_CONVERT: /* :367-399 */
      if (in->format == out->format) return ORC_INVALID_ARGUMENT;
      if (in->dtype != out->dtype) return ORC_INVALID_ARGUMENT;
      switch (in->format) {
        case ORC_FMT_GRAY: return ORC_INVALID_ARGUMENT;
        case ORC_FMT_RGB: return out->format == ORC_FMT_RGBA ? ORC_INVALID_ARGUMENT : ORC_OK;
        case ORC_FMT_RGBA: case ORC_FMT_NV12: case ORC_FMT_NV21: case ORC_FMT_YV12: case ORC_FMT_YV21: return ORC_OK;
        default: return ORC_INTERNAL;
      }
    case ORC_OP_NORMALIZE:
    case ORC_OP_DATA_TYPE_CONVERT: /* common_operator.cc:128-151 */
      if (op->inplace) return ORC_OK;
      if (in->format != out->format) return ORC_INVALID_ARGUMENT;
      if (!same_dims(in, out)) return ORC_INVALID_ARGUMENT;
      return ORC_OK;
    case ORC_OP_AUTO_CONVERT: /* image_operator.cc:503-524 */
      if (auto_requires_csc(op, in) && (!op->sub_csc || !op->sub_csc->output)) return ORC_INVALID_ARGUMENT;
      if (auto_requires_resize(op, in) && (!op->sub_resize || !op->sub_resize->output)) return ORC_INVALID_ARGUMENT;
      if (auto_requires_dtc(op, in) && (!op->sub_dtc || !op->sub_dtc->output)) return ORC_INVALID_ARGUMENT;
      return ORC_OK;
    default:
      return ORC_INTERNAL;
  }
}

static int op_create_output(orc_op* op, const orc_buffer* in);
static int op_process_impl(orc_op* op, const orc_buffer* in);

/* operator.cc:75-97 ValidateOrCreateOutput */
static int op_validate_or_create_output(orc_op* op, const orc_buffer* in) {
  int st = op->output == NULL ? ORC_INTERNAL : op_validate_output(op, in);
  if (st != ORC_OK) {
    if (op->output_assigned) return st;
    RETURN_IF_ERROR(op_create_output(op, in));
    return op_validate_output(op, in);
  }
  return ORC_OK;
}
/* operator.cc:52-57 Process */
int orc_op_process(orc_op* op, const orc_buffer* in) {
  RETURN_IF_ERROR(op_validate_input(op, in));
  RETURN_IF_ERROR(op_validate_or_create_output(op, in));
  return op_process_impl(op, in);
}

static int op_create_output(orc_op* op, const orc_buffer* in) {
  switch (op->kind) {
    case ORC_OP_CROP: /* :146-157 */
      op_own_output(op, orc_buffer_empty((size_t)(op->i[2] - op->i[0] + 1), (size_t)(op->i[3] - op->i[1] + 1),
                                         in->format, in->dtype));
      return ORC_OK;
    case ORC_OP_RESIZE: /* :225-239 */
      if (op->i[0] == -1 || op->i[1] == -1) return ORC_INVALID_ARGUMENT;
      op_own_output(op, orc_buffer_empty((size_t)op->i[0], (size_t)op->i[1], in->format, in->dtype));
      return ORC_OK;
    case ORC_OP_ROTATE: { /* :285-297 */
      int change = (op->i[0] / 90) % 2 == 1;
      op_own_output(op, orc_buffer_empty(change ? in->height : in->width, change ? in->width : in->height,
                                         in->format, in->dtype));
      return ORC_OK;
    }
    case ORC_OP_FLIP: /* :342-357 */
      if (op->i[0] && op->i[1]) {
        orc_buffer_free(op->flip_intermediate);
        op->flip_intermediate = orc_buffer_empty(in->width, in->height, in->format, in->dtype);
      }
      op_own_output(op, orc_buffer_empty(in->width, in->height, in->format, in->dtype));
      return ORC_OK;
    case ORC_OP_COLOR_SPACE_CONVERT: { /* :401-419 */
      if (!op->format_specified) return ORC_INVALID_ARGUMENT;
      int dt = is_yuv(op->i[0]) ? ORC_DT_U8 : in->dtype;
      op_own_output(op, orc_buffer_empty(in->width, in->height, op->i[0], dt));
      return ORC_OK;
    }
    case ORC_OP_NORMALIZE:
    case ORC_OP_DATA_TYPE_CONVERT: /* common_operator.cc:153-163 */
      if (!op->inplace) op_own_output(op, orc_buffer_empty(in->width, in->height, in->format, in->dtype));
      return ORC_OK;
    case ORC_OP_AUTO_CONVERT: { /* image_operator.cc:526-586 */
      orc_op* last = NULL;
      const orc_buffer* cur = in;
      /* DEVIATION (memory safety): sub-operators are rebuilt on every
       * CreateOutput instead of keeping stale output pointers. */
      orc_op_free(op->sub_csc); op->sub_csc = NULL;
      orc_op_free(op->sub_resize); op->sub_resize = NULL;
      orc_op_free(op->sub_dtc); op->sub_dtc = NULL;
      if (auto_requires_csc(op, in)) {
        op->sub_csc = orc_op_create(ORC_OP_COLOR_SPACE_CONVERT, op->output->format, 0, 0, 0, 0, 0, 0);
        RETURN_IF_ERROR(op_create_output(op->sub_csc, cur));
        cur = op->sub_csc->output;
        last = op->sub_csc;
      }
      if (auto_requires_resize(op, cur)) {
        op->sub_resize = orc_op_create(ORC_OP_RESIZE, (int)op->output->width, (int)op->output->height, 0, 0, 0, 0, 0);
        RETURN_IF_ERROR(op_create_output(op->sub_resize, cur));
        cur = op->sub_resize->output;
        last = op->sub_resize;
      }
      if (auto_requires_dtc(op, cur)) {
        op->sub_dtc = orc_op_create(ORC_OP_DATA_TYPE_CONVERT, 0, 0, 0, 0, 0, 0, 0);
        last = op->sub_dtc;
      }
      if (last) orc_op_set_output(last, op->output);
      cur = in;
      if (auto_requires_csc(op, in)) {
        if (!op->sub_csc->output) return ORC_INVALID_ARGUMENT;
        RETURN_IF_ERROR(op_validate_output(op->sub_csc, in));
        cur = op->sub_csc->output;
      }
      if (auto_requires_resize(op, cur)) {
        if (!op->sub_resize->output) return ORC_INVALID_ARGUMENT;
        RETURN_IF_ERROR(op_validate_output(op->sub_resize, cur));
        cur = op->sub_resize->output;
      }
      if (auto_requires_dtc(op, cur)) RETURN_IF_ERROR(op_validate_output(op->sub_dtc, cur));
      return ORC_OK;
    }
    default:
      return ORC_INTERNAL;
  }
}

static int op_process_impl(orc_op* op, const orc_buffer* in) {
  switch (op->kind) {
    case ORC_OP_CROP:
      return yuvop_crop(in, op->i[0], op->i[1], op->i[2], op->i[3], op->output);
    case ORC_OP_RESIZE:
      return yuvop_resize(in, op->output);
    case ORC_OP_ROTATE:
      return yuvop_rotate(in, op->i[0], op->output);
    case ORC_OP_FLIP: /* :299-309 */
      if (op->i[0] && op->i[1]) {
        if (!op->flip_intermediate) return ORC_INTERNAL; /* reference would dereference null here */
        RETURN_IF_ERROR(yuvop_flip_h(in, op->flip_intermediate));
        return yuvop_flip_v(op->flip_intermediate, op->output);
      } else if (op->i[0]) {
        return yuvop_flip_h(in, op->output);
      }
      return yuvop_flip_v(in, op->output);
    case ORC_OP_COLOR_SPACE_CONVERT:
      return yuvop_convert(in, op->output);
    case ORC_OP_NORMALIZE:
      return normalize_impl(in, op->inplace ? (orc_buffer*)in : op->output, op->f[0], op->f[1]);
    case ORC_OP_DATA_TYPE_CONVERT: /* common_operator.cc:169-175 */
      if (in->dtype != op->output->dtype) return normalize_impl(in, op->output, op->f[0], op->f[1]);
      return ORC_OK;
    case ORC_OP_AUTO_CONVERT: { /* image_operator.cc:465-491 */
      int processed = 0;
      const orc_buffer* cur = in;
      if (auto_requires_csc(op, in)) {
        RETURN_IF_ERROR(orc_op_process(op->sub_csc, in));
        cur = op->sub_csc->output;
        processed = 1;
      }
      if (auto_requires_resize(op, cur)) {
        RETURN_IF_ERROR(orc_op_process(op->sub_resize, cur));
        cur = op->sub_resize->output;
        processed = 1;
      }
      if (auto_requires_dtc(op, cur)) {
        RETURN_IF_ERROR(orc_op_process(op->sub_dtc, cur));
        processed = 1;
      }
      if (!processed) { /* Buffer::CopyFrom, buffer.cc:331-358 */
        if (op->output->dtype != in->dtype || op->output->format != in->format || !same_dims(op->output, in) ||
            is_yuv(in->format))
          return ORC_INVALID_ARGUMENT;
        memcpy(op->output->planes[0].data, in->planes[0].data, buffer_pixel_bytes(in) * buffer_num_elements(in));
      }
      return ORC_OK;
    }
    default:
      return ORC_INTERNAL;
  }
}

/* ---- processor ---------------------------------------------------------- */
struct orc_processor {
  int n;
  orc_op** ops;
};

/* image_processor.cc:25-67 Build: one operator per C++ type, empty => AutoConvert */
orc_processor* orc_processor_build(const orc_op_desc* descs, int n, int* status) {
  if (n < 0 || n > 64) {
    if (status) *status = ORC_INVALID_ARGUMENT;
    return NULL;
  }
  for (int i = 0; i < n; ++i)
    for (int j = i + 1; j < n; ++j)
      if (descs[i].kind == descs[j].kind) {
        if (status) *status = ORC_INVALID_ARGUMENT;
        return NULL;
      }
  orc_processor* p = (orc_processor*)calloc(1, sizeof(orc_processor));
  if (n == 0) {
    p->n = 1;
    p->ops = (orc_op**)calloc(1, sizeof(orc_op*));
    p->ops[0] = orc_op_create(ORC_OP_AUTO_CONVERT, 0, 0, 0, 0, 0, 0, 0);
  } else {
    p->n = n;
    p->ops = (orc_op**)calloc((size_t)n, sizeof(orc_op*));
    for (int i = 0; i < n; ++i)
      p->ops[i] = orc_op_create(descs[i].kind, descs[i].i[0], descs[i].i[1], descs[i].i[2], descs[i].i[3],
                                descs[i].f[0], descs[i].f[1], descs[i].inplace);
  }
  if (status) *status = ORC_OK;
  return p;
}
/* buffer_processor.cc:22-43 Process */
int orc_processor_process(orc_processor* p, const orc_buffer* input, orc_buffer* output) {
  if (p->n == 0) return ORC_INTERNAL;
  orc_op* lastop = p->ops[p->n - 1];
  orc_op_set_output(lastop, output);
  if (lastop->kind == ORC_OP_AUTO_CONVERT) { /* see DEVIATION in op_create_output */
    orc_op_free(lastop->sub_csc); lastop->sub_csc = NULL;
    orc_op_free(lastop->sub_resize); lastop->sub_resize = NULL;
    orc_op_free(lastop->sub_dtc); lastop->sub_dtc = NULL;
  }
  const orc_buffer* next = input;
  for (int i = 0; i < p->n; ++i) {
    RETURN_IF_ERROR(orc_op_process(p->ops[i], next));
    next = p->ops[i]->inplace ? next : p->ops[i]->output;
    if (!next) return ORC_INTERNAL;
  }
  return ORC_OK;
}
void orc_processor_free(orc_processor* p) {
  if (!p) return;
  for (int i = 0; i < p->n; ++i) orc_op_free(p->ops[i]);
  free(p->ops);
  free(p);
}

int orc_run_batch(const orc_op_desc* descs, int nops, const uint8_t* frames, size_t frame_bytes, size_t width,
                  size_t height, int format, uint8_t* outs, size_t out_bytes, int out_dtype, const int* out_dims,
                  int out_ndims, int n) {
  int st = ORC_OK;
  orc_processor* p = orc_processor_build(descs, nops, &st);
  if (!p) return st;
  for (int k = 0; k < n && st == ORC_OK; ++k) {
    orc_buffer* in = orc_buffer_from_raw(frames + (size_t)k * frame_bytes, width, height, format, ORC_DT_U8);
    orc_buffer* out = orc_buffer_from_tensor(outs + (size_t)k * out_bytes, out_dtype, out_dims, out_ndims);
    if (!in || !out) st = ORC_INTERNAL;
    else st = orc_processor_process(p, in, out);
    orc_buffer_free(in);
    orc_buffer_free(out);
  }
  orc_processor_free(p);
  return st;
}
