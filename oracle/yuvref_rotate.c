/*
 * oracle/yuvref_rotate.c -- TEST INFRASTRUCTURE ONLY (see yuvref.h header).
 *
 * Restates libyuv's rotate/mirror entry points (source/rotate.cc,
 * rotate_argb.cc, rotate_common.cc, planar_functions.cc). Pure permutations.
 * `mode` is libyuv's CLOCKWISE RotationMode; the reference maps its
 * counter-clockwise angle to it in libyuv_image_operator.cc:697-708.
 * Call sites: :719/:754 (ARGBRotate), :774 (RotatePlane), :796 (I420Rotate),
 * :858 (NV12ToI420Rotate), :1115 (ARGBMirror), :1137 (MirrorPlane),
 * :1199 (RGB24Mirror), :1256 (NV12Mirror), :1282 (I420Mirror).
 */
#include <string.h>

#include "yuvref.h"

/* Rotate a plane of `bpp`-byte elements. width/height are the SOURCE dims.
 *   90  (rotate.cc RotatePlane90: transpose reading bottom-to-top):
 *        dst[r][c] = src[H-1-c][r]       dst is H wide, W tall
 *   270 (RotatePlane270: transpose writing bottom-to-top):
 *        dst[r][c] = src[c][W-1-r]
 *   180: dst[r][c] = src[H-1-r][W-1-c] */
static int rotate_elems(const uint8_t* src, int src_stride, uint8_t* dst, int dst_stride, int width,
                        int height, int mode, int bpp) {
  if (!src || width <= 0 || height == 0 || !dst) return -1;
  if (height < 0) return -1; /* not used by the reference */
  switch (mode) {
    case ORC_ROTATE_0:
      orc_CopyPlane(src, src_stride, dst, dst_stride, width * bpp, height);
      return 0;
    case ORC_ROTATE_90:
      for (int r = 0; r < width; ++r)
        for (int c = 0; c < height; ++c)
          memcpy(dst + (ptrdiff_t)r * dst_stride + (ptrdiff_t)c * bpp,
                 src + (ptrdiff_t)(height - 1 - c) * src_stride + (ptrdiff_t)r * bpp, (size_t)bpp);
      return 0;
    case ORC_ROTATE_270:
      for (int r = 0; r < width; ++r)
        for (int c = 0; c < height; ++c)
          memcpy(dst + (ptrdiff_t)r * dst_stride + (ptrdiff_t)c * bpp,
                 src + (ptrdiff_t)c * src_stride + (ptrdiff_t)(width - 1 - r) * bpp, (size_t)bpp);
      return 0;
    case ORC_ROTATE_180:
      for (int r = 0; r < height; ++r)
        for (int c = 0; c < width; ++c)
          memcpy(dst + (ptrdiff_t)r * dst_stride + (ptrdiff_t)c * bpp,
                 src + (ptrdiff_t)(height - 1 - r) * src_stride + (ptrdiff_t)(width - 1 - c) * bpp, (size_t)bpp);
      return 0;
    default:
      return -1;
  }
}

int orc_RotatePlane(const uint8_t* src, int src_stride, uint8_t* dst, int dst_stride, int width, int height,
                    int mode) {
  return rotate_elems(src, src_stride, dst, dst_stride, width, height, mode, 1);
}
int orc_ARGBRotate(const uint8_t* src, int src_stride, uint8_t* dst, int dst_stride, int width, int height,
                   int mode) {
  return rotate_elems(src, src_stride, dst, dst_stride, width, height, mode, 4);
}
int orc_I420Rotate(const uint8_t* src_y, int src_stride_y, const uint8_t* src_u, int src_stride_u,
                   const uint8_t* src_v, int src_stride_v, uint8_t* dst_y, int dst_stride_y, uint8_t* dst_u,
                   int dst_stride_u, uint8_t* dst_v, int dst_stride_v, int width, int height, int mode) {
  int halfwidth = (width + 1) >> 1;
  int halfheight = (height + 1) >> 1;
  if (!src_y || !src_u || !src_v || width <= 0 || height == 0 || !dst_y || !dst_u || !dst_v) return -1;
  if (rotate_elems(src_y, src_stride_y, dst_y, dst_stride_y, width, height, mode, 1)) return -1;
  if (rotate_elems(src_u, src_stride_u, dst_u, dst_stride_u, halfwidth, halfheight, mode, 1)) return -1;
  return rotate_elems(src_v, src_stride_v, dst_v, dst_stride_v, halfwidth, halfheight, mode, 1);
}
/* rotate.cc NV12ToI420Rotate: rotate Y; split + rotate the interleaved plane
 * (SplitUVPlane / SplitRotateUV90/180/270): first byte of each pair -> dst_u. */
int orc_NV12ToI420Rotate(const uint8_t* src_y, int src_stride_y, const uint8_t* src_uv, int src_stride_uv,
                         uint8_t* dst_y, int dst_stride_y, uint8_t* dst_u, int dst_stride_u, uint8_t* dst_v,
                         int dst_stride_v, int width, int height, int mode) {
  int hw = (width + 1) >> 1;
  int hh = (height + 1) >> 1;
  if (!src_y || !src_uv || width <= 0 || height == 0 || !dst_y || !dst_u || !dst_v) return -1;
  if (rotate_elems(src_y, src_stride_y, dst_y, dst_stride_y, width, height, mode, 1)) return -1;
  for (int sr = 0; sr < hh; ++sr) {
    for (int sc = 0; sc < hw; ++sc) {
      int dr, dc;
      switch (mode) {
        case ORC_ROTATE_0: dr = sr; dc = sc; break;
        case ORC_ROTATE_90: dr = sc; dc = hh - 1 - sr; break;   /* dst[r][c] = src[H-1-c][r] */
        case ORC_ROTATE_270: dr = hw - 1 - sc; dc = sr; break;  /* dst[r][c] = src[c][W-1-r] */
        case ORC_ROTATE_180: dr = hh - 1 - sr; dc = hw - 1 - sc; break;
        default: return -1;
      }
      const uint8_t* p = src_uv + (ptrdiff_t)sr * src_stride_uv + 2 * sc;
      dst_u[(ptrdiff_t)dr * dst_stride_u + dc] = p[0];
      dst_v[(ptrdiff_t)dr * dst_stride_v + dc] = p[1];
    }
  }
  return 0;
}

/* Horizontal mirror of `bpp`-byte elements: dst[r][c] = src[r][W-1-c]. */
static void mirror_elems(const uint8_t* src, int src_stride, uint8_t* dst, int dst_stride, int width, int height,
                         int bpp) {
  if (height < 0) { /* planar_functions.cc: negative height also flips vertically */
    height = -height;
    src = src + (ptrdiff_t)(height - 1) * src_stride;
    src_stride = -src_stride;
  }
  for (int r = 0; r < height; ++r) {
    for (int c = 0; c < width; ++c)
      memcpy(dst + (ptrdiff_t)r * dst_stride + (ptrdiff_t)c * bpp,
             src + (ptrdiff_t)r * src_stride + (ptrdiff_t)(width - 1 - c) * bpp, (size_t)bpp);
  }
}
void orc_MirrorPlane(const uint8_t* src, int src_stride, uint8_t* dst, int dst_stride, int width, int height) {
  mirror_elems(src, src_stride, dst, dst_stride, width, height, 1);
}
int orc_ARGBMirror(const uint8_t* src, int src_stride, uint8_t* dst, int dst_stride, int width, int height) {
  if (!src || !dst || width <= 0 || height == 0) return -1;
  mirror_elems(src, src_stride, dst, dst_stride, width, height, 4);
  return 0;
}
int orc_RGB24Mirror(const uint8_t* src, int src_stride, uint8_t* dst, int dst_stride, int width, int height) {
  if (!src || !dst || width <= 0 || height == 0) return -1;
  mirror_elems(src, src_stride, dst, dst_stride, width, height, 3);
  return 0;
}
int orc_NV12Mirror(const uint8_t* src_y, int src_stride_y, const uint8_t* src_uv, int src_stride_uv,
                   uint8_t* dst_y, int dst_stride_y, uint8_t* dst_uv, int dst_stride_uv, int width, int height) {
  int halfwidth = (width + 1) >> 1;
  int halfheight = (height + 1) >> 1;
  if (!src_y || !src_uv || !dst_uv || width <= 0 || height == 0) return -1;
  if (dst_y) mirror_elems(src_y, src_stride_y, dst_y, dst_stride_y, width, height, 1);
  mirror_elems(src_uv, src_stride_uv, dst_uv, dst_stride_uv, halfwidth, halfheight, 2);
  return 0;
}
int orc_I420Mirror(const uint8_t* src_y, int src_stride_y, const uint8_t* src_u, int src_stride_u,
                   const uint8_t* src_v, int src_stride_v, uint8_t* dst_y, int dst_stride_y, uint8_t* dst_u,
                   int dst_stride_u, uint8_t* dst_v, int dst_stride_v, int width, int height) {
  int halfwidth = (width + 1) >> 1;
  int halfheight = (height + 1) >> 1;
  if (!src_y || !src_u || !src_v || !dst_u || !dst_v || width <= 0 || height == 0) return -1;
  if (dst_y) mirror_elems(src_y, src_stride_y, dst_y, dst_stride_y, width, height, 1);
  mirror_elems(src_u, src_stride_u, dst_u, dst_stride_u, halfwidth, halfheight, 1);
  mirror_elems(src_v, src_stride_v, dst_v, dst_stride_v, halfwidth, halfheight, 1);
  return 0;
}
