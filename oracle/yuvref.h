/*
 * oracle/yuvref.h -- TEST INFRASTRUCTURE ONLY (CPU oracle; never linked into
 * the product libraries).
 *
 * A plain-C restatement of the arithmetic the reference's band/buffer path
 * delegates to libyuv, plus the reference's own operator-chain semantics.
 *
 * Where the algorithm lives: NOT in /root/reference. The reference fetches
 *   @libyuv = github.com/mrsnu/libyuv tag v1.0.0
 *   (band/workspace_repo.bzl:60-66, sha256 27ea6dde...), min API level 1747
 *   (band/buffer/libyuv_image_operator.cc:1197-1210).
 * That archive is absent (no network), so every function below restates the
 * PUBLISHED libyuv *_C row/scale/rotate algorithms (x86 flavour of the
 * platform-dependent macros) and cites (a) the libyuv entry point it restates
 * and (b) the reference call site that reaches it.
 *
 * PARITY PINNING: pinned against the reference's own fixtures for the ARGB
 * bilinear scaler (hippo_resize_256), crop (hippo_crop_256, 3x3 KAT), RGB->gray
 * (+-1 vs float luma) and rotate round trip (band/test/image_operation_test.cc).
 * YUV->RGB, Normalize numerics, plane-scaler special cases and rotate direction
 * are NOT pinned by any reference test: for those this file is the frozen spec
 * ("parity unpinned").
 *
 * VECTOR ROWS: the three full-frame passes of the timed CPU baseline (4:2:0 ->
 * RGB / RGBA, RGB24 -> ARGB, InterpolateRow) also exist as AVX2 rows when the
 * library is built with AVX2 (the default, -march=x86-64-v3). The scalar forms
 * stay the definition; the vector forms are compared with them on all 2^24
 * (y, u, v) triples (orc_yuv_rows_selftest) and through whole chains against a
 * no-AVX2 build of the same sources (tests/test_oracle_golden.py).
 */
#ifndef BAND_ORACLE_YUVREF_H_
#define BAND_ORACLE_YUVREF_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- enums: numeric values follow band/common.h:115-161 ---------------- */
enum {
  ORC_FMT_GRAY = 0,
  ORC_FMT_RGB = 1,
  ORC_FMT_RGBA = 2,
  ORC_FMT_YV12 = 3,
  ORC_FMT_YV21 = 4,
  ORC_FMT_NV21 = 5,
  ORC_FMT_NV12 = 6,
  ORC_FMT_RAW = 7
};
enum { /* band/common.h:115-128 DataType */
  ORC_DT_NOTYPE = 0,
  ORC_DT_F32 = 1,
  ORC_DT_I32 = 2,
  ORC_DT_U8 = 3,
  ORC_DT_I64 = 4,
  ORC_DT_I16 = 7,
  ORC_DT_I8 = 9,
  ORC_DT_F64 = 11
};
enum { /* absl::StatusCode values used by the path */
  ORC_OK = 0,
  ORC_UNKNOWN = 2,
  ORC_INVALID_ARGUMENT = 3,
  ORC_INTERNAL = 13
};
enum { ORC_FILTER_NONE = 0, ORC_FILTER_LINEAR = 1, ORC_FILTER_BILINEAR = 2, ORC_FILTER_BOX = 3 };
enum { ORC_ROTATE_0 = 0, ORC_ROTATE_90 = 90, ORC_ROTATE_180 = 180, ORC_ROTATE_270 = 270 };

/* ---- libyuv conversions (convert_argb.cc / convert.cc / row_common.cc) -- */
int orc_NV12ToRAW(const uint8_t* src_y, int src_stride_y, const uint8_t* src_uv, int src_stride_uv,
                  uint8_t* dst_raw, int dst_stride_raw, int width, int height);
int orc_NV21ToRAW(const uint8_t* src_y, int src_stride_y, const uint8_t* src_vu, int src_stride_vu,
                  uint8_t* dst_raw, int dst_stride_raw, int width, int height);
int orc_NV12ToABGR(const uint8_t* src_y, int src_stride_y, const uint8_t* src_uv, int src_stride_uv,
                   uint8_t* dst_abgr, int dst_stride_abgr, int width, int height);
int orc_NV21ToABGR(const uint8_t* src_y, int src_stride_y, const uint8_t* src_vu, int src_stride_vu,
                   uint8_t* dst_abgr, int dst_stride_abgr, int width, int height);
int orc_I420ToRAW(const uint8_t* src_y, int src_stride_y, const uint8_t* src_u, int src_stride_u,
                  const uint8_t* src_v, int src_stride_v, uint8_t* dst_raw, int dst_stride_raw,
                  int width, int height);
int orc_I420ToABGR(const uint8_t* src_y, int src_stride_y, const uint8_t* src_u, int src_stride_u,
                   const uint8_t* src_v, int src_stride_v, uint8_t* dst_abgr, int dst_stride_abgr,
                   int width, int height);
int orc_NV12ToI420(const uint8_t* src_y, int src_stride_y, const uint8_t* src_uv, int src_stride_uv,
                   uint8_t* dst_y, int dst_stride_y, uint8_t* dst_u, int dst_stride_u, uint8_t* dst_v,
                   int dst_stride_v, int width, int height);
int orc_NV21ToI420(const uint8_t* src_y, int src_stride_y, const uint8_t* src_vu, int src_stride_vu,
                   uint8_t* dst_y, int dst_stride_y, uint8_t* dst_u, int dst_stride_u, uint8_t* dst_v,
                   int dst_stride_v, int width, int height);
int orc_I420ToNV12(const uint8_t* src_y, int src_stride_y, const uint8_t* src_u, int src_stride_u,
                   const uint8_t* src_v, int src_stride_v, uint8_t* dst_y, int dst_stride_y,
                   uint8_t* dst_uv, int dst_stride_uv, int width, int height);
int orc_I420ToNV21(const uint8_t* src_y, int src_stride_y, const uint8_t* src_u, int src_stride_u,
                   const uint8_t* src_v, int src_stride_v, uint8_t* dst_y, int dst_stride_y,
                   uint8_t* dst_vu, int dst_stride_vu, int width, int height);
int orc_RGB24ToARGB(const uint8_t* src, int src_stride, uint8_t* dst, int dst_stride, int width, int height);
int orc_ARGBToRGB24(const uint8_t* src, int src_stride, uint8_t* dst, int dst_stride, int width, int height);
int orc_ABGRToARGB(const uint8_t* src, int src_stride, uint8_t* dst, int dst_stride, int width, int height);
int orc_RAWToJ400(const uint8_t* src_raw, int src_stride_raw, uint8_t* dst_yj, int dst_stride_yj, int width, int height);
int orc_ARGBToJ400(const uint8_t* src_argb, int src_stride_argb, uint8_t* dst_yj, int dst_stride_yj, int width, int height);
int orc_RAWToI420(const uint8_t* src_raw, int src_stride_raw, uint8_t* dst_y, int dst_stride_y, uint8_t* dst_u,
                  int dst_stride_u, uint8_t* dst_v, int dst_stride_v, int width, int height);
int orc_ABGRToI420(const uint8_t* src_abgr, int src_stride_abgr, uint8_t* dst_y, int dst_stride_y, uint8_t* dst_u,
                   int dst_stride_u, uint8_t* dst_v, int dst_stride_v, int width, int height);
int orc_ABGRToNV12(const uint8_t* src_abgr, int src_stride_abgr, uint8_t* dst_y, int dst_stride_y,
                   uint8_t* dst_uv, int dst_stride_uv, int width, int height);
int orc_ABGRToNV21(const uint8_t* src_abgr, int src_stride_abgr, uint8_t* dst_y, int dst_stride_y,
                   uint8_t* dst_vu, int dst_stride_vu, int width, int height);
void orc_CopyPlane(const uint8_t* src, int src_stride, uint8_t* dst, int dst_stride, int width, int height);
void orc_SwapUVPlane(const uint8_t* src_uv, int src_stride_uv, uint8_t* dst_vu, int dst_stride_vu, int width, int height);
void orc_MergeUVPlane(const uint8_t* src_u, int src_stride_u, const uint8_t* src_v, int src_stride_v,
                      uint8_t* dst_uv, int dst_stride_uv, int width, int height);
void orc_SplitUVPlane(const uint8_t* src_uv, int src_stride_uv, uint8_t* dst_u, int dst_stride_u,
                      uint8_t* dst_v, int dst_stride_v, int width, int height);
int orc_I420Copy(const uint8_t* src_y, int src_stride_y, const uint8_t* src_u, int src_stride_u,
                 const uint8_t* src_v, int src_stride_v, uint8_t* dst_y, int dst_stride_y, uint8_t* dst_u,
                 int dst_stride_u, uint8_t* dst_v, int dst_stride_v, int width, int height);
/* single-pixel helper exported for exhaustive tests (row_common.cc YuvPixel, kYuvI601Constants) */
void orc_YuvPixel(uint8_t y, uint8_t u, uint8_t v, uint8_t* r, uint8_t* g, uint8_t* b);
long orc_yuv_rows_selftest(void); /* vector rows vs scalar pixel on all 2^24 triples: differing values (0) */

/* ---- libyuv scalers (scale.cc / scale_argb.cc / scale_uv.cc / scale_common.cc) */
int orc_FixedDiv(int num, int div);
int orc_FixedDiv1(int num, int div);
int orc_ScaleFilterReduce(int src_width, int src_height, int dst_width, int dst_height, int filtering);
void orc_ScaleSlope(int src_width, int src_height, int dst_width, int dst_height, int filtering,
                    int* x, int* y, int* dx, int* dy);
int orc_ARGBScale(const uint8_t* src_argb, int src_stride_argb, int src_width, int src_height,
                  uint8_t* dst_argb, int dst_stride_argb, int dst_width, int dst_height, int filtering);
void orc_ScalePlane(const uint8_t* src, int src_stride, int src_width, int src_height, uint8_t* dst,
                    int dst_stride, int dst_width, int dst_height, int filtering);
int orc_I420Scale(const uint8_t* src_y, int src_stride_y, const uint8_t* src_u, int src_stride_u,
                  const uint8_t* src_v, int src_stride_v, int src_width, int src_height, uint8_t* dst_y,
                  int dst_stride_y, uint8_t* dst_u, int dst_stride_u, uint8_t* dst_v, int dst_stride_v,
                  int dst_width, int dst_height, int filtering);
int orc_UVScale(const uint8_t* src_uv, int src_stride_uv, int src_width, int src_height, uint8_t* dst_uv,
                int dst_stride_uv, int dst_width, int dst_height, int filtering);
int orc_NV12Scale(const uint8_t* src_y, int src_stride_y, const uint8_t* src_uv, int src_stride_uv,
                  int src_width, int src_height, uint8_t* dst_y, int dst_stride_y, uint8_t* dst_uv,
                  int dst_stride_uv, int dst_width, int dst_height, int filtering);

/* ---- libyuv rotate / mirror (rotate.cc / rotate_argb.cc / planar_functions.cc) */
/* mode is libyuv's CLOCKWISE RotationMode (0/90/180/270). */
int orc_RotatePlane(const uint8_t* src, int src_stride, uint8_t* dst, int dst_stride, int width, int height, int mode);
int orc_ARGBRotate(const uint8_t* src, int src_stride, uint8_t* dst, int dst_stride, int width, int height, int mode);
int orc_I420Rotate(const uint8_t* src_y, int src_stride_y, const uint8_t* src_u, int src_stride_u,
                   const uint8_t* src_v, int src_stride_v, uint8_t* dst_y, int dst_stride_y, uint8_t* dst_u,
                   int dst_stride_u, uint8_t* dst_v, int dst_stride_v, int width, int height, int mode);
int orc_NV12ToI420Rotate(const uint8_t* src_y, int src_stride_y, const uint8_t* src_uv, int src_stride_uv,
                         uint8_t* dst_y, int dst_stride_y, uint8_t* dst_u, int dst_stride_u, uint8_t* dst_v,
                         int dst_stride_v, int width, int height, int mode);
void orc_MirrorPlane(const uint8_t* src, int src_stride, uint8_t* dst, int dst_stride, int width, int height);
int orc_ARGBMirror(const uint8_t* src, int src_stride, uint8_t* dst, int dst_stride, int width, int height);
int orc_RGB24Mirror(const uint8_t* src, int src_stride, uint8_t* dst, int dst_stride, int width, int height);
int orc_NV12Mirror(const uint8_t* src_y, int src_stride_y, const uint8_t* src_uv, int src_stride_uv,
                   uint8_t* dst_y, int dst_stride_y, uint8_t* dst_uv, int dst_stride_uv, int width, int height);
int orc_I420Mirror(const uint8_t* src_y, int src_stride_y, const uint8_t* src_u, int src_stride_u,
                   const uint8_t* src_v, int src_stride_v, uint8_t* dst_y, int dst_stride_y, uint8_t* dst_u,
                   int dst_stride_u, uint8_t* dst_v, int dst_stride_v, int width, int height);

#ifdef __cplusplus
}
#endif
#endif /* BAND_ORACLE_YUVREF_H_ */
