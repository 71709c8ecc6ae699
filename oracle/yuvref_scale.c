/*
 * oracle/yuvref_scale.c -- TEST INFRASTRUCTURE ONLY (see yuvref.h header).
 *
 * Restates libyuv's C scalers: source/scale_common.cc (ScaleSlope,
 * ScaleFilterReduce, FixedDiv, *FilterCols_C, InterpolateRow_C,
 * ScalePlaneVertical), scale_argb.cc (ScaleARGB dispatch), scale.cc
 * (ScalePlane dispatch, Down2/Down34/Down38, Up2) and scale_uv.cc (ScaleUV).
 * Row-by-row like libyuv (materialised row buffers, incremental y clamp), on
 * purpose structurally different from the CUDA kernels' closed-form per-pixel
 * evaluation. x86 flavour of the platform-dependent BLENDER macro.
 *
 * Reference call sites: libyuv_image_operator.cc:1175,:1219 (ARGBScale),
 * :1373 (ScalePlane), :387 (I420Scale), :422 (NV12Scale); always
 * kFilterBilinear.
 */
#include <stdlib.h>
#include <string.h>

#include "yuvref.h"
#ifdef __AVX2__
#include <immintrin.h>
#endif

#define ABS(v) ((v) < 0 ? -(v) : (v))

int orc_FixedDiv(int num, int div) { return (int)(((int64_t)num << 16) / div); }
int orc_FixedDiv1(int num, int div) { return (int)((((int64_t)num << 16) - 0x00010001) / (div - 1)); }

/* scale_common.cc ScaleFilterReduce */
int orc_ScaleFilterReduce(int src_width, int src_height, int dst_width, int dst_height, int filtering) {
  if (src_width < 0) src_width = -src_width;
  if (src_height < 0) src_height = -src_height;
  if (filtering == ORC_FILTER_BOX) {
    if (dst_width * 2 >= src_width || dst_height * 2 >= src_height) filtering = ORC_FILTER_BILINEAR;
  }
  if (filtering == ORC_FILTER_BILINEAR) {
    if (src_height == 1) filtering = ORC_FILTER_LINEAR;
    if (dst_height == src_height || dst_height * 3 == src_height) filtering = ORC_FILTER_LINEAR;
    if (src_width == 1) filtering = ORC_FILTER_NONE;
  }
  if (filtering == ORC_FILTER_LINEAR) {
    if (src_width == 1) filtering = ORC_FILTER_NONE;
    if (dst_width == src_width || dst_width * 3 == src_width) filtering = ORC_FILTER_NONE;
  }
  return filtering;
}

#define CENTERSTART(dx, s) ((dx) < 0 ? -((-(dx) >> 1) + (s)) : (((dx) >> 1) + (s)))

/* scale_common.cc ScaleSlope */
void orc_ScaleSlope(int src_width, int src_height, int dst_width, int dst_height, int filtering, int* x,
                    int* y, int* dx, int* dy) {
  if (dst_width == 1 && src_width >= 32768) dst_width = src_width;
  if (dst_height == 1 && src_height >= 32768) dst_height = src_height;
  if (filtering == ORC_FILTER_BOX) {
    *dx = orc_FixedDiv(ABS(src_width), dst_width);
    *dy = orc_FixedDiv(src_height, dst_height);
    *x = 0;
    *y = 0;
  } else if (filtering == ORC_FILTER_BILINEAR) {
    if (dst_width <= ABS(src_width)) {
      *dx = orc_FixedDiv(ABS(src_width), dst_width);
      *x = CENTERSTART(*dx, -32768);
    } else if (dst_width > 1) {
      *dx = orc_FixedDiv1(ABS(src_width), dst_width);
      *x = 0;
    }
    if (dst_height <= src_height) {
      *dy = orc_FixedDiv(src_height, dst_height);
      *y = CENTERSTART(*dy, -32768);
    } else if (dst_height > 1) {
      *dy = orc_FixedDiv1(src_height, dst_height);
      *y = 0;
    }
  } else if (filtering == ORC_FILTER_LINEAR) {
    if (dst_width <= ABS(src_width)) {
      *dx = orc_FixedDiv(ABS(src_width), dst_width);
      *x = CENTERSTART(*dx, -32768);
    } else if (dst_width > 1) {
      *dx = orc_FixedDiv1(ABS(src_width), dst_width);
      *x = 0;
    }
    *dy = orc_FixedDiv(src_height, dst_height);
    *y = *dy >> 1;
  } else {
    *dx = orc_FixedDiv(ABS(src_width), dst_width);
    *dy = orc_FixedDiv(src_height, dst_height);
    *x = CENTERSTART(*dx, 0);
    *y = CENTERSTART(*dy, 0);
  }
  if (src_width < 0) {
    *x += (dst_width - 1) * *dx;
    *dx = -*dx;
  }
}

/* row_common.cc InterpolateRow_C (+HalfRow_C): 8-bit fraction, +128 rounding. */
static void interpolate_row(uint8_t* dst, const uint8_t* src, ptrdiff_t src_stride, int width, int yf) {
  int y1 = yf, y0 = 256 - yf;
  const uint8_t* src1 = src + src_stride;
  if (y1 == 0) {
    memcpy(dst, src, (size_t)width);
    return;
  }
  int x = 0;
#ifdef __AVX2__
  /* 32 bytes at a time in 16-bit lanes: a * y0 + b * y1 + 128 <= 65408 fits; avg_epu8 IS (a + b + 1) >> 1 */
  if (y1 == 128) {
    for (; x + 32 <= width; x += 32)
      _mm256_storeu_si256((__m256i*)(dst + x), _mm256_avg_epu8(_mm256_loadu_si256((const __m256i*)(src + x)),
                                                               _mm256_loadu_si256((const __m256i*)(src1 + x))));
  } else {
    const __m256i w0 = _mm256_set1_epi16((short)y0), w1 = _mm256_set1_epi16((short)y1), rnd = _mm256_set1_epi16(128);
    const __m256i z = _mm256_setzero_si256();
    for (; x + 32 <= width; x += 32) {
      const __m256i a = _mm256_loadu_si256((const __m256i*)(src + x)), b = _mm256_loadu_si256((const __m256i*)(src1 + x));
      const __m256i lo = _mm256_srli_epi16(_mm256_add_epi16(_mm256_add_epi16(_mm256_mullo_epi16(_mm256_unpacklo_epi8(a, z), w0),
                                                                             _mm256_mullo_epi16(_mm256_unpacklo_epi8(b, z), w1)), rnd), 8);
      const __m256i hi = _mm256_srli_epi16(_mm256_add_epi16(_mm256_add_epi16(_mm256_mullo_epi16(_mm256_unpackhi_epi8(a, z), w0),
                                                                             _mm256_mullo_epi16(_mm256_unpackhi_epi8(b, z), w1)), rnd), 8);
      _mm256_storeu_si256((__m256i*)(dst + x), _mm256_packus_epi16(lo, hi)); /* unpack / pack pair up inside each 128-bit lane */
    }
  }
#endif
  if (y1 == 128) {
    for (; x < width; ++x) dst[x] = (uint8_t)((src[x] + src1[x] + 1) >> 1);
    return;
  }
  for (; x < width; ++x) dst[x] = (uint8_t)((src[x] * y0 + src1[x] * y1 + 128) >> 8);
}

/* Column filters. blender 0: ScaleFilterCols_C (8-bit plane, x86 BLENDER:
 * a + (((f>>9)*(b-a) + 0x40) >> 7)); blender 1: ScaleARGBFilterCols_C /
 * ScaleUVFilterCols_C per channel ((a*(0x7f^f) + b*f) >> 7, no rounding). */
static void filter_cols(uint8_t* dst, const uint8_t* src, int dst_width, int x, int dx, int bpp, int blender) {
  for (int j = 0; j < dst_width; ++j) {
    int xi = x >> 16;
    const uint8_t* a = src + (ptrdiff_t)xi * bpp;
    const uint8_t* b = a + bpp;
    /* libyuv always loads src[xi + 1]; with a zero fraction that value is
     * multiplied by 0, so the oracle skips the load (it may lie one pixel past
     * the row at the right edge). */
    if (blender == 0) {
      int f = x & 0xffff;
      for (int c = 0; c < bpp; ++c) {
        int av = a[c], bv = (f >> 9) ? b[c] : av;
        dst[c] = (uint8_t)(av + ((((f >> 9) * (bv - av)) + 0x40) >> 7));
      }
    } else {
      int f = (x >> 9) & 0x7f;
      for (int c = 0; c < bpp; ++c) dst[c] = (uint8_t)((a[c] * (0x7f ^ f) + (f ? b[c] * f : 0)) >> 7);
    }
    dst += bpp;
    x += dx;
  }
}
/* ScaleCols_C / ScaleARGBCols_C / ScaleUVCols_C: point sample. */
static void point_cols(uint8_t* dst, const uint8_t* src, int dst_width, int x, int dx, int bpp) {
  for (int j = 0; j < dst_width; ++j) {
    memcpy(dst, src + (ptrdiff_t)(x >> 16) * bpp, (size_t)bpp);
    dst += bpp;
    x += dx;
  }
}

/* scale_common.cc ScalePlaneVertical (note its own max_y = ((h-1)<<16) - 1). */
static void scale_vertical(int src_height, int dst_width, int dst_height, int src_stride, int dst_stride,
                           const uint8_t* src, uint8_t* dst, int x, int y, int dy, int bpp, int filtering) {
  int dst_width_bytes = dst_width * bpp;
  const int max_y = (src_height > 1) ? ((src_height - 1) << 16) - 1 : 0;
  src += (x >> 16) * bpp;
  for (int j = 0; j < dst_height; ++j) {
    if (y > max_y) y = max_y;
    int yi = y >> 16;
    int yf = filtering ? ((y >> 8) & 255) : 0;
    interpolate_row(dst, src + (ptrdiff_t)yi * src_stride, src_stride, dst_width_bytes, yf);
    dst += dst_stride;
    y += dy;
  }
}

/* Scale*BilinearDown (scale.cc / scale_argb.cc / scale_uv.cc). */
static void scale_bilinear_down(int src_width, int src_height, int dst_width, int dst_height, int src_stride,
                                int dst_stride, const uint8_t* src, uint8_t* dst, int x, int dx, int y, int dy,
                                int filtering, int bpp, int blender) {
  uint8_t* row = (uint8_t*)malloc((size_t)src_width * bpp + 64);
  const int max_y = (src_height - 1) << 16;
  if (y > max_y) y = max_y;
  for (int j = 0; j < dst_height; ++j) {
    int yi = y >> 16;
    const uint8_t* s = src + (ptrdiff_t)yi * src_stride;
    if (filtering == ORC_FILTER_LINEAR) {
      filter_cols(dst, s, dst_width, x, dx, bpp, blender);
    } else {
      int yf = (y >> 8) & 255;
      interpolate_row(row, s, src_stride, src_width * bpp, yf);
      filter_cols(dst, row, dst_width, x, dx, bpp, blender);
    }
    dst += dst_stride;
    y += dy;
    if (y > max_y) y = max_y;
  }
  free(row);
}

/* Scale*BilinearUp: columns first into a 2-row ping-pong, then InterpolateRow. */
static void scale_bilinear_up(int src_width, int src_height, int dst_width, int dst_height, int src_stride,
                              int dst_stride, const uint8_t* src_base, uint8_t* dst, int x, int dx, int y,
                              int dy, int filtering, int bpp, int blender) {
  (void)src_width;
  const int max_y = (src_height - 1) << 16;
  if (y > max_y) y = max_y;
  int yi = y >> 16;
  const uint8_t* src = src_base + (ptrdiff_t)yi * src_stride;
  const int row_size = (dst_width * bpp + 31) & ~31;
  uint8_t* row = (uint8_t*)malloc((size_t)row_size * 2);
  uint8_t* rowptr = row;
  int rowstride = row_size;
  int lasty = yi;
  if (filtering) filter_cols(rowptr, src, dst_width, x, dx, bpp, blender);
  else point_cols(rowptr, src, dst_width, x, dx, bpp);
  if (src_height > 1) src += src_stride;
  if (filtering) filter_cols(rowptr + rowstride, src, dst_width, x, dx, bpp, blender);
  else point_cols(rowptr + rowstride, src, dst_width, x, dx, bpp);
  if (src_height > 2) src += src_stride;
  for (int j = 0; j < dst_height; ++j) {
    yi = y >> 16;
    if (yi != lasty) {
      if (y > max_y) {
        y = max_y;
        yi = y >> 16;
        src = src_base + (ptrdiff_t)yi * src_stride;
      }
      if (yi != lasty) {
        if (filtering) filter_cols(rowptr, src, dst_width, x, dx, bpp, blender);
        else point_cols(rowptr, src, dst_width, x, dx, bpp);
        rowptr += rowstride;
        rowstride = -rowstride;
        lasty = yi;
        if ((y + 65536) < max_y) src += src_stride;
      }
    }
    if (filtering == ORC_FILTER_LINEAR) {
      interpolate_row(dst, rowptr, 0, dst_width * bpp, 0);
    } else {
      int yf = (y >> 8) & 255;
      interpolate_row(dst, rowptr, rowstride, dst_width * bpp, yf);
    }
    dst += dst_stride;
    y += dy;
  }
  free(row);
}

/* Scale*Simple: point sampling. */
static void scale_simple(int dst_width, int dst_height, int src_stride, int dst_stride, const uint8_t* src,
                         uint8_t* dst, int x, int dx, int y, int dy, int bpp) {
  for (int j = 0; j < dst_height; ++j) {
    point_cols(dst, src + (ptrdiff_t)(y >> 16) * src_stride, dst_width, x, dx, bpp);
    dst += dst_stride;
    y += dy;
  }
}

/* ScaleARGBDown2 / ScaleARGBDownEven / ScaleUVDown2 / ScaleUVDownEven. For a
 * bilinear request these always run the *Box row (2x2 average at
 * (x>>16, y>>16), stepping dx>>16 / dy>>16 source pixels). */
static void scale_down_even_box(int dst_width, int dst_height, int src_stride, int dst_stride,
                                const uint8_t* src, uint8_t* dst, int x, int dx, int y, int dy, int bpp) {
  int col_step = dx >> 16;
  ptrdiff_t row_stride = (ptrdiff_t)(dy >> 16) * src_stride;
  src += (ptrdiff_t)(y >> 16) * src_stride + (ptrdiff_t)(x >> 16) * bpp;
  for (int j = 0; j < dst_height; ++j) {
    const uint8_t* s = src;
    const uint8_t* t = src + src_stride;
    uint8_t* d = dst;
    for (int i = 0; i < dst_width; ++i) {
      for (int c = 0; c < bpp; ++c) d[c] = (uint8_t)((s[c] + s[bpp + c] + t[c] + t[bpp + c] + 2) >> 2);
      s += (ptrdiff_t)col_step * bpp;
      t += (ptrdiff_t)col_step * bpp;
      d += bpp;
    }
    src += row_stride;
    dst += dst_stride;
  }
}

/* Shared ScaleARGB / ScaleUV dispatcher (scale_argb.cc ScaleARGB, scale_uv.cc
 * ScaleUV; no clipping). allow_up2: ScaleUV's 2x special cases. */
static void scale_uv_up2(int src_width, int src_height, int dst_width, int dst_height, int src_stride,
                         int dst_stride, const uint8_t* src, uint8_t* dst, int bilinear, int bpp);

static void scale_packed(const uint8_t* src, int src_stride, int src_width, int src_height, uint8_t* dst,
                         int dst_stride, int dst_width, int dst_height, int filtering, int bpp) {
  int x = 0, y = 0, dx = 0, dy = 0;
  filtering = orc_ScaleFilterReduce(src_width, src_height, dst_width, dst_height, filtering);
  orc_ScaleSlope(src_width, src_height, dst_width, dst_height, filtering, &x, &y, &dx, &dy);
  if (((dx | dy) & 0xffff) == 0) {
    if (!dx || !dy) {
      filtering = ORC_FILTER_NONE;
    } else {
      if (!(dx & 0x10000) && !(dy & 0x10000)) {
        /* Down2 (dx == 2.0) and DownEven share the same 2x2 box arithmetic
         * for a bilinear request; Down4Box needs kFilterBox (never passed). */
        scale_down_even_box(dst_width, dst_height, src_stride, dst_stride, src, dst, x, dx, y, dy, bpp);
        return;
      }
      if ((dx & 0x10000) && (dy & 0x10000)) {
        filtering = ORC_FILTER_NONE;
        if (dx == 0x10000 && dy == 0x10000) {
          orc_CopyPlane(src + (ptrdiff_t)(y >> 16) * src_stride + (ptrdiff_t)(x >> 16) * bpp, src_stride, dst,
                        dst_stride, dst_width * bpp, dst_height);
          return;
        }
      }
    }
  }
  if (dx == 0x10000 && (x & 0xffff) == 0) {
    scale_vertical(src_height, dst_width, dst_height, src_stride, dst_stride, src, dst, x, y, dy, bpp,
                   filtering);
    return;
  }
  if (bpp == 2) { /* scale_uv.cc only: exact-2x up-samplers */
    if ((dst_width + 1) / 2 == src_width && filtering == ORC_FILTER_LINEAR) {
      scale_uv_up2(src_width, src_height, dst_width, dst_height, src_stride, dst_stride, src, dst, 0, bpp);
      return;
    }
    if ((dst_height + 1) / 2 == src_height && (dst_width + 1) / 2 == src_width &&
        (filtering == ORC_FILTER_BILINEAR || filtering == ORC_FILTER_BOX)) {
      scale_uv_up2(src_width, src_height, dst_width, dst_height, src_stride, dst_stride, src, dst, 1, bpp);
      return;
    }
  }
  if (filtering && dy < 65536) {
    scale_bilinear_up(src_width, src_height, dst_width, dst_height, src_stride, dst_stride, src, dst, x, dx, y,
                      dy, filtering, bpp, 1);
    return;
  }
  if (filtering) {
    scale_bilinear_down(src_width, src_height, dst_width, dst_height, src_stride, dst_stride, src, dst, x, dx,
                        y, dy, filtering, bpp, 1);
    return;
  }
  scale_simple(dst_width, dst_height, src_stride, dst_stride, src, dst, x, dx, y, dy, bpp);
}

int orc_ARGBScale(const uint8_t* src_argb, int src_stride_argb, int src_width, int src_height,
                  uint8_t* dst_argb, int dst_stride_argb, int dst_width, int dst_height, int filtering) {
  if (!src_argb || src_width == 0 || src_height == 0 || src_width > 32768 || src_height > 32768 ||
      !dst_argb || dst_width <= 0 || dst_height <= 0)
    return -1;
  scale_packed(src_argb, src_stride_argb, src_width, src_height, dst_argb, dst_stride_argb, dst_width,
               dst_height, filtering, 4);
  return 0;
}
int orc_UVScale(const uint8_t* src_uv, int src_stride_uv, int src_width, int src_height, uint8_t* dst_uv,
                int dst_stride_uv, int dst_width, int dst_height, int filtering) {
  if (!src_uv || src_width <= 0 || src_height == 0 || src_width > 32768 || src_height > 32768 || !dst_uv ||
      dst_width <= 0 || dst_height <= 0)
    return -1;
  scale_packed(src_uv, src_stride_uv, src_width, src_height, dst_uv, dst_stride_uv, dst_width, dst_height,
               filtering, 2);
  return 0;
}

/* ---- exact-2x up-samplers (scale.cc ScalePlaneUp2_Linear/_Bilinear,
 * scale_uv.cc ScaleUVLinearUp2/BilinearUp2; rows ScaleRowUp2_Linear_C /
 * ScaleRowUp2_Bilinear_C through the *_Any_C edge wrappers). libyuv >= Feb 2021;
 * VERSION-DEPENDENT, unpinned. bpp = 1 (plane) or 2 (UV). --------------------- */
static void row_up2_linear_any(const uint8_t* src, uint8_t* dst, int dst_width, int bpp) {
  int work_width = (dst_width - 1) & ~1;
  for (int c = 0; c < bpp; ++c) dst[c] = src[c];
  /* C(src, dst + 1, work_width): src_width = work_width >> 1 */
  for (int x = 0; x < (work_width >> 1); ++x) {
    for (int c = 0; c < bpp; ++c) {
      int s0 = src[x * bpp + c], s1 = src[(x + 1) * bpp + c];
      dst[(1 + 2 * x) * bpp + c] = (uint8_t)((s0 * 3 + s1 * 1 + 2) >> 2);
      dst[(2 + 2 * x) * bpp + c] = (uint8_t)((s0 * 1 + s1 * 3 + 2) >> 2);
    }
  }
  for (int c = 0; c < bpp; ++c) dst[(dst_width - 1) * bpp + c] = src[((dst_width - 1) / 2) * bpp + c];
}
static void row_up2_bilinear_any(const uint8_t* sa, ptrdiff_t src_stride, uint8_t* da, ptrdiff_t dst_stride,
                                 int dst_width, int bpp) {
  const uint8_t* sb = sa + src_stride;
  uint8_t* db = da + dst_stride;
  int work_width = (dst_width - 1) & ~1;
  int last = (dst_width - 1) / 2;
  /* NOTE: when dst_stride == 0 (first/last output row) db aliases da and the
   * second store wins, exactly as in libyuv. */
  for (int c = 0; c < bpp; ++c) {
    uint8_t v0 = (uint8_t)((3 * sa[c] + sb[c] + 2) >> 2);
    uint8_t v1 = (uint8_t)((sa[c] + 3 * sb[c] + 2) >> 2);
    da[c] = v0;
    db[c] = v1;
  }
  for (int x = 0; x < (work_width >> 1); ++x) {
    for (int c = 0; c < bpp; ++c) {
      int s0 = sa[x * bpp + c], s1 = sa[(x + 1) * bpp + c];
      int t0 = sb[x * bpp + c], t1 = sb[(x + 1) * bpp + c];
      uint8_t d0 = (uint8_t)((s0 * 9 + s1 * 3 + t0 * 3 + t1 * 1 + 8) >> 4);
      uint8_t d1 = (uint8_t)((s0 * 3 + s1 * 9 + t0 * 1 + t1 * 3 + 8) >> 4);
      uint8_t e0 = (uint8_t)((s0 * 3 + s1 * 1 + t0 * 9 + t1 * 3 + 8) >> 4);
      uint8_t e1 = (uint8_t)((s0 * 1 + s1 * 3 + t0 * 3 + t1 * 9 + 8) >> 4);
      da[(1 + 2 * x) * bpp + c] = d0;
      da[(2 + 2 * x) * bpp + c] = d1;
      db[(1 + 2 * x) * bpp + c] = e0;
      db[(2 + 2 * x) * bpp + c] = e1;
    }
  }
  for (int c = 0; c < bpp; ++c) {
    uint8_t v0 = (uint8_t)((3 * sa[last * bpp + c] + sb[last * bpp + c] + 2) >> 2);
    uint8_t v1 = (uint8_t)((sa[last * bpp + c] + 3 * sb[last * bpp + c] + 2) >> 2);
    da[(dst_width - 1) * bpp + c] = v0;
    db[(dst_width - 1) * bpp + c] = v1;
  }
}
static void scale_uv_up2(int src_width, int src_height, int dst_width, int dst_height, int src_stride,
                         int dst_stride, const uint8_t* src, uint8_t* dst, int bilinear, int bpp) {
  (void)src_width;
  if (!bilinear) {
    if (dst_height == 1) {
      row_up2_linear_any(src + (ptrdiff_t)((src_height - 1) / 2) * src_stride, dst, dst_width, bpp);
    } else {
      int dy = orc_FixedDiv(src_height - 1, dst_height - 1);
      int y = (1 << 15) - 1;
      for (int i = 0; i < dst_height; ++i) {
        row_up2_linear_any(src + (ptrdiff_t)(y >> 16) * src_stride, dst, dst_width, bpp);
        dst += dst_stride;
        y += dy;
      }
    }
    return;
  }
  row_up2_bilinear_any(src, 0, dst, 0, dst_width, bpp);
  dst += dst_stride;
  for (int x = 0; x < src_height - 1; ++x) {
    row_up2_bilinear_any(src, src_stride, dst, dst_stride, dst_width, bpp);
    src += src_stride;
    dst += 2 * (ptrdiff_t)dst_stride;
  }
  if (!(dst_height & 1)) row_up2_bilinear_any(src, 0, dst, 0, dst_width, bpp);
}

/* ---- scale.cc fixed-ratio plane reducers -------------------------------- */
/* ScalePlaneDown2 with ScaleRowDown2Box_C (bilinear request => Box). */
static void plane_down2_box(int dst_width, int dst_height, int src_stride, int dst_stride, const uint8_t* src,
                            uint8_t* dst) {
  for (int y = 0; y < dst_height; ++y) {
    const uint8_t* s = src;
    const uint8_t* t = src + src_stride;
    for (int x = 0; x < dst_width; ++x) {
      dst[x] = (uint8_t)((s[0] + s[1] + t[0] + t[1] + 2) >> 2);
      s += 2;
      t += 2;
    }
    src += 2 * (ptrdiff_t)src_stride;
    dst += dst_stride;
  }
}
/* ScaleRowDown34_0_Box_C (rows 3:1) / ScaleRowDown34_1_Box_C (rows 1:1). */
static void row_down34_box(const uint8_t* s, ptrdiff_t stride, uint8_t* d, int dst_width, int one_one) {
  const uint8_t* t = s + stride;
  for (int x = 0; x < dst_width; x += 3) {
    uint8_t a0 = (uint8_t)((s[0] * 3 + s[1] * 1 + 2) >> 2);
    uint8_t a1 = (uint8_t)((s[1] * 1 + s[2] * 1 + 1) >> 1);
    uint8_t a2 = (uint8_t)((s[2] * 1 + s[3] * 3 + 2) >> 2);
    uint8_t b0 = (uint8_t)((t[0] * 3 + t[1] * 1 + 2) >> 2);
    uint8_t b1 = (uint8_t)((t[1] * 1 + t[2] * 1 + 1) >> 1);
    uint8_t b2 = (uint8_t)((t[2] * 1 + t[3] * 3 + 2) >> 2);
    if (one_one) {
      d[0] = (uint8_t)((a0 + b0 + 1) >> 1);
      d[1] = (uint8_t)((a1 + b1 + 1) >> 1);
      d[2] = (uint8_t)((a2 + b2 + 1) >> 1);
    } else {
      d[0] = (uint8_t)((a0 * 3 + b0 + 2) >> 2);
      d[1] = (uint8_t)((a1 * 3 + b1 + 2) >> 2);
      d[2] = (uint8_t)((a2 * 3 + b2 + 2) >> 2);
    }
    d += 3;
    s += 4;
    t += 4;
  }
}
static void plane_down34(int dst_width, int dst_height, int src_stride, int dst_stride, const uint8_t* src,
                         uint8_t* dst) {
  const int filter_stride = src_stride; /* bilinear => not linear */
  int y;
  for (y = 0; y < dst_height - 2; y += 3) {
    row_down34_box(src, filter_stride, dst, dst_width, 0);
    src += src_stride;
    dst += dst_stride;
    row_down34_box(src, filter_stride, dst, dst_width, 1);
    src += src_stride;
    dst += dst_stride;
    row_down34_box(src + src_stride, -filter_stride, dst, dst_width, 0);
    src += 2 * (ptrdiff_t)src_stride;
    dst += dst_stride;
  }
  if ((dst_height % 3) == 2) {
    row_down34_box(src, filter_stride, dst, dst_width, 0);
    src += src_stride;
    dst += dst_stride;
    row_down34_box(src, 0, dst, dst_width, 1);
  } else if ((dst_height % 3) == 1) {
    row_down34_box(src, 0, dst, dst_width, 0);
  }
}
/* ScaleRowDown38_3_Box_C (8x3 -> 3x1) / ScaleRowDown38_2_Box_C (8x2 -> 3x1). */
static void row_down38_box(const uint8_t* s, ptrdiff_t stride, uint8_t* d, int dst_width, int rows3) {
  for (int i = 0; i < dst_width; i += 3) {
    if (rows3) {
      d[0] = (uint8_t)((s[0] + s[1] + s[2] + s[stride + 0] + s[stride + 1] + s[stride + 2] + s[stride * 2 + 0] +
                        s[stride * 2 + 1] + s[stride * 2 + 2]) * (65536 / 9) >> 16);
      d[1] = (uint8_t)((s[3] + s[4] + s[5] + s[stride + 3] + s[stride + 4] + s[stride + 5] + s[stride * 2 + 3] +
                        s[stride * 2 + 4] + s[stride * 2 + 5]) * (65536 / 9) >> 16);
      d[2] = (uint8_t)((s[6] + s[7] + s[stride + 6] + s[stride + 7] + s[stride * 2 + 6] + s[stride * 2 + 7]) *
                       (65536 / 6) >> 16);
    } else {
      d[0] = (uint8_t)((s[0] + s[1] + s[2] + s[stride + 0] + s[stride + 1] + s[stride + 2]) * (65536 / 6) >> 16);
      d[1] = (uint8_t)((s[3] + s[4] + s[5] + s[stride + 3] + s[stride + 4] + s[stride + 5]) * (65536 / 6) >> 16);
      d[2] = (uint8_t)((s[6] + s[7] + s[stride + 6] + s[stride + 7]) * (65536 / 4) >> 16);
    }
    s += 8;
    d += 3;
  }
}
static void plane_down38(int dst_width, int dst_height, int src_stride, int dst_stride, const uint8_t* src,
                         uint8_t* dst) {
  const int filter_stride = src_stride;
  int y;
  for (y = 0; y < dst_height - 2; y += 3) {
    row_down38_box(src, filter_stride, dst, dst_width, 1);
    src += 3 * (ptrdiff_t)src_stride;
    dst += dst_stride;
    row_down38_box(src, filter_stride, dst, dst_width, 1);
    src += 3 * (ptrdiff_t)src_stride;
    dst += dst_stride;
    row_down38_box(src, filter_stride, dst, dst_width, 0);
    src += 2 * (ptrdiff_t)src_stride;
    dst += dst_stride;
  }
  if ((dst_height % 3) == 2) {
    row_down38_box(src, filter_stride, dst, dst_width, 1);
    src += 3 * (ptrdiff_t)src_stride;
    dst += dst_stride;
    row_down38_box(src, 0, dst, dst_width, 1);
  } else if ((dst_height % 3) == 1) {
    row_down38_box(src, 0, dst, dst_width, 1);
  }
}

/* scale.cc ScalePlane (8-bit). */
void orc_ScalePlane(const uint8_t* src, int src_stride, int src_width, int src_height, uint8_t* dst,
                    int dst_stride, int dst_width, int dst_height, int filtering) {
  filtering = orc_ScaleFilterReduce(src_width, src_height, dst_width, dst_height, filtering);
  if (dst_width == src_width && dst_height == src_height) {
    orc_CopyPlane(src, src_stride, dst, dst_stride, dst_width, dst_height);
    return;
  }
  if (dst_width == src_width && filtering != ORC_FILTER_BOX) {
    int dy = 0, y = 0;
    if (dst_height <= src_height) {
      dy = orc_FixedDiv(src_height, dst_height);
      y = CENTERSTART(dy, -32768);
    } else if (src_height > 1 && dst_height > 1) {
      dy = orc_FixedDiv1(src_height, dst_height);
    }
    scale_vertical(src_height, dst_width, dst_height, src_stride, dst_stride, src, dst, 0, y, dy, 1, filtering);
    return;
  }
  if (dst_width <= ABS(src_width) && dst_height <= src_height) {
    if (4 * dst_width == 3 * src_width && 4 * dst_height == 3 * src_height && filtering == ORC_FILTER_BILINEAR) {
      plane_down34(dst_width, dst_height, src_stride, dst_stride, src, dst);
      return;
    }
    if (2 * dst_width == src_width && 2 * dst_height == src_height && filtering == ORC_FILTER_BILINEAR) {
      plane_down2_box(dst_width, dst_height, src_stride, dst_stride, src, dst);
      return;
    }
    if (8 * dst_width == 3 * src_width && 8 * dst_height == 3 * src_height && filtering == ORC_FILTER_BILINEAR) {
      plane_down38(dst_width, dst_height, src_stride, dst_stride, src, dst);
      return;
    }
    /* 1/4: only for Box/None, never reached from a bilinear request. */
  }
  if ((dst_width + 1) / 2 == src_width && filtering == ORC_FILTER_LINEAR) {
    scale_uv_up2(src_width, src_height, dst_width, dst_height, src_stride, dst_stride, src, dst, 0, 1);
    return;
  }
  if ((dst_height + 1) / 2 == src_height && (dst_width + 1) / 2 == src_width &&
      (filtering == ORC_FILTER_BILINEAR || filtering == ORC_FILTER_BOX)) {
    scale_uv_up2(src_width, src_height, dst_width, dst_height, src_stride, dst_stride, src, dst, 1, 1);
    return;
  }
  {
    int x = 0, y = 0, dx = 0, dy = 0;
    orc_ScaleSlope(src_width, src_height, dst_width, dst_height, filtering, &x, &y, &dx, &dy);
    if (filtering && dst_height > src_height) {
      scale_bilinear_up(src_width, src_height, dst_width, dst_height, src_stride, dst_stride, src, dst, x, dx, y,
                        dy, filtering, 1, 0);
      return;
    }
    if (filtering) {
      scale_bilinear_down(src_width, src_height, dst_width, dst_height, src_stride, dst_stride, src, dst, x, dx,
                          y, dy, filtering, 1, 0);
      return;
    }
    scale_simple(dst_width, dst_height, src_stride, dst_stride, src, dst, x, dx, y, dy, 1);
  }
}

/* scale.cc I420Scale / scale_uv.cc NV12Scale: chroma at SUBSAMPLE(v,1,1). */
int orc_I420Scale(const uint8_t* src_y, int src_stride_y, const uint8_t* src_u, int src_stride_u,
                  const uint8_t* src_v, int src_stride_v, int src_width, int src_height, uint8_t* dst_y,
                  int dst_stride_y, uint8_t* dst_u, int dst_stride_u, uint8_t* dst_v, int dst_stride_v,
                  int dst_width, int dst_height, int filtering) {
  int shw = (src_width + 1) >> 1, shh = (src_height + 1) >> 1;
  int dhw = (dst_width + 1) >> 1, dhh = (dst_height + 1) >> 1;
  if (!src_y || !src_u || !src_v || src_width <= 0 || src_height == 0 || src_width > 32768 ||
      src_height > 32768 || !dst_y || !dst_u || !dst_v || dst_width <= 0 || dst_height <= 0)
    return -1;
  orc_ScalePlane(src_y, src_stride_y, src_width, src_height, dst_y, dst_stride_y, dst_width, dst_height, filtering);
  orc_ScalePlane(src_u, src_stride_u, shw, shh, dst_u, dst_stride_u, dhw, dhh, filtering);
  orc_ScalePlane(src_v, src_stride_v, shw, shh, dst_v, dst_stride_v, dhw, dhh, filtering);
  return 0;
}
int orc_NV12Scale(const uint8_t* src_y, int src_stride_y, const uint8_t* src_uv, int src_stride_uv,
                  int src_width, int src_height, uint8_t* dst_y, int dst_stride_y, uint8_t* dst_uv,
                  int dst_stride_uv, int dst_width, int dst_height, int filtering) {
  int shw = (src_width + 1) >> 1, shh = (src_height + 1) >> 1;
  int dhw = (dst_width + 1) >> 1, dhh = (dst_height + 1) >> 1;
  if (!src_y || !src_uv || src_width <= 0 || src_height == 0 || src_width > 32768 || src_height > 32768 ||
      !dst_y || !dst_uv || dst_width <= 0 || dst_height <= 0)
    return -1;
  orc_ScalePlane(src_y, src_stride_y, src_width, src_height, dst_y, dst_stride_y, dst_width, dst_height, filtering);
  orc_UVScale(src_uv, src_stride_uv, shw, shh, dst_uv, dst_stride_uv, dhw, dhh, filtering);
  return 0;
}
