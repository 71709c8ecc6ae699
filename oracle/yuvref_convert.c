/*
 * oracle/yuvref_convert.c -- TEST INFRASTRUCTURE ONLY (see yuvref.h header).
 *
 * Restates libyuv's C colour-conversion paths (source/row_common.cc,
 * convert_argb.cc, convert.cc, convert_from.cc, planar_functions.cc), x86
 * flavour, as called from band/buffer/libyuv_image_operator.cc.
 */
#include <stdlib.h>
#include <string.h>

#include "yuvref.h"

/* ---------------------------------------------------------------------------
 * YuvPixel: libyuv row_common.cc `YuvPixel` with kYuvI601Constants (x86):
 *   YG=18997 YB=-1160 UB=128 UG=25 VG=52 VR=102, 6-bit fixed point.
 * NV12ToRAW/NV21ToRAW/I420ToRAW call the RGB24 row with the U/V-swapped
 * constant table (kYvuI601Constants); the net effect is R,G,B byte order with
 * the same arithmetic, restated here directly.
 * Reached from libyuv_image_operator.cc:102 (NV12ToRAW), :190 (NV21ToRAW),
 * :278 (I420ToRAW), :114/:202/:291 (…ToABGR).
 * ------------------------------------------------------------------------- */
static inline uint8_t clamp255(int32_t v) {
  v = v < 0 ? 0 : v;
  return (uint8_t)(v > 255 ? 255 : v);
}

static inline void yuv_pixel(uint8_t y, uint8_t u, uint8_t v, uint8_t* r, uint8_t* g, uint8_t* b) {
  uint32_t y32 = (uint32_t)y * 0x0101u;
  int32_t y1 = (int32_t)((uint32_t)(y32 * 18997u) >> 16) + (-1160);
  int32_t ui = (int32_t)u - 128; /* int8 wrap of (int8)u - 0x80 == u - 128 */
  int32_t vi = (int32_t)v - 128;
  int32_t b16 = y1 + ui * 128;
  int32_t g16 = y1 - (ui * 25 + vi * 52);
  int32_t r16 = y1 + vi * 102;
  *b = clamp255(b16 >> 6);
  *g = clamp255(g16 >> 6);
  *r = clamp255(r16 >> 6);
}
/* the exported form (tests call it for the 2^24-triple table); the walkers below use the inline one: an exported function
 * of a -fPIC library is neither inlined nor vectorised (it may be interposed), and the timed CPU baseline spent most of a
 * frame in that call */
void orc_YuvPixel(uint8_t y, uint8_t u, uint8_t v, uint8_t* r, uint8_t* g, uint8_t* b) { yuv_pixel(y, u, v, r, g, b); }

#ifdef __AVX2__
#include <immintrin.h>
/* Sixteen pixels of the same arithmetic in 16-bit lanes (what libyuv's own AVX2 rows do; used by the TIMED CPU baseline so that
 * it is not a scalar straw man -- bench.py --impl reference). Exactness against yuv_pixel():
 *   y1  = mulhi_epu16(y * 257, 18997) - 1160               : the unsigned 16 x 16 -> high 16 product IS (y32 * 18997) >> 16;
 *   g16, r16 stay inside int16 for every input; b16 = y1 + 128 (u - 128) can reach 34092 -- the saturating add clips it to
 *   32767, and both 32767 >> 6 and 34092 >> 6 exceed 255, so the final clamp gives 255 either way;
 *   >> 6 is an arithmetic shift in both forms, packus saturates to 0..255 like clamp255().
 * orc_yuv_rows_selftest() below compares the two forms on all 2^24 (y, u, v) triples (tests/test_oracle_golden.py).
 * uv16: 8 chroma pairs as bytes u0 v0 u1 v1 ... (one pair per two pixels). Writes r, g, b for 16 pixels. */
static inline void yuv16_avx2(const uint8_t* py, __m128i uv16, __m128i* r8, __m128i* g8, __m128i* b8) {
  const __m256i y = _mm256_cvtepu8_epi16(_mm_loadu_si128((const __m128i*)py));
  const __m256i y1 = _mm256_sub_epi16(_mm256_mulhi_epu16(_mm256_mullo_epi16(y, _mm256_set1_epi16(257)),
                                                          _mm256_set1_epi16((short)18997)),
                                      _mm256_set1_epi16(1160));
  const __m256i uv = _mm256_broadcastsi128_si256(uv16);
  /* lane 0: pairs 0..3 for pixels 0..7, lane 1: pairs 4..7 for pixels 8..15; every pair feeds two 16-bit lanes */
  const __m256i mu = _mm256_setr_epi8(0, -1, 0, -1, 2, -1, 2, -1, 4, -1, 4, -1, 6, -1, 6, -1,
                                      8, -1, 8, -1, 10, -1, 10, -1, 12, -1, 12, -1, 14, -1, 14, -1);
  const __m256i mv = _mm256_setr_epi8(1, -1, 1, -1, 3, -1, 3, -1, 5, -1, 5, -1, 7, -1, 7, -1,
                                      9, -1, 9, -1, 11, -1, 11, -1, 13, -1, 13, -1, 15, -1, 15, -1);
  const __m256i c128 = _mm256_set1_epi16(128);
  const __m256i ui = _mm256_sub_epi16(_mm256_shuffle_epi8(uv, mu), c128);
  const __m256i vi = _mm256_sub_epi16(_mm256_shuffle_epi8(uv, mv), c128);
  const __m256i b16 = _mm256_adds_epi16(y1, _mm256_slli_epi16(ui, 7));
  const __m256i g16 = _mm256_sub_epi16(y1, _mm256_add_epi16(_mm256_mullo_epi16(ui, _mm256_set1_epi16(25)),
                                                            _mm256_mullo_epi16(vi, _mm256_set1_epi16(52))));
  const __m256i r16 = _mm256_add_epi16(y1, _mm256_mullo_epi16(vi, _mm256_set1_epi16(102)));
  const __m256i r = _mm256_srai_epi16(r16, 6), g = _mm256_srai_epi16(g16, 6), b = _mm256_srai_epi16(b16, 6);
  /* packus works inside 128-bit lanes: (lo lane of r | lo lane of g ...) -> bring pixels 0..15 of one channel together */
  const __m256i rg = _mm256_permute4x64_epi64(_mm256_packus_epi16(r, g), 0xD8); /* r0..15 | g0..15 */
  const __m256i bb = _mm256_permute4x64_epi64(_mm256_packus_epi16(b, b), 0xD8); /* b0..15 | b0..15 */
  *r8 = _mm256_castsi256_si128(rg);
  *g8 = _mm256_extracti128_si256(rg, 1);
  *b8 = _mm256_castsi256_si128(bb);
}
/* 16 pixels r, g, b -> 48 interleaved bytes */
static inline void store_rgb48(uint8_t* d, __m128i r, __m128i g, __m128i b) {
  const __m128i r0 = _mm_setr_epi8(0, -1, -1, 1, -1, -1, 2, -1, -1, 3, -1, -1, 4, -1, -1, 5);
  const __m128i g0 = _mm_setr_epi8(-1, 0, -1, -1, 1, -1, -1, 2, -1, -1, 3, -1, -1, 4, -1, -1);
  const __m128i b0 = _mm_setr_epi8(-1, -1, 0, -1, -1, 1, -1, -1, 2, -1, -1, 3, -1, -1, 4, -1);
  const __m128i r1 = _mm_setr_epi8(-1, -1, 6, -1, -1, 7, -1, -1, 8, -1, -1, 9, -1, -1, 10, -1);
  const __m128i g1 = _mm_setr_epi8(5, -1, -1, 6, -1, -1, 7, -1, -1, 8, -1, -1, 9, -1, -1, 10);
  const __m128i b1 = _mm_setr_epi8(-1, 5, -1, -1, 6, -1, -1, 7, -1, -1, 8, -1, -1, 9, -1, -1);
  const __m128i r2 = _mm_setr_epi8(-1, 11, -1, -1, 12, -1, -1, 13, -1, -1, 14, -1, -1, 15, -1, -1);
  const __m128i g2 = _mm_setr_epi8(-1, -1, 11, -1, -1, 12, -1, -1, 13, -1, -1, 14, -1, -1, 15, -1);
  const __m128i b2 = _mm_setr_epi8(10, -1, -1, 11, -1, -1, 12, -1, -1, 13, -1, -1, 14, -1, -1, 15);
  _mm_storeu_si128((__m128i*)d, _mm_or_si128(_mm_or_si128(_mm_shuffle_epi8(r, r0), _mm_shuffle_epi8(g, g0)), _mm_shuffle_epi8(b, b0)));
  _mm_storeu_si128((__m128i*)(d + 16), _mm_or_si128(_mm_or_si128(_mm_shuffle_epi8(r, r1), _mm_shuffle_epi8(g, g1)), _mm_shuffle_epi8(b, b1)));
  _mm_storeu_si128((__m128i*)(d + 32), _mm_or_si128(_mm_or_si128(_mm_shuffle_epi8(r, r2), _mm_shuffle_epi8(g, g2)), _mm_shuffle_epi8(b, b2)));
}
static inline void store_rgba64(uint8_t* d, __m128i r, __m128i g, __m128i b) {
  const __m128i a = _mm_set1_epi8(-1);
  const __m128i rg_lo = _mm_unpacklo_epi8(r, g), rg_hi = _mm_unpackhi_epi8(r, g);
  const __m128i ba_lo = _mm_unpacklo_epi8(b, a), ba_hi = _mm_unpackhi_epi8(b, a);
  _mm_storeu_si128((__m128i*)d, _mm_unpacklo_epi16(rg_lo, ba_lo));
  _mm_storeu_si128((__m128i*)(d + 16), _mm_unpackhi_epi16(rg_lo, ba_lo));
  _mm_storeu_si128((__m128i*)(d + 32), _mm_unpacklo_epi16(rg_hi, ba_hi));
  _mm_storeu_si128((__m128i*)(d + 48), _mm_unpackhi_epi16(rg_hi, ba_hi));
}
/* the chroma pairs of pixels x .. x+15 as u0 v0 u1 v1 ...; uv_pix 1: two planes, 2: interleaved (either order) */
static inline __m128i load_uv16(const uint8_t* pu, const uint8_t* pv, int x, int uv_pix) {
  if (uv_pix == 1) return _mm_unpacklo_epi8(_mm_loadl_epi64((const __m128i*)(pu + (x >> 1))), _mm_loadl_epi64((const __m128i*)(pv + (x >> 1))));
  if (pv == pu + 1) return _mm_loadu_si128((const __m128i*)(pu + x));                /* U first: x >> 1 pairs of 2 bytes */
  const __m128i vu = _mm_loadu_si128((const __m128i*)(pv + x));                      /* V first: swap the bytes of every pair */
  return _mm_shuffle_epi8(vu, _mm_setr_epi8(1, 0, 3, 2, 5, 4, 7, 6, 9, 8, 11, 10, 13, 12, 15, 14));
}
/* -> number of leading pixels of the row done 16 at a time (0 when this layout has no vector form) */
static int yuv_row_avx2(const uint8_t* py, const uint8_t* pu, const uint8_t* pv, int uv_pix, uint8_t* d, int width, int bpp) {
  if (uv_pix == 2 && pv != pu + 1 && pu != pv + 1) return 0; /* two separate pixel-stride-2 buffers: scalar */
  int x = 0;
  for (; x + 16 <= width; x += 16) {
    __m128i r, g, b;
    yuv16_avx2(py + x, load_uv16(pu, pv, x, uv_pix), &r, &g, &b);
    if (bpp == 3) store_rgb48(d + 3 * x, r, g, b);
    else store_rgba64(d + 4 * x, r, g, b);
  }
  return x;
}
#else
static int yuv_row_avx2(const uint8_t* py, const uint8_t* pu, const uint8_t* pv, int uv_pix, uint8_t* d, int width, int bpp) {
  (void)py; (void)pu; (void)pv; (void)uv_pix; (void)d; (void)width; (void)bpp;
  return 0;
}
#endif

/* Both forms over all 2^24 (y, u, v) triples -> number of differing channel values (0 expected; also 0 without AVX2, where
 * nothing is compared). */
long orc_yuv_rows_selftest(void) {
  long bad = 0;
  uint8_t ys[16], us[8], vs[8], out[64], r, g, b;
  for (int u = 0; u < 256; ++u)
    for (int v = 0; v < 256; ++v)
      for (int y0 = 0; y0 < 256; y0 += 16) {
        for (int k = 0; k < 16; ++k) ys[k] = (uint8_t)(y0 + k);
        for (int k = 0; k < 8; ++k) { us[k] = (uint8_t)u; vs[k] = (uint8_t)(v ^ (k & 1 ? 0x55 : 0)); }
        if (yuv_row_avx2(ys, us, vs, 1, out, 16, 3) != 16) return 0;
        for (int k = 0; k < 16; ++k) {
          yuv_pixel(ys[k], us[k >> 1], vs[k >> 1], &r, &g, &b);
          bad += (out[3 * k] != r) + (out[3 * k + 1] != g) + (out[3 * k + 2] != b);
        }
      }
  return bad;
}

/* one row, BPP and the chroma pixel stride as compile-time constants so that the compiler can unroll and vectorise */
#define ORC_ROW_FN(NAME, BPP, UVPIX)                                                                                  \
  static void NAME(const uint8_t* restrict py, const uint8_t* restrict pu, const uint8_t* restrict pv,                \
                   uint8_t* restrict d, int width) {                                                                  \
    for (int x = 0; x < width; ++x) {                                                                                 \
      uint8_t r, g, b;                                                                                                \
      yuv_pixel(py[x], pu[(x >> 1) * (UVPIX)], pv[(x >> 1) * (UVPIX)], &r, &g, &b);                                   \
      d[(BPP) * x] = r;                                                                                               \
      d[(BPP) * x + 1] = g;                                                                                           \
      d[(BPP) * x + 2] = b;                                                                                           \
      if ((BPP) == 4) d[(BPP) * x + 3] = 255;                                                                         \
    }                                                                                                                 \
  }
ORC_ROW_FN(row_rgb_planar, 3, 1)
ORC_ROW_FN(row_rgb_semi, 3, 2)
ORC_ROW_FN(row_rgba_planar, 4, 1)
ORC_ROW_FN(row_rgba_semi, 4, 2)

/* Generic 4:2:0 -> packed RGB(A) walker. Chroma is nearest: pixel (x,y) uses
 * sample (x>>1, y>>1) (NV12ToRGB24Row_C / I422ToRGB24Row_C + the `if (y & 1)`
 * chroma-row advance of the *Matrix functions in convert_argb.cc). */
static int yuv420_to_packed(const uint8_t* src_y, int sy, const uint8_t* src_u, const uint8_t* src_v,
                            int suv, int uv_pix, uint8_t* dst, int dstride, int width, int height,
                            int bpp) {
  if (!src_y || !src_u || !src_v || !dst || width <= 0 || height == 0) return -1;
  if (height < 0) { /* negative height = inverted image (libyuv convention) */
    height = -height;
    dst = dst + (ptrdiff_t)(height - 1) * dstride;
    dstride = -dstride;
  }
  for (int y = 0; y < height; ++y) {
    const uint8_t* py = src_y + (ptrdiff_t)y * sy;
    const uint8_t* pu = src_u + (ptrdiff_t)(y >> 1) * suv;
    const uint8_t* pv = src_v + (ptrdiff_t)(y >> 1) * suv;
    uint8_t* d = dst + (ptrdiff_t)y * dstride;
    {  /* sixteen pixels at a time where the layout allows, the rest of the row with the scalar form */
      const int done = yuv_row_avx2(py, pu, pv, uv_pix, d, width, bpp);
      if (done) {
        for (int x = done; x < width; ++x) {
          uint8_t r, g, b;
          yuv_pixel(py[x], pu[(x >> 1) * uv_pix], pv[(x >> 1) * uv_pix], &r, &g, &b);
          d[bpp * x] = r; d[bpp * x + 1] = g; d[bpp * x + 2] = b;
          if (bpp == 4) d[bpp * x + 3] = 255;
        }
        continue;
      }
    }
    if (bpp == 3 && uv_pix == 1) row_rgb_planar(py, pu, pv, d, width);
    else if (bpp == 3 && uv_pix == 2) row_rgb_semi(py, pu, pv, d, width);
    else if (bpp == 4 && uv_pix == 1) row_rgba_planar(py, pu, pv, d, width);
    else if (bpp == 4 && uv_pix == 2) row_rgba_semi(py, pu, pv, d, width);
    else {
      for (int x = 0; x < width; ++x) {
        uint8_t r, g, b;
        yuv_pixel(py[x], pu[(x >> 1) * uv_pix], pv[(x >> 1) * uv_pix], &r, &g, &b);
        d[0] = r;
        d[1] = g;
        d[2] = b;
        if (bpp == 4) d[3] = 255;
        d += bpp;
      }
    }
  }
  return 0;
}

int orc_NV12ToRAW(const uint8_t* src_y, int src_stride_y, const uint8_t* src_uv, int src_stride_uv,
                  uint8_t* dst_raw, int dst_stride_raw, int width, int height) {
  return yuv420_to_packed(src_y, src_stride_y, src_uv, src_uv ? src_uv + 1 : 0, src_stride_uv, 2, dst_raw,
                          dst_stride_raw, width, height, 3);
}
int orc_NV21ToRAW(const uint8_t* src_y, int src_stride_y, const uint8_t* src_vu, int src_stride_vu,
                  uint8_t* dst_raw, int dst_stride_raw, int width, int height) {
  return yuv420_to_packed(src_y, src_stride_y, src_vu ? src_vu + 1 : 0, src_vu, src_stride_vu, 2, dst_raw,
                          dst_stride_raw, width, height, 3);
}
/* libyuv "ABGR" = bytes R,G,B,A in memory = Band kRGBA. */
int orc_NV12ToABGR(const uint8_t* src_y, int src_stride_y, const uint8_t* src_uv, int src_stride_uv,
                   uint8_t* dst_abgr, int dst_stride_abgr, int width, int height) {
  return yuv420_to_packed(src_y, src_stride_y, src_uv, src_uv ? src_uv + 1 : 0, src_stride_uv, 2, dst_abgr,
                          dst_stride_abgr, width, height, 4);
}
int orc_NV21ToABGR(const uint8_t* src_y, int src_stride_y, const uint8_t* src_vu, int src_stride_vu,
                   uint8_t* dst_abgr, int dst_stride_abgr, int width, int height) {
  return yuv420_to_packed(src_y, src_stride_y, src_vu ? src_vu + 1 : 0, src_vu, src_stride_vu, 2, dst_abgr,
                          dst_stride_abgr, width, height, 4);
}
int orc_I420ToRAW(const uint8_t* src_y, int src_stride_y, const uint8_t* src_u, int src_stride_u,
                  const uint8_t* src_v, int src_stride_v, uint8_t* dst_raw, int dst_stride_raw, int width,
                  int height) {
  if (src_stride_u != src_stride_v) return -1; /* the reference always passes equal strides (:278-283) */
  return yuv420_to_packed(src_y, src_stride_y, src_u, src_v, src_stride_u, 1, dst_raw, dst_stride_raw, width,
                          height, 3);
}
int orc_I420ToABGR(const uint8_t* src_y, int src_stride_y, const uint8_t* src_u, int src_stride_u,
                   const uint8_t* src_v, int src_stride_v, uint8_t* dst_abgr, int dst_stride_abgr, int width,
                   int height) {
  if (src_stride_u != src_stride_v) return -1;
  return yuv420_to_packed(src_y, src_stride_y, src_u, src_v, src_stride_u, 1, dst_abgr, dst_stride_abgr,
                          width, height, 4);
}

/* ---- planar_functions.cc: CopyPlane / Split / Merge / Swap -------------- */
void orc_CopyPlane(const uint8_t* src, int src_stride, uint8_t* dst, int dst_stride, int width, int height) {
  if (width <= 0 || height == 0) return;
  if (height < 0) { /* negative height: vertical flip (libyuv_image_operator.cc:899-904) */
    height = -height;
    dst = dst + (ptrdiff_t)(height - 1) * dst_stride;
    dst_stride = -dst_stride;
  }
  if (src == dst && src_stride == dst_stride) return;
  for (int y = 0; y < height; ++y) {
    memcpy(dst, src, (size_t)width);
    src += src_stride;
    dst += dst_stride;
  }
}
void orc_SplitUVPlane(const uint8_t* src_uv, int src_stride_uv, uint8_t* dst_u, int dst_stride_u,
                      uint8_t* dst_v, int dst_stride_v, int width, int height) {
  if (width <= 0 || height == 0) return;
  if (height < 0) {
    height = -height;
    dst_u = dst_u + (ptrdiff_t)(height - 1) * dst_stride_u;
    dst_v = dst_v + (ptrdiff_t)(height - 1) * dst_stride_v;
    dst_stride_u = -dst_stride_u;
    dst_stride_v = -dst_stride_v;
  }
  for (int y = 0; y < height; ++y) {
    for (int x = 0; x < width; ++x) {
      dst_u[x] = src_uv[2 * x];
      dst_v[x] = src_uv[2 * x + 1];
    }
    src_uv += src_stride_uv;
    dst_u += dst_stride_u;
    dst_v += dst_stride_v;
  }
}
void orc_MergeUVPlane(const uint8_t* src_u, int src_stride_u, const uint8_t* src_v, int src_stride_v,
                      uint8_t* dst_uv, int dst_stride_uv, int width, int height) {
  if (width <= 0 || height == 0) return;
  if (height < 0) {
    height = -height;
    dst_uv = dst_uv + (ptrdiff_t)(height - 1) * dst_stride_uv;
    dst_stride_uv = -dst_stride_uv;
  }
  for (int y = 0; y < height; ++y) {
    for (int x = 0; x < width; ++x) {
      dst_uv[2 * x] = src_u[x];
      dst_uv[2 * x + 1] = src_v[x];
    }
    src_u += src_stride_u;
    src_v += src_stride_v;
    dst_uv += dst_stride_uv;
  }
}
void orc_SwapUVPlane(const uint8_t* src_uv, int src_stride_uv, uint8_t* dst_vu, int dst_stride_vu, int width,
                     int height) {
  if (width <= 0 || height == 0) return;
  if (height < 0) {
    height = -height;
    src_uv = src_uv + (ptrdiff_t)(height - 1) * src_stride_uv;
    src_stride_uv = -src_stride_uv;
  }
  for (int y = 0; y < height; ++y) {
    for (int x = 0; x < width; ++x) {
      uint8_t u = src_uv[2 * x], v = src_uv[2 * x + 1];
      dst_vu[2 * x] = v;
      dst_vu[2 * x + 1] = u;
    }
    src_uv += src_stride_uv;
    dst_vu += dst_stride_vu;
  }
}
int orc_I420Copy(const uint8_t* src_y, int src_stride_y, const uint8_t* src_u, int src_stride_u,
                 const uint8_t* src_v, int src_stride_v, uint8_t* dst_y, int dst_stride_y, uint8_t* dst_u,
                 int dst_stride_u, uint8_t* dst_v, int dst_stride_v, int width, int height) {
  int halfwidth = (width + 1) >> 1;
  int halfheight = (height + 1) >> 1;
  if (!src_u || !src_v || !dst_u || !dst_v || width <= 0 || height == 0) return -1;
  if (height < 0) { /* negative height inverts the SOURCE (convert.cc I420Copy) */
    height = -height;
    halfheight = (height + 1) >> 1;
    src_y = src_y + (ptrdiff_t)(height - 1) * src_stride_y;
    src_u = src_u + (ptrdiff_t)(halfheight - 1) * src_stride_u;
    src_v = src_v + (ptrdiff_t)(halfheight - 1) * src_stride_v;
    src_stride_y = -src_stride_y;
    src_stride_u = -src_stride_u;
    src_stride_v = -src_stride_v;
  }
  if (dst_y) orc_CopyPlane(src_y, src_stride_y, dst_y, dst_stride_y, width, height);
  orc_CopyPlane(src_u, src_stride_u, dst_u, dst_stride_u, halfwidth, halfheight);
  orc_CopyPlane(src_v, src_stride_v, dst_v, dst_stride_v, halfwidth, halfheight);
  return 0;
}

/* ---- convert.cc / convert_from.cc: 4:2:0 repacks -------------------------
 * libyuv_image_operator.cc:131 (NV12ToI420), :219 (NV21ToI420), :307
 * (I420ToNV12), :325 (I420ToNV21). Pure byte moves. */
int orc_NV12ToI420(const uint8_t* src_y, int src_stride_y, const uint8_t* src_uv, int src_stride_uv,
                   uint8_t* dst_y, int dst_stride_y, uint8_t* dst_u, int dst_stride_u, uint8_t* dst_v,
                   int dst_stride_v, int width, int height) {
  int halfwidth = (width + 1) >> 1;
  int halfheight = (height + 1) >> 1;
  if (!src_uv || !dst_u || !dst_v || width <= 0 || height == 0) return -1;
  if (height < 0) return -1; /* not used by the reference */
  if (dst_y) orc_CopyPlane(src_y, src_stride_y, dst_y, dst_stride_y, width, height);
  orc_SplitUVPlane(src_uv, src_stride_uv, dst_u, dst_stride_u, dst_v, dst_stride_v, halfwidth, halfheight);
  return 0;
}
int orc_NV21ToI420(const uint8_t* src_y, int src_stride_y, const uint8_t* src_vu, int src_stride_vu,
                   uint8_t* dst_y, int dst_stride_y, uint8_t* dst_u, int dst_stride_u, uint8_t* dst_v,
                   int dst_stride_v, int width, int height) {
  return orc_NV12ToI420(src_y, src_stride_y, src_vu, src_stride_vu, dst_y, dst_stride_y, dst_v, dst_stride_v,
                        dst_u, dst_stride_u, width, height);
}
int orc_I420ToNV12(const uint8_t* src_y, int src_stride_y, const uint8_t* src_u, int src_stride_u,
                   const uint8_t* src_v, int src_stride_v, uint8_t* dst_y, int dst_stride_y, uint8_t* dst_uv,
                   int dst_stride_uv, int width, int height) {
  int halfwidth = (width + 1) / 2;
  int halfheight = (height + 1) / 2;
  if (!src_y || !src_u || !src_v || !dst_uv || width <= 0 || height == 0) return -1;
  if (height < 0) return -1;
  if (dst_y) orc_CopyPlane(src_y, src_stride_y, dst_y, dst_stride_y, width, height);
  orc_MergeUVPlane(src_u, src_stride_u, src_v, src_stride_v, dst_uv, dst_stride_uv, halfwidth, halfheight);
  return 0;
}
int orc_I420ToNV21(const uint8_t* src_y, int src_stride_y, const uint8_t* src_u, int src_stride_u,
                   const uint8_t* src_v, int src_stride_v, uint8_t* dst_y, int dst_stride_y, uint8_t* dst_vu,
                   int dst_stride_vu, int width, int height) {
  return orc_I420ToNV12(src_y, src_stride_y, src_v, src_stride_v, src_u, src_stride_u, dst_y, dst_stride_y,
                        dst_vu, dst_stride_vu, width, height);
}

/* ---- packed shuffles (row_common.cc RGB24ToARGBRow_C, ARGBToRGB24Row_C,
 * ARGBShuffleRow_C with kShuffleMaskABGRToARGB) ----------------------------
 * libyuv_image_operator.cc:455, :482/:678, :513. */
int orc_RGB24ToARGB(const uint8_t* src, int src_stride, uint8_t* dst, int dst_stride, int width, int height) {
  if (!src || !dst || width <= 0 || height <= 0) return -1;
  for (int y = 0; y < height; ++y) {
    const uint8_t* s = src + (ptrdiff_t)y * src_stride;
    uint8_t* d = dst + (ptrdiff_t)y * dst_stride;
    int x = 0;
#ifdef __AVX2__
    { /* four pixels per shuffle (a 16-byte load needs 16 readable bytes: stop six pixels before the end of the row) */
      const __m128i m = _mm_setr_epi8(0, 1, 2, -1, 3, 4, 5, -1, 6, 7, 8, -1, 9, 10, 11, -1);
      const __m128i a = _mm_set1_epi32((int)0xFF000000u);
      for (; x + 6 <= width; x += 4, s += 12, d += 16)
        _mm_storeu_si128((__m128i*)d, _mm_or_si128(_mm_shuffle_epi8(_mm_loadu_si128((const __m128i*)s), m), a));
    }
#endif
    for (; x < width; ++x) {
      d[0] = s[0];
      d[1] = s[1];
      d[2] = s[2];
      d[3] = 255u;
      d += 4;
      s += 3;
    }
  }
  return 0;
}
int orc_ARGBToRGB24(const uint8_t* src, int src_stride, uint8_t* dst, int dst_stride, int width, int height) {
  if (!src || !dst || width <= 0 || height <= 0) return -1;
  for (int y = 0; y < height; ++y) {
    const uint8_t* s = src + (ptrdiff_t)y * src_stride;
    uint8_t* d = dst + (ptrdiff_t)y * dst_stride;
    for (int x = 0; x < width; ++x) {
      d[0] = s[0];
      d[1] = s[1];
      d[2] = s[2];
      d += 3;
      s += 4;
    }
  }
  return 0;
}
int orc_ABGRToARGB(const uint8_t* src, int src_stride, uint8_t* dst, int dst_stride, int width, int height) {
  if (!src || !dst || width <= 0 || height <= 0) return -1;
  for (int y = 0; y < height; ++y) {
    const uint8_t* s = src + (ptrdiff_t)y * src_stride;
    uint8_t* d = dst + (ptrdiff_t)y * dst_stride;
    for (int x = 0; x < width; ++x) {
      uint8_t b0 = s[0], b1 = s[1], b2 = s[2], b3 = s[3];
      d[0] = b2;
      d[1] = b1;
      d[2] = b0;
      d[3] = b3;
      d += 4;
      s += 4;
    }
  }
  return 0;
}

/* ---- J400 luma (row_common.cc RGBToYJ, 8-bit form of libyuv >= ~1741):
 *   YJ = (77 R + 150 G + 29 B + 128) >> 8
 * RAWToJ400: bytes are R,G,B (libyuv_image_operator.cc:526);
 * ARGBToJ400: bytes are B,G,R,A (:611, after ABGRToARGB :513). */
static inline uint8_t rgb_to_yj(uint8_t r, uint8_t g, uint8_t b) {
  return (uint8_t)((77 * r + 150 * g + 29 * b + 128) >> 8);
}
int orc_RAWToJ400(const uint8_t* src_raw, int src_stride_raw, uint8_t* dst_yj, int dst_stride_yj, int width,
                  int height) {
  if (!src_raw || !dst_yj || width <= 0 || height <= 0) return -1;
  for (int y = 0; y < height; ++y) {
    const uint8_t* s = src_raw + (ptrdiff_t)y * src_stride_raw;
    uint8_t* d = dst_yj + (ptrdiff_t)y * dst_stride_yj;
    for (int x = 0; x < width; ++x) d[x] = rgb_to_yj(s[3 * x], s[3 * x + 1], s[3 * x + 2]);
  }
  return 0;
}
int orc_ARGBToJ400(const uint8_t* src_argb, int src_stride_argb, uint8_t* dst_yj, int dst_stride_yj, int width,
                   int height) {
  if (!src_argb || !dst_yj || width <= 0 || height <= 0) return -1;
  for (int y = 0; y < height; ++y) {
    const uint8_t* s = src_argb + (ptrdiff_t)y * src_stride_argb;
    uint8_t* d = dst_yj + (ptrdiff_t)y * dst_stride_yj;
    for (int x = 0; x < width; ++x) d[x] = rgb_to_yj(s[4 * x + 2], s[4 * x + 1], s[4 * x]);
  }
  return 0;
}

/* ---- RGB -> 4:2:0 (row_common.cc MAKEROWY: RGBToY, RGB2xToU/V; the default
 * C build, i.e. neither LIBYUV_RGBTOU_TRUNCATE nor LIBYUV_ARGBTOUV_PAVGB).
 * VERSION/PLATFORM-DEPENDENT in libyuv (x86 SIMD truncates); "next" row N2.
 * libyuv_image_operator.cc:570 (RAWToI420), :626/:643/:661 (ABGRTo*). */
static inline uint8_t rgb_to_y(uint8_t r, uint8_t g, uint8_t b) {
  return (uint8_t)((66 * r + 129 * g + 25 * b + 0x1080) >> 8);
}
static inline uint8_t rgb2x_to_u(int r, int g, int b) { return (uint8_t)((56 * b - 37 * g - 19 * r + 0x8080) >> 8); }
static inline uint8_t rgb2x_to_v(int r, int g, int b) { return (uint8_t)((56 * r - 47 * g - 9 * b + 0x8080) >> 8); }

/* bytes at offsets R,G,B of each bpp-byte pixel */
static void rgbx_to_y_row(const uint8_t* s, uint8_t* dy, int width, int bpp) {
  for (int x = 0; x < width; ++x) dy[x] = rgb_to_y(s[bpp * x], s[bpp * x + 1], s[bpp * x + 2]);
}
static void rgbx_to_uv_row(const uint8_t* s, int stride, uint8_t* du, uint8_t* dv, int width, int bpp) {
  const uint8_t* t = s + stride;
  int x;
  for (x = 0; x < width - 1; x += 2) {
    int ar = (s[0] + s[bpp] + t[0] + t[bpp] + 1) >> 1;
    int ag = (s[1] + s[bpp + 1] + t[1] + t[bpp + 1] + 1) >> 1;
    int ab = (s[2] + s[bpp + 2] + t[2] + t[bpp + 2] + 1) >> 1;
    *du++ = rgb2x_to_u(ar, ag, ab);
    *dv++ = rgb2x_to_v(ar, ag, ab);
    s += 2 * bpp;
    t += 2 * bpp;
  }
  if (width & 1) {
    int ar = s[0] + t[0], ag = s[1] + t[1], ab = s[2] + t[2];
    *du = rgb2x_to_u(ar, ag, ab);
    *dv = rgb2x_to_v(ar, ag, ab);
  }
}
static int rgbx_to_i420(const uint8_t* src, int sstride, uint8_t* dst_y, int dsy, uint8_t* dst_u, int dsu,
                        uint8_t* dst_v, int dsv, int width, int height, int bpp) {
  if (!src || !dst_y || !dst_u || !dst_v || width <= 0 || height <= 0) return -1;
  int y;
  for (y = 0; y < height - 1; y += 2) {
    rgbx_to_uv_row(src, sstride, dst_u, dst_v, width, bpp);
    rgbx_to_y_row(src, dst_y, width, bpp);
    rgbx_to_y_row(src + sstride, dst_y + dsy, width, bpp);
    src += 2 * (ptrdiff_t)sstride;
    dst_y += 2 * (ptrdiff_t)dsy;
    dst_u += dsu;
    dst_v += dsv;
  }
  if (height & 1) {
    rgbx_to_uv_row(src, 0, dst_u, dst_v, width, bpp);
    rgbx_to_y_row(src, dst_y, width, bpp);
  }
  return 0;
}
int orc_RAWToI420(const uint8_t* src_raw, int src_stride_raw, uint8_t* dst_y, int dst_stride_y, uint8_t* dst_u,
                  int dst_stride_u, uint8_t* dst_v, int dst_stride_v, int width, int height) {
  return rgbx_to_i420(src_raw, src_stride_raw, dst_y, dst_stride_y, dst_u, dst_stride_u, dst_v, dst_stride_v,
                      width, height, 3);
}
int orc_ABGRToI420(const uint8_t* src_abgr, int src_stride_abgr, uint8_t* dst_y, int dst_stride_y,
                   uint8_t* dst_u, int dst_stride_u, uint8_t* dst_v, int dst_stride_v, int width, int height) {
  return rgbx_to_i420(src_abgr, src_stride_abgr, dst_y, dst_stride_y, dst_u, dst_stride_u, dst_v,
                      dst_stride_v, width, height, 4);
}
static int abgr_to_nv(const uint8_t* src, int sstride, uint8_t* dst_y, int dsy, uint8_t* dst_c, int dsc,
                      int width, int height, int v_first) {
  if (!src || !dst_y || !dst_c || width <= 0 || height <= 0) return -1;
  int halfwidth = (width + 1) >> 1;
  uint8_t* row_u = (uint8_t*)malloc((size_t)halfwidth * 2);
  uint8_t* row_v = row_u + halfwidth;
  int y;
  for (y = 0; y < height; y += 2) {
    int last = (y == height - 1);
    rgbx_to_uv_row(src, last ? 0 : sstride, row_u, row_v, width, 4);
    for (int x = 0; x < halfwidth; ++x) {
      dst_c[2 * x] = v_first ? row_v[x] : row_u[x];
      dst_c[2 * x + 1] = v_first ? row_u[x] : row_v[x];
    }
    rgbx_to_y_row(src, dst_y, width, 4);
    if (!last) rgbx_to_y_row(src + sstride, dst_y + dsy, width, 4);
    src += 2 * (ptrdiff_t)sstride;
    dst_y += 2 * (ptrdiff_t)dsy;
    dst_c += dsc;
  }
  free(row_u);
  return 0;
}
int orc_ABGRToNV12(const uint8_t* src_abgr, int src_stride_abgr, uint8_t* dst_y, int dst_stride_y,
                   uint8_t* dst_uv, int dst_stride_uv, int width, int height) {
  return abgr_to_nv(src_abgr, src_stride_abgr, dst_y, dst_stride_y, dst_uv, dst_stride_uv, width, height, 0);
}
int orc_ABGRToNV21(const uint8_t* src_abgr, int src_stride_abgr, uint8_t* dst_y, int dst_stride_y,
                   uint8_t* dst_vu, int dst_stride_vu, int width, int height) {
  return abgr_to_nv(src_abgr, src_stride_abgr, dst_y, dst_stride_y, dst_vu, dst_stride_vu, width, height, 1);
}
