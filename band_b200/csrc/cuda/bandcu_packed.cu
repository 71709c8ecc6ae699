// band_b200/csrc/cuda/bandcu_packed.cu
//
// Bilinear Resize of PACKED sources (gray / RGB / RGBA) without index maps, optionally behind a per-fetch
// ColorSpaceConvert (RGBA -> RGB, RGB -> gray, RGBA -> gray) and in front of Normalize: the C5 sweep of
// BASELINE.json (AutoConvert = CSC -> Resize -> DataTypeConvert, image_operator.cc:465-491) and every
// ResizeRgb / ResizeRgba / ResizeGray call (libyuv_image_operator.cc:1146-1229,1367-1380) that takes libyuv's
// general bilinear paths:
//   ScaleARGBBilinearDown / ScalePlaneBilinearDown : rows blended first, then columns
//   ScaleARGBBilinearUp   / ScalePlaneBilinearUp   : columns first, then rows
// with the ARGB column filter ((a*(0x7f^f) + b*f) >> 7, truncating) for 3/4-channel images and the x86 plane
// filter (a + (((f>>9)*(b-a) + 0x40) >> 7)) for gray. Same arithmetic as fused_generic_kernel (the parity tests run
// both); what differs is the cost per pixel: ncu showed the generic kernel issue-bound at ~240 instructions per
// output pixel on the 64^2 -> 2048^2 case. Here one thread produces FOUR adjacent output pixels of a row: the row
// terms are shared, all index arithmetic is 32-bit, sources are read through the read-only path, and the four
// results leave as 32/128-bit stores.
#include <climits>
#include <cstring>

#include "bandcu_device.cuh"
#include "bandcu_internal.h"

namespace bandcu {

struct PackedParams {
  int sw, sh, stride0;
  int dw, dh;
  int x0, y0, dx, dy, max_y, linear;
  long long dst_row_stride;
  int vec_ok;          // outputs are aligned for 32/128-bit stores
  int up_rows;         // packed_up_rows_kernel: consecutive output rows per thread
  const float* lut;    // Normalize table (f32 outputs) or nullptr
  BatchArgs batch;
};

// One source pixel after the per-fetch conversion, channels packed c0 | c1<<8 | c2<<16 | c3<<24.
template <int SRCCH, int CVT>
__device__ __forceinline__ uint32_t packed_fetch(const uint8_t* __restrict__ row, int c) {
  if (SRCCH == 1) return __ldg(row + c);
  if (SRCCH == 4) {
    const uint32_t w = __ldg(reinterpret_cast<const uint32_t*>(row) + c);  // frames are 4-byte aligned (checked on the host)
    if (CVT == BANDCU_CVT_RGBA2GRAY) return rgb_to_yj(w & 0xFF, (w >> 8) & 0xFF, (w >> 16) & 0xFF);
    if (CVT == BANDCU_CVT_RGBA2RGB) return w & 0x00FFFFFFu;
    return w;
  }
  const uint8_t* q = row + c * 3;
  const uint32_t R = __ldg(q), G = __ldg(q + 1), B = __ldg(q + 2);
  if (CVT == BANDCU_CVT_RGB2GRAY) return rgb_to_yj(R, G, B);
  return R | (G << 8) | (B << 16);
}

template <bool PLANE>
__device__ __forceinline__ uint32_t packed_hblend(uint32_t a, uint32_t b, int x) {
  if (PLANE) return blend_cols1_round7(a, b, x & 0xffff);
  return blend_cols4_trunc7(a, b, (x >> 9) & 0x7f);
}

// Four adjacent output pixels (channels packed c0 | c1<<8 | ...) -> one output row at column ox0: 32/128-bit stores
// when the row is aligned and the four pixels exist, byte stores otherwise. f32 outputs go through the Normalize table.
template <int OUTCH, typename OT>
__device__ __forceinline__ void packed_store4(uint8_t* orow, int ox0, const uint32_t (&px)[4], int dw, bool vec_ok,
                                              const float* lut) {
  const bool full4 = (ox0 + 4 <= dw) && vec_ok;
  if (sizeof(OT) == 1) {
    uint8_t* o = orow + (size_t)ox0 * OUTCH;
    if (full4) {
      if (OUTCH == 1) {
        *reinterpret_cast<uint32_t*>(o) = (px[0] & 0xFF) | ((px[1] & 0xFF) << 8) | ((px[2] & 0xFF) << 16) | (px[3] << 24);
      } else if (OUTCH == 3) {  // 12 bytes: R0 G0 B0 R1 | G1 B1 R2 G2 | B2 R3 G3 B3
        uint32_t* o4 = reinterpret_cast<uint32_t*>(o);
        o4[0] = (px[0] & 0x00FFFFFFu) | (px[1] << 24);
        o4[1] = ((px[1] >> 8) & 0x0000FFFFu) | (px[2] << 16);
        o4[2] = ((px[2] >> 16) & 0x000000FFu) | (px[3] << 8);
      } else {
        *reinterpret_cast<uint4*>(o) = make_uint4(px[0], px[1], px[2], px[3]);
      }
    } else {
      for (int k = 0; k < 4 && ox0 + k < dw; ++k)
        for (int c = 0; c < OUTCH; ++c) o[k * OUTCH + c] = (uint8_t)((px[k] >> (8 * c)) & 0xFF);
    }
  } else {
    float* o = reinterpret_cast<float*>(orow) + (size_t)ox0 * OUTCH;
    if (full4) {
      float v[4 * OUTCH];
#pragma unroll
      for (int k = 0; k < 4; ++k)
#pragma unroll
        for (int c = 0; c < OUTCH; ++c) v[k * OUTCH + c] = lut[(px[k] >> (8 * c)) & 0xFF];
#pragma unroll
      for (int q = 0; q < OUTCH; ++q)
        reinterpret_cast<float4*>(o)[q] = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
    } else {
      for (int k = 0; k < 4 && ox0 + k < dw; ++k)
        for (int c = 0; c < OUTCH; ++c) o[k * OUTCH + c] = lut[(px[k] >> (8 * c)) & 0xFF];
    }
  }
}

enum { kPackedDown = 0, kPackedUp = 1, kPackedBox2 = 2 };

template <int SRCCH, int CVT, int OUTCH, int MODE, typename OT>
__global__ void __launch_bounds__(256) packed_bilinear_kernel(const __grid_constant__ PackedParams P) {
  constexpr bool UP = (MODE == kPackedUp);
  constexpr bool PLANE = (OUTCH == 1);
  __shared__ float lut[256];
  if (sizeof(OT) == 4) {
    const int t = threadIdx.y * blockDim.x + threadIdx.x;
    if (t < 256) lut[t] = P.lut[t];
    __syncthreads();
  }
  const int ox0 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
  const int oy = blockIdx.y * blockDim.y + threadIdx.y;
  if (ox0 >= P.dw || oy >= P.dh) return;
  const FramePtrs f = frame_ptrs(P.batch, blockIdx.z);

  int y = P.y0 + oy * P.dy;
  if (MODE != kPackedBox2 && y > P.max_y) y = P.max_y;
  // Box2 (ScaleARGBRowDown2Box / DownEvenBox / ScaleRowDown2Box): 2x2 average at (y0>>16 + oy*(dy>>16), ...)
  const int yi = MODE == kPackedBox2 ? (P.y0 >> 16) + oy * (P.dy >> 16) : (y >> 16);
  const int yf = (MODE == kPackedBox2 || P.linear) ? 0 : ((y >> 8) & 255);
  const int r0 = min(max(yi, 0), P.sh - 1), r1 = min(max(yi + 1, 0), P.sh - 1);
  const uint8_t* row0 = f.p0 + (size_t)r0 * P.stride0;
  const uint8_t* row1 = f.p0 + (size_t)r1 * P.stride0;

  uint32_t px[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int x = P.x0 + (ox0 + k) * P.dx;
    const int xi = MODE == kPackedBox2 ? (P.x0 >> 16) + (ox0 + k) * (P.dx >> 16) : (x >> 16);
    const int c0 = min(max(xi, 0), P.sw - 1), c1 = min(max(xi + 1, 0), P.sw - 1);
    uint32_t a = packed_fetch<SRCCH, CVT>(row0, c0), b = packed_fetch<SRCCH, CVT>(row0, c1);
    if (MODE == kPackedBox2) {
      px[k] = box2x2_4(a, b, packed_fetch<SRCCH, CVT>(row1, c0), packed_fetch<SRCCH, CVT>(row1, c1));
    } else if (UP) {  // columns, then rows
      uint32_t top = packed_hblend<PLANE>(a, b, x);
      if (yf) {
        const uint32_t a1 = packed_fetch<SRCCH, CVT>(row1, c0), b1 = packed_fetch<SRCCH, CVT>(row1, c1);
        top = interp_rows4(top, packed_hblend<PLANE>(a1, b1, x), yf);
      }
      px[k] = top;
    } else {   // rows, then columns
      if (yf) {
        a = interp_rows4(a, packed_fetch<SRCCH, CVT>(row1, c0), yf);
        b = interp_rows4(b, packed_fetch<SRCCH, CVT>(row1, c1), yf);
      }
      px[k] = packed_hblend<PLANE>(a, b, x);
    }
  }

  packed_store4<OUTCH, OT>(f.out + (long long)oy * P.dst_row_stride, ox0, px, P.dw, P.vec_ok != 0, lut);
}

// Scale*BilinearUp, row-streaming form. libyuv filters the two source rows yi, yi+1 horizontally into row buffers
// and blends them per output row; an up-scale by F reuses each filtered row for ~F output rows. ncu on the
// pixel-per-thread version showed the ALU pipe at 86 % with 32 instructions per output pixel while DRAM sat at 30 %.
// Here the unit of work is the 32-bit OUTPUT WORD: lane l of a warp owns words l, l+32, .. of a row segment (so every
// store instruction writes 128 contiguous bytes, 512 for f32 outputs) and walks P.up_rows consecutive output rows.
// For the current source row pair the thread keeps, per word, the packed terms of InterpolateRow_C
//     (a*(256-f) + b*f + 128) >> 8  ==  ((b-a)*f + (a<<8) + 128) >> 8      (exact in wrap-around lane arithmetic)
// i.e. d = b-a and base = (a<<8)+0x00800080 for the even and the odd bytes, so one output word costs two IMADs, a
// shift and a LOP3. The terms are rebuilt only when yi advances (the old bottom row becomes the new top row).
// Same index clamps and column filters as packed_bilinear_kernel<.., kPackedUp, ..>; the parity tests compare both
// with the oracle.

// Horizontally filtered output word `w` of source row `row` (output bytes 4w .. 4w+3 of that row).
template <int SRCCH, int CVT, int OUTCH>
__device__ __forceinline__ uint32_t up_filter_word(const uint8_t* __restrict__ row, int w, const PackedParams& P) {
  constexpr bool PLANE = (OUTCH == 1);
  auto pixel = [&](int ox) {
    ox = min(ox, P.dw - 1);  // bytes past the end of the row are never stored
    const int x = P.x0 + ox * P.dx;
    const int xi = x >> 16;
    const int c0 = min(max(xi, 0), P.sw - 1), c1 = min(max(xi + 1, 0), P.sw - 1);
    return packed_hblend<PLANE>(packed_fetch<SRCCH, CVT>(row, c0), packed_fetch<SRCCH, CVT>(row, c1), x);
  };
  if (OUTCH == 4) return pixel(w);
  if (OUTCH == 1)
    return (pixel(4 * w) & 0xFF) | ((pixel(4 * w + 1) & 0xFF) << 8) | ((pixel(4 * w + 2) & 0xFF) << 16) | (pixel(4 * w + 3) << 24);
  // RGB: the word straddles pixels p and p+1; r = first channel of p inside the word
  const int q = 4 * w, p = q / 3, r = q - 3 * p;
  const uint32_t A = pixel(p), B = pixel(p + 1);
  return __byte_perm(A, B, r == 0 ? 0x4210u : r == 1 ? 0x5421u : 0x6542u);
}

struct UpTerms {
  uint32_t d_lo, d_hi, base_lo, base_hi;
};
__device__ __forceinline__ UpTerms up_terms(uint32_t a, uint32_t b) {
  const uint32_t a_lo = a & 0x00FF00FFu, a_hi = (a >> 8) & 0x00FF00FFu;
  UpTerms t;
  t.d_lo = (b & 0x00FF00FFu) - a_lo;
  t.d_hi = ((b >> 8) & 0x00FF00FFu) - a_hi;
  t.base_lo = (a_lo << 8) + 0x00800080u;
  t.base_hi = (a_hi << 8) + 0x00800080u;
  return t;
}

template <int SRCCH, int CVT, int OUTCH, typename OT, int NWT>
__global__ void __launch_bounds__(128) packed_up_rows_kernel(const __grid_constant__ PackedParams P) {
  __shared__ float lut[256];
  if (sizeof(OT) == 4) {
    const int t = threadIdx.y * blockDim.x + threadIdx.x;
    for (int i = t; i < 256; i += blockDim.x * blockDim.y) lut[i] = P.lut[i];
    __syncthreads();
  }
  const int row_bytes = P.dw * OUTCH;
  const int row_words = (row_bytes + 3) >> 2;             // output words per row (the last one may be partial)
  const int full_words = row_bytes >> 2;
  const int w0 = blockIdx.x * (32 * NWT) + threadIdx.x;  // this lane's words: w0 + 32 j
  const int oy_begin = (blockIdx.y * blockDim.y + threadIdx.y) * P.up_rows;
  if (w0 >= row_words || oy_begin >= P.dh) return;
  const int oy_end = min(P.dh, oy_begin + P.up_rows);
  const FramePtrs f = frame_ptrs(P.batch, blockIdx.z);
  // how many of this lane's words exist / are complete (the compiler predicates on these instead of branching)
  const int n_mine = min(NWT, (row_words - w0 + 31) >> 5);
  const int n_full = max(0, min(NWT, (full_words - w0 + 31) >> 5));

  uint32_t h1[NWT];   // filtered bottom row (becomes the top row when yi advances by one)
  UpTerms t[NWT];
  int cur_yi = INT_MIN, have1 = -1;
  int y_raw = P.y0 + oy_begin * P.dy;
  uint8_t* orow = f.out + (long long)oy_begin * P.dst_row_stride;
  for (int oy = oy_begin; oy < oy_end; ++oy, orow += P.dst_row_stride, y_raw += P.dy) {
    const int y = min(y_raw, P.max_y);
    const int yi = y >> 16;
    const uint32_t yf = P.linear ? 0u : (uint32_t)((y >> 8) & 255);
    if (yi != cur_yi) {
      const int r0 = min(max(yi, 0), P.sh - 1), r1 = min(max(yi + 1, 0), P.sh - 1);
      const uint8_t* row0 = f.p0 + (size_t)r0 * P.stride0;
      const uint8_t* row1 = f.p0 + (size_t)r1 * P.stride0;
      const bool reuse = (r0 == have1);
#pragma unroll
      for (int j = 0; j < NWT; ++j) {
        if (j < n_mine) {
          const int w = w0 + 32 * j;
          const uint32_t a = reuse ? h1[j] : up_filter_word<SRCCH, CVT, OUTCH>(row0, w, P);
          const uint32_t b = (r1 == r0) ? a : up_filter_word<SRCCH, CVT, OUTCH>(row1, w, P);
          h1[j] = b;
          t[j] = up_terms(a, b);
        }
      }
      cur_yi = yi;
      have1 = r1;
    }
#pragma unroll
    for (int j = 0; j < NWT; ++j) {
      const uint32_t lo = t[j].d_lo * yf + t[j].base_lo, hi = t[j].d_hi * yf + t[j].base_hi;
      const uint32_t v = __byte_perm(lo, hi, 0x7351);  // ((lo >> 8) & 0x00FF00FF) | (hi & 0xFF00FF00)
      const int w = w0 + 32 * j;
      if (sizeof(OT) == 1) {
        if (j < n_full) reinterpret_cast<uint32_t*>(orow)[w] = v;
        else if (j < n_mine) for (int k = 4 * w; k < row_bytes; ++k) orow[k] = (uint8_t)(v >> (8 * (k - 4 * w)));
      } else {
        float* o = reinterpret_cast<float*>(orow) + 4 * w;
        if (j < n_full)
          *reinterpret_cast<float4*>(o) = make_float4(lut[v & 0xFF], lut[(v >> 8) & 0xFF], lut[(v >> 16) & 0xFF], lut[v >> 24]);
        else if (j < n_mine) for (int k = 4 * w; k < row_bytes; ++k) o[k - 4 * w] = lut[(v >> (8 * (k - 4 * w))) & 0xFF];
      }
    }
  }
}

// u8 outputs with 16-byte aligned rows: lane l owns the FOUR CONSECUTIVE words 4l .. 4l+3 of a 512-byte row segment and
// stores them as one 128-bit store (the word-interleaved form above spends an address computation, a bounds test and a
// store instruction per word: ncu counted 37 instructions per output word on 64^2 -> 2048^2, the ALU pipe at 74 %
// with DRAM at 36 %). Same terms, same row walk.
template <int SRCCH, int CVT, int OUTCH>
__global__ void __launch_bounds__(128) packed_up_rows_x16_kernel(const __grid_constant__ PackedParams P) {
  const int row_bytes = P.dw * OUTCH;
  const int row_words = (row_bytes + 3) >> 2;
  const int full_words = row_bytes >> 2;
  const int w0 = (blockIdx.x * 32 + threadIdx.x) * 4;
  const int oy_begin = (blockIdx.y * blockDim.y + threadIdx.y) * P.up_rows;
  if (w0 >= row_words || oy_begin >= P.dh) return;
  const int oy_end = min(P.dh, oy_begin + P.up_rows);
  const FramePtrs f = frame_ptrs(P.batch, blockIdx.z);
  const int n_mine = min(4, row_words - w0);
  const int n_full = max(0, min(4, full_words - w0));

  uint32_t h1[4];
  UpTerms t[4];
  int cur_yi = INT_MIN, have1 = -1;
  int y_raw = P.y0 + oy_begin * P.dy;
  uint8_t* o = f.out + (long long)oy_begin * P.dst_row_stride + 4 * (size_t)w0;
  for (int oy = oy_begin; oy < oy_end; ++oy, o += P.dst_row_stride, y_raw += P.dy) {
    const int y = min(y_raw, P.max_y);
    const int yi = y >> 16;
    const uint32_t yf = P.linear ? 0u : (uint32_t)((y >> 8) & 255);
    if (yi != cur_yi) {
      const int r0 = min(max(yi, 0), P.sh - 1), r1 = min(max(yi + 1, 0), P.sh - 1);
      const uint8_t* row0 = f.p0 + (size_t)r0 * P.stride0;
      const uint8_t* row1 = f.p0 + (size_t)r1 * P.stride0;
      const bool reuse = (r0 == have1);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        if (j < n_mine) {
          const uint32_t a = reuse ? h1[j] : up_filter_word<SRCCH, CVT, OUTCH>(row0, w0 + j, P);
          const uint32_t b = (r1 == r0) ? a : up_filter_word<SRCCH, CVT, OUTCH>(row1, w0 + j, P);
          h1[j] = b;
          t[j] = up_terms(a, b);
        }
      }
      cur_yi = yi;
      have1 = r1;
    }
    uint32_t v[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) v[j] = __byte_perm(t[j].d_lo * yf + t[j].base_lo, t[j].d_hi * yf + t[j].base_hi, 0x7351);
    if (n_full == 4) {
      *reinterpret_cast<uint4*>(o) = make_uint4(v[0], v[1], v[2], v[3]);
    } else {  // the lane that holds the end of the row
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        if (j < n_full) reinterpret_cast<uint32_t*>(o)[j] = v[j];
        else if (j < n_mine) for (int k = 4 * (w0 + j); k < row_bytes; ++k) o[k - 4 * w0] = (uint8_t)(v[j] >> (8 * (k - 4 * (w0 + j))));
      }
    }
  }
}

// RGB outputs of SMALL-ratio up-scales (factor < 8): the row-streaming kernels above re-derive the column terms of a pixel (16.16
// position, two clamped columns, 7-bit weight) every time a source row is filtered, and an RGB output WORD needs two pixels, so the
// column filter -- run once per factor output rows -- dominates (640x480 -> 1280x960: 0.20 of the HBM roofline). Here a thread owns
// FOUR PIXELS (twelve output bytes) and keeps their column terms in registers for all its rows: filtering a source row costs
// two fetches and one packed blend per pixel. Rows are walked exactly as above (terms rebuilt when yi advances, the old bottom
// row becomes the new top row); the four blended pixel words are packed into three output words with PRMT.
template <int SRCCH, int CVT, int OUTCH>
__global__ void __launch_bounds__(128) packed_up_rows_px4_kernel(const __grid_constant__ PackedParams P) {
  constexpr bool PLANE = OUTCH == 1;
  constexpr int NT = PLANE ? 1 : 4;  // blended units per thread: gray packs its four pixels into ONE word
  const int ox0 = (blockIdx.x * 32 + threadIdx.x) * 4;
  const int oy_begin = (blockIdx.y * blockDim.y + threadIdx.y) * P.up_rows;
  if (ox0 >= P.dw || oy_begin >= P.dh) return;
  const int oy_end = min(P.dh, oy_begin + P.up_rows);
  const FramePtrs f = frame_ptrs(P.batch, blockIdx.z);
  int c0[4], c1[4], xw[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int ox = min(ox0 + k, P.dw - 1);  // pixels past the end of the row are never stored
    const int x = P.x0 + ox * P.dx;
    const int xi = x >> 16;
    c0[k] = min(max(xi, 0), P.sw - 1);
    c1[k] = min(max(xi + 1, 0), P.sw - 1);
    xw[k] = x;  // packed_hblend takes the 16.16 position (7-bit truncating ARGB filter / rounding plane filter)
  }
  // column filter of one source row for this thread's four pixels. Results stay in LANE form -- channels 0, 2 in the 16-bit
  // lanes of .x, channels 1, 3 in those of .y -- which is what the truncating blend produces and what the row terms consume:
  // packing them into bytes only for up_terms() to split them again cost ~8 of ~50 instructions per pixel and source row
  // (ncu: the px4 kernel is ALU-bound at small factors, profiles/r02_ncu_px4_nv21_up2x.txt).
  auto filter = [&](const uint8_t* __restrict__ row, uint2 (&out)[NT]) {
    if (PLANE) {
      uint32_t px[4];
#pragma unroll
      for (int k = 0; k < 4; ++k)
        px[k] = packed_hblend<true>(packed_fetch<SRCCH, CVT>(row, c0[k]), packed_fetch<SRCCH, CVT>(row, c1[k]), xw[k]);
      out[0] = make_uint2((px[0] & 0xFF) | ((px[2] & 0xFF) << 16), (px[1] & 0xFF) | ((px[3] & 0xFF) << 16));
    } else {
#pragma unroll
      for (int k = 0; k < NT; ++k) {
        const uint32_t a = packed_fetch<SRCCH, CVT>(row, c0[k]), b = packed_fetch<SRCCH, CVT>(row, c1[k]);
        const uint32_t f7 = (uint32_t)((xw[k] >> 9) & 0x7f), w0 = 0x7fu ^ f7;  // blend_cols4_trunc7 without its final pack
        const uint32_t lo = (a & 0x00FF00FFu) * w0 + (b & 0x00FF00FFu) * f7;
        const uint32_t hi = ((a >> 8) & 0x00FF00FFu) * w0 + ((b >> 8) & 0x00FF00FFu) * f7;
        out[k] = make_uint2((lo >> 7) & 0x00FF00FFu, (hi >> 7) & 0x00FF00FFu);
      }
    }
  };
  auto lane_terms = [](const uint2 a, const uint2 b) {
    UpTerms t;
    t.d_lo = b.x - a.x;
    t.d_hi = b.y - a.y;
    t.base_lo = (a.x << 8) + 0x00800080u;
    t.base_hi = (a.y << 8) + 0x00800080u;
    return t;
  };
  uint2 h1[NT];
  UpTerms t[NT];
  int cur_yi = INT_MIN, have1 = -1;
  int y_raw = P.y0 + oy_begin * P.dy;
  uint8_t* o = f.out + (long long)oy_begin * P.dst_row_stride + OUTCH * (size_t)ox0;
  const bool whole = ox0 + 4 <= P.dw && P.vec_ok;
  for (int oy = oy_begin; oy < oy_end; ++oy, o += P.dst_row_stride, y_raw += P.dy) {
    const int y = min(y_raw, P.max_y);
    const int yi = y >> 16;
    const uint32_t yf = P.linear ? 0u : (uint32_t)((y >> 8) & 255);
    if (yi != cur_yi) {
      const int r0 = min(max(yi, 0), P.sh - 1), r1 = min(max(yi + 1, 0), P.sh - 1);
      uint2 a[NT], b[NT];
      if (r0 == have1) {
#pragma unroll
        for (int k = 0; k < NT; ++k) a[k] = h1[k];
      } else {
        filter(f.p0 + (size_t)r0 * P.stride0, a);
      }
      if (r1 == r0) {
#pragma unroll
        for (int k = 0; k < NT; ++k) b[k] = a[k];
      } else {
        filter(f.p0 + (size_t)r1 * P.stride0, b);
      }
#pragma unroll
      for (int k = 0; k < NT; ++k) {
        h1[k] = b[k];
        t[k] = lane_terms(a[k], b[k]);
      }
      cur_yi = yi;
      have1 = r1;
    }
    uint32_t v[NT];
#pragma unroll
    for (int k = 0; k < NT; ++k) v[k] = __byte_perm(t[k].d_lo * yf + t[k].base_lo, t[k].d_hi * yf + t[k].base_hi, 0x7351);
    if (PLANE) {
      if (whole) *reinterpret_cast<uint32_t*>(o) = v[0];
      else for (int k = 0; k < 4 && ox0 + k < P.dw; ++k) o[k] = (uint8_t)(v[0] >> (8 * k));
    } else if (whole) {
      uint32_t* ow = reinterpret_cast<uint32_t*>(o);
      ow[0] = __byte_perm(v[0], v[NT > 1 ? 1 : 0], 0x4210);
      ow[1] = __byte_perm(v[NT > 1 ? 1 : 0], v[NT > 2 ? 2 : 0], 0x5421);
      ow[2] = __byte_perm(v[NT > 2 ? 2 : 0], v[NT > 3 ? 3 : 0], 0x6542);
    } else {
      for (int k = 0; k < NT && ox0 + k < P.dw; ++k) {
        o[3 * k] = (uint8_t)v[k]; o[3 * k + 1] = (uint8_t)(v[k] >> 8); o[3 * k + 2] = (uint8_t)(v[k] >> 16);
      }
    }
  }
}

typedef void (*PackedKernel)(PackedParams);
// Small sources (the 64^2 end of BASELINE's RGBA sweep): Scale*BilinearUp as libyuv does it -- every source row filtered
// horizontally ONCE, then output rows blended from pairs of filtered rows -- with the whole frame in shared memory.
// The row-streaming kernels above filter a source row pair per thread strip straight from global memory: for a 64x64
// source and a 2x - 4x up-scale that is a chain of dependent L2 round trips per thread and a handful of warps per frame
// (ncu-free arithmetic: 7.9 M frames/s on RGBA 64^2 -> gray 128^2 = 0.04 of the HBM roofline). Here ONE CTA owns one
// frame: (1) the source is staged with 128-bit loads, converted on the way (RGBA -> gray bytes / RGB words);
// (2) pass 1 writes hf[r][ox] = column filter of source row r at output column ox for all sh rows (gray: four pixels
// per thread packed into one word; RGB / RGBA: one pixel word each); (3) pass 2 blends hf[yi], hf[yi+1] per output row
// with the packed InterpolateRow arithmetic and stores whole words, lanes along the output row.
// Same clamps and blenders as packed_bilinear_kernel<.., kPackedUp, ..>; the parity tests compare with the oracle.
template <int SRCCH, int CVT, int OUTCH>
__global__ void __launch_bounds__(256) packed_small_up_kernel(const __grid_constant__ PackedParams P) {
  constexpr bool PLANE = (OUTCH == 1);
  constexpr int SB = PLANE ? 1 : 4;           // bytes per staged source element and per filtered element
  extern __shared__ __align__(16) uint8_t sm[];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;  // warps walk rows, lanes walk along a row: no divisions
  const FramePtrs f = frame_ptrs(P.batch, blockIdx.x);
  const int src_pitch = (P.sw * SB + 15) & ~15;        // staged source row pitch
  const int hf_pitch = (P.dw * SB + 3) & ~3;           // filtered row pitch (bytes)
  uint8_t* ssrc = sm;
  uint8_t* hf = sm + P.sh * src_pitch;
  // ---- stage + convert the source frame
  if (SRCCH == 4) {
    const int n4 = P.sw >> 2;  // groups of four pixels (sw % 4 == 0, 16-byte aligned rows: checked on the host)
    for (int r = wid; r < P.sh; r += 8)
    for (int g = lane; g < n4; g += 32) {
      const uint4 v = __ldg(reinterpret_cast<const uint4*>(f.p0 + (size_t)r * P.stride0) + g);
      if (PLANE) {
        const uint32_t y0 = rgb_to_yj(v.x & 0xFF, (v.x >> 8) & 0xFF, (v.x >> 16) & 0xFF), y1 = rgb_to_yj(v.y & 0xFF, (v.y >> 8) & 0xFF, (v.y >> 16) & 0xFF);
        const uint32_t y2 = rgb_to_yj(v.z & 0xFF, (v.z >> 8) & 0xFF, (v.z >> 16) & 0xFF), y3 = rgb_to_yj(v.w & 0xFF, (v.w >> 8) & 0xFF, (v.w >> 16) & 0xFF);
        *reinterpret_cast<uint32_t*>(ssrc + r * src_pitch + 4 * g) = y0 | (y1 << 8) | (y2 << 16) | (y3 << 24);
      } else {
        const uint32_t m = CVT == BANDCU_CVT_RGBA2RGB ? 0x00FFFFFFu : 0xFFFFFFFFu;
        *reinterpret_cast<uint4*>(ssrc + r * src_pitch + 16 * g) = make_uint4(v.x & m, v.y & m, v.z & m, v.w & m);
      }
    }
  } else {  // gray or RGB sources: element-wise through the read-only path
    for (int r = wid; r < P.sh; r += 8)
    for (int c = lane; c < P.sw; c += 32) {
      const uint32_t v = packed_fetch<SRCCH, CVT>(f.p0 + (size_t)r * P.stride0, c);
      if (PLANE) ssrc[r * src_pitch + c] = (uint8_t)v;
      else *reinterpret_cast<uint32_t*>(ssrc + r * src_pitch + 4 * c) = v;
    }
  }
  __syncthreads();
  // ---- pass 1: horizontal filter of every source row
  auto hpixel = [&](const uint8_t* row, int ox) {
    const int x = P.x0 + ox * P.dx;
    const int xi = x >> 16;
    const int c0 = min(max(xi, 0), P.sw - 1), c1 = min(max(xi + 1, 0), P.sw - 1);
    if (PLANE) return packed_hblend<true>((uint32_t)row[c0], (uint32_t)row[c1], x);
    return packed_hblend<false>(reinterpret_cast<const uint32_t*>(row)[c0], reinterpret_cast<const uint32_t*>(row)[c1], x);
  };
  if (PLANE) {
    const int wpr = hf_pitch >> 2;  // words per filtered row
    for (int r = wid; r < P.sh; r += 8)
    for (int w = lane; w < wpr; w += 32) {
      const uint8_t* row = ssrc + r * src_pitch;
      const int ox = 4 * w;
      uint32_t v = hpixel(row, min(ox, P.dw - 1)) & 0xFF;
      v |= (hpixel(row, min(ox + 1, P.dw - 1)) & 0xFF) << 8;
      v |= (hpixel(row, min(ox + 2, P.dw - 1)) & 0xFF) << 16;
      v |= hpixel(row, min(ox + 3, P.dw - 1)) << 24;
      *reinterpret_cast<uint32_t*>(hf + r * hf_pitch + 4 * w) = v;
    }
  } else {
    for (int r = wid; r < P.sh; r += 8)
      for (int ox = lane; ox < P.dw; ox += 32)
        *reinterpret_cast<uint32_t*>(hf + r * hf_pitch + 4 * ox) = hpixel(ssrc + r * src_pitch, ox);
  }
  __syncthreads();
  // ---- pass 2: vertical blend, whole output words, lanes along the row
  const int row_bytes = P.dw * OUTCH;
  const int row_words = (row_bytes + 3) >> 2, full_words = row_bytes >> 2;
  // work item = (output row, group of NG words): gray / RGBA one word, RGB three words = four pixels
  constexpr int NG = OUTCH == 3 ? 3 : 1;
  const int gpr = (row_words + NG - 1) / NG;
  for (int oy = wid; oy < P.dh; oy += 8) {
    const int y = min(P.y0 + oy * P.dy, P.max_y);
    const int yi = y >> 16;
    const int yf = P.linear ? 0 : ((y >> 8) & 255);
    const int r0 = min(max(yi, 0), P.sh - 1), r1 = min(max(yi + 1, 0), P.sh - 1);
    const uint32_t* h0 = reinterpret_cast<const uint32_t*>(hf + r0 * hf_pitch);
    const uint32_t* h1 = reinterpret_cast<const uint32_t*>(hf + r1 * hf_pitch);
    uint8_t* orow = f.out + (long long)oy * P.dst_row_stride;
   for (int g = lane; g < gpr; g += 32) {
    uint32_t v[NG];
    if (OUTCH == 3) {
      uint32_t px[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int ox = min(4 * g + k, P.dw - 1);
        px[k] = interp_rows4(h0[ox], h1[ox], yf);
      }
      v[0] = __byte_perm(px[0], px[1], 0x4210);
      v[1] = __byte_perm(px[1], px[2], 0x5421);
      v[2] = __byte_perm(px[2], px[3], 0x6542);
    } else {
      v[0] = interp_rows4(h0[g], h1[g], yf);
    }
#pragma unroll
    for (int q = 0; q < NG; ++q) {
      const int w = NG * g + q;
      if (w < full_words && P.vec_ok) reinterpret_cast<uint32_t*>(orow)[w] = v[q];
      else if (w < row_words) {  // rows that are not word-aligned, and the partial last word of a row: this word's bytes only
        const int kend = min(row_bytes, 4 * w + 4);
        for (int k = 4 * w; k < kend; ++k) orow[k] = (uint8_t)(v[q] >> (8 * (k - 4 * w)));
      }
    }
   }
  }
}
template <int SRCCH, int CVT, int OUTCH>
static PackedKernel packed_small_up_of() { return packed_small_up_kernel<SRCCH, CVT, OUTCH>; }
static PackedKernel packed_pick_small_up(const BandCuPlanDesc& d) {
  if (d.src_kind == BANDCU_SRC_GRAY && d.cvt == BANDCU_CVT_NONE) return packed_small_up_of<1, BANDCU_CVT_NONE, 1>();
  if (d.src_kind == BANDCU_SRC_RGB && d.cvt == BANDCU_CVT_NONE) return packed_small_up_of<3, BANDCU_CVT_NONE, 3>();
  if (d.src_kind == BANDCU_SRC_RGB && d.cvt == BANDCU_CVT_RGB2GRAY) return packed_small_up_of<3, BANDCU_CVT_RGB2GRAY, 1>();
  if (d.src_kind == BANDCU_SRC_RGBA && d.cvt == BANDCU_CVT_NONE) return packed_small_up_of<4, BANDCU_CVT_NONE, 4>();
  if (d.src_kind == BANDCU_SRC_RGBA && d.cvt == BANDCU_CVT_RGBA2GRAY) return packed_small_up_of<4, BANDCU_CVT_RGBA2GRAY, 1>();
  if (d.src_kind == BANDCU_SRC_RGBA && d.cvt == BANDCU_CVT_RGBA2RGB) return packed_small_up_of<4, BANDCU_CVT_RGBA2RGB, 3>();
  return nullptr;
}

struct PackedPlan {
  float* lut_dev;
  int up_rows_mode;  // BANDCU_PACKED_UP_ROWS when the plan was created: -1 default, 0 off, 2 word-interleaved form
  int small_up;      // BANDCU_PACKED_SMALL_UP != 0 when the plan was created: shared-memory kernel for small sources
};

__global__ void packed_build_lut_kernel(float* lut, float mean, float sd) {
  const int i = threadIdx.x;
  if (i < 256) lut[i] = normalize_elem<float>((float)i, mean, sd);
}

static bool packed_is_identity(const BandCuAffine& m) {
  return m.r0 == 0 && m.rr == 1 && m.rc == 0 && m.c0 == 0 && m.cr == 0 && m.cc == 1;
}

bool packed_plan_supported(const BandCuPlanDesc& d) {
  const char* off = getenv("BANDCU_DISABLE_PACKED");
  if (off && off[0] == '1') return false;
  if (d.src_kind != BANDCU_SRC_GRAY && d.src_kind != BANDCU_SRC_RGB && d.src_kind != BANDCU_SRC_RGBA) return false;
  if (!packed_is_identity(d.luma_map)) return false;
  // four ADJACENT output pixels per thread only pays while a warp's source footprint stays compact: up-scaling,
  // down-scaling by at most 2, and the exact 2x box; larger ratios stay on the one-pixel-per-thread kernel
  if (d.scale_mode == BANDCU_SCALE_BILINEAR_DOWN) { if (d.dx > 0x20000) return false; }
  else if (d.scale_mode == BANDCU_SCALE_BOX2) { if (d.dx != 0x20000) return false; }
  else if (d.scale_mode != BANDCU_SCALE_BILINEAR_UP) return false;
  if ((d.channels == 1) != (d.flavor == BANDCU_FLAVOR_PLANE)) return false;
  if (d.normalize ? d.out_dtype != BANDCU_DT_F32 : d.out_dtype != BANDCU_DT_U8) return false;
  if ((long long)d.x0 + (long long)(d.dst_width + 4) * d.dx > 0x7fffffffLL) return false;
  if ((long long)d.y0 + (long long)(d.dst_height + 1) * d.dy > 0x7fffffffLL) return false;
  return true;
}

int packed_plan_create(const BandCuPlanDesc& d, PackedPlan** out) {
  PackedPlan* p = new PackedPlan();
  p->lut_dev = nullptr;
  const char* ur = getenv("BANDCU_PACKED_UP_ROWS");  // read here, once per plan (never on the launch path)
  p->up_rows_mode = (ur && ur[0] >= '0' && ur[0] <= '9') ? ur[0] - '0' : -1;
  const char* su = getenv("BANDCU_PACKED_SMALL_UP");
  p->small_up = (su && su[0] == '0') ? 0 : 1;
  if (d.normalize) {
    cudaError_t e = cudaMalloc((void**)&p->lut_dev, 256 * sizeof(float));
    if (e != cudaSuccess) { delete p; return (int)e; }
    packed_build_lut_kernel<<<1, 256>>>(p->lut_dev, d.mean, d.std_);
    count_launch();
    e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { cudaFree(p->lut_dev); delete p; return (int)e; }
  }
  *out = p;
  return BANDCU_OK;
}

void packed_plan_destroy(PackedPlan* p) {
  if (!p) return;
  if (p->lut_dev) (void)cudaFree(p->lut_dev);
  delete p;
}

template <int SRCCH, int CVT, int OUTCH>
static PackedKernel packed_pick2(int mode, bool f32, int up_rows) {
  if (mode == kPackedUp) {
    const bool rows = up_rows != 0, narrow = up_rows == 1;
    constexpr int NW = OUTCH == 3 ? 3 : 4;
    if (up_rows == 3) return packed_up_rows_x16_kernel<SRCCH, CVT, OUTCH>;  // u8, 16-byte aligned rows
    if (up_rows == 4) {  // RGB outputs, factor < 8: four pixels per thread, column terms in registers
      if (OUTCH == 3 || OUTCH == 1) return packed_up_rows_px4_kernel<SRCCH, CVT, (OUTCH == 4 ? 3 : OUTCH)>;
      return nullptr;
    }
    if (rows && narrow) return f32 ? packed_up_rows_kernel<SRCCH, CVT, OUTCH, float, 1> : packed_up_rows_kernel<SRCCH, CVT, OUTCH, uint8_t, 1>;
    if (rows) return f32 ? packed_up_rows_kernel<SRCCH, CVT, OUTCH, float, NW> : packed_up_rows_kernel<SRCCH, CVT, OUTCH, uint8_t, NW>;
    return f32 ? packed_bilinear_kernel<SRCCH, CVT, OUTCH, kPackedUp, float> : packed_bilinear_kernel<SRCCH, CVT, OUTCH, kPackedUp, uint8_t>;
  }
  if (mode == kPackedBox2) return f32 ? packed_bilinear_kernel<SRCCH, CVT, OUTCH, kPackedBox2, float> : packed_bilinear_kernel<SRCCH, CVT, OUTCH, kPackedBox2, uint8_t>;
  return f32 ? packed_bilinear_kernel<SRCCH, CVT, OUTCH, kPackedDown, float> : packed_bilinear_kernel<SRCCH, CVT, OUTCH, kPackedDown, uint8_t>;
}
static PackedKernel packed_pick(const BandCuPlanDesc& d, int up_rows) {
  const int up = d.scale_mode == BANDCU_SCALE_BILINEAR_UP ? kPackedUp : d.scale_mode == BANDCU_SCALE_BOX2 ? kPackedBox2 : kPackedDown;
  const bool f32 = d.normalize != 0;
  if (d.src_kind == BANDCU_SRC_GRAY && d.cvt == BANDCU_CVT_NONE) return packed_pick2<1, BANDCU_CVT_NONE, 1>(up, f32, up_rows);
  if (d.src_kind == BANDCU_SRC_RGB && d.cvt == BANDCU_CVT_NONE) return packed_pick2<3, BANDCU_CVT_NONE, 3>(up, f32, up_rows);
  if (d.src_kind == BANDCU_SRC_RGB && d.cvt == BANDCU_CVT_RGB2GRAY) return packed_pick2<3, BANDCU_CVT_RGB2GRAY, 1>(up, f32, up_rows);
  if (d.src_kind == BANDCU_SRC_RGBA && d.cvt == BANDCU_CVT_NONE) return packed_pick2<4, BANDCU_CVT_NONE, 4>(up, f32, up_rows);
  if (d.src_kind == BANDCU_SRC_RGBA && d.cvt == BANDCU_CVT_RGBA2GRAY) return packed_pick2<4, BANDCU_CVT_RGBA2GRAY, 1>(up, f32, up_rows);
  if (d.src_kind == BANDCU_SRC_RGBA && d.cvt == BANDCU_CVT_RGBA2RGB) return packed_pick2<4, BANDCU_CVT_RGBA2RGB, 3>(up, f32, up_rows);
  return nullptr;
}

int launch_packed(PackedPlan* p, const BandCuPlanDesc& d, const BatchArgs& batch, const TableHints& hints, int n,
                  cudaStream_t stream) {
  bool in4, out16;
  if (batch.table) {
    if (!hints.valid) return BANDCU_E_UNSUPPORTED;
    in4 = hints.aligned16;
    out16 = hints.out_aligned16;
  } else {
    in4 = (reinterpret_cast<uintptr_t>(batch.first.p0) & 3) == 0 && (n == 1 || batch.in_stride % 4 == 0);
    out16 = (reinterpret_cast<uintptr_t>(batch.first.out) & 15) == 0 && (n == 1 || batch.out_stride % 16 == 0);
  }
  if (d.src_kind == BANDCU_SRC_RGBA && !(in4 && d.stride0 % 4 == 0)) return BANDCU_E_UNSUPPORTED;  // 32-bit pixel loads
  PackedParams P;
  memset(&P, 0, sizeof(P));
  P.sw = d.vw; P.sh = d.vh; P.stride0 = d.stride0;
  P.dw = d.dst_width; P.dh = d.dst_height;
  P.x0 = d.x0; P.y0 = d.y0; P.dx = d.dx; P.dy = d.dy; P.max_y = d.max_y; P.linear = d.linear;
  P.dst_row_stride = d.dst_row_stride;
  P.vec_ok = (out16 && d.dst_row_stride % 16 == 0) ? 1 : 0;
  P.lut = p->lut_dev;
  dim3 block(32, 8, 1);
  dim3 grid((d.dst_width + 127) / 128, (d.dst_height + 7) / 8, 1);
  // row-streaming up-scaler: needs word (u8) / 16-byte (f32) aligned output rows
  int up_rows = 0;
  if (d.scale_mode == BANDCU_SCALE_BILINEAR_UP && p->up_rows_mode != 0) {
    bool out4;
    if (batch.table) out4 = hints.out_aligned16;
    else out4 = (reinterpret_cast<uintptr_t>(batch.first.out) & 3) == 0 && (n == 1 || batch.out_stride % 4 == 0);
    const bool ok = d.normalize ? P.vec_ok != 0 : (out4 && d.dst_row_stride % 4 == 0);
    if (ok) {
      const int row_words = (d.dst_width * d.channels + 3) / 4;
      const int nwt = d.channels == 3 ? 3 : 4;
      up_rows = row_words <= 64 ? 1 : 2;
      // u8 rows of at least 1 KB that start 16-byte aligned: four consecutive words per lane, 128-bit stores
      if (!d.normalize && P.vec_ok && row_words >= 128 && p->up_rows_mode != 2) up_rows = 3;
      // RGB outputs of small-ratio up-scales: the column filter dominates, keep its terms in registers (four pixels per thread)
      // (measured: at factors >= 8 the 128-bit-store form wins for RGB and gray alike)
      if (!d.normalize && p->up_rows_mode < 0 && (d.channels == 3 || d.channels == 1) && d.dst_height < 8 * d.vh) {
        up_rows = 4;
        P.vec_ok = 1;  // word stores: `ok` above established 4-byte aligned rows
      }
      const int per_warp = up_rows == 3 ? 128 : up_rows == 4 ? 32 * d.channels : 32 * (up_rows == 1 ? 1 : nwt);
      // rows per thread: long enough to amortise the column filter of a source row pair (~4 source rows), short
      // enough to leave >= ~8 threads per lane of the machine
      const long long col_threads = (long long)n * ((row_words + per_warp - 1) / per_warp) * 32;
      int rpt = (int)((4LL * d.dst_height + d.vh - 1) / d.vh);
      rpt = rpt < 16 ? 16 : rpt > 256 ? 256 : rpt;
      // ... and at least 2 (gray) / 3 (RGB, RGBA) waves of resident threads (64 registers: 1024 threads per SM). The target used to
      // be 8 waves: on small batches of large outputs that cut a thread's strip to 16 rows, and the column filter of a source
      // row pair -- rebuilt at the top of every strip -- was amortised over half a source row's worth of output
      // (ncu: 8.9 instructions per output byte on RGBA 64^2 -> gray 2048^2; 0.52 -> 0.80 of the HBM roofline with 2 waves)
      const long long want_threads = 148LL * 1024 * (d.channels == 1 ? 2 : 3);
      while (rpt > 16 && col_threads * ((d.dst_height + rpt - 1) / rpt) < want_threads) rpt >>= 1;
      P.up_rows = rpt;
      block = dim3(32, 4, 1);  // each thread: its words of P.up_rows rows
      grid = dim3((row_words + per_warp - 1) / per_warp, (d.dst_height + 4 * rpt - 1) / (4 * rpt), 1);
    }
  }
  // small sources: the whole frame and its column-filtered rows fit shared memory -> one CTA per frame
  if (d.scale_mode == BANDCU_SCALE_BILINEAR_UP && !d.normalize && p->up_rows_mode != 0 && p->small_up != 0) {
    const int sb = d.channels == 1 ? 1 : 4;
    const size_t smem = (size_t)d.vh * ((d.vw * sb + 15) & ~15) + (size_t)d.vh * ((d.dst_width * sb + 3) & ~3);
    const bool src_ok = d.src_kind != BANDCU_SRC_RGBA ||
                        ((d.vw & 3) == 0 && d.stride0 % 16 == 0 &&
                         (batch.table ? hints.aligned16
                                      : ((reinterpret_cast<uintptr_t>(batch.first.p0) & 15) == 0 && (n == 1 || batch.in_stride % 16 == 0))));
    bool out4;
    if (batch.table) out4 = hints.out_aligned16;
    else out4 = (reinterpret_cast<uintptr_t>(batch.first.out) & 3) == 0 && (n == 1 || batch.out_stride % 4 == 0);
    PackedKernel small = packed_pick_small_up(d);
    // one CTA per frame only pays while a frame is a few tens of KB: measured +5 % on RGBA 64^2 -> RGB 128^2 and +2 % on
    // -> gray 256^2, but -27 % on -> RGB 256^2 (196 KB of output per CTA) and -38 % on -> gray 1024^2
    const size_t out_bytes = (size_t)d.dst_width * d.dst_height * d.channels;
    if (small && src_ok && smem <= 48 * 1024 && out_bytes <= 64 * 1024) {
      P.vec_ok = (out4 && d.dst_row_stride % 4 == 0) ? 1 : 0;
      cudaError_t e = cudaFuncSetAttribute(small, cudaFuncAttributeMaxDynamicSharedMemorySize, 48 * 1024);  // one value for every plan
      if (e != cudaSuccess) return (int)e;
      for (int done = 0; done < n;) {
        const int chunk = n - done > (1 << 30) ? (1 << 30) : n - done;
        P.batch = batch;
        if (P.batch.table) {
          P.batch.table += done;
        } else {
          P.batch.first.p0 += (long long)done * batch.in_stride;
          P.batch.first.out += (long long)done * batch.out_stride;
        }
        small<<<chunk, 256, smem, stream>>>(P);
        count_launch();
        done += chunk;
      }
      return (int)cudaGetLastError();
    }
  }
  PackedKernel kern = packed_pick(d, up_rows);
  if (!kern) return BANDCU_E_UNSUPPORTED;
  for (int done = 0; done < n;) {
    const int chunk = n - done > 65535 ? 65535 : n - done;
    P.batch = batch;
    if (P.batch.table) {
      P.batch.table += done;
    } else {
      P.batch.first.p0 += (long long)done * batch.in_stride;
      P.batch.first.out += (long long)done * batch.out_stride;
    }
    grid.z = chunk;
    kern<<<grid, block, 0, stream>>>(P);
    count_launch();
    done += chunk;
  }
  return (int)cudaGetLastError();
}

}  // namespace bandcu
