// band_b200/csrc/cuda/bandcu_device.cuh
//
// Device-side building blocks shared by the generic and the staged kernels:
// the integer arithmetic of the reference's libyuv CPU path, evaluated lazily
// per output element. Each helper names the libyuv routine whose result it
// must equal bit for bit (libyuv is an un-vendored dependency of the reference:
// band/workspace_repo.bzl:60-66; call sites in
// band/buffer/libyuv_image_operator.cc are cited at the kernels).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "bandcu.h"

namespace bandcu {

struct FramePtrs {
  const uint8_t* p0;
  const uint8_t* pu;
  const uint8_t* pv;
  uint8_t* out;
};

struct BatchArgs {
  FramePtrs first;          // strided mode
  long long in_stride;      // bytes between consecutive frames' planes
  long long out_stride;     // bytes between consecutive outputs
  const BandCuFrame* table; // non-null: per-frame pointer table (device memory)
};

__device__ __forceinline__ FramePtrs frame_ptrs(const BatchArgs& b, int f) {
  FramePtrs p;
  if (b.table) {
    BandCuFrame t = b.table[f];
    p.p0 = (const uint8_t*)t.p0;
    p.pu = (const uint8_t*)t.pu;
    p.pv = (const uint8_t*)t.pv;
    p.out = (uint8_t*)t.out;
  } else {
    long long o = (long long)f * b.in_stride;
    p.p0 = b.first.p0 + o;
    p.pu = b.first.pu ? b.first.pu + o : nullptr;
    p.pv = b.first.pv ? b.first.pv + o : nullptr;
    p.out = b.first.out + (long long)f * b.out_stride;
  }
  return p;
}

// ---- libyuv row_common.cc YuvPixel, kYuvI601Constants (x86): 6-bit fixed point
//   y1 = ((Y*0x0101*18997) >> 16) - 1160
//   B = clamp((y1 + 128*(U-128)) >> 6)            = clamp((yy + 128U         - 17544) >> 6)
//   G = clamp((y1 - 25*(U-128) - 52*(V-128)) >> 6) = clamp((yy - 25U - 52V    +  8696) >> 6)
//   R = clamp((y1 + 102*(V-128)) >> 6)            = clamp((yy + 102V         - 14216) >> 6)
// with yy = (Y * 4882229) >> 16. Returned packed R | G<<8 | B<<16.
__device__ __forceinline__ int yuv_luma_term(int y) { return (int)(((unsigned)y * 4882229u) >> 16); }
struct ChromaTerms {
  int cb, cg, cr;
};
__device__ __forceinline__ ChromaTerms yuv_chroma_terms(int u, int v) {
  ChromaTerms t;
  t.cb = 128 * u - 17544;
  t.cg = 8696 - 25 * u - 52 * v;
  t.cr = 102 * v - 14216;
  return t;
}
// clamp((a) >> 6) to [0,255] == min(max(a,0), 16383) >> 6
__device__ __forceinline__ int clamp_shift6(int a) { return min(max(a, 0), 16383) >> 6; }
__device__ __forceinline__ uint32_t yuv_to_rgb_packed(int yy, const ChromaTerms& t) {
  uint32_t r = (uint32_t)clamp_shift6(yy + t.cr);
  uint32_t g = (uint32_t)clamp_shift6(yy + t.cg);
  uint32_t b = (uint32_t)clamp_shift6(yy + t.cb);
  return r | (g << 8) | (b << 16);
}

// ---- libyuv row_common.cc RGBToYJ (8-bit form): (77R + 150G + 29B + 128) >> 8
__device__ __forceinline__ uint32_t rgb_to_yj(uint32_t r, uint32_t g, uint32_t b) {
  return (77u * r + 150u * g + 29u * b + 128u) >> 8;
}

// ---- SWAR helpers on 4 packed u8 channels ------------------------------------
// InterpolateRow_C: (a*(256-f) + b*f + 128) >> 8 per channel. Each 16-bit lane
// stays <= 255*256+128 = 65408, so two channels share one 32-bit multiply.
__device__ __forceinline__ uint32_t interp_rows4(uint32_t a, uint32_t b, int f) {
  if (f == 0) return a;
  uint32_t w0 = 256u - (uint32_t)f, w1 = (uint32_t)f;
  uint32_t lo = (a & 0x00FF00FFu) * w0 + (b & 0x00FF00FFu) * w1 + 0x00800080u;
  uint32_t hi = ((a >> 8) & 0x00FF00FFu) * w0 + ((b >> 8) & 0x00FF00FFu) * w1 + 0x00800080u;
  return ((lo >> 8) & 0x00FF00FFu) | (hi & 0xFF00FF00u);
}
// ScaleARGBFilterCols_C / ScaleUVFilterCols_C: (a*(0x7f^f) + b*f) >> 7, f = 7 bits,
// NO rounding and weights that sum to 127. Lanes stay <= 255*127 = 32385.
__device__ __forceinline__ uint32_t blend_cols4_trunc7(uint32_t a, uint32_t b, int f) {
  uint32_t w0 = (uint32_t)(0x7f ^ f), w1 = (uint32_t)f;
  uint32_t lo = (a & 0x00FF00FFu) * w0 + (b & 0x00FF00FFu) * w1;
  uint32_t hi = ((a >> 8) & 0x00FF00FFu) * w0 + ((b >> 8) & 0x00FF00FFu) * w1;
  return ((lo >> 7) & 0x00FF00FFu) | ((hi << 1) & 0xFF00FF00u);
}
// ScaleFilterCols_C, x86 BLENDER: a + (((f>>9)*(b-a) + 0x40) >> 7), single channel.
__device__ __forceinline__ uint32_t blend_cols1_round7(uint32_t a, uint32_t b, int f16) {
  int fa = f16 >> 9;
  return (uint32_t)((int)a + ((fa * ((int)b - (int)a) + 0x40) >> 7));
}
// ScaleARGBRowDown2Box_C / ScaleARGBRowDownEvenBox_C / ScaleRowDown2Box_C.
__device__ __forceinline__ uint32_t box2x2_4(uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  uint32_t lo = (a & 0x00FF00FFu) + (b & 0x00FF00FFu) + (c & 0x00FF00FFu) + (d & 0x00FF00FFu) + 0x00020002u;
  uint32_t hi = ((a >> 8) & 0x00FF00FFu) + ((b >> 8) & 0x00FF00FFu) + ((c >> 8) & 0x00FF00FFu) +
                ((d >> 8) & 0x00FF00FFu) + 0x00020002u;
  return ((lo >> 2) & 0x00FF00FFu) | ((hi << 6) & 0xFF00FF00u);
}

// Normalize element (common_operator.cc:35): out = static_cast<Out>((in - mean)) / std
// -- the cast binds to the difference, the division runs in float, the quotient
// converts to Out on assignment. IEEE round-to-nearest sub and div.
template <typename OT>
__device__ __forceinline__ OT normalize_elem(float in, float mean, float sd) {
  float diff = __fsub_rn(in, mean);
  OT t = (OT)diff;
  return (OT)__fdiv_rn((float)t, sd);
}
template <>
__device__ __forceinline__ float normalize_elem<float>(float in, float mean, float sd) {
  return __fdiv_rn(__fsub_rn(in, mean), sd);
}

}  // namespace bandcu
