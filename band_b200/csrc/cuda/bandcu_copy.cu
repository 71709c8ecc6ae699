// band_b200/csrc/cuda/bandcu_copy.cu
//
// Same-size conversions: a ColorSpaceConvert with no Resize behind it (optionally behind a Crop), the most common
// single call into the reference's image operator (libyuv_image_operator.cc:1709-1727 -> NV12ToRAW / NV21ToRAW /
// I420ToRAW / ..ToARGB-family :92-300, ARGBToRAW / RAWToJ400 / ARGBToJ400-style packed conversions, and the plain
// luma copy of YUV -> gray). The pixel-per-thread kernel spends ~240 instructions and six byte-wide memory
// instructions per output pixel on these (measured: 1080p NV21 -> RGB at 0.77 TB/s, 12 % of the HBM copy rate); here
// a thread owns FOUR adjacent pixels of TWO rows (one chroma row), reads and writes 32-bit words only, converts pixel
// pairs in packed 16-bit lanes (the identities of bandcu_fast.cu) and assembles the output words with PRMT.
//
// Arithmetic is unchanged (BT.601 6-bit fixed point YuvPixel; RGBToYJ); the parity tests compare with the oracle.
#include <cstring>

#include "bandcu_device.cuh"
#include "bandcu_internal.h"

namespace bandcu {

struct CopyParams {
  BatchArgs batch;
  int W, H;                 // output (= virtual image) size
  int stride0, stride_uv;
  long long luma_off, chroma_off;  // crop origin folded into the plane pointers
  long long dst_row_stride;
  int out16;                // output rows are 16-byte aligned (RGBA: one 128-bit store per thread and row)
  float mean, std_;         // f32 variant: Normalize behind the conversion
};

// ---- packed LEFT|RIGHT pixel pairs: both lanes share one chroma sample ----
struct CChroma {
  uint32_t kr, kg, kb;
};
__device__ __forceinline__ CChroma c_chroma_terms(uint32_t u, uint32_t v) {
  const uint32_t u2 = u * 0x00010001u, v2 = v * 0x00010001u;
  CChroma c;
  c.kr = __vadd2(v2 * 102u, 0xC878C878u /* -14216 */);
  const uint32_t t = u2 * 25u + v2 * 52u;
  c.kg = __vadd2(~t, 0x21F921F9u /* 8696 + 1 */);
  c.kb = u2 << 7;
  return c;
}
__device__ __forceinline__ void c_luma_pair_to_rgb(uint32_t yL, uint32_t yR, const CChroma& c, uint32_t& R2, uint32_t& G2,
                                                   uint32_t& B2) {
  const uint32_t yy = __byte_perm(yL * 4882229u, yR * 4882229u, 0x7632);
  const uint32_t r = __viaddmin_s16x2_relu(yy, c.kr, 0x3FFF3FFFu);
  const uint32_t g = __viaddmin_s16x2_relu(yy, c.kg, 0x3FFF3FFFu);
  uint32_t b = __vmaxu2(__viaddmin_u16x2(yy, c.kb, 0x84878487u), 0x44884488u);
  b = __vadd2(b, 0xBB78BB78u);
  R2 = (r >> 6) & 0x00FF00FFu;
  G2 = (g >> 6) & 0x00FF00FFu;
  B2 = (b >> 6) & 0x00FF00FFu;
}

// Four pixels of one row (luma word y4 = Y0 Y1 Y2 Y3) with the chroma of pairs A (px 0,1) and B (px 2,3) -> o[].
template <int OUTCH>
__device__ __forceinline__ void c_row4(uint32_t y4, const CChroma& ca, const CChroma& cb, uint32_t (&o)[4]) {
  uint32_t ra, ga, ba, rb, gb, bb;
  c_luma_pair_to_rgb(y4 & 0xFF, (y4 >> 8) & 0xFF, ca, ra, ga, ba);
  c_luma_pair_to_rgb((y4 >> 16) & 0xFF, y4 >> 24, cb, rb, gb, bb);
  const uint32_t rga = __byte_perm(ra, ga, 0x6240), rgb = __byte_perm(rb, gb, 0x6240);  // R0 G0 R1 G1
  if (OUTCH == 3) {
    o[0] = __byte_perm(rga, ba, 0x2410);                             // R0 G0 B0 R1
    o[1] = __byte_perm(__byte_perm(rga, ba, 0x0063), rgb, 0x5410);   // G1 B1 R2 G2
    o[2] = __byte_perm(rgb, bb, 0x6324);                             // B2 R3 G3 B3
  } else {
    const uint32_t bfa = ba | 0xFF00FF00u, bfb = bb | 0xFF00FF00u;   // B0 FF B1 FF
    o[0] = __byte_perm(rga, bfa, 0x5410);
    o[1] = __byte_perm(rga, bfa, 0x7632);
    o[2] = __byte_perm(rgb, bfb, 0x5410);
    o[3] = __byte_perm(rgb, bfb, 0x7632);
  }
}

enum { kCopyUV = 0, kCopyVU = 1, kCopyPlanar = 2 };

// The two chroma samples of a four-pixel group: interleaved word `w` = (first, second) x 2, or planar 16-bit halves.
template <int CH>
__device__ __forceinline__ void c_group_chroma(uint32_t w, uint32_t uu, uint32_t vv, CChroma& ca, CChroma& cb) {
  if (CH == kCopyPlanar) {
    ca = c_chroma_terms(uu & 0xFF, vv & 0xFF);
    cb = c_chroma_terms((uu >> 8) & 0xFF, (vv >> 8) & 0xFF);
  } else {
    const uint32_t a0 = w & 0xFF, b0 = (w >> 8) & 0xFF, a1 = (w >> 16) & 0xFF, b1 = w >> 24;
    ca = CH == kCopyVU ? c_chroma_terms(b0, a0) : c_chroma_terms(a0, b0);
    cb = CH == kCopyVU ? c_chroma_terms(b1, a1) : c_chroma_terms(a1, b1);
  }
}
template <int OUTCH>
__device__ __forceinline__ void c_store4(uint8_t* op, const uint32_t (&o)[4], bool out16) {
  if (OUTCH == 4 && out16) {
    *reinterpret_cast<uint4*>(op) = make_uint4(o[0], o[1], o[2], o[3]);
  } else {
#pragma unroll
    for (int k = 0; k < OUTCH; ++k) reinterpret_cast<uint32_t*>(op)[k] = o[k];
  }
}

// NPX = pixels per thread and row: 4 (32-bit accesses) or 8 (64-bit loads, 64/128-bit stores; frames whose planes,
// strides and outputs are 8-byte aligned -- 16 for RGBA outputs -- and whose width is a multiple of 8).
template <int CH, int OUTCH, int NPX>
__global__ void __launch_bounds__(256) yuv_copy_rows_kernel(const __grid_constant__ CopyParams P) {
  const int x = (blockIdx.x * 32 + threadIdx.x) * NPX;
  const int y = (blockIdx.y * 8 + threadIdx.y) * 2;
  if (x >= P.W || y >= P.H) return;
  const FramePtrs f = frame_ptrs(P.batch, blockIdx.z);
  const uint8_t* yp = f.p0 + P.luma_off + (long long)y * P.stride0 + x;
  const bool two = y + 1 < P.H;
  const long long coff = P.chroma_off + (long long)(y >> 1) * P.stride_uv;
  uint8_t* op = f.out + (long long)y * P.dst_row_stride + (size_t)x * OUTCH;
  if (NPX == 4) {
    const uint32_t y0 = __ldg(reinterpret_cast<const uint32_t*>(yp));
    const uint32_t y1 = two ? __ldg(reinterpret_cast<const uint32_t*>(yp + P.stride0)) : y0;
    uint32_t w = 0, uu = 0, vv = 0;
    if (CH == kCopyPlanar) {
      uu = __ldg(reinterpret_cast<const uint16_t*>(f.pu + coff + (x >> 1)));
      vv = __ldg(reinterpret_cast<const uint16_t*>(f.pv + coff + (x >> 1)));
    } else {
      w = __ldg(reinterpret_cast<const uint32_t*>((CH == kCopyVU ? f.pv : f.pu) + coff + x));  // (first, second) pairs
    }
    CChroma ca, cb;
    c_group_chroma<CH>(w, uu, vv, ca, cb);
    uint32_t o[4];
    c_row4<OUTCH>(y0, ca, cb, o);
    c_store4<OUTCH>(op, o, P.out16 != 0);
    if (!two) return;
    c_row4<OUTCH>(y1, ca, cb, o);
    c_store4<OUTCH>(op + P.dst_row_stride, o, P.out16 != 0);
    return;
  }
  // eight pixels: two four-pixel groups
  const uint2 y0 = __ldg(reinterpret_cast<const uint2*>(yp));
  const uint2 y1 = two ? __ldg(reinterpret_cast<const uint2*>(yp + P.stride0)) : y0;
  uint2 w = make_uint2(0u, 0u);
  uint32_t uu = 0, vv = 0;
  if (CH == kCopyPlanar) {
    uu = __ldg(reinterpret_cast<const uint32_t*>(f.pu + coff + (x >> 1)));
    vv = __ldg(reinterpret_cast<const uint32_t*>(f.pv + coff + (x >> 1)));
  } else {
    w = __ldg(reinterpret_cast<const uint2*>((CH == kCopyVU ? f.pv : f.pu) + coff + x));
  }
  CChroma c0a, c0b, c1a, c1b;
  c_group_chroma<CH>(w.x, uu & 0xFFFFu, vv & 0xFFFFu, c0a, c0b);
  c_group_chroma<CH>(w.y, uu >> 16, vv >> 16, c1a, c1b);
#pragma unroll
  for (int row = 0; row < 2; ++row) {
    if (row == 1 && !two) break;
    const uint2 yy = row == 0 ? y0 : y1;
    uint32_t a[4], b[4];
    c_row4<OUTCH>(yy.x, c0a, c0b, a);
    c_row4<OUTCH>(yy.y, c1a, c1b, b);
    uint8_t* o = op + (long long)row * P.dst_row_stride;
    if (OUTCH == 4) {
      reinterpret_cast<uint4*>(o)[0] = make_uint4(a[0], a[1], a[2], a[3]);
      reinterpret_cast<uint4*>(o)[1] = make_uint4(b[0], b[1], b[2], b[3]);
    } else {
      reinterpret_cast<uint2*>(o)[0] = make_uint2(a[0], a[1]);
      reinterpret_cast<uint2*>(o)[1] = make_uint2(a[2], b[0]);
      reinterpret_cast<uint2*>(o)[2] = make_uint2(b[1], b[2]);
    }
  }
}

// ColorSpaceConvert(RGB) -> Normalize into float32 with NO Resize in between (a model fed at camera resolution; the
// reference runs NV12ToRAW and then NormalizeFromTo<uint8, float>, common_operator.cc:24-40). Same four-pixel, two-row
// threads as above; the twelve bytes of a row group go through a 256-entry table built by the CTA with the element
// arithmetic of the other kernels (IEEE subtract and divide) and leave as three 128-bit stores: a warp writes 1.5 KB
// of contiguous floats per row. The pixel-per-thread kernel was the only path for this chain before.
template <int CH>
__global__ void __launch_bounds__(256) yuv_copy_rows_f32_kernel(const __grid_constant__ CopyParams P) {
  __shared__ float lut[256];
  lut[threadIdx.y * 32 + threadIdx.x] = normalize_elem<float>((float)(threadIdx.y * 32 + threadIdx.x), P.mean, P.std_);
  __syncthreads();
  const int x = (blockIdx.x * 32 + threadIdx.x) * 4;
  const int y = (blockIdx.y * 8 + threadIdx.y) * 2;
  if (x >= P.W || y >= P.H) return;
  const FramePtrs f = frame_ptrs(P.batch, blockIdx.z);
  const uint8_t* yp = f.p0 + P.luma_off + (long long)y * P.stride0 + x;
  const bool two = y + 1 < P.H;
  const long long coff = P.chroma_off + (long long)(y >> 1) * P.stride_uv;
  const uint32_t y0 = __ldg(reinterpret_cast<const uint32_t*>(yp));
  const uint32_t y1 = two ? __ldg(reinterpret_cast<const uint32_t*>(yp + P.stride0)) : y0;
  uint32_t w = 0, uu = 0, vv = 0;
  if (CH == kCopyPlanar) {
    uu = __ldg(reinterpret_cast<const uint16_t*>(f.pu + coff + (x >> 1)));
    vv = __ldg(reinterpret_cast<const uint16_t*>(f.pv + coff + (x >> 1)));
  } else {
    w = __ldg(reinterpret_cast<const uint32_t*>((CH == kCopyVU ? f.pv : f.pu) + coff + x));
  }
  CChroma ca, cb;
  c_group_chroma<CH>(w, uu, vv, ca, cb);
  float4* op = reinterpret_cast<float4*>(f.out + (long long)y * P.dst_row_stride) + 3 * (x >> 2);
#pragma unroll
  for (int row = 0; row < 2; ++row) {
    if (row == 1 && !two) break;
    uint32_t o[4];
    c_row4<3>(row == 0 ? y0 : y1, ca, cb, o);
    float4* q = reinterpret_cast<float4*>(reinterpret_cast<uint8_t*>(op) + (long long)row * P.dst_row_stride);
#pragma unroll
    for (int k = 0; k < 3; ++k)
      q[k] = make_float4(lut[o[k] & 0xFF], lut[(o[k] >> 8) & 0xFF], lut[(o[k] >> 16) & 0xFF], lut[o[k] >> 24]);
  }
}

// Packed sources, four pixels per thread: RGBA -> RGB (drop alpha), RGBA -> gray, RGB -> gray (RGBToYJ), and the
// luma plane of a 4:2:0 frame -> gray (a strided copy). SRCB / DSTB = bytes per pixel.
enum { kPkRgbaToRgb = 0, kPkRgbaToGray = 1, kPkRgbToGray = 2, kPkLumaCopy = 3 };
// One four-pixel group: SW source words -> DW output words.
template <int MODE>
__device__ __forceinline__ void pk_group(const uint32_t* p, uint32_t* o) {
  if (MODE == kPkLumaCopy) {
    o[0] = p[0];
  } else if (MODE == kPkRgbToGray) {
    const uint32_t a = p[0], b = p[1], c = p[2];  // R0G0B0R1 G1B1R2G2 B2R3G3B3
    const uint32_t g0 = rgb_to_yj(a & 0xFF, (a >> 8) & 0xFF, (a >> 16) & 0xFF);
    const uint32_t g1 = rgb_to_yj(a >> 24, b & 0xFF, (b >> 8) & 0xFF);
    const uint32_t g2 = rgb_to_yj((b >> 16) & 0xFF, b >> 24, c & 0xFF);
    const uint32_t g3 = rgb_to_yj((c >> 8) & 0xFF, (c >> 16) & 0xFF, c >> 24);
    o[0] = g0 | (g1 << 8) | (g2 << 16) | (g3 << 24);
  } else if (MODE == kPkRgbaToGray) {
    uint32_t g[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) g[k] = rgb_to_yj(p[k] & 0xFF, (p[k] >> 8) & 0xFF, (p[k] >> 16) & 0xFF);
    o[0] = g[0] | (g[1] << 8) | (g[2] << 16) | (g[3] << 24);
  } else {
    o[0] = __byte_perm(p[0], p[1], 0x4210);  // R0 G0 B0 R1
    o[1] = __byte_perm(p[1], p[2], 0x5421);  // G1 B1 R2 G2
    o[2] = __byte_perm(p[2], p[3], 0x6542);  // B2 R3 G3 B3
  }
}
// NPX = 4: 32-bit accesses. NPX = 16: four groups per thread, 128-bit loads and stores (16-byte granular frames).
template <int MODE, int NPX>
__global__ void __launch_bounds__(256) packed_copy_rows_kernel(const __grid_constant__ CopyParams P) {
  const int x = (blockIdx.x * 32 + threadIdx.x) * NPX;
  const int y = blockIdx.y * 8 + threadIdx.y;
  if (x >= P.W || y >= P.H) return;
  const FramePtrs f = frame_ptrs(P.batch, blockIdx.z);
  constexpr int SRCB = MODE == kPkRgbToGray ? 3 : MODE == kPkLumaCopy ? 1 : 4;
  constexpr int DSTB = MODE == kPkRgbaToRgb ? 3 : 1;
  constexpr int SW = SRCB, DW = DSTB, G = NPX / 4;  // words per four-pixel group in / out
  const uint8_t* sp = f.p0 + P.luma_off + (long long)y * P.stride0 + (size_t)x * SRCB;
  uint8_t* op = f.out + (long long)y * P.dst_row_stride + (size_t)x * DSTB;
  uint32_t in[SW * G], out[DW * G];
  if (G == 4) {
#pragma unroll
    for (int k = 0; k < SW; ++k) {
      const uint4 q = __ldg(reinterpret_cast<const uint4*>(sp) + k);
      in[4 * k] = q.x; in[4 * k + 1] = q.y; in[4 * k + 2] = q.z; in[4 * k + 3] = q.w;
    }
  } else {
#pragma unroll
    for (int k = 0; k < SW; ++k) in[k] = __ldg(reinterpret_cast<const uint32_t*>(sp) + k);
  }
#pragma unroll
  for (int g = 0; g < G; ++g) pk_group<MODE>(in + g * SW, out + g * DW);
  if (G == 4) {
#pragma unroll
    for (int k = 0; k < DW; ++k)
      reinterpret_cast<uint4*>(op)[k] = make_uint4(out[4 * k], out[4 * k + 1], out[4 * k + 2], out[4 * k + 3]);
  } else {
#pragma unroll
    for (int k = 0; k < DW; ++k) reinterpret_cast<uint32_t*>(op)[k] = out[k];
  }
}

static bool copy_is_translation(const BandCuAffine& m) {
  return m.rr == 1 && m.rc == 0 && m.cr == 0 && m.cc == 1 && m.r0 >= 0 && m.c0 >= 0;
}

typedef void (*CopyKernel)(const CopyParams);

// BANDCU_E_UNSUPPORTED: not a same-size conversion this file covers, or the batch is not word-aligned.
int launch_copy_rows(const BandCuPlanDesc& d, const BatchArgs& batch, const TableHints& hints, int n, cudaStream_t stream) {
  static const bool off = getenv("BANDCU_DISABLE_COPY_ROWS") && getenv("BANDCU_DISABLE_COPY_ROWS")[0] == '1';
  if (off) return BANDCU_E_UNSUPPORTED;
  if (d.scale_mode != BANDCU_SCALE_COPY) return BANDCU_E_UNSUPPORTED;
  // Normalize behind the conversion: only 4:2:0 -> RGB -> float32 has a row kernel
  const bool f32 = d.normalize != 0;
  if (f32 && !(d.out_dtype == BANDCU_DT_F32 && d.src_kind == BANDCU_SRC_YUV420 && d.cvt == BANDCU_CVT_YUV2RGB))
    return BANDCU_E_UNSUPPORTED;
  if (d.dst_width != d.vw || d.dst_height != d.vh || (d.vw & 3)) return BANDCU_E_UNSUPPORTED;
  if (!copy_is_translation(d.luma_map) || !copy_is_translation(d.pre_map) || !copy_is_translation(d.chroma_map))
    return BANDCU_E_UNSUPPORTED;
  CopyParams P;
  memset(&P, 0, sizeof(P));
  P.batch = batch;
  P.W = d.vw; P.H = d.vh;
  P.stride0 = d.stride0; P.stride_uv = d.stride_uv;
  P.dst_row_stride = d.dst_row_stride;
  P.mean = d.mean;
  P.std_ = d.std_;
  CopyKernel kern = nullptr;
  int src_bpp = 1, rows_per_thread = 1, px_per_thread = 4, pk_mode = -1;
  bool need_chroma = false;
  if (d.src_kind == BANDCU_SRC_YUV420 && (d.cvt == BANDCU_CVT_YUV2RGB || d.cvt == BANDCU_CVT_YUV2RGBA)) {
    if ((d.pre_map.r0 & 1) || (d.pre_map.c0 & 1)) return BANDCU_E_UNSUPPORTED;
    // a pixel pair (2k, 2k+1) and a row pair must share their chroma sample: even luma origin inside the converted image
    // (a crop BEFORE the conversion re-bases chroma at origin/2, so any origin keeps the pairing)
    if (d.pixel_stride_uv != 1 && d.pixel_stride_uv != 2) return BANDCU_E_UNSUPPORTED;
    need_chroma = true;
    rows_per_thread = 2;
  } else if (d.src_kind == BANDCU_SRC_YUV420 && d.cvt == BANDCU_CVT_YUV2GRAY) {
    pk_mode = kPkLumaCopy;
  } else if (d.src_kind == BANDCU_SRC_RGBA && d.cvt == BANDCU_CVT_RGBA2RGB) {
    pk_mode = kPkRgbaToRgb; src_bpp = 4;
  } else if (d.src_kind == BANDCU_SRC_RGBA && d.cvt == BANDCU_CVT_RGBA2GRAY) {
    pk_mode = kPkRgbaToGray; src_bpp = 4;
  } else if (d.src_kind == BANDCU_SRC_RGB && d.cvt == BANDCU_CVT_RGB2GRAY) {
    pk_mode = kPkRgbToGray; src_bpp = 3;
  } else {
    return BANDCU_E_UNSUPPORTED;
  }
  P.luma_off = (long long)d.luma_map.r0 * d.stride0 + (long long)d.luma_map.c0 * src_bpp;
  P.chroma_off = (long long)(d.chroma_map.r0 + d.pre_map.r0 / 2) * d.stride_uv +
                 (long long)(d.chroma_map.c0 + d.pre_map.c0 / 2) * d.pixel_stride_uv;
  // 32-bit accesses everywhere: plane pointers, strides, crop offsets and outputs must be word aligned
  if ((d.stride0 & 3) || (P.luma_off & 3) || (d.dst_row_stride & 3)) return BANDCU_E_UNSUPPORTED;
  bool in4, out4, out16;
  if (batch.table) {
    if (!hints.valid) return BANDCU_E_UNSUPPORTED;
    in4 = hints.planes_aligned4;
    out4 = out16 = hints.out_aligned16;
  } else {
    uintptr_t bits = reinterpret_cast<uintptr_t>(batch.first.p0) | (n == 1 ? 0 : (uintptr_t)batch.in_stride);
    if (need_chroma) bits |= reinterpret_cast<uintptr_t>(batch.first.pu) | reinterpret_cast<uintptr_t>(batch.first.pv);
    const uintptr_t obits = reinterpret_cast<uintptr_t>(batch.first.out) | (n == 1 ? 0 : (uintptr_t)batch.out_stride);
    in4 = (bits & 3) == 0;
    out4 = (obits & 3) == 0;
    out16 = (obits & 15) == 0;
  }
  if (need_chroma) {
    const bool planar = d.pixel_stride_uv == 1;
    // planar: 16-bit chroma loads at byte offset x/2 (x % 4 == 0); interleaved: 32-bit loads at byte offset x
    if ((d.stride_uv & (planar ? 1 : 3)) || (P.chroma_off & (planar ? 1 : 3))) return BANDCU_E_UNSUPPORTED;
    int ch;
    if (planar) {
      ch = kCopyPlanar;
    } else if (batch.table) {
      if (!hints.semi_planar) return BANDCU_E_UNSUPPORTED;
      ch = hints.vu ? kCopyVU : kCopyUV;
    } else if (batch.first.pv == batch.first.pu + 1) {
      ch = kCopyUV;
    } else if (batch.first.pu == batch.first.pv + 1) {
      ch = kCopyVU;
    } else {
      return BANDCU_E_UNSUPPORTED;
    }
    // interleaved: the pair base (the lower of pu / pv) must be word aligned; the other pointer is odd by construction
    if (!planar && !batch.table) {
      const uint8_t* base = ch == kCopyVU ? batch.first.pv : batch.first.pu;
      in4 = ((reinterpret_cast<uintptr_t>(batch.first.p0) | reinterpret_cast<uintptr_t>(base) |
              (n == 1 ? 0 : (uintptr_t)batch.in_stride)) & 3) == 0;
    } else if (!planar) {
      in4 = hints.aligned16;
    }
    const bool rgba = d.cvt == BANDCU_CVT_YUV2RGBA;
    // eight pixels per thread when everything is 8-byte granular (planar chroma: 4-byte loads at x/2; RGBA: 16-byte stores)
    if (f32) {
      if (!out16 || (d.dst_row_stride & 15)) return BANDCU_E_UNSUPPORTED;  // 128-bit stores
    } else if ((d.vw & 7) == 0 && !batch.table && in4 && out4) {
      uintptr_t ib = reinterpret_cast<uintptr_t>(batch.first.p0) | (uintptr_t)d.stride0 | (uintptr_t)P.luma_off |
                     (n == 1 ? 0 : (uintptr_t)batch.in_stride);
      if (planar) {
        const uintptr_t cb = reinterpret_cast<uintptr_t>(batch.first.pu) | reinterpret_cast<uintptr_t>(batch.first.pv) |
                             (uintptr_t)d.stride_uv | (uintptr_t)P.chroma_off;
        if (cb & 3) ib |= 7;
      } else {
        ib |= reinterpret_cast<uintptr_t>(ch == kCopyVU ? batch.first.pv : batch.first.pu) | (uintptr_t)d.stride_uv |
              (uintptr_t)P.chroma_off;
      }
      const uintptr_t ob = reinterpret_cast<uintptr_t>(batch.first.out) | (uintptr_t)d.dst_row_stride |
                           (n == 1 ? 0 : (uintptr_t)batch.out_stride);
      if ((ib & 7) == 0 && (ob & (rgba ? 15 : 7)) == 0) px_per_thread = 8;
    }
    const bool p8 = px_per_thread == 8;
#define BANDCU_COPY_PICK(C) \
    (rgba ? (p8 ? yuv_copy_rows_kernel<C, 4, 8> : yuv_copy_rows_kernel<C, 4, 4>) \
          : (p8 ? yuv_copy_rows_kernel<C, 3, 8> : yuv_copy_rows_kernel<C, 3, 4>))
    if (f32) kern = ch == kCopyUV ? yuv_copy_rows_f32_kernel<kCopyUV> : ch == kCopyVU ? yuv_copy_rows_f32_kernel<kCopyVU>
                                                                                      : yuv_copy_rows_f32_kernel<kCopyPlanar>;
    else if (ch == kCopyUV) kern = BANDCU_COPY_PICK(kCopyUV);
    else if (ch == kCopyVU) kern = BANDCU_COPY_PICK(kCopyVU);
    else kern = BANDCU_COPY_PICK(kCopyPlanar);
#undef BANDCU_COPY_PICK
  }
  if (pk_mode >= 0) {
    // sixteen pixels per thread when rows, planes and outputs are 16-byte granular
    if ((d.vw & 15) == 0 && !batch.table) {
      const uintptr_t bits = reinterpret_cast<uintptr_t>(batch.first.p0) | (uintptr_t)d.stride0 | (uintptr_t)P.luma_off |
                             reinterpret_cast<uintptr_t>(batch.first.out) | (uintptr_t)d.dst_row_stride |
                             (n == 1 ? 0 : ((uintptr_t)batch.in_stride | (uintptr_t)batch.out_stride));
      if ((bits & 15) == 0) px_per_thread = 16;
    }
    const bool w16 = px_per_thread == 16;
    switch (pk_mode) {
      case kPkLumaCopy: kern = w16 ? packed_copy_rows_kernel<kPkLumaCopy, 16> : packed_copy_rows_kernel<kPkLumaCopy, 4>; break;
      case kPkRgbaToRgb: kern = w16 ? packed_copy_rows_kernel<kPkRgbaToRgb, 16> : packed_copy_rows_kernel<kPkRgbaToRgb, 4>; break;
      case kPkRgbaToGray: kern = w16 ? packed_copy_rows_kernel<kPkRgbaToGray, 16> : packed_copy_rows_kernel<kPkRgbaToGray, 4>; break;
      default: kern = w16 ? packed_copy_rows_kernel<kPkRgbToGray, 16> : packed_copy_rows_kernel<kPkRgbToGray, 4>; break;
    }
  }
  if (!in4 || !out4) return BANDCU_E_UNSUPPORTED;
  P.out16 = (out16 && (d.dst_row_stride & 15) == 0) ? 1 : 0;
  const dim3 block(32, 8, 1);
  const int rows = (d.vh + rows_per_thread - 1) / rows_per_thread;
  dim3 grid((d.vw / px_per_thread + 31) / 32, (rows + 7) / 8, 1);
  for (int done = 0; done < n;) {
    const int chunk = n - done > 65535 ? 65535 : n - done;
    CopyParams Q = P;
    if (Q.batch.table) {
      Q.batch.table += done;
    } else {
      const long long o = (long long)done * Q.batch.in_stride;
      Q.batch.first.p0 += o;
      if (Q.batch.first.pu) Q.batch.first.pu += o;
      if (Q.batch.first.pv) Q.batch.first.pv += o;
      Q.batch.first.out += (long long)done * Q.batch.out_stride;
    }
    grid.z = chunk;
    kern<<<grid, block, 0, stream>>>(Q);
    count_launch();
    done += chunk;
  }
  return (int)cudaGetLastError();
}

}  // namespace bandcu
