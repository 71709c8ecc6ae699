// band_b200/csrc/cuda/bandcu_generic.cu
//
// Generic fused kernel: one thread per output pixel, every scaler code path,
// every source layout and any Crop/Rotate/Flip index map. It is the
// always-correct path (and the one C3's rotated 4K YV12 chain runs on); the
// staged fast kernel in bandcu_fast.cu overrides it for the NV12/NV21 ->
// bilinear-down hot chain.
//
// Replaces, fused into one pass: LibyuvImageOperator::ColorSpaceConvert
// (libyuv_image_operator.cc:1709-1727), ::Resize (:1627-1646), ::Crop
// (:1609-1625), ::Rotate (:1648-1667), ::Flip* (:1669-1707) with a packed
// output, and Normalize/DataTypeConvert (common_operator.cc:24-175).
#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "bandcu_device.cuh"
#include "bandcu_internal.h"

namespace bandcu {

struct GenericParams {
  BandCuPlanDesc d;
  BatchArgs batch;
  int transposed;  // 1: threadIdx.x walks output ROWS (the index map swaps axes: consecutive output rows are
                   // consecutive source columns, so this keeps the warp's source reads inside a few sectors)
};

__device__ __forceinline__ int map_row(const BandCuAffine& m, int r, int c) { return m.r0 + m.rr * r + m.rc * c; }
__device__ __forceinline__ int map_col(const BandCuAffine& m, int r, int c) { return m.c0 + m.cr * r + m.cc * c; }

// Fetch one pixel of the virtual image (after index maps and colour
// conversion) as packed channels c0 | c1<<8 | c2<<16 | c3<<24.
// SRC / CVT: compile-time source kind and conversion (-1 = read them from the descriptor); IDENT = 1 when all
// three index maps are the identity. launch_generic() picks the instantiation; the arithmetic is the same.
template <int SRC, int CVT, int IDENT>
__device__ __forceinline__ uint32_t fetch_px(const BandCuPlanDesc& d, const FramePtrs& f, int r, int c) {
  r = min(max(r, 0), d.vh - 1);
  c = min(max(c, 0), d.vw - 1);
  const int row = IDENT ? r : map_row(d.luma_map, r, c);
  const int col = IDENT ? c : map_col(d.luma_map, r, c);
  const uint8_t* p = f.p0 + (long long)row * d.stride0;
  const int src_kind = SRC < 0 ? d.src_kind : SRC;
  const int cvt = CVT < 0 ? d.cvt : CVT;
  switch (src_kind) {
    case BANDCU_SRC_GRAY:
      return p[col];
    case BANDCU_SRC_UV: {
      const uint8_t* q = p + col * 2;
      return (uint32_t)q[0] | ((uint32_t)q[1] << 8);
    }
    case BANDCU_SRC_RGB: {
      const uint8_t* q = p + col * 3;
      uint32_t R = q[0], G = q[1], B = q[2];
      if (cvt == BANDCU_CVT_RGB2GRAY) return rgb_to_yj(R, G, B);
      return R | (G << 8) | (B << 16);
    }
    case BANDCU_SRC_RGBA: {
      const uint8_t* q = p + col * 4;
      uint32_t R, G, B, A;
      if ((reinterpret_cast<uintptr_t>(q) & 3) == 0) {  // one 32-bit load (always, for 4-byte aligned frames)
        const uint32_t w = *reinterpret_cast<const uint32_t*>(q);
        if (cvt == BANDCU_CVT_NONE) return w;
        if (cvt == BANDCU_CVT_RGBA2RGB) return w & 0x00FFFFFFu;
        R = w & 0xFF; G = (w >> 8) & 0xFF; B = (w >> 16) & 0xFF; A = w >> 24;
      } else {
        R = q[0]; G = q[1]; B = q[2]; A = q[3];
      }
      if (cvt == BANDCU_CVT_RGBA2GRAY) return rgb_to_yj(R, G, B);
      if (cvt == BANDCU_CVT_RGBA2RGB) return R | (G << 8) | (B << 16);
      return R | (G << 8) | (B << 16) | (A << 24);
    }
    default: {  // BANDCU_SRC_YUV420
      const int Y = p[col];
      if (cvt == BANDCU_CVT_YUV2GRAY) return (uint32_t)Y;
      // chroma is NEAREST: sample (r1>>1, c1>>1) of the image the conversion ran on
      const int r1 = IDENT ? r : map_row(d.pre_map, r, c), c1 = IDENT ? c : map_col(d.pre_map, r, c);
      const int cr = IDENT ? (r1 >> 1) : map_row(d.chroma_map, r1 >> 1, c1 >> 1);
      const int cc = IDENT ? (c1 >> 1) : map_col(d.chroma_map, r1 >> 1, c1 >> 1);
      const long long off = (long long)cr * d.stride_uv + (long long)cc * d.pixel_stride_uv;
      const int U = f.pu[off], V = f.pv[off];
      uint32_t rgb = yuv_to_rgb_packed(yuv_luma_term(Y), yuv_chroma_terms(U, V));
      if (cvt == BANDCU_CVT_YUV2RGBA) rgb |= 0xFF000000u;
      return rgb;
    }
  }
}

// One output pixel of the Resize stage (packed channels).
template <int SRC, int CVT, int IDENT>
__device__ __forceinline__ uint32_t scale_px(const BandCuPlanDesc& d, const FramePtrs& f, int oy, int ox) {
  const bool plane = (d.flavor == BANDCU_FLAVOR_PLANE);
  switch (d.scale_mode) {
    case BANDCU_SCALE_COPY:
      return fetch_px<SRC, CVT, IDENT>(d, f, oy, ox);
    case BANDCU_SCALE_POINT: {  // ScaleARGBSimple / ScalePlaneSimple
      long long y = (long long)d.y0 + (long long)oy * d.dy;
      long long x = (long long)d.x0 + (long long)ox * d.dx;
      return fetch_px<SRC, CVT, IDENT>(d, f, (int)(y >> 16), (int)(x >> 16));
    }
    case BANDCU_SCALE_BILINEAR_DOWN: {  // ScaleARGBBilinearDown / ScalePlaneBilinearDown
      long long y = (long long)d.y0 + (long long)oy * d.dy;
      if (y > d.max_y) y = d.max_y;
      const int yi = (int)(y >> 16);
      const int yf = d.linear ? 0 : (int)((y >> 8) & 255);
      const long long x = (long long)d.x0 + (long long)ox * d.dx;
      const int xi = (int)(x >> 16);
      uint32_t a = fetch_px<SRC, CVT, IDENT>(d, f, yi, xi);
      uint32_t b = fetch_px<SRC, CVT, IDENT>(d, f, yi, xi + 1);
      if (yf) {
        a = interp_rows4(a, fetch_px<SRC, CVT, IDENT>(d, f, yi + 1, xi), yf);
        b = interp_rows4(b, fetch_px<SRC, CVT, IDENT>(d, f, yi + 1, xi + 1), yf);
      }
      if (plane) return blend_cols1_round7(a, b, (int)(x & 0xffff));
      return blend_cols4_trunc7(a, b, (int)((x >> 9) & 0x7f));
    }
    case BANDCU_SCALE_BILINEAR_UP: {  // Scale*BilinearUp: columns first, then rows
      long long y = (long long)d.y0 + (long long)oy * d.dy;
      if (y > d.max_y) y = d.max_y;
      const int yi = (int)(y >> 16);
      const int yf = d.linear ? 0 : (int)((y >> 8) & 255);
      const long long x = (long long)d.x0 + (long long)ox * d.dx;
      const int xi = (int)(x >> 16);
      const int xf7 = (int)((x >> 9) & 0x7f), xf16 = (int)(x & 0xffff);
      uint32_t a0 = fetch_px<SRC, CVT, IDENT>(d, f, yi, xi), b0 = fetch_px<SRC, CVT, IDENT>(d, f, yi, xi + 1);
      uint32_t top = plane ? blend_cols1_round7(a0, b0, xf16) : blend_cols4_trunc7(a0, b0, xf7);
      if (!yf) return top;
      uint32_t a1 = fetch_px<SRC, CVT, IDENT>(d, f, yi + 1, xi), b1 = fetch_px<SRC, CVT, IDENT>(d, f, yi + 1, xi + 1);
      uint32_t bot = plane ? blend_cols1_round7(a1, b1, xf16) : blend_cols4_trunc7(a1, b1, xf7);
      return interp_rows4(top, bot, yf);
    }
    case BANDCU_SCALE_BOX2: {  // 2x2 box at (y0>>16 + oy*(dy>>16), x0>>16 + ox*(dx>>16))
      const int r = (d.y0 >> 16) + oy * (d.dy >> 16);
      const int c = (d.x0 >> 16) + ox * (d.dx >> 16);
      return box2x2_4(fetch_px<SRC, CVT, IDENT>(d, f, r, c), fetch_px<SRC, CVT, IDENT>(d, f, r, c + 1), fetch_px<SRC, CVT, IDENT>(d, f, r + 1, c),
                      fetch_px<SRC, CVT, IDENT>(d, f, r + 1, c + 1));
    }
    case BANDCU_SCALE_VERTICAL: {  // ScalePlaneVertical (its own max_y, passed in d.max_y)
      long long y = (long long)d.y0 + (long long)oy * d.dy;
      if (y > d.max_y) y = d.max_y;
      const int yi = (int)(y >> 16);
      const int yf = d.linear ? 0 : (int)((y >> 8) & 255);  // d.linear here = "filtering == None"
      const int c = ox + (d.x0 >> 16);
      uint32_t a = fetch_px<SRC, CVT, IDENT>(d, f, yi, c);
      if (!yf) return a;
      return interp_rows4(a, fetch_px<SRC, CVT, IDENT>(d, f, yi + 1, c), yf);
    }
    case BANDCU_SCALE_DOWN34: {  // ScalePlaneDown34 with ScaleRowDown34_{0,1}_Box_C
      const int g = oy / 3, k = oy - 3 * g;
      const int m = ox / 3, j = ox - 3 * m;
      int rs, rt;  // s row, t row
      if (k == 0) { rs = 4 * g; rt = 4 * g + 1; }
      else if (k == 1) { rs = 4 * g + 1; rt = 4 * g + 2; }
      else { rs = 4 * g + 3; rt = 4 * g + 2; }
      int hs, ht;
      const int c = 4 * m;
      if (j == 0) {
        hs = ((int)fetch_px<SRC, CVT, IDENT>(d, f, rs, c) * 3 + (int)fetch_px<SRC, CVT, IDENT>(d, f, rs, c + 1) + 2) >> 2;
        ht = ((int)fetch_px<SRC, CVT, IDENT>(d, f, rt, c) * 3 + (int)fetch_px<SRC, CVT, IDENT>(d, f, rt, c + 1) + 2) >> 2;
      } else if (j == 1) {
        hs = ((int)fetch_px<SRC, CVT, IDENT>(d, f, rs, c + 1) + (int)fetch_px<SRC, CVT, IDENT>(d, f, rs, c + 2) + 1) >> 1;
        ht = ((int)fetch_px<SRC, CVT, IDENT>(d, f, rt, c + 1) + (int)fetch_px<SRC, CVT, IDENT>(d, f, rt, c + 2) + 1) >> 1;
      } else {
        hs = ((int)fetch_px<SRC, CVT, IDENT>(d, f, rs, c + 2) + (int)fetch_px<SRC, CVT, IDENT>(d, f, rs, c + 3) * 3 + 2) >> 2;
        ht = ((int)fetch_px<SRC, CVT, IDENT>(d, f, rt, c + 2) + (int)fetch_px<SRC, CVT, IDENT>(d, f, rt, c + 3) * 3 + 2) >> 2;
      }
      return (k == 1) ? (uint32_t)((hs + ht + 1) >> 1) : (uint32_t)((hs * 3 + ht + 2) >> 2);
    }
    case BANDCU_SCALE_DOWN38: {  // ScalePlaneDown38 with ScaleRowDown38_{3,2}_Box_C
      const int g = oy / 3, k = oy - 3 * g;
      const int m = ox / 3, j = ox - 3 * m;
      const int r0 = 8 * g + 3 * k, nr = (k == 2) ? 2 : 3;
      const int c0 = 8 * m + 3 * j, nc = (j == 2) ? 2 : 3;
      int sum = 0;
      for (int rr = 0; rr < nr; ++rr)
        for (int cc = 0; cc < nc; ++cc) sum += (int)fetch_px<SRC, CVT, IDENT>(d, f, r0 + rr, c0 + cc);
      const int n = nr * nc;
      const int mul = (n == 9) ? (65536 / 9) : (n == 6) ? (65536 / 6) : (65536 / 4);
      return (uint32_t)((sum * mul) >> 16) & 0xFFu;
    }
    case BANDCU_SCALE_UP2_LINEAR:
    case BANDCU_SCALE_UP2_BILINEAR: {  // Scale{Plane,UV}Up2_{Linear,Bilinear} + *_Any_C edges, per channel
      const int dw = d.dst_width, dh = d.dst_height, sh = d.vh;
      int cn, cf;  // near / far source column
      if (ox == 0) { cn = cf = 0; }
      else if (ox == dw - 1) { cn = cf = (dw - 1) / 2; }
      else { const int xx = (ox - 1) >> 1; if (ox & 1) { cn = xx; cf = xx + 1; } else { cn = xx + 1; cf = xx; } }
      uint32_t out = 0;
      if (d.scale_mode == BANDCU_SCALE_UP2_LINEAR) {
        int row;
        if (dh == 1) row = (sh - 1) / 2;
        else row = (int)(((long long)32767 + (long long)oy * d.dy) >> 16);  // d.dy = FixedDiv(sh-1, dh-1)
        const uint32_t pn = fetch_px<SRC, CVT, IDENT>(d, f, row, cn), pf = fetch_px<SRC, CVT, IDENT>(d, f, row, cf);
        for (int k = 0; k < d.channels; ++k) {
          const int n = (pn >> (8 * k)) & 0xFF, fr = (pf >> (8 * k)) & 0xFF;
          out |= (uint32_t)((3 * n + fr + 2) >> 2) << (8 * k);
        }
        return out;
      }
      int rn, rf;  // near / far source row
      if (oy == 0) { rn = rf = 0; }
      else if (oy == dh - 1 && !(dh & 1)) { rn = rf = sh - 1; }
      else { const int yy = (oy - 1) >> 1; if (oy & 1) { rn = yy; rf = yy + 1; } else { rn = yy + 1; rf = yy; } }
      const uint32_t pa = fetch_px<SRC, CVT, IDENT>(d, f, rn, cn), pb = fetch_px<SRC, CVT, IDENT>(d, f, rn, cf);
      const uint32_t pc = fetch_px<SRC, CVT, IDENT>(d, f, rf, cn), pd = fetch_px<SRC, CVT, IDENT>(d, f, rf, cf);
      for (int k = 0; k < d.channels; ++k) {
        const int A = (pa >> (8 * k)) & 0xFF, B = (pb >> (8 * k)) & 0xFF;
        const int Cc = (pc >> (8 * k)) & 0xFF, D = (pd >> (8 * k)) & 0xFF;
        out |= (uint32_t)((9 * A + 3 * B + 3 * Cc + D + 8) >> 4) << (8 * k);
      }
      return out;
    }
    default:
      return 0;
  }
}

template <typename OT>
__device__ __forceinline__ void store_norm(uint8_t* out_row, int ox, int ch, uint32_t px, float mean, float sd) {
  OT* o = reinterpret_cast<OT*>(out_row) + (long long)ox * ch;
  for (int k = 0; k < ch; ++k) o[k] = normalize_elem<OT>((float)((px >> (8 * k)) & 0xFFu), mean, sd);
}

template <int SRC, int CVT, int IDENT>
__global__ void __launch_bounds__(256) fused_generic_kernel(const __grid_constant__ GenericParams P) {
  const BandCuPlanDesc& d = P.d;
  const int ta = blockIdx.x * blockDim.x + threadIdx.x;
  const int tb = blockIdx.y * blockDim.y + threadIdx.y;
  const int ox = P.transposed ? tb : ta;
  const int oy = P.transposed ? ta : tb;
  if (ox >= d.dst_width || oy >= d.dst_height) return;
  const FramePtrs f = frame_ptrs(P.batch, blockIdx.z);
  const uint32_t px = scale_px<SRC, CVT, IDENT>(d, f, oy, ox);
  uint8_t* out_row = f.out + (long long)oy * d.dst_row_stride;
  const int ch = d.channels;
  if (!d.normalize) {
    uint8_t* o = out_row + (long long)ox * ch;
    if (ch == 4) {
      if ((reinterpret_cast<uintptr_t>(o) & 3) == 0) *reinterpret_cast<uint32_t*>(o) = px;
      else { o[0] = px & 0xFF; o[1] = (px >> 8) & 0xFF; o[2] = (px >> 16) & 0xFF; o[3] = px >> 24; }
    } else {
      for (int k = 0; k < ch; ++k) o[k] = (uint8_t)((px >> (8 * k)) & 0xFFu);
    }
    return;
  }
  switch (d.out_dtype) {
    case BANDCU_DT_F32: store_norm<float>(out_row, ox, ch, px, d.mean, d.std_); break;
    case BANDCU_DT_U8: store_norm<uint8_t>(out_row, ox, ch, px, d.mean, d.std_); break;
    case BANDCU_DT_I8: store_norm<int8_t>(out_row, ox, ch, px, d.mean, d.std_); break;
    case BANDCU_DT_I16: store_norm<int16_t>(out_row, ox, ch, px, d.mean, d.std_); break;
    case BANDCU_DT_I32: store_norm<int32_t>(out_row, ox, ch, px, d.mean, d.std_); break;
    default: break;
  }
}

// Plane scalers (ScalePlane on gray / Y / planar U, V; UVScale on interleaved chroma: ResizeYv / ResizeNv,
// libyuv_image_operator.cc:376-435) on the paths no staged kernel covers -- point sampling, ratios above 2, 3/4, 3/8, the 2x
// up-samplers: FOUR adjacent output pixels per thread through the same scale_px(), stored as one word (gray) or two (UV).
// One pixel per thread meant one byte store per thread (a warp wrote 32 bytes) and the row terms of the scaler -- y, its clamp,
// the row pointers -- computed once per output BYTE; here they are shared by four pixels and a warp stores 128 / 256 bytes.
template <int SRC>
__global__ void __launch_bounds__(256) fused_plane_x4_kernel(const __grid_constant__ GenericParams P) {
  const BandCuPlanDesc& d = P.d;
  constexpr int CH = SRC == BANDCU_SRC_UV ? 2 : 1;
  const int ox0 = (blockIdx.x * 32 + threadIdx.x) * 4;
  const int oy = blockIdx.y * blockDim.y + threadIdx.y;
  if (ox0 >= d.dst_width || oy >= d.dst_height) return;
  const FramePtrs f = frame_ptrs(P.batch, blockIdx.z);
  uint32_t px[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) px[k] = scale_px<SRC, BANDCU_CVT_NONE, 1>(d, f, oy, min(ox0 + k, d.dst_width - 1));
  uint8_t* o = f.out + (long long)oy * d.dst_row_stride + (long long)ox0 * CH;
  const int nvalid = min(4, d.dst_width - ox0);
  if (nvalid == 4 && (reinterpret_cast<uintptr_t>(o) & 3) == 0) {
    if (CH == 1) {
      *reinterpret_cast<uint32_t*>(o) = (px[0] & 0xFF) | ((px[1] & 0xFF) << 8) | ((px[2] & 0xFF) << 16) | (px[3] << 24);
    } else {
      reinterpret_cast<uint32_t*>(o)[0] = (px[0] & 0xFFFF) | (px[1] << 16);
      reinterpret_cast<uint32_t*>(o)[1] = (px[2] & 0xFFFF) | (px[3] << 16);
    }
    return;
  }
  for (int k = 0; k < nvalid; ++k)
    for (int c = 0; c < CH; ++c) o[k * CH + c] = (uint8_t)(px[k] >> (8 * c));
}

static bool is_identity_map(const BandCuAffine& m) {
  return m.r0 == 0 && m.rr == 1 && m.rc == 0 && m.c0 == 0 && m.cr == 0 && m.cc == 1;
}
typedef void (*GenericKernel)(GenericParams);
// Specialised instantiations for every (source kind, conversion) pair the plan compiler emits, with and without
// index maps; anything else runs the fully run-time variant.
static GenericKernel pick_generic(int src, int cvt, bool ident) {
#define BANDCU_PAIR(S, C)                                                                                   \
  if (src == (S) && cvt == (C)) return ident ? fused_generic_kernel<(S), (C), 1> : fused_generic_kernel<(S), (C), 0>;
  BANDCU_PAIR(BANDCU_SRC_GRAY, BANDCU_CVT_NONE)
  BANDCU_PAIR(BANDCU_SRC_UV, BANDCU_CVT_NONE)
  BANDCU_PAIR(BANDCU_SRC_RGB, BANDCU_CVT_NONE)
  BANDCU_PAIR(BANDCU_SRC_RGB, BANDCU_CVT_RGB2GRAY)
  BANDCU_PAIR(BANDCU_SRC_RGBA, BANDCU_CVT_NONE)
  BANDCU_PAIR(BANDCU_SRC_RGBA, BANDCU_CVT_RGBA2GRAY)
  BANDCU_PAIR(BANDCU_SRC_RGBA, BANDCU_CVT_RGBA2RGB)
  BANDCU_PAIR(BANDCU_SRC_YUV420, BANDCU_CVT_YUV2RGB)
  BANDCU_PAIR(BANDCU_SRC_YUV420, BANDCU_CVT_YUV2RGBA)
  BANDCU_PAIR(BANDCU_SRC_YUV420, BANDCU_CVT_YUV2GRAY)
#undef BANDCU_PAIR
  return fused_generic_kernel<-1, -1, 0>;
}

static int launch_packed_map_tiles(const BandCuPlanDesc& d, const BatchArgs& batch, int n, cudaStream_t stream);

int launch_generic(const BandCuPlanDesc& d, const BatchArgs& batch, int n, cudaStream_t stream) {
  {
    const int e = launch_packed_map_tiles(d, batch, n, stream);
    if (e != BANDCU_E_UNSUPPORTED) return e;
  }
  GenericParams P;
  P.d = d;
  P.batch = batch;
  P.transposed = (d.luma_map.rc != 0 && d.luma_map.rr == 0) ? 1 : 0;
  dim3 block(32, 8, 1);
  dim3 grid((d.dst_width + 31) / 32, (d.dst_height + 7) / 8, 1);
  if (P.transposed) grid = dim3((d.dst_height + 31) / 32, (d.dst_width + 7) / 8, 1);
  const bool ident = is_identity_map(d.luma_map) && is_identity_map(d.pre_map) && is_identity_map(d.chroma_map);
  void (*kern)(GenericParams) = pick_generic(d.src_kind, d.cvt, ident);
  {  // u8 planes without index maps: four pixels per thread
    static const bool x4_off = getenv("BANDCU_DISABLE_PLANE_X4") && getenv("BANDCU_DISABLE_PLANE_X4")[0] == '1';
    const bool gray = d.src_kind == BANDCU_SRC_GRAY && d.channels == 1, uv = d.src_kind == BANDCU_SRC_UV && d.channels == 2;
    if (!x4_off && (gray || uv) && d.cvt == BANDCU_CVT_NONE && ident && !d.normalize && d.out_dtype == BANDCU_DT_U8 &&
        d.scale_mode != BANDCU_SCALE_COPY) {
      kern = gray ? fused_plane_x4_kernel<BANDCU_SRC_GRAY> : fused_plane_x4_kernel<BANDCU_SRC_UV>;
      grid = dim3((d.dst_width + 127) / 128, (d.dst_height + 7) / 8, 1);
    }
  }
  // gridDim.z is limited to 65535: chunk very large batches
  for (int done = 0; done < n;) {
    int chunk = n - done > 65535 ? 65535 : n - done;
    GenericParams Q = P;
    if (Q.batch.table) {
      Q.batch.table += done;
    } else {
      long long o = (long long)done * Q.batch.in_stride;
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

// ---------------------------------------------------------------------------
// 4:2:0 -> 4:2:0 index-mapped copy: CropNv/CropYv (:950-1037), RotateYv/RotateNv
// (:786-887), YUV<->YUV ColorSpaceConvert (:125-162,:213-250,:302-367), Flip*
// (:1232-1363). Pure permutation of luma and chroma samples.
// ---------------------------------------------------------------------------
struct YuvMapParams {
  BandCuYuvMapDesc d;
  BandCuYuvPtrs p;                       // frame 0
  long long src_step, dst_step;          // bytes between consecutive frames (every plane pointer advances by it)
};
// blockIdx.z = 2 * frame + plane group (0: luma, 1: chroma)
__global__ void __launch_bounds__(256) yuv_map_kernel(const __grid_constant__ YuvMapParams P) {
  const BandCuYuvMapDesc& d = P.d;
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int r = blockIdx.y * blockDim.y + threadIdx.y;
  const long long so_f = (long long)(blockIdx.z >> 1) * P.src_step, do_f = (long long)(blockIdx.z >> 1) * P.dst_step;
  if ((blockIdx.z & 1) == 0) {
    if (c >= d.dst_width || r >= d.dst_height || !P.p.dy) return;
    const int sr = map_row(d.luma_map, r, c), sc = map_col(d.luma_map, r, c);
    ((uint8_t*)P.p.dy)[do_f + (long long)r * d.dst_stride_y + c] =
        ((const uint8_t*)P.p.sy)[so_f + (long long)sr * d.src_stride_y + sc];
  } else {
    const int cw = d.chroma_width_override > 0 ? d.chroma_width_override : (d.dst_width + 1) / 2;
    const int chh = (d.dst_height + 1) / 2;
    if (c >= cw || r >= chh) return;
    const int sr = map_row(d.chroma_map, r, c), sc = map_col(d.chroma_map, r, c);
    const long long so = so_f + (long long)sr * d.src_stride_uv + (long long)sc * d.src_pixel_stride_uv;
    const long long dof = do_f + (long long)r * d.dst_stride_uv + (long long)c * d.dst_pixel_stride_uv;
    ((uint8_t*)P.p.du)[dof] = ((const uint8_t*)P.p.su)[so];
    ((uint8_t*)P.p.dv)[dof] = ((const uint8_t*)P.p.sv)[so];
  }
}
// Tile-staged form of the same copy (the byte-per-thread kernel above moved 0.45 TB/s on a quarter-turn of 1080p NV21
// and 0.6-1.0 TB/s on flips and crops: one byte per load and store instruction, and under a quarter-turn every lane of a
// load touches its own sector). A CTA owns 32 output rows x 64 output bytes of one plane. The maps are signed
// permutations, so the tile's samples form one source rectangle: it is copied into shared memory row by row
// (coalesced byte loads, any alignment), then every thread gathers the four bytes of an output word from the staged
// rectangle and stores the word. The staged pitch is odd, so the gather is conflict-free whether consecutive output
// bytes walk a staged row (crop, flips) or a staged column (quarter-turns). E = bytes per element: 1 for luma and
// planar chroma, 2 for interleaved chroma pairs (copied as pairs from the lower of the two pointers).
template <int N>
struct IntTag {
  static constexpr int value = N;
};
// Tile: 32 output rows x kPlaneTew(E) output elements (64, 32, 32, 32 elements = 64, 64, 96, 128 bytes). Staged either
// way round the map turns it: at most 64 rows of at most 128 bytes.
__host__ __device__ constexpr int plane_tew(int E) { return E == 1 ? 64 : 32; }
// Quarter-turns of one- and two-byte elements use larger tiles (plane_turn_tile below): 64 output elements wide, 128 / 64 rows.
__host__ __device__ inline bool plane_big_turn(const PlaneMap& pl) { return pl.m.rr == 0 && pl.esize <= 2; }
__host__ __device__ inline int plane_tile_w(const PlaneMap& pl) { return plane_big_turn(pl) ? 64 : plane_tew(pl.esize); }
__host__ __device__ inline int plane_tile_h(const PlaneMap& pl) { return plane_big_turn(pl) ? (pl.esize == 1 ? 128 : 64) : 32; }
template <int E>
__device__ __forceinline__ void plane_map_tile(const PlaneMap& pl, long long src_off, long long dst_off, uint8_t* stage) {
  constexpr int TEW = plane_tew(E), TR = 32, WPR = TEW * E / 4;  // words per output tile row
  const int c_lo = blockIdx.x * TEW, r_lo = blockIdx.y * TR;
  if (c_lo >= pl.w || r_lo >= pl.h) return;
  const int c_hi = min(c_lo + TEW, pl.w) - 1, r_hi = min(r_lo + TR, pl.h) - 1;
  // source rectangle of the (clamped) tile: corners of a monotone map
  const int ra = map_row(pl.m, r_lo, c_lo), rb = map_row(pl.m, r_hi, c_hi);
  const int ca = map_col(pl.m, r_lo, c_lo), cb = map_col(pl.m, r_hi, c_hi);
  const int sr0 = min(ra, rb), sc0 = min(ca, cb);
  const int srows = abs(ra - rb) + 1, sbytes = (abs(ca - cb) + 1) * E;
  const int pitch = sbytes | 1;
  const uint8_t* src = pl.src + src_off + (long long)sr0 * pl.src_stride + (long long)sc0 * E;
  const int lane = threadIdx.x, wid = threadIdx.y;
  // all of a thread's loads are issued before its first shared-memory store (predicated, fully unrolled). One-byte
  // elements stage 32 rows x 64 bytes (straight maps) or 64 rows x 32 bytes (quarter-turns): two shapes, 8 slots each.
  auto stage_rect = [&](auto nr_tag, auto nb_tag) {
    constexpr int NR = decltype(nr_tag)::value, NB = decltype(nb_tag)::value;
    uint8_t v[NR * NB];
#pragma unroll
    for (int i = 0; i < NR; ++i) {
      const int r = wid + 8 * i;
      const uint8_t* srow = src + (long long)r * pl.src_stride + lane;
#pragma unroll
      for (int j = 0; j < NB; ++j)
        if (r < srows && lane + 32 * j < sbytes) v[i * NB + j] = __ldg(srow + 32 * j);
    }
#pragma unroll
    for (int i = 0; i < NR; ++i) {
      const int r = wid + 8 * i;
#pragma unroll
      for (int j = 0; j < NB; ++j)
        if (r < srows && lane + 32 * j < sbytes) stage[r * pitch + lane + 32 * j] = v[i * NB + j];
    }
  };
  if (E == 1) {
    if (pl.m.rr == 0) stage_rect(IntTag<8>(), IntTag<1>());
    else stage_rect(IntTag<4>(), IntTag<2>());
  } else {
    stage_rect(IntTag<4>(), IntTag<(E == 2 ? 2 : E == 3 ? 3 : 4)>());
  }
  __syncthreads();
  // staged offset of output element (r, c): row term + column term (each mapped coordinate depends on one of them)
  const int row_mul = pl.m.rr * pitch + pl.m.cr * E, col_mul = pl.m.rc * pitch + pl.m.cc * E;
  const int base = (pl.m.r0 - sr0) * pitch + (pl.m.c0 - sc0) * E;
  uint8_t* dst = pl.dst + dst_off;
  const int row_bytes = pl.w * E;
  for (int i = wid * 32 + lane; i < TR * WPR; i += 256) {
    const int ry = i / WPR, wq = i - ry * WPR;
    const int r = r_lo + ry, b0 = c_lo * E + wq * 4;  // first output byte of this word
    if (r > r_hi || b0 >= row_bytes) continue;
    const int rterm = base + row_mul * r;
    uint32_t v[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int b = min(b0 + k, row_bytes - 1);
      const int c = E == 1 ? b : E == 2 ? (b >> 1) : E == 4 ? (b >> 2) : (int)(((unsigned)b * 43691u) >> 17);
      v[k] = stage[rterm + col_mul * c + (b - c * E)];
    }
    uint8_t* o = dst + (long long)r * pl.dst_stride + b0;
    if (pl.vec && b0 + 4 <= row_bytes) {
      *reinterpret_cast<uint32_t*>(o) = v[0] | (v[1] << 8) | (v[2] << 16) | (v[3] << 24);
    } else {
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if (b0 + k < row_bytes) o[k] = (uint8_t)v[k];
    }
  }
}
// Quarter-turn of a plane of one- or two-byte elements. The staged rectangle is 64 source rows x 128 bytes (8 KB): a
// thread keeps eight 32-bit loads in flight (byte loads in two batches when the rectangle is not word aligned) -- the
// 32-row tiles above hold 8 bytes per thread in flight, ~16 KB per SM, which is the 1.2 TB/s they measure. Bytes go
// into shared memory at an odd pitch, so the gather of an output word (four staged rows, one column) is
// conflict-free; each thread keeps its word column, so the four column terms are computed once.
template <int E>
__device__ __forceinline__ void plane_turn_tile(const PlaneMap& pl, long long src_off, long long dst_off, uint8_t* stage) {
  constexpr int TEW = 64, TR = E == 1 ? 128 : 64, WPR = TEW * E / 4, RSTEP = 256 / WPR;
  const int c_lo = blockIdx.x * TEW, r_lo = blockIdx.y * TR;
  if (c_lo >= pl.w || r_lo >= pl.h) return;
  const int c_hi = min(c_lo + TEW, pl.w) - 1, r_hi = min(r_lo + TR, pl.h) - 1;
  const int ra = map_row(pl.m, r_lo, c_lo), rb = map_row(pl.m, r_hi, c_hi);
  const int ca = map_col(pl.m, r_lo, c_lo), cb = map_col(pl.m, r_hi, c_hi);
  const int sr0 = min(ra, rb), sc0 = min(ca, cb);
  const int srows = abs(ra - rb) + 1, sbytes = (abs(ca - cb) + 1) * E;  // <= 64 rows x 128 bytes
  const int pitch = sbytes | 1;
  const uint8_t* src = pl.src + src_off + (long long)sr0 * pl.src_stride + (long long)sc0 * E;
  const int lane = threadIdx.x, wid = threadIdx.y;
  if (((reinterpret_cast<uintptr_t>(src) | (uintptr_t)pl.src_stride | (uintptr_t)sbytes) & 3) == 0) {
    const int wprs = sbytes >> 2;  // <= 32 words per staged row: lane = word
    uint32_t v[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int r = wid + 8 * i;
      if (r < srows && lane < wprs) v[i] = __ldg(reinterpret_cast<const uint32_t*>(src + (long long)r * pl.src_stride) + lane);
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int r = wid + 8 * i;
      if (r < srows && lane < wprs) {
        uint8_t* q = stage + r * pitch + 4 * lane;
        q[0] = (uint8_t)v[i]; q[1] = (uint8_t)(v[i] >> 8); q[2] = (uint8_t)(v[i] >> 16); q[3] = (uint8_t)(v[i] >> 24);
      }
    }
  } else {
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      uint8_t v[16];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int r = wid + 8 * (4 * half + i);
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (r < srows && lane + 32 * j < sbytes) v[4 * i + j] = __ldg(src + (long long)r * pl.src_stride + lane + 32 * j);
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int r = wid + 8 * (4 * half + i);
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (r < srows && lane + 32 * j < sbytes) stage[r * pitch + lane + 32 * j] = v[4 * i + j];
      }
    }
  }
  __syncthreads();
  const int row_mul = pl.m.rr * pitch + pl.m.cr * E, col_mul = pl.m.rc * pitch + pl.m.cc * E;
  const int base = (pl.m.r0 - sr0) * pitch + (pl.m.c0 - sc0) * E;
  const int row_bytes = pl.w * E;
  const int tid = wid * 32 + lane;
  const int wq = tid & (WPR - 1), b0 = c_lo * E + wq * 4;
  if (b0 >= row_bytes) return;
  int cterm[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int b = min(b0 + k, row_bytes - 1);
    const int c = E == 1 ? b : (b >> 1);
    cterm[k] = base + col_mul * c + (b - c * E);
  }
  const bool whole = pl.vec && b0 + 4 <= row_bytes;
  int r = r_lo + tid / WPR;
  uint8_t* o = pl.dst + dst_off + (long long)r * pl.dst_stride + b0;
  const long long ostep = (long long)RSTEP * pl.dst_stride;
  for (; r <= r_hi; r += RSTEP, o += ostep) {
    const int rterm = row_mul * r;
    const uint32_t v0 = stage[rterm + cterm[0]], v1 = stage[rterm + cterm[1]], v2 = stage[rterm + cterm[2]],
                   v3 = stage[rterm + cterm[3]];
    if (whole) {
      *reinterpret_cast<uint32_t*>(o) = v0 | (v1 << 8) | (v2 << 16) | (v3 << 24);
    } else {
      o[0] = (uint8_t)v0;
      if (b0 + 1 < row_bytes) o[1] = (uint8_t)v1;
      if (b0 + 2 < row_bytes) o[2] = (uint8_t)v2;
      if (b0 + 3 < row_bytes) o[3] = (uint8_t)v3;
    }
  }
}
// Every present plane is a quarter-turn of one- or two-byte elements (4:2:0 frames, gray, UV): blockIdx.z = 3 * frame + plane
__global__ void __launch_bounds__(256) plane_turn_kernel(const __grid_constant__ PlaneMapParams P) {
  __shared__ uint8_t stage[64 * 129 + 64];  // 64 rows x 129 bytes
  const PlaneMap& pl = P.pl[blockIdx.z % 3];
  if (!pl.src) return;
  const long long so = (long long)(blockIdx.z / 3) * P.src_step, dof = (long long)(blockIdx.z / 3) * P.dst_step;
  if (pl.esize == 1) plane_turn_tile<1>(pl, so, dof, stage);
  else plane_turn_tile<2>(pl, so, dof, stage);
}
// blockIdx.z = 3 * frame + plane
__global__ void __launch_bounds__(256) plane_map_tile_kernel(const __grid_constant__ PlaneMapParams P) {
  __shared__ uint8_t stage[64 * 33 > 32 * 129 ? 64 * 33 : 32 * 129];
  const PlaneMap& pl = P.pl[blockIdx.z % 3];
  if (!pl.src) return;
  const long long so = (long long)(blockIdx.z / 3) * P.src_step, dof = (long long)(blockIdx.z / 3) * P.dst_step;
  switch (pl.esize) {
    case 1: plane_map_tile<1>(pl, so, dof, stage); break;
    case 2: plane_map_tile<2>(pl, so, dof, stage); break;
    case 3: plane_map_tile<3>(pl, so, dof, stage); break;
    default: plane_map_tile<4>(pl, so, dof, stage); break;
  }
}

// Straight maps (crop, flips, half-turn) need no staging: 16 output bytes are 16 CONSECUTIVE source bytes, ascending or
// descending. The staged kernel above keeps 8 - 16 bytes per thread in flight (~16 KB per SM: 1.2 TB/s by Little's
// law, which is what it measured on flips and crops); here a thread owns one 16-byte column of FOUR rows and issues
// its four loads (128-bit when the source chunk is aligned, 32-bit or byte loads otherwise) before the first store.
// A horizontal flip reverses the chunk: words swapped, bytes (E = 1) or byte pairs (E = 2) reversed by PRMT.
// blockIdx.z = 3 * frame + plane; grid.x covers 512-byte column groups, grid.y 32-row groups.
__global__ void __launch_bounds__(256) plane_rows16_kernel(const __grid_constant__ PlaneMapParams P) {
  const PlaneMap& pl = P.pl[blockIdx.z % 3];
  if (!pl.src) return;
  const int E = pl.esize;
  const int row_bytes = pl.w * E;
  const int b0 = (blockIdx.x * 32 + threadIdx.x) * 16;
  const int r0 = (blockIdx.y * 8 + threadIdx.y) * 4;
  if (b0 >= row_bytes || r0 >= pl.h) return;
  const uint8_t* src = pl.src + (long long)(blockIdx.z / 3) * P.src_step;
  uint8_t* dst = pl.dst + (long long)(blockIdx.z / 3) * P.dst_step;
  const bool rev = pl.m.cc < 0;
  const bool whole = b0 + 16 <= row_bytes;
  if (whole) {
    // lowest source byte of the chunk (row-relative): ascending from c0 * E + b0, or the mirror image of [b0, b0 + 16)
    const int sb = !rev ? pl.m.c0 * E + b0 : (pl.m.c0 + 1) * E - b0 - 16;
    const uint32_t perm = !rev ? 0x3210u : (E == 1 ? 0x0123u : E == 2 ? 0x1032u : 0x3210u);
    uint4 v[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int r = min(r0 + j, pl.h - 1);
      const uint8_t* a = src + (long long)(pl.m.r0 + pl.m.rr * r) * pl.src_stride + sb;
      const uintptr_t al = reinterpret_cast<uintptr_t>(a);
      if ((al & 15) == 0) {
        v[j] = __ldg(reinterpret_cast<const uint4*>(a));
      } else if ((al & 3) == 0) {
        const uint32_t* q = reinterpret_cast<const uint32_t*>(a);
        v[j] = make_uint4(__ldg(q), __ldg(q + 1), __ldg(q + 2), __ldg(q + 3));
      } else {
        uint32_t w[4];
#pragma unroll
        for (int k = 0; k < 4; ++k)
          w[k] = (uint32_t)__ldg(a + 4 * k) | ((uint32_t)__ldg(a + 4 * k + 1) << 8) | ((uint32_t)__ldg(a + 4 * k + 2) << 16) |
                 ((uint32_t)__ldg(a + 4 * k + 3) << 24);
        v[j] = make_uint4(w[0], w[1], w[2], w[3]);
      }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (r0 + j >= pl.h) break;
      uint4 o = v[j];
      if (rev) o = make_uint4(__byte_perm(v[j].w, 0u, perm), __byte_perm(v[j].z, 0u, perm), __byte_perm(v[j].y, 0u, perm),
                              __byte_perm(v[j].x, 0u, perm));
      *reinterpret_cast<uint4*>(dst + (long long)(r0 + j) * pl.dst_stride + b0) = o;
    }
    return;
  }
  for (int j = 0; j < 4 && r0 + j < pl.h; ++j) {  // the partial chunk at the end of a row: byte by byte
    const int r = r0 + j;
    const uint8_t* srow = src + (long long)(pl.m.r0 + pl.m.rr * r) * pl.src_stride;
    uint8_t* o = dst + (long long)r * pl.dst_stride;
    for (int b = b0; b < row_bytes; ++b) {
      const int c = E == 1 ? b : E == 2 ? (b >> 1) : E == 4 ? (b >> 2) : (int)(((unsigned)b * 43691u) >> 17);
      o[b] = __ldg(srow + (long long)(pl.m.c0 + pl.m.cc * c) * E + (b - c * E));
    }
  }
}

static bool plane_map_ok(const BandCuAffine& m) {
  const bool straight = (m.rr == 1 || m.rr == -1) && (m.cc == 1 || m.cc == -1) && m.rc == 0 && m.cr == 0;
  const bool swapped = m.rr == 0 && m.cc == 0 && (m.rc == 1 || m.rc == -1) && (m.cr == 1 || m.cr == -1);
  return straight || swapped;
}

static int run_plane_maps(PlaneMapParams& P, int n, cudaStream_t stream);

// BANDCU_E_UNSUPPORTED: maps that are not unit signed permutations, or a chroma repack (pixel strides differ)
static int launch_plane_map_tiles(const BandCuYuvMapDesc& d, const BandCuYuvPtrs& p, long long src_step, long long dst_step,
                                  int n, cudaStream_t stream) {
  static const bool off = getenv("BANDCU_DISABLE_PLANE_TILES") && getenv("BANDCU_DISABLE_PLANE_TILES")[0] == '1';
  if (off) return BANDCU_E_UNSUPPORTED;
  if (!plane_map_ok(d.luma_map) || !plane_map_ok(d.chroma_map)) return BANDCU_E_UNSUPPORTED;
  if (d.src_pixel_stride_uv != d.dst_pixel_stride_uv) return BANDCU_E_UNSUPPORTED;
  const int ps = d.src_pixel_stride_uv;
  if (ps != 1 && ps != 2) return BANDCU_E_UNSUPPORTED;
  const int cw = d.chroma_width_override > 0 ? d.chroma_width_override : (d.dst_width + 1) / 2;
  const int chh = (d.dst_height + 1) / 2;
  PlaneMapParams P;
  memset(&P, 0, sizeof(P));
  P.src_step = src_step;
  P.dst_step = dst_step;
  auto aligned = [&](const void* ptr, int stride) {
    return ((reinterpret_cast<uintptr_t>(ptr) | (uintptr_t)stride | (n == 1 ? 0 : (uintptr_t)dst_step)) & 3) == 0;
  };
  if (p.dy) {
    P.pl[0] = PlaneMap{(const uint8_t*)p.sy, (uint8_t*)p.dy, d.src_stride_y, d.dst_stride_y, d.dst_width, d.dst_height, 1,
                       aligned(p.dy, d.dst_stride_y) ? 1 : 0, d.luma_map};
  }
  if (ps == 1) {
    P.pl[1] = PlaneMap{(const uint8_t*)p.su, (uint8_t*)p.du, d.src_stride_uv, d.dst_stride_uv, cw, chh, 1,
                       aligned(p.du, d.dst_stride_uv) ? 1 : 0, d.chroma_map};
    P.pl[2] = PlaneMap{(const uint8_t*)p.sv, (uint8_t*)p.dv, d.src_stride_uv, d.dst_stride_uv, cw, chh, 1,
                       aligned(p.dv, d.dst_stride_uv) ? 1 : 0, d.chroma_map};
  } else {
    // interleaved pairs keep their byte order only if source and destination agree on which of U / V comes first
    const long long sd = (const uint8_t*)p.sv - (const uint8_t*)p.su, dd = (const uint8_t*)p.dv - (const uint8_t*)p.du;
    if (sd != dd || (sd != 1 && sd != -1)) return BANDCU_E_UNSUPPORTED;
    const uint8_t* s0 = sd == 1 ? (const uint8_t*)p.su : (const uint8_t*)p.sv;
    uint8_t* d0 = dd == 1 ? (uint8_t*)p.du : (uint8_t*)p.dv;
    P.pl[1] = PlaneMap{s0, d0, d.src_stride_uv, d.dst_stride_uv, cw, chh, 2, aligned(d0, d.dst_stride_uv) ? 1 : 0,
                       d.chroma_map};
  }
  return run_plane_maps(P, n, stream);
}

static int run_plane_maps(PlaneMapParams& P, int n, cudaStream_t stream) {
  {  // every plane straight, destination rows 16-byte aligned, no 3-byte elements under a horizontal flip: no staging
    bool rows16 = true;
    int bx = 1, by = 1;
    for (int k = 0; k < 3; ++k) {
      const PlaneMap& pl = P.pl[k];
      if (!pl.src) continue;
      const uintptr_t bits = reinterpret_cast<uintptr_t>(pl.dst) | (uintptr_t)pl.dst_stride | (n == 1 ? 0 : (uintptr_t)P.dst_step);
      if (pl.m.rr == 0 || (bits & 15) != 0 || (pl.esize == 3 && pl.m.cc < 0)) rows16 = false;
      bx = max(bx, (pl.w * pl.esize + 511) / 512);
      by = max(by, (pl.h + 31) / 32);
    }
    if (rows16) {
      for (int done = 0; done < n;) {
        const int chunk = n - done > 21845 ? 21845 : n - done;
        PlaneMapParams Q = P;
        for (int k = 0; k < 3; ++k) {
          if (!Q.pl[k].src) continue;
          Q.pl[k].src += (long long)done * P.src_step;
          Q.pl[k].dst += (long long)done * P.dst_step;
        }
        plane_rows16_kernel<<<dim3(bx, by, 3 * chunk), dim3(32, 8, 1), 0, stream>>>(Q);
        count_launch();
        done += chunk;
      }
      return (int)cudaGetLastError();
    }
  }
  {  // quarter-turns of word-aligned planes: word-granular transposes (bandcu_turn.cu)
    const int e = launch_plane_turns(P, n, stream);
    if (e != BANDCU_E_UNSUPPORTED) return e;
  }
  bool turn = true;  // every present plane a quarter-turn of 1- / 2-byte elements: the large-tile kernel
  int tx = 1, ty = 1;
  for (int k = 0; k < 3; ++k) {
    if (!P.pl[k].src) continue;
    if (!plane_big_turn(P.pl[k])) turn = false;
    tx = max(tx, (P.pl[k].w + plane_tile_w(P.pl[k]) - 1) / plane_tile_w(P.pl[k]));
    ty = max(ty, (P.pl[k].h + plane_tile_h(P.pl[k]) - 1) / plane_tile_h(P.pl[k]));
  }
  int gx = 1, gy = 1;
  for (int k = 0; k < 3; ++k) {
    if (!P.pl[k].src) continue;
    const int tew = plane_tew(P.pl[k].esize);
    gx = max(gx, (P.pl[k].w + tew - 1) / tew);
    gy = max(gy, (P.pl[k].h + 31) / 32);
  }
  const long long src_step = P.src_step, dst_step = P.dst_step;
  const dim3 block(32, 8, 1);
  for (int done = 0; done < n;) {
    const int chunk = n - done > 21845 ? 21845 : n - done;  // gridDim.z <= 65535
    PlaneMapParams Q = P;
    for (int k = 0; k < 3; ++k) {
      if (!Q.pl[k].src) continue;
      Q.pl[k].src += (long long)done * src_step;
      Q.pl[k].dst += (long long)done * dst_step;
    }
    if (turn) plane_turn_kernel<<<dim3(tx, ty, 3 * chunk), block, 0, stream>>>(Q);
    else plane_map_tile_kernel<<<dim3(gx, gy, 3 * chunk), block, 0, stream>>>(Q);
    count_launch();
    done += chunk;
  }
  return (int)cudaGetLastError();
}

// Crop / Rotate / Flip of a packed image (gray, UV pairs, RGB, RGBA) with nothing else fused: the plane copy above with
// 1 .. 4 bytes per element (the pixel-per-thread kernel rotated 1080p RGB at 0.24 TB/s: three byte loads and three
// byte stores per pixel, the loads a row apart under a quarter-turn).
static int launch_packed_map_tiles(const BandCuPlanDesc& d, const BatchArgs& batch, int n, cudaStream_t stream) {
  static const bool off = getenv("BANDCU_DISABLE_PLANE_TILES") && getenv("BANDCU_DISABLE_PLANE_TILES")[0] == '1';
  if (off || batch.table) return BANDCU_E_UNSUPPORTED;
  if (d.scale_mode != BANDCU_SCALE_COPY || d.normalize || d.cvt != BANDCU_CVT_NONE) return BANDCU_E_UNSUPPORTED;
  if (d.src_kind == BANDCU_SRC_YUV420 || d.dst_width != d.vw || d.dst_height != d.vh) return BANDCU_E_UNSUPPORTED;
  if (d.channels < 1 || d.channels > 4 || !plane_map_ok(d.luma_map)) return BANDCU_E_UNSUPPORTED;
  if (d.dst_row_stride > 0x7fffffffLL) return BANDCU_E_UNSUPPORTED;
  PlaneMapParams P;
  memset(&P, 0, sizeof(P));
  P.src_step = n == 1 ? 0 : batch.in_stride;
  P.dst_step = n == 1 ? 0 : batch.out_stride;
  const bool vec = ((reinterpret_cast<uintptr_t>(batch.first.out) | (uintptr_t)d.dst_row_stride | (uintptr_t)P.dst_step) & 3) == 0;
  P.pl[0] = PlaneMap{batch.first.p0, batch.first.out, d.stride0, (int)d.dst_row_stride, d.vw, d.vh, d.channels, vec ? 1 : 0,
                     d.luma_map};
  return run_plane_maps(P, n, stream);
}


// Chroma REPACKS with untouched geometry (ColorSpaceConvert between 4:2:0 layouts: NV12ToI420, I420ToNV12 / NV21,
// SwapUVPlane, libyuv_image_operator.cc:125-162,:213-250,:302-367): the luma plane is a straight copy and the chroma
// samples only change their interleaving. The byte-per-thread kernel above moved one byte per load and store; here a
// thread moves 16 luma bytes (blockIdx.z even) or EIGHT chroma samples of one row (odd): one 128-bit or two 64-bit loads,
// four PRMTs to split or merge the U / V bytes, one 128-bit or two 64-bit stores.
struct YuvRepackParams {
  const uint8_t *sy, *su, *sv;  // su / sv: plane bases; interleaved sources are read from the lower of the two
  uint8_t *dy, *du, *dv;
  int src_stride_y, src_stride_uv, dst_stride_y, dst_stride_uv;
  int width, height, cw, ch;
  int src_il, dst_il;      // 0 planar, 1 interleaved U first, 2 interleaved V first
  long long src_step, dst_step;
};
__global__ void __launch_bounds__(256) yuv_repack_rows_kernel(const __grid_constant__ YuvRepackParams P) {
  const int f = blockIdx.z >> 1;
  const long long so = (long long)f * P.src_step, dof = (long long)f * P.dst_step;
  const int x = blockIdx.x * blockDim.x + threadIdx.x, r = blockIdx.y * blockDim.y + threadIdx.y;
  if ((blockIdx.z & 1) == 0) {
    if (!P.dy || x * 16 >= P.width || r >= P.height) return;
    const uint4 v = __ldg(reinterpret_cast<const uint4*>(P.sy + so + (long long)r * P.src_stride_y) + x);
    reinterpret_cast<uint4*>(P.dy + dof + (long long)r * P.dst_stride_y)[x] = v;
    return;
  }
  if (x * 8 >= P.cw || r >= P.ch) return;
  uint2 u, v;  // eight U bytes, eight V bytes
  if (P.src_il) {
    const uint8_t* base = (P.src_il == 1 ? P.su : P.sv) + so + (long long)r * P.src_stride_uv;
    const uint4 w = __ldg(reinterpret_cast<const uint4*>(base) + x);   // (first, second) x 8
    const uint2 first = make_uint2(__byte_perm(w.x, w.y, 0x6420), __byte_perm(w.z, w.w, 0x6420));
    const uint2 second = make_uint2(__byte_perm(w.x, w.y, 0x7531), __byte_perm(w.z, w.w, 0x7531));
    u = P.src_il == 1 ? first : second;
    v = P.src_il == 1 ? second : first;
  } else {
    u = __ldg(reinterpret_cast<const uint2*>(P.su + so + (long long)r * P.src_stride_uv) + x);
    v = __ldg(reinterpret_cast<const uint2*>(P.sv + so + (long long)r * P.src_stride_uv) + x);
  }
  if (P.dst_il) {
    const uint2 first = P.dst_il == 1 ? u : v, second = P.dst_il == 1 ? v : u;
    uint8_t* base = (P.dst_il == 1 ? P.du : P.dv) + dof + (long long)r * P.dst_stride_uv;
    reinterpret_cast<uint4*>(base)[x] = make_uint4(__byte_perm(first.x, second.x, 0x5140), __byte_perm(first.x, second.x, 0x7362),
                                                   __byte_perm(first.y, second.y, 0x5140), __byte_perm(first.y, second.y, 0x7362));
  } else {
    reinterpret_cast<uint2*>(P.du + dof + (long long)r * P.dst_stride_uv)[x] = u;
    reinterpret_cast<uint2*>(P.dv + dof + (long long)r * P.dst_stride_uv)[x] = v;
  }
}
static bool yuv_map_is_identity(const BandCuAffine& m) {
  return m.r0 == 0 && m.rr == 1 && m.rc == 0 && m.c0 == 0 && m.cr == 0 && m.cc == 1;
}
// BANDCU_E_UNSUPPORTED: index maps, widths that are not whole vectors, unaligned planes.
static int launch_yuv_repack_rows(const BandCuYuvMapDesc& d, const BandCuYuvPtrs& p, long long src_step, long long dst_step, int n,
                                  cudaStream_t stream) {
  static const bool off = getenv("BANDCU_DISABLE_COPY_ROWS") && getenv("BANDCU_DISABLE_COPY_ROWS")[0] == '1';
  if (off || !yuv_map_is_identity(d.luma_map) || !yuv_map_is_identity(d.chroma_map) || d.chroma_width_override > 0)
    return BANDCU_E_UNSUPPORTED;
  YuvRepackParams P;
  memset(&P, 0, sizeof(P));
  auto layout = [](int ps, const void* u, const void* v) {
    if (ps == 1) return 0;
    const long long dd = (const uint8_t*)v - (const uint8_t*)u;
    return ps == 2 && dd == 1 ? 1 : ps == 2 && dd == -1 ? 2 : -1;
  };
  P.src_il = layout(d.src_pixel_stride_uv, p.su, p.sv);
  P.dst_il = layout(d.dst_pixel_stride_uv, p.du, p.dv);
  if (P.src_il < 0 || P.dst_il < 0) return BANDCU_E_UNSUPPORTED;
  P.width = d.dst_width; P.height = d.dst_height;
  P.cw = (d.dst_width + 1) / 2; P.ch = (d.dst_height + 1) / 2;
  if ((P.cw & 7) || (p.dy && (P.width & 15))) return BANDCU_E_UNSUPPORTED;
  const uintptr_t steps = n == 1 ? 0 : ((uintptr_t)src_step | (uintptr_t)dst_step);
  uintptr_t bits = steps | (uintptr_t)d.src_stride_uv | (uintptr_t)d.dst_stride_uv;
  if (p.dy) bits |= reinterpret_cast<uintptr_t>(p.sy) | reinterpret_cast<uintptr_t>(p.dy) | (uintptr_t)d.src_stride_y | (uintptr_t)d.dst_stride_y;
  // interleaved planes move as 16-byte vectors from the lower pointer, planar ones as 8-byte vectors
  const uintptr_t sb = P.src_il ? reinterpret_cast<uintptr_t>(P.src_il == 1 ? p.su : p.sv)
                                : ((reinterpret_cast<uintptr_t>(p.su) | reinterpret_cast<uintptr_t>(p.sv)) << 1);
  const uintptr_t db = P.dst_il ? reinterpret_cast<uintptr_t>(P.dst_il == 1 ? p.du : p.dv)
                                : ((reinterpret_cast<uintptr_t>(p.du) | reinterpret_cast<uintptr_t>(p.dv)) << 1);
  if ((bits | sb | db) & 15) return BANDCU_E_UNSUPPORTED;
  P.sy = (const uint8_t*)p.sy; P.su = (const uint8_t*)p.su; P.sv = (const uint8_t*)p.sv;
  P.dy = (uint8_t*)p.dy; P.du = (uint8_t*)p.du; P.dv = (uint8_t*)p.dv;
  P.src_stride_y = d.src_stride_y; P.src_stride_uv = d.src_stride_uv;
  P.dst_stride_y = d.dst_stride_y; P.dst_stride_uv = d.dst_stride_uv;
  P.src_step = src_step; P.dst_step = dst_step;
  const dim3 block(32, 8, 1);
  const int gx = std::max((P.width / 16 + 31) / 32, (P.cw / 8 + 31) / 32), gy = (P.height + 7) / 8;
  for (int done = 0; done < n;) {
    const int chunk = n - done > 32767 ? 32767 : n - done;
    YuvRepackParams Q = P;
    const long long so = (long long)done * src_step, dof = (long long)done * dst_step;
    Q.sy += so; Q.su += so; Q.sv += so;
    if (Q.dy) Q.dy += dof;
    Q.du += dof; Q.dv += dof;
    yuv_repack_rows_kernel<<<dim3(gx, gy, 2 * chunk), block, 0, stream>>>(Q);
    count_launch();
    done += chunk;
  }
  return (int)cudaGetLastError();
}

int launch_yuv_map(const BandCuYuvMapDesc& d, const BandCuYuvPtrs& p, long long src_step, long long dst_step, int n,
                   cudaStream_t stream) {
  {
    const int e = launch_plane_map_tiles(d, p, src_step, dst_step, n, stream);
    if (e != BANDCU_E_UNSUPPORTED) return e;
  }
  {
    const int e = launch_yuv_repack_rows(d, p, src_step, dst_step, n, stream);
    if (e != BANDCU_E_UNSUPPORTED) return e;
  }
  YuvMapParams P;
  P.d = d;
  P.src_step = src_step;
  P.dst_step = dst_step;
  dim3 block(32, 8, 1);
  for (int done = 0; done < n;) {  // gridDim.z <= 65535
    const int chunk = n - done > 32767 ? 32767 : n - done;
    P.p = p;
    const long long so = (long long)done * src_step, dof = (long long)done * dst_step;
    P.p.sy = (const uint8_t*)p.sy + so; P.p.su = (const uint8_t*)p.su + so; P.p.sv = (const uint8_t*)p.sv + so;
    if (p.dy) P.p.dy = (uint8_t*)p.dy + dof;
    P.p.du = (uint8_t*)p.du + dof; P.p.dv = (uint8_t*)p.dv + dof;
    dim3 grid((d.dst_width + 31) / 32, (d.dst_height + 7) / 8, 2 * chunk);
    yuv_map_kernel<<<grid, block, 0, stream>>>(P);
    count_launch();
    done += chunk;
  }
  return (int)cudaGetLastError();
}

// ---------------------------------------------------------------------------
// RGB / RGBA -> 4:2:0: RAWToI420 / ABGRToI420 / ABGRToNV12 / ABGRToNV21
// (libyuv_image_operator.cc:570,:626,:643,:661). libyuv row_common.cc, default C build:
//   Y = (66 R + 129 G + 25 B + 0x1080) >> 8
//   per 2x2 block (edge samples repeated): aX = (sum of the four X + 1) >> 1   ("2x" average)
//   U = (56 aB - 37 aG - 19 aR + 0x8080) >> 8 ;  V = (56 aR - 47 aG - 9 aB + 0x8080) >> 8
// One thread per chroma sample: it also writes the block's (up to) four luma samples.
// ---------------------------------------------------------------------------
struct RgbToYuvParams {
  BandCuRgbToYuvDesc d;
  BandCuRgbToYuvPtrs p;                  // frame 0
  long long src_step, dst_step;          // bytes between consecutive frames; blockIdx.z = frame
};
__global__ void __launch_bounds__(256) rgb_to_yuv420_kernel(const __grid_constant__ RgbToYuvParams P) {
  const BandCuRgbToYuvDesc& d = P.d;
  const int cx = blockIdx.x * blockDim.x + threadIdx.x;
  const int cy = blockIdx.y * blockDim.y + threadIdx.y;
  const int cw = (d.width + 1) >> 1, chh = (d.height + 1) >> 1;
  if (cx >= cw || cy >= chh) return;
  const long long do_f = (long long)blockIdx.z * P.dst_step;
  const uint8_t* src = (const uint8_t*)P.p.src + (long long)blockIdx.z * P.src_step;
  uint8_t* dy = (uint8_t*)P.p.dy + do_f;
  const int x0 = 2 * cx, y0 = 2 * cy;
  const int x1 = min(x0 + 1, d.width - 1), y1 = min(y0 + 1, d.height - 1);
  int sr = 0, sg = 0, sb = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int r = (k & 2) ? y1 : y0, c = (k & 1) ? x1 : x0;
    const uint8_t* q = src + (long long)map_row(d.src_map, r, c) * d.src_stride +
                       (long long)map_col(d.src_map, r, c) * d.src_channels;
    const int R = q[0], G = q[1], B = q[2];
    sr += R; sg += G; sb += B;
    const bool fresh = !((k & 1) && x1 == x0) && !((k & 2) && y1 == y0);  // a repeated edge sample is not a pixel
    if (fresh) dy[(long long)r * d.dst_stride_y + c] = (uint8_t)((66 * R + 129 * G + 25 * B + 0x1080) >> 8);
  }
  const int ar = (sr + 1) >> 1, ag = (sg + 1) >> 1, ab = (sb + 1) >> 1;
  const long long o = do_f + (long long)cy * d.dst_stride_uv + (long long)cx * d.dst_pixel_stride_uv;
  ((uint8_t*)P.p.du)[o] = (uint8_t)((56 * ab - 37 * ag - 19 * ar + 0x8080) >> 8);
  ((uint8_t*)P.p.dv)[o] = (uint8_t)((56 * ar - 47 * ag - 9 * ab + 0x8080) >> 8);
}
// Row form of the same conversion for untouched, word-aligned sources (the common case: a decoded RGBA / RGB frame
// going to the encoder or to a 4:2:0 model input). The one-thread-per-chroma-sample kernel above issues twelve byte
// loads and up to six byte stores per 2x2 block (0.50 of the measured HBM bandwidth on 1080p RGBA -> NV21). Here a
// thread owns FOUR pixels of TWO rows: two 128-bit loads (RGBA) or six 32-bit loads (RGB, re-cut into pixel words by
// funnel shifts), luma as one dot product per pixel (IDP4A: 66 R + 129 G + 25 B), the 2x2 channel sums in packed 16-bit
// lanes, and word stores: four luma bytes per row, and both chroma samples of the thread (one word when the chroma is
// interleaved, two half-words when it is planar). A warp reads 512 contiguous bytes per row and writes 128.
// CH: 3 / 4 source bytes per pixel. UVMODE: 0 planar, 1 interleaved U first (NV12), 2 interleaved V first (NV21).
template <int CH, int UVMODE>
__global__ void __launch_bounds__(256) rgb_to_yuv420_rows_kernel(const __grid_constant__ RgbToYuvParams P) {
  const BandCuRgbToYuvDesc& d = P.d;
  const int x4 = blockIdx.x * blockDim.x + threadIdx.x;   // group of four pixels
  const int cy = blockIdx.y * blockDim.y + threadIdx.y;   // chroma row = pair of luma rows
  if (x4 * 4 >= d.width || cy * 2 >= d.height) return;
  const long long do_f = (long long)blockIdx.z * P.dst_step;
  const uint8_t* src = (const uint8_t*)P.p.src + (long long)blockIdx.z * P.src_step +
                       (long long)(d.src_map.r0 + 2 * cy) * d.src_stride + (long long)(d.src_map.c0 + 4 * x4) * CH;
  uint32_t px[2][4];
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    const uint8_t* q = src + (long long)r * d.src_stride;
    if (CH == 4) {
      const uint4 v = __ldg(reinterpret_cast<const uint4*>(q));
      px[r][0] = v.x; px[r][1] = v.y; px[r][2] = v.z; px[r][3] = v.w;
    } else {
      const uint32_t* w = reinterpret_cast<const uint32_t*>(q);
      const uint32_t w0 = __ldg(w), w1 = __ldg(w + 1), w2 = __ldg(w + 2);  // R0 G0 B0 R1 | G1 B1 R2 G2 | B2 R3 G3 B3
      px[r][0] = w0;
      px[r][1] = __funnelshift_r(w0, w1, 24);
      px[r][2] = __funnelshift_r(w1, w2, 16);
      px[r][3] = w2 >> 8;
    }
  }
  // luma: (66 R + 129 G + 25 B + 0x1080) >> 8, byte 3 of a pixel word (alpha or the next pixel's red) weighs 0
  uint8_t* dy = (uint8_t*)P.p.dy + do_f + (long long)(2 * cy) * d.dst_stride_y + 4 * x4;
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    uint32_t y[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) y[k] = __dp4a(px[r][k], 0x00198142u, 0x1080u) >> 8;
    *reinterpret_cast<uint32_t*>(dy + (long long)r * d.dst_stride_y) = y[0] | (y[1] << 8) | (y[2] << 16) | (y[3] << 24);
  }
  // chroma: sums of R | B and of G over each 2x2 block in 16-bit lanes, "2x" average (sum + 1) >> 1
  uint32_t uv[2][2];
#pragma unroll
  for (int c = 0; c < 2; ++c) {
    const uint32_t a = px[0][2 * c], b = px[0][2 * c + 1], e = px[1][2 * c], f = px[1][2 * c + 1];
    const uint32_t rb = (a & 0x00FF00FFu) + (b & 0x00FF00FFu) + (e & 0x00FF00FFu) + (f & 0x00FF00FFu);
    const uint32_t g = ((a >> 8) & 0xFFu) + ((b >> 8) & 0xFFu) + ((e >> 8) & 0xFFu) + ((f >> 8) & 0xFFu);
    const int ar = (int)(((rb & 0xFFFFu) + 1u) >> 1), ab = (int)(((rb >> 16) + 1u) >> 1), ag = (int)((g + 1u) >> 1);
    uv[c][0] = (uint32_t)((56 * ab - 37 * ag - 19 * ar + 0x8080) >> 8) & 0xFFu;
    uv[c][1] = (uint32_t)((56 * ar - 47 * ag - 9 * ab + 0x8080) >> 8) & 0xFFu;
  }
  const long long co = do_f + (long long)cy * d.dst_stride_uv;
  if (UVMODE == 0) {
    *reinterpret_cast<uint16_t*>((uint8_t*)P.p.du + co + 2 * x4) = (uint16_t)(uv[0][0] | (uv[1][0] << 8));
    *reinterpret_cast<uint16_t*>((uint8_t*)P.p.dv + co + 2 * x4) = (uint16_t)(uv[0][1] | (uv[1][1] << 8));
  } else {
    const int f = UVMODE == 1 ? 0 : 1, s2 = 1 - f;   // first / second byte of a pair
    uint8_t* base = (uint8_t*)(UVMODE == 1 ? P.p.du : P.p.dv);
    *reinterpret_cast<uint32_t*>(base + co + 4 * x4) = uv[0][f] | (uv[0][s2] << 8) | (uv[1][f] << 16) | (uv[1][s2] << 24);
  }
}

// BANDCU_E_UNSUPPORTED: a source behind an index map other than a word-aligned crop, odd sizes, unaligned planes.
static int launch_rgb_to_yuv_rows(const BandCuRgbToYuvDesc& d, const BandCuRgbToYuvPtrs& p, long long src_step, long long dst_step,
                                  int n, cudaStream_t stream) {
  static const bool off = getenv("BANDCU_DISABLE_COPY_ROWS") && getenv("BANDCU_DISABLE_COPY_ROWS")[0] == '1';
  if (off) return BANDCU_E_UNSUPPORTED;
  const BandCuAffine& m = d.src_map;
  if (!(m.rr == 1 && m.rc == 0 && m.cr == 0 && m.cc == 1 && m.r0 >= 0 && m.c0 >= 0)) return BANDCU_E_UNSUPPORTED;
  if ((d.width & 3) || (d.height & 1)) return BANDCU_E_UNSUPPORTED;
  const int CH = d.src_channels;
  const uintptr_t steps = n == 1 ? 0 : ((uintptr_t)src_step | (uintptr_t)dst_step);
  const uintptr_t sbits = reinterpret_cast<uintptr_t>(p.src) | (uintptr_t)d.src_stride | (uintptr_t)((long long)m.c0 * CH) | steps;
  if (sbits & (CH == 4 ? 15 : 3)) return BANDCU_E_UNSUPPORTED;
  if ((reinterpret_cast<uintptr_t>(p.dy) | (uintptr_t)d.dst_stride_y | steps) & 3) return BANDCU_E_UNSUPPORTED;
  int uvmode;
  if (d.dst_pixel_stride_uv == 1) {
    uvmode = 0;
    if ((reinterpret_cast<uintptr_t>(p.du) | reinterpret_cast<uintptr_t>(p.dv) | (uintptr_t)d.dst_stride_uv | steps) & 1)
      return BANDCU_E_UNSUPPORTED;
  } else if (d.dst_pixel_stride_uv == 2) {
    const long long dd = (const uint8_t*)p.dv - (const uint8_t*)p.du;
    if (dd != 1 && dd != -1) return BANDCU_E_UNSUPPORTED;
    uvmode = dd == 1 ? 1 : 2;
    const void* first = dd == 1 ? p.du : p.dv;
    if ((reinterpret_cast<uintptr_t>(first) | (uintptr_t)d.dst_stride_uv | steps) & 3) return BANDCU_E_UNSUPPORTED;
  } else {
    return BANDCU_E_UNSUPPORTED;
  }
  RgbToYuvParams P;
  P.d = d;
  P.src_step = src_step;
  P.dst_step = dst_step;
  const dim3 block(32, 8, 1);
  const dim3 grid_xy((d.width / 4 + 31) / 32, (d.height / 2 + 7) / 8, 1);
  for (int done = 0; done < n;) {
    const int chunk = n - done > 65535 ? 65535 : n - done;
    const long long so = (long long)done * src_step, dof = (long long)done * dst_step;
    P.p.src = (const uint8_t*)p.src + so;
    P.p.dy = (uint8_t*)p.dy + dof; P.p.du = (uint8_t*)p.du + dof; P.p.dv = (uint8_t*)p.dv + dof;
    const dim3 grid(grid_xy.x, grid_xy.y, chunk);
    if (CH == 4) {
      if (uvmode == 0) rgb_to_yuv420_rows_kernel<4, 0><<<grid, block, 0, stream>>>(P);
      else if (uvmode == 1) rgb_to_yuv420_rows_kernel<4, 1><<<grid, block, 0, stream>>>(P);
      else rgb_to_yuv420_rows_kernel<4, 2><<<grid, block, 0, stream>>>(P);
    } else {
      if (uvmode == 0) rgb_to_yuv420_rows_kernel<3, 0><<<grid, block, 0, stream>>>(P);
      else if (uvmode == 1) rgb_to_yuv420_rows_kernel<3, 1><<<grid, block, 0, stream>>>(P);
      else rgb_to_yuv420_rows_kernel<3, 2><<<grid, block, 0, stream>>>(P);
    }
    count_launch();
    done += chunk;
  }
  return (int)cudaGetLastError();
}

int launch_rgb_to_yuv(const BandCuRgbToYuvDesc& d, const BandCuRgbToYuvPtrs& p, long long src_step, long long dst_step,
                      int n, cudaStream_t stream) {
  {
    const int e = launch_rgb_to_yuv_rows(d, p, src_step, dst_step, n, stream);
    if (e != BANDCU_E_UNSUPPORTED) return e;
  }
  RgbToYuvParams P;
  P.d = d;
  P.src_step = src_step;
  P.dst_step = dst_step;
  dim3 block(32, 8, 1);
  for (int done = 0; done < n;) {
    const int chunk = n - done > 65535 ? 65535 : n - done;
    const long long so = (long long)done * src_step, dof = (long long)done * dst_step;
    P.p.src = (const uint8_t*)p.src + so;
    P.p.dy = (uint8_t*)p.dy + dof; P.p.du = (uint8_t*)p.du + dof; P.p.dv = (uint8_t*)p.dv + dof;
    dim3 grid(((d.width + 1) / 2 + 31) / 32, ((d.height + 1) / 2 + 7) / 8, chunk);
    rgb_to_yuv420_kernel<<<grid, block, 0, stream>>>(P);
    count_launch();
    done += chunk;
  }
  return (int)cudaGetLastError();
}

// ---------------------------------------------------------------------------
// Element-wise Normalize / DataTypeConvert over any dtype pair
// (common_operator.cc:24-65): out[i] = static_cast<Out>((in[i] - mean)) / std.
// The difference is computed in float for integer/float inputs and in double for
// a double input (usual arithmetic conversions against the float mean).
// ---------------------------------------------------------------------------
template <typename IT, typename OT>
__global__ void __launch_bounds__(256) normalize_kernel(const IT* __restrict__ in, OT* __restrict__ out, long long n,
                                                        float mean, float sd) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long step = (long long)gridDim.x * blockDim.x;
  for (; i < n; i += step) {
    if constexpr (sizeof(IT) == 8) {  // double input
      double diff = __dsub_rn((double)in[i], (double)mean);
      OT t = (OT)diff;
      out[i] = (OT)__fdiv_rn((float)t, sd);
    } else {
      out[i] = normalize_elem<OT>((float)in[i], mean, sd);
    }
  }
}
template <typename IT>
static int normalize_from(const void* in, void* out, int out_dtype, long long n, float mean, float sd,
                          cudaStream_t stream) {
  int block = 256;
  long long want = (n + block - 1) / block;
  int grid = (int)(want > 148LL * 16 ? 148LL * 16 : (want < 1 ? 1 : want));
  switch (out_dtype) {
    case BANDCU_DT_F32: normalize_kernel<IT, float><<<grid, block, 0, stream>>>((const IT*)in, (float*)out, n, mean, sd); break;
    case BANDCU_DT_U8: normalize_kernel<IT, uint8_t><<<grid, block, 0, stream>>>((const IT*)in, (uint8_t*)out, n, mean, sd); break;
    case BANDCU_DT_I8: normalize_kernel<IT, int8_t><<<grid, block, 0, stream>>>((const IT*)in, (int8_t*)out, n, mean, sd); break;
    case BANDCU_DT_I16: normalize_kernel<IT, int16_t><<<grid, block, 0, stream>>>((const IT*)in, (int16_t*)out, n, mean, sd); break;
    case BANDCU_DT_I32: normalize_kernel<IT, int32_t><<<grid, block, 0, stream>>>((const IT*)in, (int32_t*)out, n, mean, sd); break;
    default: return BANDCU_E_UNSUPPORTED;
  }
  count_launch();
  return (int)cudaGetLastError();
}
int launch_normalize(const void* in, int in_dtype, void* out, int out_dtype, long long n, float mean, float sd,
                     cudaStream_t stream) {
  switch (in_dtype) {
    case BANDCU_DT_U8: return normalize_from<uint8_t>(in, out, out_dtype, n, mean, sd, stream);
    case BANDCU_DT_I8: return normalize_from<int8_t>(in, out, out_dtype, n, mean, sd, stream);
    case BANDCU_DT_I16: return normalize_from<int16_t>(in, out, out_dtype, n, mean, sd, stream);
    case BANDCU_DT_I32: return normalize_from<int32_t>(in, out, out_dtype, n, mean, sd, stream);
    case BANDCU_DT_F32: return normalize_from<float>(in, out, out_dtype, n, mean, sd, stream);
    case BANDCU_DT_F64: return normalize_from<double>(in, out, out_dtype, n, mean, sd, stream);
    default: return BANDCU_E_UNSUPPORTED;
  }
}

}  // namespace bandcu
