// bandcu_planes.cu -- down-scaling one u8 plane the way libyuv's plane scalers walk it: a whole source row pair is blended
// ONCE (InterpolateRow, four pixels per 32-bit word), then the column filter picks its two taps per output pixel out of
// that blended row (ScalePlaneBilinearDown = InterpolateRow + ScaleFilterCols, scale.cc; ScaleUVBilinearDown =
// InterpolateRow + ScaleUVFilterCols, scale_uv.cc) -- the same order of operations, so the same bits.
//
// Who calls it: Resize of a 4:2:0 BUFFER (ResizeYv / ResizeNv, libyuv_image_operator.cc:376-435: I420Scale / NV12Scale run
// ScalePlane on Y and planar U, V and UVScale on interleaved chroma) and Resize of a gray image (:1367-1380).
// Before: the 4-pixels-per-thread packed kernel (gray, ratios <= 2) and the pixel-per-thread kernel (everything else) fetched
// four BYTES per output pixel and blended one pixel per packed word: 1080p -> 720p NV21 ran at 0.25 of the HBM roofline
// (ncu launch list: 476 us for 256 luma planes, 257 us for 256 chroma planes).
//
// One WARP owns an output row of one plane of one frame (eight rows per CTA, no CTA barrier):
//   stage   : the blended source row (word loads of rows yi and yi+1, interp_rows4 on four pixels at once) -> the warp's
//             slice of shared memory;
//   columns : four adjacent output pixels per thread, taps read as bytes (plane) / 16-bit pairs (UV) from shared memory,
//             plane blender (x86 ScaleFilterCols_C rounding) or the 7-bit truncating ARGB/UV blender, one 32-bit store
//             (plane) or two (UV).
// BANDCU_SCALE_BOX2 (exact even ratios: ScaleRowDown2Box / ScaleUVRowDown2Box / DownEvenBox) stages the two rows unblended
// and sums the 2x2 box per pixel.
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdlib>

#include "bandcu.h"
#include "bandcu_device.cuh"
#include "bandcu_internal.h"

namespace bandcu {
namespace {

struct PlaneRowsParams {
  BatchArgs batch;
  int sw, sh;      // source plane, pixels
  int stride0;     // bytes between source rows
  int dw, dh;
  int x0, y0, dx, dy, max_y, linear;
  int box;         // 1: BANDCU_SCALE_BOX2
  long long dst_row_stride;
  int rows_per_cta;  // consecutive output rows per WARP
  int row_words;   // 32-bit words covering one source row (the last one may be partial)
  int full_words;  // words that lie entirely inside the row
};

// the partial last word of a row, assembled from bytes (the bytes behind the last pixel of the last row need not exist)
__device__ __noinline__ uint32_t row_tail_word(const uint8_t* __restrict__ row, int w, int nbytes) {
  uint32_t v = 0;
  for (int k = 4 * w; k < nbytes; ++k) v |= (uint32_t)__ldg(row + k) << (8 * (k - 4 * w));
  return v;
}

template <int CH>
__global__ void __launch_bounds__(256) plane_rows_down_kernel(const __grid_constant__ PlaneRowsParams P) {
  extern __shared__ __align__(16) uint32_t sm[];
  const FramePtrs f = frame_ptrs(P.batch, blockIdx.y);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int pitch = P.row_words + 1;
  // one WARP per output row, no CTA barrier: a row's loads are all in flight together and eight rows per CTA overlap
  // (a CTA-wide stage -> barrier -> filter loop ran one load round trip per row, back to back: 449 us per 256 luma planes)
  uint32_t* top = sm + wid * (P.box ? 2 * pitch : pitch);
  uint32_t* bot = top + pitch;  // box mode only
  const int oy_begin = (blockIdx.x * 8 + wid) * P.rows_per_cta;
  const int oy_end = min(P.dh, oy_begin + P.rows_per_cta);
  for (int oy = oy_begin; oy < oy_end; ++oy) {
    int r0, r1, yf = 0;
    if (P.box) {
      const int r = (P.y0 >> 16) + oy * (P.dy >> 16);
      r0 = min(max(r, 0), P.sh - 1);
      r1 = min(max(r + 1, 0), P.sh - 1);
    } else {
      long long y = (long long)P.y0 + (long long)oy * P.dy;
      if (y > P.max_y) y = P.max_y;
      const int yi = (int)(y >> 16);
      yf = P.linear ? 0 : (int)((y >> 8) & 255);
      r0 = min(max(yi, 0), P.sh - 1);
      r1 = min(max(yi + 1, 0), P.sh - 1);
    }
    const uint8_t* s0 = f.p0 + (long long)r0 * P.stride0;
    const uint8_t* s1 = f.p0 + (long long)r1 * P.stride0;
    const uint32_t* w0 = reinterpret_cast<const uint32_t*>(s0);
    const uint32_t* w1 = reinterpret_cast<const uint32_t*>(s1);
    if (P.box) {
#pragma unroll 4
      for (int w = lane; w < P.full_words; w += 32) {
        top[w] = __ldg(w0 + w);
        bot[w] = __ldg(w1 + w);
      }
    } else if (yf) {
#pragma unroll 4
      for (int w = lane; w < P.full_words; w += 32) top[w] = interp_rows4(__ldg(w0 + w), __ldg(w1 + w), yf);
    } else {
#pragma unroll 4
      for (int w = lane; w < P.full_words; w += 32) top[w] = __ldg(w0 + w);
    }
    if (P.full_words < P.row_words && lane == 0) {  // one partial word at the end of the row
      const int w = P.full_words, nbytes = P.sw * CH;
      const uint32_t a = row_tail_word(s0, w, nbytes), b = row_tail_word(s1, w, nbytes);
      if (P.box) { top[w] = a; bot[w] = b; }
      else top[w] = yf ? interp_rows4(a, b, yf) : a;
    }
    __syncwarp();
    uint8_t* orow = f.out + (long long)oy * P.dst_row_stride;
    const bool words_ok = ((reinterpret_cast<uintptr_t>(orow) & 3) == 0);
    for (int ox0 = 4 * lane; ox0 < P.dw; ox0 += 4 * 32) {
      uint32_t px[4];
      const bool whole = ox0 + 4 <= P.dw;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int ox = whole ? ox0 + k : min(ox0 + k, P.dw - 1);
        if (P.box) {
          const int c = (P.x0 >> 16) + ox * (P.dx >> 16);
          const int c0 = min(max(c, 0), P.sw - 1), c1 = min(max(c + 1, 0), P.sw - 1);
          if (CH == 1) {
            const uint8_t *t = reinterpret_cast<const uint8_t*>(top), *b = reinterpret_cast<const uint8_t*>(bot);
            px[k] = box2x2_4(t[c0], t[c1], b[c0], b[c1]);
          } else {
            const uint16_t *t = reinterpret_cast<const uint16_t*>(top), *b = reinterpret_cast<const uint16_t*>(bot);
            px[k] = box2x2_4(t[c0], t[c1], b[c0], b[c1]);
          }
        } else {
          const int x = P.x0 + ox * P.dx;  // 0 <= x < 2^31: checked on the host
          const int xi = x >> 16;
          const int c0 = min(xi, P.sw - 1), c1 = min(xi + 1, P.sw - 1);
          if (CH == 1) {
            const uint8_t* t = reinterpret_cast<const uint8_t*>(top);
            px[k] = blend_cols1_round7(t[c0], t[c1], x & 0xffff);
          } else {
            const uint16_t* t = reinterpret_cast<const uint16_t*>(top);
            px[k] = blend_cols4_trunc7(t[c0], t[c1], (x >> 9) & 0x7f);
          }
        }
      }
      uint8_t* o = orow + (long long)ox0 * CH;
      if (whole && words_ok) {
        if (CH == 1) {
          *reinterpret_cast<uint32_t*>(o) = (px[0] & 0xFF) | ((px[1] & 0xFF) << 8) | ((px[2] & 0xFF) << 16) | (px[3] << 24);
        } else {
          reinterpret_cast<uint32_t*>(o)[0] = (px[0] & 0xFFFF) | (px[1] << 16);
          reinterpret_cast<uint32_t*>(o)[1] = (px[2] & 0xFFFF) | (px[3] << 16);
        }
      } else {
        for (int k = 0; k < 4 && ox0 + k < P.dw; ++k)
          for (int c = 0; c < CH; ++c) o[k * CH + c] = (uint8_t)(px[k] >> (8 * c));
      }
    }
    __syncwarp();  // the next row overwrites this warp's staged row
  }
}

// ---- exact 2x up-samplers --------------------------------------------------------------------------------------------------
// ScalePlaneUp2_Bilinear / ScaleUVBilinearUp2 (scale.cc, scale_uv.cc; + their *_Any_C edge rules): every output pixel is
// (9A + 3B + 3C + D + 8) >> 4 over the 2x2 source neighbourhood nearest to it. Per output row the vertical half
// v[c] = 3 * near[c] + far[c] (<= 1020, exact in 16 bits) is staged once per SOURCE column, and an output pixel is
// (3 * v[cn] + v[cf] + 8) >> 4 = the same sum. ScalePlaneUp2_Linear / ScaleUVLinearUp2: rows are point-sampled, the pixel is
// (3 * n + f + 2) >> 2 -- the same walk with v = the source pixel. The pixel-per-thread kernel evaluated the four taps, the
// edge rules and a per-channel loop for every output pixel (about a hundred instructions each: 0.04-0.06 of the HBM roofline
// on RGBA 64^2 -> gray 128^2 and on 540p -> 1080p NV21).
// K: 0 gray source, 1 interleaved UV (two channels, SWAR on 16-bit lanes), 2 RGBA -> gray on the fetch, 3 RGB -> gray.
struct PlaneUp2Params {
  BatchArgs batch;
  int sw, sh, stride0;
  int dw, dh;
  int dy;        // UP2_LINEAR: FixedDiv(sh - 1, dh - 1)
  int bilinear;  // 1: 9-3-3-1; 0: linear
  long long dst_row_stride;
  int rows_per_warp;
};

template <int K>
__device__ __forceinline__ uint32_t up2_fetch(const uint8_t* __restrict__ row, int c) {
  if (K == 0) return __ldg(row + c);
  if (K == 2) {
    const uint32_t w = __ldg(reinterpret_cast<const uint32_t*>(row) + c);
    return rgb_to_yj(w & 0xFF, (w >> 8) & 0xFF, (w >> 16) & 0xFF);
  }
  return rgb_to_yj(__ldg(row + 3 * c), __ldg(row + 3 * c + 1), __ldg(row + 3 * c + 2));
}

template <int K>
__global__ void __launch_bounds__(256) plane_up2_kernel(const __grid_constant__ PlaneUp2Params P) {
  extern __shared__ __align__(16) uint32_t sm[];
  constexpr int CH = K == 1 ? 2 : 1;
  const FramePtrs f = frame_ptrs(P.batch, blockIdx.y);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  uint32_t* v = sm + wid * (P.sw + 2) + 1;  // v[-1] and v[sw]: the clamped neighbours of the first / last column
  const int oy_begin = (blockIdx.x * 8 + wid) * P.rows_per_warp;
  const int oy_end = min(P.dh, oy_begin + P.rows_per_warp);
  const uint32_t rnd = P.bilinear ? 8u : 2u;
  const int sh_out = P.bilinear ? 4 : 2;
  for (int oy = oy_begin; oy < oy_end; ++oy) {
    int rn, rf;
    if (P.bilinear) {
      if (oy == 0) { rn = rf = 0; }
      else if (oy == P.dh - 1 && !(P.dh & 1)) { rn = rf = P.sh - 1; }
      else { const int yy = (oy - 1) >> 1; if (oy & 1) { rn = yy; rf = yy + 1; } else { rn = yy + 1; rf = yy; } }
      rn = min(max(rn, 0), P.sh - 1);
      rf = min(max(rf, 0), P.sh - 1);
    } else {
      rn = P.dh == 1 ? (P.sh - 1) / 2 : (int)(((long long)32767 + (long long)oy * P.dy) >> 16);
      rn = rf = min(max(rn, 0), P.sh - 1);
    }
    const uint8_t* sn = f.p0 + (long long)rn * P.stride0;
    const uint8_t* sf = f.p0 + (long long)rf * P.stride0;
    // K == 1 keeps its two channels in the 16-bit lanes of one word: u | v << 16
    auto fetch = [&](const uint8_t* row, int c) -> uint32_t {
      if (K == 1) {
        const uint32_t q = __ldg(reinterpret_cast<const uint16_t*>(row) + c);
        return (q & 0xFF) | ((q & 0xFF00) << 8);
      }
      return up2_fetch<K>(row, c);
    };
    if (P.bilinear) {
#pragma unroll 4
      for (int c = lane; c < P.sw; c += 32) v[c] = 3u * fetch(sn, c) + fetch(sf, c);
    } else {
#pragma unroll 4
      for (int c = lane; c < P.sw; c += 32) v[c] = fetch(sn, c);
    }
    __syncwarp();
    if (lane == 0) { v[-1] = v[0]; v[P.sw] = v[P.sw - 1]; }
    __syncwarp();
    uint8_t* orow = f.out + (long long)oy * P.dst_row_stride;
    const bool words_ok = ((reinterpret_cast<uintptr_t>(orow) & 3) == 0);
    const uint32_t mask = CH == 2 ? 0x00FF00FFu : 0xFFu;
    const uint32_t rr = CH == 2 ? rnd * 0x00010001u : rnd;
    for (int ox0 = 4 * lane; ox0 < P.dw; ox0 += 4 * 32) {
      uint32_t px[4];
      const int j = ox0 >> 1;
      if (ox0 + 4 < P.dw) {  // the row's last pixel (its own edge rule) is not among these four
        const uint32_t a = v[j - 1], b = v[j], c = v[j + 1], e = v[j + 2];
        px[0] = ((3u * b + a + rr) >> sh_out) & mask;
        px[1] = ((3u * b + c + rr) >> sh_out) & mask;
        px[2] = ((3u * c + b + rr) >> sh_out) & mask;
        px[3] = ((3u * c + e + rr) >> sh_out) & mask;
      } else {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int ox = min(ox0 + k, P.dw - 1);
          int cn, cf;
          if (ox == 0) { cn = cf = 0; }
          else if (ox == P.dw - 1) { cn = cf = (P.dw - 1) / 2; }
          else { const int xx = (ox - 1) >> 1; if (ox & 1) { cn = xx; cf = xx + 1; } else { cn = xx + 1; cf = xx; } }
          cn = min(cn, P.sw - 1);
          cf = min(cf, P.sw - 1);
          px[k] = ((3u * v[cn] + v[cf] + rr) >> sh_out) & mask;
        }
      }
      uint8_t* o = orow + (long long)ox0 * CH;
      if (ox0 + 4 <= P.dw && words_ok) {
        if (CH == 1) {
          *reinterpret_cast<uint32_t*>(o) = px[0] | (px[1] << 8) | (px[2] << 16) | (px[3] << 24);
        } else {  // u | v << 16  ->  u, v bytes
          reinterpret_cast<uint32_t*>(o)[0] = (px[0] & 0xFF) | ((px[0] >> 8) & 0xFF00) | ((px[1] & 0xFF) << 16) | ((px[1] & 0xFF0000) << 8);
          reinterpret_cast<uint32_t*>(o)[1] = (px[2] & 0xFF) | ((px[2] >> 8) & 0xFF00) | ((px[3] & 0xFF) << 16) | ((px[3] & 0xFF0000) << 8);
        }
      } else {
        for (int k = 0; k < 4 && ox0 + k < P.dw; ++k) {
          o[k * CH] = (uint8_t)px[k];
          if (CH == 2) o[k * CH + 1] = (uint8_t)(px[k] >> 16);
        }
      }
    }
    __syncwarp();
  }
}

bool is_identity(const BandCuAffine& m) {
  return m.r0 == 0 && m.rr == 1 && m.rc == 0 && m.c0 == 0 && m.cr == 0 && m.cc == 1;
}

}  // namespace

// BANDCU_E_UNSUPPORTED: not a plane down-scale this file covers, or the source rows are not word-aligned.
static int launch_plane_up2(const BandCuPlanDesc& d, const BatchArgs& batch, const TableHints& hints, int n, cudaStream_t stream) {
  if (d.normalize || d.out_dtype != BANDCU_DT_U8) return BANDCU_E_UNSUPPORTED;
  if (!is_identity(d.luma_map) || !is_identity(d.pre_map) || !is_identity(d.chroma_map)) return BANDCU_E_UNSUPPORTED;
  int kind = -1, align = 1;
  if (d.src_kind == BANDCU_SRC_GRAY && d.cvt == BANDCU_CVT_NONE && d.channels == 1) kind = 0;
  else if (d.src_kind == BANDCU_SRC_UV && d.cvt == BANDCU_CVT_NONE && d.channels == 2) { kind = 1; align = 2; }
  else if (d.src_kind == BANDCU_SRC_RGBA && d.cvt == BANDCU_CVT_RGBA2GRAY && d.channels == 1) { kind = 2; align = 4; }
  else if (d.src_kind == BANDCU_SRC_RGB && d.cvt == BANDCU_CVT_RGB2GRAY && d.channels == 1) kind = 3;
  if (kind < 0) return BANDCU_E_UNSUPPORTED;
  if (align > 1) {
    if (d.stride0 & (align - 1)) return BANDCU_E_UNSUPPORTED;
    if (batch.table) {
      if (!hints.valid || !hints.planes_aligned4) return BANDCU_E_UNSUPPORTED;
    } else if ((reinterpret_cast<uintptr_t>(batch.first.p0) & (align - 1)) || (n > 1 && (batch.in_stride & (align - 1)))) {
      return BANDCU_E_UNSUPPORTED;
    }
  }
  PlaneUp2Params P;
  P.batch = batch;
  P.sw = d.vw; P.sh = d.vh; P.stride0 = d.stride0;
  P.dw = d.dst_width; P.dh = d.dst_height;
  P.dy = d.dy;
  P.bilinear = d.scale_mode == BANDCU_SCALE_UP2_BILINEAR ? 1 : 0;
  P.dst_row_stride = d.dst_row_stride;
  const size_t smem = (size_t)8 * (d.vw + 2) * sizeof(uint32_t);
  constexpr int kSmemCap = 160 * 1024;  // sources up to 5118 pixels wide
  if (smem > (size_t)kSmemCap) return BANDCU_E_UNSUPPORTED;
  int R = 4;  // consecutive output rows per warp share their two source rows through L1
  while (R > 1 && (long long)n * ((d.dst_height + 8 * R - 1) / (8 * R)) < 148LL * 16) R >>= 1;
  P.rows_per_warp = R;
  void (*kern)(PlaneUp2Params) = kind == 0 ? plane_up2_kernel<0> : kind == 1 ? plane_up2_kernel<1> : kind == 2 ? plane_up2_kernel<2> : plane_up2_kernel<3>;
  if (smem > 48 * 1024) {  // the limit is an attribute of the kernel: one value for every caller
    const cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemCap);
    if (e != cudaSuccess) return (int)e;
  }
  for (int done = 0; done < n;) {
    const int chunk = n - done > 65535 ? 65535 : n - done;
    PlaneUp2Params Q = P;
    if (Q.batch.table) {
      Q.batch.table += done;
    } else {
      Q.batch.first.p0 += (long long)done * batch.in_stride;
      Q.batch.first.out += (long long)done * batch.out_stride;
    }
    dim3 grid((d.dst_height + 8 * R - 1) / (8 * R), chunk, 1);
    kern<<<grid, 256, smem, stream>>>(Q);
    count_launch();
    done += chunk;
  }
  return (int)cudaGetLastError();
}

int launch_plane_rows(const BandCuPlanDesc& d, const BatchArgs& batch, const TableHints& hints, int n, cudaStream_t stream) {
  if (d.scale_mode != BANDCU_SCALE_BILINEAR_DOWN && d.scale_mode != BANDCU_SCALE_BOX2 && d.scale_mode != BANDCU_SCALE_UP2_BILINEAR &&
      d.scale_mode != BANDCU_SCALE_UP2_LINEAR)
    return BANDCU_E_UNSUPPORTED;
  static const bool off = getenv("BANDCU_DISABLE_PLANE_ROWS") && getenv("BANDCU_DISABLE_PLANE_ROWS")[0] == '1';
  if (off) return BANDCU_E_UNSUPPORTED;
  if (d.scale_mode == BANDCU_SCALE_UP2_BILINEAR || d.scale_mode == BANDCU_SCALE_UP2_LINEAR)
    return launch_plane_up2(d, batch, hints, n, stream);
  const bool gray = d.src_kind == BANDCU_SRC_GRAY && d.channels == 1 && d.flavor == BANDCU_FLAVOR_PLANE;
  const bool uv = d.src_kind == BANDCU_SRC_UV && d.channels == 2 && d.flavor == BANDCU_FLAVOR_ARGB;
  if (!gray && !uv) return BANDCU_E_UNSUPPORTED;
  if (d.cvt != BANDCU_CVT_NONE || d.normalize || d.out_dtype != BANDCU_DT_U8) return BANDCU_E_UNSUPPORTED;
  if (!is_identity(d.luma_map) || !is_identity(d.pre_map) || !is_identity(d.chroma_map)) return BANDCU_E_UNSUPPORTED;
  if (d.dx <= 0 || d.dy <= 0 || d.x0 < 0) return BANDCU_E_UNSUPPORTED;
  if ((long long)d.x0 + (long long)d.dst_width * d.dx > 0x7fffffffLL) return BANDCU_E_UNSUPPORTED;  // 32-bit column positions
  if (d.scale_mode == BANDCU_SCALE_BOX2 && (((d.dx | d.dy) & 0xffff) != 0)) return BANDCU_E_UNSUPPORTED;
  const int ch = gray ? 1 : 2;
  if (d.stride0 & 3) return BANDCU_E_UNSUPPORTED;
  if (batch.table) {
    if (!hints.valid || !hints.planes_aligned4) return BANDCU_E_UNSUPPORTED;
  } else {
    if ((reinterpret_cast<uintptr_t>(batch.first.p0) & 3) || (n > 1 && (batch.in_stride & 3))) return BANDCU_E_UNSUPPORTED;
  }
  PlaneRowsParams P;
  P.batch = batch;
  P.sw = d.vw; P.sh = d.vh; P.stride0 = d.stride0;
  P.dw = d.dst_width; P.dh = d.dst_height;
  P.x0 = d.x0; P.y0 = d.y0; P.dx = d.dx; P.dy = d.dy; P.max_y = d.max_y; P.linear = d.linear;
  P.box = d.scale_mode == BANDCU_SCALE_BOX2 ? 1 : 0;
  P.dst_row_stride = d.dst_row_stride;
  const int row_bytes = d.vw * ch;
  P.row_words = (row_bytes + 3) / 4;
  P.full_words = row_bytes / 4;
  const size_t smem = (size_t)8 * (P.box ? 2 : 1) * (P.row_words + 1) * sizeof(uint32_t);  // one staged row (pair) per warp
  if (smem > 48 * 1024) return BANDCU_E_UNSUPPORTED;
  // consecutive rows per warp (they share source rows through L1): 2 while that still leaves a few waves over 148 SMs
  int R = 2;
  while (R > 1 && (long long)n * ((d.dst_height + 8 * R - 1) / (8 * R)) < 148LL * 16) R >>= 1;
  P.rows_per_cta = R;
  void (*kern)(PlaneRowsParams) = gray ? plane_rows_down_kernel<1> : plane_rows_down_kernel<2>;
  for (int done = 0; done < n;) {
    const int chunk = n - done > 65535 ? 65535 : n - done;
    PlaneRowsParams Q = P;
    if (Q.batch.table) {
      Q.batch.table += done;
    } else {
      Q.batch.first.p0 += (long long)done * batch.in_stride;
      Q.batch.first.out += (long long)done * batch.out_stride;
    }
    dim3 grid((d.dst_height + 8 * R - 1) / (8 * R), chunk, 1);
    kern<<<grid, 256, smem, stream>>>(Q);
    count_launch();
    done += chunk;
  }
  return (int)cudaGetLastError();
}

}  // namespace bandcu
