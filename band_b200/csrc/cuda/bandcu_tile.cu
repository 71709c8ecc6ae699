// band_b200/csrc/cuda/bandcu_tile.cu
//
// Tile-staged kernel for 4:2:0 -> RGB -> bilinear down-scale [-> Normalize] behind ANY composed Crop / Rotate / Flip
// index map and for any 4:2:0 layout (planar YV12/YV21 or semi-planar NV12/NV21): BASELINE config C3
// (4K YV12 -> Rotate 90 -> Crop -> ColorSpaceConvert -> Resize 512x512 -> Normalize) and every 4:2:0 chain the
// row-staged kernel of bandcu_fast.cu cannot take (planar chroma, index maps).
//
// Replaces, fused: RotateYv / RotateNv (libyuv_image_operator.cc:786-887), CropYv / CropNv (:950-1037), Flip*
// (:1232-1363), ConvertFromYv / ConvertFromNv12 / ConvertFromNv21 -> kRGB (:269-300, :92-111, :180-199), ResizeRgb
// (:1146-1187, ARGBScale general bilinear-down path) and Normalize (common_operator.cc:24-40).
//
// Why tiles: with a quarter-turn in the index map an output ROW walks a source COLUMN. The pixel-per-thread generic
// kernel then gathers single bytes from rows 3840 bytes apart (ncu: 0.39 of the HBM roofline on C3, L1/L2 gather
// bound). Here a CTA owns a TR x TC output tile. The index maps are affine, so the source samples of a tile lie in
// an axis-aligned rectangle of the luma plane and one of each chroma plane; the CTA copies those rectangles into
// shared memory with fully coalesced 32-bit loads (every fetched sector is used), then each thread evaluates its
// output pixels from shared memory with the packed LEFT|RIGHT 16-bit-lane arithmetic of the row-staged kernel and
// stores with lanes along the output row (384 contiguous bytes per warp for f32). Affine maps turn into three
// integers per plane: address = A + B*r + C*c.
#include <cstdlib>
#include <cstring>
#include <vector>

#include "bandcu_device.cuh"
#include "bandcu_internal.h"

namespace bandcu {

// Source rectangles of one output tile (the same for every frame of a plan): computed once on the host at plan creation
// instead of by every CTA (three affine bounding boxes: ~170 of the ~1060 instructions a warp spent per tile).
struct __align__(16) TileEntry {
  int l_r0, l_rows, l_first;  // luma: first source row, rows, first source byte of a row
  int c_r0, c_rows, c_first;  // chroma plane(s): the same, in plane bytes
  int pad[2];
};

struct TileParams {
  BandCuPlanDesc d;
  BatchArgs batch;
  const TileEntry* geom;  // [tiles_y][tiles_x]
  int TR, TC;          // output rows / columns per tile
  int tiles_x, tiles_y;
  int interleaved;     // 1: semi-planar chroma (one staged plane, U/V one byte apart)
  int v_first;         // interleaved: V is the first byte of a pair (NV21)
  int lut_f32;         // output goes through the Normalize table
  const float* lut;
  int l_pitch, c_pitch;  // staged row pitches (bytes), the worst case over all tiles: launch constants
  int l_rows_max, c_rows_max;  // staged rows, worst case (plane offsets in shared memory are launch constants too)
  // Crop / Rotate / Flip compose into SIGNED PERMUTATIONS, so every mapped coordinate depends on the virtual row
  // alone or on the virtual column alone and a shared-memory offset splits into f(row) + g(col):
  //   luma(r,c)   = AL(tile) + lr_mul * r + lc_mul * c
  //   chroma(r,c) = AC(tile) + cr_mul * ((cr_0 + cr_s * r) >> 1) + cc_mul * ((cc_0 + cc_s * c) >> 1)
  int lr_mul, lc_mul;
  int cr_0, cr_s, cr_mul, cc_0, cc_s, cc_mul;
  int out_vec;         // output rows allow 128-bit (f32) / 32-bit (u8) stores
  // skip != 0: the luma plane is staged per SAMPLE, not per rectangle: slot 2i+k holds the source row of the k-th
  // (top/left, bottom/right) sample of the tile's i-th position along the axis that maps to source rows. A down-scale
  // by s >= 2 along that axis touches only 2 of every s source rows: C3 stages 64 luma rows per tile instead of 134.
  int skip;
  int l_row_0, l_row_mul;  // source luma row = l_row_0 + l_row_mul * (virtual index along the row axis)
  // skip_c: the same for the chroma plane(s), once the ratio along that axis reaches 6 (chroma is half resolution):
  // source chroma row = c_row_0 + c_row_mul * ((q_0 + q_s * virtual index) >> 1), (q_0, q_s) = that axis's pre-map terms
  int skip_c;
  int c_row_0, c_row_mul;
  int l_lpr_shift, c_lpr_shift;  // log2(threads serving one staged row) for the luma / chroma planes
  int off_terms;       // offset of the per-tile term tables in dynamic shared memory
};


struct Rect {
  int r0, r1, c0, c1;  // inclusive
};

__host__ __device__ __forceinline__ int t_map_row(const BandCuAffine& m, int r, int c) { return m.r0 + m.rr * r + m.rc * c; }
__host__ __device__ __forceinline__ int t_map_col(const BandCuAffine& m, int r, int c) { return m.c0 + m.cr * r + m.cc * c; }

// Bounding box of an affine map over the corners of [ra,rb] x [ca,cb] (monotone in each coordinate).
__host__ __device__ __forceinline__ Rect map_rect(const BandCuAffine& m, int ra, int rb, int ca, int cb) {
  const int r00 = t_map_row(m, ra, ca), r01 = t_map_row(m, ra, cb), r10 = t_map_row(m, rb, ca), r11 = t_map_row(m, rb, cb);
  const int c00 = t_map_col(m, ra, ca), c01 = t_map_col(m, ra, cb), c10 = t_map_col(m, rb, ca), c11 = t_map_col(m, rb, cb);
  Rect q;
  q.r0 = min(min(r00, r01), min(r10, r11));
  q.r1 = max(max(r00, r01), max(r10, r11));
  q.c0 = min(min(c00, c01), min(c10, c11));
  q.c1 = max(max(c00, c01), max(c10, c11));
  return q;
}

// Virtual-image rectangle (clamped like fetch_px) sampled by output rows [oy0,oy1] x columns [ox0,ox1].
__host__ __device__ __forceinline__ Rect virtual_rect(const BandCuPlanDesc& d, int oy0, int oy1, int ox0, int ox1) {
  long long ya = (long long)d.y0 + (long long)oy0 * d.dy, yb = (long long)d.y0 + (long long)oy1 * d.dy;
  if (ya > d.max_y) ya = d.max_y;
  if (yb > d.max_y) yb = d.max_y;
  const long long xa = (long long)d.x0 + (long long)ox0 * d.dx, xb = (long long)d.x0 + (long long)ox1 * d.dx;
  Rect v;
  v.r0 = min(max((int)(ya >> 16), 0), d.vh - 1);
  v.r1 = min(max((int)(yb >> 16) + 1, 0), d.vh - 1);
  v.c0 = min(max((int)(xa >> 16), 0), d.vw - 1);
  v.c1 = min(max((int)(xb >> 16) + 1, 0), d.vw - 1);
  return v;
}

// The three rectangles of a tile: luma (plane 0 coordinates), chroma (chroma-plane sample coordinates).
__host__ __device__ __forceinline__ void tile_rects(const BandCuPlanDesc& d, const Rect& v, Rect* luma, Rect* chroma) {
  *luma = map_rect(d.luma_map, v.r0, v.r1, v.c0, v.c1);
  const Rect p = map_rect(d.pre_map, v.r0, v.r1, v.c0, v.c1);  // where the colour conversion saw these pixels
  *chroma = map_rect(d.chroma_map, p.r0 >> 1, p.r1 >> 1, p.c0 >> 1, p.c1 >> 1);
}

// Staged plane geometry. A staged row holds the plane bytes [first_byte & ~(G-1), last_byte | (G-1)] (G = staging
// granule: 16 bytes when every plane row starts 16-byte aligned, else 4) at a pitch of an ODD number of granules, so
// lanes that walk source ROWS (as they do under a quarter-turn) spread over the banks.
template <int G>
__host__ __device__ __forceinline__ int staged_pitch(int first_byte, int last_byte) {
  const int granules = ((last_byte / G) - (first_byte / G) + 1) | 1;
  return granules * G;
}

// G == 16: cp.async (LDGSTS) 16-byte copies, no register staging -- a thread's copies are all in flight at once.
// G == 4 : 32-bit loads through registers (any word-aligned layout).
// 2^lpr_shift threads serve one staged row (the next power of two >= granules per row, a launch constant), so the
// loop has no division: (source, destination) advance by a fixed number of rows per pass.
// Per-sample luma staging (skip mode): staged row `slot` comes from source row row_of(slot).
template <int G, typename RowOf>
__device__ __forceinline__ void stage_rows(uint8_t* smem_dst, const uint8_t* __restrict__ plane, int stride, int nslots,
                                           int first_byte, int pitch, int lpr_shift, int row_bytes, RowOf row_of) {
  const int gpr = min(pitch / G, (row_bytes - (first_byte / G) * G + G - 1) / G);
  const int lpr = 1 << lpr_shift;
  const int rows_per_pass = (int)blockDim.x >> lpr_shift;
  const int g0 = (int)threadIdx.x & (lpr - 1);
  if (g0 >= gpr) return;
  const uint8_t* col0 = plane + (first_byte / G) * G + g0 * G;
  const uint32_t dst0 = (uint32_t)__cvta_generic_to_shared(smem_dst) + (uint32_t)(g0 * G);
  for (int r = (int)threadIdx.x >> lpr_shift; r < nslots; r += rows_per_pass) {
    const uint8_t* src = col0 + (long long)row_of(r) * stride;
    const uint32_t dst = dst0 + (uint32_t)(r * pitch);
    for (int g = g0, o = 0; g < gpr; g += lpr, o += lpr * G) {
      if (G == 16) {
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + o), "l"(src + o) : "memory");
      } else {
        const uint32_t v = __ldg(reinterpret_cast<const uint32_t*>(src + o));
        asm volatile("st.shared.u32 [%0], %1;" ::"r"(dst + o), "r"(v) : "memory");
      }
    }
  }
}

template <int G>
__device__ __forceinline__ void stage_plane(uint8_t* smem_dst, const uint8_t* __restrict__ plane, int stride,
                                            int row0, int nrows, int first_byte, int pitch, int lpr_shift,
                                            int row_bytes) {
  // granules per staged row: the launch-wide pitch, but never past the end of the plane row (the pitch is the worst
  // case over all tiles and odd; a tile at the right edge must not read into memory behind the last row)
  const int gpr = min(pitch / G, (row_bytes - (first_byte / G) * G + G - 1) / G);
  const int lpr = 1 << lpr_shift;
  const int rows_per_pass = (int)blockDim.x >> lpr_shift;
  int r = (int)threadIdx.x >> lpr_shift;
  const int g0 = (int)threadIdx.x & (lpr - 1);
  if (g0 >= gpr) return;
  const uint8_t* src = plane + (long long)(row0 + r) * stride + (first_byte / G) * G + g0 * G;
  uint32_t dst = (uint32_t)__cvta_generic_to_shared(smem_dst) + (uint32_t)(r * pitch + g0 * G);
  const long long src_step = (long long)rows_per_pass * stride;
  const uint32_t dst_step = (uint32_t)(rows_per_pass * pitch);
  for (; r < nrows; r += rows_per_pass, src += src_step, dst += dst_step) {
    for (int g = g0, o = 0; g < gpr; g += lpr, o += lpr * G) {  // one trip unless a row has more than 32 granules
      if (G == 16) {
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + o), "l"(src + o) : "memory");
      } else {
        const uint32_t v = __ldg(reinterpret_cast<const uint32_t*>(src + o));
        asm volatile("st.shared.u32 [%0], %1;" ::"r"(dst + o), "r"(v) : "memory");
      }
    }
  }
}

// ---- packed LEFT|RIGHT pixel-pair arithmetic (same identities as bandcu_fast.cu; values instead of addresses) ----
struct TChroma {
  uint32_t kr, kg, kb;
};
__device__ __forceinline__ TChroma t_chroma_terms(uint32_t u2, uint32_t v2) {
  TChroma c;
  c.kr = __vadd2(v2 * 102u, 0xC878C878u /* -14216 */);
  const uint32_t t = u2 * 25u + v2 * 52u;
  c.kg = __vadd2(~t, 0x21F921F9u /* 8696 + 1 */);
  c.kb = u2 << 7;
  return c;
}
__device__ __forceinline__ void t_luma_pair_to_rgb(uint32_t yL, uint32_t yR, const TChroma& c, uint32_t& R2, uint32_t& G2,
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
__device__ __forceinline__ uint32_t t_vblend2(uint32_t top, uint32_t bot, uint32_t w0, uint32_t w1) {
  // bytes 1 and 3 of the sums are the two results: one PRMT instead of shift + mask
  return __byte_perm(top * w0 + bot * w1 + 0x00800080u, 0u, 0x4341);
}

struct __align__(16) RowTerms {   // everything an output ROW contributes (32 bytes)
  int lA, lB;       // luma offsets of virtual rows rA, rB
  int kA, kB;       // chroma offsets
  uint32_t w0, w1;  // InterpolateRow weights
  int pad[2];
};
struct __align__(16) ColTerms {   // everything an output COLUMN contributes (32 bytes)
  int lA, lB, kA, kB;
  uint32_t wp;      // packed 7-bit column weights
  int pad[3];
};
// Clamped virtual rows (columns) sampled by output row oy (column ox): fetch_px's clamps.
__device__ __forceinline__ void virt_rows(const BandCuPlanDesc& d, int oy, int* rA, int* rB) {
  long long y = (long long)d.y0 + (long long)oy * d.dy;
  if (y > d.max_y) y = d.max_y;
  const int yi = (int)(y >> 16);
  *rA = min(max(yi, 0), d.vh - 1);
  *rB = min(max(yi + 1, 0), d.vh - 1);
}
__device__ __forceinline__ void virt_cols(const BandCuPlanDesc& d, int ox, int* cA, int* cB) {
  const int xi = (int)(((long long)d.x0 + (long long)ox * d.dx) >> 16);
  *cA = min(max(xi, 0), d.vw - 1);
  *cB = min(max(xi + 1, 0), d.vw - 1);
}
// `slot_pitch` != 0: this axis selects the staged luma row (skip mode): offsets are slot * pitch for tile position i.
__device__ __forceinline__ RowTerms row_terms(const TileParams& P, int oy, int i, int slot_pitch, int slot_pitch_c) {
  const BandCuPlanDesc& d = P.d;
  long long y = (long long)d.y0 + (long long)oy * d.dy;
  if (y > d.max_y) y = d.max_y;
  const int yi = (int)(y >> 16);
  RowTerms t;
  t.w1 = d.linear ? 0u : (uint32_t)((y >> 8) & 255);
  t.w0 = 256u - t.w1;
  const int rA = min(max(yi, 0), d.vh - 1), rB = min(max(yi + 1, 0), d.vh - 1);
  t.lA = slot_pitch ? (2 * i) * slot_pitch : P.lr_mul * rA;
  t.lB = slot_pitch ? (2 * i + 1) * slot_pitch : P.lr_mul * rB;
  t.kA = slot_pitch_c ? (2 * i) * slot_pitch_c : P.cr_mul * ((P.cr_0 + P.cr_s * rA) >> 1);
  t.kB = slot_pitch_c ? (2 * i + 1) * slot_pitch_c : P.cr_mul * ((P.cr_0 + P.cr_s * rB) >> 1);
  return t;
}
__device__ __forceinline__ ColTerms col_terms(const TileParams& P, int ox, int i, int slot_pitch, int slot_pitch_c) {
  const BandCuPlanDesc& d = P.d;
  const long long x = (long long)d.x0 + (long long)ox * d.dx;
  const int xi = (int)(x >> 16);
  const uint32_t xf = (uint32_t)((x >> 9) & 0x7f);
  ColTerms t;
  t.wp = xf | ((0x7fu ^ xf) << 16);
  const int cA = min(max(xi, 0), d.vw - 1), cB = min(max(xi + 1, 0), d.vw - 1);
  t.lA = slot_pitch ? (2 * i) * slot_pitch : P.lc_mul * cA;
  t.lB = slot_pitch ? (2 * i + 1) * slot_pitch : P.lc_mul * cB;
  t.kA = slot_pitch_c ? (2 * i) * slot_pitch_c : P.cc_mul * ((P.cc_0 + P.cc_s * cA) >> 1);
  t.kB = slot_pitch_c ? (2 * i + 1) * slot_pitch_c : P.cc_mul * ((P.cc_0 + P.cc_s * cB) >> 1);
  return t;
}
// Chroma of the LEFT|RIGHT sample pair at byte offsets oL, oR. Planar: four byte loads. Interleaved (IL): the staged
// plane holds (first, second) byte pairs at even offsets -- one 16-bit load per sample, and two PRMTs (selectors sel_u,
// sel_v pick U and V for NV12 or NV21) build the packed lanes.
template <bool IL>
__device__ __forceinline__ TChroma t_chroma_at(const uint8_t* su, const uint8_t* sv, int oL, int oR, uint32_t sel_u,
                                               uint32_t sel_v) {
  if (IL) {
    const uint32_t wL = *reinterpret_cast<const uint16_t*>(su + oL), wR = *reinterpret_cast<const uint16_t*>(su + oR);
    return t_chroma_terms(__byte_perm(wL, wR, sel_u), __byte_perm(wL, wR, sel_v));
  }
  return t_chroma_terms(su[oL] | ((uint32_t)su[oR] << 16), sv[oL] | ((uint32_t)sv[oR] << 16));
}
// One output pixel from the staged planes: raw products, the 8-bit results sit in bits 23..30.
template <bool IL>
__device__ __forceinline__ void tile_pixel(const uint8_t* sl, const uint8_t* su, const uint8_t* sv, uint32_t sel_u,
                                           uint32_t sel_v, const RowTerms& R, const ColTerms& Cn, uint32_t& r, uint32_t& g,
                                           uint32_t& b) {
  uint32_t tr, tg, tb, br, bg, bb;
  {
    const TChroma ct = t_chroma_at<IL>(su, sv, R.kA + Cn.kA, R.kA + Cn.kB, sel_u, sel_v);
    t_luma_pair_to_rgb(sl[R.lA + Cn.lA], sl[R.lA + Cn.lB], ct, tr, tg, tb);
  }
  if (R.w1) {
    const TChroma cb = t_chroma_at<IL>(su, sv, R.kB + Cn.kA, R.kB + Cn.kB, sel_u, sel_v);
    t_luma_pair_to_rgb(sl[R.lB + Cn.lA], sl[R.lB + Cn.lB], cb, br, bg, bb);
  } else {
    br = tr; bg = tg; bb = tb;
  }
  r = t_vblend2(tr, br, R.w0, R.w1) * Cn.wp;
  g = t_vblend2(tg, bg, R.w0, R.w1) * Cn.wp;
  b = t_vblend2(tb, bb, R.w0, R.w1) * Cn.wp;
}

// SWAP: the luma map exchanges the axes (quarter-turns). Lanes then run along the output COLUMN so that a warp's
// shared-memory reads walk consecutive source bytes (lanes along the output row would stride by whole staged rows:
// 4-way bank conflicts measured, L1 at 84 %); each thread produces FOUR adjacent pixels of its output row and stores
// them as 48 contiguous bytes (three 128-bit stores for f32). Without SWAP lanes run along the output row and a
// warp stores 384 contiguous bytes per row.
// The row and column terms of the tile (32 + 32 entries) are computed once by two warps while the cp.async copies
// are in flight and read back as broadcasts.
template <typename OT, int G, bool SWAP, bool IL>
__global__ void __launch_bounds__(256) tile_yuv_bilinear_down_kernel(const __grid_constant__ TileParams P) {
  extern __shared__ __align__(16) uint8_t smem[];
  __shared__ float lut[256];
  const BandCuPlanDesc& d = P.d;
  // f32 outputs: the Normalize table; one-byte Normalize outputs (int8 / uint8 models): 32-bit entries holding the byte
  if (sizeof(OT) == 4 || P.lut) lut[threadIdx.x] = P.lut[threadIdx.x];  // blockDim.x == 256
  const int oy0 = blockIdx.y * P.TR, ox0 = blockIdx.x * P.TC;
  const int oy1 = min(oy0 + P.TR, d.dst_height) - 1, ox1 = min(ox0 + P.TC, d.dst_width) - 1;
  const FramePtrs f = frame_ptrs(P.batch, blockIdx.z);

  const TileEntry te = P.geom[blockIdx.y * P.tiles_x + blockIdx.x];
  const int ps = d.pixel_stride_uv;
  const int l_first = te.l_first, c_first = te.c_first;
  const int l_rows = te.l_rows, c_rows = te.c_rows;
  const int c_row_bytes = ((d.src_width + 1) >> 1) * ps;  // bytes of one chroma plane row
  uint8_t* sl = smem;
  uint8_t* su = sl + P.l_rows_max * P.l_pitch;
  uint8_t* sv;
  if (P.skip) {
    const int npos = SWAP ? (ox1 - ox0 + 1) : (oy1 - oy0 + 1);
    stage_rows<G>(sl, f.p0, d.stride0, 2 * npos, l_first, P.l_pitch, P.l_lpr_shift, d.src_width, [&](int slot) {
      int a, b;
      if (SWAP) virt_cols(d, ox0 + (slot >> 1), &a, &b);
      else virt_rows(d, oy0 + (slot >> 1), &a, &b);
      return P.l_row_0 + P.l_row_mul * ((slot & 1) ? b : a);
    });
  } else {
    stage_plane<G>(sl, f.p0, d.stride0, te.l_r0, l_rows, l_first, P.l_pitch, P.l_lpr_shift, d.src_width);
  }
  const int npos_c = SWAP ? (ox1 - ox0 + 1) : (oy1 - oy0 + 1);
  auto chroma_row_of = [&](int slot) {
    int a, b, q;
    if (SWAP) { virt_cols(d, ox0 + (slot >> 1), &a, &b); q = (P.cc_0 + P.cc_s * ((slot & 1) ? b : a)) >> 1; }
    else { virt_rows(d, oy0 + (slot >> 1), &a, &b); q = (P.cr_0 + P.cr_s * ((slot & 1) ? b : a)) >> 1; }
    return P.c_row_0 + P.c_row_mul * q;
  };
  auto stage_chroma = [&](uint8_t* dst, const uint8_t* plane) {
    if (P.skip_c) stage_rows<G>(dst, plane, d.stride_uv, 2 * npos_c, c_first, P.c_pitch, P.c_lpr_shift, c_row_bytes, chroma_row_of);
    else stage_plane<G>(dst, plane, d.stride_uv, te.c_r0, c_rows, c_first, P.c_pitch, P.c_lpr_shift, c_row_bytes);
  };
  // IL: su stays the base of the (first, second) byte pairs; byte 2 of a zero-extended 16-bit load is the zero lane filler
  const uint32_t sel_u = P.v_first ? 0x2521u : 0x2420u, sel_v = P.v_first ? 0x2420u : 0x2521u;
  if (IL) {
    stage_chroma(su, P.v_first ? f.pv : f.pu);
    sv = su;
  } else {
    sv = su + P.c_rows_max * P.c_pitch;
    stage_chroma(su, f.pu);
    stage_chroma(sv, f.pv);
  }
  if (G == 16) asm volatile("cp.async.commit_group;" ::: "memory");

  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;  // 8 warps
  RowTerms* rt = reinterpret_cast<RowTerms*>(smem + P.off_terms);        // [32], 32 bytes each
  ColTerms* ct = reinterpret_cast<ColTerms*>(smem + P.off_terms + 1024);  // [32], 32 bytes each
  if (wid == 0) rt[lane] = row_terms(P, min(oy0 + lane, oy1), min(lane, oy1 - oy0), (P.skip && !SWAP) ? P.l_pitch : 0,
                                 (P.skip_c && !SWAP) ? P.c_pitch : 0);
  else if (wid == 1) ct[lane] = col_terms(P, min(ox0 + lane, ox1), min(lane, ox1 - ox0), (P.skip && SWAP) ? P.l_pitch : 0,
                                      (P.skip_c && SWAP) ? P.c_pitch : 0);
  if (G == 16) asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncthreads();

  // tile-dependent constants of the two affine address forms
  const int AL = (P.skip ? 0 : (d.luma_map.r0 - te.l_r0) * P.l_pitch) + d.luma_map.c0 - (l_first / G) * G;
  const int AC = (P.skip_c ? 0 : (d.chroma_map.r0 - te.c_r0) * P.c_pitch) + d.chroma_map.c0 * ps - (c_first / G) * G;
  const uint8_t* gl = sl + AL;
  const uint8_t* gu = su + AC;
  const uint8_t* gv = sv + AC;

  // The 8-bit results of the tile go through a [32][25]-word shared array (odd pitch: the compute layout's writes and
  // the row-major reads are both conflict-free) and leave as row-major 16-byte (f32) / 4-byte (u8) pieces: a warp
  // writes 1.3 output rows of the tile per store instruction. (Storing from the quarter-turn compute layout -- each
  // lane its own output row -- cost 32 L1 wavefronts per store instruction, a third of the kernel's L1 data-pipe
  // load: profiles/r01_ncu_full_tile_v6_uncoalesced_stores.txt.)
  uint32_t* os = reinterpret_cast<uint32_t*>(smem + P.off_terms + 2048);
  if (!SWAP) {
    // lane = output column inside the tile (TC == 32); warp w owns rows w, w+8, ...
    if (ox0 + lane <= ox1) {
      const ColTerms Cn = ct[lane];
      uint8_t* ob = reinterpret_cast<uint8_t*>(os) + lane * 3;
      for (int ry = wid; oy0 + ry <= oy1; ry += 8) {
        const RowTerms R = rt[ry];
        uint32_t r, g, b;
        tile_pixel<IL>(gl, gu, gv, sel_u, sel_v, R, Cn, r, g, b);
        ob[ry * 100] = (uint8_t)(r >> 23); ob[ry * 100 + 1] = (uint8_t)(g >> 23); ob[ry * 100 + 2] = (uint8_t)(b >> 23);
      }
    }
  } else {
    // lane = output row inside the tile (TR == 32); warp w owns columns 4w .. 4w+3
    const int oy = oy0 + lane;
    const int cx0 = wid * 4;
    if (oy <= oy1 && ox0 + cx0 <= ox1) {
      const RowTerms R = rt[lane];
      uint32_t px[12];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const ColTerms Cn = ct[min(cx0 + j, 31)];
        uint32_t r, g, b;
        tile_pixel<IL>(gl, gu, gv, sel_u, sel_v, R, Cn, r, g, b);
        px[3 * j] = (r >> 23) & 0xFF; px[3 * j + 1] = (g >> 23) & 0xFF; px[3 * j + 2] = (b >> 23) & 0xFF;
      }
#pragma unroll
      for (int q = 0; q < 3; ++q)
        os[lane * 25 + wid * 3 + q] = px[4 * q] | (px[4 * q + 1] << 8) | (px[4 * q + 2] << 16) | (px[4 * q + 3] << 24);
    }
  }
  __syncthreads();
  const int nrows = oy1 - oy0 + 1, ncols = ox1 - ox0 + 1;
  uint8_t* obase = f.out + (long long)oy0 * d.dst_row_stride + (size_t)ox0 * (3 * sizeof(OT));
  if (P.out_vec && ncols == 32) {  // 24 aligned pieces per row (ox0 is a multiple of 32)
    for (int idx = threadIdx.x; idx < nrows * 24; idx += 256) {
      const int row = idx / 24, q = idx - row * 24;
      uint32_t w = os[row * 25 + q];
      uint8_t* o = obase + (long long)row * d.dst_row_stride;
      if (sizeof(OT) == 1 && P.lut)
        w = (__float_as_uint(lut[w & 0xFF]) & 0xFF) | ((__float_as_uint(lut[(w >> 8) & 0xFF]) & 0xFF) << 8) |
            ((__float_as_uint(lut[(w >> 16) & 0xFF]) & 0xFF) << 16) | (__float_as_uint(lut[w >> 24]) << 24);
      if (sizeof(OT) == 1) reinterpret_cast<uint32_t*>(o)[q] = w;
      else reinterpret_cast<float4*>(o)[q] = make_float4(lut[w & 0xFF], lut[(w >> 8) & 0xFF], lut[(w >> 16) & 0xFF], lut[w >> 24]);
    }
  } else {  // right-edge tiles, unaligned outputs: element by element, still row-major
    const uint8_t* ob = reinterpret_cast<const uint8_t*>(os);
    for (int e = threadIdx.x; e < nrows * 96; e += 256) {
      const int row = e / 96, c = e - row * 96;
      if (c >= 3 * ncols) continue;
      const uint32_t v = ob[row * 100 + c];
      uint8_t* o = obase + (long long)row * d.dst_row_stride;
      if (sizeof(OT) == 1) o[c] = P.lut ? (uint8_t)__float_as_uint(lut[v]) : (uint8_t)v;
      else reinterpret_cast<float*>(o)[c] = lut[v];
    }
  }
}

template <typename T>
__global__ void tile_build_byte_lut_kernel(uint32_t* lut, float mean, float sd) {
  const int i = threadIdx.x;
  if (i < 256) lut[i] = (uint32_t)(uint8_t)normalize_elem<T>((float)i, mean, sd);
}
__global__ void tile_build_lut_kernel(float* lut, float mean, float sd) {
  const int i = threadIdx.x;
  if (i < 256) lut[i] = normalize_elem<float>((float)i, mean, sd);
}

// Worst-case staged geometry of a TR x TC tiling (evaluated on the host over every tile: a few thousand at most).
struct TileGeom {
  int l_pitch, c_pitch, l_rows, c_rows;
  size_t stage_bytes;  // luma + chroma plane(s)
};
template <int G>
static TileGeom tile_geometry(const BandCuPlanDesc& d, int TR, int TC, bool interleaved, bool skip, bool skip_c) {
  const int tiles_x = (d.dst_width + TC - 1) / TC, tiles_y = (d.dst_height + TR - 1) / TR;
  TileGeom g = {0, 0, 0, 0, 0};
  const int ps = d.pixel_stride_uv;
  for (int ty = 0; ty < tiles_y; ++ty) {
    for (int tx = 0; tx < tiles_x; ++tx) {
      const int oy0 = ty * TR, ox0 = tx * TC;
      const int oy1 = (oy0 + TR < d.dst_height ? oy0 + TR : d.dst_height) - 1;
      const int ox1 = (ox0 + TC < d.dst_width ? ox0 + TC : d.dst_width) - 1;
      const Rect v = virtual_rect(d, oy0, oy1, ox0, ox1);
      Rect L, C;
      tile_rects(d, v, &L, &C);
      g.l_pitch = max(g.l_pitch, staged_pitch<G>(L.c0, L.c1));
      g.c_pitch = max(g.c_pitch, staged_pitch<G>(C.c0 * ps, C.c1 * ps + (interleaved ? 1 : 0)));
      g.l_rows = max(g.l_rows, L.r1 - L.r0 + 1);
      g.c_rows = max(g.c_rows, C.r1 - C.r0 + 1);
    }
  }
  if (skip) g.l_rows = 2 * (d.luma_map.rr != 0 ? TR : TC);  // two sampled source rows per tile position
  if (skip_c) g.c_rows = 2 * (d.luma_map.rr != 0 ? TR : TC);
  g.stage_bytes = (size_t)g.l_rows * g.l_pitch + (size_t)(interleaved ? 1 : 2) * g.c_rows * g.c_pitch;
  g.stage_bytes = (g.stage_bytes + 15) / 16 * 16;
  return g;
}

struct TilePlan {
  float* lut_dev;
  TileEntry* geom_dev;
  int TR, TC;
  TileGeom g16, g4;
  bool skip, skip_c;
};

static bool is_permutation(const BandCuAffine& m) {
  const bool straight = m.rr != 0 && m.cc != 0 && m.rc == 0 && m.cr == 0;
  const bool swapped = m.rr == 0 && m.cc == 0 && m.rc != 0 && m.cr != 0;
  return straight || swapped;
}

bool tile_plan_supported(const BandCuPlanDesc& d) {
  const char* off = getenv("BANDCU_DISABLE_TILE");
  if (off && off[0] == '1') return false;
  if (d.src_kind != BANDCU_SRC_YUV420 || d.cvt != BANDCU_CVT_YUV2RGB || d.channels != 3) return false;
  if (d.flavor != BANDCU_FLAVOR_ARGB || d.scale_mode != BANDCU_SCALE_BILINEAR_DOWN) return false;
  if (d.pixel_stride_uv != 1 && d.pixel_stride_uv != 2) return false;
  if (d.normalize ? (d.out_dtype != BANDCU_DT_F32 && d.out_dtype != BANDCU_DT_I8 && d.out_dtype != BANDCU_DT_U8)
                  : d.out_dtype != BANDCU_DT_U8)
    return false;
  if (!is_permutation(d.luma_map) || !is_permutation(d.pre_map) || !is_permutation(d.chroma_map)) return false;
  // 32-bit word staging: rows must start word aligned and whole words must stay inside the plane rows
  if (d.src_width <= 0 || d.src_width % 4 != 0 || d.stride0 % 4 != 0 || d.stride_uv % 4 != 0) return false;
  if ((((d.src_width + 1) / 2) * d.pixel_stride_uv) % 4 != 0) return false;
  return true;
}

int tile_plan_create(const BandCuPlanDesc& d, TilePlan** out) {
  const bool inter = d.pixel_stride_uv == 2;
  const int TR = 32, TC = 32;  // lanes run along one tile edge: both stay 32
  // per-sample luma staging pays once the down-scale along the axis that maps to source rows reaches 2
  const char* ns = getenv("BANDCU_TILE_NO_SKIP");
  const bool skip = !(ns && ns[0] == '1') && (d.luma_map.rr != 0 ? d.dy : d.dx) >= 0x20000;
  // chroma too once the ratio reaches 6 (below that nearly every half-resolution row is sampled and the per-slot row
  // arithmetic costs more than it saves: measured at 4.8), if chroma rows follow the same virtual axis as luma rows (they do for every
  // chain the plan compiler emits; checked, not assumed)
  const bool luma_rows_from_vrows = d.luma_map.rr != 0;
  const bool chroma_rows_from_vrows = d.pre_map.rr != 0 ? d.chroma_map.rr != 0 : d.chroma_map.rc != 0;
  const bool skip_c = skip && luma_rows_from_vrows == chroma_rows_from_vrows && (luma_rows_from_vrows ? d.dy : d.dx) >= 0x60000;
  const TileGeom g16 = tile_geometry<16>(d, TR, TC, inter, skip, skip_c), g4 = tile_geometry<4>(d, TR, TC, inter, skip, skip_c);
  if (g16.stage_bytes + 2048 + 3200 > 100 * 1024) return BANDCU_E_UNSUPPORTED;  // very large ratios: keep >= 2 CTAs per SM
  TilePlan* p = new TilePlan();
  p->lut_dev = nullptr;
  p->geom_dev = nullptr;
  {
    const int tiles_x = (d.dst_width + TC - 1) / TC, tiles_y = (d.dst_height + TR - 1) / TR;
    std::vector<TileEntry> geom((size_t)tiles_x * tiles_y);
    for (int ty = 0; ty < tiles_y; ++ty) {
      for (int tx = 0; tx < tiles_x; ++tx) {
        const int oy0 = ty * TR, ox0 = tx * TC;
        const int oy1 = (oy0 + TR < d.dst_height ? oy0 + TR : d.dst_height) - 1;
        const int ox1 = (ox0 + TC < d.dst_width ? ox0 + TC : d.dst_width) - 1;
        Rect L, C;
        tile_rects(d, virtual_rect(d, oy0, oy1, ox0, ox1), &L, &C);
        TileEntry& e = geom[(size_t)ty * tiles_x + tx];
        memset(&e, 0, sizeof(e));
        e.l_r0 = L.r0; e.l_rows = L.r1 - L.r0 + 1; e.l_first = L.c0;
        e.c_r0 = C.r0; e.c_rows = C.r1 - C.r0 + 1; e.c_first = C.c0 * d.pixel_stride_uv;
      }
    }
    cudaError_t e = cudaMalloc((void**)&p->geom_dev, geom.size() * sizeof(TileEntry));
    if (e == cudaSuccess) e = cudaMemcpy(p->geom_dev, geom.data(), geom.size() * sizeof(TileEntry), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) { if (p->geom_dev) cudaFree(p->geom_dev); delete p; return (int)e; }
  }
  p->TR = TR; p->TC = TC;
  p->g16 = g16; p->g4 = g4;
  p->skip = skip;
  p->skip_c = skip_c;
  if (d.normalize) {
    cudaError_t e = cudaMalloc((void**)&p->lut_dev, 256 * sizeof(float));
    if (e != cudaSuccess) { cudaFree(p->geom_dev); delete p; return (int)e; }
    if (d.out_dtype == BANDCU_DT_I8) tile_build_byte_lut_kernel<int8_t><<<1, 256>>>((uint32_t*)p->lut_dev, d.mean, d.std_);
    else if (d.out_dtype == BANDCU_DT_U8) tile_build_byte_lut_kernel<uint8_t><<<1, 256>>>((uint32_t*)p->lut_dev, d.mean, d.std_);
    else tile_build_lut_kernel<<<1, 256>>>(p->lut_dev, d.mean, d.std_);
    count_launch();
    e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { cudaFree(p->lut_dev); cudaFree(p->geom_dev); delete p; return (int)e; }
  }
  *out = p;
  return BANDCU_OK;
}

void tile_plan_destroy(TilePlan* p) {
  if (!p) return;
  if (p->lut_dev) (void)cudaFree(p->lut_dev);
  if (p->geom_dev) (void)cudaFree(p->geom_dev);
  delete p;
}

typedef void (*TileKernel)(const TileParams);
template <typename OT>
static TileKernel pick_tile_kernel(bool g16, bool swap, bool il) {
  if (il) {
    if (g16) return swap ? tile_yuv_bilinear_down_kernel<OT, 16, true, true> : tile_yuv_bilinear_down_kernel<OT, 16, false, true>;
    return swap ? tile_yuv_bilinear_down_kernel<OT, 4, true, true> : tile_yuv_bilinear_down_kernel<OT, 4, false, true>;
  }
  if (g16) return swap ? tile_yuv_bilinear_down_kernel<OT, 16, true, false> : tile_yuv_bilinear_down_kernel<OT, 16, false, false>;
  return swap ? tile_yuv_bilinear_down_kernel<OT, 4, true, false> : tile_yuv_bilinear_down_kernel<OT, 4, false, false>;
}

int launch_tile(TilePlan* p, const BandCuPlanDesc& d, const BatchArgs& batch, const TableHints& hints, int n,
                cudaStream_t stream) {
  TileParams P;
  memset(&P, 0, sizeof(P));
  P.d = d;
  const bool f32_out = d.normalize && d.out_dtype == BANDCU_DT_F32;
  P.batch = batch;
  P.TR = p->TR; P.TC = p->TC;
  P.tiles_x = (d.dst_width + P.TC - 1) / P.TC;
  P.tiles_y = (d.dst_height + P.TR - 1) / P.TR;
  P.lut = p->lut_dev;
  P.geom = p->geom_dev;
  bool aligned4, aligned16 = (d.stride0 % 16 == 0) && (d.stride_uv % 16 == 0) && (d.src_width % 16 == 0) &&
                             ((((d.src_width + 1) / 2) * d.pixel_stride_uv) % 16 == 0);
  if (d.pixel_stride_uv == 2) {
    if (batch.table) {
      if (!hints.valid || !hints.semi_planar) return BANDCU_E_UNSUPPORTED;
      P.v_first = hints.vu ? 1 : 0;
      aligned4 = hints.aligned16;
      aligned16 = aligned16 && hints.aligned16;
    } else {
      if (batch.first.pv == batch.first.pu + 1) P.v_first = 0;
      else if (batch.first.pu == batch.first.pv + 1) P.v_first = 1;
      else return BANDCU_E_UNSUPPORTED;
      const uint8_t* uv0 = P.v_first ? batch.first.pv : batch.first.pu;
      const uintptr_t bits = reinterpret_cast<uintptr_t>(batch.first.p0) | reinterpret_cast<uintptr_t>(uv0) |
                             (n == 1 ? 0 : (uintptr_t)batch.in_stride);
      aligned4 = (bits & 3) == 0;
      aligned16 = aligned16 && (bits & 15) == 0;
    }
    P.interleaved = 1;
  } else {
    if (batch.table) {
      if (!hints.valid || !hints.planes_aligned4) return BANDCU_E_UNSUPPORTED;
      aligned4 = true;
      aligned16 = aligned16 && hints.planes_aligned16;
    } else {
      const uintptr_t bits = reinterpret_cast<uintptr_t>(batch.first.p0) | reinterpret_cast<uintptr_t>(batch.first.pu) |
                             reinterpret_cast<uintptr_t>(batch.first.pv) | (n == 1 ? 0 : (uintptr_t)batch.in_stride);
      aligned4 = (bits & 3) == 0;
      aligned16 = aligned16 && (bits & 15) == 0;
    }
  }
  if (!aligned4) return BANDCU_E_UNSUPPORTED;
  const TileGeom& g = aligned16 ? p->g16 : p->g4;
  P.l_pitch = g.l_pitch; P.c_pitch = g.c_pitch;
  P.l_rows_max = g.l_rows; P.c_rows_max = g.c_rows;
  // separable address forms (signed permutations: each mapped coordinate depends on one virtual coordinate)
  const int ps = d.pixel_stride_uv;
  P.skip = p->skip ? 1 : 0;
  P.skip_c = p->skip_c ? 1 : 0;
  P.c_row_0 = d.chroma_map.r0;
  P.c_row_mul = d.chroma_map.rr != 0 ? d.chroma_map.rr : d.chroma_map.rc;
  P.l_row_0 = d.luma_map.r0;
  P.l_row_mul = d.luma_map.rr != 0 ? d.luma_map.rr : d.luma_map.rc;
  P.lr_mul = d.luma_map.rr * g.l_pitch + d.luma_map.cr;
  P.lc_mul = d.luma_map.rc * g.l_pitch + d.luma_map.cc;
  const int BC = d.chroma_map.rr * g.c_pitch + d.chroma_map.cr * ps;  // per chroma-index row
  const int CC = d.chroma_map.rc * g.c_pitch + d.chroma_map.cc * ps;  // per chroma-index column
  if (d.pre_map.rr != 0) {  // conversion-image row <- virtual row
    P.cr_0 = d.pre_map.r0; P.cr_s = d.pre_map.rr; P.cr_mul = BC;
    P.cc_0 = d.pre_map.c0; P.cc_s = d.pre_map.cc; P.cc_mul = CC;
  } else {                  // conversion-image row <- virtual column
    P.cr_0 = d.pre_map.c0; P.cr_s = d.pre_map.cr; P.cr_mul = CC;
    P.cc_0 = d.pre_map.r0; P.cc_s = d.pre_map.rc; P.cc_mul = BC;
  }
  const bool swap = d.luma_map.rr == 0;
  // threads per staged row (2^shift): the split of the 256 threads over (rows, granules) that costs a thread the fewest
  // instructions: passes * (row set-up + trips * copy). Per-sample rows pay ~30 instructions of set-up (the row's
  // source index), rectangle rows ~6; a copy ~4. C3: 64 sampled luma rows of 11 granules -> 4 threads per row, one pass,
  // three copies each (was 16 threads per row, four passes).
  auto shift_for = [](int granules, int rows, int setup) {
    int best = 0, best_cost = 1 << 30;
    for (int sh = 0; sh <= 5; ++sh) {
      const int lpr = 1 << sh, rows_per_pass = 256 >> sh;
      const int cost = ((rows + rows_per_pass - 1) / rows_per_pass) * (setup + ((granules + lpr - 1) / lpr) * 4);
      if (cost <= best_cost) { best_cost = cost; best = sh; }  // ties: more threads per row (longer contiguous runs)
    }
    return best;
  };
  const int G = aligned16 ? 16 : 4;
  P.l_lpr_shift = shift_for(g.l_pitch / G, g.l_rows, p->skip ? 30 : 6);
  P.c_lpr_shift = shift_for(g.c_pitch / G, g.c_rows, p->skip_c ? 30 : 6);
  P.off_terms = (int)g.stage_bytes;
  const size_t smem = g.stage_bytes + 2048 + 3200;  // + row / column term tables + the [32][25]-word output tile
  {
    const int need = f32_out ? 15 : 3;
    bool ok;
    if (batch.table) ok = hints.out_aligned16;
    else ok = (reinterpret_cast<uintptr_t>(batch.first.out) & need) == 0 && (n == 1 || (batch.out_stride & need) == 0);
    P.out_vec = (ok && (d.dst_row_stride & need) == 0) ? 1 : 0;
  }
  TileKernel kern = f32_out ? pick_tile_kernel<float>(aligned16, swap, P.interleaved != 0) : pick_tile_kernel<uint8_t>(aligned16, swap, P.interleaved != 0);
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return (int)e;
  for (int done = 0; done < n;) {
    const int chunk = n - done > 65535 ? 65535 : n - done;
    TileParams Q = P;
    if (Q.batch.table) {
      Q.batch.table += done;
    } else {
      const long long o = (long long)done * Q.batch.in_stride;
      Q.batch.first.p0 += o;
      Q.batch.first.pu += o;
      Q.batch.first.pv += o;
      Q.batch.first.out += (long long)done * Q.batch.out_stride;
    }
    kern<<<dim3(P.tiles_x, P.tiles_y, chunk), 256, smem, stream>>>(Q);
    count_launch();
    done += chunk;
  }
  return (int)cudaGetLastError();
}

}  // namespace bandcu
