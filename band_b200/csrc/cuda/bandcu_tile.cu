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
#include <algorithm>
#include <vector>

#include "bandcu_device.cuh"
#include "bandcu_internal.h"

namespace bandcu {

// Source rectangles of one output tile (the same for every frame of a plan): computed once on the host at plan creation
// instead of by every CTA (three affine bounding boxes: ~170 of the ~1060 instructions a warp spent per tile).
struct __align__(16) TileEntry {
  int l_r0;          // luma: first source row of the rectangle
  int l_rows_gpr;    // rows | granules per staged row << 16 (never past the end of the plane row)
  int l_col0;        // first staged byte of a row, rounded down to the staging granule
  int AL;            // tile constant of the staged-luma address form (see TileParams)
  int c_r0, c_rows_gpr, c_col0;  // chroma plane(s): the same, in plane bytes
  int AC;
};

// Everything an output ROW / COLUMN contributes to a sample address and to the blend. The same for every frame and
// every tile row (column), so the host computes them once per plan -- [dst_height] and [dst_width] entries -- and the
// kernel reads its entry with one 128-bit and one 64-bit load (L2-resident: 24 KB for a 512x512 output). The first
// version rebuilt 32 + 32 entries per CTA behind a block barrier: ~60 instructions on 64 threads and a dependency
// between the staging copies and the arithmetic.
struct __align__(32) RowTerms {
  int lA, lB;       // luma offsets of virtual rows rA, rB
  int kA, kB;       // chroma offsets
  uint32_t w0, w1;  // InterpolateRow weights
  int pad[2];
};
struct __align__(32) ColTerms {
  int lA, lB, kA, kB;
  uint32_t wp;      // packed 7-bit column weights, pre-shifted by one: the blended byte lands in bits 24..31
  int pad[3];
};

struct TileParams {
  BandCuPlanDesc d;
  BatchArgs batch;
  const TileEntry* geom;  // [tiles_y][tiles_x]
  const RowTerms* rt;     // [dst_height]
  const ColTerms* ct;     // [dst_width]
  const int* lrow;        // skip:   [2 * positions] source luma row of every staged slot
  const int* crow;        // skip_c: [2 * positions] source chroma row of every staged slot
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
  int off_u, off_v;    // offsets of the staged chroma plane(s) in dynamic shared memory
  int off_out;         // offset of the [32][25]-word output tile
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
// loops have no division: (source, destination) advance by a fixed number of rows per pass. Destinations are 32-bit
// shared-space addresses throughout.
template <int G>
__device__ __forceinline__ void copy_granules(uint32_t dst, const uint8_t* src, int g0, int gpr, int lpr) {
  for (int g = g0, o = 0; g < gpr; g += lpr, o += lpr * G) {  // one trip unless a row has more granules than lpr
    if (G == 16) {
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + o), "l"(src + o) : "memory");
    } else {
      const uint32_t v = __ldg(reinterpret_cast<const uint32_t*>(src + o));
      asm volatile("st.shared.u32 [%0], %1;" ::"r"(dst + o), "r"(v) : "memory");
    }
  }
}
// Per-sample staging (skip mode): staged row `slot` comes from source row rows[slot] -- a host-built table
// (TileParams::lrow / crow), so a slot costs one 32-bit load instead of the 64-bit fixed-point stepping.
template <int G>
__device__ __forceinline__ void stage_slots(uint32_t smem_dst, const uint8_t* __restrict__ plane, int stride, int nslots,
                                            int col0_bytes, int gpr, int pitch, int lpr_shift,
                                            const int* __restrict__ rows) {
  const int lpr = 1 << lpr_shift;
  const int rows_per_pass = (int)blockDim.x >> lpr_shift;
  const int g0 = (int)threadIdx.x & (lpr - 1);
  if (g0 >= gpr) return;
  const uint8_t* col0 = plane + col0_bytes + g0 * G;
  const uint32_t dst0 = smem_dst + (uint32_t)(g0 * G);
  for (int r = (int)threadIdx.x >> lpr_shift; r < nslots; r += rows_per_pass)
    copy_granules<G>(dst0 + (uint32_t)(r * pitch), col0 + (long long)__ldg(rows + r) * stride, g0, gpr, lpr);
}

template <int G>
__device__ __forceinline__ void stage_plane(uint32_t smem_dst, const uint8_t* __restrict__ plane, int stride,
                                            int row0, int nrows, int col0_bytes, int gpr, int pitch, int lpr_shift) {
  const int lpr = 1 << lpr_shift;
  const int rows_per_pass = (int)blockDim.x >> lpr_shift;
  int r = (int)threadIdx.x >> lpr_shift;
  const int g0 = (int)threadIdx.x & (lpr - 1);
  if (g0 >= gpr) return;
  const uint8_t* src = plane + (long long)(row0 + r) * stride + col0_bytes + g0 * G;
  uint32_t dst = smem_dst + (uint32_t)(r * pitch + g0 * G);
  const long long src_step = (long long)rows_per_pass * stride;
  const uint32_t dst_step = (uint32_t)(rows_per_pass * pitch);
  for (; r < nrows; r += rows_per_pass, src += src_step, dst += dst_step) copy_granules<G>(dst, src, g0, gpr, lpr);
}

// ---- packed LEFT|RIGHT pixel-pair arithmetic (same identities as bandcu_fast.cu) ----
__device__ __forceinline__ uint32_t t_lds_u8(uint32_t a) {
  uint32_t v;
  asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(a));
  return v;
}
__device__ __forceinline__ uint32_t t_lds_u16(uint32_t a) {
  uint16_t v;
  asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(a));
  return v;
}
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
  // bytes 1 and 3 of the sums are the two results: one PRMT instead of shift + mask. w1 == 0 returns `top` exactly
  // ((256 t + 128) >> 8), whatever `bot` holds: no branch on the fraction.
  return __byte_perm(top * w0 + bot * w1 + 0x00800080u, 0u, 0x4341);
}

// Clamped virtual rows (columns) sampled by output row oy (column ox): fetch_px's clamps.
__host__ __device__ __forceinline__ void virt_rows(const BandCuPlanDesc& d, int oy, int* rA, int* rB) {
  long long y = (long long)d.y0 + (long long)oy * d.dy;
  if (y > d.max_y) y = d.max_y;
  const int yi = (int)(y >> 16);
  *rA = min(max(yi, 0), d.vh - 1);
  *rB = min(max(yi + 1, 0), d.vh - 1);
}
__host__ __device__ __forceinline__ void virt_cols(const BandCuPlanDesc& d, int ox, int* cA, int* cB) {
  const int xi = (int)(((long long)d.x0 + (long long)ox * d.dx) >> 16);
  *cA = min(max(xi, 0), d.vw - 1);
  *cB = min(max(xi + 1, 0), d.vw - 1);
}
// Host side (plan tables). `slot_pitch` != 0: this axis selects the staged luma row (skip mode): offsets are
// slot * pitch for tile position i = index % 32.
static RowTerms row_terms(const TileParams& P, int oy, int i, int slot_pitch, int slot_pitch_c) {
  const BandCuPlanDesc& d = P.d;
  long long y = (long long)d.y0 + (long long)oy * d.dy;
  if (y > d.max_y) y = d.max_y;
  RowTerms t;
  memset(&t, 0, sizeof(t));
  t.w1 = d.linear ? 0u : (uint32_t)((y >> 8) & 255);
  t.w0 = 256u - t.w1;
  int rA, rB;
  virt_rows(d, oy, &rA, &rB);
  t.lA = slot_pitch ? (2 * i) * slot_pitch : P.lr_mul * rA;
  t.lB = slot_pitch ? (2 * i + 1) * slot_pitch : P.lr_mul * rB;
  t.kA = slot_pitch_c ? (2 * i) * slot_pitch_c : P.cr_mul * ((P.cr_0 + P.cr_s * rA) >> 1);
  t.kB = slot_pitch_c ? (2 * i + 1) * slot_pitch_c : P.cr_mul * ((P.cr_0 + P.cr_s * rB) >> 1);
  return t;
}
static ColTerms col_terms(const TileParams& P, int ox, int i, int slot_pitch, int slot_pitch_c) {
  const BandCuPlanDesc& d = P.d;
  const long long x = (long long)d.x0 + (long long)ox * d.dx;
  const uint32_t xf = (uint32_t)((x >> 9) & 0x7f);
  ColTerms t;
  memset(&t, 0, sizeof(t));
  // ScaleARGBFilterCols_C on a packed pair: (L | R << 16) * (xf | (0x7f ^ xf) << 16) carries L*(0x7f^xf) + R*xf in its
  // upper half (bits 23..30 hold the 8-bit result); shifted left once more the result is the top byte.
  t.wp = (xf << 1) | ((0x7fu ^ xf) << 17);
  int cA, cB;
  virt_cols(d, ox, &cA, &cB);
  t.lA = slot_pitch ? (2 * i) * slot_pitch : P.lc_mul * cA;
  t.lB = slot_pitch ? (2 * i + 1) * slot_pitch : P.lc_mul * cB;
  t.kA = slot_pitch_c ? (2 * i) * slot_pitch_c : P.cc_mul * ((P.cc_0 + P.cc_s * cA) >> 1);
  t.kB = slot_pitch_c ? (2 * i + 1) * slot_pitch_c : P.cc_mul * ((P.cc_0 + P.cc_s * cB) >> 1);
  return t;
}
// Chroma of the LEFT|RIGHT sample pair at shared addresses aL, aR of the U plane (planar: V sits vd bytes further;
// interleaved (IL): the staged plane holds (first, second) byte pairs at even offsets -- one 16-bit load per sample,
// and two PRMTs with selectors sel_u, sel_v pick U and V for NV12 or NV21).
template <bool IL>
__device__ __forceinline__ TChroma t_chroma_at(uint32_t aL, uint32_t aR, uint32_t vd, uint32_t sel_u, uint32_t sel_v) {
  if (IL) {
    const uint32_t wL = t_lds_u16(aL), wR = t_lds_u16(aR);
    return t_chroma_terms(__byte_perm(wL, wR, sel_u), __byte_perm(wL, wR, sel_v));
  }
  return t_chroma_terms(t_lds_u8(aL) | (t_lds_u8(aR) << 16), t_lds_u8(aL + vd) | (t_lds_u8(aR + vd) << 16));
}
// LEFT and RIGHT samples fall into the same chroma cell (xi even: half of all columns): one sample feeds both lanes.
template <bool IL>
__device__ __forceinline__ TChroma t_chroma_at_one(uint32_t a, uint32_t vd, uint32_t sel_u, uint32_t sel_v) {
  if (IL) {
    const uint32_t w = t_lds_u16(a);
    return t_chroma_terms(__byte_perm(w, w, sel_u), __byte_perm(w, w, sel_v));
  }
  return t_chroma_terms(t_lds_u8(a) * 0x10001u, t_lds_u8(a + vd) * 0x10001u);
}
// One output pixel from the staged planes: the four luma and four chroma sample addresses (Top/Bottom x Left/Right). Raw
// products: the 8-bit results are the top bytes. one_col / one_row: the two columns / rows share their chroma sample;
// callers pass them only where the flag is warp-uniform (a property of the output column under a quarter-turn, of the
// output row otherwise), so the branch never diverges.
template <bool IL>
__device__ __forceinline__ void tile_pixel(uint32_t lTL, uint32_t lTR, uint32_t lBL, uint32_t lBR, uint32_t kTL, uint32_t kTR,
                                           uint32_t kBL, uint32_t kBR, bool one_col, bool one_row, uint32_t vd, uint32_t sel_u,
                                           uint32_t sel_v, uint32_t w0, uint32_t w1, uint32_t wp, uint32_t& r, uint32_t& g,
                                           uint32_t& b) {
  uint32_t tr, tg, tb, br, bg, bb;
  const TChroma ct = one_col ? t_chroma_at_one<IL>(kTL, vd, sel_u, sel_v) : t_chroma_at<IL>(kTL, kTR, vd, sel_u, sel_v);
  t_luma_pair_to_rgb(t_lds_u8(lTL), t_lds_u8(lTR), ct, tr, tg, tb);
  TChroma cb = ct;
  if (!one_row) cb = one_col ? t_chroma_at_one<IL>(kBL, vd, sel_u, sel_v) : t_chroma_at<IL>(kBL, kBR, vd, sel_u, sel_v);
  t_luma_pair_to_rgb(t_lds_u8(lBL), t_lds_u8(lBR), cb, br, bg, bb);
  r = t_vblend2(tr, br, w0, w1) * wp;
  g = t_vblend2(tg, bg, w0, w1) * wp;
  b = t_vblend2(tb, bb, w0, w1) * wp;
}

// SWAP: the luma map exchanges the axes (quarter-turns). Lanes then run along the output COLUMN so that a warp's
// shared-memory reads walk consecutive source bytes (lanes along the output row would stride by whole staged rows:
// 4-way bank conflicts measured, L1 at 84 %); each thread produces FOUR adjacent pixels of its output row.
// Without SWAP lanes run along the output row. Either way the tile's 8-bit results go through a [32][25]-word shared
// array and leave row-major (see below).
template <typename OT, int G, bool SWAP, bool IL>
__global__ void __launch_bounds__(256) tile_yuv_bilinear_down_kernel(const __grid_constant__ TileParams P) {
  extern __shared__ __align__(16) uint8_t smem[];
  __shared__ uint32_t lut[256];
  const BandCuPlanDesc& d = P.d;
  // f32 outputs: the Normalize table; one-byte Normalize outputs (int8 / uint8 models): 32-bit entries holding the byte
  if (sizeof(OT) == 4 || P.lut) lut[threadIdx.x] = __ldg(reinterpret_cast<const uint32_t*>(P.lut) + threadIdx.x);  // blockDim.x == 256
  const int oy0 = blockIdx.y * 32, ox0 = blockIdx.x * 32;  // TR == TC == 32
  const int nrows = min(32, d.dst_height - oy0), ncols = min(32, d.dst_width - ox0);
  const FramePtrs f = frame_ptrs(P.batch, blockIdx.z);

  const TileEntry te = P.geom[blockIdx.y * P.tiles_x + blockIdx.x];
  const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(smem);
  const uint32_t sl = sbase, su = sbase + (uint32_t)P.off_u;
  const int npos = SWAP ? ncols : nrows, pos0 = SWAP ? ox0 : oy0;
  if (P.skip) stage_slots<G>(sl, f.p0, d.stride0, 2 * npos, te.l_col0, te.l_rows_gpr >> 16, P.l_pitch, P.l_lpr_shift, P.lrow + 2 * pos0);
  else stage_plane<G>(sl, f.p0, d.stride0, te.l_r0, te.l_rows_gpr & 0xFFFF, te.l_col0, te.l_rows_gpr >> 16, P.l_pitch, P.l_lpr_shift);
  auto stage_chroma = [&](uint32_t dst, const uint8_t* plane) {
    if (P.skip_c) stage_slots<G>(dst, plane, d.stride_uv, 2 * npos, te.c_col0, te.c_rows_gpr >> 16, P.c_pitch, P.c_lpr_shift, P.crow + 2 * pos0);
    else stage_plane<G>(dst, plane, d.stride_uv, te.c_r0, te.c_rows_gpr & 0xFFFF, te.c_col0, te.c_rows_gpr >> 16, P.c_pitch, P.c_lpr_shift);
  };
  // IL: su is the base of the (first, second) byte pairs; byte 2 of a zero-extended 16-bit load is the zero lane filler
  const uint32_t sel_u = P.v_first ? 0x2521u : 0x2420u, sel_v = P.v_first ? 0x2420u : 0x2521u;
  if (IL) {
    stage_chroma(su, P.v_first ? f.pv : f.pu);
  } else {
    stage_chroma(su, f.pu);
    stage_chroma(sbase + (uint32_t)P.off_v, f.pv);
  }
  if (G == 16) asm volatile("cp.async.commit_group;" ::: "memory");

  // the tile's 32 row and 32 column terms move from the plan's tables into shared memory while the copies fly (two
  // warps, two 16-byte loads each); the arithmetic then reads them as broadcasts
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;  // 8 warps
  uint4* terms = reinterpret_cast<uint4*>(smem + P.off_out + 3200);  // [32] RowTerms then [32] ColTerms, 32 bytes each
  if (wid < 2) {
    const uint4* src = wid == 0 ? reinterpret_cast<const uint4*>(P.rt + min(oy0 + lane, d.dst_height - 1))
                                : reinterpret_cast<const uint4*>(P.ct + min(ox0 + lane, d.dst_width - 1));
    const uint4 a = __ldg(src), b2 = __ldg(src + 1);
    terms[wid * 64 + lane * 2] = a;
    terms[wid * 64 + lane * 2 + 1] = b2;
  }
  const RowTerms* rt = reinterpret_cast<const RowTerms*>(terms);
  const ColTerms* ct = reinterpret_cast<const ColTerms*>(terms + 64);
  const uint32_t vd = (uint32_t)(P.off_v - P.off_u);
  const uint32_t gl = sl + (uint32_t)te.AL, gu = su + (uint32_t)te.AC;
  uint32_t* os = reinterpret_cast<uint32_t*>(smem + P.off_out);
  if (G == 16) asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncthreads();
  if (!SWAP) {
    // lane = output column inside the tile; warp w owns rows w, w+8, ...
    if (lane < ncols) {
      const ColTerms Cn = ct[lane];
      const uint32_t lL = gl + (uint32_t)Cn.lA, lR = gl + (uint32_t)Cn.lB, kL = gu + (uint32_t)Cn.kA, kR = gu + (uint32_t)Cn.kB;
      uint8_t* ob = reinterpret_cast<uint8_t*>(os) + lane * 3;
      for (int ry = wid; ry < nrows; ry += 8) {
        const RowTerms R = rt[ry];  // warp-uniform: broadcast
        uint32_t r, g, b;
        tile_pixel<IL>(lL + R.lA, lR + R.lA, lL + R.lB, lR + R.lB, kL + R.kA, kR + R.kA, kL + R.kB, kR + R.kB, false, R.kA == R.kB,
                       vd, sel_u, sel_v, R.w0, R.w1, Cn.wp, r, g, b);
        ob[ry * 100] = (uint8_t)(r >> 24); ob[ry * 100 + 1] = (uint8_t)(g >> 24); ob[ry * 100 + 2] = (uint8_t)(b >> 24);
      }
    }
  } else {
    // lane = output row inside the tile; warp w owns columns 4w .. 4w+3
    const int cx0 = wid * 4;
    if (lane < nrows && cx0 < ncols) {
      const RowTerms R = rt[lane];
      const uint32_t lT = gl + (uint32_t)R.lA, lB = gl + (uint32_t)R.lB, kT = gu + (uint32_t)R.kA, kB = gu + (uint32_t)R.kB;
      uint32_t px[12];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const ColTerms Cn = ct[cx0 + j];  // warp-uniform: broadcast (entries past the image repeat the last column)
        tile_pixel<IL>(lT + Cn.lA, lT + Cn.lB, lB + Cn.lA, lB + Cn.lB, kT + Cn.kA, kT + Cn.kB, kB + Cn.kA, kB + Cn.kB,
                       Cn.kA == Cn.kB, false, vd, sel_u, sel_v, R.w0, R.w1, Cn.wp, px[3 * j], px[3 * j + 1], px[3 * j + 2]);
      }
      // top bytes of twelve raw products -> three words (three PRMTs per word)
#pragma unroll
      for (int q = 0; q < 3; ++q) {
        const uint32_t lo = __byte_perm(px[4 * q], px[4 * q + 1], 0x0073), hi = __byte_perm(px[4 * q + 2], px[4 * q + 3], 0x0073);
        os[lane * 25 + wid * 3 + q] = __byte_perm(lo, hi, 0x5410);
      }
    }
  }
  __syncthreads();
  // The 8-bit results of the tile went through a [32][25]-word shared array (odd pitch: the compute layout's writes and
  // the row-major reads are both conflict-free) and leave as row-major 16-byte (f32) / 4-byte (u8) pieces: a warp
  // writes 1.3 output rows of the tile per store instruction. (Storing from the quarter-turn compute layout -- each
  // lane its own output row -- cost 32 L1 wavefronts per store instruction, a third of the kernel's L1 data-pipe
  // load: profiles/r01_ncu_full_tile_v6_uncoalesced_stores.txt.)
  uint8_t* obase = f.out + (long long)oy0 * d.dst_row_stride + (size_t)ox0 * (3 * sizeof(OT));
  if (P.out_vec && ncols == 32) {  // 24 aligned pieces per row (ox0 is a multiple of 32)
    // piece (row, q) of thread t: t = 24 row + q once, then every pass adds 256 = 10 rows + 16 pieces
    int row = (int)threadIdx.x / 24, q = (int)threadIdx.x - row * 24;
    const uint32_t stride = (uint32_t)d.dst_row_stride;
    for (; row < nrows; row += 10, q += 16) {
      if (q >= 24) { q -= 24; ++row; if (row >= nrows) break; }
      uint32_t wv = os[row * 25 + q];
      uint8_t* o = obase + row * stride;
      if (sizeof(OT) == 1 && P.lut)
        wv = (lut[wv & 0xFF] & 0xFF) | ((lut[(wv >> 8) & 0xFF] & 0xFF) << 8) | ((lut[(wv >> 16) & 0xFF] & 0xFF) << 16) |
             (lut[wv >> 24] << 24);
      if (sizeof(OT) == 1) reinterpret_cast<uint32_t*>(o)[q] = wv;
      else reinterpret_cast<uint4*>(o)[q] = make_uint4(lut[wv & 0xFF], lut[(wv >> 8) & 0xFF], lut[(wv >> 16) & 0xFF], lut[wv >> 24]);
    }
  } else {  // right-edge tiles, unaligned outputs: element by element, still row-major
    const uint8_t* ob = reinterpret_cast<const uint8_t*>(os);
    for (int e = threadIdx.x; e < nrows * 96; e += 256) {
      const int row = e / 96, c = e - row * 96;
      if (c >= 3 * ncols) continue;
      const uint32_t v = ob[row * 100 + c];
      uint8_t* o = obase + (long long)row * d.dst_row_stride;
      if (sizeof(OT) == 1) o[c] = P.lut ? (uint8_t)lut[v] : (uint8_t)v;
      else reinterpret_cast<uint32_t*>(o)[c] = lut[v];
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

// Everything a launch needs that depends on the staging granule G (16 when the batch is 16-byte granular, else 4): the
// worst-case geometry, and the device tables built from it once per plan.
struct TileTables {
  TileGeom g;
  TileEntry* geom_dev = nullptr;
  RowTerms* rt_dev = nullptr;
  ColTerms* ct_dev = nullptr;
  int* lrow_dev = nullptr;
  int* crow_dev = nullptr;
};

struct TilePlan {
  float* lut_dev;
  int TR, TC;
  TileTables t16, t4;
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

// The launch parameters that follow from the plan and the geometry alone (no batch pointers): separable address forms
// (signed permutations: each mapped coordinate depends on one virtual coordinate), staging split, shared-memory map.
static void fill_geometry_params(TileParams& P, const BandCuPlanDesc& d, const TileGeom& g, int G, bool skip, bool skip_c) {
  P.d = d;
  P.TR = 32; P.TC = 32;
  P.tiles_x = (d.dst_width + 31) / 32;
  P.tiles_y = (d.dst_height + 31) / 32;
  P.interleaved = d.pixel_stride_uv == 2 ? 1 : 0;
  P.l_pitch = g.l_pitch; P.c_pitch = g.c_pitch;
  P.l_rows_max = g.l_rows; P.c_rows_max = g.c_rows;
  const int ps = d.pixel_stride_uv;
  P.skip = skip ? 1 : 0;
  P.skip_c = skip_c ? 1 : 0;
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
  // threads per staged row (2^shift): the split of the 256 threads over (rows, granules) that costs a thread the fewest
  // instructions: passes * (row set-up + trips * copy). Per-sample rows pay ~8 instructions of set-up (table load and
  // row address), rectangle rows ~6; a copy ~4.
  auto shift_for = [](int granules, int rows, int setup) {
    int best = 0, best_cost = 1 << 30;
    for (int sh = 0; sh <= 5; ++sh) {
      const int lpr = 1 << sh, rows_per_pass = 256 >> sh;
      const int cost = ((rows + rows_per_pass - 1) / rows_per_pass) * (setup + ((granules + lpr - 1) / lpr) * 4);
      if (cost <= best_cost) { best_cost = cost; best = sh; }  // ties: more threads per row (longer contiguous runs)
    }
    return best;
  };
  P.l_lpr_shift = shift_for(g.l_pitch / G, g.l_rows, skip ? 8 : 6);
  P.c_lpr_shift = shift_for(g.c_pitch / G, g.c_rows, skip_c ? 8 : 6);
  P.off_u = g.l_rows * g.l_pitch;
  P.off_v = P.off_u + (P.interleaved ? 0 : g.c_rows * g.c_pitch);
  P.off_out = (int)g.stage_bytes;
}

template <typename T>
static cudaError_t upload(T** dev, const std::vector<T>& host) {
  *dev = nullptr;
  if (host.empty()) return cudaSuccess;
  cudaError_t e = cudaMalloc((void**)dev, host.size() * sizeof(T));
  if (e == cudaSuccess) e = cudaMemcpy(*dev, host.data(), host.size() * sizeof(T), cudaMemcpyHostToDevice);
  return e;
}

static void free_tables(TileTables& t) {
  if (t.geom_dev) (void)cudaFree(t.geom_dev);
  if (t.rt_dev) (void)cudaFree(t.rt_dev);
  if (t.ct_dev) (void)cudaFree(t.ct_dev);
  if (t.lrow_dev) (void)cudaFree(t.lrow_dev);
  if (t.crow_dev) (void)cudaFree(t.crow_dev);
  t = TileTables();
}

template <int G>
static cudaError_t build_tables(const BandCuPlanDesc& d, bool skip, bool skip_c, TileTables* out) {
  const bool inter = d.pixel_stride_uv == 2;
  out->g = tile_geometry<G>(d, 32, 32, inter, skip, skip_c);
  TileParams P;
  memset(&P, 0, sizeof(P));
  fill_geometry_params(P, d, out->g, G, skip, skip_c);
  const bool swap = d.luma_map.rr == 0;
  const int ps = d.pixel_stride_uv;
  // per-tile source rectangles and the tile constants of the two affine address forms
  std::vector<TileEntry> geom((size_t)P.tiles_x * P.tiles_y);
  for (int ty = 0; ty < P.tiles_y; ++ty) {
    for (int tx = 0; tx < P.tiles_x; ++tx) {
      const int oy0 = ty * 32, ox0 = tx * 32;
      const int oy1 = (oy0 + 32 < d.dst_height ? oy0 + 32 : d.dst_height) - 1;
      const int ox1 = (ox0 + 32 < d.dst_width ? ox0 + 32 : d.dst_width) - 1;
      Rect L, C;
      tile_rects(d, virtual_rect(d, oy0, oy1, ox0, ox1), &L, &C);
      TileEntry& e = geom[(size_t)ty * P.tiles_x + tx];
      memset(&e, 0, sizeof(e));
      // granules per staged row: the launch-wide pitch, but never past the end of the plane row (the pitch is the
      // worst case over all tiles and odd; a tile at the right edge must not read into memory behind the last row)
      const int c_row_bytes = ((d.src_width + 1) >> 1) * ps;
      const int l_first = L.c0, c_first = C.c0 * ps;
      const int l_gpr = std::min(P.l_pitch / G, (d.src_width - (l_first / G) * G + G - 1) / G);
      const int c_gpr = std::min(P.c_pitch / G, (c_row_bytes - (c_first / G) * G + G - 1) / G);
      e.l_r0 = L.r0; e.l_rows_gpr = (L.r1 - L.r0 + 1) | (l_gpr << 16); e.l_col0 = (l_first / G) * G;
      e.c_r0 = C.r0; e.c_rows_gpr = (C.r1 - C.r0 + 1) | (c_gpr << 16); e.c_col0 = (c_first / G) * G;
      e.AL = (skip ? 0 : (d.luma_map.r0 - e.l_r0) * P.l_pitch) + d.luma_map.c0 - e.l_col0;
      e.AC = (skip_c ? 0 : (d.chroma_map.r0 - e.c_r0) * P.c_pitch) + d.chroma_map.c0 * ps - e.c_col0;
    }
  }
  // row / column terms of every output row / column (tile position = index % 32)
  std::vector<RowTerms> rt(d.dst_height);
  std::vector<ColTerms> ct(d.dst_width);
  for (int oy = 0; oy < d.dst_height; ++oy)
    rt[oy] = row_terms(P, oy, oy & 31, (skip && !swap) ? P.l_pitch : 0, (skip_c && !swap) ? P.c_pitch : 0);
  for (int ox = 0; ox < d.dst_width; ++ox)
    ct[ox] = col_terms(P, ox, ox & 31, (skip && swap) ? P.l_pitch : 0, (skip_c && swap) ? P.c_pitch : 0);
  // per-sample staging: the source row of every slot (two per position along the axis that maps to source rows)
  std::vector<int> lrow, crow;
  const int npos = swap ? d.dst_width : d.dst_height;
  if (skip || skip_c) {
    for (int p = 0; p < npos; ++p) {
      int a, b;
      if (swap) virt_cols(d, p, &a, &b);
      else virt_rows(d, p, &a, &b);
      if (skip) {
        lrow.push_back(P.l_row_0 + P.l_row_mul * a);
        lrow.push_back(P.l_row_0 + P.l_row_mul * b);
      }
      if (skip_c) {
        const int q0 = swap ? P.cc_0 : P.cr_0, qs = swap ? P.cc_s : P.cr_s;
        crow.push_back(P.c_row_0 + P.c_row_mul * ((q0 + qs * a) >> 1));
        crow.push_back(P.c_row_0 + P.c_row_mul * ((q0 + qs * b) >> 1));
      }
    }
  }
  cudaError_t e = upload(&out->geom_dev, geom);
  if (e == cudaSuccess) e = upload(&out->rt_dev, rt);
  if (e == cudaSuccess) e = upload(&out->ct_dev, ct);
  if (e == cudaSuccess) e = upload(&out->lrow_dev, lrow);
  if (e == cudaSuccess) e = upload(&out->crow_dev, crow);
  return e;
}

static size_t tile_smem_bytes(const TileGeom& g) { return g.stage_bytes + 3200 + 2048; }  // + the [32][25]-word output tile + row / column terms

int tile_plan_create(const BandCuPlanDesc& d, TilePlan** out) {
  // per-sample luma staging pays once the down-scale along the axis that maps to source rows reaches 2
  const char* ns = getenv("BANDCU_TILE_NO_SKIP");
  const bool skip = !(ns && ns[0] == '1') && (d.luma_map.rr != 0 ? d.dy : d.dx) >= 0x20000;
  // chroma too once the ratio reaches 6 (below that nearly every half-resolution row is sampled and the per-slot row
  // arithmetic costs more than it saves: measured at 4.8), if chroma rows follow the same virtual axis as luma rows (they do for every
  // chain the plan compiler emits; checked, not assumed)
  const bool luma_rows_from_vrows = d.luma_map.rr != 0;
  const bool chroma_rows_from_vrows = d.pre_map.rr != 0 ? d.chroma_map.rr != 0 : d.chroma_map.rc != 0;
  const bool skip_c = skip && luma_rows_from_vrows == chroma_rows_from_vrows && (luma_rows_from_vrows ? d.dy : d.dx) >= 0x60000;
  {
    const TileGeom probe = tile_geometry<16>(d, 32, 32, d.pixel_stride_uv == 2, skip, skip_c);
    if (tile_smem_bytes(probe) > 100 * 1024) return BANDCU_E_UNSUPPORTED;  // very large ratios: keep >= 2 CTAs per SM
  }
  TilePlan* p = new TilePlan();
  p->lut_dev = nullptr;
  p->TR = 32; p->TC = 32;
  p->skip = skip;
  p->skip_c = skip_c;
  cudaError_t e = build_tables<16>(d, skip, skip_c, &p->t16);
  if (e == cudaSuccess) e = build_tables<4>(d, skip, skip_c, &p->t4);
  if (e != cudaSuccess) { free_tables(p->t16); free_tables(p->t4); delete p; return (int)e; }
  if (d.normalize) {
    e = cudaMalloc((void**)&p->lut_dev, 256 * sizeof(float));
    if (e != cudaSuccess) { free_tables(p->t16); free_tables(p->t4); delete p; return (int)e; }
    if (d.out_dtype == BANDCU_DT_I8) tile_build_byte_lut_kernel<int8_t><<<1, 256>>>((uint32_t*)p->lut_dev, d.mean, d.std_);
    else if (d.out_dtype == BANDCU_DT_U8) tile_build_byte_lut_kernel<uint8_t><<<1, 256>>>((uint32_t*)p->lut_dev, d.mean, d.std_);
    else tile_build_lut_kernel<<<1, 256>>>(p->lut_dev, d.mean, d.std_);
    count_launch();
    e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { cudaFree(p->lut_dev); free_tables(p->t16); free_tables(p->t4); delete p; return (int)e; }
  }
  *out = p;
  return BANDCU_OK;
}

void tile_plan_destroy(TilePlan* p) {
  if (!p) return;
  if (p->lut_dev) (void)cudaFree(p->lut_dev);
  free_tables(p->t16);
  free_tables(p->t4);
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
  const bool f32_out = d.normalize && d.out_dtype == BANDCU_DT_F32;
  int v_first = 0;
  bool aligned4, aligned16 = (d.stride0 % 16 == 0) && (d.stride_uv % 16 == 0) && (d.src_width % 16 == 0) &&
                             ((((d.src_width + 1) / 2) * d.pixel_stride_uv) % 16 == 0);
  if (d.pixel_stride_uv == 2) {
    if (batch.table) {
      if (!hints.valid || !hints.semi_planar) return BANDCU_E_UNSUPPORTED;
      v_first = hints.vu ? 1 : 0;
      aligned4 = hints.aligned16;
      aligned16 = aligned16 && hints.aligned16;
    } else {
      if (batch.first.pv == batch.first.pu + 1) v_first = 0;
      else if (batch.first.pu == batch.first.pv + 1) v_first = 1;
      else return BANDCU_E_UNSUPPORTED;
      const uint8_t* uv0 = v_first ? batch.first.pv : batch.first.pu;
      const uintptr_t bits = reinterpret_cast<uintptr_t>(batch.first.p0) | reinterpret_cast<uintptr_t>(uv0) |
                             (n == 1 ? 0 : (uintptr_t)batch.in_stride);
      aligned4 = (bits & 3) == 0;
      aligned16 = aligned16 && (bits & 15) == 0;
    }
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
  const TileTables& t = aligned16 ? p->t16 : p->t4;
  fill_geometry_params(P, d, t.g, aligned16 ? 16 : 4, p->skip, p->skip_c);
  P.batch = batch;
  P.v_first = v_first;
  P.lut = p->lut_dev;
  P.geom = t.geom_dev;
  P.rt = t.rt_dev;
  P.ct = t.ct_dev;
  P.lrow = t.lrow_dev;
  P.crow = t.crow_dev;
  const bool swap = d.luma_map.rr == 0;
  const size_t smem = tile_smem_bytes(t.g);
  {
    const int need = f32_out ? 15 : 3;
    bool ok;
    if (batch.table) ok = hints.out_aligned16;
    else ok = (reinterpret_cast<uintptr_t>(batch.first.out) & need) == 0 && (n == 1 || (batch.out_stride & need) == 0);
    P.out_vec = (ok && (d.dst_row_stride & need) == 0) ? 1 : 0;
  }
  TileKernel kern = f32_out ? pick_tile_kernel<float>(aligned16, swap, P.interleaved != 0) : pick_tile_kernel<uint8_t>(aligned16, swap, P.interleaved != 0);
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024);  // one value for every plan (G = 16 and G = 4 tiles share kernels)
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
