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

#include "bandcu_device.cuh"
#include "bandcu_internal.h"

namespace bandcu {

struct TileParams {
  BandCuPlanDesc d;
  BatchArgs batch;
  int TR, TC;          // output rows / columns per tile
  int tiles_x, tiles_y;
  int interleaved;     // 1: semi-planar chroma (one staged plane, U/V one byte apart)
  int v_first;         // interleaved: V is the first byte of a pair (NV21)
  int lut_f32;         // output goes through the Normalize table
  const float* lut;
};

struct TilePlan {
  float* lut_dev;
  int TR, TC;
  size_t smem;
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
// G == 4 : 32-bit loads through registers (any word-aligned layout). No divisions inside the loop: (row, granule)
// advance incrementally by blockDim.x.
template <int G>
__device__ __forceinline__ void stage_plane(uint8_t* smem_dst, const uint8_t* __restrict__ plane, int stride,
                                            int row0, int nrows, int first_byte, int pitch) {
  const int gpr = pitch / G;                       // granules per staged row
  const int total = nrows * gpr;
  const int step_r = (int)blockDim.x / gpr, step_g = (int)blockDim.x - step_r * gpr;
  int r = (int)threadIdx.x / gpr, g = (int)threadIdx.x - r * gpr;
  const uint8_t* src0 = plane + (long long)row0 * stride + (first_byte / G) * G;
  for (int i = threadIdx.x; i < total; i += blockDim.x) {
    const uint8_t* src = src0 + (long long)r * stride + g * G;
    uint8_t* dst = smem_dst + r * pitch + g * G;
    if (G == 16) {
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src) : "memory");
    } else {
      *reinterpret_cast<uint32_t*>(dst) = __ldg(reinterpret_cast<const uint32_t*>(src));
    }
    r += step_r;
    g += step_g;
    if (g >= gpr) { g -= gpr; ++r; }
  }
}

// ---- packed LEFT|RIGHT pixel-pair arithmetic (same identities as bandcu_fast.cu; values instead of addresses) ----
struct TChroma {
  uint32_t kr, kg, kb;
};
__device__ __forceinline__ TChroma t_chroma_pair(uint32_t uL, uint32_t vL, uint32_t uR, uint32_t vR) {
  const uint32_t u2 = uL | (uR << 16), v2 = vL | (vR << 16);
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
  return ((top * w0 + bot * w1 + 0x00800080u) >> 8) & 0x00FF00FFu;
}

template <typename OT, int G>
__global__ void __launch_bounds__(256) tile_yuv_bilinear_down_kernel(const __grid_constant__ TileParams P) {
  extern __shared__ __align__(16) uint8_t smem[];
  __shared__ float lut[256];
  const BandCuPlanDesc& d = P.d;
  if (sizeof(OT) == 4) {
    lut[threadIdx.x] = P.lut[threadIdx.x];  // blockDim.x == 256
  }
  const int tile = blockIdx.x;
  const int ty = tile / P.tiles_x, tx = tile - ty * P.tiles_x;
  const int oy0 = ty * P.TR, ox0 = tx * P.TC;
  const int oy1 = min(oy0 + P.TR, d.dst_height) - 1, ox1 = min(ox0 + P.TC, d.dst_width) - 1;
  const FramePtrs f = frame_ptrs(P.batch, blockIdx.y);

  const Rect v = virtual_rect(d, oy0, oy1, ox0, ox1);
  Rect L, C;
  tile_rects(d, v, &L, &C);
  // luma plane
  const int l_first = L.c0, l_pitch = staged_pitch<G>(L.c0, L.c1);
  const int l_rows = L.r1 - L.r0 + 1;
  uint8_t* sl = smem;
  stage_plane<G>(sl, f.p0, d.stride0, L.r0, l_rows, l_first, l_pitch);
  // chroma plane(s)
  const int ps = d.pixel_stride_uv;
  const int c_first = C.c0 * ps, c_last = C.c1 * ps + (P.interleaved ? 1 : 0);
  const int c_pitch = staged_pitch<G>(c_first, c_last);
  const int c_rows = C.r1 - C.r0 + 1;
  uint8_t* su = sl + l_rows * l_pitch;
  uint8_t* sv;
  if (P.interleaved) {
    const uint8_t* base = P.v_first ? f.pv : f.pu;
    stage_plane<G>(su, base, d.stride_uv, C.r0, c_rows, c_first, c_pitch);
    sv = su + (P.v_first ? 0 : 1);
    su = su + (P.v_first ? 1 : 0);
  } else {
    sv = su + c_rows * c_pitch;
    stage_plane<G>(su, f.pu, d.stride_uv, C.r0, c_rows, c_first, c_pitch);
    stage_plane<G>(sv, f.pv, d.stride_uv, C.r0, c_rows, c_first, c_pitch);
  }
  if (G == 16) asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
  __syncthreads();

  // affine address forms: luma(r,c) = AL + BL r + CL c over virtual coordinates; chroma over (r1>>1, c1>>1)
  const int BL = d.luma_map.rr * l_pitch + d.luma_map.cr, CL = d.luma_map.rc * l_pitch + d.luma_map.cc;
  const int AL = (d.luma_map.r0 - L.r0) * l_pitch + d.luma_map.c0 - (l_first / G) * G;
  const int BC = d.chroma_map.rr * c_pitch + d.chroma_map.cr * ps, CC = d.chroma_map.rc * c_pitch + d.chroma_map.cc * ps;
  const int AC = (d.chroma_map.r0 - C.r0) * c_pitch + d.chroma_map.c0 * ps - (c_first / G) * G;
  auto luma_at = [&](int r, int c) -> uint32_t { return sl[AL + BL * r + CL * c]; };
  auto chroma_off = [&](int r, int c) -> int {
    const int r1 = t_map_row(d.pre_map, r, c) >> 1, c1 = t_map_col(d.pre_map, r, c) >> 1;
    return AC + BC * r1 + CC * c1;
  };

  const int lx = threadIdx.x & 31, ly = threadIdx.x >> 5;  // 8 warps: lanes along the output row
  for (int cx = lx; cx < P.TC; cx += 32) {
    const int ox = ox0 + cx;
    if (ox > ox1) break;
    const long long x = (long long)d.x0 + (long long)ox * d.dx;
    const int xi = (int)(x >> 16);
    const uint32_t xf = (uint32_t)((x >> 9) & 0x7f);
    const uint32_t wp = xf | ((0x7fu ^ xf) << 16);
    const int cA = min(max(xi, 0), d.vw - 1), cB = min(max(xi + 1, 0), d.vw - 1);
    for (int ry = ly; ry < P.TR; ry += 8) {
      const int oy = oy0 + ry;
      if (oy > oy1) break;
      long long y = (long long)d.y0 + (long long)oy * d.dy;
      if (y > d.max_y) y = d.max_y;
      const int yi = (int)(y >> 16);
      const uint32_t w1 = d.linear ? 0u : (uint32_t)((y >> 8) & 255), w0 = 256u - w1;
      const int rA = min(max(yi, 0), d.vh - 1), rB = min(max(yi + 1, 0), d.vh - 1);
      uint32_t tr, tg, tb, br, bg, bb;
      {
        const int oL = chroma_off(rA, cA), oR = chroma_off(rA, cB);
        const TChroma ct = t_chroma_pair(su[oL], sv[oL], su[oR], sv[oR]);
        t_luma_pair_to_rgb(luma_at(rA, cA), luma_at(rA, cB), ct, tr, tg, tb);
      }
      if (w1) {
        const int oL = chroma_off(rB, cA), oR = chroma_off(rB, cB);
        const TChroma cb = t_chroma_pair(su[oL], sv[oL], su[oR], sv[oR]);
        t_luma_pair_to_rgb(luma_at(rB, cA), luma_at(rB, cB), cb, br, bg, bb);
      } else {
        br = tr; bg = tg; bb = tb;
      }
      const uint32_t r = t_vblend2(tr, br, w0, w1) * wp;
      const uint32_t g = t_vblend2(tg, bg, w0, w1) * wp;
      const uint32_t b = t_vblend2(tb, bb, w0, w1) * wp;
      uint8_t* o = f.out + (long long)oy * d.dst_row_stride + (size_t)ox * (3 * sizeof(OT));
      if (sizeof(OT) == 1) {
        o[0] = (uint8_t)(r >> 23); o[1] = (uint8_t)(g >> 23); o[2] = (uint8_t)(b >> 23);
      } else {
        float* of = reinterpret_cast<float*>(o);
        of[0] = lut[(r >> 23) & 0xFF]; of[1] = lut[(g >> 23) & 0xFF]; of[2] = lut[(b >> 23) & 0xFF];
      }
    }
  }
}

__global__ void tile_build_lut_kernel(float* lut, float mean, float sd) {
  const int i = threadIdx.x;
  if (i < 256) lut[i] = normalize_elem<float>((float)i, mean, sd);
}

// Worst-case shared memory of a TR x TC tile: evaluated on the host over every tile (cheap: a few thousand tiles).
static size_t tile_smem_bytes(const BandCuPlanDesc& d, int TR, int TC, bool interleaved) {
  const int tiles_x = (d.dst_width + TC - 1) / TC, tiles_y = (d.dst_height + TR - 1) / TR;
  size_t worst = 0;
  for (int ty = 0; ty < tiles_y; ++ty) {
    for (int tx = 0; tx < tiles_x; ++tx) {
      const int oy0 = ty * TR, ox0 = tx * TC;
      const int oy1 = (oy0 + TR < d.dst_height ? oy0 + TR : d.dst_height) - 1;
      const int ox1 = (ox0 + TC < d.dst_width ? ox0 + TC : d.dst_width) - 1;
      const Rect v = virtual_rect(d, oy0, oy1, ox0, ox1);
      Rect L, C;
      tile_rects(d, v, &L, &C);
      const int ps = d.pixel_stride_uv;
      // sized for the 16-byte granule (never smaller than the 4-byte one)
      const size_t lb = (size_t)(L.r1 - L.r0 + 1) * staged_pitch<16>(L.c0, L.c1);
      const size_t cb = (size_t)(C.r1 - C.r0 + 1) * staged_pitch<16>(C.c0 * ps, C.c1 * ps + (interleaved ? 1 : 0));
      const size_t total = lb + (interleaved ? cb : 2 * cb);
      if (total > worst) worst = total;
    }
  }
  return worst;
}

bool tile_plan_supported(const BandCuPlanDesc& d) {
  const char* off = getenv("BANDCU_DISABLE_TILE");
  if (off && off[0] == '1') return false;
  if (d.src_kind != BANDCU_SRC_YUV420 || d.cvt != BANDCU_CVT_YUV2RGB || d.channels != 3) return false;
  if (d.flavor != BANDCU_FLAVOR_ARGB || d.scale_mode != BANDCU_SCALE_BILINEAR_DOWN) return false;
  if (d.pixel_stride_uv != 1 && d.pixel_stride_uv != 2) return false;
  if (d.normalize ? d.out_dtype != BANDCU_DT_F32 : d.out_dtype != BANDCU_DT_U8) return false;
  // 32-bit word staging: rows must start word aligned and whole words must stay inside the plane rows
  if (d.src_width <= 0 || d.src_width % 4 != 0 || d.stride0 % 4 != 0 || d.stride_uv % 4 != 0) return false;
  if ((((d.src_width + 1) / 2) * d.pixel_stride_uv) % 4 != 0) return false;
  return true;
}

int tile_plan_create(const BandCuPlanDesc& d, TilePlan** out) {
  // the chroma layout (planar / interleaved) is a property of the pointers, known per batch; size for the larger
  const bool inter = d.pixel_stride_uv == 2;
  int TR = 32, TC = 32;
  size_t smem = tile_smem_bytes(d, TR, TC, inter);
  const size_t cap = 100 * 1024;  // keep >= 2 CTAs per SM
  while (smem > cap && (TR > 8 || TC > 8)) {
    if (TR >= TC && TR > 8) TR >>= 1; else if (TC > 8) TC >>= 1; else TR >>= 1;
    smem = tile_smem_bytes(d, TR, TC, inter);
  }
  if (smem > 200 * 1024 || TC < 32) return BANDCU_E_UNSUPPORTED;  // lanes run along the output row: keep TC == 32
  TilePlan* p = new TilePlan();
  p->lut_dev = nullptr;
  p->TR = TR; p->TC = TC; p->smem = smem;
  if (d.normalize) {
    cudaError_t e = cudaMalloc((void**)&p->lut_dev, 256 * sizeof(float));
    if (e != cudaSuccess) { delete p; return (int)e; }
    tile_build_lut_kernel<<<1, 256>>>(p->lut_dev, d.mean, d.std_);
    count_launch();
    e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { cudaFree(p->lut_dev); delete p; return (int)e; }
  }
  *out = p;
  return BANDCU_OK;
}

void tile_plan_destroy(TilePlan* p) {
  if (!p) return;
  if (p->lut_dev) (void)cudaFree(p->lut_dev);
  delete p;
}

int launch_tile(TilePlan* p, const BandCuPlanDesc& d, const BatchArgs& batch, const TableHints& hints, int n,
                cudaStream_t stream) {
  TileParams P;
  memset(&P, 0, sizeof(P));
  P.d = d;
  P.batch = batch;
  P.TR = p->TR; P.TC = p->TC;
  P.tiles_x = (d.dst_width + P.TC - 1) / P.TC;
  P.tiles_y = (d.dst_height + P.TR - 1) / P.TR;
  P.lut = p->lut_dev;
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
  void (*kern)(const TileParams);
  if (aligned16) kern = d.normalize ? tile_yuv_bilinear_down_kernel<float, 16> : tile_yuv_bilinear_down_kernel<uint8_t, 16>;
  else kern = d.normalize ? tile_yuv_bilinear_down_kernel<float, 4> : tile_yuv_bilinear_down_kernel<uint8_t, 4>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p->smem);
  if (e != cudaSuccess) return (int)e;
  const int tiles = P.tiles_x * P.tiles_y;
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
    kern<<<dim3(tiles, chunk, 1), 256, p->smem, stream>>>(Q);
    count_launch();
    done += chunk;
  }
  return (int)cudaGetLastError();
}

}  // namespace bandcu
