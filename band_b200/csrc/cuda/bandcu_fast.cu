// band_b200/csrc/cuda/bandcu_fast.cu
//
// The hot kernel: NV12/NV21 frame batch -> RGB -> bilinear down-scale ->
// [Normalize] fused into one pass, written straight into the tensor.
//
// It evaluates exactly what the reference materialises stage by stage:
//   ColorSpaceConvert(kRGB)   libyuv_image_operator.cc:92-111 / :180-199  (NV12ToRAW / NV21ToRAW)
//   Resize(w,h)               :1146-1187  (RGB24ToARGB -> ARGBScale bilinear -> ARGBToRGB24)
//   Normalize(mean,std)       common_operator.cc:24-40
// for the general ScaleARGBBilinearDown path (non-integer ratios), i.e. every
// BASELINE config with a 4:2:0 semi-planar source. Integer results are bit
// exact; the float stage is a 256-entry table built with IEEE sub/div.
//
// Design (sm_100a). ncu on the first version showed the kernel is NOT limited by
// HBM (it never reads the rows the resize skips: 2.06 MB of DRAM traffic per
// frame against 3.71 MB algorithmic) but by instruction issue and by one block
// barrier per work item. Hence:
//  * persistent, warp-specialised CTAs. Warp 0 is the producer: it posts one
//    cp.async.bulk (TMA engine) per needed source row into an S-deep ring of
//    shared-memory stages (mbarrier complete_tx). The other warps consume:
//    wait(full[s]) -> compute -> arrive(empty[s]). No __syncthreads in the loop;
//    warps drift apart freely and the ring keeps HBM requests in flight;
//  * one work item = R output rows of one frame; consumer warps are split into
//    G row groups so one staged item feeds G*ceil(dw/32) warps;
//  * one thread per output column: the x-dependent terms (byte offsets, packed
//    7-bit weights) live in registers across rows;
//  * packed 16-bit arithmetic on the LEFT|RIGHT source pixel pair: BT.601
//    clamps as one VIADDMNMX.S16x2.RELU (DPX) per channel pair, vertical blend
//    as two IMADs on both lanes, and the horizontal blend of a pair as ONE
//    32-bit multiply ((L|R<<16) * (w1|w0<<16) >> 23);
//  * results leave as direct global stores (a warp writes 384 contiguous bytes
//    per row); L2 merges them into full sectors.
#include <cstdlib>
#include <cstring>

#include "bandcu_device.cuh"
#include "bandcu_internal.h"

namespace bandcu {

constexpr int kMaxStages = 4;
constexpr int kMaxRows = 16;

struct FastParams {
  int W, H;            // source luma dims
  int stride0, stride_uv;
  int x0, y0, dx, dy, max_y, linear;
  int dw, dh;
  int R;               // output rows per work item
  int S;               // ring depth
  int cols;            // consumer threads per row group (multiple of 32)
  int passes;          // column passes per row
  int row_pitch;       // bytes per staged source row slot
  int stage_bytes;     // R * 4 * row_pitch
  long long dst_row_stride;
  int groups_per_frame;
  long long n_items;
  int use_bulk;        // 1: cp.async.bulk staging, 0: producer-warp byte copies
  long long luma_off, chroma_off;  // Crop folded into the frame: byte offsets of the crop origin in the luma / chroma planes
  int v_off;           // planar chroma: byte offset of the V row inside a chroma slot (the U row sits at 0)
  int out_row_bytes;   // dw * 3 * sizeof(OT)
  int out_buf_bytes;   // BULKOUT: R * out_row_bytes (one of two output staging buffers)
  int off_out;         // BULKOUT: byte offset of the output staging buffers in shared memory
  const void* lut;     // 256 x float (normalize) or nullptr
  BatchArgs batch;
};

struct FastPlan {
  void* lut_dev;
  int num_sms;
  int R, S;
  int bulk_store;  // results leave through shared memory + cp.async.bulk stores (outputs in pinned host memory)
  int no_bulk;     // BANDCU_FAST_NO_BULK=1 when the plan was created: stage with producer-warp byte copies
  // occupancy of the last launch configuration (kernel, block, shared memory): asked of the runtime once, not per launch
  // (cudaFuncSetAttribute + cudaOccupancyMaxActiveBlocksPerMultiprocessor cost ~2 us, a fifth of a single-frame call)
  const void* occ_kern;
  int occ_block, occ_per_sm;
  size_t occ_smem;
};

// ---- PTX wrappers ------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// try_wait parks the warp in hardware until the phase completes or the time limit passes; with the default limit a
// DRAM-bound launch still executed more wait-loop instructions than arithmetic (ncu: 214 K warp-instructions per frame on the
// north star against 93 K of real work), which costs issue slots and -- under the 1000 W cap of a sustained run -- clock.
// A 20 us suspend hint keeps the wake-up on completion and removes most of the polling.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n"
      "@p bra WAIT_DONE;\n"
      "bra WAIT_LOOP;\n"
      "WAIT_DONE:\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(parity), "r"(20000u)
      : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(smem_dst)),
               "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

__device__ __forceinline__ void bulk_s2g(void* gdst, uint32_t smem_src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void named_bar_sync(int id, int threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

// Source rows one output row samples (ScaleARGBBilinearDown row stepping).
struct RowSel {
  int yi, yf;      // luma row and 8-bit vertical fraction
  int ct, cb;      // chroma rows for luma rows yi and yi+1
};
__device__ __forceinline__ RowSel row_select(const FastParams& P, int oy) {
  long long y = (long long)P.y0 + (long long)oy * P.dy;
  if (y > P.max_y) y = P.max_y;
  RowSel s;
  s.yi = (int)(y >> 16);
  s.yf = P.linear ? 0 : (int)((y >> 8) & 255);
  s.ct = s.yi >> 1;
  s.cb = (s.yi + 1) >> 1;
  return s;
}

// (frame, row group) of the work items a CTA visits: item = blockIdx.x + k * gridDim.x. One division at the start;
// afterwards the pair advances by a constant step (the per-item 64-bit division was 4 % of all instructions).
struct ItemWalk {
  int f, g, step_f, step_g, gpf;
  __device__ __forceinline__ explicit ItemWalk(const FastParams& P) {
    gpf = P.groups_per_frame;
    f = (int)blockIdx.x / gpf;
    g = (int)blockIdx.x - f * gpf;
    step_f = (int)gridDim.x / gpf;
    step_g = (int)gridDim.x - step_f * gpf;
  }
  __device__ __forceinline__ void next() {
    f += step_f;
    g += step_g;
    if (g >= gpf) { g -= gpf; ++f; }
  }
};

// Shared-memory map: barriers | per-stage row info | Normalize table | ring of stages.
constexpr int kOffFull = 0;                              // kMaxStages x u64
constexpr int kOffEmpty = 8 * kMaxStages;                // kMaxStages x u64
constexpr int kOffRowInfo = 16 * kMaxStages;             // kMaxStages x kMaxRows x u32
constexpr int kOffLut = kOffRowInfo + 4 * kMaxStages * kMaxRows;  // replicated 256 x f32 table
constexpr int kLutBytes = 256 * 4 * 8;  // 256 entries x kLutCopies
constexpr int kOffStages = ((kOffLut + kLutBytes + 127) / 128) * 128;

// Chroma layout of the source (template parameter CH).
enum { kChromaUV = 0 /* NV12: U,V interleaved */, kChromaVU = 1 /* NV21 */, kChromaPlanar = 2 /* YV12 / YV21 */ };

// Producer (warp 0): stage the source rows of one work item. Slot layout per output row j:
//   [4j+0] luma yi   [4j+1] luma yi+1 (only if yf)   [4j+2] chroma ct   [4j+3] chroma cb (only if yf && cb != ct)
// rowinfo[j] = yf | 0x100 when the bottom luma row has its own chroma row (slot 3).
// Planar chroma: a chroma slot holds the U row at offset 0 and the V row at offset P.v_off (two copies).
template <int CH>
__device__ __forceinline__ void produce_item(const FastParams& P, int f, int g, uint8_t* stage, uint32_t* rowinfo,
                                             uint64_t* full) {
  const int lane = threadIdx.x;
  FramePtrs fp = frame_ptrs(P.batch, f);
  fp.p0 += P.luma_off;
  fp.pu += P.chroma_off;
  fp.pv += P.chroma_off;
  const uint8_t* uv = CH == kChromaVU ? fp.pv : fp.pu;  // first byte of each interleaved pair; planar: the U plane
  const int hb = (P.W + 1) >> 1;                         // planar: bytes per chroma row of one plane
  const int oy0 = g * P.R;
  const int nrows = min(P.R, P.dh - oy0);
  const int ybytes = P.W, cbytes = ((P.W + 1) >> 1) << 1;
  if (lane < nrows) {
    const RowSel s = row_select(P, oy0 + lane);
    rowinfo[lane] = (uint32_t)s.yf | ((s.yf && s.cb != s.ct) ? 0x100u : 0u);
  }
  __syncwarp();
  if (P.use_bulk) {
    if (lane == 0) {
      uint32_t total = 0;
      for (int j = 0; j < nrows; ++j) {
        const RowSel s = row_select(P, oy0 + j);
        total += ybytes + cbytes;
        if (s.yf) total += ybytes + (s.cb != s.ct ? cbytes : 0);
      }
      mbar_expect_tx(full, total);
    }
    __syncwarp();
    for (int k = lane; k < nrows * 4; k += 32) {
      const int j = k >> 2, slot = k & 3;
      const RowSel s = row_select(P, oy0 + j);
      uint8_t* dst = stage + (size_t)k * P.row_pitch;
      if (slot == 0) bulk_g2s(dst, fp.p0 + (long long)s.yi * P.stride0, ybytes, full);
      else if (slot == 1) { if (s.yf) bulk_g2s(dst, fp.p0 + (long long)(s.yi + 1) * P.stride0, ybytes, full); }
      else if (slot == 2 || (s.yf && s.cb != s.ct)) {
        const long long off = (long long)(slot == 2 ? s.ct : s.cb) * P.stride_uv;
        if (CH == kChromaPlanar) {
          bulk_g2s(dst, fp.pu + off, hb, full);
          bulk_g2s(dst + P.v_off, fp.pv + off, hb, full);
        } else {
          bulk_g2s(dst, uv + off, cbytes, full);
        }
      }
    }
  } else {
    // byte-granular staging by the producer warp (any alignment / odd widths)
    for (int k = 0; k < nrows * 4; ++k) {
      const int j = k >> 2, slot = k & 3;
      const RowSel s = row_select(P, oy0 + j);
      const uint8_t* src;
      int nb;
      if (slot == 0) { src = fp.p0 + (long long)s.yi * P.stride0; nb = ybytes; }
      else if (slot == 1) { if (!s.yf) continue; src = fp.p0 + (long long)(s.yi + 1) * P.stride0; nb = ybytes; }
      else if (slot == 2) { src = uv + (long long)s.ct * P.stride_uv; nb = cbytes; }
      else { if (!s.yf || s.cb == s.ct) continue; src = uv + (long long)s.cb * P.stride_uv; nb = cbytes; }
      uint8_t* dst = stage + (size_t)k * P.row_pitch;
      if (CH == kChromaPlanar && slot >= 2) {
        const uint8_t* srcv = fp.pv + (src - fp.pu);
        for (int i = lane; i < hb; i += 32) { dst[i] = src[i]; dst[P.v_off + i] = srcv[i]; }
      } else {
        for (int i = lane; i < nb; i += 32) dst[i] = src[i];
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(full);  // release: the warp's stores above are visible to waiters
  }
}

// ---- packed pixel arithmetic ---------------------------------------------------
// Everything below works on the LEFT and RIGHT source pixels of an output column at once: a register holds
// value_left | value_right << 16 (8-bit values or 16-bit terms in 16-bit lanes).
//
// libyuv YuvPixel (kYuvI601Constants):  yy = (Y * 0x0101 * 18997) >> 16
//   R = clamp((yy + 102 V           - 14216) >> 6)
//   G = clamp((yy - 25 U - 52 V     +  8696) >> 6)
//   B = clamp((yy + 128 U           - 17544) >> 6)
// clamp(x >> 6, 0, 255) == clamp(x, 0, 16383) >> 6. The R and G sums fit signed 16 bits, so one
// VIADDMNMX.S16x2.RELU does add + upper clamp + lower clamp for both pixels. The B sum can exceed 32767, so it
// runs biased by +17544 in unsigned lanes.
__device__ __forceinline__ uint32_t lds_u8(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ uint32_t lds_u16(uint32_t addr) {
  uint16_t v;
  asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ uint32_t lds_u32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ float lds_f32(uint32_t addr) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
  return v;
}

// Chroma-only terms of the pixel pair; shared by the top and the bottom luma row when both use one chroma row.
struct ChromaPair {
  uint32_t kr;  // 102 V - 14216            (s16 lanes)
  uint32_t kg;  // 8696 - 25 U - 52 V       (s16 lanes)
  uint32_t kb;  // 128 U                    (u16 lanes)
};
template <int CH>
__device__ __forceinline__ ChromaPair chroma_pair(uint32_t a_left, uint32_t a_right, uint32_t v_off) {
  uint32_t u2, v2;  // U_left | U_right << 16, V_left | V_right << 16
  if (CH == kChromaPlanar) {
    u2 = lds_u8(a_left) | (lds_u8(a_right) << 16);
    v2 = lds_u8(a_left + v_off) | (lds_u8(a_right + v_off) << 16);
  } else {
    const uint32_t w = __byte_perm(lds_u16(a_left), lds_u16(a_right), 0x5410);
    u2 = __byte_perm(w, 0u, CH == kChromaVU ? 0x4341 : 0x4240);
    v2 = __byte_perm(w, 0u, CH == kChromaVU ? 0x4240 : 0x4341);
  }
  ChromaPair c;
  c.kr = __vadd2(v2 * 102u, 0xC878C878u /* -14216 */);
  const uint32_t t = u2 * 25u + v2 * 52u;  // lanes <= 19635
  c.kg = __vadd2(~t, 0x21F921F9u /* 8696 + 1: ~t == -t - 1 per lane */);
  c.kb = u2 << 7;
  return c;
}
// luma pair at byte address a (left pixel) and a + 1 (right pixel) -> 8-bit R, G, B pairs
__device__ __forceinline__ void luma_pair_to_rgb(uint32_t a, const ChromaPair& c, uint32_t& R2, uint32_t& G2,
                                                 uint32_t& B2) {
  const uint32_t pL = lds_u8(a) * 4882229u;
  const uint32_t pR = lds_u8(a + 1) * 4882229u;
  const uint32_t yy = __byte_perm(pL, pR, 0x7632);  // (pL >> 16) | (pR >> 16) << 16, lanes <= 18997
  const uint32_t r = __viaddmin_s16x2_relu(yy, c.kr, 0x3FFF3FFFu);
  const uint32_t g = __viaddmin_s16x2_relu(yy, c.kg, 0x3FFF3FFFu);
  uint32_t b = __vmaxu2(__viaddmin_u16x2(yy, c.kb, 0x84878487u /* 17544 + 16383 */), 0x44884488u /* 17544 */);
  b = __vadd2(b, 0xBB78BB78u /* -17544 */);
  R2 = (r >> 6) & 0x00FF00FFu;
  G2 = (g >> 6) & 0x00FF00FFu;
  B2 = (b >> 6) & 0x00FF00FFu;
}
// InterpolateRow_C on both lanes: (top*(256-f) + bottom*f + 128) >> 8; f == 0 keeps top whatever bottom holds.
__device__ __forceinline__ uint32_t vblend2(uint32_t top, uint32_t bot, uint32_t w0, uint32_t w1) {
  // bytes 1 and 3 of the sums are the two results: one PRMT instead of shift + mask
  return __byte_perm(top * w0 + bot * w1 + 0x00800080u, 0u, 0x4341);
}
// ScaleARGBFilterCols_C on a packed pair: (L*(0x7f^xf) + R*xf) >> 7, truncating. wp = xf | (0x7f^xf) << 16, so
// (L | R<<16) * wp = L*xf + (L*(0x7f^xf) + R*xf) << 16 (+ a term beyond bit 31); L*xf < 65536 never carries.
// Returns the product: the 8-bit result sits in bits 23..30 (bits 24..31 with the weights shifted left once, as the kernel passes them).
__device__ __forceinline__ uint32_t hblend_pair_raw(uint32_t v2, uint32_t wp) { return v2 * wp; }

constexpr int kLutCopies = 8;  // the Normalize table is replicated so that lanes l and l' (l != l' mod 8) never collide

// BULKOUT: the tensors live in PINNED HOST memory. Direct 12-byte stores would cross PCIe as 32-byte write
// TLPs (measured: 5x slower than a copy); instead an item's rows are assembled in one of two shared-memory
// output buffers and leave as cp.async.bulk stores (whole rows, maximum-payload TLPs), which overlap the
// producer's bulk loads of the next items on the other direction of the link (scripts/pcie_tma_probe.cu:
// 49 GB/s in + 26 GB/s out at once, against 52 GB/s total when a copy engine drains a staging buffer).
template <typename OT, int CH, bool BULKOUT>
__global__ void __launch_bounds__(1024) fast_nv_bilinear_down_kernel(const __grid_constant__ FastParams P) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + kOffFull);
  uint64_t* empty = reinterpret_cast<uint64_t*>(smem + kOffEmpty);
  uint32_t* rowinfo = reinterpret_cast<uint32_t*>(smem + kOffRowInfo);
  float* lut = reinterpret_cast<float*>(smem + kOffLut);
  uint8_t* stages = smem + kOffStages;
  const int consumer_warps = P.cols >> 5;

  if (threadIdx.x == 0) {
    for (int s = 0; s < P.S; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], consumer_warps);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (P.lut) {  // lut[v * kLutCopies + k] = table[v]
    for (int i = threadIdx.x; i < 256 * kLutCopies; i += blockDim.x)
      lut[i] = reinterpret_cast<const float*>(P.lut)[i / kLutCopies];
  }
  __syncthreads();

  if (threadIdx.x < 32) {
    // ---------------- producer warp ----------------
    int s = 0;
    uint32_t wrap = 0;  // how many times the ring has wrapped
    ItemWalk w(P);
    for (long long item = blockIdx.x; item < P.n_items; item += gridDim.x, w.next()) {
      if (wrap) mbar_wait(&empty[s], (wrap - 1) & 1);  // consumers released the previous use of this stage
      produce_item<CH>(P, w.f, w.g, stages + (size_t)s * P.stage_bytes, rowinfo + s * kMaxRows, &full[s]);
      if (++s == P.S) { s = 0; ++wrap; }
    }
    return;
  }

  // ---------------- consumer warps: one thread per output column ----------------
  const int col = threadIdx.x - 32;
  const int lane = threadIdx.x & 31;
  const uint32_t sbase = smem_u32(smem);
  const uint32_t lut_lane = sbase + kOffLut + (uint32_t)(lane & (kLutCopies - 1)) * 4u;
  const uint32_t pitch = (uint32_t)P.row_pitch;
  int s = 0;
  uint32_t wrap = 0;
  uint32_t ob = 0;  // BULKOUT: which output buffer this item fills
  const long long orow_stride = BULKOUT ? (long long)P.out_row_bytes : P.dst_row_stride;
  ItemWalk walk(P);
  for (long long item = blockIdx.x; item < P.n_items; item += gridDim.x, walk.next()) {
    const uint32_t stage = sbase + kOffStages + (uint32_t)s * (uint32_t)P.stage_bytes;
    const uint32_t ri = sbase + kOffRowInfo + (uint32_t)s * (kMaxRows * 4u);
    const int f = walk.f, g = walk.g;
    const int oy0 = g * P.R;
    const int nrows = min(P.R, P.dh - oy0);
    uint8_t* const gdst = frame_ptrs(P.batch, f).out + (long long)oy0 * P.dst_row_stride;
    uint8_t* gout = gdst;
    if (BULKOUT) {
      // the bulk store issued two items ago has finished READING this buffer
      if (col == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
      named_bar_sync(1, P.cols);
      gout = smem + P.off_out + ob * (uint32_t)P.out_buf_bytes;
    }
    mbar_wait(&full[s], wrap & 1);

    for (int pass = 0; pass < P.passes; ++pass) {
      const int ox = pass * P.cols + col;
      if (ox < P.dw) {
        const int x = P.x0 + ox * P.dx;
        const uint32_t xi = (uint32_t)(x >> 16);
        const uint32_t xf = (uint32_t)((x >> 9) & 0x7f);
        // column weights pre-shifted by one bit: the blended byte then lands in bits 24..31 of the product (a plain shift
        // extracts it; no mask) -- 2 L xf < 65536 still cannot carry into the upper half
        const uint32_t wp = (xf << 1) | ((0x7fu ^ xf) << 17);
        // byte offsets inside one staged row group: luma pair, left / right chroma pair (slot 2 = top chroma row)
        const uint32_t csh = CH == kChromaPlanar ? 0u : 1u;  // planar rows hold one byte per chroma sample
        const uint32_t cL = 2u * pitch + ((xi >> 1) << csh), cR = 2u * pitch + (((xi + 1u) >> 1) << csh);
        const uint32_t v_off = (uint32_t)P.v_off;
        uint32_t row = stage;                                   // slot 0 of output row j
        uint8_t* o = gout + (size_t)ox * (3 * sizeof(OT));      // output pixel of row j
#pragma unroll 2
        for (int j = 0; j < nrows; ++j, row += 4u * pitch, o += orow_stride) {
          const uint32_t info = lds_u32(ri + 4u * (uint32_t)j);
          const uint32_t w1 = info & 0xFFu, w0 = 256u - w1;
          const ChromaPair ctop = chroma_pair<CH>(row + cL, row + cR, v_off);
          ChromaPair cbot = ctop;
          if (info & 0x100u) cbot = chroma_pair<CH>(row + cL + pitch, row + cR + pitch, v_off);  // slot 3: warp-uniform
          uint32_t tr, tg, tb, br, bg, bb;
          luma_pair_to_rgb(row + xi, ctop, tr, tg, tb);
          luma_pair_to_rgb(row + xi + pitch, cbot, br, bg, bb);
          const uint32_t r = hblend_pair_raw(vblend2(tr, br, w0, w1), wp);
          const uint32_t gq = hblend_pair_raw(vblend2(tg, bg, w0, w1), wp);
          const uint32_t b = hblend_pair_raw(vblend2(tb, bb, w0, w1), wp);
          if (sizeof(OT) == 1) {
            if (P.lut) {  // Normalize into a one-byte type (int8 / uint8 model inputs): the table holds the byte
              o[0] = (uint8_t)lds_u32(lut_lane + (r >> 24) * (4u * kLutCopies));
              o[1] = (uint8_t)lds_u32(lut_lane + (gq >> 24) * (4u * kLutCopies));
              o[2] = (uint8_t)lds_u32(lut_lane + (b >> 24) * (4u * kLutCopies));
            } else {
              o[0] = (uint8_t)(r >> 24); o[1] = (uint8_t)(gq >> 24); o[2] = (uint8_t)(b >> 24);
            }
          } else {
            // table index v = raw >> 24; byte offset v * 4 * kLutCopies + this lane's copy: one shift and one multiply-add
            float* of = reinterpret_cast<float*>(o);
            of[0] = lds_f32(lut_lane + (r >> 24) * (4u * kLutCopies));
            of[1] = lds_f32(lut_lane + (gq >> 24) * (4u * kLutCopies));
            of[2] = lds_f32(lut_lane + (b >> 24) * (4u * kLutCopies));
          }
        }
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty[s]);
    if (++s == P.S) { s = 0; ++wrap; }
    if (BULKOUT) {
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> visible to the bulk engine
      named_bar_sync(1, P.cols);
      if (col == 0) {
        const uint32_t src = sbase + (uint32_t)P.off_out + ob * (uint32_t)P.out_buf_bytes;
        if (P.dst_row_stride == (long long)P.out_row_bytes) {
          bulk_s2g(gdst, src, (uint32_t)(nrows * P.out_row_bytes));
        } else {
          for (int j = 0; j < nrows; ++j)
            bulk_s2g(gdst + (long long)j * P.dst_row_stride, src + (uint32_t)(j * P.out_row_bytes), (uint32_t)P.out_row_bytes);
        }
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
      ob ^= 1u;
    }
  }
  if (BULKOUT && col == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");  // stores complete before the CTA retires
}

// Builds the Normalize table with the same device arithmetic as the generic kernel.
// One-byte Normalize outputs: the 32-bit table entry holds the output byte (same element arithmetic, then the cast).
template <typename T>
__global__ void build_byte_lut_kernel(uint32_t* lut, float mean, float sd) {
  const int i = threadIdx.x;
  if (i < 256) lut[i] = (uint32_t)(uint8_t)normalize_elem<T>((float)i, mean, sd);
}
__global__ void build_lut_kernel(float* lut, float mean, float sd) {
  int i = threadIdx.x;
  if (i < 256) lut[i] = normalize_elem<float>((float)i, mean, sd);
}

// Crop only (a translation): the crop origin becomes a byte offset into each plane and the kernel runs unchanged on
// the cropped width. CropNv / CropYv take the chroma origin at (x0/2, y0/2) (libyuv_image_operator.cc:950-1037), which
// is what the plan's chroma map carries; a crop AFTER the conversion (pre_map translation) samples chroma at
// ((y0+r)>>1, (x0+c)>>1), the same thing only for even origins.
static bool is_translation(const BandCuAffine& m) { return m.rr == 1 && m.rc == 0 && m.cr == 0 && m.cc == 1 && m.r0 >= 0 && m.c0 >= 0; }
static bool crop_only(const BandCuPlanDesc& d) {
  return is_translation(d.luma_map) && is_translation(d.pre_map) && is_translation(d.chroma_map) &&
         (d.pre_map.r0 & 1) == 0 && (d.pre_map.c0 & 1) == 0;
}


static int env_int(const char* name, int lo, int hi, int dflt) {
  const char* v = getenv(name);
  if (!v || !*v) return dflt;
  const int x = atoi(v);
  return (x < lo || x > hi) ? dflt : x;
}

bool fast_plan_supported(const BandCuPlanDesc& d) {
  const char* off = getenv("BANDCU_DISABLE_FAST");
  if (off && off[0] == '1') return false;
  if (d.src_kind != BANDCU_SRC_YUV420 || (d.pixel_stride_uv != 2 && d.pixel_stride_uv != 1)) return false;
  if (d.cvt != BANDCU_CVT_YUV2RGB || d.channels != 3) return false;
  if (d.flavor != BANDCU_FLAVOR_ARGB || d.scale_mode != BANDCU_SCALE_BILINEAR_DOWN) return false;
  if (!crop_only(d)) return false;  // identity maps included
  if (d.normalize && d.out_dtype != BANDCU_DT_F32 && d.out_dtype != BANDCU_DT_I8 && d.out_dtype != BANDCU_DT_U8) return false;
  if (d.vw < 2 || d.vw > 8192 || d.dst_width > 4096) return false;
  // 32-bit x stepping inside the kernel
  if ((long long)d.x0 + (long long)d.dst_width * d.dx > 0x7fffffffLL) return false;
  return true;
}

int fast_plan_create(const BandCuPlanDesc& d, FastPlan** out) {
  FastPlan* p = new FastPlan();
  memset(p, 0, sizeof(*p));
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e == cudaSuccess) e = cudaDeviceGetAttribute(&p->num_sms, cudaDevAttrMultiProcessorCount, dev);
  if (e != cudaSuccess) { delete p; return (int)e; }
  if (d.normalize) {
    e = cudaMalloc(&p->lut_dev, 256 * sizeof(float));
    if (e != cudaSuccess) { delete p; return (int)e; }
    if (d.out_dtype == BANDCU_DT_I8) build_byte_lut_kernel<int8_t><<<1, 256>>>((uint32_t*)p->lut_dev, d.mean, d.std_);
    else if (d.out_dtype == BANDCU_DT_U8) build_byte_lut_kernel<uint8_t><<<1, 256>>>((uint32_t*)p->lut_dev, d.mean, d.std_);
    else build_lut_kernel<<<1, 256>>>((float*)p->lut_dev, d.mean, d.std_);
    count_launch();
    e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { cudaFree(p->lut_dev); delete p; return (int)e; }
  }
  p->R = env_int("BANDCU_FAST_ROWS", 1, kMaxRows, 0);  // 0: chosen per launch from the source row length
  p->S = env_int("BANDCU_FAST_STAGES", 2, kMaxStages, 2);
  p->no_bulk = env_int("BANDCU_FAST_NO_BULK", 0, 1, 0);  // read here, once: nothing on the launch path touches the environment
  *out = p;
  return BANDCU_OK;
}

void fast_plan_destroy(FastPlan* p) {
  if (!p) return;
  if (p->lut_dev) (void)cudaFree(p->lut_dev);
  delete p;
}

typedef void (*FastKernel)(const FastParams);
template <typename OT>
static FastKernel pick_kernel(int ch, bool bulk_out) {
  if (bulk_out) {
    return ch == kChromaPlanar ? fast_nv_bilinear_down_kernel<OT, kChromaPlanar, true>
           : ch == kChromaVU   ? fast_nv_bilinear_down_kernel<OT, kChromaVU, true>
                               : fast_nv_bilinear_down_kernel<OT, kChromaUV, true>;
  }
  return ch == kChromaPlanar ? fast_nv_bilinear_down_kernel<OT, kChromaPlanar, false>
         : ch == kChromaVU   ? fast_nv_bilinear_down_kernel<OT, kChromaVU, false>
                             : fast_nv_bilinear_down_kernel<OT, kChromaUV, false>;
}

void fast_plan_set_bulk_store(FastPlan* p, int on) {
  if (p) p->bulk_store = on ? 1 : 0;
}

int launch_fast(FastPlan* p, const BandCuPlanDesc& d, const BatchArgs& batch, const TableHints& hints, int n,
                cudaStream_t stream) {
  int ch;
  if (d.pixel_stride_uv == 1) {
    ch = kChromaPlanar;
    if (batch.table ? !hints.valid : (!batch.first.pu || !batch.first.pv)) return BANDCU_E_UNSUPPORTED;
  } else if (batch.table) {
    if (!hints.valid || !hints.semi_planar) return BANDCU_E_UNSUPPORTED;
    ch = hints.vu ? kChromaVU : kChromaUV;
  } else {
    const uint8_t* pu = batch.first.pu;
    const uint8_t* pv = batch.first.pv;
    if (pv == pu + 1) ch = kChromaUV;        // NV12: U first
    else if (pu == pv + 1) ch = kChromaVU;   // NV21: V first
    else return BANDCU_E_UNSUPPORTED;
  }
  const bool vu = ch == kChromaVU;
  FastParams P;
  memset(&P, 0, sizeof(P));
  P.W = d.vw; P.H = d.vh;
  P.stride0 = d.stride0; P.stride_uv = d.stride_uv;
  P.x0 = d.x0; P.y0 = d.y0; P.dx = d.dx; P.dy = d.dy; P.max_y = d.max_y; P.linear = d.linear;
  P.dw = d.dst_width; P.dh = d.dst_height;
  P.dst_row_stride = d.dst_row_stride;
  const int cbytes = ((d.vw + 1) >> 1) << 1;
  P.row_pitch = ((cbytes + 15) / 16) * 16 + 16;  // +16: xi+1 may touch one byte past the row when xf == 0
  P.v_off = ((((d.vw + 1) >> 1) + 15) / 16) * 16;
  P.luma_off = (long long)d.luma_map.r0 * d.stride0 + d.luma_map.c0;
  P.chroma_off = (long long)(d.chroma_map.r0 + d.pre_map.r0 / 2) * d.stride_uv +
                 (long long)(d.chroma_map.c0 + d.pre_map.c0 / 2) * d.pixel_stride_uv;
  // geometry: rows per item, row groups, ring depth; shrink until the ring fits shared memory
  // rows per work item: 4 for wide sources (1080p: 31 KB per stage), 8 for narrow ones (640 wide: the same stage size,
  // half the barrier traffic; measured +11 % on 640x480 -> 224, -25 % on 1080p)
  // wide rows with >= 288 output columns: 6 (two resident CTAs still hold >= 20 consumer warps per SM; +4-5 % on 300 and 513 wide
  // outputs, -3 % on 224 wide ones, which stay at 4)
  P.R = p->R ? p->R : (P.row_pitch <= 1024 ? 8 : (d.dst_width >= 288 ? 6 : 4));
  P.S = p->S;
  // small batches (the single-frame latency case): fewer rows per item so that every SM gets work
  while (P.R > 1 && (long long)n * ((d.dst_height + P.R - 1) / P.R) < 2LL * p->num_sms) P.R >>= 1;
  P.passes = (d.dst_width + 511) / 512;
  P.cols = (((d.dst_width + P.passes - 1) / P.passes) + 31) / 32 * 32;
  const bool f32_out = d.normalize && d.out_dtype == BANDCU_DT_F32;
  const int esize = f32_out ? 4 : 1;
  P.out_row_bytes = d.dst_width * 3 * esize;
  // bulk stores need 16-byte granularity everywhere (row length, row stride, every output pointer)
  bool out_aligned;
  if (batch.table) out_aligned = hints.valid && hints.out_aligned16;
  else out_aligned = ((reinterpret_cast<uintptr_t>(batch.first.out) & 15) == 0) && (n == 1 || batch.out_stride % 16 == 0);
  const bool bulk_out = p->bulk_store && out_aligned && (P.out_row_bytes % 16 == 0) && (d.dst_row_stride % 16 == 0);
  const size_t smem_cap = 226 * 1024;
  auto need = [&](int S, int R) {
    size_t b = kOffStages + (size_t)S * R * 4 * P.row_pitch;
    if (bulk_out) b = ((b + 127) / 128) * 128 + 2 * (size_t)R * P.out_row_bytes;
    return b;
  };
  while (P.S > 2 && need(P.S, P.R) > smem_cap) --P.S;
  while (P.R > 1 && need(P.S, P.R) > smem_cap) --P.R;
  P.stage_bytes = P.R * 4 * P.row_pitch;
  P.out_buf_bytes = P.R * P.out_row_bytes;
  P.off_out = (int)(((kOffStages + (size_t)P.S * P.stage_bytes + 127) / 128) * 128);
  const size_t smem = need(P.S, P.R);
  if (smem > smem_cap) return BANDCU_E_UNSUPPORTED;
  const int block = 32 + P.cols;
  P.groups_per_frame = (d.dst_height + P.R - 1) / P.R;
  P.n_items = (long long)n * P.groups_per_frame;
  P.lut = p->lut_dev;
  P.batch = batch;
  bool aligned = (d.vw % 16 == 0) && (d.stride0 % 16 == 0) && (d.stride_uv % 16 == 0) && (P.luma_off % 16 == 0) &&
                 (P.chroma_off % 16 == 0);
  if (ch == kChromaPlanar) {
    aligned = aligned && (d.vw % 32 == 0);  // each plane's chroma row is its own 16-byte granular copy
    if (batch.table) aligned = aligned && hints.planes_aligned16;
    else aligned = aligned && ((reinterpret_cast<uintptr_t>(batch.first.p0) | reinterpret_cast<uintptr_t>(batch.first.pu) |
                                reinterpret_cast<uintptr_t>(batch.first.pv)) & 15) == 0 && (n == 1 || batch.in_stride % 16 == 0);
  } else if (batch.table) {
    aligned = aligned && hints.aligned16;
  } else {
    const uint8_t* uv0 = vu ? batch.first.pv : batch.first.pu;
    aligned = aligned && ((reinterpret_cast<uintptr_t>(batch.first.p0) & 15) == 0) &&
              ((reinterpret_cast<uintptr_t>(uv0) & 15) == 0) && (n == 1 || batch.in_stride % 16 == 0);
  }
  P.use_bulk = aligned ? 1 : 0;
  if (p->no_bulk) P.use_bulk = 0;
  // persistent grid: every SM holds as many CTAs as shared memory, threads and registers allow
  FastKernel kern = f32_out ? pick_kernel<float>(ch, bulk_out) : pick_kernel<uint8_t>(ch, bulk_out);
  int per_sm = 1;
  if (p->occ_kern == (const void*)kern && p->occ_block == block && p->occ_smem == smem) {
    per_sm = p->occ_per_sm;
  } else {
    // the attribute belongs to the kernel, not to this plan: every plan raises it to the same cap, so that a plan with a
    // smaller ring never lowers it under another plan's cached geometry
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_cap);
    if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, block, smem);
    if (e != cudaSuccess) return (int)e;
    p->occ_kern = (const void*)kern; p->occ_block = block; p->occ_smem = smem; p->occ_per_sm = per_sm;
  }
  if (per_sm < 1) return BANDCU_E_UNSUPPORTED;
  long long want = (long long)p->num_sms * per_sm;
  int grid = (int)(P.n_items < want ? P.n_items : want);
  if (grid < 1) grid = 1;
  kern<<<grid, block, smem, stream>>>(P);
  count_launch();
  return (int)cudaGetLastError();
}

}  // namespace bandcu
