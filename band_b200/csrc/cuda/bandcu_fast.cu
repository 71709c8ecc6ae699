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
// Design (sm_100a, HBM-bound):
//  * persistent CTAs, one work item = R output rows of one frame;
//  * the 2 luma rows + <=2 chroma rows each output row samples are staged into
//    shared memory by the TMA engine (cp.async.bulk + mbarrier complete_tx),
//    double-buffered so the next item's rows stream in while this one computes;
//    rows the resize never samples are never read from HBM;
//  * one thread per output column: x-dependent terms (xi, xf, chroma byte
//    offsets) live in registers across the R rows;
//  * SWAR arithmetic: R|B share one 32-bit lane pair for the vertical and the
//    horizontal blend (no cross-lane carry: <= 65408 and <= 32385);
//  * results go through a shared-memory row stage and leave as 128-bit
//    coalesced stores (NHWC rows are 16-byte multiples for the named shapes).
#include <cstring>

#include "bandcu_device.cuh"
#include "bandcu_internal.h"

namespace bandcu {

struct FastParams {
  int W, H;            // source luma dims
  int stride0, stride_uv;
  int x0, y0, dx, dy, max_y, linear;
  int dw, dh;
  int R;               // output rows per work item
  int row_pitch;       // bytes per staged source row slot
  int out_row_bytes;   // dw * 3 * sizeof(OT)
  long long dst_row_stride;
  int groups_per_frame;
  long long n_items;
  int use_bulk;        // 1: cp.async.bulk staging, 0: cooperative loads
  int in_stage_bytes, out_stage_bytes;
  const void* lut;     // 256 x OT (normalize) or nullptr
  BatchArgs batch;
};

struct FastPlan {
  void* lut_dev;
  int lut_dtype;
  int num_sms;
  int block;
  int R;
};

// ---- PTX wrappers ------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra WAIT_DONE;\n"
      "bra WAIT_LOOP;\n"
      "WAIT_DONE:\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(parity)
      : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(smem_dst)),
               "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

// Source rows one output row samples.
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

// Stage the source rows of one work item. Slot layout per output row j:
//   [4j+0] luma yi   [4j+1] luma yi+1 (only if yf)   [4j+2] chroma ct   [4j+3] chroma cb (only if yf && cb != ct)
template <bool VU>
__device__ __forceinline__ void stage_item(const FastParams& P, long long item, uint8_t* stage, uint64_t* bar) {
  const int f = (int)(item / P.groups_per_frame);
  const int g = (int)(item - (long long)f * P.groups_per_frame);
  const FramePtrs fp = frame_ptrs(P.batch, f);
  const uint8_t* uv = VU ? fp.pv : fp.pu;  // first byte of each interleaved pair
  const int oy0 = g * P.R;
  const int nrows = min(P.R, P.dh - oy0);
  const int ybytes = P.W, cbytes = ((P.W + 1) >> 1) << 1;
  if (P.use_bulk) {
    // warp 0: lane 0 posts the byte count, then lanes issue one row copy each
    if (threadIdx.x < 32) {
      const int lane = threadIdx.x;
      if (lane == 0) {
        uint32_t total = 0;
        for (int j = 0; j < nrows; ++j) {
          RowSel s = row_select(P, oy0 + j);
          total += ybytes + cbytes;
          if (s.yf) total += ybytes + (s.cb != s.ct ? cbytes : 0);
        }
        mbar_expect_tx(bar, total);
      }
      __syncwarp();
      for (int k = lane; k < nrows * 4; k += 32) {
        const int j = k >> 2, slot = k & 3;
        RowSel s = row_select(P, oy0 + j);
        uint8_t* dst = stage + (size_t)k * P.row_pitch;
        if (slot == 0) bulk_g2s(dst, fp.p0 + (long long)s.yi * P.stride0, ybytes, bar);
        else if (slot == 1) { if (s.yf) bulk_g2s(dst, fp.p0 + (long long)(s.yi + 1) * P.stride0, ybytes, bar); }
        else if (slot == 2) bulk_g2s(dst, uv + (long long)s.ct * P.stride_uv, cbytes, bar);
        else { if (s.yf && s.cb != s.ct) bulk_g2s(dst, uv + (long long)s.cb * P.stride_uv, cbytes, bar); }
      }
    }
  } else {
    // cooperative byte-granular staging (any alignment / odd widths)
    for (int k = 0; k < nrows * 4; ++k) {
      const int j = k >> 2, slot = k & 3;
      RowSel s = row_select(P, oy0 + j);
      const uint8_t* src;
      int nb;
      if (slot == 0) { src = fp.p0 + (long long)s.yi * P.stride0; nb = ybytes; }
      else if (slot == 1) { if (!s.yf) continue; src = fp.p0 + (long long)(s.yi + 1) * P.stride0; nb = ybytes; }
      else if (slot == 2) { src = uv + (long long)s.ct * P.stride_uv; nb = cbytes; }
      else { if (!s.yf || s.cb == s.ct) continue; src = uv + (long long)s.cb * P.stride_uv; nb = cbytes; }
      uint8_t* dst = stage + (size_t)k * P.row_pitch;
      for (int i = threadIdx.x; i < nb; i += blockDim.x) dst[i] = src[i];
    }
  }
}

// Copy `nbytes` from shared (same 16-byte phase as dst) to global.
__device__ __forceinline__ void copy_out(const uint8_t* s, uint8_t* g, int nbytes) {
  const int head = min((int)((16 - (reinterpret_cast<uintptr_t>(g) & 15)) & 15), nbytes);
  const int mid = (nbytes - head) >> 4;
  const int tail = nbytes - head - (mid << 4);
  if ((int)threadIdx.x < head) g[threadIdx.x] = s[threadIdx.x];
  const uint4* s4 = reinterpret_cast<const uint4*>(s + head);
  uint4* g4 = reinterpret_cast<uint4*>(g + head);
  for (int i = threadIdx.x; i < mid; i += blockDim.x) g4[i] = s4[i];
  const int toff = head + (mid << 4);
  if ((int)threadIdx.x < tail) g[toff + threadIdx.x] = s[toff + threadIdx.x];
}

template <typename OT, bool VU>
__global__ void __launch_bounds__(512) fast_nv_bilinear_down_kernel(const __grid_constant__ FastParams P) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem);            // 2 barriers
  OT* lut = reinterpret_cast<OT*>(smem + 16);                     // 256 entries (<= 1024 B)
  uint8_t* in_stage0 = smem + 16 + 1024;
  uint8_t* out_stage0 = in_stage0 + 2 * (size_t)P.in_stage_bytes;

  if (threadIdx.x == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (P.lut) {
    for (int i = threadIdx.x; i < 256; i += blockDim.x) lut[i] = reinterpret_cast<const OT*>(P.lut)[i];
  }
  __syncthreads();

  uint32_t phase0 = 0, phase1 = 0;
  long long item = blockIdx.x;
  if (item < P.n_items) stage_item<VU>(P, item, in_stage0, &bars[0]);

  for (int it = 0; item < P.n_items; item += gridDim.x, ++it) {
    const int st = it & 1;
    uint8_t* in_stage = in_stage0 + (size_t)st * P.in_stage_bytes;
    uint8_t* out_stage = out_stage0 + (size_t)st * P.out_stage_bytes;
    const long long next = item + gridDim.x;
    if (next < P.n_items)
      stage_item<VU>(P, next, in_stage0 + (size_t)(st ^ 1) * P.in_stage_bytes, &bars[st ^ 1]);
    if (P.use_bulk) {
      if (st == 0) { mbar_wait(&bars[0], phase0); phase0 ^= 1; }
      else { mbar_wait(&bars[1], phase1); phase1 ^= 1; }
    } else {
      __syncthreads();
    }

    const int f = (int)(item / P.groups_per_frame);
    const int g = (int)(item - (long long)f * P.groups_per_frame);
    const int oy0 = g * P.R;
    const int nrows = min(P.R, P.dh - oy0);
    const FramePtrs fp = frame_ptrs(P.batch, f);
    uint8_t* gout = fp.out + (long long)oy0 * P.dst_row_stride;
    const bool contiguous = (P.dst_row_stride == (long long)P.out_row_bytes);
    const int phase = (int)(reinterpret_cast<uintptr_t>(gout) & 15);

    for (int ox = threadIdx.x; ox < P.dw; ox += blockDim.x) {
      const int x = P.x0 + ox * P.dx;
      const int xi = x >> 16;
      const uint32_t xf = (uint32_t)((x >> 9) & 0x7f);
      const uint32_t wx0 = 0x7fu ^ xf, wx1 = xf;
      const int cL = (xi >> 1) << 1, cR = ((xi + 1) >> 1) << 1;
#pragma unroll 2
      for (int j = 0; j < nrows; ++j) {
        const RowSel s = row_select(P, oy0 + j);
        const uint8_t* Yt = in_stage + (size_t)(4 * j) * P.row_pitch;
        const uint8_t* Ct = Yt + 2 * (size_t)P.row_pitch;
        // top row: two luma samples, their (nearest) chroma pairs
        const int yyL = yuv_luma_term(Yt[xi]), yyR = yuv_luma_term(Yt[xi + 1]);
        const uint32_t uvL = *reinterpret_cast<const uint16_t*>(Ct + cL);
        const uint32_t uvR = *reinterpret_cast<const uint16_t*>(Ct + cR);
        const ChromaTerms tL = VU ? yuv_chroma_terms(uvL >> 8, uvL & 0xFF) : yuv_chroma_terms(uvL & 0xFF, uvL >> 8);
        const ChromaTerms tR = VU ? yuv_chroma_terms(uvR >> 8, uvR & 0xFF) : yuv_chroma_terms(uvR & 0xFF, uvR >> 8);
        // packed lanes: rb = R | B<<16 ; gg = G_left | G_right<<16
        uint32_t rbL = (uint32_t)clamp_shift6(yyL + tL.cr) | ((uint32_t)clamp_shift6(yyL + tL.cb) << 16);
        uint32_t rbR = (uint32_t)clamp_shift6(yyR + tR.cr) | ((uint32_t)clamp_shift6(yyR + tR.cb) << 16);
        uint32_t gg = (uint32_t)clamp_shift6(yyL + tL.cg) | ((uint32_t)clamp_shift6(yyR + tR.cg) << 16);
        if (s.yf) {
          const uint8_t* Yb = Yt + P.row_pitch;
          const uint8_t* Cb = (s.cb != s.ct) ? Yt + 3 * (size_t)P.row_pitch : Ct;
          const int zzL = yuv_luma_term(Yb[xi]), zzR = yuv_luma_term(Yb[xi + 1]);
          const uint32_t vwL = *reinterpret_cast<const uint16_t*>(Cb + cL);
          const uint32_t vwR = *reinterpret_cast<const uint16_t*>(Cb + cR);
          const ChromaTerms bL = VU ? yuv_chroma_terms(vwL >> 8, vwL & 0xFF) : yuv_chroma_terms(vwL & 0xFF, vwL >> 8);
          const ChromaTerms bR = VU ? yuv_chroma_terms(vwR >> 8, vwR & 0xFF) : yuv_chroma_terms(vwR & 0xFF, vwR >> 8);
          const uint32_t rbL2 = (uint32_t)clamp_shift6(zzL + bL.cr) | ((uint32_t)clamp_shift6(zzL + bL.cb) << 16);
          const uint32_t rbR2 = (uint32_t)clamp_shift6(zzR + bR.cr) | ((uint32_t)clamp_shift6(zzR + bR.cb) << 16);
          const uint32_t gg2 = (uint32_t)clamp_shift6(zzL + bL.cg) | ((uint32_t)clamp_shift6(zzR + bR.cg) << 16);
          // InterpolateRow_C on two 16-bit lanes at once
          const uint32_t w1 = (uint32_t)s.yf, w0 = 256u - w1;
          rbL = ((rbL * w0 + rbL2 * w1 + 0x00800080u) >> 8) & 0x00FF00FFu;
          rbR = ((rbR * w0 + rbR2 * w1 + 0x00800080u) >> 8) & 0x00FF00FFu;
          gg = ((gg * w0 + gg2 * w1 + 0x00800080u) >> 8) & 0x00FF00FFu;
        }
        // ScaleARGBFilterCols_C: (a*(0x7f^xf) + b*xf) >> 7, truncating
        const uint32_t rb = rbL * wx0 + rbR * wx1;  // lanes <= 32385
        const uint32_t r = (rb >> 7) & 0xFFu;
        const uint32_t b = rb >> 23;
        const uint32_t gq = ((gg & 0xFFFFu) * wx0 + (gg >> 16) * wx1) >> 7;
        if (sizeof(OT) == 1) {
          uint8_t* o = out_stage + phase + (size_t)j * P.out_row_bytes + ox * 3;
          o[0] = (uint8_t)r; o[1] = (uint8_t)gq; o[2] = (uint8_t)b;
        } else {
          OT* o = reinterpret_cast<OT*>(out_stage + phase + (size_t)j * P.out_row_bytes) + ox * 3;
          o[0] = lut[r]; o[1] = lut[gq]; o[2] = lut[b];
        }
      }
    }
    __syncthreads();  // out_stage complete; every thread is done reading in_stage
    if (contiguous) {
      copy_out(out_stage + phase, gout, nrows * P.out_row_bytes);
    } else {
      for (int j = 0; j < nrows; ++j) {
        // per-row phase: rows are staged back to back, so only copy vectorised when phases agree
        const uint8_t* srow = out_stage + phase + (size_t)j * P.out_row_bytes;
        uint8_t* grow = gout + (long long)j * P.dst_row_stride;
        if (((reinterpret_cast<uintptr_t>(srow) ^ reinterpret_cast<uintptr_t>(grow)) & 15) == 0) copy_out(srow, grow, P.out_row_bytes);
        else for (int i = threadIdx.x; i < P.out_row_bytes; i += blockDim.x) grow[i] = srow[i];
      }
    }
  }
}

// Builds the Normalize table with the same device arithmetic as the generic kernel.
template <typename OT>
__global__ void build_lut_kernel(OT* lut, float mean, float sd) {
  int i = threadIdx.x;
  if (i < 256) lut[i] = normalize_elem<OT>((float)i, mean, sd);
}

static bool is_identity(const BandCuAffine& m) {
  return m.r0 == 0 && m.rr == 1 && m.rc == 0 && m.c0 == 0 && m.cr == 0 && m.cc == 1;
}

bool fast_plan_supported(const BandCuPlanDesc& d) {
  const char* off = getenv("BANDCU_DISABLE_FAST");
  if (off && off[0] == '1') return false;
  if (d.src_kind != BANDCU_SRC_YUV420 || d.pixel_stride_uv != 2) return false;
  if (d.cvt != BANDCU_CVT_YUV2RGB || d.channels != 3) return false;
  if (d.flavor != BANDCU_FLAVOR_ARGB || d.scale_mode != BANDCU_SCALE_BILINEAR_DOWN) return false;
  if (!is_identity(d.luma_map) || !is_identity(d.pre_map) || !is_identity(d.chroma_map)) return false;
  if (d.normalize ? (d.out_dtype != BANDCU_DT_F32) : false) return false;
  if (d.vw < 2 || d.vw > 8192 || d.dst_width > 4096) return false;
  // 32-bit x/y stepping inside the kernel
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
    build_lut_kernel<float><<<1, 256>>>((float*)p->lut_dev, d.mean, d.std_);
    count_launch();
    e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { cudaFree(p->lut_dev); delete p; return (int)e; }
  }
  // block: smallest multiple of 32 covering dst_width in ceil(dw/512) column passes
  const int passes = (d.dst_width + 511) / 512;
  const int per = (d.dst_width + passes - 1) / passes;
  p->block = ((per + 31) / 32) * 32;
  if (p->block < 64) p->block = 64;
  p->R = 4;
  const char* r = getenv("BANDCU_FAST_ROWS");
  if (r && atoi(r) > 0 && atoi(r) <= 16) p->R = atoi(r);
  *out = p;
  return BANDCU_OK;
}

void fast_plan_destroy(FastPlan* p) {
  if (!p) return;
  if (p->lut_dev) (void)cudaFree(p->lut_dev);
  delete p;
}

template <typename OT>
static int launch_fast_t(FastPlan* p, FastParams& P, bool vu, int grid, size_t smem, cudaStream_t stream) {
  auto kern = vu ? fast_nv_bilinear_down_kernel<OT, true> : fast_nv_bilinear_down_kernel<OT, false>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return (int)e;
  kern<<<grid, p->block, smem, stream>>>(P);
  count_launch();
  return (int)cudaGetLastError();
}

int launch_fast(FastPlan* p, const BandCuPlanDesc& d, const BatchArgs& batch, const TableHints& hints, int n,
                cudaStream_t stream) {
  bool vu;
  if (batch.table) {
    if (!hints.valid || !hints.semi_planar) return BANDCU_E_UNSUPPORTED;
    vu = hints.vu;
  } else {
    const uint8_t* pu = batch.first.pu;
    const uint8_t* pv = batch.first.pv;
    if (pv == pu + 1) vu = false;       // NV12: U first
    else if (pu == pv + 1) vu = true;   // NV21: V first
    else return BANDCU_E_UNSUPPORTED;
  }
  FastParams P;
  memset(&P, 0, sizeof(P));
  P.W = d.vw; P.H = d.vh;
  P.stride0 = d.stride0; P.stride_uv = d.stride_uv;
  P.x0 = d.x0; P.y0 = d.y0; P.dx = d.dx; P.dy = d.dy; P.max_y = d.max_y; P.linear = d.linear;
  P.dw = d.dst_width; P.dh = d.dst_height;
  P.R = p->R;
  const int esz = d.normalize ? 4 : 1;
  P.out_row_bytes = d.dst_width * 3 * esz;
  P.dst_row_stride = d.dst_row_stride;
  const int cbytes = ((d.vw + 1) >> 1) << 1;
  P.row_pitch = ((cbytes + 15) / 16) * 16 + 16;  // +16: xi+1 may touch one byte past the row when xf == 0
  P.in_stage_bytes = P.R * 4 * P.row_pitch;
  P.out_stage_bytes = ((P.R * P.out_row_bytes + 15) / 16) * 16 + 32;
  P.groups_per_frame = (d.dst_height + P.R - 1) / P.R;
  P.n_items = (long long)n * P.groups_per_frame;
  P.lut = p->lut_dev;
  P.batch = batch;
  bool aligned = (d.vw % 16 == 0) && (d.stride0 % 16 == 0) && (d.stride_uv % 16 == 0);
  if (batch.table) {
    aligned = aligned && hints.aligned16;
  } else {
    const uint8_t* uv0 = vu ? batch.first.pv : batch.first.pu;
    aligned = aligned && ((reinterpret_cast<uintptr_t>(batch.first.p0) & 15) == 0) &&
              ((reinterpret_cast<uintptr_t>(uv0) & 15) == 0) && (n == 1 || batch.in_stride % 16 == 0);
  }
  P.use_bulk = aligned ? 1 : 0;
  const char* nb = getenv("BANDCU_FAST_NO_BULK");
  if (nb && nb[0] == '1') P.use_bulk = 0;
  const size_t smem = 16 + 1024 + 2 * (size_t)P.in_stage_bytes + 2 * (size_t)P.out_stage_bytes;
  if (smem > 200 * 1024) return BANDCU_E_UNSUPPORTED;
  // persistent grid: as many CTAs per SM as shared memory / threads allow
  int per_sm = (int)((220 * 1024) / (smem + 1024));
  int by_threads = 2048 / p->block;
  if (per_sm > by_threads) per_sm = by_threads;
  if (per_sm < 1) per_sm = 1;
  if (per_sm > 8) per_sm = 8;
  long long want = (long long)p->num_sms * per_sm;
  int grid = (int)(P.n_items < want ? P.n_items : want);
  if (grid < 1) grid = 1;
  return d.normalize ? launch_fast_t<float>(p, P, vu, grid, smem, stream)
                     : launch_fast_t<uint8_t>(p, P, vu, grid, smem, stream);
}

}  // namespace bandcu
