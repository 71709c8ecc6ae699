// band_b200/csrc/cuda/bandcu_turn.cu
//
// Quarter-turns (Rotate 90 / 270, alone or composed with flips) of 4:2:0 planes, gray, interleaved chroma pairs, RGB
// and RGBA images as WORD-granular transposes. Replaces RotatePlane / I420Rotate / NV12ToI420Rotate + MergeUVPlane /
// ARGBRotate behind LibyuvImageOperator::Rotate (/root/reference/band/buffer/libyuv_image_operator.cc:697-887): a pure
// permutation, so the roofline is HBM (bytes in + bytes out).
//
// The byte-staged kernels in bandcu_generic.cu scatter every source byte into shared memory with its own store and
// gather every output byte with its own load (eight byte-wide shared accesses, shifts and ORs per output word):
// 0.42 of the measured HBM bandwidth on 1080p NV21, 0.20 on RGB. Here
//   * a CTA owns one output tile of 32 words x (32 * EPW) rows, EPW = elements per word (4, 2, 1 for 1-, 2-, 4-byte
//     elements); its source tile is 32 words x (32 * EPW) rows as well;
//   * load: lanes along the source row (128 contiguous bytes per warp and row), 4 - 16 word loads per thread in flight,
//     each word stored ONCE into a [rows][33]-word shared array indexed by the OUTPUT column it feeds;
//   * turn: a thread reads the EPW words of EPW consecutive output columns at one source word column (lane split chosen
//     so that the 32 lanes hit 32 banks), transposes the EPW x EPW block in registers (PRMT), and stores EPW words to
//     EPW output rows; the lanes of one store cover 32 - 128 contiguous bytes of one output row;
//   * RGB (3-byte elements): source rows are loaded as words and re-cut into one word per pixel with two shuffles and a
//     funnel shift, turned as 4-byte elements, and four pixels are packed back into three words on the way out.
// Needs word-aligned planes (pointers, strides, tile origins); anything else stays on the byte-staged kernels.
#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "bandcu_device.cuh"
#include "bandcu_internal.h"

namespace bandcu {
namespace {

constexpr int kPitch = 33;  // words per shared row

// What a turn kernel needs to know about one plane, precomputed on the host.
struct TurnPlane {
  const uint8_t* src;
  uint8_t* dst;
  int src_stride, dst_stride;
  int w, h;          // output elements
  int src_w;         // source elements per row that may be read (= h of the output)
  int r0, rc;        // source row    = r0 + rc * (output column)
  int c0, cr;        // source column = c0 + cr * (output row)
  int esize;
};
struct TurnParams {
  TurnPlane pl[3];   // the planes that exist, packed: blockIdx.z = np * frame + plane
  int np;
  long long src_step, dst_step;
};

// EPW x EPW element transpose of EPW words a[k] (k = consecutive output columns = shared rows): result o[b] holds, for
// source element b of the word, the EPW elements of output columns k = 0 .. EPW-1 in ascending order.
template <int E>
__device__ __forceinline__ void turn_block(const uint32_t (&a)[4], uint32_t (&o)[4]) {
  if (E == 1) {
    const uint32_t t0 = __byte_perm(a[0], a[1], 0x5140), t1 = __byte_perm(a[2], a[3], 0x5140);
    const uint32_t t2 = __byte_perm(a[0], a[1], 0x7362), t3 = __byte_perm(a[2], a[3], 0x7362);
    o[0] = __byte_perm(t0, t1, 0x5410);
    o[1] = __byte_perm(t0, t1, 0x7632);
    o[2] = __byte_perm(t2, t3, 0x5410);
    o[3] = __byte_perm(t2, t3, 0x7632);
  } else if (E == 2) {
    o[0] = __byte_perm(a[0], a[1], 0x5410);
    o[1] = __byte_perm(a[0], a[1], 0x7632);
  } else {
    o[0] = a[0];
  }
}

// E: 1, 2, 4 = element bytes; 3 = RGB (turned as 4-byte pixels, four shared rows per block like E == 1).
template <int E>
__device__ __forceinline__ void turn_tile(const TurnPlane& pl, long long src_off, long long dst_off, uint32_t* sm) {
  constexpr int EB = E;                        // element bytes in memory
  constexpr int EPW = E == 1 ? 4 : E == 2 ? 2 : 1;  // source elements per loaded "unit" word after unpacking
  constexpr int KB = E == 3 ? 4 : EPW;         // shared rows per turned block (output columns per stored group)
  constexpr int NU = 32 * KB;                  // output columns (= source rows) per tile
  constexpr int NV = 32 * EPW;                 // output rows (= source elements along a row) per tile
  const int c_lo = blockIdx.x * NU, r_lo = blockIdx.y * NV;
  if (c_lo >= pl.w || r_lo >= pl.h) return;
  const int lane = threadIdx.x, wid = threadIdx.y;
  const uint8_t* src = pl.src + src_off;
  // source element of output row r: c0 + cr * r. The tile's lowest source element:
  const int se_lo = pl.cr > 0 ? pl.c0 + r_lo : pl.c0 - (r_lo + NV - 1);
  constexpr int JB = KB == 4 ? 2 : KB == 2 ? 1 : 0, JL = 1 << JB, IL = 32 >> JB;  // lanes along j / along i
  constexpr int NIG = 32 / IL;                                                   // i groups (j groups = 32 / JL)
  const int i_low = lane >> JB, j_low = lane & (JL - 1);
  uint8_t* dst = pl.dst + dst_off;
  if (c_lo + NU <= pl.w && r_lo + NV <= pl.h && se_lo >= 0) {
    // ---- interior tile: every word whole, no guards. All of a thread's loads are issued before its first shared store
    // (NU / 8 = 4 - 16 words in flight per thread), source and destination pointers advance by constants.
    constexpr int NL = NU / 8;
    uint32_t v[NL];
    const uint8_t* row = src + (long long)(pl.r0 + pl.rc * (c_lo + wid)) * pl.src_stride + (long long)se_lo * EB;
    const long long rstep = (long long)(8 * pl.rc) * pl.src_stride;
    if (E == 3) {
      const int wq = (3 * lane) >> 2, sh = ((3 * lane) & 3) * 8;
#pragma unroll
      for (int k = 0; k < NL; ++k) v[k] = lane < 24 ? __ldg(reinterpret_cast<const uint32_t*>(row + k * rstep) + lane) : 0u;
#pragma unroll
      for (int k = 0; k < NL; ++k) {
        const uint32_t lo = __shfl_sync(0xffffffffu, v[k], wq), hi = __shfl_sync(0xffffffffu, v[k], min(wq + 1, 23));
        sm[(wid + 8 * k) * kPitch + lane] = __funnelshift_r(lo, hi, sh);
      }
    } else {
#pragma unroll
      for (int k = 0; k < NL; ++k) v[k] = __ldg(reinterpret_cast<const uint32_t*>(row + k * rstep) + lane);
#pragma unroll
      for (int k = 0; k < NL; ++k) sm[(wid + 8 * k) * kPitch + lane] = v[k];
    }
    __syncthreads();
    const long long ostep = (long long)pl.cr * pl.dst_stride;  // from source element b to b + 1: one output row down or up
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const int q = wid * 4 + t;
      const int i = (q % NIG) * IL + i_low, j = (q / NIG) * JL + j_low;
      uint32_t a[4] = {0, 0, 0, 0}, o[4];
#pragma unroll
      for (int k = 0; k < KB; ++k) a[k] = sm[(KB * i + k) * kPitch + j];
      const int e0 = EPW * j;
      const int r = r_lo + (pl.cr > 0 ? e0 : NV - 1 - e0);
      uint8_t* p = dst + (long long)r * pl.dst_stride + (long long)(c_lo + KB * i) * EB;
      if (E == 3) {
        uint32_t* pw = reinterpret_cast<uint32_t*>(p);
        pw[0] = __byte_perm(a[0], a[1], 0x4210);
        pw[1] = __byte_perm(a[1], a[2], 0x5421);
        pw[2] = __byte_perm(a[2], a[3], 0x6542);
      } else {
        turn_block<E>(a, o);
#pragma unroll
        for (int b = 0; b < EPW; ++b) *reinterpret_cast<uint32_t*>(p + b * ostep) = o[b];
      }
    }
    return;
  }
  // ---- edge tile: element-wise guards
  // load: shared row u <- the source row that feeds output column c_lo + u; lane j = unit word j of the tile
#pragma unroll 4
  for (int u = wid; u < NU; u += 8) {
    const int c = c_lo + u;
    uint32_t v = 0;
    if (c < pl.w) {
      const uint8_t* row = src + (long long)(pl.r0 + pl.rc * c) * pl.src_stride;
      if (E == 3) {
        // 32 pixels = 96 bytes = 24 words from byte offset se_lo * 3 (a multiple of 96, see the host checks); pixel q
        // starts at byte 3 q: words q*3/4 and the next one, funnel-shifted
        const int e0 = se_lo;  // first source pixel of the tile (may be negative in a clipped edge tile)
        uint32_t wv = 0;
        if (lane < 24) {
          const int b = e0 * 3 + lane * 4;  // byte offset of this word in the row
          if (b >= 0 && b + 4 <= pl.src_w * 3) wv = __ldg(reinterpret_cast<const uint32_t*>(row + b));
          else {
            for (int k = 0; k < 4; ++k)
              if (b + k >= 0 && b + k < pl.src_w * 3) wv |= (uint32_t)__ldg(row + b + k) << (8 * k);
          }
        }
        const int wq = (3 * lane) >> 2, sh = ((3 * lane) & 3) * 8;
        const uint32_t lo = __shfl_sync(0xffffffffu, wv, wq), hi = __shfl_sync(0xffffffffu, wv, min(wq + 1, 23));
        v = __funnelshift_r(lo, hi, sh);
      } else {
        const int e = se_lo + lane * EPW;  // first source element of this word
        if (e >= 0 && e + EPW <= pl.src_w) v = __ldg(reinterpret_cast<const uint32_t*>(row + (long long)e * EB));
        else {
          for (int k = 0; k < EPW * EB; ++k) {
            const int ek = e + k / EB;
            if (ek >= 0 && ek < pl.src_w) v |= (uint32_t)__ldg(row + (long long)e * EB + k) << (8 * k);
          }
        }
      }
    }
    sm[u * kPitch + lane] = v;
  }
  __syncthreads();
  // ---- turn and store. Lane split (i_low, j_low): j_low takes log2(KB) bits so that the KB-row stride of i and the
  // unit stride of j fill all 32 banks (pitch 33 = 1 mod 32: bank = KB * i + k + j).
#pragma unroll
  for (int t = 0; t < 4; ++t) {
    const int q = wid * 4 + t;
    const int i = (q % NIG) * IL + i_low;   // output word group: output columns c_lo + KB * i ..
    const int j = (q / NIG) * JL + j_low;   // unit word column of the source tile
    uint32_t a[4] = {0, 0, 0, 0}, o[4];
#pragma unroll
    for (int k = 0; k < KB; ++k) a[k] = sm[(KB * i + k) * kPitch + j];
    const int c = c_lo + KB * i;            // first output column of the group
    if (c >= pl.w) continue;
    if (E == 3) {
      // a[k] = pixel (R,G,B,x) of output column c + k, all for source pixel j: one output row, twelve bytes
      const int e = j;                                        // source element inside the tile
      const int v = pl.cr > 0 ? e : NV - 1 - e;               // output row inside the tile
      const int r = r_lo + v;
      if (r >= pl.h) continue;
      uint8_t* p = dst + (long long)r * pl.dst_stride + (long long)c * 3;
      if (c + 4 <= pl.w) {
        uint32_t* pw = reinterpret_cast<uint32_t*>(p);
        pw[0] = __byte_perm(a[0], a[1], 0x4210);
        pw[1] = __byte_perm(a[1], a[2], 0x5421);
        pw[2] = __byte_perm(a[2], a[3], 0x6542);
      } else {
        for (int k = 0; k < 4 && c + k < pl.w; ++k) {
          p[3 * k] = (uint8_t)a[k]; p[3 * k + 1] = (uint8_t)(a[k] >> 8); p[3 * k + 2] = (uint8_t)(a[k] >> 16);
        }
      }
      continue;
    }
    turn_block<E>(a, o);
#pragma unroll
    for (int b = 0; b < EPW; ++b) {
      const int e = EPW * j + b;                              // source element inside the tile
      const int v = pl.cr > 0 ? e : NV - 1 - e;               // output row inside the tile
      const int r = r_lo + v;
      if (r >= pl.h) continue;
      uint8_t* p = dst + (long long)r * pl.dst_stride + (long long)c * EB;
      if (c + EPW <= pl.w) {
        *reinterpret_cast<uint32_t*>(p) = o[b];
      } else {
        for (int k = 0; k < (pl.w - c) * EB; ++k) p[k] = (uint8_t)(o[b] >> (8 * k));
      }
    }
  }
}

template <int E>
__global__ void __launch_bounds__(256) plane_quarter_turn_kernel(const __grid_constant__ TurnParams P) {
  __shared__ uint32_t sm[(E == 1 || E == 3 ? 128 : E == 2 ? 64 : 32) * kPitch];
  const int f = blockIdx.z / P.np;
  const TurnPlane& pl = P.pl[blockIdx.z - f * P.np];
  turn_tile<E>(pl, (long long)f * P.src_step, (long long)f * P.dst_step, sm);
}
// 4:2:0 frames mix element sizes (luma 1 byte, interleaved chroma 2): one launch covers both
__global__ void __launch_bounds__(256) plane_quarter_turn_mixed_kernel(const __grid_constant__ TurnParams P) {
  __shared__ uint32_t sm[128 * kPitch];
  const int f = blockIdx.z / P.np;
  const TurnPlane& pl = P.pl[blockIdx.z - f * P.np];
  const long long so = (long long)f * P.src_step, dof = (long long)f * P.dst_step;
  if (pl.esize == 1) turn_tile<1>(pl, so, dof, sm);
  else turn_tile<2>(pl, so, dof, sm);
}

}  // namespace

int launch_plane_turns(const PlaneMapParams& in, int n, cudaStream_t stream) {
  static const bool off = getenv("BANDCU_DISABLE_WORD_TURNS") && getenv("BANDCU_DISABLE_WORD_TURNS")[0] == '1';
  if (off) return BANDCU_E_UNSUPPORTED;
  TurnParams P;
  memset(&P, 0, sizeof(P));
  P.src_step = in.src_step;
  P.dst_step = in.dst_step;
  int gx = 1, gy = 1, esizes = 0;
  for (int k = 0; k < 3; ++k) {
    const PlaneMap& pm = in.pl[k];
    if (!pm.src) continue;
    const BandCuAffine& m = pm.m;
    if (!(m.rr == 0 && m.cc == 0 && (m.rc == 1 || m.rc == -1) && (m.cr == 1 || m.cr == -1))) return BANDCU_E_UNSUPPORTED;
    const int E = pm.esize;
    if (E < 1 || E > 4) return BANDCU_E_UNSUPPORTED;
    // word granularity: plane pointers, strides and frame steps; the source element range [c0 - (h-1), c0] or
    // [c0, c0 + h-1] must start on a word (cr > 0: c0 * E; cr < 0: (c0 + 1) * E, tiles are cut from the top of the range)
    const uintptr_t bits = reinterpret_cast<uintptr_t>(pm.src) | reinterpret_cast<uintptr_t>(pm.dst) | (uintptr_t)pm.src_stride |
                           (uintptr_t)pm.dst_stride | (n == 1 ? 0 : ((uintptr_t)in.src_step | (uintptr_t)in.dst_step));
    if (bits & 3) return BANDCU_E_UNSUPPORTED;
    const int edge = m.cr > 0 ? m.c0 : m.c0 + 1;
    if ((edge * E) & 3) return BANDCU_E_UNSUPPORTED;
    if (E == 3 && (edge % 4) != 0) return BANDCU_E_UNSUPPORTED;  // 32-pixel tiles must start on a 96-byte = word boundary
    TurnPlane& t = P.pl[P.np++];
    t.src = pm.src; t.dst = pm.dst;
    t.src_stride = pm.src_stride; t.dst_stride = pm.dst_stride;
    t.w = pm.w; t.h = pm.h;
    t.r0 = m.r0; t.rc = m.rc; t.c0 = m.c0; t.cr = m.cr;
    t.esize = E;
    // elements of a source row that may be read: up to the highest element the map touches
    t.src_w = (m.cr > 0 ? m.c0 + pm.h : m.c0 + 1);
    if (m.cr < 0 && m.c0 - (pm.h - 1) < 0) return BANDCU_E_UNSUPPORTED;
    const int epw = E == 1 ? 4 : E == 2 ? 2 : 1, kb = E == 3 ? 4 : epw;
    gx = std::max(gx, (pm.w + 32 * kb - 1) / (32 * kb));
    gy = std::max(gy, (pm.h + 32 * epw - 1) / (32 * epw));
    esizes |= 1 << E;
  }
  if (esizes == 0) return BANDCU_E_UNSUPPORTED;
  const bool mixed = (esizes & ~((1 << 1) | (1 << 2))) == 0 && esizes == ((1 << 1) | (1 << 2));
  if (!mixed && (esizes & (esizes - 1)) != 0) return BANDCU_E_UNSUPPORTED;  // other mixes do not occur
  const dim3 block(32, 8, 1);
  for (int done = 0; done < n;) {
    const int chunk = n - done > 21845 ? 21845 : n - done;  // gridDim.z <= 65535
    TurnParams Q = P;
    for (int k = 0; k < P.np; ++k) {
      Q.pl[k].src += (long long)done * P.src_step;
      Q.pl[k].dst += (long long)done * P.dst_step;
    }
    const dim3 grid(gx, gy, P.np * chunk);
    if (mixed) plane_quarter_turn_mixed_kernel<<<grid, block, 0, stream>>>(Q);
    else if (esizes == (1 << 1)) plane_quarter_turn_kernel<1><<<grid, block, 0, stream>>>(Q);
    else if (esizes == (1 << 2)) plane_quarter_turn_kernel<2><<<grid, block, 0, stream>>>(Q);
    else if (esizes == (1 << 3)) plane_quarter_turn_kernel<3><<<grid, block, 0, stream>>>(Q);
    else plane_quarter_turn_kernel<4><<<grid, block, 0, stream>>>(Q);
    count_launch();
    done += chunk;
  }
  return (int)cudaGetLastError();
}

}  // namespace bandcu
