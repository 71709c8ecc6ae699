// bandcu_noalloc.cu -- a two-stage chain whose intermediate image lives in a file-scope device array instead of
// cudaMalloc'ed plan scratch (bandcu_two_stage_run_noalloc, include/bandcu.h).
//
// No counterpart in the reference; added for callers that cannot allocate device
// memory (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc). The ordinary route for a
// chain that cannot be one launch -- e.g. 4:2:0 -> RGBA, then the bilinear up-scale of that image -- is
// FusedPlan's scratch image (bandcu_malloc, fused_plan.cc EnsureCapacity) plus one BandCuPlan per stage (each may own a
// cudaMalloc'ed Normalize table). Here both stages run on the pixel-per-thread kernel (fused_generic_kernel, which needs
// no plan object and computes the same values as every staged kernel), and the second stage reads the image the first
// one wrote into g_noalloc_scratch below.
//
// Rules of the array:
//   * BANDCU_NOALLOC_SCRATCH_BYTES (8 MiB: one 1080p RGBA image) per GPU that loads libbandcu.so, held until exit;
//   * in HOST code the name g_noalloc_scratch is a stand-in, not a device address: the address comes from
//     cudaGetSymbolAddress, once per device with that device current (scratch_for_current_device), and THAT pointer is
//     what the kernels receive;
//   * every call shares the array and finds the previous call's bytes in it. Nothing has to start at zero: stage 1
//     writes every pixel stage 2 may read (checked below: stage 2's index map must stay inside stage 1's image);
//   * calls that overlap on the GPU (two streams, two host threads) race on it. One call at a time per GPU.
// The entry point calls no cudaMalloc / cudaFree / cudaMallocHost and no allocating library routine.
#include <cuda_runtime.h>

#include <cstdio>
#include <mutex>

#include "bandcu.h"
#include "bandcu_device.cuh"
#include "bandcu_internal.h"

namespace bandcu {
namespace {

static __device__ __align__(256) unsigned char g_noalloc_scratch[BANDCU_NOALLOC_SCRATCH_BYTES];

constexpr int kMaxDevices = 64;
std::mutex g_mu;
unsigned char* g_scratch_ptr[kMaxDevices] = {};  // device address of g_noalloc_scratch on device i (0: not resolved yet)

// Device address of the array on the CURRENT device; resolved once per device. cudaGetSymbolAddress allocates nothing:
// the array is part of the module image that the runtime loads with the library.
int scratch_for_current_device(unsigned char** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return report_error((int)e, "two_stage_run_noalloc: cudaGetDevice");
  if (dev < 0 || dev >= kMaxDevices) return report_error(BANDCU_E_UNSUPPORTED, "two_stage_run_noalloc: device index above 63");
  std::lock_guard<std::mutex> lock(g_mu);
  if (!g_scratch_ptr[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess) return report_error((int)e, "two_stage_run_noalloc: cudaGetSymbolAddress(g_noalloc_scratch)");
    if (!p) return report_error(BANDCU_E_INVALID, "two_stage_run_noalloc: cudaGetSymbolAddress returned a null address");
    g_scratch_ptr[dev] = static_cast<unsigned char*>(p);
  }
  *out = g_scratch_ptr[dev];
  return BANDCU_OK;
}

int channels_of_source(int src_kind) {
  switch (src_kind) {
    case BANDCU_SRC_GRAY: return 1;
    case BANDCU_SRC_RGB: return 3;
    case BANDCU_SRC_RGBA: return 4;
    default: return 0;
  }
}

// every pixel the second stage can fetch lies inside the w x h image of the first stage (an affine map takes its extremes
// at the corners of the virtual image)
bool map_stays_inside(const BandCuAffine& m, int vw, int vh, int w, int h) {
  for (int k = 0; k < 4; ++k) {
    const long long r = (k & 1) ? vh - 1 : 0, c = (k & 2) ? vw - 1 : 0;
    const long long row = m.r0 + m.rr * r + m.rc * c, col = m.c0 + m.cr * r + m.cc * c;
    if (row < 0 || row >= h || col < 0 || col >= w) return false;
  }
  return true;
}

}  // namespace
}  // namespace bandcu

using namespace bandcu;

extern "C" int bandcu_two_stage_run_noalloc(const BandCuPlanDesc* first, const BandCuPlanDesc* second,
                                            const BandCuFrame* frame, void* stream) {
  if (!first || !second || !frame) return report_error(BANDCU_E_INVALID, "two_stage_run_noalloc: null argument");
  if (!frame->p0 || !frame->out) return report_error(BANDCU_E_INVALID, "two_stage_run_noalloc: null plane / output pointer");
  if (first->src_kind == BANDCU_SRC_YUV420 && first->cvt != BANDCU_CVT_YUV2GRAY && (!frame->pu || !frame->pv))
    return report_error(BANDCU_E_INVALID, "two_stage_run_noalloc: null chroma pointer");
  // stage 1 leaves a u8 image with 1, 3 or 4 channels
  if (first->normalize || first->out_dtype != BANDCU_DT_U8)
    return report_error(BANDCU_E_INVALID, "two_stage_run_noalloc: the first stage must store u8 pixels (no Normalize)");
  if (first->dst_width <= 0 || first->dst_height <= 0 || second->dst_width <= 0 || second->dst_height <= 0 ||
      second->vw <= 0 || second->vh <= 0)
    return report_error(BANDCU_E_INVALID, "two_stage_run_noalloc: non-positive image dimension");
  const int ch = channels_of_source(second->src_kind);
  if (ch == 0 || ch != first->channels)
    return report_error(BANDCU_E_INVALID,
                        "two_stage_run_noalloc: the second stage's source must be the gray / RGB / RGBA image the first stage stores");
  if (first->dst_row_stride < (long long)first->dst_width * ch || second->stride0 != first->dst_row_stride)
    return report_error(BANDCU_E_INVALID,
                        "two_stage_run_noalloc: second->stride0 must equal first->dst_row_stride (>= dst_width * channels)");
  const long long need = first->dst_row_stride * (long long)first->dst_height;
  if (need > (long long)BANDCU_NOALLOC_SCRATCH_BYTES) {
    char msg[256];
    snprintf(msg, sizeof(msg),
             "two_stage_run_noalloc: the intermediate image needs %lld bytes, the static scratch holds %lld "
             "(BANDCU_NOALLOC_SCRATCH_BYTES); use the plan API for larger frames",
             need, (long long)BANDCU_NOALLOC_SCRATCH_BYTES);
    return report_error(BANDCU_E_INVALID, msg);
  }
  if (!map_stays_inside(second->luma_map, second->vw, second->vh, first->dst_width, first->dst_height))
    return report_error(BANDCU_E_INVALID, "two_stage_run_noalloc: the second stage's index map leaves the first stage's image");
  unsigned char* scratch = nullptr;
  int e = scratch_for_current_device(&scratch);
  if (e != BANDCU_OK) return e;
  cudaStream_t s = (cudaStream_t)stream;
  BatchArgs b;
  b.in_stride = b.out_stride = 0;
  b.table = nullptr;
  // stage 1: caller's planes -> the array
  b.first.p0 = (const uint8_t*)frame->p0;
  b.first.pu = (const uint8_t*)frame->pu;
  b.first.pv = (const uint8_t*)frame->pv;
  b.first.out = scratch;
  e = launch_generic(*first, b, 1, s);
  if (e != 0) return report_error(e, "two_stage_run_noalloc: first stage launch");
  // stage 2: the array -> caller's output (same stream: ordered behind stage 1)
  b.first.p0 = scratch;
  b.first.pu = b.first.pv = nullptr;
  b.first.out = (uint8_t*)frame->out;
  e = launch_generic(*second, b, 1, s);
  if (e != 0) return report_error(e, "two_stage_run_noalloc: second stage launch");
  return BANDCU_OK;
}
