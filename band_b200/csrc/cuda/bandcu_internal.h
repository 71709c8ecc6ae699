// band_b200/csrc/cuda/bandcu_internal.h -- host-side declarations shared by the
// .cu translation units of libbandcu.so (not part of the C-ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "bandcu.h"

namespace bandcu {
struct BatchArgs;

void count_launch();
// records the calling thread's bandcu_last_error_string() and returns `code` (cudaError_t or BANDCU_E_*)
int report_error(int code, const char* what);
int launch_generic(const BandCuPlanDesc& d, const BatchArgs& batch, int n, cudaStream_t stream);
int launch_yuv_map(const BandCuYuvMapDesc& d, const BandCuYuvPtrs& p, long long src_step, long long dst_step, int n,
                   cudaStream_t stream);
int launch_rgb_to_yuv(const BandCuRgbToYuvDesc& d, const BandCuRgbToYuvPtrs& p, long long src_step, long long dst_step,
                      int n, cudaStream_t stream);
int launch_normalize(const void* in, int in_dtype, void* out, int out_dtype, long long n, float mean, float sd,
                     cudaStream_t stream);

// Fast (staged) path for the NV12/NV21 -> RGB -> bilinear-down [-> Normalize] chain.
struct FastPlan;
bool fast_plan_supported(const BandCuPlanDesc& d);
int fast_plan_create(const BandCuPlanDesc& d, FastPlan** out);
void fast_plan_destroy(FastPlan* p);
// results leave through shared memory + bulk stores (tensors in pinned host memory) when the batch allows it
void fast_plan_set_bulk_store(FastPlan* p, int on);
// What the host knows about a pointer-table batch (the table itself lives in device memory).
struct TableHints {
  bool valid = false;    // false: strided batch, derive everything from BatchArgs::first
  bool semi_planar = false;  // every frame has |pu - pv| == 1 with the same order
  bool vu = false;       // V first (NV21)
  bool aligned16 = false;  // every luma / chroma base pointer is 16-byte aligned
  bool out_aligned16 = false;  // every output pointer is 16-byte aligned
  bool planes_aligned4 = false;  // every p0 / pu / pv pointer is 4-byte aligned
  bool planes_aligned16 = false;  // ... 16-byte aligned
};
// returns BANDCU_E_UNSUPPORTED if this particular batch (layout) cannot use it
int launch_fast(FastPlan* p, const BandCuPlanDesc& d, const BatchArgs& batch, const TableHints& hints, int n,
                cudaStream_t stream);
// Tile-staged 4:2:0 -> RGB -> bilinear-down [-> Normalize] behind any Crop/Rotate/Flip index map, planar or
// semi-planar chroma (bandcu_tile.cu).
struct TilePlan;
bool tile_plan_supported(const BandCuPlanDesc& d);
int tile_plan_create(const BandCuPlanDesc& d, TilePlan** out);  // BANDCU_E_UNSUPPORTED: tiles do not fit shared memory
void tile_plan_destroy(TilePlan* p);
int launch_tile(TilePlan* p, const BandCuPlanDesc& d, const BatchArgs& batch, const TableHints& hints, int n,
                cudaStream_t stream);
// Packed (gray / RGB / RGBA) bilinear Resize without index maps: four output pixels per thread.
struct PackedPlan;
bool packed_plan_supported(const BandCuPlanDesc& d);
int packed_plan_create(const BandCuPlanDesc& d, PackedPlan** out);
void packed_plan_destroy(PackedPlan* p);
int launch_packed(PackedPlan* p, const BandCuPlanDesc& d, const BatchArgs& batch, const TableHints& hints, int n,
                  cudaStream_t stream);
// Same-size conversions (ColorSpaceConvert without a Resize, optionally behind a Crop): four pixels per thread, 32-bit
// accesses (bandcu_copy.cu). BANDCU_E_UNSUPPORTED: another chain, or a batch that is not word aligned.
int launch_copy_rows(const BandCuPlanDesc& d, const BatchArgs& batch, const TableHints& hints, int n, cudaStream_t stream);
// Down-scaling one u8 plane (gray / Y / U / V, or interleaved chroma): blended row pair in shared memory, then the column
// filter (bandcu_planes.cu). BANDCU_E_UNSUPPORTED for anything else.
int launch_plane_rows(const BandCuPlanDesc& d, const BatchArgs& batch, const TableHints& hints, int n, cudaStream_t stream);
// One plane of an index-mapped copy (Crop / Rotate / Flip of a 4:2:0 plane or of a packed image).
struct PlaneMap {
  const uint8_t* src;
  uint8_t* dst;
  int src_stride, dst_stride;
  int w, h;        // output elements
  int esize;       // bytes per element: 1 .. 4
  int vec;         // output rows are word aligned
  BandCuAffine m;  // output element (r, c) -> source element (row, col)
};
struct PlaneMapParams {
  PlaneMap pl[3];
  long long src_step, dst_step;
};
// Quarter-turns as word-granular register / shared-memory transposes (bandcu_turn.cu). BANDCU_E_UNSUPPORTED: some plane
// is not a quarter-turn or misses the word alignment the kernel needs; the byte-staged kernels take it.
int launch_plane_turns(const PlaneMapParams& P, int n, cudaStream_t stream);
}  // namespace bandcu
