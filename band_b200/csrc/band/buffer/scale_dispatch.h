// band/buffer/scale_dispatch.h -- host-side choice of the scaler code path.
//
// The reference's Resize hands every request to libyuv with kFilterBilinear
// (/root/reference/band/buffer/libyuv_image_operator.cc:1175,1219,1373); libyuv
// then picks one of several row/column algorithms whose ROUNDING differs
// (SURVEY Appendix A.2-A.4). The fused kernels implement each algorithm; this
// file restates libyuv's dispatch rules (scale_common.cc ScaleFilterReduce /
// ScaleSlope, scale_argb.cc ScaleARGB, scale.cc ScalePlane) so the plan
// compiler selects the same one for a given (src dims, dst dims, flavour).
#ifndef BAND_BUFFER_SCALE_DISPATCH_H_
#define BAND_BUFFER_SCALE_DISPATCH_H_

namespace band {
namespace cuda {

// Filled into BandCuPlanDesc {flavor, scale_mode, linear, x0, y0, dx, dy, max_y}.
struct ScaleSpec {
  int flavor = 0;  // BANDCU_FLAVOR_*
  int mode = 0;    // BANDCU_SCALE_*
  int linear = 0;  // 1: no vertical blend (filter reduced below bilinear)
  int x0 = 0, y0 = 0, dx = 0, dy = 0;  // 16.16 source stepping
  int max_y = 0;
};

// 16.16 helpers (libyuv scale_common.cc FixedDiv_C / FixedDiv1_C).
int FixedDiv(int num, int div);
int FixedDiv1(int num, int div);

// ARGBScale(..., kFilterBilinear) on a src_w x src_h image of 4-byte (or, for
// RGB, 3 used bytes of 4) pixels -> dst_w x dst_h.
ScaleSpec DispatchArgbScale(int src_w, int src_h, int dst_w, int dst_h);
// UVScale(..., kFilterBilinear) on an interleaved 2-channel chroma plane (inside NV12Scale,
// libyuv_image_operator.cc:422): ScaleARGB's rules plus the exact-2x up-samplers of scale_uv.cc.
ScaleSpec DispatchUvScale(int src_w, int src_h, int dst_w, int dst_h);
// ScalePlane(..., kFilterBilinear) on an 8-bit plane.
ScaleSpec DispatchPlaneScale(int src_w, int src_h, int dst_w, int dst_h);

}  // namespace cuda
}  // namespace band
#endif  // BAND_BUFFER_SCALE_DISPATCH_H_
