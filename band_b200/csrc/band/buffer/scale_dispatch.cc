// band/buffer/scale_dispatch.cc -- see scale_dispatch.h.
#include "band/buffer/scale_dispatch.h"

#include <cstdint>

#include "bandcu.h"

namespace band {
namespace cuda {
namespace {

enum Filter { kNone = 0, kLinear = 1, kBilinear = 2 };

// ScaleFilterReduce for a bilinear request: a dimension that is copied 1:1, is
// a single sample, or shrinks by exactly 3 (centre sample lands on a pixel)
// needs no blend along that axis.
Filter ReduceBilinear(int sw, int sh, int dw, int dh) {
  Filter f = kBilinear;
  if (sh == 1 || dh == sh || dh * 3 == sh) f = kLinear;
  if (sw == 1) f = kNone;
  if (f == kLinear && (dw == sw || dw * 3 == sw)) f = kNone;
  return f;
}

struct Slope {
  int x = 0, y = 0, dx = 0, dy = 0;
};

// Start and step along one axis when that axis is filtered: centred sampling
// when shrinking, edge-to-edge when enlarging.
void FilteredAxis(int src, int dst, int* start, int* step) {
  if (dst <= src) {
    *step = FixedDiv(src, dst);
    *start = (*step >> 1) - 32768;
  } else if (dst > 1) {
    *step = FixedDiv1(src, dst);
    *start = 0;
  }
}

// ScaleSlope (positive widths only: the reference never mirrors through the scaler).
Slope ComputeSlope(int sw, int sh, int dw, int dh, Filter f) {
  if (dw == 1 && sw >= 32768) dw = sw;
  if (dh == 1 && sh >= 32768) dh = sh;
  Slope s;
  if (f == kBilinear) {
    FilteredAxis(sw, dw, &s.x, &s.dx);
    FilteredAxis(sh, dh, &s.y, &s.dy);
  } else if (f == kLinear) {
    FilteredAxis(sw, dw, &s.x, &s.dx);
    s.dy = FixedDiv(sh, dh);
    s.y = s.dy >> 1;
  } else {
    s.dx = FixedDiv(sw, dw);
    s.dy = FixedDiv(sh, dh);
    s.x = s.dx >> 1;
    s.y = s.dy >> 1;
  }
  return s;
}

ScaleSpec Make(int flavor, int mode, const Slope& s, int linear, int max_y) {
  ScaleSpec out;
  out.flavor = flavor;
  out.mode = mode;
  out.linear = linear;
  out.x0 = s.x;
  out.y0 = s.y;
  out.dx = s.dx;
  out.dy = s.dy;
  out.max_y = max_y;
  return out;
}

// ScalePlaneVertical clamps one 1/65536th short of the last row.
int VerticalMaxY(int sh) { return sh > 1 ? ((sh - 1) << 16) - 1 : 0; }

}  // namespace

int FixedDiv(int num, int div) { return static_cast<int>((static_cast<int64_t>(num) << 16) / div); }
int FixedDiv1(int num, int div) {
  return static_cast<int>(((static_cast<int64_t>(num) << 16) - 0x00010001) / (div - 1));
}

namespace {
// Shared by ScaleARGB (4 bytes per pixel) and ScaleUV (2 bytes per pixel).
ScaleSpec DispatchPacked(int sw, int sh, int dw, int dh, bool uv) {
  const int kF = BANDCU_FLAVOR_ARGB;
  Filter f = ReduceBilinear(sw, sh, dw, dh);
  const Slope s = ComputeSlope(sw, sh, dw, dh, f);
  const int max_y = (sh - 1) << 16;
  if (((s.dx | s.dy) & 0xffff) == 0 && s.dx != 0 && s.dy != 0) {
    const bool dx_even = (s.dx & 0x10000) == 0, dy_even = (s.dy & 0x10000) == 0;
    // even integer ratios on both axes: 2x2 box at the slope's origin, stepping whole pixels
    if (dx_even && dy_even) return Make(kF, BANDCU_SCALE_BOX2, s, 0, max_y);
    // odd integer ratios on both axes: the centre sample is exact, no filtering
    if (!dx_even && !dy_even) f = kNone;
  }
  // unit horizontal step starting on a pixel: rows are blended whole
  if (s.dx == 0x10000 && (s.x & 0xffff) == 0)
    return Make(kF, BANDCU_SCALE_VERTICAL, s, f == kNone ? 1 : 0, VerticalMaxY(sh));
  if (uv) {  // scale_uv.cc only: exact 2x up-samplers
    if ((dw + 1) / 2 == sw && f == kLinear) {
      Slope u;
      u.dy = dh > 1 ? FixedDiv(sh - 1, dh - 1) : 0;
      return Make(kF, BANDCU_SCALE_UP2_LINEAR, u, 1, max_y);
    }
    if ((dh + 1) / 2 == sh && (dw + 1) / 2 == sw && f == kBilinear) {
      Slope u;
      return Make(kF, BANDCU_SCALE_UP2_BILINEAR, u, 0, max_y);
    }
  }
  if (f != kNone && s.dy < 65536) return Make(kF, BANDCU_SCALE_BILINEAR_UP, s, f == kLinear ? 1 : 0, max_y);
  if (f != kNone) return Make(kF, BANDCU_SCALE_BILINEAR_DOWN, s, f == kLinear ? 1 : 0, max_y);
  return Make(kF, BANDCU_SCALE_POINT, s, 1, max_y);
}
}  // namespace

ScaleSpec DispatchArgbScale(int sw, int sh, int dw, int dh) { return DispatchPacked(sw, sh, dw, dh, false); }
ScaleSpec DispatchUvScale(int sw, int sh, int dw, int dh) { return DispatchPacked(sw, sh, dw, dh, true); }

ScaleSpec DispatchPlaneScale(int sw, int sh, int dw, int dh) {
  const int kF = BANDCU_FLAVOR_PLANE;
  const Filter f = ReduceBilinear(sw, sh, dw, dh);
  const int max_y = (sh - 1) << 16;
  if (dw == sw && dh == sh) {
    Slope s;
    s.dx = s.dy = 0x10000;
    return Make(kF, BANDCU_SCALE_COPY, s, 1, max_y);
  }
  if (dw == sw) {  // rows only
    Slope s;
    s.dx = 0x10000;
    if (dh <= sh) {
      s.dy = FixedDiv(sh, dh);
      s.y = (s.dy >> 1) - 32768;
    } else if (sh > 1 && dh > 1) {
      s.dy = FixedDiv1(sh, dh);
    }
    return Make(kF, BANDCU_SCALE_VERTICAL, s, f == kNone ? 1 : 0, VerticalMaxY(sh));
  }
  if (dw <= sw && dh <= sh && f == kBilinear) {
    Slope s;
    if (4 * dw == 3 * sw && 4 * dh == 3 * sh) return Make(kF, BANDCU_SCALE_DOWN34, s, 0, max_y);
    if (2 * dw == sw && 2 * dh == sh) {
      s.dx = s.dy = 2 << 16;
      return Make(kF, BANDCU_SCALE_BOX2, s, 0, max_y);
    }
    if (8 * dw == 3 * sw && 8 * dh == 3 * sh) return Make(kF, BANDCU_SCALE_DOWN38, s, 0, max_y);
  }
  if ((dw + 1) / 2 == sw && f == kLinear) {  // exact 2x wider, rows copied
    Slope s;
    s.dy = dh > 1 ? FixedDiv(sh - 1, dh - 1) : 0;
    return Make(kF, BANDCU_SCALE_UP2_LINEAR, s, 1, max_y);
  }
  if ((dh + 1) / 2 == sh && (dw + 1) / 2 == sw && f == kBilinear) {
    Slope s;
    return Make(kF, BANDCU_SCALE_UP2_BILINEAR, s, 0, max_y);
  }
  const Slope s = ComputeSlope(sw, sh, dw, dh, f);
  if (f != kNone && dh > sh) return Make(kF, BANDCU_SCALE_BILINEAR_UP, s, f == kLinear ? 1 : 0, max_y);
  if (f != kNone) return Make(kF, BANDCU_SCALE_BILINEAR_DOWN, s, f == kLinear ? 1 : 0, max_y);
  return Make(kF, BANDCU_SCALE_POINT, s, 1, max_y);
}

}  // namespace cuda
}  // namespace band
