// band/buffer/fused_plan.cc -- see fused_plan.h.
#include "band/buffer/fused_plan.h"

#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "absl/strings/str_format.h"
#include "band/buffer/operator.h"
#include "band/logger.h"

namespace band {
namespace cuda {
namespace {

constexpr BandCuAffine kIdentity = {0, 1, 0, 0, 0, 1};
// host-only marker in View::cvt: the packed source becomes a 4:2:0 image (own kernel, not a per-fetch conversion)
constexpr int kCvtRgbToYuv = 1000;

bool IsIdentity(const BandCuAffine& m) {
  return m.r0 == 0 && m.rr == 1 && m.rc == 0 && m.c0 == 0 && m.cr == 0 && m.cc == 1;
}

// outer(inner(r, c))
BandCuAffine Compose(const BandCuAffine& o, const BandCuAffine& i) {
  BandCuAffine m;
  m.r0 = o.r0 + o.rr * i.r0 + o.rc * i.c0;
  m.rr = o.rr * i.rr + o.rc * i.cr;
  m.rc = o.rr * i.rc + o.rc * i.cc;
  m.c0 = o.c0 + o.cr * i.r0 + o.cc * i.c0;
  m.cr = o.cr * i.rr + o.cc * i.cr;
  m.cc = o.cr * i.rc + o.cc * i.cc;
  return m;
}

absl::Status CudaError(const char* what) {
  return absl::InternalError(absl::StrFormat("%s: %s", what, bandcu_last_error_string()));
}

int Channels(BufferFormat f) { return static_cast<int>(Buffer::GetNumPixelElements(f)); }

// The contiguous single-allocation layout Buffer::CreateFromRaw describes.
ImageLayout ContiguousLayout(BufferFormat format, int w, int h, DataType dtype) {
  ImageLayout l;
  l.format = format;
  l.dtype = dtype;
  l.width = w;
  l.height = h;
  if (Buffer::IsYUV(format)) {
    const size_t uvw = (static_cast<size_t>(w) + 1) / 2;
    const bool nv = format == BufferFormat::kNV12 || format == BufferFormat::kNV21;
    l.num_planes = 3;
    l.row_stride[0] = w;
    l.pixel_stride[0] = 1;
    l.row_stride[1] = l.row_stride[2] = nv ? uvw * 2 : uvw;
    l.pixel_stride[1] = l.pixel_stride[2] = nv ? 2 : 1;
  } else {
    const size_t ch = Channels(format);
    l.num_planes = 1;
    l.row_stride[0] = static_cast<size_t>(w) * ch * GetDataTypeBytes(dtype);
    l.pixel_stride[0] = ch * GetDataTypeBytes(dtype);
  }
  return l;
}

size_t ContiguousBytes(const ImageLayout& l) {
  if (Buffer::IsYUV(l.format))
    return Buffer::GetBufferByteSize({static_cast<size_t>(l.width), static_cast<size_t>(l.height)}, l.format);
  return l.row_stride[0] * l.height;
}

// Plane pointers of a contiguous frame at `base` (same rules as Buffer::CreateFromRaw).
SlotPointers ContiguousPointers(unsigned char* base, const ImageLayout& l) {
  SlotPointers p;
  p.y = base;
  if (!Buffer::IsYUV(l.format)) return p;
  const size_t wh = static_cast<size_t>(l.width) * l.height;
  const size_t uv = ((static_cast<size_t>(l.width) + 1) / 2) * ((static_cast<size_t>(l.height) + 1) / 2);
  unsigned char* c0 = base + wh;
  switch (l.format) {
    case BufferFormat::kNV21: p.v = c0; p.u = c0 + 1; break;
    case BufferFormat::kNV12: p.u = c0; p.v = c0 + 1; break;
    case BufferFormat::kYV21: p.u = c0; p.v = c0 + uv; break;
    default: p.v = c0; p.u = c0 + uv; break;  // kYV12
  }
  return p;
}

// Plane order of a Buffer (libyuv_image_operator.cc:51-88 GetYuvDataFromBuffer):
// NV21/YV12 store {Y,V,U}; NV12/YV21 store {Y,U,V}.
SlotPointers BufferPointers(const Buffer& b) {
  SlotPointers p;
  p.y = const_cast<unsigned char*>(b[0].data);
  if (Buffer::IsYUV(b.GetBufferFormat()) && b.GetNumPlanes() == 3) {
    const bool v_first = b.GetBufferFormat() == BufferFormat::kNV21 || b.GetBufferFormat() == BufferFormat::kYV12;
    p.v = const_cast<unsigned char*>(b[v_first ? 1 : 2].data);
    p.u = const_cast<unsigned char*>(b[v_first ? 2 : 1].data);
  }
  return p;
}

int SrcKind(BufferFormat f) {
  switch (f) {
    case BufferFormat::kGrayScale: return BANDCU_SRC_GRAY;
    case BufferFormat::kRGB: return BANDCU_SRC_RGB;
    case BufferFormat::kRGBA: return BANDCU_SRC_RGBA;
    default: return BANDCU_SRC_YUV420;
  }
}

}  // namespace

ImageLayout ImageLayout::Of(const Buffer& b) {
  ImageLayout l;
  l.format = b.GetBufferFormat();
  l.dtype = b.GetDataType();
  if (b.GetDimension().size() >= 2) {
    l.width = static_cast<int>(b.GetDimension()[0]);
    l.height = static_cast<int>(b.GetDimension()[1]);
  }
  l.orientation = static_cast<int>(b.GetOrientation());
  l.num_planes = static_cast<int>(std::min<size_t>(b.GetNumPlanes(), 3));
  for (int k = 0; k < l.num_planes; ++k) {
    l.row_stride[k] = b[k].row_stride_bytes;
    l.pixel_stride[k] = b[k].pixel_stride_bytes;
  }
  return l;
}

bool ImageLayout::operator==(const ImageLayout& o) const {
  if (format != o.format || dtype != o.dtype || width != o.width || height != o.height || num_planes != o.num_planes ||
      orientation != o.orientation)
    return false;
  for (int k = 0; k < num_planes; ++k)
    if (row_stride[k] != o.row_stride[k] || pixel_stride[k] != o.pixel_stride[k]) return false;
  return true;
}

// What the next operator of the chain sees, and how it derives from `src`.
struct ChainLowering::View {
  int slot = kSlotInput;
  ImageLayout src;
  BufferFormat format = BufferFormat::kRaw;
  DataType dtype = DataType::kUInt8;
  int w = 0, h = 0;          // dims the next operator sees
  int pre_w = 0, pre_h = 0;  // dims of the image the Resize stage samples
  BandCuAffine luma = kIdentity, pre = kIdentity, chroma = kIdentity;
  int cvt = BANDCU_CVT_NONE;
  bool scaled = false;
  ScaleSpec scale;
  bool normalized = false;
  float mean = 0.f, std = 1.f;
  // Normalize applied to already materialised memory (any dtype, linear index,
  // common_operator.cc:24-40) instead of the fused u8 store stage
  bool elementwise = false;
  DataType ew_in_dtype = DataType::kUInt8;
  // the view IS the image in memory. A Crop whose box starts at (0, 0) leaves identity maps but a smaller image: its rows
  // are not contiguous in the source, so linear-memory stages (element-wise Normalize, plane scalers) must not take it
  // for materialised memory (found by scripts/fuzz_chains.py, seed 19022).
  bool Untouched() const {
    return IsIdentity(luma) && IsIdentity(pre) && IsIdentity(chroma) && cvt == BANDCU_CVT_NONE && !scaled &&
           !normalized && w == src.width && h == src.height;
  }
  bool IsYuv() const { return Buffer::IsYUV(format); }
};

// ---------------------------------------------------------------------------
// ChainLowering
// ---------------------------------------------------------------------------
ChainLowering::ChainLowering(const Buffer& input) : plan_(new FusedPlan()), view_(new View()) {
  plan_->input_ = ImageLayout::Of(input);
  // diagnostic switches are sampled here, once per plan: nothing on the launch path reads the environment
  plan_->zero_copy_allowed_ = getenv("BAND_NO_ZERO_COPY") == nullptr;
  if (const char* c = getenv("BAND_CHUNK_FRAMES"))
    if (atoi(c) > 0) plan_->chunk_frames_ = atoi(c);
  plan_->register_host_ = HostRegistry::Get().Enabled();
  if (plan_->register_host_) HostRegistry::Get().AddUser();
  ResetView(kSlotInput, plan_->input_);
}

ChainLowering::~ChainLowering() = default;

void ChainLowering::ResetView(int slot, const ImageLayout& layout) {
  *view_ = View();
  view_->slot = slot;
  view_->src = layout;
  view_->format = layout.format;
  view_->dtype = layout.dtype;
  view_->w = view_->pre_w = layout.width;
  view_->h = view_->pre_h = layout.height;
}

// Close the current segment into device scratch; the next stage starts from it.
absl::Status ChainLowering::Flush() {
  const DataType dt = view_->dtype;
  ImageLayout l = ContiguousLayout(view_->format, view_->w, view_->h, dt);
  const int slot = static_cast<int>(plan_->scratch_.size());
  plan_->scratch_.push_back(l);
  RETURN_IF_ERROR(EmitInto(slot, l));
  ResetView(slot, l);
  return absl::OkStatus();
}

absl::Status ChainLowering::EmitInto(int dst_slot, const ImageLayout& dst) {
  View& v = *view_;
  // YV12 <-> YV21: both sides' U and V planes are resolved by name (BufferPointers), like the reference's *Yv routines
  auto planar = [](BufferFormat f) { return f == BufferFormat::kYV12 || f == BufferFormat::kYV21; };
  const bool same_layout = dst.format == v.format || (planar(dst.format) && planar(v.format));
  if (!same_layout || dst.width != v.w || dst.height != v.h)
    return absl::InvalidArgumentError(absl::StrFormat(
        "output buffer is %s %dx%d but the chain produces %s %dx%d", ToString(dst.format), dst.width, dst.height,
        ToString(v.format), v.w, v.h));
  if (dst.dtype != v.dtype)
    return absl::InvalidArgumentError(absl::StrFormat("output buffer holds %s but the chain produces %s",
                                                      ToString(dst.dtype), ToString(v.dtype)));
  Segment seg;
  seg.src_slot = v.slot;
  seg.dst_slot = dst_slot;
  if (v.elementwise) {  // Normalize over linear memory: W*H*channels elements, strides ignored
    seg.kind = Segment::Kind::kElementwise;
    seg.in_dtype = static_cast<int>(v.ew_in_dtype);
    seg.out_dtype = static_cast<int>(v.dtype);
    seg.count = static_cast<long long>(v.w) * v.h * Channels(v.format);
    seg.mean = v.mean;
    seg.std = v.std;
    plan_->segments_.push_back(seg);
    return absl::OkStatus();
  }
  if (v.cvt == kCvtRgbToYuv) {  // packed RGB / RGBA source -> 4:2:0 image
    if (dst.num_planes != 3) return absl::InvalidArgumentError("YUV output buffer needs three planes");
    if (v.src.dtype != DataType::kUInt8)
      return absl::UnimplementedError("image stages on non-uint8 pixels are not supported");
    seg.kind = Segment::Kind::kRgbToYuv;
    BandCuRgbToYuvDesc& d = seg.rgb2yuv;
    d.src_channels = Channels(v.src.format);
    d.src_stride = static_cast<int>(v.src.row_stride[0]);
    d.src_map = v.luma;
    d.width = v.w;
    d.height = v.h;
    d.dst_stride_y = static_cast<int>(dst.row_stride[0]);
    d.dst_stride_uv = static_cast<int>(dst.row_stride[1]);
    d.dst_pixel_stride_uv = static_cast<int>(dst.pixel_stride[1]);
    plan_->segments_.push_back(seg);
    return absl::OkStatus();
  }
  if (v.IsYuv() && v.scaled) {  // Resize in the YUV domain: every plane through its own libyuv scaler
    if (dst.num_planes != 3) return absl::InvalidArgumentError("YUV output buffer needs three planes");
    const bool nv = v.format == BufferFormat::kNV12 || v.format == BufferFormat::kNV21;
    const size_t want_ps = nv ? 2 : 1;
    if (v.src.pixel_stride[1] != want_ps || dst.pixel_stride[1] != want_ps || v.src.pixel_stride[2] != want_ps ||
        dst.pixel_stride[2] != want_ps)
      return absl::UnimplementedError("Resize of a 4:2:0 buffer with non-native chroma pixel strides is not supported");
    seg.kind = Segment::Kind::kYuvScale;
    seg.interleaved_uv = nv;
    auto plane_desc = [&](int src_kind, int channels, const ScaleSpec& spec, int sw, int sh, int dw, int dh,
                          size_t src_stride, size_t dst_stride) {
      BandCuPlanDesc d{};
      d.src_kind = src_kind;
      d.stride0 = static_cast<int>(src_stride);
      d.vw = sw;
      d.vh = sh;
      d.luma_map = d.pre_map = d.chroma_map = kIdentity;
      d.cvt = BANDCU_CVT_NONE;
      d.flavor = spec.flavor;
      d.scale_mode = spec.mode;
      d.linear = spec.linear;
      d.x0 = spec.x0; d.y0 = spec.y0; d.dx = spec.dx; d.dy = spec.dy; d.max_y = spec.max_y;
      d.dst_width = dw;
      d.dst_height = dh;
      d.channels = channels;
      d.out_dtype = BANDCU_DT_U8;
      d.dst_row_stride = static_cast<long long>(dst_stride);
      return d;
    };
    const int sw = v.pre_w, sh = v.pre_h, dw = v.w, dh = v.h;
    const int shw = (sw + 1) / 2, shh = (sh + 1) / 2, dhw = (dw + 1) / 2, dhh = (dh + 1) / 2;
    BandCuPlanDesc dy = plane_desc(BANDCU_SRC_GRAY, 1, DispatchPlaneScale(sw, sh, dw, dh), sw, sh, dw, dh,
                                   v.src.row_stride[0], dst.row_stride[0]);
    if (bandcu_plan_create(&dy, &seg.plane_plan[0]) != BANDCU_OK) return CudaError("bandcu_plan_create");
    if (nv) {
      BandCuPlanDesc duv = plane_desc(BANDCU_SRC_UV, 2, DispatchUvScale(shw, shh, dhw, dhh), shw, shh, dhw, dhh,
                                      v.src.row_stride[1], dst.row_stride[1]);
      if (bandcu_plan_create(&duv, &seg.plane_plan[1]) != BANDCU_OK) return CudaError("bandcu_plan_create");
    } else {
      for (int k = 1; k <= 2; ++k) {
        BandCuPlanDesc dc = plane_desc(BANDCU_SRC_GRAY, 1, DispatchPlaneScale(shw, shh, dhw, dhh), shw, shh, dhw, dhh,
                                       v.src.row_stride[k], dst.row_stride[k]);
        if (bandcu_plan_create(&dc, &seg.plane_plan[k]) != BANDCU_OK) return CudaError("bandcu_plan_create");
      }
    }
    plan_->segments_.push_back(seg);
    return absl::OkStatus();
  }
  if (v.IsYuv()) {  // 4:2:0 in, 4:2:0 out: pure permutation of luma and chroma samples
    if (dst.num_planes != 3) return absl::InvalidArgumentError("YUV output buffer needs three planes");
    seg.kind = Segment::Kind::kYuvMap;
    BandCuYuvMapDesc& d = seg.yuv;
    d.src_stride_y = static_cast<int>(v.src.row_stride[0]);
    d.src_stride_uv = static_cast<int>(v.src.row_stride[1]);
    d.src_pixel_stride_uv = static_cast<int>(v.src.pixel_stride[1]);
    d.dst_stride_y = static_cast<int>(dst.row_stride[0]);
    d.dst_stride_uv = static_cast<int>(dst.row_stride[1]);
    d.dst_pixel_stride_uv = static_cast<int>(dst.pixel_stride[1]);
    d.dst_width = v.w;
    d.dst_height = v.h;
    d.chroma_width_override = 0;
    d.luma_map = v.luma;
    d.chroma_map = v.chroma;
    plan_->segments_.push_back(seg);
    return absl::OkStatus();
  }
  if (v.src.dtype != DataType::kUInt8)
    return absl::UnimplementedError("image stages on non-uint8 pixels are not supported");
  seg.kind = Segment::Kind::kPixel;
  BandCuPlanDesc& d = seg.pixel;
  d.src_kind = SrcKind(v.src.format);
  d.stride0 = static_cast<int>(v.src.row_stride[0]);
  d.stride_uv = static_cast<int>(v.src.row_stride[1]);
  d.pixel_stride_uv = static_cast<int>(v.src.pixel_stride[1]);
  d.src_width = v.src.width;
  d.src_height = v.src.height;
  d.vw = v.pre_w;
  d.vh = v.pre_h;
  d.luma_map = v.luma;
  d.pre_map = v.pre;
  d.chroma_map = v.chroma;
  d.cvt = v.cvt;
  if (v.scaled) {
    d.flavor = v.scale.flavor;
    d.scale_mode = v.scale.mode;
    d.linear = v.scale.linear;
    d.x0 = v.scale.x0;
    d.y0 = v.scale.y0;
    d.dx = v.scale.dx;
    d.dy = v.scale.dy;
    d.max_y = v.scale.max_y;
  } else {
    d.flavor = v.format == BufferFormat::kGrayScale ? BANDCU_FLAVOR_PLANE : BANDCU_FLAVOR_ARGB;
    d.scale_mode = BANDCU_SCALE_COPY;
    d.dx = d.dy = 0x10000;
  }
  d.dst_width = v.w;
  d.dst_height = v.h;
  d.channels = Channels(v.format);
  d.out_dtype = static_cast<int>(v.dtype);  // BANDCU_DT_* follow band::DataType
  d.normalize = v.normalized ? 1 : 0;
  d.mean = v.mean;
  d.std_ = v.std;
  // tensors are contiguous NHWC (Buffer::CreateFromTensor's row stride ignores the
  // element size, buffer.cc:137-141); uint8 images may carry row padding
  d.dst_row_stride = (dst.dtype == DataType::kUInt8 && dst.row_stride[0] >= static_cast<size_t>(v.w) * d.channels)
                         ? static_cast<long long>(dst.row_stride[0])
                         : static_cast<long long>(v.w) * d.channels * static_cast<long long>(GetDataTypeBytes(dst.dtype));
  if (bandcu_plan_create(&d, &seg.plan) != BANDCU_OK) return CudaError("bandcu_plan_create");
  plan_->segments_.push_back(seg);
  return absl::OkStatus();
}

// new view pixel (r,c) := old view pixel luma(r,c); 4:2:0 views move their chroma planes by `chroma`
absl::Status ChainLowering::ApplyIndexMap(const BandCuAffine& luma, const BandCuAffine& chroma, int new_w,
                                          int new_h) {
  if (view_->dtype != DataType::kUInt8)
    return absl::UnimplementedError("Crop/Rotate/Flip on non-uint8 pixels are not supported");
  if (view_->scaled || view_->normalized || view_->cvt == kCvtRgbToYuv) RETURN_IF_ERROR(Flush());
  View& v = *view_;
  terminal_noop_ = false;
  v.luma = Compose(v.luma, luma);
  if (v.IsYuv()) {
    v.chroma = Compose(v.chroma, chroma);
  } else if (Buffer::IsYUV(v.src.format)) {
    v.pre = Compose(v.pre, luma);  // position inside the image the colour conversion ran on
  }
  v.w = v.pre_w = new_w;
  v.h = v.pre_h = new_h;
  return absl::OkStatus();
}

// Crop: inclusive box; chroma planes start at (x0/2, y0/2)
// (libyuv_image_operator.cc:910-1037 CropPlane/CropNv/CropYv).
absl::Status ChainLowering::Crop(int x0, int y0, int x1, int y1) {
  const BandCuAffine luma = {y0, 1, 0, x0, 0, 1};
  const BandCuAffine chroma = {y0 / 2, 1, 0, x0 / 2, 0, 1};
  return ApplyIndexMap(luma, chroma, x1 - x0 + 1, y1 - y0 + 1);
}

// Rotate: COUNTER-clockwise (libyuv_image_operator.cc:697-708 maps 90 -> kRotate270);
// chroma planes rotate on their own with half dims ((W+1)/2,(H+1)/2) (:786-887).
absl::Status ChainLowering::Rotate(int angle) {
  const int W = view_->w, H = view_->h;
  const int hw = (W + 1) / 2, hh = (H + 1) / 2;
  switch (angle) {
    case 0:
      return absl::OkStatus();
    case 90:  // dst[r][c] = src[c][W-1-r], dst is H wide and W tall
      return ApplyIndexMap({0, 0, 1, W - 1, -1, 0}, {0, 0, 1, hw - 1, -1, 0}, H, W);
    case 180:
      return ApplyIndexMap({H - 1, -1, 0, W - 1, 0, -1}, {hh - 1, -1, 0, hw - 1, 0, -1}, W, H);
    case 270:  // dst[r][c] = src[H-1-c][r]
      return ApplyIndexMap({H - 1, 0, -1, 0, 1, 0}, {hh - 1, 0, -1, 0, 1, 0}, H, W);
    default:
      return absl::InvalidArgumentError(absl::StrFormat("Rotate: angle %d is not a multiple of 90", angle));
  }
}

// Flip (libyuv_image_operator.cc:1108-1363): horizontal mirrors columns, vertical
// reverses rows; both = horizontal then vertical.
absl::Status ChainLowering::Flip(bool horizontal, bool vertical) {
  const int W = view_->w, H = view_->h;
  const int hw = (W + 1) / 2, hh = (H + 1) / 2;
  BandCuAffine luma = kIdentity, chroma = kIdentity;
  if (horizontal) {
    luma.c0 = W - 1; luma.cc = -1;
    chroma.c0 = hw - 1; chroma.cc = -1;
  }
  if (vertical) {
    luma.r0 = H - 1; luma.rr = -1;
    chroma.r0 = hh - 1; chroma.rr = -1;
  }
  return ApplyIndexMap(luma, chroma, W, H);
}

// ColorSpaceConvert (libyuv_image_operator.cc:1709-1727 and the ConvertFrom* it
// dispatches to): a per-fetch conversion of the segment's source pixels.
absl::Status ChainLowering::ConvertColor(BufferFormat to) {
  if (view_->dtype != DataType::kUInt8)
    return absl::UnimplementedError("ColorSpaceConvert on non-uint8 pixels is not supported");
  if (view_->scaled || view_->normalized || view_->cvt != BANDCU_CVT_NONE) RETURN_IF_ERROR(Flush());
  View& v = *view_;
  const BufferFormat from = v.format;
  terminal_noop_ = false;
  if (Buffer::IsYUV(from)) {
    if (Buffer::IsYUV(to)) {  // repack (NV12ToI420 / I420ToNV12 / SwapUVPlane ...): same samples, new layout
      v.format = to;
      return absl::OkStatus();
    }
    switch (to) {
      case BufferFormat::kRGB: v.cvt = BANDCU_CVT_YUV2RGB; break;
      case BufferFormat::kRGBA: v.cvt = BANDCU_CVT_YUV2RGBA; break;
      case BufferFormat::kGrayScale: v.cvt = BANDCU_CVT_YUV2GRAY; break;
      default:
        return absl::InternalError(absl::StrFormat("Format %s is not supported.", ToString(to)));
    }
    v.format = to;
    v.pre = kIdentity;  // the conversion runs on the image as seen here
    return absl::OkStatus();
  }
  if (from == BufferFormat::kRGB && to == BufferFormat::kGrayScale) {
    v.cvt = BANDCU_CVT_RGB2GRAY;
  } else if (from == BufferFormat::kRGBA && to == BufferFormat::kRGB) {
    v.cvt = BANDCU_CVT_RGBA2RGB;
  } else if (from == BufferFormat::kRGBA && to == BufferFormat::kGrayScale) {
    v.cvt = BANDCU_CVT_RGBA2GRAY;
  } else if ((from == BufferFormat::kRGB || from == BufferFormat::kRGBA) && Buffer::IsYUV(to)) {
    // RAWToI420 / ABGRTo{NV12,NV21,I420}: every target layout receives the same Y/U/V samples (the reference's
    // RGB -> NV route goes through a temporary I420 image and a sample-preserving repack, :536-583)
    v.cvt = kCvtRgbToYuv;
  } else {
    return absl::InternalError(absl::StrFormat("Format %s -> %s is not supported.", ToString(from), ToString(to)));
  }
  v.format = to;
  return absl::OkStatus();
}

// Resize (libyuv_image_operator.cc:1627-1646): always kFilterBilinear; the code
// path libyuv would take for these dims is chosen by scale_dispatch.cc.
absl::Status ChainLowering::Scale(int dst_width, int dst_height) {
  if (view_->dtype != DataType::kUInt8)
    return absl::UnimplementedError("Resize on non-uint8 pixels is not supported");
  if (dst_width <= 0 || dst_height <= 0) return absl::InvalidArgumentError("Resize: non-positive output dimension");
  if (view_->IsYuv()) {
    // ResizeYv / ResizeNv (libyuv_image_operator.cc:376-435): the planes of a MATERIALISED 4:2:0 image go through
    // ScalePlane (Y, planar U and V) or UVScale (interleaved chroma); pending index maps / repacks are closed first
    if (!view_->Untouched() || view_->format != view_->src.format) RETURN_IF_ERROR(Flush());
    View& y = *view_;
    terminal_noop_ = false;
    y.scaled = true;
    y.pre_w = y.w;
    y.pre_h = y.h;
    y.w = dst_width;
    y.h = dst_height;
    return absl::OkStatus();
  }
  if (view_->scaled || view_->normalized) RETURN_IF_ERROR(Flush());
  {
    const View& cur = *view_;
    // RGB / RGBA -> gray in front of an up-scale (RAWToJ400 / ARGBToJ400 then ScalePlane): fused, every source pixel would be
    // converted once per fetch, (factor)^2 times over; the gray image is a quarter of the source and is materialised first.
    // (Outputs below 256 KB stay fused: there the second launch costs more than the conversions it saves -- measured on
    // RGBA 64^2 -> gray 256^2.)
    if ((cur.cvt == BANDCU_CVT_RGB2GRAY || cur.cvt == BANDCU_CVT_RGBA2GRAY) &&
        static_cast<long long>(dst_width) * dst_height >= 256 * 1024 &&
        DispatchPlaneScale(cur.w, cur.h, dst_width, dst_height).mode == BANDCU_SCALE_BILINEAR_UP)
      RETURN_IF_ERROR(Flush());
  }
  {
    // An UP-scale of a 4:2:0 source into RGB / RGBA (ConvertFromNv* then ResizeRgb's ARGBScale up path,
    // libyuv_image_operator.cc:92-300,1146-1187): fused, every output pixel would convert its four source pixels again --
    // (scale factor)^2 conversions per source pixel on the pixel-per-thread kernel. Materialising the converted image
    // once (the same-size conversion kernel) and up-scaling it with the row-streaming packed kernels is what the
    // reference does stage by stage, moves W*H*3 extra bytes against an output (factor)^2 times that size, and keeps
    // both launches on vectorised kernels. Values are identical either way.
    const View& cur = *view_;
    const bool yuv_to_packed = Buffer::IsYUV(cur.src.format) && (cur.cvt == BANDCU_CVT_YUV2RGB || cur.cvt == BANDCU_CVT_YUV2RGBA);
    if (yuv_to_packed && DispatchArgbScale(cur.w, cur.h, dst_width, dst_height).mode == BANDCU_SCALE_BILINEAR_UP) {
      // The scratch image is RGBA even when the chain asked for RGB: the up-scaler then fetches a source pixel with one
      // 32-bit load instead of three byte loads and drops the alpha byte per fetch (ARGBToRGB24 semantics) -- the column
      // filter of a small-ratio up-scale is bound by exactly those fetches. Same R, G, B values either way.
      const bool widen = cur.cvt == BANDCU_CVT_YUV2RGB;
      if (widen) {
        view_->cvt = BANDCU_CVT_YUV2RGBA;
        view_->format = BufferFormat::kRGBA;
      }
      RETURN_IF_ERROR(Flush());
      if (widen) {
        view_->cvt = BANDCU_CVT_RGBA2RGB;
        view_->format = BufferFormat::kRGB;
      }
    }
  }
  View& v = *view_;
  terminal_noop_ = false;
  v.scale = v.format == BufferFormat::kGrayScale ? DispatchPlaneScale(v.w, v.h, dst_width, dst_height)
                                                 : DispatchArgbScale(v.w, v.h, dst_width, dst_height);
  v.scaled = true;
  v.pre_w = v.w;
  v.pre_h = v.h;
  v.w = dst_width;
  v.h = dst_height;
  return absl::OkStatus();
}

// Normalize / DataTypeConvert (common_operator.cc:24-175):
//   out[i] = static_cast<Out>(in[i] - mean) / std   over W*H*channels elements.
// On uint8 pixels produced by this segment it is the store stage of the fused
// kernel; on anything already in memory it is an element-wise launch.
absl::Status ChainLowering::Normalize(float mean, float std, DataType out_dtype, bool inplace) {
  if (view_->IsYuv()) return absl::InvalidArgumentError("Normalize: YUV buffers are not supported");
  switch (out_dtype) {
    case DataType::kUInt8: case DataType::kInt8: case DataType::kInt16: case DataType::kInt32: case DataType::kFloat32:
      break;
    default:
      // the reference logs "unsupported data type" and writes nothing (common_operator.cc:60-64)
      return absl::InvalidArgumentError(absl::StrFormat("Normalize: unsupported output data type %s", ToString(out_dtype)));
  }
  if (inplace) {
    // in == out buffer, same element type; the chain output is never written
    if (!view_->Untouched()) RETURN_IF_ERROR(Flush());
    View& v = *view_;
    Segment seg;
    seg.kind = Segment::Kind::kElementwise;
    seg.src_slot = seg.dst_slot = v.slot;
    seg.in_dtype = seg.out_dtype = static_cast<int>(v.dtype);
    seg.count = static_cast<long long>(v.w) * v.h * Channels(v.format);
    seg.mean = mean;
    seg.std = std;
    if (v.slot == kSlotInput) plan_->writes_input_ = true;
    plan_->segments_.push_back(seg);
    terminal_noop_ = true;
    return absl::OkStatus();
  }
  terminal_noop_ = false;
  const bool in_memory = view_->Untouched();
  if (!in_memory && view_->dtype == DataType::kUInt8 && !view_->normalized) {
    View& v = *view_;
    v.normalized = true;
    v.mean = mean;
    v.std = std;
    v.dtype = out_dtype;
    return absl::OkStatus();
  }
  if (!in_memory) RETURN_IF_ERROR(Flush());
  View& v = *view_;
  v.elementwise = true;
  v.normalized = true;  // any later stage starts a new segment
  v.ew_in_dtype = v.dtype;
  v.dtype = out_dtype;
  v.mean = mean;
  v.std = std;
  return absl::OkStatus();
}

absl::Status ChainLowering::PassThrough() {
  terminal_noop_ = true;
  return absl::OkStatus();
}

absl::Status ChainLowering::LowerNested(const IBufferOperator& op, const Buffer& input) {
  return op.Lower(*this, input);
}

absl::StatusOr<std::unique_ptr<FusedPlan>> ChainLowering::Finish(const Buffer* output) {
  if (terminal_noop_) {
    plan_->writes_output_ = false;
  } else {
    if (output == nullptr) return absl::InternalError("Null output buffer");
    plan_->output_ = ImageLayout::Of(*output);
    RETURN_IF_ERROR(EmitInto(kSlotOutput, plan_->output_));
  }
  for (const ImageLayout& l : plan_->scratch_) plan_->scratch_bytes_.push_back((ContiguousBytes(l) + 255) / 256 * 256);
  plan_->scratch_mem_.assign(plan_->scratch_.size(), nullptr);
  return std::move(plan_);
}

// ---------------------------------------------------------------------------
// FusedPlan: staging + launch
// ---------------------------------------------------------------------------
namespace {

constexpr size_t kAlign = 256;
constexpr int kMaxPassFrames = 2048;  // frames per pass: bounds the staging areas
size_t AlignUp(size_t v) { return (v + kAlign - 1) / kAlign * kAlign; }

// Bytes a plane occupies from its first to its last addressed byte.
size_t PlaneExtent(const ImageLayout& l, int k) {
  if (Buffer::IsYUV(l.format)) {
    if (k == 0) return static_cast<size_t>(l.height - 1) * l.row_stride[0] + l.width;
    const size_t cw = (static_cast<size_t>(l.width) + 1) / 2, ch = (static_cast<size_t>(l.height) + 1) / 2;
    return (ch - 1) * l.row_stride[k] + (cw - 1) * l.pixel_stride[k] + 1;
  }
  const size_t ch = Channels(l.format), esz = GetDataTypeBytes(l.dtype);
  if (l.format == BufferFormat::kRaw) return l.row_stride[0];
  if (l.dtype != DataType::kUInt8) return static_cast<size_t>(l.width) * l.height * ch * esz;  // contiguous tensor
  const size_t row = static_cast<size_t>(l.width) * ch;
  const size_t stride = l.row_stride[0] >= row ? l.row_stride[0] : row;
  return static_cast<size_t>(l.height - 1) * stride + row;
}

// Host byte ranges of one frame; planes that overlap or touch are merged (an NV12 frame from CreateFromRaw is
// one range; Android Image.Planes may be three). Planes that merely sit NEAR each other stay separate: the bytes
// between them belong to the caller (and may be unmapped).
struct HostRanges {
  int count = 0;
  const unsigned char* lo[3] = {nullptr, nullptr, nullptr};
  size_t bytes[3] = {0, 0, 0};
  int range_of_plane[3] = {0, 0, 0};
  size_t Total() const {
    size_t t = 0;
    for (int i = 0; i < count; ++i) t += AlignUp(bytes[i]);
    return t;
  }
};

HostRanges RangesOf(const Buffer& b, const ImageLayout& l) {
  HostRanges r;
  int order[3] = {0, 1, 2};
  const int np = l.num_planes;
  for (int i = 0; i < np; ++i)
    for (int j = i + 1; j < np; ++j)
      if (b[order[j]].data < b[order[i]].data) std::swap(order[i], order[j]);
  for (int i = 0; i < np; ++i) {
    const int k = order[i];
    const unsigned char* lo = b[k].data;
    const unsigned char* hi = lo + PlaneExtent(l, k);
    if (r.count > 0) {
      const int c = r.count - 1;
      const unsigned char* cur_hi = r.lo[c] + r.bytes[c];
      if (lo <= cur_hi) {  // overlaps or is exactly adjacent: one copy
        if (hi > cur_hi) r.bytes[c] = static_cast<size_t>(hi - r.lo[c]);
        r.range_of_plane[k] = c;
        continue;
      }
    }
    r.lo[r.count] = lo;
    r.bytes[r.count] = static_cast<size_t>(hi - lo);
    r.range_of_plane[k] = r.count;
    ++r.count;
  }
  return r;
}

// The bytes of a frame an operator chain may WRITE: per plane `rows` rows of `row_bytes` bytes, `stride` apart.
// libyuv writes `width` bytes per row and nothing else (row padding, gaps between planes and the bytes between the
// samples of a pixel-strided plane stay the caller's), so downloads into host images copy exactly these rectangles.
struct PlaneRects {
  int count = 0;
  const unsigned char* lo[3] = {nullptr, nullptr, nullptr};
  size_t row_bytes[3] = {0, 0, 0}, stride[3] = {0, 0, 0}, rows[3] = {0, 0, 0};
  bool exact = true;  // false: a plane whose written bytes are not whole row segments (split pixel-stride-2 chroma)
  // every rectangle is unpadded and they sit back to back: the frame is one contiguous block of written bytes
  bool Dense() const {
    if (!exact) return false;
    for (int i = 0; i < count; ++i) {
      if (rows[i] > 1 && stride[i] != row_bytes[i]) return false;
      if (i > 0 && lo[i] != lo[i - 1] + rows[i - 1] * row_bytes[i - 1]) return false;
    }
    return true;
  }
};

PlaneRects RectsOf(const Buffer& b, const ImageLayout& l) {
  PlaneRects r;
  auto add = [&](const unsigned char* lo, size_t row_bytes, size_t stride, size_t rows) {
    r.lo[r.count] = lo; r.row_bytes[r.count] = row_bytes; r.stride[r.count] = stride; r.rows[r.count] = rows;
    ++r.count;
  };
  if (Buffer::IsYUV(l.format) && l.num_planes == 3) {
    const size_t cw = (static_cast<size_t>(l.width) + 1) / 2, ch = (static_cast<size_t>(l.height) + 1) / 2;
    add(b[0].data, l.width, l.row_stride[0], l.height);
    if (l.pixel_stride[1] == 1 && l.pixel_stride[2] == 1) {
      const int first = b[1].data <= b[2].data ? 1 : 2;
      add(b[first].data, cw, l.row_stride[first], ch);
      add(b[3 - first].data, cw, l.row_stride[3 - first], ch);
    } else if (l.pixel_stride[1] == 2 && l.pixel_stride[2] == 2 && l.row_stride[1] == l.row_stride[2] &&
               (b[1].data + 1 == b[2].data || b[2].data + 1 == b[1].data)) {
      add(std::min(b[1].data, b[2].data), 2 * cw, l.row_stride[1], ch);  // interleaved pairs
    } else {
      r.exact = false;
    }
    return r;
  }
  const size_t ch = Channels(l.format), esz = GetDataTypeBytes(l.dtype);
  if (l.format == BufferFormat::kRaw) {
    add(b[0].data, l.row_stride[0], l.row_stride[0], 1);
  } else if (l.dtype != DataType::kUInt8) {  // contiguous tensor (strides ignored like the reference)
    const size_t total = static_cast<size_t>(l.width) * l.height * ch * esz;
    add(b[0].data, total, total, 1);
  } else {
    const size_t row = static_cast<size_t>(l.width) * ch;
    add(b[0].data, row, l.row_stride[0] >= row ? l.row_stride[0] : row, l.height);
  }
  return r;
}

// Device pointers of a frame whose host ranges were packed at `base`.
SlotPointers StagedPointers(const Buffer& b, const ImageLayout& l, const HostRanges& r, unsigned char* base) {
  unsigned char* range_base[3];
  size_t off = 0;
  for (int i = 0; i < r.count; ++i) {
    range_base[i] = base + off;
    off += AlignUp(r.bytes[i]);
  }
  auto at = [&](int k) { return range_base[r.range_of_plane[k]] + (b[k].data - r.lo[r.range_of_plane[k]]); };
  SlotPointers p;
  p.y = at(0);
  if (Buffer::IsYUV(l.format) && l.num_planes == 3) {
    const bool v_first = l.format == BufferFormat::kNV21 || l.format == BufferFormat::kYV12;
    p.v = at(v_first ? 1 : 2);
    p.u = at(v_first ? 2 : 1);
  }
  return p;
}

}  // namespace

FusedPlan::~FusedPlan() {
  if (stream_) (void)Synchronize();
  ReleaseHeldHost();
  if (register_host_) HostRegistry::Get().RemoveUser();
  for (Segment& s : segments_) {
    if (s.plan) bandcu_plan_destroy(s.plan);
    for (BandCuPlan* p : s.plane_plan)
      if (p) bandcu_plan_destroy(p);
  }
  for (unsigned char* p : scratch_mem_)
    if (p) bandcu_free(p);
  if (in_stage_) bandcu_free(in_stage_);
  if (out_stage_) bandcu_free(out_stage_);
  for (void* st : {stream_, copy_in_, copy_out_})
    if (st) bandcu_stream_destroy(st);
  for (void* ev : events_)
    if (ev) bandcu_event_destroy(ev);
  if (fence_a_) bandcu_event_destroy(fence_a_);
  if (fence_b_) bandcu_event_destroy(fence_b_);
}

void* FusedPlan::ComputeStream() { return EnsureStreams().ok() ? stream_ : nullptr; }

absl::Status FusedPlan::Synchronize() {
  if (!stream_) return absl::OkStatus();
  if (bandcu_stream_sync(stream_) != BANDCU_OK) return CudaError("stream sync");
  if (bandcu_stream_sync(copy_out_) != BANDCU_OK) return CudaError("stream sync");
  in_flight_ = false;
  ReleaseHeldHost();
  return absl::OkStatus();
}

bool FusedPlan::HostFrameIsMapped(const Buffer& frame) {
  if (!register_host_) return frame.IsPinnedHost();
  // Ranges this cache registered are HELD for the duration of the pass (the runtime reports them as pinned too, but an
  // unheld range could be evicted by another processor's registration while this pass's kernel still reads it).
  const HostRanges r = RangesOf(frame, input_);
  auto acquire_all = [&](bool existing_only) {
    const size_t first = held_host_.size();
    for (int i = 0; i < r.count; ++i) {
      if (!HostRegistry::Get().Acquire(r.lo[i], r.bytes[i], &held_host_, existing_only)) {
        std::vector<HostRegistry::Entry*> undo(held_host_.begin() + first, held_host_.end());
        held_host_.resize(first);
        HostRegistry::Get().Release(&undo);
        return false;
      }
    }
    return true;
  };
  if (acquire_all(/*existing_only=*/true)) return true;
  if (frame.IsPinnedHost()) return true;  // page-locked by the caller (BandHostAlloc, cudaHostRegister of their own)
  return acquire_all(/*existing_only=*/false);
}

void FusedPlan::ReleaseHeldHost() {
  if (!held_host_.empty()) HostRegistry::Get().Release(&held_host_);
}

absl::Status FusedPlan::Record(const Buffer* const* inputs, Buffer* const* outputs, int n, void** graph_exec) {
  if (graph_exec == nullptr || n <= 0) return absl::InvalidArgumentError("Record: bad arguments");
  *graph_exec = nullptr;
  // one ordinary pass first: it creates the streams, sizes every plan-owned allocation and proves the batch valid,
  // so that the recorded pass consists of kernel launches only
  RETURN_IF_ERROR(Run(inputs, outputs, n, /*synchronize=*/true));
  if (last_pass_staged_)
    return absl::InvalidArgumentError(
        "RecordBatch: this batch moves through staging copies (pageable or unaligned host memory, or a chain that is not "
        "one row-staged launch over pinned frames); record device-resident frames and tensors instead");
  if (n > kMaxPassFrames) return absl::InvalidArgumentError("RecordBatch: more frames than one pass holds");
  if (bandcu_stream_begin_capture(stream_) != BANDCU_OK) return CudaError("bandcu_stream_begin_capture");
  absl::Status st = Run(inputs, outputs, n, /*synchronize=*/false);
  void* exec = nullptr;
  const int rc = bandcu_stream_end_capture(stream_, &exec);  // always close the capture, also after a failed Run
  in_flight_ = false;                                         // nothing was enqueued: the launches went into the graph
  if (!st.ok()) {
    if (exec) bandcu_graph_destroy(exec);
    return st;
  }
  if (rc != BANDCU_OK) return CudaError("bandcu_stream_end_capture");
  *graph_exec = exec;
  return absl::OkStatus();
}

bool FusedPlan::IsSingleFastLaunch() const {
  return segments_.size() == 1 && segments_[0].kind == Segment::Kind::kPixel && segments_[0].plan &&
         bandcu_plan_is_fast(segments_[0].plan) == 1;
}

absl::Status FusedPlan::EnsureStreams() {
  if (stream_) return absl::OkStatus();
  if (bandcu_stream_create(&stream_) != BANDCU_OK) return CudaError("bandcu_stream_create");
  if (bandcu_stream_create(&copy_in_) != BANDCU_OK) return CudaError("bandcu_stream_create");
  if (bandcu_stream_create(&copy_out_) != BANDCU_OK) return CudaError("bandcu_stream_create");
  return absl::OkStatus();
}

void* FusedPlan::Event(size_t index) {
  while (events_.size() <= index) {
    void* ev = nullptr;
    if (bandcu_event_create_sync(&ev) != BANDCU_OK) return nullptr;
    events_.push_back(ev);
  }
  return events_[index];
}

absl::Status FusedPlan::EnsureCapacity(int frames, size_t in_bytes, size_t out_bytes) {
  if (in_bytes > in_stage_bytes_) {
    if (in_stage_) bandcu_free(in_stage_);
    in_stage_ = nullptr;
    in_stage_bytes_ = 0;
    if (bandcu_malloc(reinterpret_cast<void**>(&in_stage_), in_bytes) != BANDCU_OK) return CudaError("input staging");
    in_stage_bytes_ = in_bytes;
  }
  if (out_bytes > out_stage_bytes_) {
    if (out_stage_) bandcu_free(out_stage_);
    out_stage_ = nullptr;
    out_stage_bytes_ = 0;
    if (bandcu_malloc(reinterpret_cast<void**>(&out_stage_), out_bytes) != BANDCU_OK) return CudaError("output staging");
    out_stage_bytes_ = out_bytes;
  }
  if (frames > capacity_) {
    for (size_t k = 0; k < scratch_mem_.size(); ++k) {
      if (scratch_mem_[k]) bandcu_free(scratch_mem_[k]);
      scratch_mem_[k] = nullptr;
    }
    capacity_ = 0;
    for (size_t k = 0; k < scratch_mem_.size(); ++k) {
      if (bandcu_malloc(reinterpret_cast<void**>(&scratch_mem_[k]), scratch_bytes_[k] * frames) != BANDCU_OK)
        return CudaError("scratch image");
    }
    capacity_ = frames;
  }
  return absl::OkStatus();
}

SlotPointers FusedPlan::Resolve(int slot, int frame) const {
  if (slot == kSlotInput) return in_ptrs_[frame];
  if (slot == kSlotOutput) return out_ptrs_[frame];
  return ContiguousPointers(scratch_mem_[slot] + static_cast<size_t>(frame) * scratch_bytes_[slot], scratch_[slot]);
}

// Launches every segment for frames [first, last) on the compute stream.
absl::Status FusedPlan::RunSegments(int first, int last, bool rerun) {
  const int m = last - first;
  for (Segment& seg : segments_) {
    switch (seg.kind) {
      case Segment::Kind::kPixel: {
        if (rerun) {  // the shim still holds this segment's previous batch (pointers, strides or table, alignment hints)
          if (bandcu_plan_rerun(seg.plan, stream_) != BANDCU_OK) return CudaError("bandcu_plan_rerun");
          break;
        }
        frames_.resize(m);
        for (int f = 0; f < m; ++f) {
          const SlotPointers s = Resolve(seg.src_slot, first + f), d = Resolve(seg.dst_slot, first + f);
          frames_[f] = BandCuFrame{s.y, s.u, s.v, d.y};
        }
        if (bandcu_plan_run(seg.plan, frames_.data(), m, stream_) != BANDCU_OK) return CudaError("bandcu_plan_run");
        break;
      }
      case Segment::Kind::kYuvMap: {
        // frames whose planes all advance by one stride per frame (batches carved from one allocation, the scratch
        // images) go out as one launch
        for (int f = first; f < last;) {
          const SlotPointers s = Resolve(seg.src_slot, f), d = Resolve(seg.dst_slot, f);
          long long ss = 0, ds = 0;
          int g = f + 1;
          for (; g < last; ++g) {
            const SlotPointers sg = Resolve(seg.src_slot, g), dg = Resolve(seg.dst_slot, g);
            if (g == f + 1) { ss = sg.y - s.y; ds = dg.u - d.u; }
            const long long k = g - f;
            if (sg.y - s.y != ss * k || sg.u - s.u != ss * k || sg.v - s.v != ss * k || dg.u - d.u != ds * k ||
                dg.v - d.v != ds * k || (d.y ? dg.y - d.y != ds * k : dg.y != nullptr))
              break;
          }
          const BandCuYuvPtrs p = {s.y, s.u, s.v, d.y, d.u, d.v};
          if (bandcu_yuv_map_strided(&seg.yuv, &p, ss, ds, g - f, stream_) != BANDCU_OK) return CudaError("bandcu_yuv_map");
          f = g;
        }
        break;
      }
      case Segment::Kind::kRgbToYuv: {
        for (int f = first; f < last;) {
          const SlotPointers s = Resolve(seg.src_slot, f), d = Resolve(seg.dst_slot, f);
          long long ss = 0, ds = 0;
          int g = f + 1;
          for (; g < last; ++g) {
            const SlotPointers sg = Resolve(seg.src_slot, g), dg = Resolve(seg.dst_slot, g);
            if (g == f + 1) { ss = sg.y - s.y; ds = dg.y - d.y; }
            const long long k = g - f;
            if (sg.y - s.y != ss * k || dg.y - d.y != ds * k || dg.u - d.u != ds * k || dg.v - d.v != ds * k) break;
          }
          const BandCuRgbToYuvPtrs p = {s.y, d.y, d.u, d.v};
          if (bandcu_rgb_to_yuv420_strided(&seg.rgb2yuv, &p, ss, ds, g - f, stream_) != BANDCU_OK)
            return CudaError("bandcu_rgb_to_yuv420");
          f = g;
        }
        break;
      }
      case Segment::Kind::kYuvScale: {
        frames_.resize(m);
        for (int k = 0; k < 3; ++k) {
          if (!seg.plane_plan[k]) continue;
          if (rerun) {
            if (bandcu_plan_rerun(seg.plane_plan[k], stream_) != BANDCU_OK) return CudaError("bandcu_plan_rerun");
            continue;
          }
          for (int f = 0; f < m; ++f) {
            const SlotPointers s = Resolve(seg.src_slot, first + f), d = Resolve(seg.dst_slot, first + f);
            const unsigned char* sp = k == 0 ? s.y : (seg.interleaved_uv ? std::min(s.u, s.v) : (k == 1 ? s.u : s.v));
            unsigned char* dp = k == 0 ? d.y : (seg.interleaved_uv ? std::min(d.u, d.v) : (k == 1 ? d.u : d.v));
            frames_[f] = BandCuFrame{sp, nullptr, nullptr, dp};
          }
          if (bandcu_plan_run(seg.plane_plan[k], frames_.data(), m, stream_) != BANDCU_OK)
            return CudaError("bandcu_plan_run");
        }
        break;
      }
      case Segment::Kind::kElementwise: {
        if (seg.count == 0) break;  // kRaw buffers have no pixel elements: silent no-op like the reference
        const size_t ib = GetDataTypeBytes(static_cast<DataType>(seg.in_dtype)) * seg.count;
        const size_t ob = GetDataTypeBytes(static_cast<DataType>(seg.out_dtype)) * seg.count;
        for (int f = first; f < last;) {  // frames that sit back to back go out as one launch
          const SlotPointers s = Resolve(seg.src_slot, f), d = Resolve(seg.dst_slot, f);
          int g = f + 1;
          while (g < last && Resolve(seg.src_slot, g).y == s.y + ib * (g - f) &&
                 Resolve(seg.dst_slot, g).y == d.y + ob * (g - f))
            ++g;
          if (bandcu_normalize(s.y, seg.in_dtype, d.y, seg.out_dtype, seg.count * (g - f), seg.mean, seg.std,
                               stream_) != BANDCU_OK)
            return CudaError("bandcu_normalize");
          f = g;
        }
        break;
      }
    }
  }
  return absl::OkStatus();
}

namespace {
// Frames that sit exactly back to back in PAGEABLE host memory can move as one copy. Pinned memory is
// never merged across frames: a copy must not span two cudaMallocHost allocations, and neighbouring
// allocations cannot be told from one big one here. Frames with gaps are not merged either (the gap may
// be unmapped).
bool UniformStep(const std::vector<HostRanges>& r, size_t* step) {
  if (r.empty() || r[0].count != 1) return false;
  if (r.size() == 1) {
    *step = AlignUp(r[0].bytes[0]);
    return true;
  }
  const ptrdiff_t d = r[1].lo[0] - r[0].lo[0];
  if (d != static_cast<ptrdiff_t>(r[0].bytes[0]) || (d & 15)) return false;
  if (bandcu_pointer_kind(r[0].lo[0]) != 0 || bandcu_pointer_kind(r.back().lo[0]) != 0) return false;
  for (size_t f = 1; f < r.size(); ++f)
    if (r[f].count != 1 || r[f].bytes[0] != r[0].bytes[0] || r[f].lo[0] - r[0].lo[0] != d * static_cast<ptrdiff_t>(f))
      return false;
  *step = static_cast<size_t>(d);
  return true;
}
}  // namespace

absl::Status FusedPlan::RunPass(const Buffer* const* inputs, Buffer* const* outputs, int m, bool in_host,
                                bool out_host, bool synchronize) {
  // Zero-copy input: a chain that lowers to the single staged launch reads every needed source row exactly
  // once, through the TMA engine. When the frames sit in PINNED host memory (mapped into the device address
  // space under UVA) the kernel pulls those rows straight over PCIe: rows the resize never samples never
  // cross the bus and there is no staging copy at all. Pageable frames are staged through HBM as before
  // (or registered on first sight when the plan's registration cache is on, see HostRegistry).
  const bool reads_in_place = segments_.size() == 1 && segments_[0].kind == Segment::Kind::kPixel && segments_[0].plan &&
                              bandcu_plan_staging_kind(segments_[0].plan) == 1;  // the row-staged kernel only: the tile-staged
  // kernel's ~150-byte row fragments are poor PCIe reads (measured on C3: 2.4 K frames/s in place vs 3.8 K staged)
  bool zero_copy = in_host && !writes_input_ && reads_in_place && zero_copy_allowed_;
  for (int f = 0; zero_copy && f < m; ++f)
    if (!HostFrameIsMapped(*inputs[f])) zero_copy = false;
  const bool in_staged = in_host && !zero_copy;
  // Zero-copy output: the staged kernel assembles rows in shared memory and pushes them into PINNED host tensors
  // with bulk stores, so results cross PCIe while the next source rows are being pulled (both directions of the
  // link busy at once) and neither a staging buffer nor a D2H copy exists. Needs 16-byte granularity.
  bool zc_out = out_host && writes_output_ && IsSingleFastLaunch() && zero_copy_allowed_ &&
                output_.num_planes == 1 && output_.row_stride[0] % 16 == 0 &&
                (static_cast<size_t>(segments_[0].pixel.dst_width) * segments_[0].pixel.channels *
                 GetDataTypeBytes(output_.dtype)) % 16 == 0;
  for (int f = 0; zc_out && f < m; ++f) {
    const unsigned char* p = static_cast<const Buffer&>(*outputs[f])[0].data;
    if ((reinterpret_cast<uintptr_t>(p) & 15) != 0 || !outputs[f]->IsPinnedHost()) zc_out = false;
  }
  if (IsSingleFastLaunch() && bandcu_plan_set_host_output(segments_[0].plan, zc_out ? 1 : 0) != BANDCU_OK)
    return CudaError("bandcu_plan_set_host_output");
  if (zc_out) out_host = false;  // from here on the tensors are addressed like device memory
  std::vector<HostRanges> in_r, out_r;
  std::vector<PlaneRects> out_rect;
  size_t in_step = 0, out_step = 0;
  bool in_uniform = false, out_uniform = false;
  bool out_preload = false;  // some output's written bytes are not whole row segments: keep the caller's bytes around them
  if (in_staged) {
    in_r.reserve(m);
    for (int f = 0; f < m; ++f) in_r.push_back(RangesOf(*inputs[f], input_));
    in_uniform = UniformStep(in_r, &in_step);
    if (!in_uniform) {
      for (const HostRanges& r : in_r) in_step = std::max(in_step, r.Total());
    }
  }
  const bool out_staged = out_host && writes_output_;
  if (out_staged) {
    out_r.reserve(m);
    out_rect.reserve(m);
    bool dense = true;
    for (int f = 0; f < m; ++f) {
      out_r.push_back(RangesOf(*outputs[f], output_));
      out_rect.push_back(RectsOf(*outputs[f], output_));
      if (!out_rect.back().Dense()) dense = false;
      if (!out_rect.back().exact) out_preload = true;
    }
    out_uniform = dense && UniformStep(out_r, &out_step);
    if (!out_uniform) {
      out_step = 0;
      for (const HostRanges& r : out_r) out_step = std::max(out_step, r.Total());
    }
  }
  RETURN_IF_ERROR(EnsureCapacity(m, in_step * m + kAlign, out_step * m + kAlign));
  in_ptrs_.resize(m);
  out_ptrs_.assign(m, SlotPointers());
  for (int f = 0; f < m; ++f) {
    in_ptrs_[f] = in_staged ? StagedPointers(*inputs[f], input_, in_r[f], in_stage_ + in_step * f)
                          : BufferPointers(*inputs[f]);
    if (writes_output_)
      out_ptrs_[f] = out_host ? StagedPointers(*outputs[f], output_, out_r[f], out_stage_ + out_step * f)
                              : BufferPointers(*outputs[f]);
  }
  // device address, inside frame f's output staging area, of the host output byte `p`
  auto staged_out = [&](int f, const unsigned char* p) -> unsigned char* {
    size_t off = 0;
    const HostRanges& r = out_r[f];
    for (int i = 0; i < r.count; ++i) {
      if (p >= r.lo[i] && p < r.lo[i] + r.bytes[i]) return out_stage_ + out_step * f + off + (p - r.lo[i]);
      off += AlignUp(r.bytes[i]);
    }
    return nullptr;
  };
  const bool download = out_staged || (in_staged && writes_input_);
  const bool staged = in_staged || download;
  last_pass_staged_ = staged;
  const int chunk = staged ? chunk_frames_ : m;
  size_t ev = 0;
  if (staged && in_flight_) {
    // an earlier asynchronous pass may still use the staging areas: its kernels read in_stage_, its
    // downloads read out_stage_ (and in_stage_ when an in-place stage wrote the staged input back)
    if (!fence_a_ && bandcu_event_create_sync(&fence_a_) != BANDCU_OK) return CudaError("fence event");
    if (!fence_b_ && bandcu_event_create_sync(&fence_b_) != BANDCU_OK) return CudaError("fence event");
    if (bandcu_event_record(fence_a_, stream_) != BANDCU_OK || bandcu_stream_wait_event(copy_in_, fence_a_) != BANDCU_OK ||
        bandcu_event_record(fence_b_, copy_out_) != BANDCU_OK || bandcu_stream_wait_event(stream_, fence_b_) != BANDCU_OK ||
        bandcu_stream_wait_event(copy_in_, fence_b_) != BANDCU_OK)
      return CudaError("fence");
  }
  for (int a = 0; a < m; a += chunk) {
    const int b = std::min(m, a + chunk);
    if (in_staged) {
      if (in_uniform) {
        const size_t bytes = in_step * (b - a - 1) + in_r[a].bytes[0];
        if (bandcu_memcpy_h2d(in_stage_ + in_step * a, in_r[a].lo[0], bytes, copy_in_) != BANDCU_OK)
          return CudaError("upload");
      } else {
        for (int f = a; f < b; ++f) {
          size_t off = 0;
          for (int i = 0; i < in_r[f].count; ++i) {
            if (bandcu_memcpy_h2d(in_stage_ + in_step * f + off, in_r[f].lo[i], in_r[f].bytes[i], copy_in_) != BANDCU_OK)
              return CudaError("upload");
            off += AlignUp(in_r[f].bytes[i]);
          }
        }
      }
    }
    if (out_preload) {  // the staging image starts as a copy of the caller's bytes, so a whole-range download restores them
      for (int f = a; f < b; ++f) {
        size_t off = 0;
        for (int i = 0; i < out_r[f].count; ++i) {
          if (bandcu_memcpy_h2d(out_stage_ + out_step * f + off, out_r[f].lo[i], out_r[f].bytes[i], copy_in_) != BANDCU_OK)
            return CudaError("upload");
          off += AlignUp(out_r[f].bytes[i]);
        }
      }
    }
    if (in_staged || out_preload) {
      void* e = Event(ev++);
      if (!e || bandcu_event_record(e, copy_in_) != BANDCU_OK || bandcu_stream_wait_event(stream_, e) != BANDCU_OK)
        return CudaError("upload event");
    }
    RETURN_IF_ERROR(RunSegments(a, b));
    if (download) {
      void* e = Event(ev++);
      if (!e || bandcu_event_record(e, stream_) != BANDCU_OK || bandcu_stream_wait_event(copy_out_, e) != BANDCU_OK)
        return CudaError("compute event");
      if (out_staged) {
        if (out_uniform) {
          const size_t bytes = out_step * (b - a - 1) + out_r[a].bytes[0];
          if (bandcu_memcpy_d2h(const_cast<unsigned char*>(out_r[a].lo[0]), out_stage_ + out_step * a, bytes,
                                copy_out_) != BANDCU_OK)
            return CudaError("download");
        } else {
          for (int f = a; f < b; ++f) {
            const PlaneRects& pr = out_rect[f];
            if (pr.Dense() || !pr.exact) {  // whole ranges: nothing but written (or preloaded) bytes inside
              size_t off = 0;
              for (int i = 0; i < out_r[f].count; ++i) {
                if (bandcu_memcpy_d2h(const_cast<unsigned char*>(out_r[f].lo[i]), out_stage_ + out_step * f + off,
                                      out_r[f].bytes[i], copy_out_) != BANDCU_OK)
                  return CudaError("download");
                off += AlignUp(out_r[f].bytes[i]);
              }
              continue;
            }
            // padded rows / planes with gaps: exactly the row segments the chain writes, pitch to pitch
            for (int i = 0; i < pr.count; ++i) {
              unsigned char* src = staged_out(f, pr.lo[i]);
              if (!src) return absl::InternalError("output plane outside its staging range");
              if (bandcu_memcpy2d_d2h(const_cast<unsigned char*>(pr.lo[i]), pr.stride[i], src, pr.stride[i],
                                      pr.row_bytes[i], pr.rows[i], copy_out_) != BANDCU_OK)
                return CudaError("download");
            }
          }
        }
      }
      if (in_staged && writes_input_) {  // an in-place stage changed the (staged) chain input: copy it back
        for (int f = a; f < b; ++f) {
          size_t off = 0;
          for (int i = 0; i < in_r[f].count; ++i) {
            if (bandcu_memcpy_d2h(const_cast<unsigned char*>(in_r[f].lo[i]), in_stage_ + in_step * f + off,
                                  in_r[f].bytes[i], copy_out_) != BANDCU_OK)
              return CudaError("download");
            off += AlignUp(in_r[f].bytes[i]);
          }
        }
      }
    }
  }
  in_flight_ = true;
  if (!synchronize) return absl::OkStatus();
  if (bandcu_stream_sync(stream_) != BANDCU_OK) return CudaError("stream sync");
  if (download && bandcu_stream_sync(copy_out_) != BANDCU_OK) return CudaError("stream sync");
  in_flight_ = false;
  ReleaseHeldHost();
  return absl::OkStatus();
}

absl::Status FusedPlan::Rerun(bool synchronize) {
  if (!rerunnable_ || last_frames_ <= 0) return absl::FailedPreconditionError("Rerun: no replayable batch");
  RETURN_IF_ERROR(RunSegments(0, last_frames_, /*rerun=*/true));
  in_flight_ = true;
  if (!synchronize) return absl::OkStatus();
  if (bandcu_stream_sync(stream_) != BANDCU_OK) return CudaError("stream sync");
  in_flight_ = false;
  return absl::OkStatus();
}

absl::Status FusedPlan::Run(const Buffer* const* inputs, Buffer* const* outputs, int n, bool synchronize) {
  rerunnable_ = false;
  if (n <= 0) return absl::OkStatus();
  if (inputs == nullptr || (writes_output_ && outputs == nullptr)) return absl::InvalidArgumentError("null frame list");
  const MemorySpace in_space = inputs[0] ? inputs[0]->GetMemorySpace() : MemorySpace::kHost;
  const MemorySpace out_space = (writes_output_ && outputs[0]) ? outputs[0]->GetMemorySpace() : MemorySpace::kHost;
  for (int f = 0; f < n; ++f) {
    if (inputs[f] == nullptr) return absl::InvalidArgumentError("null input buffer");
    if (!(ImageLayout::Of(*inputs[f]) == input_) || inputs[f]->GetMemorySpace() != in_space)
      return absl::InvalidArgumentError("every input of a batch must share format, dimensions, strides and memory space");
    if (!writes_output_) continue;
    if (outputs[f] == nullptr) return absl::InternalError("Null output buffer");
    if (!(ImageLayout::Of(*outputs[f]) == output_) || outputs[f]->GetMemorySpace() != out_space)
      return absl::InvalidArgumentError("every output of a batch must share format, dimensions, strides and memory space");
  }
  if (in_space == MemorySpace::kVirtual || out_space == MemorySpace::kVirtual)
    return absl::InternalError("a shape-only buffer cannot be processed");
  RETURN_IF_ERROR(EnsureStreams());
  for (int base = 0; base < n; base += kMaxPassFrames) {
    const int m = std::min(kMaxPassFrames, n - base);
    RETURN_IF_ERROR(RunPass(inputs + base, writes_output_ ? outputs + base : nullptr, m,
                            in_space == MemorySpace::kHost, out_space == MemorySpace::kHost, synchronize));
  }
  // one pass that moved nothing through staging and holds no registered host range: the same launches may be replayed
  rerunnable_ = n <= kMaxPassFrames && !last_pass_staged_ && !register_host_;
  last_frames_ = n;
  return absl::OkStatus();
}

absl::Status RunSingleOperator(IBufferOperator& op, const Buffer& input) {
  Buffer* out = op.GetOutput();
  std::shared_ptr<FusedPlan>& cached = op.single_plan_;
  const bool reusable = cached && cached->input_layout() == ImageLayout::Of(input) &&
                        (!cached->writes_output() || (out && cached->output_layout() == ImageLayout::Of(*out)));
  if (!reusable) {
    cached.reset();
    ChainLowering lowering(input);
    RETURN_IF_ERROR(lowering.LowerNested(op, input));
    absl::StatusOr<std::unique_ptr<FusedPlan>> plan = lowering.Finish(out);
    if (!plan.ok()) return plan.status();
    cached = std::move(plan).value();
  }
  const Buffer* in = &input;
  return cached->Run(&in, &out, 1);
}

absl::Status SetDevice(int device) {
  if (bandcu_set_device(device) != BANDCU_OK) return CudaError("bandcu_set_device");
  return absl::OkStatus();
}

}  // namespace cuda
}  // namespace band
