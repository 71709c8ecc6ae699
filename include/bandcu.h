/*
 * include/bandcu.h -- C-ABI of the CUDA shim (libbandcu.so).
 *
 * This is the thin boundary the C++ host layer (band_b200/csrc/band/...) calls
 * instead of libyuv: plain pointers, sizes and POD structs; no C++ types, no
 * exceptions, no torch types. It replaces, for the GPU backing, what
 * /root/reference/band/buffer/libyuv_image_operator.h:28-60
 * (LibyuvImageOperator::{Crop,Resize,Rotate,FlipHorizontally,FlipVertically,
 * ColorSpaceConvert}) and band/buffer/common_operator.cc:24-40
 * (NormalizeFromTo) provide on the CPU -- but as ONE fused launch per frame
 * batch instead of one materialised pass per operator.
 *
 * Every function returns 0 on success or a non-zero cudaError_t / BANDCU_E_*
 * code; bandcu_last_error_string() describes the last failure on the calling
 * thread. There is no CPU fallback: without a CUDA device every compute entry
 * point fails.
 */
#ifndef BAND_INCLUDE_BANDCU_H_
#define BAND_INCLUDE_BANDCU_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BANDCU_API __attribute__((visibility("default")))

enum {
  BANDCU_OK = 0,
  BANDCU_E_INVALID = 10001,     /* malformed descriptor / null pointer */
  BANDCU_E_UNSUPPORTED = 10002, /* combination the kernels do not implement */
};

/* Source pixel layouts. YUV420 covers NV12/NV21/YV12/YV21 uniformly: separate
 * U and V base pointers plus a chroma pixel stride (2 = interleaved, 1 =
 * planar), exactly how GetYuvDataFromBuffer (libyuv_image_operator.cc:51-88)
 * describes them. */
enum {
  BANDCU_SRC_GRAY = 0,
  BANDCU_SRC_RGB = 1,
  BANDCU_SRC_RGBA = 2,
  BANDCU_SRC_YUV420 = 3,
  BANDCU_SRC_UV = 4, /* one interleaved chroma plane as a 2-channel image (UVScale inside NV12Scale, :422) */
};

/* Per-pixel conversion applied when a source pixel is fetched. */
enum {
  BANDCU_CVT_NONE = 0,      /* keep the source channels (GRAY:1, RGB:3, RGBA:4) */
  BANDCU_CVT_YUV2RGB = 1,   /* NV12ToRAW / NV21ToRAW / I420ToRAW  (:102,:190,:278) */
  BANDCU_CVT_YUV2RGBA = 2,  /* ...ToABGR, A = 255                 (:114,:202,:291) */
  BANDCU_CVT_YUV2GRAY = 3,  /* Y-plane copy                       (:163-170)       */
  BANDCU_CVT_RGB2GRAY = 4,  /* RAWToJ400                          (:526)           */
  BANDCU_CVT_RGBA2GRAY = 5, /* ABGRToARGB + ARGBToJ400            (:592-619)       */
  BANDCU_CVT_RGBA2RGB = 6,  /* ARGBToRGB24 on RGBA bytes          (:675-687)       */
};

/* Which libyuv scaler family the Resize stage follows. */
enum {
  BANDCU_FLAVOR_ARGB = 0,  /* ARGBScale: RGB / RGBA (libyuv_image_operator.cc:1175,:1219) */
  BANDCU_FLAVOR_PLANE = 1, /* ScalePlane: gray (:1373) */
};

/* Scaler code path, chosen on the host by the libyuv dispatch rules
 * (scale_argb.cc ScaleARGB / scale.cc ScalePlane). */
enum {
  BANDCU_SCALE_COPY = 0,          /* no Resize stage */
  BANDCU_SCALE_POINT = 1,         /* Scale*Simple */
  BANDCU_SCALE_BILINEAR_DOWN = 2, /* Scale*BilinearDown (rows then columns) */
  BANDCU_SCALE_BILINEAR_UP = 3,   /* Scale*BilinearUp (columns then rows) */
  BANDCU_SCALE_BOX2 = 4,          /* ScaleARGBDown2/DownEven box, ScalePlaneDown2 box */
  BANDCU_SCALE_VERTICAL = 5,      /* ScalePlaneVertical */
  BANDCU_SCALE_DOWN34 = 6,        /* ScalePlaneDown34 (plane only) */
  BANDCU_SCALE_DOWN38 = 7,        /* ScalePlaneDown38 (plane only) */
  BANDCU_SCALE_UP2_LINEAR = 8,    /* ScalePlaneUp2_Linear / ScaleUVLinearUp2 (plane and UV only) */
  BANDCU_SCALE_UP2_BILINEAR = 9,  /* ScalePlaneUp2_Bilinear / ScaleUVBilinearUp2 (plane and UV only) */
};

/* Output element types (Normalize / DataTypeConvert targets,
 * common_operator.cc:42-65). Values follow band::DataType (common.h:115-128). */
enum {
  BANDCU_DT_F32 = 1,
  BANDCU_DT_I32 = 2,
  BANDCU_DT_U8 = 3,
  BANDCU_DT_I16 = 7,
  BANDCU_DT_I8 = 9,
  BANDCU_DT_F64 = 11,
};

/* Integer affine index map: row = r0 + rr*r + rc*c ; col = c0 + cr*r + cc*c.
 * Crop / Rotate / Flip compose into these on the host. */
typedef struct BandCuAffine {
  int r0, rr, rc;
  int c0, cr, cc;
} BandCuAffine;

/* A fused plan: [index maps] -> [pixel conversion] -> [Resize] -> [Normalize]
 * evaluated lazily per output element; results are identical to running the
 * reference's operator chain stage by stage. */
typedef struct BandCuPlanDesc {
  /* source planes */
  int src_kind;        /* BANDCU_SRC_* */
  int stride0;         /* bytes between rows of plane 0 (Y or packed) */
  int stride_uv;       /* bytes between chroma rows */
  int pixel_stride_uv; /* 1 planar, 2 interleaved */
  /* virtual image seen by the arithmetic stages (after all index maps) */
  int vw, vh;
  BandCuAffine luma_map;   /* virtual (r,c) -> plane-0 (row, col) */
  BandCuAffine pre_map;    /* virtual (r,c) -> coordinate (r1,c1) in the image the colour conversion ran on */
  BandCuAffine chroma_map; /* (r1>>1, c1>>1) -> chroma plane (row, col) */
  int cvt;                 /* BANDCU_CVT_* */
  /* Resize stage */
  int flavor;     /* BANDCU_FLAVOR_* */
  int scale_mode; /* BANDCU_SCALE_* */
  int linear;     /* filter reduced to kFilterLinear: no vertical blend */
  int x0, y0, dx, dy; /* 16.16 */
  int max_y;          /* clamp for y (mode dependent) */
  int dst_width, dst_height;
  /* output */
  int channels;       /* 1, 2 (UV), 3 or 4 channels per output pixel */
  int out_dtype;      /* BANDCU_DT_* */
  int normalize;      /* 0: store u8 results; 1: out = (T)((T)(v - mean) / std) */
  float mean, std_;
  long long dst_row_stride; /* bytes between output rows (contiguous NHWC: dst_width*channels*sizeof(T)) */
  /* luma dimensions of the SOURCE planes (before any index map); 0 = unknown, which only disables the tile-staged
   * kernel (it stages whole 32-bit words and must know they stay inside the plane rows) */
  int src_width, src_height;
} BandCuPlanDesc;

typedef struct BandCuPlan BandCuPlan;

/* Per-frame pointers (device memory). pu/pv unused for packed sources. */
typedef struct BandCuFrame {
  const void* p0;
  const void* pu;
  const void* pv;
  void* out;
} BandCuFrame;

/* ---- plans ---------------------------------------------------------------- */
BANDCU_API int bandcu_plan_create(const BandCuPlanDesc* desc, BandCuPlan** plan);
BANDCU_API int bandcu_plan_destroy(BandCuPlan* plan);
/* 1 if the plan runs on the staged/specialised fast kernel, 0 for the generic one. */
BANDCU_API int bandcu_plan_is_fast(const BandCuPlan* plan);
/* 1 if the plan has a staged kernel that reads every needed source byte exactly once, in bulk (row-staged: 1,
 * tile-staged: 2), so that pinned host frames can be read in place over PCIe; 0 otherwise. */
BANDCU_API int bandcu_plan_staging_kind(const BandCuPlan* plan);
/* Output hand-off of the staged kernel. 0 (default): direct global stores (tensors in HBM). 1: the tensors of the
 * following runs live in PINNED HOST memory; rows are assembled in shared memory and leave as cp.async.bulk stores,
 * so results cross PCIe inside the kernel while it pulls the next source rows (no staging buffer, no D2H copy).
 * Ignored by plans that do not run on the staged kernel (bandcu_plan_is_fast() == 0). */
BANDCU_API int bandcu_plan_set_host_output(BandCuPlan* plan, int on);
/* Run n frames whose pointers advance by a fixed stride (bytes). */
BANDCU_API int bandcu_plan_run_strided(BandCuPlan* plan, const BandCuFrame* first, long long in_stride,
                                       long long out_stride, int n, void* stream);
/* Run n frames with arbitrary device pointers (host array; uploaded internally). */
BANDCU_API int bandcu_plan_run(BandCuPlan* plan, const BandCuFrame* frames, int n, void* stream);
/* Launch the plan's previous batch again: same frame pointers, same count (bandcu_plan_run / _run_strided must have
 * succeeded on this plan before). Skips the per-frame host work -- stride detection, alignment hints, pointer-table
 * comparison -- which bounds the rate of batches of very small frames. */
BANDCU_API int bandcu_plan_rerun(BandCuPlan* plan, void* stream);
/* number of kernel launches issued by this library on the calling process so far */
BANDCU_API long long bandcu_launch_count(void);

/* ---- a two-stage chain that allocates nothing -------------------------------
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 *
 * Runs `first` on the frame's planes into an intermediate u8 image, then `second` (whose source is that image:
 * src_kind GRAY / RGB / RGBA == first->channels, stride0 == first->dst_row_stride) into frame->out -- what the plan
 * compiler does with plan-owned cudaMalloc'ed scratch for chains that cannot be one launch, e.g. 4:2:0 -> RGBA followed
 * by the bilinear up-scale of that image. The intermediate image lives in a file-scope device array of
 * BANDCU_NOALLOC_SCRATCH_BYTES (bandcu_noalloc.cu): every GPU that loads libbandcu.so holds those 8 MiB until exit.
 * first->dst_row_stride * first->dst_height above that size is rejected with BANDCU_E_INVALID and a message; so is a
 * second stage whose index map leaves the intermediate image. Both stages run on the pixel-per-thread kernel (values
 * identical to bandcu_plan_run's). The call makes no cudaMalloc / cudaFree / cudaMallocHost and no allocating library
 * call; the first call on a device resolves the array's address (cudaGetSymbolAddress) -- make that one outside stream
 * capture. ALL calls share the array, which keeps the previous call's bytes (nothing needs to start at zero: stage 1
 * writes every pixel stage 2 reads); calls that overlap on one GPU (two streams, two threads) RACE on it -- one call at
 * a time per GPU. frame->p0/pu/pv/out are device pointers as for bandcu_plan_run. */
#define BANDCU_NOALLOC_SCRATCH_BYTES (8u << 20)
BANDCU_API int bandcu_two_stage_run_noalloc(const BandCuPlanDesc* first, const BandCuPlanDesc* second,
                                            const BandCuFrame* frame, void* stream);

/* ---- 4:2:0 -> 4:2:0 index-mapped copy (Crop/Rotate/Flip/ColorSpaceConvert with a
 * YUV output: CropNv/CropYv :950-1037, RotateYv/RotateNv :786-887, YUV<->YUV
 * repacks :125-162,:213-250,:302-367, Flip* :1232-1363) -------------------- */
typedef struct BandCuYuvMapDesc {
  int src_stride_y, src_stride_uv, src_pixel_stride_uv;
  int dst_stride_y, dst_stride_uv, dst_pixel_stride_uv;
  int dst_width, dst_height;   /* luma dims of the output; chroma dims = (d+1)/2 */
  int chroma_width_override;   /* >0: copy this many chroma samples per row instead of (dst_width+1)/2 */
  BandCuAffine luma_map;       /* dst (r,c) -> src luma (row, col) */
  BandCuAffine chroma_map;     /* dst chroma (r,c) -> src chroma (row, col) */
} BandCuYuvMapDesc;
typedef struct BandCuYuvPtrs {
  const void *sy, *su, *sv;
  void *dy, *du, *dv;
} BandCuYuvPtrs;
BANDCU_API int bandcu_yuv_map(const BandCuYuvMapDesc* desc, const BandCuYuvPtrs* ptrs, void* stream);
/* n frames whose plane pointers all advance by src_frame_stride / dst_frame_stride bytes per frame: one launch. */
BANDCU_API int bandcu_yuv_map_strided(const BandCuYuvMapDesc* desc, const BandCuYuvPtrs* first, long long src_frame_stride,
                                      long long dst_frame_stride, int n, void* stream);

/* ---- RGB / RGBA -> 4:2:0 (ColorSpaceConvert to NV12/NV21/YV12/YV21: RAWToI420 :570, ABGRToNV12 :626,
 * ABGRToNV21 :643, ABGRToI420 :661; libyuv's default C rows RGBToY / RGB2xToU / RGB2xToV with the 2x2
 * chroma average). One launch per frame; a pending Crop/Rotate/Flip of the packed source folds into src_map. */
typedef struct BandCuRgbToYuvDesc {
  int src_channels;            /* 3 (RGB) or 4 (RGBA: bytes R,G,B,A) */
  int src_stride;              /* bytes between source rows */
  BandCuAffine src_map;        /* converted-image (r,c) -> source (row, col) */
  int width, height;           /* dims of the converted image */
  int dst_stride_y, dst_stride_uv, dst_pixel_stride_uv;
} BandCuRgbToYuvDesc;
typedef struct BandCuRgbToYuvPtrs {
  const void* src;
  void *dy, *du, *dv;
} BandCuRgbToYuvPtrs;
BANDCU_API int bandcu_rgb_to_yuv420(const BandCuRgbToYuvDesc* desc, const BandCuRgbToYuvPtrs* ptrs, void* stream);
BANDCU_API int bandcu_rgb_to_yuv420_strided(const BandCuRgbToYuvDesc* desc, const BandCuRgbToYuvPtrs* first,
                                            long long src_frame_stride, long long dst_frame_stride, int n, void* stream);

/* ---- element-wise Normalize / DataTypeConvert on an arbitrary dtype pair
 * (common_operator.cc:24-65); n elements, contiguous. ----------------------- */
BANDCU_API int bandcu_normalize(const void* in, int in_dtype, void* out, int out_dtype, long long n, float mean,
                                float std_, void* stream);

/* ---- device plumbing ------------------------------------------------------ */
BANDCU_API int bandcu_device_count(int* count);
BANDCU_API int bandcu_set_device(int device);
BANDCU_API int bandcu_malloc(void** ptr, size_t bytes);
BANDCU_API int bandcu_free(void* ptr);
BANDCU_API int bandcu_host_alloc(void** ptr, size_t bytes); /* pinned */
BANDCU_API int bandcu_host_free(void* ptr);
BANDCU_API int bandcu_memcpy_h2d(void* dst, const void* src, size_t bytes, void* stream);
BANDCU_API int bandcu_memcpy_d2h(void* dst, const void* src, size_t bytes, void* stream);
BANDCU_API int bandcu_memcpy_d2d(void* dst, const void* src, size_t bytes, void* stream);
/* `height` rows of `width` bytes, device pitch -> host pitch: downloads into padded host images write only the bytes
 * libyuv would write (row padding stays the caller's). */
BANDCU_API int bandcu_memcpy2d_d2h(void* dst, size_t dst_pitch, const void* src, size_t src_pitch, size_t width,
                                   size_t height, void* stream);
/* Page-lock and map caller-owned host memory (cudaHostRegister, portable across devices) so kernels can read it in
 * place; `ptr` and `bytes` must be multiples of the 4 KB page. */
BANDCU_API int bandcu_host_register(void* ptr, size_t bytes);
BANDCU_API int bandcu_host_unregister(void* ptr);
/* 1 when kernels on the current device may dereference a registered host pointer as it is (UVA, no separate device
 * alias: cudaDevAttrCanUseHostPointerForRegisteredMem), 0 when not, <0 on error. */
BANDCU_API int bandcu_registered_host_pointers_usable(void);
BANDCU_API int bandcu_memset(void* dst, int value, size_t bytes, void* stream);
BANDCU_API int bandcu_stream_create(void** stream);
BANDCU_API int bandcu_stream_destroy(void* stream);
BANDCU_API int bandcu_stream_sync(void* stream);
/* ---- CUDA graphs: record what a stream is given between begin/end, replay it with one launch ----
 * (launch-bound cases, e.g. the single-frame latency configuration). Capture is thread-local and relaxed: allocation
 * calls made by the recording thread stay legal. A pointer-table batch cannot be recorded (its table upload would be
 * replayed from plan-owned memory that later runs rewrite): bandcu_plan_run() fails while its stream is capturing. */
BANDCU_API int bandcu_stream_begin_capture(void* stream);
BANDCU_API int bandcu_stream_end_capture(void* stream, void** graph_exec);
BANDCU_API int bandcu_graph_launch(void* graph_exec, void* stream);
BANDCU_API int bandcu_graph_destroy(void* graph_exec);
BANDCU_API int bandcu_event_create(void** ev);
BANDCU_API int bandcu_event_create_sync(void** ev); /* ordering only (timing disabled) */
BANDCU_API int bandcu_event_destroy(void* ev);
BANDCU_API int bandcu_event_record(void* ev, void* stream);
BANDCU_API int bandcu_event_elapsed_ms(void* start, void* stop, float* ms);
BANDCU_API int bandcu_stream_wait_event(void* stream, void* ev);
/* 1 if ptr is device (or managed) memory, 0 if host, <0 on error. */
BANDCU_API int bandcu_pointer_is_device(const void* ptr);
/* 0 pageable host memory, 1 pinned / registered host memory, 2 device or managed memory, <0 on error. */
BANDCU_API int bandcu_pointer_kind(const void* ptr);
/* ordinal of the device a device pointer was allocated on; -1 for host memory or on error. */
BANDCU_API int bandcu_pointer_device(const void* ptr);
BANDCU_API const char* bandcu_last_error_string(void);

#ifdef __cplusplus
}
#endif
#endif /* BAND_INCLUDE_BANDCU_H_ */
