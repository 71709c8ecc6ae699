// band_b200/csrc/cuda/bandcu_api.cu -- extern "C" surface of libbandcu.so
// (declared in include/bandcu.h). POD in, int out; no C++ types or exceptions
// cross this boundary.
#include <atomic>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <new>

#include "bandcu_device.cuh"
#include "bandcu_internal.h"

namespace bandcu {

static std::atomic<long long> g_launches{0};
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

static thread_local char g_err[512] = "";
static int fail(int code, const char* what) {
  if (code >= 10000) {
    snprintf(g_err, sizeof(g_err), "bandcu: %s", what);
  } else {
    snprintf(g_err, sizeof(g_err), "bandcu: %s: %s (%s)", what, cudaGetErrorName((cudaError_t)code),
             cudaGetErrorString((cudaError_t)code));
    (void)cudaGetLastError();  // clear the sticky-less error state
  }
  return code;
}
// the same thread-local message for the entry points that live in other files (bandcu_noalloc.cu)
int report_error(int code, const char* what) { return fail(code, what); }
#define CU_TRY(expr, what)                      \
  do {                                          \
    cudaError_t _e = (expr);                    \
    if (_e != cudaSuccess) return fail((int)_e, what); \
  } while (0)

}  // namespace bandcu

using namespace bandcu;

struct BandCuPlan {
  BandCuPlanDesc d;
  FastPlan* fast;
  TilePlan* tile;
  PackedPlan* packed;
  // pointer-table staging for bandcu_plan_run()
  BandCuFrame* table_dev;
  BandCuFrame* table_host;  // pinned
  int table_cap;
  cudaEvent_t table_free;   // recorded after the launch that consumed the table
  bool table_pending;
  cudaStream_t table_stream;  // stream the upload was ordered on
  int table_valid_n;        // >0: table_dev holds exactly table_host[0..n) (reused when the next call repeats it)
  // the previous successful batch, for bandcu_plan_rerun
  BatchArgs last_b;
  TableHints last_hints;
  int last_n;
};

static int validate_desc(const BandCuPlanDesc* d) {
  if (!d) return BANDCU_E_INVALID;
  if (d->src_kind < BANDCU_SRC_GRAY || d->src_kind > BANDCU_SRC_UV) return BANDCU_E_INVALID;
  if (d->vw <= 0 || d->vh <= 0 || d->dst_width <= 0 || d->dst_height <= 0) return BANDCU_E_INVALID;
  if (d->channels < 1 || d->channels > 4) return BANDCU_E_INVALID;
  if (d->scale_mode < BANDCU_SCALE_COPY || d->scale_mode > BANDCU_SCALE_UP2_BILINEAR) return BANDCU_E_INVALID;
  const bool up2 = d->scale_mode == BANDCU_SCALE_UP2_LINEAR || d->scale_mode == BANDCU_SCALE_UP2_BILINEAR;
  if (d->scale_mode >= BANDCU_SCALE_DOWN34 && !(up2 && d->src_kind == BANDCU_SRC_UV) &&
      (d->flavor != BANDCU_FLAVOR_PLANE || d->channels != 1))
    return BANDCU_E_UNSUPPORTED;
  if (d->flavor == BANDCU_FLAVOR_PLANE && d->channels != 1) return BANDCU_E_UNSUPPORTED;
  if (d->normalize) {
    switch (d->out_dtype) {
      case BANDCU_DT_F32: case BANDCU_DT_U8: case BANDCU_DT_I8: case BANDCU_DT_I16: case BANDCU_DT_I32: break;
      default: return BANDCU_E_UNSUPPORTED;
    }
  }
  // channel count must agree with the conversion
  int ch;
  switch (d->cvt) {
    case BANDCU_CVT_NONE:
      ch = d->src_kind == BANDCU_SRC_GRAY ? 1 : d->src_kind == BANDCU_SRC_RGB ? 3 : d->src_kind == BANDCU_SRC_RGBA ? 4
           : d->src_kind == BANDCU_SRC_UV ? 2 : -1;
      break;
    case BANDCU_CVT_YUV2RGB: ch = d->src_kind == BANDCU_SRC_YUV420 ? 3 : -1; break;
    case BANDCU_CVT_YUV2RGBA: ch = d->src_kind == BANDCU_SRC_YUV420 ? 4 : -1; break;
    case BANDCU_CVT_YUV2GRAY: ch = d->src_kind == BANDCU_SRC_YUV420 ? 1 : -1; break;
    case BANDCU_CVT_RGB2GRAY: ch = d->src_kind == BANDCU_SRC_RGB ? 1 : -1; break;
    case BANDCU_CVT_RGBA2GRAY: ch = d->src_kind == BANDCU_SRC_RGBA ? 1 : -1; break;
    case BANDCU_CVT_RGBA2RGB: ch = d->src_kind == BANDCU_SRC_RGBA ? 3 : -1; break;
    default: ch = -1;
  }
  if (ch != d->channels) return BANDCU_E_INVALID;
  return BANDCU_OK;
}

extern "C" {

int bandcu_plan_create(const BandCuPlanDesc* desc, BandCuPlan** plan) {
  if (!plan) return fail(BANDCU_E_INVALID, "plan_create: null out pointer");
  *plan = nullptr;
  int v = validate_desc(desc);
  if (v != BANDCU_OK) return fail(v, "plan_create: invalid or unsupported descriptor");
  BandCuPlan* p = new (std::nothrow) BandCuPlan();
  if (!p) return fail(BANDCU_E_INVALID, "plan_create: out of host memory");
  memset(p, 0, sizeof(*p));
  p->d = *desc;
  if (fast_plan_supported(p->d)) {
    int e = fast_plan_create(p->d, &p->fast);
    if (e != BANDCU_OK) {
      delete p;
      return fail(e, "plan_create: fast plan setup");
    }
  }
  if (tile_plan_supported(p->d)) {  // also built next to a fast plan: batches the row-staged kernel refuses land here
    int e = tile_plan_create(p->d, &p->tile);
    if (e != BANDCU_OK && e != BANDCU_E_UNSUPPORTED) {
      if (p->fast) fast_plan_destroy(p->fast);
      delete p;
      return fail(e, "plan_create: tile plan setup");
    }
  }
  if (!p->fast && packed_plan_supported(p->d)) {
    int e = packed_plan_create(p->d, &p->packed);
    if (e != BANDCU_OK) {
      delete p;
      return fail(e, "plan_create: packed plan setup");
    }
  }
  *plan = p;
  return BANDCU_OK;
}

int bandcu_plan_destroy(BandCuPlan* plan) {
  if (!plan) return BANDCU_OK;
  if (plan->table_pending) (void)cudaEventSynchronize(plan->table_free);
  if (plan->table_free) (void)cudaEventDestroy(plan->table_free);
  if (plan->table_dev) (void)cudaFree(plan->table_dev);
  if (plan->table_host) (void)cudaFreeHost(plan->table_host);
  if (plan->fast) fast_plan_destroy(plan->fast);
  if (plan->tile) tile_plan_destroy(plan->tile);
  if (plan->packed) packed_plan_destroy(plan->packed);
  delete plan;
  return BANDCU_OK;
}

int bandcu_plan_is_fast(const BandCuPlan* plan) { return plan && plan->fast ? 1 : 0; }

int bandcu_plan_staging_kind(const BandCuPlan* plan) { return !plan ? 0 : plan->fast ? 1 : plan->tile ? 2 : 0; }

int bandcu_plan_set_host_output(BandCuPlan* plan, int on) {
  if (!plan) return fail(BANDCU_E_INVALID, "plan_set_host_output: null plan");
  if (plan->fast) fast_plan_set_bulk_store(plan->fast, on);
  return BANDCU_OK;
}

// BANDCU_TRACE=1: one stderr line per batch naming the kernel family that ran it (diagnostics only).
static void trace_path(const char* what, const BandCuPlan* plan, int n) {
  static const bool on = getenv("BANDCU_TRACE") && getenv("BANDCU_TRACE")[0] == '1';
  if (on) fprintf(stderr, "[bandcu] %s kernel: %d frame(s), plan fast=%d tile=%d packed=%d\n", what, n, plan->fast != nullptr,
                  plan->tile != nullptr, plan->packed != nullptr);
}

static int run_batch(BandCuPlan* plan, const BatchArgs& b, const TableHints& hints, int n, cudaStream_t s) {
  if (plan->fast) {
    int e = launch_fast(plan->fast, plan->d, b, hints, n, s);
    if (e == BANDCU_OK) { trace_path("row-staged", plan, n); return BANDCU_OK; }
    if (e != BANDCU_E_UNSUPPORTED) return fail(e, "plan_run: fast kernel launch");
    // this batch's pointers/strides do not meet the staged kernel's alignment
    // rules: the generic kernel computes the identical result.
  }
  if (plan->tile) {
    int e = launch_tile(plan->tile, plan->d, b, hints, n, s);
    if (e == BANDCU_OK) { trace_path("tile-staged", plan, n); return BANDCU_OK; }
    if (e != BANDCU_E_UNSUPPORTED) return fail(e, "plan_run: tile kernel launch");
  }
  {  // u8 plane down-scales (Resize of a 4:2:0 buffer, of a gray image): ahead of the packed kernel, which covers gray only
    int e = launch_plane_rows(plan->d, b, hints, n, s);
    if (e == BANDCU_OK) { trace_path("plane-rows", plan, n); return BANDCU_OK; }
    if (e != BANDCU_E_UNSUPPORTED) return fail(e, "plan_run: plane-rows kernel launch");
  }
  if (plan->packed) {
    int e = launch_packed(plan->packed, plan->d, b, hints, n, s);
    if (e == BANDCU_OK) { trace_path("packed", plan, n); return BANDCU_OK; }
    if (e != BANDCU_E_UNSUPPORTED) return fail(e, "plan_run: packed kernel launch");
  }
  {
    int e = launch_copy_rows(plan->d, b, hints, n, s);
    if (e == BANDCU_OK) { trace_path("copy-rows", plan, n); return BANDCU_OK; }
    if (e != BANDCU_E_UNSUPPORTED) return fail(e, "plan_run: copy-rows kernel launch");
  }
  int e = launch_generic(plan->d, b, n, s);
  if (e != 0) return fail(e, "plan_run: generic kernel launch");
  trace_path("generic", plan, n);
  return BANDCU_OK;
}

int bandcu_plan_run_strided(BandCuPlan* plan, const BandCuFrame* first, long long in_stride, long long out_stride,
                            int n, void* stream) {
  if (!plan || !first || n < 0) return fail(BANDCU_E_INVALID, "plan_run_strided: bad arguments");
  if (n == 0) return BANDCU_OK;
  if (!first->p0 || !first->out) return fail(BANDCU_E_INVALID, "plan_run_strided: null plane/output pointer");
  if (plan->d.src_kind == BANDCU_SRC_YUV420 && plan->d.cvt != BANDCU_CVT_YUV2GRAY && (!first->pu || !first->pv))
    return fail(BANDCU_E_INVALID, "plan_run_strided: null chroma pointer");
  BatchArgs b;
  b.first.p0 = (const uint8_t*)first->p0;
  b.first.pu = (const uint8_t*)first->pu;
  b.first.pv = (const uint8_t*)first->pv;
  b.first.out = (uint8_t*)first->out;
  b.in_stride = in_stride;
  b.out_stride = out_stride;
  b.table = nullptr;
  const int e = run_batch(plan, b, TableHints(), n, (cudaStream_t)stream);
  if (e == BANDCU_OK) { plan->last_b = b; plan->last_hints = TableHints(); plan->last_n = n; }
  else plan->last_n = 0;
  return e;
}

int bandcu_plan_run(BandCuPlan* plan, const BandCuFrame* frames, int n, void* stream) {
  if (!plan || !frames || n < 0) return fail(BANDCU_E_INVALID, "plan_run: bad arguments");
  if (n == 0) return BANDCU_OK;
  cudaStream_t s = (cudaStream_t)stream;
  if (n == 1) return bandcu_plan_run_strided(plan, frames, 0, 0, 1, stream);
  // uniform stride? then no table is needed
  {
    long long is = (const char*)frames[1].p0 - (const char*)frames[0].p0;
    long long os = (const char*)frames[1].out - (const char*)frames[0].out;
    bool uniform = true;
    for (int i = 1; i < n && uniform; ++i) {
      uniform = ((const char*)frames[i].p0 - (const char*)frames[0].p0 == is * i) &&
                ((const char*)frames[i].out - (const char*)frames[0].out == os * i) &&
                (!frames[0].pu || (const char*)frames[i].pu - (const char*)frames[0].pu == is * i) &&
                (!frames[0].pv || (const char*)frames[i].pv - (const char*)frames[0].pv == is * i);
    }
    if (uniform) return bandcu_plan_run_strided(plan, frames, is, os, n, stream);
  }
  {  // a table upload cannot be part of a recorded graph (see bandcu_stream_begin_capture)
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(s, &cs) == cudaSuccess && cs != cudaStreamCaptureStatusNone)
      return fail(BANDCU_E_UNSUPPORTED, "plan_run: a pointer-table batch cannot be recorded into a CUDA graph");
  }
  // what the staged kernel needs to know about the table, established on the host
  TableHints hints;
  hints.valid = true;
  hints.semi_planar = frames[0].pu && frames[0].pv;
  hints.aligned16 = true;
  hints.out_aligned16 = true;
  hints.planes_aligned4 = true;
  hints.planes_aligned16 = true;
  for (int i = 0; i < n; ++i) {
    const uintptr_t bits = reinterpret_cast<uintptr_t>(frames[i].p0) | reinterpret_cast<uintptr_t>(frames[i].pu) |
                           reinterpret_cast<uintptr_t>(frames[i].pv);
    if (bits & 3) hints.planes_aligned4 = false;
    if (bits & 15) hints.planes_aligned16 = false;
    if (reinterpret_cast<uintptr_t>(frames[i].p0) & 15) hints.aligned16 = false;
    if (reinterpret_cast<uintptr_t>(frames[i].out) & 15) hints.out_aligned16 = false;
  }
  if (hints.semi_planar) {
    const long long d0 = (const char*)frames[0].pv - (const char*)frames[0].pu;
    hints.semi_planar = (d0 == 1 || d0 == -1);
    hints.vu = (d0 == -1);
    for (int i = 0; i < n && hints.semi_planar; ++i) {
      hints.semi_planar = frames[i].pu && frames[i].pv && ((const char*)frames[i].pv - (const char*)frames[i].pu == d0);
      const void* uv0 = hints.vu ? frames[i].pv : frames[i].pu;
      if ((reinterpret_cast<uintptr_t>(frames[i].p0) & 15) || (reinterpret_cast<uintptr_t>(uv0) & 15))
        hints.aligned16 = false;
    }
  }
  // same pointers as the previous call: the table already in device memory is still right
  const bool same_table = plan->table_valid_n == n && plan->table_host && plan->table_stream == s &&
                          memcmp(plan->table_host, frames, sizeof(BandCuFrame) * n) == 0;
  if (same_table) {
    BatchArgs b;
    memset(&b, 0, sizeof(b));
    b.table = plan->table_dev;
    int e = run_batch(plan, b, hints, n, s);
    if (e != BANDCU_OK) return e;
    // this launch reads the table too: a later rewrite must wait for it, not only for the launch that uploaded it
    CU_TRY(cudaEventRecord(plan->table_free, s), "plan_run: event record");
    plan->table_pending = true;
    plan->last_b = b; plan->last_hints = hints; plan->last_n = n;
    return BANDCU_OK;
  }
  if (plan->table_pending) {  // previous launch may still read the table
    CU_TRY(cudaEventSynchronize(plan->table_free), "plan_run: wait for pointer table");
    plan->table_pending = false;
  }
  plan->table_valid_n = 0;
  if (n > plan->table_cap) {
    if (plan->table_dev) (void)cudaFree(plan->table_dev);
    if (plan->table_host) (void)cudaFreeHost(plan->table_host);
    plan->table_dev = nullptr;
    plan->table_host = nullptr;
    plan->table_cap = 0;
    int cap = n < 64 ? 64 : n;
    CU_TRY(cudaMalloc((void**)&plan->table_dev, sizeof(BandCuFrame) * cap), "plan_run: table alloc");
    CU_TRY(cudaMallocHost((void**)&plan->table_host, sizeof(BandCuFrame) * cap), "plan_run: pinned table alloc");
    plan->table_cap = cap;
  }
  if (!plan->table_free) CU_TRY(cudaEventCreateWithFlags(&plan->table_free, cudaEventDisableTiming), "plan_run: event");
  memcpy(plan->table_host, frames, sizeof(BandCuFrame) * n);
  CU_TRY(cudaMemcpyAsync(plan->table_dev, plan->table_host, sizeof(BandCuFrame) * n, cudaMemcpyHostToDevice, s),
         "plan_run: table upload");
  BatchArgs b;
  memset(&b, 0, sizeof(b));
  b.table = plan->table_dev;
  int e = run_batch(plan, b, hints, n, s);
  if (e != BANDCU_OK) { plan->last_n = 0; return e; }
  CU_TRY(cudaEventRecord(plan->table_free, s), "plan_run: event record");
  plan->table_pending = true;
  plan->table_valid_n = n;
  plan->table_stream = s;
  plan->last_b = b; plan->last_hints = hints; plan->last_n = n;
  return BANDCU_OK;
}

int bandcu_plan_rerun(BandCuPlan* plan, void* stream) {
  if (!plan || plan->last_n <= 0) return fail(BANDCU_E_INVALID, "plan_rerun: no previous batch");
  cudaStream_t s = (cudaStream_t)stream;
  if (plan->last_b.table) {
    // the device table still holds the previous batch only if nothing rewrote it and the stream is the one it was ordered on
    if (plan->table_valid_n != plan->last_n || plan->table_stream != s) return fail(BANDCU_E_INVALID, "plan_rerun: pointer table changed");
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(s, &cs) == cudaSuccess && cs != cudaStreamCaptureStatusNone)
      return fail(BANDCU_E_UNSUPPORTED, "plan_rerun: a pointer-table batch cannot be recorded into a CUDA graph");
  }
  const int e = run_batch(plan, plan->last_b, plan->last_hints, plan->last_n, s);
  if (e != BANDCU_OK) return e;
  if (plan->last_b.table) {
    CU_TRY(cudaEventRecord(plan->table_free, s), "plan_rerun: event record");
    plan->table_pending = true;
  }
  return BANDCU_OK;
}

long long bandcu_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

int bandcu_yuv_map(const BandCuYuvMapDesc* desc, const BandCuYuvPtrs* ptrs, void* stream) {
  if (!desc || !ptrs || !ptrs->sy || !ptrs->su || !ptrs->sv || !ptrs->du || !ptrs->dv)
    return fail(BANDCU_E_INVALID, "yuv_map: null argument");
  if (desc->dst_width <= 0 || desc->dst_height <= 0) return fail(BANDCU_E_INVALID, "yuv_map: bad dims");
  int e = launch_yuv_map(*desc, *ptrs, 0, 0, 1, (cudaStream_t)stream);
  if (e != 0) return fail(e, "yuv_map: launch");
  return BANDCU_OK;
}

int bandcu_yuv_map_strided(const BandCuYuvMapDesc* desc, const BandCuYuvPtrs* first, long long src_frame_stride,
                           long long dst_frame_stride, int n, void* stream) {
  if (!desc || !first || !first->sy || !first->su || !first->sv || !first->du || !first->dv || n < 0)
    return fail(BANDCU_E_INVALID, "yuv_map_strided: bad argument");
  if (desc->dst_width <= 0 || desc->dst_height <= 0) return fail(BANDCU_E_INVALID, "yuv_map_strided: bad dims");
  if (n == 0) return BANDCU_OK;
  int e = launch_yuv_map(*desc, *first, src_frame_stride, dst_frame_stride, n, (cudaStream_t)stream);
  if (e != 0) return fail(e, "yuv_map_strided: launch");
  return BANDCU_OK;
}

int bandcu_rgb_to_yuv420(const BandCuRgbToYuvDesc* desc, const BandCuRgbToYuvPtrs* ptrs, void* stream) {
  if (!desc || !ptrs || !ptrs->src || !ptrs->dy || !ptrs->du || !ptrs->dv)
    return fail(BANDCU_E_INVALID, "rgb_to_yuv420: null argument");
  if (desc->width <= 0 || desc->height <= 0 || (desc->src_channels != 3 && desc->src_channels != 4))
    return fail(BANDCU_E_INVALID, "rgb_to_yuv420: bad descriptor");
  int e = launch_rgb_to_yuv(*desc, *ptrs, 0, 0, 1, (cudaStream_t)stream);
  if (e != 0) return fail(e, "rgb_to_yuv420: launch");
  return BANDCU_OK;
}

int bandcu_rgb_to_yuv420_strided(const BandCuRgbToYuvDesc* desc, const BandCuRgbToYuvPtrs* first, long long src_frame_stride,
                                 long long dst_frame_stride, int n, void* stream) {
  if (!desc || !first || !first->src || !first->dy || !first->du || !first->dv || n < 0)
    return fail(BANDCU_E_INVALID, "rgb_to_yuv420_strided: bad argument");
  if (desc->width <= 0 || desc->height <= 0 || (desc->src_channels != 3 && desc->src_channels != 4))
    return fail(BANDCU_E_INVALID, "rgb_to_yuv420_strided: bad descriptor");
  if (n == 0) return BANDCU_OK;
  int e = launch_rgb_to_yuv(*desc, *first, src_frame_stride, dst_frame_stride, n, (cudaStream_t)stream);
  if (e != 0) return fail(e, "rgb_to_yuv420_strided: launch");
  return BANDCU_OK;
}

int bandcu_normalize(const void* in, int in_dtype, void* out, int out_dtype, long long n, float mean, float std_,
                     void* stream) {
  if (!in || !out || n < 0) return fail(BANDCU_E_INVALID, "normalize: bad arguments");
  if (n == 0) return BANDCU_OK;
  int e = launch_normalize(in, in_dtype, out, out_dtype, n, mean, std_, (cudaStream_t)stream);
  if (e >= 10000) return fail(e, "normalize: unsupported dtype pair");
  if (e != 0) return fail(e, "normalize: launch");
  return BANDCU_OK;
}

int bandcu_device_count(int* count) {
  if (!count) return fail(BANDCU_E_INVALID, "device_count: null");
  CU_TRY(cudaGetDeviceCount(count), "cudaGetDeviceCount");
  return BANDCU_OK;
}
int bandcu_set_device(int device) {
  CU_TRY(cudaSetDevice(device), "cudaSetDevice");
  return BANDCU_OK;
}
int bandcu_malloc(void** ptr, size_t bytes) {
  if (!ptr) return fail(BANDCU_E_INVALID, "malloc: null");
  CU_TRY(cudaMalloc(ptr, bytes ? bytes : 1), "cudaMalloc");
  return BANDCU_OK;
}
int bandcu_free(void* ptr) {
  if (ptr) CU_TRY(cudaFree(ptr), "cudaFree");
  return BANDCU_OK;
}
int bandcu_host_alloc(void** ptr, size_t bytes) {
  if (!ptr) return fail(BANDCU_E_INVALID, "host_alloc: null");
  CU_TRY(cudaMallocHost(ptr, bytes ? bytes : 1), "cudaMallocHost");
  return BANDCU_OK;
}
int bandcu_host_free(void* ptr) {
  if (ptr) CU_TRY(cudaFreeHost(ptr), "cudaFreeHost");
  return BANDCU_OK;
}
int bandcu_memcpy_h2d(void* dst, const void* src, size_t bytes, void* stream) {
  CU_TRY(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, (cudaStream_t)stream), "memcpy h2d");
  return BANDCU_OK;
}
int bandcu_memcpy_d2h(void* dst, const void* src, size_t bytes, void* stream) {
  CU_TRY(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, (cudaStream_t)stream), "memcpy d2h");
  return BANDCU_OK;
}
int bandcu_memcpy_d2d(void* dst, const void* src, size_t bytes, void* stream) {
  CU_TRY(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, (cudaStream_t)stream), "memcpy d2d");
  return BANDCU_OK;
}
int bandcu_memcpy2d_d2h(void* dst, size_t dst_pitch, const void* src, size_t src_pitch, size_t width, size_t height,
                        void* stream) {
  if (width == 0 || height == 0) return BANDCU_OK;
  CU_TRY(cudaMemcpy2DAsync(dst, dst_pitch, src, src_pitch, width, height, cudaMemcpyDeviceToHost, (cudaStream_t)stream),
         "memcpy2d d2h");
  return BANDCU_OK;
}
int bandcu_host_register(void* ptr, size_t bytes) {
  if (!ptr || !bytes) return fail(BANDCU_E_INVALID, "host_register: null range");
  CU_TRY(cudaHostRegister(ptr, bytes, cudaHostRegisterPortable | cudaHostRegisterMapped), "cudaHostRegister");
  return BANDCU_OK;
}
int bandcu_host_unregister(void* ptr) {
  if (ptr) CU_TRY(cudaHostUnregister(ptr), "cudaHostUnregister");
  return BANDCU_OK;
}
int bandcu_registered_host_pointers_usable(void) {
  int dev = 0, v = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&v, cudaDevAttrCanUseHostPointerForRegisteredMem, dev) != cudaSuccess) {
    (void)cudaGetLastError();
    return -1;
  }
  return v ? 1 : 0;
}
int bandcu_memset(void* dst, int value, size_t bytes, void* stream) {
  CU_TRY(cudaMemsetAsync(dst, value, bytes, (cudaStream_t)stream), "memset");
  return BANDCU_OK;
}
int bandcu_stream_create(void** stream) {
  if (!stream) return fail(BANDCU_E_INVALID, "stream_create: null");
  cudaStream_t s;
  CU_TRY(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking), "cudaStreamCreate");
  *stream = (void*)s;
  return BANDCU_OK;
}
int bandcu_stream_destroy(void* stream) {
  if (stream) CU_TRY(cudaStreamDestroy((cudaStream_t)stream), "cudaStreamDestroy");
  return BANDCU_OK;
}
int bandcu_stream_sync(void* stream) {
  CU_TRY(cudaStreamSynchronize((cudaStream_t)stream), "cudaStreamSynchronize");
  return BANDCU_OK;
}
int bandcu_stream_begin_capture(void* stream) {
  if (!stream) return fail(BANDCU_E_INVALID, "begin_capture: the default stream cannot be captured");
  CU_TRY(cudaStreamBeginCapture((cudaStream_t)stream, cudaStreamCaptureModeRelaxed), "cudaStreamBeginCapture");
  return BANDCU_OK;
}
int bandcu_stream_end_capture(void* stream, void** graph_exec) {
  if (!graph_exec) return fail(BANDCU_E_INVALID, "end_capture: null out pointer");
  *graph_exec = nullptr;
  cudaGraph_t graph = nullptr;
  CU_TRY(cudaStreamEndCapture((cudaStream_t)stream, &graph), "cudaStreamEndCapture");
  cudaGraphExec_t exec = nullptr;
  cudaError_t e = cudaGraphInstantiate(&exec, graph, 0);
  (void)cudaGraphDestroy(graph);
  if (e != cudaSuccess) return fail((int)e, "cudaGraphInstantiate");
  *graph_exec = (void*)exec;
  return BANDCU_OK;
}
int bandcu_graph_launch(void* graph_exec, void* stream) {
  if (!graph_exec) return fail(BANDCU_E_INVALID, "graph_launch: null graph");
  CU_TRY(cudaGraphLaunch((cudaGraphExec_t)graph_exec, (cudaStream_t)stream), "cudaGraphLaunch");
  count_launch();
  return BANDCU_OK;
}
int bandcu_graph_destroy(void* graph_exec) {
  if (graph_exec) CU_TRY(cudaGraphExecDestroy((cudaGraphExec_t)graph_exec), "cudaGraphExecDestroy");
  return BANDCU_OK;
}
int bandcu_event_create(void** ev) {
  if (!ev) return fail(BANDCU_E_INVALID, "event_create: null");
  cudaEvent_t e;
  CU_TRY(cudaEventCreate(&e), "cudaEventCreate");
  *ev = (void*)e;
  return BANDCU_OK;
}
int bandcu_event_create_sync(void** ev) {
  if (!ev) return fail(BANDCU_E_INVALID, "event_create_sync: null");
  cudaEvent_t e;
  CU_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming), "cudaEventCreateWithFlags");
  *ev = (void*)e;
  return BANDCU_OK;
}
int bandcu_stream_wait_event(void* stream, void* ev) {
  CU_TRY(cudaStreamWaitEvent((cudaStream_t)stream, (cudaEvent_t)ev, 0), "cudaStreamWaitEvent");
  return BANDCU_OK;
}
int bandcu_event_destroy(void* ev) {
  if (ev) CU_TRY(cudaEventDestroy((cudaEvent_t)ev), "cudaEventDestroy");
  return BANDCU_OK;
}
int bandcu_event_record(void* ev, void* stream) {
  CU_TRY(cudaEventRecord((cudaEvent_t)ev, (cudaStream_t)stream), "cudaEventRecord");
  return BANDCU_OK;
}
int bandcu_event_elapsed_ms(void* start, void* stop, float* ms) {
  if (!ms) return fail(BANDCU_E_INVALID, "event_elapsed: null");
  CU_TRY(cudaEventSynchronize((cudaEvent_t)stop), "cudaEventSynchronize");
  CU_TRY(cudaEventElapsedTime(ms, (cudaEvent_t)start, (cudaEvent_t)stop), "cudaEventElapsedTime");
  return BANDCU_OK;
}
int bandcu_pointer_is_device(const void* ptr) {
  cudaPointerAttributes a;
  cudaError_t e = cudaPointerGetAttributes(&a, ptr);
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    return -1;
  }
  return (a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged) ? 1 : 0;
}
int bandcu_pointer_kind(const void* ptr) {
  cudaPointerAttributes a;
  cudaError_t e = cudaPointerGetAttributes(&a, ptr);
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    return -1;
  }
  if (a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged) return 2;
  return a.type == cudaMemoryTypeHost ? 1 : 0;
}
int bandcu_pointer_device(const void* ptr) {
  cudaPointerAttributes a;
  cudaError_t e = cudaPointerGetAttributes(&a, ptr);
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    return -1;
  }
  return (a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged) ? a.device : -1;
}
const char* bandcu_last_error_string(void) { return g_err; }

}  // extern "C"
