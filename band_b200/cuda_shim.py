"""ctypes binding of the CUDA shim's C-ABI (include/bandcu.h -> lib/libbandcu.so).

Thin and literal: one Python callable per exported symbol, POD structs mirrored
field for field. There is no fallback: if the library is missing or no CUDA
device is present, calls raise.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libbandcu.so")

SRC_GRAY, SRC_RGB, SRC_RGBA, SRC_YUV420, SRC_UV = 0, 1, 2, 3, 4
CVT_NONE, CVT_YUV2RGB, CVT_YUV2RGBA, CVT_YUV2GRAY, CVT_RGB2GRAY, CVT_RGBA2GRAY, CVT_RGBA2RGB = range(7)
FLAVOR_ARGB, FLAVOR_PLANE = 0, 1
(SCALE_COPY, SCALE_POINT, SCALE_BILINEAR_DOWN, SCALE_BILINEAR_UP, SCALE_BOX2, SCALE_VERTICAL, SCALE_DOWN34,
 SCALE_DOWN38, SCALE_UP2_LINEAR, SCALE_UP2_BILINEAR) = range(10)
DT_F32, DT_I32, DT_U8, DT_I16, DT_I8, DT_F64 = 1, 2, 3, 7, 9, 11


class Affine(C.Structure):
    _fields_ = [("r0", C.c_int), ("rr", C.c_int), ("rc", C.c_int), ("c0", C.c_int), ("cr", C.c_int), ("cc", C.c_int)]

    @staticmethod
    def identity():
        return Affine(0, 1, 0, 0, 0, 1)


class PlanDesc(C.Structure):
    _fields_ = [
        ("src_kind", C.c_int), ("stride0", C.c_int), ("stride_uv", C.c_int), ("pixel_stride_uv", C.c_int),
        ("vw", C.c_int), ("vh", C.c_int),
        ("luma_map", Affine), ("pre_map", Affine), ("chroma_map", Affine),
        ("cvt", C.c_int),
        ("flavor", C.c_int), ("scale_mode", C.c_int), ("linear", C.c_int),
        ("x0", C.c_int), ("y0", C.c_int), ("dx", C.c_int), ("dy", C.c_int), ("max_y", C.c_int),
        ("dst_width", C.c_int), ("dst_height", C.c_int),
        ("channels", C.c_int), ("out_dtype", C.c_int), ("normalize", C.c_int),
        ("mean", C.c_float), ("std_", C.c_float),
        ("dst_row_stride", C.c_longlong),
        ("src_width", C.c_int), ("src_height", C.c_int),
    ]


class Frame(C.Structure):
    _fields_ = [("p0", C.c_void_p), ("pu", C.c_void_p), ("pv", C.c_void_p), ("out", C.c_void_p)]


class YuvMapDesc(C.Structure):
    _fields_ = [
        ("src_stride_y", C.c_int), ("src_stride_uv", C.c_int), ("src_pixel_stride_uv", C.c_int),
        ("dst_stride_y", C.c_int), ("dst_stride_uv", C.c_int), ("dst_pixel_stride_uv", C.c_int),
        ("dst_width", C.c_int), ("dst_height", C.c_int), ("chroma_width_override", C.c_int),
        ("luma_map", Affine), ("chroma_map", Affine),
    ]


class YuvPtrs(C.Structure):
    _fields_ = [("sy", C.c_void_p), ("su", C.c_void_p), ("sv", C.c_void_p), ("dy", C.c_void_p), ("du", C.c_void_p),
                ("dv", C.c_void_p)]


class RgbToYuvDesc(C.Structure):
    _fields_ = [("src_channels", C.c_int), ("src_stride", C.c_int), ("src_map", Affine), ("width", C.c_int),
                ("height", C.c_int), ("dst_stride_y", C.c_int), ("dst_stride_uv", C.c_int),
                ("dst_pixel_stride_uv", C.c_int)]


class RgbToYuvPtrs(C.Structure):
    _fields_ = [("src", C.c_void_p), ("dy", C.c_void_p), ("du", C.c_void_p), ("dv", C.c_void_p)]


class BandCuError(RuntimeError):
    pass


_lib = None


def lib():
    """Load libbandcu.so (built in-tree by __graft_entry__.build()). Fails loudly."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise BandCuError("%s is missing: run `python -c 'import __graft_entry__ as g; g.build()'`" % LIB_PATH)
    L = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    vp, ci, ll = C.c_void_p, C.c_int, C.c_longlong
    L.bandcu_plan_create.argtypes = [C.POINTER(PlanDesc), C.POINTER(vp)]
    L.bandcu_plan_destroy.argtypes = [vp]
    L.bandcu_plan_is_fast.argtypes = [vp]
    L.bandcu_plan_run_strided.argtypes = [vp, C.POINTER(Frame), ll, ll, ci, vp]
    L.bandcu_plan_run.argtypes = [vp, C.POINTER(Frame), ci, vp]
    L.bandcu_launch_count.restype = ll
    L.bandcu_yuv_map.argtypes = [C.POINTER(YuvMapDesc), C.POINTER(YuvPtrs), vp]
    L.bandcu_rgb_to_yuv420.argtypes = [C.POINTER(RgbToYuvDesc), C.POINTER(RgbToYuvPtrs), vp]
    L.bandcu_yuv_map_strided.argtypes = [C.POINTER(YuvMapDesc), C.POINTER(YuvPtrs), C.c_longlong, C.c_longlong, C.c_int, vp]
    L.bandcu_rgb_to_yuv420_strided.argtypes = [C.POINTER(RgbToYuvDesc), C.POINTER(RgbToYuvPtrs), C.c_longlong, C.c_longlong,
                                               C.c_int, vp]
    L.bandcu_normalize.argtypes = [vp, ci, vp, ci, ll, C.c_float, C.c_float, vp]
    L.bandcu_device_count.argtypes = [C.POINTER(ci)]
    L.bandcu_set_device.argtypes = [ci]
    L.bandcu_malloc.argtypes = [C.POINTER(vp), C.c_size_t]
    L.bandcu_free.argtypes = [vp]
    L.bandcu_host_alloc.argtypes = [C.POINTER(vp), C.c_size_t]
    L.bandcu_host_free.argtypes = [vp]
    for name in ("bandcu_memcpy_h2d", "bandcu_memcpy_d2h", "bandcu_memcpy_d2d"):
        getattr(L, name).argtypes = [vp, vp, C.c_size_t, vp]
    L.bandcu_memset.argtypes = [vp, ci, C.c_size_t, vp]
    L.bandcu_memcpy2d_d2h.argtypes = [vp, C.c_size_t, vp, C.c_size_t, C.c_size_t, C.c_size_t, vp]
    L.bandcu_host_register.argtypes = [vp, C.c_size_t]
    L.bandcu_host_unregister.argtypes = [vp]
    L.bandcu_stream_create.argtypes = [C.POINTER(vp)]
    L.bandcu_stream_destroy.argtypes = [vp]
    L.bandcu_stream_sync.argtypes = [vp]
    L.bandcu_event_create.argtypes = [C.POINTER(vp)]
    L.bandcu_event_create_sync.argtypes = [C.POINTER(vp)]
    L.bandcu_stream_wait_event.argtypes = [vp, vp]
    L.bandcu_event_destroy.argtypes = [vp]
    L.bandcu_event_record.argtypes = [vp, vp]
    L.bandcu_event_elapsed_ms.argtypes = [vp, vp, C.POINTER(C.c_float)]
    L.bandcu_pointer_is_device.argtypes = [vp]
    L.bandcu_pointer_kind.argtypes = [vp]
    L.bandcu_pointer_device.argtypes = [vp]
    L.bandcu_plan_rerun.argtypes = [vp, vp]
    L.bandcu_two_stage_run_noalloc.argtypes = [C.POINTER(PlanDesc), C.POINTER(PlanDesc), C.POINTER(Frame), vp]
    L.bandcu_last_error_string.restype = C.c_char_p
    _lib = L
    return L


NOALLOC_SCRATCH_BYTES = 8 << 20   # BANDCU_NOALLOC_SCRATCH_BYTES (include/bandcu.h)


def check(code, what=""):
    if code != 0:
        raise BandCuError("%s failed (%d): %s" % (what, code, lib().bandcu_last_error_string().decode()))


EXPORTED = [
    "bandcu_plan_create", "bandcu_plan_destroy", "bandcu_plan_is_fast", "bandcu_plan_set_host_output", "bandcu_plan_staging_kind", "bandcu_stream_begin_capture", "bandcu_stream_end_capture",
    "bandcu_graph_launch", "bandcu_graph_destroy", "bandcu_plan_run_strided", "bandcu_plan_run",
    "bandcu_launch_count", "bandcu_yuv_map", "bandcu_yuv_map_strided", "bandcu_rgb_to_yuv420", "bandcu_rgb_to_yuv420_strided", "bandcu_normalize", "bandcu_device_count", "bandcu_set_device",
    "bandcu_malloc", "bandcu_free", "bandcu_host_alloc", "bandcu_host_free", "bandcu_memcpy_h2d", "bandcu_memcpy_d2h",
    "bandcu_memcpy_d2d", "bandcu_memset", "bandcu_stream_create", "bandcu_stream_destroy", "bandcu_stream_sync",
    "bandcu_event_create", "bandcu_event_create_sync", "bandcu_stream_wait_event", "bandcu_event_destroy", "bandcu_event_record", "bandcu_event_elapsed_ms",
    "bandcu_pointer_is_device", "bandcu_pointer_kind", "bandcu_last_error_string",
    "bandcu_pointer_device", "bandcu_plan_rerun", "bandcu_memcpy2d_d2h", "bandcu_host_register", "bandcu_host_unregister", "bandcu_registered_host_pointers_usable",
    "bandcu_two_stage_run_noalloc",
]
