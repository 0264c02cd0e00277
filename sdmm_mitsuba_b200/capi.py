"""ctypes binding of the C ABI in include/sdmm_b200.h (libsdmm_b200.so).

This is the ONLY route from Python to the product: there is no CPU or PyTorch fallback.  If the
library has not been built (`python __graft_entry__.py` / `sdmm_mitsuba_b200._build.build()`),
loading fails loudly.
"""
import ctypes as C
import os

from . import _build

LIB_PATH = _build.LIB_PATH

JOINT_PLANES = 32
EM_PLANES = 81
SCALAR_WORDS = 8
QUERY_RECORD_FLOATS = 32

PCG32_DEFAULT_STATE = 0x853c49e6748fea9b
PCG32_DEFAULT_STREAM = 0xda3e39cb94b95bdb

ST_STEP_ABORTED, ST_CDF_INVALID, ST_NON_PSD, ST_COND_INVALID = 1, 2, 4, 8

# every symbol include/sdmm_b200.h declares (tests check the library exports exactly these)
SYMBOLS = [
    "sdmmb_create", "sdmmb_destroy", "sdmmb_get_error_string", "sdmmb_last_cuda_error",
    "sdmmb_num_components", "sdmmb_max_cells", "sdmmb_upload_cells", "sdmmb_download_cells",
    "sdmmb_download_derived", "sdmmb_set_samples", "sdmmb_set_samples_host", "sdmmb_em_step",
    "sdmmb_optimize_host", "sdmmb_prepare_conditioner", "sdmmb_download_conditioner", "sdmmb_sample_pdf", "sdmmb_pdf",
    "sdmmb_sample_pdf_host", "sdmmb_pdf_host", "sdmmb_set_tree", "sdmmb_find_cells", "sdmmb_find_cells_host", "sdmmb_collect_samples",
    "sdmmb_set_min_samples", "sdmmb_set_async_records", "sdmmb_publish_records", "sdmmb_query_records", "sdmmb_allgather_records",
    "sdmmb_launch_count", "sdmmb_last_estep_ms", "sdmmb_measure_peaks", "sdmmb_initialize", "sdmmb_initialize_host",
    "sdmmb_copy_cells", "sdmmb_sample_pdf_product", "sdmmb_pdf_product", "sdmmb_pt_create", "sdmmb_pt_destroy", "sdmmb_pt_render",
    "sdmmb_set_samples_borrowed", "sdmmb_set_em_overlap", "sdmmb_peer_export", "sdmmb_peer_attach", "sdmmb_peer_detach", "sdmmb_peer_wait", "sdmmb_peer_status",
    "sdmmb_host_alloc", "sdmmb_host_free", "sdmmb_sample_pdf_pruned", "sdmmb_pt_set_product",
    "sdmmb_sample_pdf_noalloc", "sdmmb_pt_pool_stats", "sdmmb_pt_rebucket",
]
NOALLOC_MAX_VERTICES = 65536   # SDMMB_NOALLOC_MAX_VERTICES
NOALLOC_MAX_CELLS = 65536      # SDMMB_NOALLOC_MAX_CELLS
E_LIMIT = -10
PEER_HANDLE_BYTES = 192  # SDMMB_PEER_HANDLE_BYTES
ITERS_INTEGRATOR = -1    # SDMMB_ITERS_INTEGRATOR: the integrator's 2|1 iteration rule per cell
E_PEER = -9



class Quad(C.Structure):   # sdmmb_quad (include/sdmm_b200_render.h)
    _fields_ = [("p0", C.c_float * 3), ("e1", C.c_float * 3), ("e2", C.c_float * 3),
                ("reflectance", C.c_float * 3), ("emission", C.c_float * 3)]


class Camera(C.Structure):   # sdmmb_camera
    _fields_ = [("origin", C.c_float * 3), ("right", C.c_float * 3), ("up", C.c_float * 3),
                ("forward", C.c_float * 3), ("fov_x_degrees", C.c_float), ("width", C.c_int), ("height", C.c_int)]


class PtConfig(C.Structure):   # sdmmb_pt_config
    _fields_ = [("aabb_min", C.c_float * 3), ("spatial_norm", C.c_float), ("root_min", C.c_float * 3),
                ("root_max", C.c_float * 3), ("max_depth", C.c_int), ("saved_per_path", C.c_int)]


class PtStats(C.Structure):   # sdmmb_pt_stats
    _fields_ = [("paths", C.c_uint64), ("guided_queries", C.c_uint64), ("mis_queries", C.c_uint64),
                ("pool_count", C.c_uint64), ("pool_overflow", C.c_uint64), ("launches", C.c_uint64)]


_lib = None


class SdmmError(RuntimeError):
    def __init__(self, code, what, detail=""):
        self.code = code
        super().__init__("%s failed: %s (%d)%s" % (what, _errstr(code), code, (": " + detail) if detail else ""))


def _errstr(code):
    try:
        return lib().sdmmb_get_error_string(code).decode()
    except Exception:
        return "?"


def lib():
    """Load libsdmm_b200.so; raises if the CUDA extension has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            "sdmm_mitsuba_b200: %s is missing -- build the CUDA extension first "
            "(python -c 'import __graft_entry__ as g; g.build()'); there is no CPU fallback" % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    vp, fp, up, bp = C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p  # raw addresses (host or device)
    i, sz = C.c_int, C.c_size_t
    L.sdmmb_create.argtypes = [C.POINTER(vp), i, i, i]
    L.sdmmb_destroy.argtypes = [vp]
    L.sdmmb_get_error_string.argtypes = [i]
    L.sdmmb_get_error_string.restype = C.c_char_p
    L.sdmmb_last_cuda_error.argtypes = [vp]
    L.sdmmb_last_cuda_error.restype = C.c_char_p
    L.sdmmb_num_components.argtypes = [vp]
    L.sdmmb_max_cells.argtypes = [vp]
    L.sdmmb_upload_cells.argtypes = [vp, i, i, fp, fp, up, vp]
    L.sdmmb_download_cells.argtypes = [vp, i, i, fp, fp, up, vp]
    L.sdmmb_download_derived.argtypes = [vp, i, i, fp, vp]
    L.sdmmb_set_samples.argtypes = [vp, i, i, fp, sz, fp, up, vp]
    L.sdmmb_set_samples_borrowed.argtypes = [vp, i, i, fp, sz, fp, up, vp]
    L.sdmmb_set_samples_host.argtypes = [vp, i, i, fp, sz, fp, up, vp]
    L.sdmmb_em_step.argtypes = [vp, i, i, i, i, fp, vp]
    L.sdmmb_optimize_host.argtypes = [vp, i, i, fp, sz, fp, up, i, i, fp, fp, up, vp]
    L.sdmmb_prepare_conditioner.argtypes = [vp, i, i, vp]
    L.sdmmb_download_conditioner.argtypes = [vp, i, i, fp, vp]
    L.sdmmb_sample_pdf.argtypes = [vp, sz, up, fp, fp, fp, fp, fp, up, bp, i, vp]
    L.sdmmb_pdf.argtypes = [vp, sz, up, fp, fp, fp, bp, i, vp]
    L.sdmmb_sample_pdf_noalloc.argtypes = [vp, sz, vp, vp, vp, vp, vp, vp, vp, vp, vp]
    L.sdmmb_sample_pdf_host.argtypes = [vp, sz, up, fp, fp, fp, fp, fp, up, bp, i, vp]
    L.sdmmb_pdf_host.argtypes = [vp, sz, up, fp, fp, fp, bp, i, vp]
    L.sdmmb_set_tree.argtypes = [vp, i, vp]
    L.sdmmb_find_cells.argtypes = [vp, sz, fp, up, vp]
    L.sdmmb_find_cells_host.argtypes = [vp, sz, fp, up, vp]
    L.sdmmb_collect_samples.argtypes = [vp, sz, fp, sz, fp, fp, sz, up, C.POINTER(C.c_uint64), vp]
    L.sdmmb_set_min_samples.argtypes = [vp, C.c_uint32]
    L.sdmmb_set_async_records.argtypes = [vp, i]
    L.sdmmb_publish_records.argtypes = [vp, i, i, vp]
    L.sdmmb_query_records.argtypes = [vp, C.POINTER(vp), C.POINTER(sz)]
    L.sdmmb_allgather_records.argtypes = [vp, vp, i, vp]
    L.sdmmb_set_em_overlap.argtypes = [vp, i]
    L.sdmmb_peer_export.argtypes = [vp, vp]
    L.sdmmb_peer_attach.argtypes = [vp, i, i, vp]
    L.sdmmb_peer_detach.argtypes = [vp]
    L.sdmmb_peer_wait.argtypes = [vp, vp]
    L.sdmmb_peer_status.argtypes = [vp, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), vp]
    L.sdmmb_launch_count.argtypes = [vp, C.POINTER(C.c_uint64)]
    L.sdmmb_last_estep_ms.argtypes = [vp, C.POINTER(C.c_float), C.POINTER(i)]
    L.sdmmb_measure_peaks.argtypes = [i, C.POINTER(C.c_float)]
    L.sdmmb_initialize.argtypes = [vp, i, i, fp, fp, sz, fp, up, fp, C.c_uint64, C.c_uint64, vp, vp]
    L.sdmmb_initialize_host.argtypes = [vp, i, i, fp, fp, sz, fp, up, fp, C.c_uint64, C.c_uint64, vp, vp]
    L.sdmmb_pt_create.argtypes = [vp, C.POINTER(vp), C.POINTER(Quad), i, C.POINTER(Camera), C.POINTER(PtConfig)]
    L.sdmmb_pt_destroy.argtypes = [vp]
    L.sdmmb_copy_cells.argtypes = [vp, i, vp, vp, vp]
    L.sdmmb_sample_pdf_pruned.argtypes = [vp, sz, vp, vp, vp, vp, i, i, vp, vp, vp, vp, vp, vp]
    L.sdmmb_host_alloc.argtypes = [C.POINTER(vp), sz]
    L.sdmmb_host_free.argtypes = [vp]
    L.sdmmb_sample_pdf_product.argtypes = [vp, sz, vp, vp, vp, vp, vp, vp, i, fp, vp, vp, vp, vp, vp, vp]
    L.sdmmb_pdf_product.argtypes = [vp, sz, vp, vp, vp, vp, vp, vp, i, fp, vp, vp, vp]
    L.sdmmb_pt_set_product.argtypes = [vp, i, fp]
    L.sdmmb_pt_pool_stats.argtypes = [vp, vp, sz, sz, sz, vp, vp, vp]
    L.sdmmb_pt_rebucket.argtypes = [vp, vp, sz, sz, vp, vp, vp, vp, vp, vp]
    L.sdmmb_pt_render.argtypes = [vp, i, C.c_uint64, i, i, vp, fp, fp, sz, sz, C.POINTER(PtStats), vp]
    for name in SYMBOLS:
        if name not in ("sdmmb_get_error_string", "sdmmb_last_cuda_error"):
            getattr(L, name).restype = i
    _lib = L
    return L


def check(rc, what, ctx=None):
    if rc != 0:
        detail = ""
        if ctx is not None and rc in (-2, -3):
            detail = lib().sdmmb_last_cuda_error(ctx).decode()
        raise SdmmError(rc, what, detail)
