"""ctypes binding of libgdpt.so (include/gdpt.h).  No compute happens in Python and there is
no fallback: if the CUDA library is missing, importing anything that needs it raises."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libgdpt.so")

PRECISION_F64 = 0
PRECISION_F32 = 1
OFFSETS_REFERENCE = 0
OFFSETS_SHIFT = 1
CREATE_NO_BVH = 1
CREATE_BVH_HOST = 2
CREATE_BVH_DEVICE = 4
DEFAULT_SEED = 0x31E241F862A1FB5E

dp = C.POINTER(C.c_double)
ip = C.POINTER(C.c_int32)


class GdptImage(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("channels", C.c_int32),
                ("data", dp), ("palette", dp), ("palette_size", C.c_int32), ("index", C.POINTER(C.c_uint8))]


class GdptSceneDesc(C.Structure):
    _fields_ = [
        ("sample_to_cam", C.c_double * 16), ("cam_to_world", C.c_double * 16),
        ("width", C.c_int32), ("height", C.c_int32), ("filter_kind", C.c_int32), ("filter_param", C.c_double),
        ("n_shapes", C.c_int32), ("shapes", ip), ("shape_offsets", C.POINTER(C.c_int64)), ("spheres", dp),
        ("n_positions", C.c_int64), ("n_triangles", C.c_int64), ("n_normals", C.c_int64), ("n_uvs", C.c_int64),
        ("positions", dp), ("indices", ip), ("normals", dp), ("uvs", dp),
        ("n_materials", C.c_int32), ("materials", dp), ("n_textures", C.c_int32), ("textures", dp),
        ("n_images3", C.c_int32), ("n_images1", C.c_int32), ("images3", C.POINTER(GdptImage)), ("images1", C.POINTER(GdptImage)),
        ("n_lights", C.c_int32), ("lights", dp),
        ("max_depth", C.c_int32), ("rr_depth", C.c_int32), ("samples_per_pixel", C.c_int32),
        ("envmap_to_world", dp), ("envmap_to_local", dp),
    ]


class GdptRenderParams(C.Structure):
    _fields_ = [("spp", C.c_int32), ("sample0", C.c_int32), ("max_depth", C.c_int32), ("rr_depth", C.c_int32),
                ("seed", C.c_uint64), ("data_cost", C.c_double), ("precision", C.c_int32),
                ("y0", C.c_int32), ("y1", C.c_int32), ("wave_samples", C.c_int32), ("offset_mode", C.c_int32),
                ("strip_count", C.c_int32), ("strip_phase", C.c_int32)]


class GdptL1Params(C.Structure):
    _fields_ = [("irls_iterations", C.c_int32), ("cg_iterations", C.c_int32), ("start", C.c_int32), ("reweight_primal", C.c_int32),
                ("alpha", C.c_double), ("epsilon", C.c_double), ("cg_tolerance", C.c_double)]


class GdptImages(C.Structure):
    _fields_ = [("primal", dp), ("dx", dp), ("dy", dp), ("poisson", dp)]


class GdptStats(C.Structure):
    _fields_ = [("ms_total", C.c_double), ("ms_trace", C.c_double), ("ms_shade", C.c_double), ("ms_poisson", C.c_double),
                ("ms_copy", C.c_double), ("kernel_launches", C.c_int64), ("rays_traced", C.c_int64), ("paths", C.c_int64),
                ("bounce_iterations", C.c_int64), ("tail_paths", C.c_int64)]


class GdptDistGeom(C.Structure):
    _fields_ = [("w", C.c_int32), ("h", C.c_int32), ("n_ranks", C.c_int32), ("rank", C.c_int32),
                ("y0", C.c_int32), ("y1", C.c_int32), ("band_w", C.c_int32)]


# every symbol include/gdpt.h declares
L1_START_FOURIER, L1_START_PRIMAL, L1_START_OUT = 0, 1, 2

EXPORTS = ["gdpt_last_error", "gdpt_device_count", "gdpt_create", "gdpt_create_ex", "gdpt_bvh_info", "gdpt_bvh_adopt", "gdpt_destroy", "gdpt_scene_bounds", "gdpt_render",
           "gdpt_render_accumulate_device", "gdpt_merge_device", "gdpt_poisson", "gdpt_poisson_device", "gdpt_get_stats",
           "gdpt_synchronize", "gdpt_set_stream", "gdpt_traversal_counts", "gdpt_measure_l2_read", "gdpt_sample_primary", "gdpt_intersect", "gdpt_bsdf", "gdpt_grad_samples", "gdpt_pcg_fill",
           "gdpt_render_path", "gdpt_render_path_accumulate_device", "gdpt_path_samples", "gdpt_sample_light", "gdpt_occluded", "gdpt_render_aux",
           "gdpt_merge_rows_device", "gdpt_poisson_dist_rows_forward", "gdpt_poisson_dist_columns", "gdpt_poisson_dist_rows_inverse",
           "gdpt_poisson_device_noalloc", "gdpt_set_option", "gdpt_host_alloc", "gdpt_host_free",
           "gdpt_multi_create", "gdpt_multi_render", "gdpt_multi_get_stats", "gdpt_multi_device_count", "gdpt_multi_destroy",
           "gdpt_smallgdpt_render", "gdpt_poisson_l1", "gdpt_poisson_l1_device"]

_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("libgdpt.so is not built (run __graft_entry__.build()); there is no CPU fallback")
        L = C.CDLL(LIB_PATH)
        vp = C.c_void_p
        L.gdpt_last_error.restype = C.c_char_p
        L.gdpt_create.argtypes = [C.POINTER(GdptSceneDesc), C.c_int, C.POINTER(vp)]
        L.gdpt_create_ex.argtypes = [C.POINTER(GdptSceneDesc), C.c_int, C.c_int, C.POINTER(vp)]
        L.gdpt_bvh_info.argtypes = [vp, C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int32), C.POINTER(vp), C.POINTER(vp)]
        L.gdpt_bvh_adopt.argtypes = [vp, vp, C.c_int64, vp, C.c_int64, C.c_int32]
        L.gdpt_destroy.argtypes = [vp]
        L.gdpt_destroy.restype = None
        L.gdpt_scene_bounds.argtypes = [vp, dp]
        L.gdpt_render.argtypes = [vp, C.POINTER(GdptRenderParams), C.POINTER(GdptImages)]
        L.gdpt_render_accumulate_device.argtypes = [vp, C.POINTER(GdptRenderParams), vp, vp, vp, vp, vp]
        L.gdpt_merge_device.argtypes = [vp, C.c_int, C.c_int, vp, vp, vp, vp, vp, vp]
        L.gdpt_poisson.argtypes = [C.c_int, C.c_int, C.c_int, dp, dp, dp, C.c_double, dp]
        L.gdpt_poisson_device.argtypes = [vp, C.c_int, C.c_int, vp, vp, vp, C.c_double, vp]
        L.gdpt_poisson_device_noalloc.argtypes = [C.c_int, C.c_int, C.c_int, vp, vp, vp, C.c_double, vp, vp]
        L.gdpt_merge_rows_device.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, vp, vp, vp, vp, vp, vp]
        L.gdpt_poisson_dist_rows_forward.argtypes = [vp, C.POINTER(GdptDistGeom), vp, vp, vp, C.c_double, vp, vp]
        L.gdpt_poisson_dist_columns.argtypes = [vp, C.POINTER(GdptDistGeom), C.c_double, vp, vp, vp]
        L.gdpt_poisson_dist_rows_inverse.argtypes = [vp, C.POINTER(GdptDistGeom), vp, vp]
        L.gdpt_get_stats.argtypes = [vp, C.POINTER(GdptStats)]
        L.gdpt_synchronize.argtypes = [vp]
        L.gdpt_set_option.argtypes = [vp, C.c_char_p, C.c_int]
        L.gdpt_multi_create.argtypes = [C.POINTER(GdptSceneDesc), C.c_int, ip, C.POINTER(vp)]
        L.gdpt_multi_render.argtypes = [vp, C.POINTER(GdptRenderParams), C.POINTER(GdptImages)]
        L.gdpt_multi_get_stats.argtypes = [vp, C.POINTER(GdptStats)]
        L.gdpt_multi_device_count.argtypes = [vp]
        L.gdpt_multi_destroy.argtypes = [vp]
        L.gdpt_multi_destroy.restype = None
        L.gdpt_smallgdpt_render.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(GdptImages), dp]
        L.gdpt_poisson_l1.argtypes = [C.c_int, C.c_int, C.c_int, dp, dp, dp, C.POINTER(GdptL1Params), dp, dp]
        L.gdpt_poisson_l1_device.argtypes = [vp, C.c_int, C.c_int, vp, vp, vp, C.POINTER(GdptL1Params), vp, dp]
        L.gdpt_host_alloc.argtypes = [C.c_size_t, C.POINTER(vp)]
        L.gdpt_host_free.argtypes = [vp]
        L.gdpt_host_free.restype = None
        L.gdpt_set_stream.argtypes = [vp, vp]
        L.gdpt_traversal_counts.argtypes = [vp, C.POINTER(GdptRenderParams), dp]
        L.gdpt_measure_l2_read.argtypes = [C.c_int, C.c_int, C.c_int, dp]
        L.gdpt_sample_primary.argtypes = [vp, C.c_int, C.c_int, dp, dp]
        L.gdpt_intersect.argtypes = [vp, C.c_int, C.c_int, dp, dp, ip, dp]
        L.gdpt_bsdf.argtypes = [vp, C.c_int, C.c_int, C.c_int, dp, dp, dp, dp]
        L.gdpt_grad_samples.argtypes = [vp, C.POINTER(GdptRenderParams), C.c_int, ip, ip, C.POINTER(C.c_uint64), dp]
        L.gdpt_render_path.argtypes = [vp, C.POINTER(GdptRenderParams), dp]
        L.gdpt_render_path_accumulate_device.argtypes = [vp, C.POINTER(GdptRenderParams), vp]
        L.gdpt_path_samples.argtypes = [vp, C.POINTER(GdptRenderParams), C.c_int, ip, ip, C.POINTER(C.c_uint64), dp]
        L.gdpt_sample_light.argtypes = [vp, C.c_int, C.c_int, dp, dp, dp]
        L.gdpt_occluded.argtypes = [vp, C.c_int, C.c_int, dp, ip]
        L.gdpt_render_aux.argtypes = [vp, C.c_int, C.c_int, dp]
        L.gdpt_pcg_fill.argtypes = [C.c_int, C.c_uint64, C.c_uint64, C.c_int, dp, C.POINTER(C.c_uint32)]
        _lib = L
    return _lib


class GdptError(RuntimeError):
    pass


def check(rc):
    if rc != 0:
        raise GdptError("libgdpt error %d: %s" % (rc, lib().gdpt_last_error().decode(errors="replace")))


def as_dp(a):
    return a.ctypes.data_as(dp) if a is not None else None


def f64(a, shape=None):
    a = np.ascontiguousarray(a, dtype=np.float64)
    return a.reshape(shape) if shape is not None else a


# ---- page-locked host arrays for image outputs (gdpt_host_alloc), recycled through a small pool: pinning 50 MB costs more
# than the copy it speeds up, so blocks go back to the pool when the array that wraps them is collected
_pinned_free = {}


def _pinned_return(ptr, nbytes):
    _pinned_free.setdefault(nbytes, []).append(ptr)


def pinned_empty(shape, dtype=np.float64):
    """An uninitialised page-locked numpy array; falls back to a pageable one when no device is usable."""
    import weakref
    shape = tuple(int(v) for v in shape)
    nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
    free = _pinned_free.get(nbytes)
    if free:
        ptr = free.pop()
    else:
        h = C.c_void_p()
        if nbytes == 0 or lib().gdpt_host_alloc(nbytes, C.byref(h)) != 0 or not h.value:
            return np.empty(shape, dtype=dtype)
        ptr = h.value
    buf = (C.c_char * nbytes).from_address(ptr)
    arr = np.frombuffer(buf, dtype=dtype).reshape(shape)
    weakref.finalize(buf, _pinned_return, ptr, nbytes)
    return arr
