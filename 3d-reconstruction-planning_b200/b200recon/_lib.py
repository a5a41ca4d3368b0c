"""ctypes binding of libb200recon.so (C ABI: include/b200recon.h).

There is NO CPU fallback: if the CUDA library is missing or no B200 is visible, the first call that
needs the device raises.  Loading the library itself does not need a GPU (the `-m "not gpu"` tests
check that every symbol of the header is exported).
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B2R_LIB") or os.path.join(os.path.dirname(_HERE), "lib", "libb200recon.so")

c_vp = C.c_void_p
c_dp = C.POINTER(C.c_double)
c_i64p = C.POINTER(C.c_int64)


class Intrinsic(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("fx", C.c_double), ("fy", C.c_double),
                ("cx", C.c_double), ("cy", C.c_double)]


class Volume(C.Structure):
    _fields_ = [("resolution", C.c_int32), ("x_begin", C.c_int32), ("x_end", C.c_int32), ("color_type", C.c_int32),
                ("voxel_length", C.c_double), ("sdf_trunc", C.c_double), ("origin", C.c_double * 3),
                ("d_tsdfw", c_vp), ("d_color", c_vp), ("d_rows", c_vp)]


class IcpParams(C.Structure):
    _fields_ = [("max_correspondence_distance", C.c_double), ("relative_fitness", C.c_double),
                ("relative_rmse", C.c_double), ("max_iteration", C.c_int32), ("estimation", C.c_int32),
                ("lambda_geometric", C.c_double)]


class IcpResult(C.Structure):
    _fields_ = [("transformation", C.c_double * 16), ("fitness", C.c_double), ("inlier_rmse", C.c_double),
                ("n_correspondences", C.c_int64), ("iterations", C.c_int32), ("converged", C.c_int32)]


class Scene(C.Structure):
    _fields_ = [("c2w", C.c_float * 12), ("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float),
                ("W", C.c_int32), ("H", C.c_int32), ("ell_c", C.c_float * 3), ("ell_r", C.c_float * 3),
                ("plane_z", C.c_float), ("has_plane", C.c_int32), ("checker", C.c_float), ("depth_scale", C.c_float),
                ("n_tris", C.c_int32), ("mesh_min", C.c_float * 3), ("mesh_max", C.c_float * 3)]


SEARCH_KNN = 0
SEARCH_HYBRID = 2
PAD_LO = 1
PAD_HI = 2

# name -> (restype, argtypes); must list every function declared in include/b200recon.h
SIGNATURES = {
    "b2r_last_error": (C.c_char_p, []),
    "b2r_version": (C.c_int, []),
    "b2r_ctx_create": (C.c_int, [C.c_int, C.POINTER(c_vp)]),
    "b2r_ctx_destroy": (C.c_int, [c_vp]),
    "b2r_ctx_launch_count": (C.c_int, [c_vp, c_i64p]),
    "b2r_ctx_sync": (C.c_int, [c_vp, c_vp]),
    "b2r_ctx_join": (C.c_int, [c_vp, c_vp]),
    "b2r_ctx_updates_mirror": (C.c_int, [c_vp, C.POINTER(C.c_uint64)]),
    "b2r_ctx_kernel_timing": (C.c_int, [c_vp, C.c_int]),
    "b2r_ctx_kernel_time": (C.c_int, [c_vp, c_dp, c_i64p]),
    "b2r_ctx_trace": (C.c_int, [c_vp, C.c_int]),
    "b2r_ctx_trace_read": (C.c_int, [c_vp, C.POINTER(C.c_uint64), C.c_int, C.POINTER(C.c_int)]),
    "b2r_depth_to_float": (C.c_int, [c_vp, c_vp, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, c_vp, c_vp]),
    "b2r_backproject": (C.c_int, [c_vp, c_vp, c_vp, C.c_int, C.POINTER(Intrinsic), c_dp, c_vp, c_vp, c_i64p, c_vp]),
    "b2r_backproject_static": (C.c_int, [C.c_int, c_vp, c_vp, C.c_int, C.POINTER(Intrinsic), c_dp, c_vp, c_vp, c_vp, c_vp]),
    "b2r_backproject_static_max_pixels": (C.c_int, [c_i64p]),
    "b2r_integrate": (C.c_int, [c_vp, C.POINTER(Volume), c_vp, c_vp, C.POINTER(Intrinsic), c_dp, c_vp, c_vp]),
    "b2r_integrate_host": (C.c_int, [c_vp, C.POINTER(Volume), c_vp, C.c_int, c_vp, C.POINTER(Intrinsic), c_dp,
                                     C.c_double, C.c_double, c_vp, c_vp]),
    "b2r_integrate_host_staged": (C.c_int, [c_vp, C.POINTER(Volume), c_vp, C.c_int, c_vp, C.POINTER(Intrinsic), c_dp,
                                            C.c_double, C.c_double, c_vp, c_vp]),
    "b2r_integrate_units": (C.c_int, [c_vp, C.POINTER(Volume), c_vp, c_vp, C.POINTER(Intrinsic), c_dp, c_dp, C.c_int, C.c_int,
                                      C.POINTER(C.c_int32), c_vp, C.c_int32, c_vp, c_vp, c_vp, c_vp]),
    "b2r_selftest_div": (C.c_int, [c_vp, C.c_int64, C.c_uint32, c_i64p, c_vp]),
    "b2r_selftest_chain": (C.c_int, [c_vp, C.c_int64, C.c_uint32, c_i64p, c_vp]),
    "b2r_volume_reset": (C.c_int, [c_vp, C.POINTER(Volume), c_vp]),
    "b2r_volume_export": (C.c_int, [c_vp, C.POINTER(Volume), c_vp, c_vp, c_vp, c_vp]),
    "b2r_volume_import": (C.c_int, [c_vp, C.POINTER(Volume), c_vp, c_vp, c_vp, c_vp]),
    "b2r_extract_mesh_count": (C.c_int, [c_vp, C.POINTER(Volume), c_i64p, c_i64p, c_vp]),
    "b2r_extract_mesh_vertices": (C.c_int, [c_vp, C.POINTER(Volume), C.c_int64, c_vp, c_vp, c_vp]),
    "b2r_mesh_seam_buffers": (C.c_int, [c_vp, C.POINTER(Volume), C.POINTER(c_vp), C.POINTER(c_vp), c_i64p]),
    "b2r_extract_mesh_triangles": (C.c_int, [c_vp, C.POINTER(Volume), c_vp, c_vp]),
    "b2r_extract_points_count": (C.c_int, [c_vp, C.POINTER(Volume), c_i64p, c_vp]),
    "b2r_extract_points_emit": (C.c_int, [c_vp, C.POINTER(Volume), c_vp, c_vp, c_vp, c_vp]),
    "b2r_extract_points_emit_units": (C.c_int, [c_vp, C.POINTER(Volume), C.c_int, C.POINTER(C.c_int32), c_vp, c_vp, c_vp,
                                                c_vp, c_vp]),
    "b2r_icp_batch": (C.c_int, [c_vp, C.c_int, c_vp, c_i64p, c_vp, c_vp, c_i64p, c_vp, c_vp, c_vp, c_dp,
                                C.POINTER(IcpParams), C.POINTER(IcpResult), c_vp, c_vp]),
    "b2r_voxel_down_sample": (C.c_int, [c_vp, c_vp, c_vp, c_vp, C.c_int64, C.c_double, c_vp, c_vp, c_vp, c_i64p, c_vp]),
    "b2r_estimate_normals": (C.c_int, [c_vp, c_vp, C.c_int64, C.c_int, C.c_double, C.c_int, c_vp, c_vp, c_vp]),
    "b2r_statistical_outlier": (C.c_int, [c_vp, c_vp, C.c_int64, C.c_int, C.c_double, c_vp, c_i64p, c_vp]),
    "b2r_color_gradient": (C.c_int, [c_vp, c_vp, c_vp, c_vp, C.c_int64, C.c_double, C.c_int, c_vp, c_vp]),
    "b2r_knn_search": (C.c_int, [c_vp, c_vp, C.c_int64, c_vp, C.c_int64, C.c_int, C.c_double, C.c_int, c_vp, c_vp, c_vp]),
    "b2r_raycast_scene": (C.c_int, [c_vp, C.POINTER(Scene), c_vp, c_vp, c_vp, c_vp]),
}

_lib = None


class B2RError(RuntimeError):
    """Raised for every non-zero return code of the C ABI (Open3D raises RuntimeError from LogError)."""


def load():
    """Load libb200recon.so (no GPU needed just to load)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise B2RError("libb200recon.so is not built (%s); run `python -c 'import __graft_entry__ as g; "
                           "g.build()'` -- there is no CPU fallback" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        for name, (rt, at) in SIGNATURES.items():
            f = getattr(L, name)  # AttributeError if the symbol is missing
            f.restype = rt
            f.argtypes = at
        _lib = L
    return _lib


def check(rc):
    if rc != 0:
        raise B2RError("%s (code %d)" % (load().b2r_last_error().decode(), rc))
