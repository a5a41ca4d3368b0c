"""ORACLE -- TEST INFRASTRUCTURE ONLY (ctypes loader for oracle/liborc.so).

CPU restatement of the Open3D 0.18.0 legacy arithmetic behind the reference's stitcher.py
(SURVEY.md Appendix A).  PARITY UNPINNED against the real Open3D wheel: it is absent from
/root/reference, not installable here, and the reference has no tests / golden vectors.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  The product package (b200recon) never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

c_dp = C.POINTER(C.c_double)
c_fp = C.POINTER(C.c_float)


def build(force=False):
    so = os.path.join(_HERE, "liborc.so")
    srcs = [os.path.join(_HERE, f) for f in os.listdir(_HERE) if f.endswith((".cpp", ".h", ".inc"))]
    if force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "liborc.so"], stdout=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        so = os.path.join(_HERE, "liborc.so")
        if not os.path.exists(so):
            build()
        L = C.CDLL(so)
        L.orc_volume_create.restype = C.c_void_p
        L.orc_volume_create.argtypes = [C.c_double, C.c_int, C.c_double, C.c_int, c_dp, C.c_int, C.c_int]
        L.orc_volume_destroy.argtypes = [C.c_void_p]
        L.orc_volume_reset.argtypes = [C.c_void_p]
        for n, rt in (("orc_volume_tsdf", c_fp), ("orc_volume_weight", c_fp), ("orc_volume_color", c_dp),
                      ("orc_volume_vertices", c_dp), ("orc_volume_vertex_colors", c_dp),
                      ("orc_volume_normals", c_dp), ("orc_volume_triangles", C.POINTER(C.c_int32))):
            getattr(L, n).restype = rt
            getattr(L, n).argtypes = [C.c_void_p]
        L.orc_volume_integrate.restype = C.c_int64
        L.orc_volume_integrate.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                           C.c_double, C.c_double, C.c_double, C.c_double, c_dp]
        L.orc_volume_extract_mesh.argtypes = [C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
        L.orc_volume_extract_points.argtypes = [C.c_void_p, C.POINTER(C.c_int64)]
        L.orc_scalable_create.restype = C.c_void_p
        L.orc_scalable_create.argtypes = [C.c_double, C.c_double, C.c_int, C.c_int, C.c_int]
        for n in ("orc_scalable_destroy", "orc_scalable_reset"):
            getattr(L, n).argtypes = [C.c_void_p]
        L.orc_scalable_num_units.restype = C.c_int64
        L.orc_scalable_num_units.argtypes = [C.c_void_p]
        L.orc_scalable_unit_keys.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_scalable_unit_arrays.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_scalable_integrate.restype = C.c_int64
        L.orc_scalable_integrate.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_double, C.c_double,
                                             C.c_double, C.c_double, c_dp, c_dp, C.POINTER(C.c_int64)]
        L.orc_scalable_extract_mesh.argtypes = [C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
        L.orc_scalable_extract_points.argtypes = [C.c_void_p, C.POINTER(C.c_int64)]
        for n, rt in (("orc_scalable_vertices", c_dp), ("orc_scalable_vertex_colors", c_dp),
                      ("orc_scalable_normals", c_dp), ("orc_scalable_triangles", C.POINTER(C.c_int32))):
            getattr(L, n).restype = rt
            getattr(L, n).argtypes = [C.c_void_p]
        L.orc_set_num_threads.argtypes = [C.c_int]
        L.orc_max_threads.restype = C.c_int
        L.orc_depth_to_float.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.c_void_p]
        L.orc_distance_multiplier.argtypes = [C.c_int, C.c_int, C.c_double, C.c_double, C.c_double, C.c_double, C.c_void_p]
        L.orc_backproject.restype = C.c_int64
        L.orc_backproject.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double,
                                      C.c_double, C.c_double, c_dp, C.c_void_p, C.c_void_p]
        L.orc_registration_icp.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int64, C.c_double,
                                           c_dp, C.c_int, C.c_double, C.c_double, C.c_int, c_dp, c_dp, c_dp,
                                           C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int)]
        L.orc_nearest.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_double, C.c_void_p, C.c_void_p]
        L.orc_raycast_scene.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_registration_colored_icp.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p,
                                                   C.c_void_p, C.c_int64, C.c_double, c_dp, C.c_double, C.c_double,
                                                   C.c_double, C.c_int, c_dp, c_dp, c_dp, C.c_void_p,
                                                   C.POINTER(C.c_int64), C.POINTER(C.c_int)]
        L.orc_voxel_down_sample.restype = C.c_int64
        L.orc_voxel_down_sample.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_double, C.c_void_p,
                                            C.c_void_p, C.c_void_p]
        L.orc_estimate_normals.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.c_double, C.c_int, C.c_void_p, C.c_void_p]
        L.orc_statistical_outlier.restype = C.c_int64
        L.orc_statistical_outlier.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.c_double, C.c_void_p, C.c_void_p]
        L.orc_color_gradient.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_double, C.c_int, C.c_void_p]
        L.orc_knn.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int, C.c_double, C.c_int, C.c_void_p,
                              C.c_void_p]
        _LIB = L
    return _LIB


def _dp(a):
    return a.ctypes.data_as(c_dp)


def _vp(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def set_num_threads(n):
    """OpenMP team size for the oracle's parallel loops (torchrun exports OMP_NUM_THREADS=1 to its workers)."""
    lib().orc_set_num_threads(int(n))


def max_threads():
    return int(lib().orc_max_threads())


# ---------------------------------------------------------------------------------- images
def depth_to_float(depth, depth_scale=1000.0, depth_trunc=3.0):
    """A.1 ConvertDepthToFloatImage."""
    depth = np.ascontiguousarray(depth)
    dt = {np.dtype(np.uint8): 0, np.dtype(np.uint16): 1, np.dtype(np.float32): 2}[depth.dtype]
    H, W = depth.shape[:2]
    out = np.empty((H, W), np.float32)
    rc = lib().orc_depth_to_float(_vp(depth), dt, H, W, depth_scale, depth_trunc, _vp(out))
    assert rc == 0
    return out


def distance_multiplier(H, W, fx, fy, cx, cy):
    out = np.empty((H, W), np.float32)
    lib().orc_distance_multiplier(H, W, fx, fy, cx, cy, _vp(out))
    return out


def backproject(depth_f32, color_u8, K, extrinsic=None):
    """A.2 CreateFromRGBDImage -> (points f64[N,3], colors f64[N,3] or None)."""
    depth_f32 = np.ascontiguousarray(depth_f32, np.float32)
    H, W = depth_f32.shape
    ext = np.ascontiguousarray(np.eye(4) if extrinsic is None else extrinsic, np.float64)
    pts = np.empty((H * W, 3), np.float64)
    cols = None
    nc = 0
    if color_u8 is not None:
        color_u8 = np.ascontiguousarray(color_u8, np.uint8)
        nc = color_u8.shape[2]
        cols = np.empty((H * W, 3), np.float64)
    n = lib().orc_backproject(_vp(depth_f32), _vp(color_u8), nc, H, W, K[0], K[1], K[2], K[3], _dp(ext),
                              _vp(pts), _vp(cols))
    return pts[:n].copy(), (None if cols is None else cols[:n].copy())


# ---------------------------------------------------------------------------------- volume
class UniformTSDFVolume:
    """A.3/A.4.  Arrays are exposed in Open3D index order: [x, y, z] (z fastest)."""

    def __init__(self, length, resolution, sdf_trunc, color_type=1, origin=(0.0, 0.0, 0.0), x_range=None):
        self.R = int(resolution)
        self.length = float(length)
        self.sdf_trunc = float(sdf_trunc)
        self.color_type = int(color_type)
        self.origin = np.ascontiguousarray(origin, np.float64)
        xb, xe = (0, self.R) if x_range is None else x_range
        self.x_range = (xb, xe)
        self._h = C.c_void_p(lib().orc_volume_create(self.length, self.R, self.sdf_trunc, self.color_type,
                                                      _dp(self.origin), xb, xe))

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_volume_destroy(self._h)
            self._h = None

    def reset(self):
        lib().orc_volume_reset(self._h)

    def integrate(self, depth_f32, color, K, extrinsic):
        depth_f32 = np.ascontiguousarray(depth_f32, np.float32)
        H, W = depth_f32.shape
        if color is not None:
            color = np.ascontiguousarray(color)
        ext = np.ascontiguousarray(extrinsic, np.float64)
        n = lib().orc_volume_integrate(self._h, _vp(depth_f32), _vp(color), H, W, K[0], K[1], K[2], K[3], _dp(ext))
        if n < 0:
            raise RuntimeError("oracle integrate failed: %d" % n)
        return int(n)

    def _shape(self):
        return (self.x_range[1] - self.x_range[0], self.R, self.R)

    def tsdf(self):
        n = int(np.prod(self._shape()))
        return np.ctypeslib.as_array(lib().orc_volume_tsdf(self._h), (n,)).reshape(self._shape())

    def weight(self):
        n = int(np.prod(self._shape()))
        return np.ctypeslib.as_array(lib().orc_volume_weight(self._h), (n,)).reshape(self._shape())

    def color(self):
        if self.color_type == 0:
            return None
        n = int(np.prod(self._shape()))
        return np.ctypeslib.as_array(lib().orc_volume_color(self._h), (n * 3,)).reshape(self._shape() + (3,))

    def extract_triangle_mesh(self):
        nv, nt = C.c_int64(0), C.c_int64(0)
        rc = lib().orc_volume_extract_mesh(self._h, C.byref(nv), C.byref(nt))
        assert rc == 0
        nv, nt = nv.value, nt.value
        V = np.ctypeslib.as_array(lib().orc_volume_vertices(self._h), (nv * 3,)).reshape(nv, 3).copy() if nv else np.zeros((0, 3))
        T = np.ctypeslib.as_array(lib().orc_volume_triangles(self._h), (nt * 3,)).reshape(nt, 3).copy() if nt else np.zeros((0, 3), np.int32)
        Cc = None
        if self.color_type != 0:
            Cc = np.ctypeslib.as_array(lib().orc_volume_vertex_colors(self._h), (nv * 3,)).reshape(nv, 3).copy() if nv else np.zeros((0, 3))
        return V, T, Cc

    def extract_point_cloud(self):
        n = C.c_int64(0)
        rc = lib().orc_volume_extract_points(self._h, C.byref(n))
        assert rc == 0
        n = n.value
        if n == 0:
            z = np.zeros((0, 3))
            return z, z.copy(), (None if self.color_type == 0 else z.copy())
        P = np.ctypeslib.as_array(lib().orc_volume_vertices(self._h), (n * 3,)).reshape(n, 3).copy()
        N = np.ctypeslib.as_array(lib().orc_volume_normals(self._h), (n * 3,)).reshape(n, 3).copy()
        Cc = None
        if self.color_type != 0:
            Cc = np.ctypeslib.as_array(lib().orc_volume_vertex_colors(self._h), (n * 3,)).reshape(n, 3).copy()
        return P, N, Cc


# ---------------------------------------------------------------------------------- ICP
class ScalableTSDFVolume:
    """Open3D ScalableTSDFVolume restated (scalable_oracle.cpp): a hash of unit_resolution^3 uniform units."""

    def __init__(self, voxel_length, sdf_trunc, color_type=1, volume_unit_resolution=16, depth_sampling_stride=4):
        self.voxel_length = float(voxel_length)
        self.sdf_trunc = float(sdf_trunc)
        self.color_type = int(color_type)
        self.res = int(volume_unit_resolution)
        self.stride = int(depth_sampling_stride)
        self._h = C.c_void_p(lib().orc_scalable_create(self.voxel_length, self.sdf_trunc, self.color_type, self.res,
                                                        self.stride))
        self.touched = 0

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_scalable_destroy(self._h)
            self._h = None

    def reset(self):
        lib().orc_scalable_reset(self._h)

    def set_bounds(self, unit_offset, n_units):
        """Restrict the hash to the cube of `n_units`^3 units starting at `unit_offset` (unit indices) -- the product's
        bounded backing cube; touches outside are counted (`clipped()`), not opened.  None removes the bound."""
        if unit_offset is None:
            lib().orc_scalable_set_bounds(self._h, None, None)
            return
        lo = (C.c_int32 * 3)(*[int(v) for v in unit_offset])
        hi = (C.c_int32 * 3)(*[int(v) + int(n_units) for v in unit_offset])
        lib().orc_scalable_set_bounds(self._h, lo, hi)

    def clipped(self):
        lib().orc_scalable_clipped.restype = C.c_int64
        return int(lib().orc_scalable_clipped(self._h))

    def integrate(self, depth_f32, color, K, extrinsic):
        depth_f32 = np.ascontiguousarray(depth_f32, np.float32)
        H, W = depth_f32.shape
        if color is not None:
            color = np.ascontiguousarray(color)
        ext = np.ascontiguousarray(extrinsic, np.float64)
        pose = np.ascontiguousarray(np.linalg.inv(ext))     # the product's host does the same inversion
        t = C.c_int64(0)
        n = lib().orc_scalable_integrate(self._h, _vp(depth_f32), _vp(color), H, W, K[0], K[1], K[2], K[3], _dp(ext),
                                         _dp(pose), C.byref(t))
        if n < 0:
            raise RuntimeError("oracle scalable integrate failed: %d" % n)
        self.touched = t.value
        return int(n)

    def unit_keys(self):
        n = lib().orc_scalable_num_units(self._h)
        out = np.zeros((n, 3), np.int32)
        if n:
            lib().orc_scalable_unit_keys(self._h, _vp(out))
        return out

    def unit_arrays(self, key):
        """(tsdf, weight, color|None) of one unit in Open3D order [x, y, z]."""
        r = self.res
        key = np.ascontiguousarray(key, np.int32)
        ts, w = np.zeros((r, r, r), np.float32), np.zeros((r, r, r), np.float32)
        col = np.zeros((r, r, r, 3), np.float64) if self.color_type else None
        rc = lib().orc_scalable_unit_arrays(self._h, _vp(key), _vp(ts), _vp(w), _vp(col))
        if rc:
            raise KeyError(tuple(key))
        return ts, w, col

    def dense_arrays(self, unit_offset, n_units):
        """The units inside the box of n_units^3 units starting at unit_offset, laid out as one dense cube
        [x, y, z] (zeros where no unit exists); also returns the number of opened units outside the box."""
        r, R = self.res, self.res * int(n_units)
        ts, w = np.zeros((R, R, R), np.float32), np.zeros((R, R, R), np.float32)
        col = np.zeros((R, R, R, 3), np.float64) if self.color_type else None
        outside = 0
        for k in self.unit_keys():
            l = k - np.asarray(unit_offset)
            if (l < 0).any() or (l >= n_units).any():
                outside += 1
                continue
            a, b, c = (l * r).tolist()
            t1, w1, c1 = self.unit_arrays(k)
            ts[a:a + r, b:b + r, c:c + r] = t1
            w[a:a + r, b:b + r, c:c + r] = w1
            if col is not None:
                col[a:a + r, b:b + r, c:c + r] = c1
        return ts, w, col, outside

    def extract_triangle_mesh(self):
        nv, nt = C.c_int64(0), C.c_int64(0)
        assert lib().orc_scalable_extract_mesh(self._h, C.byref(nv), C.byref(nt)) == 0
        nv, nt = nv.value, nt.value
        V = np.ctypeslib.as_array(lib().orc_scalable_vertices(self._h), (nv * 3,)).reshape(nv, 3).copy() if nv else np.zeros((0, 3))
        T = np.ctypeslib.as_array(lib().orc_scalable_triangles(self._h), (nt * 3,)).reshape(nt, 3).copy() if nt else np.zeros((0, 3), np.int32)
        Cc = None
        if self.color_type != 0:
            Cc = np.ctypeslib.as_array(lib().orc_scalable_vertex_colors(self._h), (nv * 3,)).reshape(nv, 3).copy() if nv else np.zeros((0, 3))
        return V, T, Cc

    def extract_point_cloud(self):
        n = C.c_int64(0)
        assert lib().orc_scalable_extract_points(self._h, C.byref(n)) == 0
        n = n.value
        if n == 0:
            z = np.zeros((0, 3))
            return z, z.copy(), (None if self.color_type == 0 else z.copy())
        P = np.ctypeslib.as_array(lib().orc_scalable_vertices(self._h), (n * 3,)).reshape(n, 3).copy()
        N = np.ctypeslib.as_array(lib().orc_scalable_normals(self._h), (n * 3,)).reshape(n, 3).copy()
        Cc = None
        if self.color_type != 0:
            Cc = np.ctypeslib.as_array(lib().orc_scalable_vertex_colors(self._h), (n * 3,)).reshape(n, 3).copy()
        return P, N, Cc


def registration_icp(src, tgt, tgt_normals, max_dist, init=None, kind=1, rel_fitness=1e-6, rel_rmse=1e-6,
                     max_iter=30):
    """A.5.  kind 0 = point-to-point, 1 = point-to-plane.  Returns dict."""
    src = np.ascontiguousarray(src, np.float64)
    tgt = np.ascontiguousarray(tgt, np.float64)
    if tgt_normals is not None:
        tgt_normals = np.ascontiguousarray(tgt_normals, np.float64)
    init = np.ascontiguousarray(np.eye(4) if init is None else init, np.float64)
    T = np.empty((4, 4), np.float64)
    fit, rmse = C.c_double(0), C.c_double(0)
    corr = np.empty((len(src),), np.int32)
    nc, it = C.c_int64(0), C.c_int(0)
    rc = lib().orc_registration_icp(_vp(src), len(src), _vp(tgt), _vp(tgt_normals), len(tgt), max_dist, _dp(init),
                                    kind, rel_fitness, rel_rmse, max_iter, _dp(T), C.byref(fit), C.byref(rmse),
                                    _vp(corr), C.byref(nc), C.byref(it))
    if rc != 0:
        raise RuntimeError("oracle registration_icp error %d" % rc)
    return dict(transformation=T, fitness=fit.value, inlier_rmse=rmse.value, corr=corr, n_corr=nc.value,
                iterations=it.value)


def registration_colored_icp(src, src_col, tgt, tgt_normals, tgt_col, max_dist, init=None, lambda_geometric=0.968,
                             rel_fitness=1e-6, rel_rmse=1e-6, max_iter=30, tgt_grad=None):
    """A.6: colour gradient of the target (Hybrid(2*max_dist, 30)) + two-residual ICP.  Returns dict."""
    a = lambda v: np.ascontiguousarray(v, np.float64)
    src, src_col, tgt, tgt_normals, tgt_col = a(src), a(src_col), a(tgt), a(tgt_normals), a(tgt_col)
    if tgt_grad is None:
        tgt_grad = color_gradient(tgt, tgt_normals, tgt_col, 2.0 * max_dist, 30)
    tgt_grad = a(tgt_grad)
    init = a(np.eye(4) if init is None else init)
    T = np.empty((4, 4), np.float64)
    fit, rmse = C.c_double(0), C.c_double(0)
    corr = np.empty((len(src),), np.int32)
    nc, it = C.c_int64(0), C.c_int(0)
    rc = lib().orc_registration_colored_icp(_vp(src), _vp(src_col), len(src), _vp(tgt), _vp(tgt_normals), _vp(tgt_col),
                                            _vp(tgt_grad), len(tgt), max_dist, _dp(init), lambda_geometric, rel_fitness,
                                            rel_rmse, max_iter, _dp(T), C.byref(fit), C.byref(rmse), _vp(corr),
                                            C.byref(nc), C.byref(it))
    if rc != 0:
        raise RuntimeError("oracle registration_colored_icp error %d" % rc)
    return dict(transformation=T, fitness=fit.value, inlier_rmse=rmse.value, corr=corr, n_corr=nc.value,
                iterations=it.value, color_gradient=tgt_grad)


# ---------------------------------------------------------------------------------- point-cloud preprocessing
SEARCH_KNN, SEARCH_RADIUS, SEARCH_HYBRID = 0, 1, 2


def voxel_down_sample(pts, voxel_size, normals=None, colors=None):
    """A.7 -> (points, normals|None, colors|None); cells in order of their first point index."""
    a = lambda v: None if v is None else np.ascontiguousarray(v, np.float64)
    pts, normals, colors = a(pts), a(normals), a(colors)
    n = len(pts)
    op = np.empty((n, 3))
    on = np.empty((n, 3)) if normals is not None else None
    oc = np.empty((n, 3)) if colors is not None else None
    m = lib().orc_voxel_down_sample(_vp(pts), _vp(normals), _vp(colors), n, float(voxel_size), _vp(op), _vp(on), _vp(oc))
    if m < 0:
        raise RuntimeError("[VoxelDownSample] voxel_size <= 0.")
    return op[:m].copy(), None if on is None else on[:m].copy(), None if oc is None else oc[:m].copy()


def estimate_normals(pts, mode=SEARCH_HYBRID, radius=0.1, max_nn=30, prev=None):
    """A.8 (fast_normal_computation=True)."""
    pts = np.ascontiguousarray(pts, np.float64)
    prev = None if prev is None else np.ascontiguousarray(prev, np.float64)
    out = np.empty_like(pts)
    lib().orc_estimate_normals(_vp(pts), len(pts), mode, float(radius), int(max_nn), _vp(prev), _vp(out))
    return out


def statistical_outlier(pts, nb_neighbors, std_ratio):
    """A.9 -> (kept indices int64 ascending, per-point mean neighbour distance)."""
    pts = np.ascontiguousarray(pts, np.float64)
    idx = np.empty((len(pts),), np.int64)
    avg = np.empty((len(pts),), np.float64)
    m = lib().orc_statistical_outlier(_vp(pts), len(pts), int(nb_neighbors), float(std_ratio), _vp(idx), _vp(avg))
    return idx[:m].copy(), avg


def color_gradient(pts, normals, colors, radius, max_nn=30):
    """A.6 InitializePointCloudForColoredICP."""
    a = lambda v: np.ascontiguousarray(v, np.float64)
    pts, normals, colors = a(pts), a(normals), a(colors)
    g = np.empty_like(pts)
    lib().orc_color_gradient(_vp(pts), _vp(normals), _vp(colors), len(pts), float(radius), int(max_nn), _vp(g))
    return g


def knn(pts, q, mode, radius, k):
    pts = np.ascontiguousarray(pts, np.float64)
    q = np.ascontiguousarray(q, np.float64)
    idx = np.empty((len(q), k), np.int32)
    d2 = np.empty((len(q), k), np.float64)
    lib().orc_knn(_vp(pts), len(pts), _vp(q), len(q), mode, float(radius), k, _vp(idx), _vp(d2))
    return idx, d2


def nearest(tgt, q, max_dist):
    tgt = np.ascontiguousarray(tgt, np.float64)
    q = np.ascontiguousarray(q, np.float64)
    idx = np.empty((len(q),), np.int32)
    d2 = np.empty((len(q),), np.float64)
    lib().orc_nearest(_vp(tgt), len(tgt), _vp(q), len(q), max_dist, _vp(idx), _vp(d2))
    return idx, d2


# ---------------------------------------------------------------------------------- scene
class _Scene(C.Structure):
    _fields_ = [("c2w", C.c_float * 12), ("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float),
                ("W", C.c_int), ("H", C.c_int), ("ell_c", C.c_float * 3), ("ell_r", C.c_float * 3),
                ("plane_z", C.c_float), ("has_plane", C.c_int), ("checker", C.c_float), ("depth_scale", C.c_float),
                ("n_tris", C.c_int), ("mesh_min", C.c_float * 3), ("mesh_max", C.c_float * 3)]


def raycast_scene(cam_to_world, K, W, H, ell_c=(0, 0, 0.02), ell_r=(0.084, 0.034, 0.020), plane_z=0.0,
                  has_plane=True, checker=0.04, depth_scale=1000.0, tris=None):
    """CPU twin of the synthetic frame source -> (depth uint16 [H,W], color uint8 [H,W,3])."""
    sc = _Scene()
    m = np.asarray(cam_to_world, np.float64)[:3, :4].astype(np.float32).reshape(-1)
    for i in range(12):
        sc.c2w[i] = float(m[i])
    sc.fx, sc.fy, sc.cx, sc.cy = [float(np.float32(v)) for v in K]
    sc.W, sc.H = W, H
    for i in range(3):
        sc.ell_c[i] = float(np.float32(ell_c[i]))
        sc.ell_r[i] = float(np.float32(ell_r[i]))
    sc.plane_z = float(np.float32(plane_z))
    sc.has_plane = int(bool(has_plane))
    sc.checker = float(np.float32(checker))
    sc.depth_scale = float(np.float32(depth_scale))
    tr = None
    sc.n_tris = 0
    if tris is not None:
        tr = np.ascontiguousarray(tris, np.float32).reshape(-1, 9)
        sc.n_tris = len(tr)
        mn = tr.reshape(-1, 3).min(0)
        mx = tr.reshape(-1, 3).max(0)
        for i in range(3):
            sc.mesh_min[i] = float(mn[i])
            sc.mesh_max[i] = float(mx[i])
    depth = np.empty((H, W), np.uint16)
    color = np.empty((H, W, 3), np.uint8)
    lib().orc_raycast_scene(C.byref(sc), _vp(tr), _vp(depth), _vp(color))
    return depth, color
