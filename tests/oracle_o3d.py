"""TEST INFRASTRUCTURE: the subset of Open3D's Python API that stitcher.py uses, backed by the CPU ORACLE
(oracle/), so that one and the same call sequence (tests/stitch_mirror.py) can run on the oracle and on
b200recon and the results be compared.  Never imported by the product."""
import types

import numpy as np

from oracle import oracle as orc


class _Image:
    def __init__(self, a):
        self.a = np.ascontiguousarray(a)


class _RGBD:
    pass


class PointCloud:
    def __init__(self):
        self._p = np.zeros((0, 3))
        self._c = None
        self._n = None

    points = property(lambda s: s._p, lambda s, v: setattr(s, "_p", np.asarray(v, np.float64).reshape(-1, 3)))
    colors = property(lambda s: s._c if s._c is not None else np.zeros((0, 3)),
                      lambda s, v: setattr(s, "_c", np.asarray(v, np.float64).reshape(-1, 3)))
    normals = property(lambda s: s._n if s._n is not None else np.zeros((0, 3)),
                       lambda s, v: setattr(s, "_n", np.asarray(v, np.float64).reshape(-1, 3)))

    def has_colors(self):
        return self._c is not None and len(self._c) == len(self._p) and len(self._p) > 0

    def has_normals(self):
        return self._n is not None and len(self._n) == len(self._p) and len(self._p) > 0

    @staticmethod
    def create_from_rgbd_image(rgbd, intrinsic, extrinsic=None):
        K = (intrinsic.fx, intrinsic.fy, intrinsic.cx, intrinsic.cy)
        P, Cc = orc.backproject(rgbd.depth, rgbd.color, K, np.eye(4) if extrinsic is None else extrinsic)
        out = PointCloud()
        out._p, out._c = P, Cc
        return out

    def voxel_down_sample(self, vs):
        p, n, c = orc.voxel_down_sample(self._p, vs, self._n if self.has_normals() else None,
                                        self._c if self.has_colors() else None)
        out = PointCloud()
        out._p, out._n, out._c = p, n, c
        return out

    def estimate_normals(self, sp):
        self._n = orc.estimate_normals(self._p, orc.SEARCH_HYBRID, sp.radius, sp.max_nn,
                                       self._n if self.has_normals() else None)

    def transform(self, T):
        T = np.asarray(T, np.float64)
        self._p = self._p @ T[:3, :3].T + T[:3, 3]
        if self.has_normals():
            self._n = self._n @ T[:3, :3].T
        return self

    def __iadd__(self, o):
        keep_c = (len(self._p) == 0 or self.has_colors()) and o.has_colors()
        keep_n = (len(self._p) == 0 or self.has_normals()) and o.has_normals()
        c = np.concatenate([self.colors, o._c]) if keep_c else None
        n = np.concatenate([self.normals, o._n]) if keep_n else None
        self._p = np.concatenate([self._p, o._p])
        self._c, self._n = c, n
        return self

    def remove_statistical_outlier(self, nb_neighbors, std_ratio):
        idx, _ = orc.statistical_outlier(self._p, nb_neighbors, std_ratio)
        out = PointCloud()
        out._p = self._p[idx]
        out._c = self._c[idx] if self.has_colors() else None
        out._n = self._n[idx] if self.has_normals() else None
        return out, idx.tolist()


class _Hybrid:
    def __init__(self, radius, max_nn):
        self.radius, self.max_nn = radius, max_nn


class _Criteria:
    def __init__(self, relative_fitness=1e-6, relative_rmse=1e-6, max_iteration=30):
        self.f, self.r, self.n = relative_fitness, relative_rmse, max_iteration


class _Result:
    pass


def _create_rgbd(color, depth, depth_scale=1000.0, depth_trunc=3.0, convert_rgb_to_intensity=True):
    assert not convert_rgb_to_intensity          # stitcher.py:42 passes False
    r = _RGBD()
    r.depth = orc.depth_to_float(depth.a, depth_scale, depth_trunc)
    r.color = color.a
    return r


def _res(d):
    r = _Result()
    r.transformation, r.fitness, r.inlier_rmse = d["transformation"], d["fitness"], d["inlier_rmse"]
    return r


def _colored(src, tgt, dist, init, est=None, crit=None):
    crit = crit or _Criteria()
    return _res(orc.registration_colored_icp(src._p, src._c, tgt._p, tgt._n, tgt._c, dist, init, 0.968, crit.f, crit.r, crit.n))


def _icp(src, tgt, dist, init, est=None, crit=None):
    crit = crit or _Criteria()
    return _res(orc.registration_icp(src._p, tgt._p, tgt._n if tgt.has_normals() else None, dist, init,
                                     getattr(est, "kind", 0), crit.f, crit.r, crit.n))


class _P2P:
    kind = 0


class _P2L:
    kind = 1


class _Intrinsic:
    def __init__(self, width, height, fx, fy, cx, cy):
        self.width, self.height, self.fx, self.fy, self.cx, self.cy = width, height, fx, fy, cx, cy


class _Mesh:
    def __init__(self, V, T, Cc):
        self.vertices, self.triangles = V, T
        self.vertex_colors = Cc if Cc is not None else np.zeros((0, 3))


class _ColorType:
    NoColor, RGB8, Gray32 = 0, 1, 2


class _UniformTSDFVolume:
    """pipelines.integration.UniformTSDFVolume on the oracle (recon_oracle.cpp)."""

    def __init__(self, length, resolution, sdf_trunc, color_type=1, origin=(0.0, 0.0, 0.0)):
        self.v = orc.UniformTSDFVolume(length, resolution, sdf_trunc, int(color_type), origin)
        self.color_type = int(color_type)

    def integrate(self, rgbd, intrinsic, extrinsic):
        K = (intrinsic.fx, intrinsic.fy, intrinsic.cx, intrinsic.cy)
        return self.v.integrate(rgbd.depth, rgbd.color if self.color_type else None, K, np.asarray(extrinsic, np.float64))

    def extract_triangle_mesh(self):
        return _Mesh(*self.v.extract_triangle_mesh())

    def extract_point_cloud(self):
        P, N, Cc = self.v.extract_point_cloud()
        out = PointCloud()
        out._p, out._n, out._c = P, N, (Cc if Cc is not None and len(Cc) == len(P) else None)
        return out


geometry = types.SimpleNamespace(Image=_Image, RGBDImage=types.SimpleNamespace(create_from_color_and_depth=_create_rgbd),
                                 PointCloud=PointCloud, KDTreeSearchParamHybrid=_Hybrid)
utility = types.SimpleNamespace(Vector3dVector=lambda a: np.asarray(a, np.float64))
camera = types.SimpleNamespace(PinholeCameraIntrinsic=_Intrinsic)
pipelines = types.SimpleNamespace(registration=types.SimpleNamespace(
    registration_colored_icp=_colored, registration_icp=_icp, compute_fpfh_feature=lambda pcd, sp: None,
    TransformationEstimationForColoredICP=lambda: None, TransformationEstimationPointToPoint=_P2P,
    TransformationEstimationPointToPlane=_P2L, ICPConvergenceCriteria=_Criteria),
    integration=types.SimpleNamespace(UniformTSDFVolume=_UniformTSDFVolume, TSDFVolumeColorType=_ColorType))
