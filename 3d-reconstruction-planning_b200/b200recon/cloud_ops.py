"""Point-cloud preprocessing of `stitch_sequence` (stitcher.py:50-71,151-159) on the GPU: the methods below
are attached to geometry.PointCloud under Open3D's names.  Each is one call into libb200recon.so (cloud.cu);
arithmetic restated in SURVEY.md A.7-A.9."""
import ctypes as C

import numpy as np

from . import _lib, runtime
from .geometry import KDTreeSearchParamHybrid, KDTreeSearchParamKNN, KDTreeSearchParamRadius, PointCloud


def search_args(search_param):
    """(mode, radius, max_nn) of an Open3D KDTreeSearchParam*."""
    if search_param is None:
        search_param = KDTreeSearchParamKNN()
    if isinstance(search_param, KDTreeSearchParamHybrid):
        return _lib.SEARCH_HYBRID, float(search_param.radius), int(search_param.max_nn)
    if isinstance(search_param, KDTreeSearchParamKNN):
        return _lib.SEARCH_KNN, 0.0, int(search_param.knn)
    if isinstance(search_param, KDTreeSearchParamRadius):
        raise NotImplementedError("KDTreeSearchParamRadius (unbounded neighbour count) is not on the reference's path; "
                                  "use KDTreeSearchParamHybrid(radius, max_nn)")
    raise TypeError("search_param must be a KDTreeSearchParam")


def voxel_down_sample(self, voxel_size):
    """Open3D PointCloud.voxel_down_sample (stitcher.py:58,151).  Cells come out ordered by their
    first point (Open3D: unordered_map order -- compare as sets)."""
    voxel_size = float(voxel_size)
    if voxel_size <= 0.0:
        raise RuntimeError("[VoxelDownSample] voxel_size <= 0.")
    out = PointCloud()
    n = len(self._p)
    if n == 0:
        return out
    c = runtime.ctx()
    t = runtime.torch()
    P = self._p.get_dev()
    N = self._n.get_dev() if self.has_normals() else None
    Cc = self._c.get_dev() if self.has_colors() else None
    oP = t.empty((n, 3), dtype=t.float64, device=c.tdev)
    oN = t.empty((n, 3), dtype=t.float64, device=c.tdev) if N is not None else None
    oC = t.empty((n, 3), dtype=t.float64, device=c.tdev) if Cc is not None else None
    m = C.c_int64(0)
    _lib.check(_lib.load().b2r_voxel_down_sample(c.handle, runtime.ptr(P), runtime.ptr(N), runtime.ptr(Cc), n, voxel_size,
                                                 runtime.ptr(oP), runtime.ptr(oN), runtime.ptr(oC), C.byref(m), c.stream()))
    out._p.set_dev(oP[: m.value].clone())
    if oN is not None:
        out._n.set_dev(oN[: m.value].clone())
    if oC is not None:
        out._c.set_dev(oC[: m.value].clone())
    return out


def estimate_normals(self, search_param=None, fast_normal_computation=True):
    """Open3D PointCloud.estimate_normals (stitcher.py:61-63), in place."""
    if not fast_normal_computation:
        raise NotImplementedError("fast_normal_computation=False (iterative eigensolver) is not on the reference's path")
    mode, radius, max_nn = search_args(search_param)
    n = len(self._p)
    if n == 0:
        return self
    c = runtime.ctx()
    t = runtime.torch()
    P = self._p.get_dev()
    prev = self._n.get_dev() if self.has_normals() else None
    out = t.empty((n, 3), dtype=t.float64, device=c.tdev)
    _lib.check(_lib.load().b2r_estimate_normals(c.handle, runtime.ptr(P), n, mode, radius, max_nn, runtime.ptr(prev),
                                                runtime.ptr(out), c.stream()))
    self._n.set_dev(out)
    return self


def remove_statistical_outlier(self, nb_neighbors, std_ratio, print_progress=False):
    """Open3D PointCloud.remove_statistical_outlier (stitcher.py:158-159) -> (cloud, kept indices)."""
    if nb_neighbors < 1 or std_ratio <= 0:
        raise RuntimeError("[RemoveStatisticalOutliers] Illegal input parameters, the number of neighbors and "
                           "standard deviation ratio must be positive.")
    n = len(self._p)
    if n == 0:
        return PointCloud(), []
    c = runtime.ctx()
    t = runtime.torch()
    P = self._p.get_dev()
    idx = t.empty((n,), dtype=t.int64, device=c.tdev)
    m = C.c_int64(0)
    _lib.check(_lib.load().b2r_statistical_outlier(c.handle, runtime.ptr(P), n, int(nb_neighbors), float(std_ratio),
                                                   runtime.ptr(idx), C.byref(m), c.stream()))
    kept = idx[: m.value]
    out = PointCloud()
    out._p.set_dev(P[kept])
    if self.has_colors():
        out._c.set_dev(self._c.get_dev()[kept])
    if self.has_normals():
        out._n.set_dev(self._n.get_dev()[kept])
    return out, kept.cpu().numpy().tolist()


def color_gradient(pcd, radius, max_nn=30):
    """InitializePointCloudForColoredICP: per-point colour gradient, CUDA tensor float64 [N,3]."""
    if not (pcd.has_normals() and pcd.has_colors()):
        raise RuntimeError("ColoredICP requires pre-computed normal vectors and colors for target PointCloud.")
    c = runtime.ctx()
    t = runtime.torch()
    n = len(pcd._p)
    g = t.zeros((n, 3), dtype=t.float64, device=c.tdev)
    if n:
        _lib.check(_lib.load().b2r_color_gradient(c.handle, runtime.ptr(pcd._p.get_dev()), runtime.ptr(pcd._n.get_dev()),
                                                  runtime.ptr(pcd._c.get_dev()), n, float(radius), int(max_nn),
                                                  runtime.ptr(g), c.stream()))
    return g


def knn_search(points, queries, search_param):
    """Neighbour lists of the exact grid search (tests): (idx int32 [Q,k] -1 padded, d2 float64 [Q,k])."""
    mode, radius, k = search_args(search_param)
    c = runtime.ctx()
    t = runtime.torch()
    P = runtime.to_device(np.asarray(points, np.float64))
    Q = runtime.to_device(np.asarray(queries, np.float64))
    idx = t.empty((len(Q), k), dtype=t.int32, device=c.tdev)
    d2 = t.empty((len(Q), k), dtype=t.float64, device=c.tdev)
    _lib.check(_lib.load().b2r_knn_search(c.handle, runtime.ptr(P), len(P), runtime.ptr(Q), len(Q), mode, radius, k,
                                          runtime.ptr(idx), runtime.ptr(d2), c.stream()))
    return idx.cpu().numpy(), d2.cpu().numpy()


PointCloud.voxel_down_sample = voxel_down_sample
PointCloud.estimate_normals = estimate_normals
PointCloud.remove_statistical_outlier = remove_statistical_outlier
