"""`o3d.pipelines.registration` subset used by stitcher.py:94-110 (+ the north_star point-to-plane
variant), backed by the batched ICP kernels of libb200recon.so."""
import ctypes as C

import numpy as np

from .. import _lib, geometry, runtime
from ..utility import Vector2iVector


class ICPConvergenceCriteria:
    """ICPConvergenceCriteria(relative_fitness=1e-6, relative_rmse=1e-6, max_iteration=30)."""

    def __init__(self, relative_fitness=1e-6, relative_rmse=1e-6, max_iteration=30):
        self.relative_fitness = float(relative_fitness)
        self.relative_rmse = float(relative_rmse)
        self.max_iteration = int(max_iteration)

    def __repr__(self):
        return "ICPConvergenceCriteria class with relative_fitness=%e, relative_rmse=%e, and max_iteration=%d" % (
            self.relative_fitness, self.relative_rmse, self.max_iteration)


class TransformationEstimation:
    _kind = -1


class TransformationEstimationPointToPoint(TransformationEstimation):
    _kind = 0

    def __init__(self, with_scaling=False):
        if with_scaling:
            raise NotImplementedError("with_scaling=True is not on the reference's path")
        self.with_scaling = False


class TransformationEstimationPointToPlane(TransformationEstimation):
    _kind = 1

    def __init__(self, kernel=None):
        if kernel is not None:
            raise NotImplementedError("robust kernels are not on the reference's path (default L2 only)")


class TransformationEstimationForColoredICP(TransformationEstimation):
    _kind = 2

    def __init__(self, lambda_geometric=0.968, kernel=None):
        if kernel is not None:
            raise NotImplementedError("robust kernels are not on the reference's path (default L2 only)")
        self.lambda_geometric = float(lambda_geometric)


class RegistrationResult:
    """Open3D RegistrationResult: transformation, fitness, inlier_rmse, correspondence_set."""

    def __init__(self):
        self.transformation = np.eye(4)
        self.fitness = 0.0
        self.inlier_rmse = 0.0
        self.correspondence_set = Vector2iVector()
        self.iterations = 0
        self.converged = False

    def __repr__(self):
        return ("RegistrationResult with fitness=%e, inlier_rmse=%e, and correspondence_set size of %d\n"
                "Access transformation to get result." % (self.fitness, self.inlier_rmse, len(self.correspondence_set)))


class Feature:
    """Open3D registration.Feature: `data` float64 [dim, N]."""

    def __init__(self):
        self.data = np.zeros((33, 0))

    def dimension(self):
        return self.data.shape[0]

    def num(self):
        return self.data.shape[1]


def compute_fpfh_feature(input, search_param):
    """stitcher.py:67-69.  The reference computes FPFH and never reads it (stitcher.py:89-90 assign,
    nothing consumes): dead work, so no kernel is spent on it -- a zero Feature of Open3D's shape
    [33, N] is returned (documented in DESIGN.md)."""
    f = Feature()
    f.data = np.zeros((33, len(input)), np.float64)
    return f


def _check_icp_args(target, max_correspondence_distance, estimation_method):
    if max_correspondence_distance <= 0.0:
        raise RuntimeError("Invalid max_correspondence_distance.")
    kind = estimation_method._kind
    if kind in (1, 2) and not target.has_normals():
        raise RuntimeError("TransformationEstimationPointToPlane and TransformationEstimationColoredICP require "
                           "pre-computed normal vectors for target PointCloud.")
    return kind


def registration_icp_batch(sources, targets, max_correspondence_distance, inits=None,
                           estimation_method=None, criteria=None, with_correspondences=True, device=None):
    """Batched Open3D `registration_icp`: pair p registers sources[p] to targets[p]; all pairs run
    concurrently in ONE sequence of kernel launches (BASELINE config 4).  Returns a list of
    RegistrationResult.  Multi-GPU: callers give each rank the pairs p with p % world == rank
    (`parallel.icp_round_robin`); no collective is involved."""
    estimation_method = estimation_method or TransformationEstimationPointToPoint()
    criteria = criteria or ICPConvergenceCriteria()
    n = len(sources)
    if len(targets) != n:
        raise RuntimeError("registration_icp_batch: %d sources but %d targets" % (n, len(targets)))
    kind = -1
    for tg in targets:
        kind = _check_icp_args(tg, float(max_correspondence_distance), estimation_method)
    if n == 0:
        return []
    c = runtime.ctx(device)
    t = runtime.torch()
    src_off = np.zeros(n + 1, np.int64)
    tgt_off = np.zeros(n + 1, np.int64)
    for p in range(n):
        src_off[p + 1] = src_off[p] + len(sources[p])
        tgt_off[p + 1] = tgt_off[p] + len(targets[p])

    def cat(arrs):
        arrs = [a for a in arrs if a.shape[0] > 0]
        if not arrs:
            return t.zeros((0, 3), dtype=t.float64, device=c.tdev)
        return arrs[0] if len(arrs) == 1 else t.cat(arrs)

    S = cat([s._p.get_dev() for s in sources]).contiguous()
    T = cat([g._p.get_dev() for g in targets]).contiguous()
    N = cat([g._n.get_dev() for g in targets]).contiguous() if kind >= 1 else None
    SC = TC = TG = None
    if kind == 2:
        # Open3D RegistrationColoredICP: both clouds need colours; the target's colour gradient is precomputed with
        # KDTreeSearchParamHybrid(radius = 2 * max_correspondence_distance, max_nn = 30)
        from .. import cloud_ops
        for s_, g_ in zip(sources, targets):
            if not (s_.has_colors() and g_.has_colors()):
                raise RuntimeError("ColoredICP requires source and target PointClouds with colors.")
        SC = cat([s_._c.get_dev() for s_ in sources]).contiguous()
        TC = cat([g_._c.get_dev() for g_ in targets]).contiguous()
        TG = cat([cloud_ops.color_gradient(g_, 2.0 * float(max_correspondence_distance), 30) for g_ in targets]).contiguous()
    init = np.ascontiguousarray(np.stack([np.eye(4) if inits is None or inits[p] is None else
                                          np.asarray(inits[p], np.float64).reshape(4, 4) for p in range(n)]))
    prm = _lib.IcpParams(float(max_correspondence_distance), criteria.relative_fitness, criteria.relative_rmse,
                         criteria.max_iteration, kind, float(getattr(estimation_method, "lambda_geometric", 0.968)))
    res = (_lib.IcpResult * n)()
    corr = t.empty((int(src_off[n]),), dtype=t.int32, device=c.tdev) if with_correspondences else None
    _lib.check(_lib.load().b2r_icp_batch(c.handle, n, runtime.ptr(S), src_off.ctypes.data_as(_lib.c_i64p),
                                         runtime.ptr(T), runtime.ptr(N), tgt_off.ctypes.data_as(_lib.c_i64p),
                                         runtime.ptr(SC), runtime.ptr(TC), runtime.ptr(TG),
                                         init.ctypes.data_as(_lib.c_dp), C.byref(prm), res,
                                         runtime.ptr(corr), c.stream()))
    corr_h = corr.cpu().numpy() if corr is not None else None
    out = []
    for p in range(n):
        r = RegistrationResult()
        r.transformation = np.array(res[p].transformation[:], np.float64).reshape(4, 4)
        r.fitness = float(res[p].fitness)
        r.inlier_rmse = float(res[p].inlier_rmse)
        r.iterations = int(res[p].iterations)
        r.converged = bool(res[p].converged)
        if corr_h is not None:
            cp = corr_h[src_off[p]:src_off[p + 1]]
            i = np.nonzero(cp >= 0)[0]
            r.correspondence_set = Vector2iVector(np.stack([i, cp[i]], 1).astype(np.int32))
        out.append(r)
    return out


def registration_icp(source, target, max_correspondence_distance, init=None, estimation_method=None, criteria=None):
    """Open3D pipelines.registration.registration_icp (stitcher.py:106-110) -> RegistrationResult
    with .transformation / .fitness / .inlier_rmse / .correspondence_set."""
    return registration_icp_batch([source], [target], max_correspondence_distance, [init], estimation_method,
                                  criteria)[0]


def evaluate_registration(source, target, max_correspondence_distance, transformation=None):
    """Open3D evaluate_registration: correspondences/fitness/rmse at a fixed transformation (zero iterations)."""
    return registration_icp_batch([source], [target], max_correspondence_distance, [transformation],
                                  TransformationEstimationPointToPoint(), ICPConvergenceCriteria(max_iteration=0))[0]


def registration_colored_icp(source, target, max_correspondence_distance, init=None, estimation_method=None,
                             criteria=None):
    """Open3D pipelines.registration.registration_colored_icp (stitcher.py:94-102): colour gradient of
    the target (b2r_color_gradient) + ICP with a geometric and a photometric residual per correspondence."""
    estimation_method = estimation_method or TransformationEstimationForColoredICP()
    if estimation_method._kind != 2:
        raise RuntimeError("registration_colored_icp needs TransformationEstimationForColoredICP")
    return registration_icp_batch([source], [target], max_correspondence_distance, [init], estimation_method,
                                  criteria)[0]
