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
