"""`o3d.camera` subset used by the reference (stitcher.py:264-267, main.py:241-244)."""
import numpy as np

from . import _lib


class PinholeCameraIntrinsic:
    """Open3D camera.PinholeCameraIntrinsic(width, height, fx, fy, cx, cy)."""

    def __init__(self, width=-1, height=-1, fx=0.0, fy=0.0, cx=0.0, cy=0.0):
        if isinstance(width, PinholeCameraIntrinsic):
            o = width
            width, height, fx, fy, cx, cy = o.width, o.height, *o.get_focal_length(), *o.get_principal_point()
        self.width = int(width)
        self.height = int(height)
        self.intrinsic_matrix = np.array([[fx, 0.0, cx], [0.0, fy, cy], [0.0, 0.0, 1.0]], dtype=np.float64)

    def set_intrinsics(self, width, height, fx, fy, cx, cy):
        self.__init__(width, height, fx, fy, cx, cy)

    def get_focal_length(self):
        return float(self.intrinsic_matrix[0, 0]), float(self.intrinsic_matrix[1, 1])

    def get_principal_point(self):
        return float(self.intrinsic_matrix[0, 2]), float(self.intrinsic_matrix[1, 2])

    def get_skew(self):
        return float(self.intrinsic_matrix[0, 1])

    def is_valid(self):
        return self.width > 0 and self.height > 0

    def _c(self):
        fx, fy = self.get_focal_length()
        cx, cy = self.get_principal_point()
        return _lib.Intrinsic(self.width, self.height, fx, fy, cx, cy)

    def __repr__(self):
        return "PinholeCameraIntrinsic with width = %d and height = %d.\nAccess intrinsics with intrinsic_matrix." % (
            self.width, self.height)
