"""`o3d.utility` subset (stitcher.py:148)."""
import numpy as np


class _VectorBase(np.ndarray):
    _dtype = np.float64
    _cols = 3

    def __new__(cls, array=None):
        if array is None:
            array = np.zeros((0, cls._cols), cls._dtype)
        a = np.ascontiguousarray(np.asarray(array, dtype=cls._dtype)).reshape(-1, cls._cols)
        return a.view(cls)

    def __repr__(self):
        return "std::vector<Eigen::Vector%d%s> with %d elements.\nUse numpy.asarray() to access data." % (
            self._cols, "d" if self._dtype == np.float64 else "i", len(self))


class Vector3dVector(_VectorBase):
    """float64 [N,3]; np.asarray(v) gives the buffer."""


class Vector3iVector(_VectorBase):
    _dtype = np.int32


class Vector2iVector(_VectorBase):
    _dtype = np.int32
    _cols = 2


class DoubleVector(np.ndarray):
    def __new__(cls, array=None):
        a = np.ascontiguousarray(np.asarray([] if array is None else array, dtype=np.float64)).reshape(-1)
        return a.view(cls)
