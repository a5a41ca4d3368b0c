"""`o3d.geometry` subset used by the reference's stitcher.py.

Arrays cross this boundary as numpy, like Open3D (points/colors/normals float64 [N,3], images
uint8/uint16/float32 [H,W(,C)]); the data lives on the B200 as torch tensors and every operation is
a kernel of libb200recon.so called through the C ABI.
"""
import ctypes as C

import numpy as np

from . import _lib, runtime
from .utility import Vector3dVector, Vector3iVector


# ----------------------------------------------------------------------------------------- images
class Image:
    """Open3D geometry.Image(ndarray) -- stitcher.py:34-35.  Keeps dtype; C-contiguous [H,W(,C)]."""

    def __init__(self, array=None):
        if array is None:
            array = np.zeros((0, 0), np.uint8)
        if isinstance(array, Image):
            array = array._a
        a = np.ascontiguousarray(array)
        if a.dtype not in (np.uint8, np.uint16, np.float32):
            raise RuntimeError("Image can only be initialized from buffer of uint8, uint16, or float!")
        if a.ndim not in (2, 3):
            raise RuntimeError("Image can only be initialized from 2D or 3D numpy array.")
        self._a = a
        self._dev = None

    @property
    def width(self):
        return self._a.shape[1] if self._a.ndim >= 2 else 0

    @property
    def height(self):
        return self._a.shape[0]

    @property
    def num_of_channels(self):
        return 1 if self._a.ndim == 2 else self._a.shape[2]

    @property
    def bytes_per_channel(self):
        return self._a.dtype.itemsize

    def __array__(self, dtype=None, copy=None):
        return self._a if dtype is None else self._a.astype(dtype)

    def is_empty(self):
        return self._a.size == 0

    def device_tensor(self):
        if self._dev is None:
            self._dev = runtime.to_device(self._a)
        return self._dev

    def __repr__(self):
        return "Image of size %dx%d, with %d channels.\nUse numpy.asarray to access buffer data." % (
            self.width, self.height, self.num_of_channels)


_DEPTH_DTYPE = {np.dtype(np.uint8): 0, np.dtype(np.uint16): 1, np.dtype(np.float32): 2}


class RGBDImage:
    """Open3D geometry.RGBDImage (color + float32-metres depth).

    `create_from_color_and_depth` is lazy: it validates, keeps the caller's host buffers and the conversion constants,
    and the float32-metres depth image is produced on the device by whoever needs it first --
    `PointCloud.create_from_rgbd_image`, `np.asarray(rgbd.depth)`, or never: `UniformTSDFVolume.integrate` hands the
    raw frame to `b2r_integrate_host_staged`, whose frame-preparation kernel does Open3D's conversion on the fly
    (one H2D copy of 5 B/pixel instead of a conversion launch + 4 B/pixel float image per frame)."""

    def __init__(self):
        self.color = Image()
        self.depth = Image()
        self._depth_dev = None  # float32 metres on device
        self._raw = None        # (raw depth ndarray, depth_scale, depth_trunc) while the depth is still unconverted

    @staticmethod
    def create_from_color_and_depth(color, depth, depth_scale=1000.0, depth_trunc=3.0, convert_rgb_to_intensity=True):
        """stitcher.py:38-42.  depth -> float32 metres: p /= scale; p >= trunc -> 0."""
        color = color if isinstance(color, Image) else Image(color)
        depth = depth if isinstance(depth, Image) else Image(depth)
        if color.height != depth.height or color.width != depth.width:
            raise RuntimeError("[CreateFromColorAndDepth] Unsupported image format.")
        if depth.num_of_channels != 1:
            raise RuntimeError("[CreateFromColorAndDepth] Unsupported image format.")
        rgbd = RGBDImage()
        if convert_rgb_to_intensity:
            a = color._a
            if a.ndim == 3 and a.shape[2] >= 3:  # Image::CreateFloatImage, weighted RGB
                if a.dtype == np.uint8:
                    g = (0.2990 * a[..., 0] + 0.5870 * a[..., 1] + 0.1140 * a[..., 2]) / 255.0
                else:
                    g = 0.2990 * a[..., 0] + 0.5870 * a[..., 1] + 0.1140 * a[..., 2]
            else:
                g = a.astype(np.float32) / (255.0 if a.dtype == np.uint8 else 1.0)
            rgbd.color = Image(g.astype(np.float32))
        else:
            rgbd.color = color
        rgbd._raw = (depth._a if depth._a.ndim == 2 else depth._a.reshape(depth.height, depth.width),
                     float(depth_scale), float(depth_trunc))
        rgbd.depth = _LazyDepthImage(rgbd)
        return rgbd

    def _depth_device(self):
        """float32 metres [H,W] on the device (Image.ConvertDepthToFloatImage, b2r_depth_to_float), made on first use."""
        if self._depth_dev is None and self._raw is not None:
            raw, scale, trunc = self._raw
            c = runtime.ctx()
            t = runtime.torch()
            H, W = raw.shape
            out = t.empty((H, W), dtype=t.float32, device=c.tdev)
            if H * W:
                d_in = runtime.to_device(raw)
                _lib.check(_lib.load().b2r_depth_to_float(c.handle, runtime.ptr(d_in), _DEPTH_DTYPE[raw.dtype], H, W,
                                                          scale, trunc, runtime.ptr(out), c.stream()))
            self._depth_dev = out
        return self._depth_dev

    def __repr__(self):
        return "RGBDImage of size \nColor image : %dx%d, with %d channels.\nDepth image : %dx%d, with %d channels." % (
            self.color.width, self.color.height, self.color.num_of_channels, self.depth.width, self.depth.height,
            self.depth.num_of_channels)


class _LazyDepthImage(Image):
    """The RGBD's depth image: float32 metres, converted on the device on first use, copied to the host only when
    somebody asks for the numpy buffer."""

    def __init__(self, rgbd):
        self._rgbd = rgbd
        self._host = None

    @property
    def _dev(self):
        return self._rgbd._depth_device()

    @property
    def _a(self):
        if self._host is None:
            self._host = self._dev.cpu().numpy()
        return self._host

    @property
    def width(self):
        return int(self._rgbd._raw[0].shape[1])

    @property
    def height(self):
        return int(self._rgbd._raw[0].shape[0])

    @property
    def num_of_channels(self):
        return 1

    @property
    def bytes_per_channel(self):
        return 4

    def device_tensor(self):
        return self._dev


def _depth_f32_device(rgbd):
    if isinstance(rgbd, RGBDImage) and rgbd._raw is not None:
        return rgbd._depth_device()
    if getattr(rgbd, "_depth_dev", None) is not None:
        return rgbd._depth_dev
    d = rgbd.depth
    if d.num_of_channels != 1 or d.bytes_per_channel != 4:
        raise RuntimeError("Unsupported image format.")
    return d.device_tensor()


# ----------------------------------------------------------------------------------------- search params
class KDTreeSearchParam:
    pass


class KDTreeSearchParamHybrid(KDTreeSearchParam):
    def __init__(self, radius, max_nn):
        self.radius = float(radius)
        self.max_nn = int(max_nn)


class KDTreeSearchParamKNN(KDTreeSearchParam):
    def __init__(self, knn=30):
        self.knn = int(knn)


class KDTreeSearchParamRadius(KDTreeSearchParam):
    def __init__(self, radius):
        self.radius = float(radius)


# ----------------------------------------------------------------------------------------- point cloud
class _DevArray:
    """float64 [N,3] attribute living on the device with a lazily synchronised numpy view."""

    def __init__(self):
        self.dev = None
        self.host = None

    def set_host(self, a):
        a = np.ascontiguousarray(np.asarray(a, dtype=np.float64)).reshape(-1, 3)
        self.host = a
        self.dev = None

    def set_dev(self, t):
        self.dev = t
        self.host = None

    def get_host(self):
        if self.host is None:
            self.host = np.zeros((0, 3)) if self.dev is None else self.dev.cpu().numpy()
        return self.host

    def get_dev(self):
        if self.dev is None:
            h = self.host if self.host is not None else np.zeros((0, 3))
            self.dev = runtime.to_device(h)
        return self.dev

    def __len__(self):
        if self.dev is not None:
            return int(self.dev.shape[0])
        return 0 if self.host is None else len(self.host)


class PointCloud:
    """Open3D geometry.PointCloud (points / colors / normals as float64 [N,3])."""

    def __init__(self, points=None):
        self._p, self._c, self._n = _DevArray(), _DevArray(), _DevArray()
        if points is not None:
            self.points = points

    # numpy-facing attributes (np.asarray(pcd.points) as in stitcher.py:146)
    @property
    def points(self):
        return Vector3dVector(self._p.get_host())

    @points.setter
    def points(self, v):
        self._p.set_host(np.asarray(v))

    @property
    def colors(self):
        return Vector3dVector(self._c.get_host())

    @colors.setter
    def colors(self, v):
        self._c.set_host(np.asarray(v))

    @property
    def normals(self):
        return Vector3dVector(self._n.get_host())

    @normals.setter
    def normals(self, v):
        self._n.set_host(np.asarray(v))

    def has_points(self):
        return len(self._p) > 0

    def has_colors(self):
        return len(self._p) > 0 and len(self._c) == len(self._p)

    def has_normals(self):
        return len(self._p) > 0 and len(self._n) == len(self._p)

    def is_empty(self):
        return not self.has_points()

    def __len__(self):
        return len(self._p)

    def __repr__(self):
        return "PointCloud with %d points." % len(self._p)

    @staticmethod
    def create_from_rgbd_image(image, intrinsic, extrinsic=None, project_valid_depth_only=True):
        """stitcher.py:45-46: row-major scan, invalid pixels dropped."""
        if not project_valid_depth_only:
            raise NotImplementedError("project_valid_depth_only=False is not on the reference's path")
        c = runtime.ctx()
        t = runtime.torch()
        depth = _depth_f32_device(image)
        H, W = int(depth.shape[0]), int(depth.shape[1])
        if (W, H) != (intrinsic.width, intrinsic.height):
            # Open3D does not check here; an intrinsic of another size would read out of bounds
            raise RuntimeError("[CreateFromRGBDImage] image size does not match the intrinsic")
        col = image.color
        dcol, nc = None, 0
        if col is not None and not col.is_empty():
            if col.bytes_per_channel == 1 and col.num_of_channels in (3, 4):
                dcol, nc = col.device_tensor(), col.num_of_channels
            elif col.bytes_per_channel == 4 and col.num_of_channels == 1:
                g = np.asarray(col)
                g8 = None  # float intensity: replicate on the host side below
            else:
                raise RuntimeError("[CreateFromRGBDImage] Unsupported image format.")
        pts = t.empty((H * W, 3), dtype=t.float64, device=c.tdev)
        cols = t.empty((H * W, 3), dtype=t.float64, device=c.tdev) if dcol is not None else None
        n = C.c_int64(0)
        _, ext = runtime.mat16(np.eye(4) if extrinsic is None else extrinsic)
        K = intrinsic._c()
        _lib.check(_lib.load().b2r_backproject(c.handle, runtime.ptr(depth), runtime.ptr(dcol), nc, C.byref(K), ext,
                                               runtime.ptr(pts), runtime.ptr(cols), C.byref(n), c.stream()))
        pcd = PointCloud()
        pcd._p.set_dev(pts[: n.value])
        if cols is not None:
            pcd._c.set_dev(cols[: n.value])
        elif col is not None and not col.is_empty() and col.bytes_per_channel == 4:
            d = depth.cpu().numpy()
            g = np.asarray(col)[d > 0].astype(np.float64)
            pcd._c.set_host(np.stack([g, g, g], 1))
        return pcd

    def transform(self, transformation):
        """stitcher.py:135 -- points T*[p,1] (divided by w), normals R*n."""
        T = np.asarray(transformation, np.float64).reshape(4, 4)
        t = runtime.torch()
        if len(self._p):
            P = self._p.get_dev()
            Tt = runtime.to_device(T)
            h = P @ Tt[:3, :3].T + Tt[:3, 3]
            w = P @ Tt[3, :3] + Tt[3, 3]
            self._p.set_dev(h / w[:, None])
            if self.has_normals():
                self._n.set_dev(self._n.get_dev() @ Tt[:3, :3].T)
        return self

    def __iadd__(self, other):
        """stitcher.py:136 -- Open3D keeps colors/normals only when both clouds have them."""
        t = runtime.torch()
        if len(other._p) == 0:      # Open3D's operator+= returns early: colours / normals of `self` survive
            return self
        n0 = len(self._p)
        keep_c = (n0 == 0 or self.has_colors()) and other.has_colors()
        keep_n = (n0 == 0 or self.has_normals()) and other.has_normals()
        newc = t.cat([self._c.get_dev(), other._c.get_dev()]) if keep_c and n0 else (other._c.get_dev() if keep_c else None)
        newn = t.cat([self._n.get_dev(), other._n.get_dev()]) if keep_n and n0 else (other._n.get_dev() if keep_n else None)
        self._p.set_dev(t.cat([self._p.get_dev(), other._p.get_dev()]) if n0 else other._p.get_dev().clone())
        self._c = _DevArray()
        self._n = _DevArray()
        if newc is not None:
            self._c.set_dev(newc)
        if newn is not None:
            self._n.set_dev(newn)
        return self

    def __add__(self, other):
        out = self.clone()
        out += other
        return out

    def clone(self):
        o = PointCloud()
        for a, b in ((self._p, o._p), (self._c, o._c), (self._n, o._n)):
            if len(a):
                b.set_dev(a.get_dev().clone())
        return o

    def __copy__(self):
        return self.clone()

    def __deepcopy__(self, memo):
        return self.clone()

    def paint_uniform_color(self, color):
        """stitcher.py:184."""
        t = runtime.torch()
        n = len(self._p)
        c = runtime.to_device(np.asarray(color, np.float64).reshape(1, 3))
        self._c.set_dev(c.expand(n, 3).contiguous())
        return self

    def select_by_index(self, indices, invert=False):
        t = runtime.torch()
        idx = runtime.to_device(np.asarray(indices, np.int64))
        n = len(self._p)
        if invert:
            m = t.ones(n, dtype=t.bool, device=idx.device)
            m[idx] = False
            idx = t.nonzero(m).squeeze(1)
        o = PointCloud()
        o._p.set_dev(self._p.get_dev()[idx])
        if self.has_colors():
            o._c.set_dev(self._c.get_dev()[idx])
        if self.has_normals():
            o._n.set_dev(self._n.get_dev()[idx])
        return o

    def get_min_bound(self):
        return self._p.get_dev().min(0).values.cpu().numpy() if len(self._p) else np.zeros(3)

    def get_max_bound(self):
        return self._p.get_dev().max(0).values.cpu().numpy() if len(self._p) else np.zeros(3)

    # voxel_down_sample / estimate_normals / remove_statistical_outlier are attached by cloud_ops.py


class TriangleMesh:
    """Open3D geometry.TriangleMesh{vertices, triangles, vertex_colors}."""

    def __init__(self):
        self._v = np.zeros((0, 3))
        self._t = np.zeros((0, 3), np.int32)
        self._vc = np.zeros((0, 3))
        self._vn = np.zeros((0, 3))

    @property
    def vertices(self):
        return Vector3dVector(self._v)

    @vertices.setter
    def vertices(self, v):
        self._v = np.ascontiguousarray(np.asarray(v, np.float64)).reshape(-1, 3)

    @property
    def triangles(self):
        return Vector3iVector(self._t)

    @triangles.setter
    def triangles(self, v):
        self._t = np.ascontiguousarray(np.asarray(v, np.int32)).reshape(-1, 3)

    @property
    def vertex_colors(self):
        return Vector3dVector(self._vc)

    @vertex_colors.setter
    def vertex_colors(self, v):
        self._vc = np.ascontiguousarray(np.asarray(v, np.float64)).reshape(-1, 3)

    @property
    def vertex_normals(self):
        return Vector3dVector(self._vn)

    def has_vertices(self):
        return len(self._v) > 0

    def has_triangles(self):
        return len(self._v) > 0 and len(self._t) > 0

    def has_vertex_colors(self):
        return len(self._v) > 0 and len(self._vc) == len(self._v)

    def compute_vertex_normals(self, normalized=True):
        v, tr = self._v, self._t
        n = np.zeros_like(v)
        if len(tr):
            fn = np.cross(v[tr[:, 1]] - v[tr[:, 0]], v[tr[:, 2]] - v[tr[:, 0]])
            for k in range(3):
                np.add.at(n, tr[:, k], fn)
            if normalized:
                ln = np.linalg.norm(n, axis=1, keepdims=True)
                n = np.where(ln > 0, n / np.maximum(ln, 1e-300), n)
        self._vn = n
        return self

    def __repr__(self):
        return "TriangleMesh with %d points and %d triangles." % (len(self._v), len(self._t))
