"""`o3d.pipelines.integration`: UniformTSDFVolume and ScalableTSDFVolume on the B200 (north_star API; SURVEY.md
A.3/A.4, rows a5/a6/f4).

    vol = UniformTSDFVolume(length, resolution, sdf_trunc, color_type, origin)
    vol.integrate(rgbd, intrinsic, extrinsic)
    mesh = vol.extract_triangle_mesh();  pcd = vol.extract_point_cloud()

Multi-GPU: pass `x_range=(x_begin, x_end)` to own one slab (a range of x planes) of the cube; `parallel.SlabVolume` wires
the slabs of all ranks together (halo exchange + gather over NCCL).
"""
import ctypes as C
import enum

import numpy as np

from .. import _lib, geometry, runtime


class TSDFVolumeColorType(enum.IntEnum):
    NoColor = 0
    RGB8 = 1
    Gray32 = 2


class TSDFVolume:
    pass


class UniformTSDFVolume(TSDFVolume):
    def __init__(self, length, resolution, sdf_trunc, color_type=TSDFVolumeColorType.RGB8, origin=(0.0, 0.0, 0.0),
                 x_range=None, device=None, context=None):
        """`x_range` / `device` / `context` are extensions (multi-GPU slabs, several slabs in one process)."""
        self.length = float(length)
        self.resolution = int(resolution)
        self.sdf_trunc = float(sdf_trunc)
        self.color_type = TSDFVolumeColorType(int(color_type))
        self.origin = np.ascontiguousarray(origin, np.float64).reshape(3)
        self.voxel_length = self.length / self.resolution
        if self.resolution <= 0 or self.length <= 0 or self.sdf_trunc <= 0:
            raise RuntimeError("UniformTSDFVolume: length, resolution and sdf_trunc must be positive")
        zb, ze = (0, self.resolution) if x_range is None else (int(x_range[0]), int(x_range[1]))
        if not (0 <= zb <= ze <= self.resolution):
            raise RuntimeError("UniformTSDFVolume: bad x_range %r" % (x_range,))
        self.x_begin, self.x_end = zb, ze
        self._ctx = context if context is not None else runtime.ctx(device)
        t = runtime.torch()
        R = self.resolution
        planes = (ze - zb) + _lib.PAD_LO + _lib.PAD_HI
        # {tsdf, weight} pairs (8 B/voxel) + packed colour (4 B/voxel), z-slowest / y-fastest
        self._tsdfw = t.zeros((planes, R, R, 2), dtype=t.float32, device=self._ctx.tdev)
        self._color = None
        if self.color_type != TSDFVolumeColorType.NoColor:
            self._color = t.zeros((planes, R, R), dtype=t.int32, device=self._ctx.tdev)
        # rows (plane, x) that may hold a voxel with weight != 0: extraction only reads cube rows next to one
        self._rows = t.zeros((planes, R), dtype=t.uint8, device=self._ctx.tdev)
        self._desc = _lib.Volume(R, zb, ze, int(self.color_type), self.voxel_length, self.sdf_trunc,
                                 (C.c_double * 3)(*self.origin), self._tsdfw.data_ptr(),
                                 0 if self._color is None else self._color.data_ptr(), self._rows.data_ptr())
        self._updates = t.zeros(1, dtype=t.int64, device=self._ctx.tdev)
        self.count_updates = False

    # ------------------------------------------------------------------ integrate
    _RAW_FRAME_PATH = True      # integrate() may hand an unconverted RGBDImage to b2r_integrate_host_staged

    def integrate(self, image, intrinsic, extrinsic):
        """Open3D UniformTSDFVolume.integrate(rgbd, intrinsic, extrinsic); extrinsic = world->camera.

        An RGBDImage fresh from `create_from_color_and_depth` (stitcher.py:38-42) still holds the caller's raw host
        frame: it goes through `b2r_integrate_host_staged` (pinned staging + one H2D DMA + conversion inside the
        frame-preparation kernel).  An RGBDImage whose float depth already exists on the device takes
        `b2r_integrate`."""
        raw = getattr(image, "_raw", None)
        if (self._RAW_FRAME_PATH and raw is not None and image._depth_dev is None
                and self.color_type != TSDFVolumeColorType.Gray32):
            d, scale, trunc = raw
            H, W = d.shape
            if (W, H) != (intrinsic.width, intrinsic.height):
                raise RuntimeError("[UniformTSDFVolume::Integrate] Unsupported image format.")
            col = None
            if self.color_type == TSDFVolumeColorType.RGB8:
                ci = image.color
                if ci.num_of_channels != 3 or ci.bytes_per_channel != 1 or (ci.width, ci.height) != (W, H):
                    raise RuntimeError("[UniformTSDFVolume::Integrate] Unsupported image format.")
                col = ci._a
            self._integrate_host_checked(d, col, intrinsic, extrinsic, scale, trunc, staged=True)
            return
        depth = geometry._depth_f32_device(image)
        H, W = int(depth.shape[0]), int(depth.shape[1])
        if (W, H) != (intrinsic.width, intrinsic.height):
            raise RuntimeError("[UniformTSDFVolume::Integrate] Unsupported image format.")
        dcol = None
        if self.color_type == TSDFVolumeColorType.RGB8:
            col = image.color
            if col.num_of_channels != 3 or col.bytes_per_channel != 1 or (col.width, col.height) != (W, H):
                raise RuntimeError("[UniformTSDFVolume::Integrate] Unsupported image format.")
            dcol = col.device_tensor()
        elif self.color_type == TSDFVolumeColorType.Gray32:
            col = image.color
            if col.num_of_channels != 1 or col.bytes_per_channel != 4 or (col.width, col.height) != (W, H):
                raise RuntimeError("[UniformTSDFVolume::Integrate] Unsupported image format.")
            dcol = col.device_tensor()
        self.integrate_device(depth, dcol, intrinsic, extrinsic)

    def integrate_device(self, depth_f32, color, intrinsic, extrinsic):
        """Same with frame buffers already resident in HBM (torch CUDA tensors)."""
        c = self._ctx
        _, ext = runtime.mat16(extrinsic)
        K = intrinsic._c()
        upd = runtime.ptr(self._updates) if self.count_updates else C.c_void_p(0)
        _lib.check(_lib.load().b2r_integrate(c.handle, C.byref(self._desc), runtime.ptr(depth_f32), runtime.ptr(color),
                                             C.byref(K), ext, upd, c.stream()))

    def integrate_host(self, depth, color, intrinsic, extrinsic, depth_scale=1000.0, depth_trunc=3.0):
        """End-to-end entry: raw HOST frame -> H2D copy + depth conversion (stitcher.py:40-41 constants) + integrate.
        depth: uint8/uint16/float32 [H,W]; colour: uint8 [H,W,3] (RGB8 volumes; ignored by NoColor volumes).

        * numpy arrays (pageable): copied into the library's pinned staging ring first
          (`b2r_integrate_host_staged`); the arrays may be reused as soon as the call returns.
        * pinned torch CPU tensors: DMA'd straight from the caller's memory (`b2r_integrate_host`, no host copy).
          The volume keeps a reference to them until the copy has completed, but the caller must not MODIFY them
          before then (`vol.updates()`, `ctx.sync()` or any synchronisation of the current stream).
        Shape, dtype and contiguity are checked against the intrinsic like Open3D's Integrate does
        ("Unsupported image format")."""
        t = runtime.torch()
        if self.color_type == TSDFVolumeColorType.Gray32:
            raise RuntimeError("[UniformTSDFVolume::Integrate] Unsupported image format. (Gray32 volumes take a float "
                               "colour image through integrate())")
        pinned = isinstance(depth, t.Tensor)
        if color is not None and pinned != isinstance(color, t.Tensor):
            raise RuntimeError("integrate_host: depth and colour must both be numpy arrays or both torch CPU tensors")
        if pinned:
            if depth.is_cuda or (color is not None and color.is_cuda):
                raise RuntimeError("integrate_host takes HOST frames; use integrate_device for CUDA tensors")
            if not depth.is_pinned() or (color is not None and self.color_type == TSDFVolumeColorType.RGB8 and not color.is_pinned()):
                depth = depth.numpy()           # pageable torch tensors: same path as numpy
                color = None if color is None else color.numpy()
                pinned = False
        self._integrate_host_checked(depth, color, intrinsic, extrinsic, depth_scale, depth_trunc, staged=not pinned)

    _HOST_DT_NP = {np.dtype(np.uint8): 0, np.dtype(np.uint16): 1, np.dtype(np.float32): 2}

    def _integrate_host_checked(self, depth, color, intrinsic, extrinsic, depth_scale, depth_trunc, staged):
        t = runtime.torch()
        c = self._ctx
        H, W = int(intrinsic.height), int(intrinsic.width)
        bad = RuntimeError("[UniformTSDFVolume::Integrate] Unsupported image format.")
        need_color = self.color_type == TSDFVolumeColorType.RGB8
        if staged:
            if not isinstance(depth, np.ndarray) or depth.dtype not in self._HOST_DT_NP or depth.shape not in ((H, W), (H, W, 1)):
                raise bad
            depth = np.ascontiguousarray(depth)
            dcode, dptr = self._HOST_DT_NP[depth.dtype], depth.ctypes.data
            cptr = 0
            if need_color:
                if not isinstance(color, np.ndarray) or color.dtype != np.uint8 or color.shape != (H, W, 3):
                    raise bad
                color = np.ascontiguousarray(color)
                cptr = color.ctypes.data
            fn = _lib.load().b2r_integrate_host_staged
        else:
            dt = {t.uint8: 0, t.uint16: 1, t.float32: 2}
            if depth.dtype not in dt or tuple(depth.shape) not in ((H, W), (H, W, 1)) or not depth.is_contiguous():
                raise bad
            dcode, dptr = dt[depth.dtype], depth.data_ptr()
            cptr = 0
            if need_color:
                if color is None or color.dtype != t.uint8 or tuple(color.shape) != (H, W, 3) or not color.is_contiguous():
                    raise bad
                cptr = color.data_ptr()
            fn = _lib.load().b2r_integrate_host
            self._hold(depth, color if need_color else None)
        _, ext = runtime.mat16(extrinsic)
        K = intrinsic._c()
        upd = runtime.ptr(self._updates) if self.count_updates else C.c_void_p(0)
        _lib.check(fn(c.handle, C.byref(self._desc), C.c_void_p(dptr), dcode, C.c_void_p(cptr), C.byref(K), ext,
                      float(depth_scale), float(depth_trunc), upd, c.stream()))

    # Pinned caller frames are read by an asynchronous DMA after integrate_host has returned: keep them alive until an
    # event recorded on the current stream BEHIND the call (the stream waits for the copy) has completed.  One event
    # per _HOLD_EVERY frames, not per frame.
    _HOLD_EVERY = 16

    def _hold(self, *tensors):
        t = runtime.torch()
        if not hasattr(self, "_held"):
            self._held, self._guards = [], []
        self._held.append(tensors)
        if len(self._held) >= self._HOLD_EVERY:
            # the event is recorded at the start of the NEXT call, i.e. behind every frame in `held`
            ev = t.cuda.Event()
            ev.record(t.cuda.current_stream(self._ctx.tdev))
            # (this call's own frame was appended above but its copy is not queued yet: it stays in the new list)
            self._guards.append((ev, self._held[:-1]))
            self._held = self._held[-1:]
            while self._guards and self._guards[0][0].query():
                self._guards.pop(0)

    def _release_held(self):
        """after a host-side synchronisation that covers every queued copy"""
        if hasattr(self, "_held"):
            self._held, self._guards = [], []

    def updates(self):
        """Number of voxel updates counted since the last reset (needs count_updates=True)."""
        self._ctx.join()
        n = int(self._updates.item())       # blocks on the current stream, which is ordered after every copy
        self._release_held()
        return n

    def __del__(self):
        # the volume's tensors go back to torch's allocator: no streaming kernel may still be writing them
        try:
            self._ctx.join()
        except Exception:
            pass

    def reset(self):
        c = self._ctx
        _lib.check(_lib.load().b2r_volume_reset(c.handle, C.byref(self._desc), c.stream()))   # joins the side stream
        self._updates.zero_()

    # ------------------------------------------------------------------ raw access (tests, persistence)
    def export_arrays(self):
        """(tsdf, weight, color) as numpy in Open3D order [x_local, y, z] (the slab's x planes); colour [.., 3] in
        0..255 or None."""
        c = self._ctx
        t = runtime.torch()
        R, nx = self.resolution, self.x_end - self.x_begin
        ts = t.empty((nx, R, R), dtype=t.float32, device=c.tdev)
        w = t.empty((nx, R, R), dtype=t.float32, device=c.tdev)
        col = t.empty((nx, R, R, 3), dtype=t.float32, device=c.tdev) if self._color is not None else None
        _lib.check(_lib.load().b2r_volume_export(c.handle, C.byref(self._desc), runtime.ptr(ts), runtime.ptr(w),
                                                 runtime.ptr(col), c.stream()))
        return ts.cpu().numpy(), w.cpu().numpy(), (None if col is None else col.cpu().numpy())

    def import_arrays(self, tsdf, weight, color=None):
        c = self._ctx
        t = runtime.torch()
        R, nx = self.resolution, self.x_end - self.x_begin
        ts = runtime.to_device(np.ascontiguousarray(tsdf, np.float32).reshape(nx, R, R))
        w = runtime.to_device(np.ascontiguousarray(weight, np.float32).reshape(nx, R, R))
        col = None if color is None else runtime.to_device(np.ascontiguousarray(color, np.float32).reshape(nx, R, R, 3))
        _lib.check(_lib.load().b2r_volume_import(c.handle, C.byref(self._desc), runtime.ptr(ts), runtime.ptr(w),
                                                 runtime.ptr(col), c.stream()))
        c.sync()

    def save(self, path):
        """Persist the slab (checkpoint/resume of the volume)."""
        ts, w, col = self.export_arrays()
        np.savez(path, tsdf=ts, weight=w, color=(np.zeros(0) if col is None else col), length=self.length,
                 resolution=self.resolution, sdf_trunc=self.sdf_trunc, color_type=int(self.color_type),
                 origin=self.origin, x_range=np.array([self.x_begin, self.x_end]))

    @staticmethod
    def load(path, device=None):
        d = np.load(path)
        v = UniformTSDFVolume(float(d["length"]), int(d["resolution"]), float(d["sdf_trunc"]),
                              TSDFVolumeColorType(int(d["color_type"])), d["origin"],
                              x_range=tuple(int(x) for x in d["x_range"]), device=device)
        v.import_arrays(d["tsdf"], d["weight"], d["color"] if d["color"].size else None)
        return v

    # ------------------------------------------------------------------ extraction
    def _mesh_count(self):
        c = self._ctx
        nv, nt = C.c_int64(0), C.c_int64(0)
        _lib.check(_lib.load().b2r_extract_mesh_count(c.handle, C.byref(self._desc), C.byref(nv), C.byref(nt),
                                                      c.stream()))
        return nv.value, nt.value

    def _mesh_vertices(self, nv, vertex_base=0):
        c = self._ctx
        t = runtime.torch()
        V = t.empty((nv, 3), dtype=t.float64, device=c.tdev)
        Cc = t.empty((nv, 3), dtype=t.float64, device=c.tdev) if self._color is not None else None
        _lib.check(_lib.load().b2r_extract_mesh_vertices(c.handle, C.byref(self._desc), int(vertex_base), runtime.ptr(V),
                                                         runtime.ptr(Cc), c.stream()))
        return V, Cc

    def _mesh_triangles(self, nt):
        c = self._ctx
        t = runtime.torch()
        T = t.empty((nt, 3), dtype=t.int32, device=c.tdev)
        _lib.check(_lib.load().b2r_extract_mesh_triangles(c.handle, C.byref(self._desc), runtime.ptr(T), c.stream()))
        return T

    def extract_triangle_mesh_device(self):
        """(vertices f64[V,3], triangles i32[T,3], colors f64[V,3]|None) as CUDA tensors."""
        nv, nt = self._mesh_count()
        V, Cc = self._mesh_vertices(nv)
        T = self._mesh_triangles(nt)
        return V, T, Cc

    def extract_triangle_mesh(self):
        """Open3D UniformTSDFVolume.extract_triangle_mesh().  Vertex ORDER differs from
        Open3D's first-touch scan order (ours is memory order); the vertex/triangle SETS are identical."""
        V, T, Cc = self.extract_triangle_mesh_device()
        m = geometry.TriangleMesh()
        m.vertices = V.cpu().numpy()
        m.triangles = T.cpu().numpy()
        if Cc is not None:
            m.vertex_colors = Cc.cpu().numpy()
        return m

    def extract_point_cloud_device(self):
        c = self._ctx
        t = runtime.torch()
        n = C.c_int64(0)
        _lib.check(_lib.load().b2r_extract_points_count(c.handle, C.byref(self._desc), C.byref(n), c.stream()))
        n = n.value
        P = t.empty((n, 3), dtype=t.float64, device=c.tdev)
        N = t.empty((n, 3), dtype=t.float64, device=c.tdev)
        Cc = t.empty((n, 3), dtype=t.float64, device=c.tdev) if self._color is not None else None
        _lib.check(_lib.load().b2r_extract_points_emit(c.handle, C.byref(self._desc), runtime.ptr(P), runtime.ptr(N),
                                                       runtime.ptr(Cc), c.stream()))
        return P, N, Cc

    def extract_point_cloud(self):
        """Open3D UniformTSDFVolume.extract_point_cloud()."""
        P, N, Cc = self.extract_point_cloud_device()
        pcd = geometry.PointCloud()
        pcd._p.set_dev(P)
        pcd._n.set_dev(N)
        if Cc is not None:
            pcd._c.set_dev(Cc)
        return pcd


class ScalableTSDFVolume(UniformTSDFVolume):
    """Open3D `ScalableTSDFVolume(voxel_length, sdf_trunc, color_type, volume_unit_resolution=16,
    depth_sampling_stride=4)` (SURVEY.md A.3 last bullet, row f4): per frame only the 16^3-voxel units within
    sdf_trunc of a (stride-4) depth sample are integrated, each with its own origin and its own float32 z-chain.

    B200 design: the units are NOT hashed.  They live at their place in a dense backing cube in HBM (a unit that was
    never opened holds weight 0, which every extraction rule of Open3D treats exactly like a missing unit), so mesh /
    point extraction, persistence and the slab machinery are the dense kernels unchanged, and "scalable" is the
    per-frame list of touched units (`b2r_integrate_units`).  Open3D's volume is unbounded; this one covers
    `resolution`^3 voxels (multiple of the unit size) whose first unit is `unit_offset` (default: cube centred on the
    world origin; `origin=` snaps down to a unit boundary).  Unit touches outside the cube are counted
    (`units_clipped()`), never silently dropped.

    Differences from Open3D: vertex / point order (memory order instead of hash-iteration order); voxels on the
    outermost layer of the backing cube produce no points (the dense extraction's border rule)."""

    _RAW_FRAME_PATH = False     # the unit kernels take the converted float depth image (b2r_integrate_units)

    def __init__(self, voxel_length, sdf_trunc, color_type=TSDFVolumeColorType.RGB8, volume_unit_resolution=16,
                 depth_sampling_stride=4, *, resolution=512, origin=None, unit_offset=None, device=None):
        self.volume_unit_resolution = int(volume_unit_resolution)
        self.depth_sampling_stride = int(depth_sampling_stride)
        ur = self.volume_unit_resolution
        if not (1 <= ur <= 16):
            raise RuntimeError("ScalableTSDFVolume: volume_unit_resolution must be in [1, 16] (got %d)" % ur)
        if self.depth_sampling_stride < 1:
            raise RuntimeError("ScalableTSDFVolume: depth_sampling_stride must be >= 1")
        if int(resolution) % ur:
            raise RuntimeError("ScalableTSDFVolume: resolution must be a multiple of volume_unit_resolution")
        self.volume_unit_length = float(voxel_length) * ur
        nu = int(resolution) // ur
        if unit_offset is None:
            if origin is None:
                unit_offset = (-(nu // 2),) * 3
            else:
                unit_offset = tuple(int(np.floor(float(o) / self.volume_unit_length + 1e-9)) for o in origin)
        self.unit_offset = np.ascontiguousarray(unit_offset, np.int32).reshape(3)
        self.units_per_side = nu
        super().__init__(float(voxel_length) * int(resolution), int(resolution), sdf_trunc, color_type,
                         self.unit_offset.astype(np.float64) * self.volume_unit_length, device=device)
        self.voxel_length = float(voxel_length)
        self._desc.voxel_length = self.voxel_length          # exactly the constructor argument, not length / R
        t = runtime.torch()
        self._unit_stamp = t.zeros(nu ** 3, dtype=t.int32, device=self._ctx.tdev)
        self._unit_list = t.empty(nu ** 3, dtype=t.int32, device=self._ctx.tdev)
        self._unit_counters = t.zeros(4, dtype=t.int64, device=self._ctx.tdev)
        self._frame_id = 0

    def integrate_device(self, depth_f32, color, intrinsic, extrinsic):
        c = self._ctx
        E, ext = runtime.mat16(extrinsic)
        P, pose = runtime.mat16(np.linalg.inv(E))      # Open3D: camera_pose = extrinsic.inverse(), float64
        K = intrinsic._c()
        upd = runtime.ptr(self._updates) if self.count_updates else C.c_void_p(0)
        self._frame_id += 1
        off = (C.c_int32 * 3)(*[int(v) for v in self.unit_offset])
        _lib.check(_lib.load().b2r_integrate_units(
            c.handle, C.byref(self._desc), runtime.ptr(depth_f32), runtime.ptr(color), C.byref(K), ext, pose,
            self.volume_unit_resolution, self.depth_sampling_stride, off, runtime.ptr(self._unit_stamp),
            self._frame_id, runtime.ptr(self._unit_list), runtime.ptr(self._unit_counters), upd, c.stream()))

    def integrate_host(self, depth, color, intrinsic, extrinsic, depth_scale=1000.0, depth_trunc=3.0):
        """Raw HOST frame -> H2D + Open3D's depth conversion + integrate (same contract as the uniform volume)."""
        img = geometry.RGBDImage.create_from_color_and_depth(
            geometry.Image(np.ascontiguousarray(color)) if color is not None else geometry.Image(np.zeros(depth.shape + (3,), np.uint8)),
            geometry.Image(np.ascontiguousarray(depth)), depth_scale=depth_scale, depth_trunc=depth_trunc,
            convert_rgb_to_intensity=False)
        self.integrate(img, intrinsic, extrinsic)

    def extract_point_cloud_device(self):
        """ScalableTSDFVolume::ExtractPointCloud: world-coordinate interpolation from the unit origins and the
        scalable GetTSDFAt for the normals (`b2r_extract_points_emit_units`)."""
        c = self._ctx
        t = runtime.torch()
        n = C.c_int64(0)
        _lib.check(_lib.load().b2r_extract_points_count(c.handle, C.byref(self._desc), C.byref(n), c.stream()))
        n = n.value
        P = t.empty((n, 3), dtype=t.float64, device=c.tdev)
        N = t.empty((n, 3), dtype=t.float64, device=c.tdev)
        Cc = t.empty((n, 3), dtype=t.float64, device=c.tdev) if self._color is not None else None
        off = (C.c_int32 * 3)(*[int(v) for v in self.unit_offset])
        _lib.check(_lib.load().b2r_extract_points_emit_units(
            c.handle, C.byref(self._desc), self.volume_unit_resolution, off, runtime.ptr(self._unit_stamp),
            runtime.ptr(P), runtime.ptr(N), runtime.ptr(Cc), c.stream()))
        return P, N, Cc

    def reset(self):
        super().reset()
        self._unit_stamp.zero_()
        self._unit_counters.zero_()
        self._frame_id = 0

    # -- unit bookkeeping (device counters, read on demand: integrate itself never synchronises)
    def _counters(self):
        self._ctx.join()
        return [int(v) for v in self._unit_counters.cpu().tolist()]

    def units_touched_last_frame(self):
        return self._counters()[0]

    def units_clipped(self):
        """unit touches that fell outside the backing cube so far (Open3D would have opened those units)"""
        return self._counters()[1]

    def units_opened(self):
        return self._counters()[2]

    def unit_keys(self):
        """global integer indices of the opened units, int32 [n, 3] (Open3D's volume_units_ keys)"""
        self._ctx.join()
        nu = self.units_per_side
        lin = runtime.torch().nonzero(self._unit_stamp).flatten().cpu().numpy()
        k = np.stack([lin // (nu * nu), (lin // nu) % nu, lin % nu], 1).astype(np.int32)
        return k + self.unit_offset[None, :]

    def save(self, path):
        raise NotImplementedError("ScalableTSDFVolume.save: persist through export_arrays()/unit_keys()")
