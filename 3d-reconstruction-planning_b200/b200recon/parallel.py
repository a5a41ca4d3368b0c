"""Multi-GPU layer: one process per GPU, `torch.distributed` for the plumbing.

  * `SlabVolume`     -- a dense R^3 TSDF cube sharded into slabs of x planes (the cube is stored x-slowest, so a
                        slab and its halo planes are contiguous, and a slab owns WHOLE (x, y) columns: Open3D's
                        float32 z-chain and the per-column planning of a frame run once per column whatever the number
                        of ranks), one per rank.  Every rank integrates
                        every frame into its own slab: NO data-path collective during integration (the
                        float32 z-chain of each slab starts with the replay of the planes below it, so every
                        slab is bit-identical to the single-GPU volume).
  * halo exchange    -- before extraction each rank sends its first B2R_PAD_HI owned planes to the rank
                        below and its last B2R_PAD_LO owned planes to the rank above (the volume is stored
                        z-slowest, so a plane is one contiguous message -- no pack kernel).
  * mesh gather      -- per-rank vertex/triangle counts -> all_gather -> global vertex bases; one more
                        seam message (the vertex-id table of the first owned plane goes to the rank below);
                        vertices / triangles are sent to rank 0.
  * `icp_round_robin`-- batched ICP pairs, pair p -> rank p % world, no collective on the data path.

The collectives are written against `torch.distributed` only (NCCL over NVLink on the GPUs; the same code
runs on the `gloo` backend with CPU tensors, which is how the host logic is tested without a GPU).
"""
import numpy as np

from . import _lib


def _dist():
    import torch.distributed as dist
    return dist


def world_info(group=None):
    """(rank, world) of the default (or given) process group; (0, 1) when torch.distributed is not initialised."""
    dist = _dist()
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(group), dist.get_world_size(group)
    return 0, 1


def slab_range(resolution, rank, world):
    """Planes [x_begin, x_end) owned by `rank`: contiguous, balanced to within one plane."""
    R = int(resolution)
    return R * rank // world, R * (rank + 1) // world


def slab_ranges(resolution, world):
    return [slab_range(resolution, r, world) for r in range(world)]


def balanced_slab_ranges(plane_work, world, min_planes=4):
    """Contiguous slabs (x planes) with (nearly) equal WORK instead of equal plane counts.  Integration only touches the
    voxels between the cameras and the surfaces, which on a table-top orbit is the lower part of the cube: equal
    slabs leave most ranks idle.  `plane_work[z]` = expected voxel updates per frame sequence in plane z (see
    `probe_plane_work`); returns [(x_begin, x_end)] per rank, every slab at least `min_planes` thick (halo)."""
    w = np.asarray(plane_work, np.float64).reshape(-1)
    R = len(w)
    if world * min_planes > R:
        raise RuntimeError("%d planes cannot give %d slabs of >= %d planes" % (R, world, min_planes))
    w = np.maximum(w, 0.0)
    w = w + (w.sum() * 1e-3 / R + 1e-30)      # a small per-plane cost: empty regions are still split, not piled on one rank
    c = np.concatenate([[0.0], np.cumsum(w)])
    cuts = [0]
    for k in range(1, world):
        z = int(np.searchsorted(c, c[-1] * k / world))
        z = max(z, cuts[-1] + min_planes)
        z = min(z, R - (world - k) * min_planes)
        cuts.append(z)
    cuts.append(R)
    return [(cuts[i], cuts[i + 1]) for i in range(world)]


def rebalance_slab_ranges(ranges, plane_work, times, min_planes=4):
    """One feedback step on top of `balanced_slab_ranges`: `times[r]` = measured integration time of rank r on
    its slab `ranges[r]` (any unit).  Updates are not the whole cost -- slabs under the surfaces still sample voxels
    they then reject, slabs near a camera have half-empty tiles -- so the per-plane work estimate of every slab is
    rescaled to the time that slab really took, and the cube is cut again into slabs of equal estimated time."""
    w = np.maximum(np.asarray(plane_work, np.float64).reshape(-1), 0.0)
    w = w + (w.sum() * 1e-3 / len(w) + 1e-30)
    cost = np.zeros_like(w)
    for (a, b), t in zip(ranges, times):
        cost[a:b] = w[a:b] * (float(t) / w[a:b].sum())
    return balanced_slab_ranges(cost, len(ranges), min_planes)


def agree_plane_work(plane_work, device=None, group=None):
    """A per-plane cost profile that was MEASURED (profile_plane_cost: kernel timings) differs from rank to rank.
    Slab boundaries cut from per-rank numbers would not tile the cube, so the ranks agree on the mean profile first
    (one all_reduce; it also averages the timing noise).  Without a process group: returned unchanged."""
    rank, world = world_info(group)
    w = np.asarray(plane_work, np.float64)
    if world == 1:
        return w
    import torch
    t = torch.as_tensor(w.copy(), device=device if device is not None else "cpu")
    _dist().all_reduce(t, group=group)
    return (t / world).cpu().numpy()


def check_same_ranges(ranges, resolution, device=None, group=None):
    """Raises unless `ranges` tiles [0, resolution) and every rank of the group holds the same list."""
    rank, world = world_info(group)
    b = [int(a) for a, _ in ranges] + [int(ranges[-1][1])]
    if b[0] != 0 or b[-1] != int(resolution) or any(int(ranges[i][1]) != int(ranges[i + 1][0]) for i in range(len(ranges) - 1)) \
            or any(b[i] > b[i + 1] for i in range(len(b) - 1)):
        raise RuntimeError("slab ranges %r do not tile [0, %d)" % (ranges, resolution))
    if world == 1:
        return
    import torch
    mine = torch.tensor(b, dtype=torch.int64, device=device if device is not None else "cpu")
    allb = [torch.zeros_like(mine) for _ in range(world)]
    _dist().all_gather(allb, mine, group=group)
    if not all(torch.equal(mine, o) for o in allb):
        raise RuntimeError("slab boundaries differ between ranks: %r" % [o.tolist() for o in allb])


def profile_plane_cost(length, resolution, sdf_trunc, origin, frames, intrinsic, plane_work, chunk=32, color_type=None,
                       device=None):
    """Measured cost profile for `balanced_slab_ranges`: times the streaming kernel of `frames` -- an iterable of
    (depth_f32 CUDA tensor, colour CUDA tensor or None, extrinsic) -- on every `chunk`-plane slab of the cube by itself,
    subtracts the fixed per-launch floor (the cheapest slab) and spreads each slab's time over its planes in
    proportion to `plane_work` (uniformly where the slab has no updates at all: planes under a surface are sampled and
    rejected, which costs time but shows up in no update count).  Setup-time calibration, a second or two; every rank
    measures the same profile on its own GPU, no collective."""
    from . import runtime
    from .pipelines import integration
    R = int(resolution)
    ct = integration.TSDFVolumeColorType.RGB8 if color_type is None else color_type
    c = runtime.ctx(device)
    t = runtime.torch()
    w = np.maximum(np.asarray(plane_work, np.float64).reshape(-1), 0.0)
    frames = list(frames)
    times, slabs = [], []
    for a in range(0, R, chunk):
        b = min(R, a + chunk)
        v = integration.UniformTSDFVolume(length, R, sdf_trunc, ct, origin, x_range=(a, b), device=device)
        v.integrate_device(frames[0][0], frames[0][1], intrinsic, frames[0][2])       # warm-up (scratch allocation)
        c.kernel_timing(True)
        for d, col, ext in frames:
            v.integrate_device(d, col, intrinsic, ext)
        c.kernel_timing(False)
        ms, n = c.kernel_time()
        times.append(ms / max(n, 1))
        slabs.append((a, b))
        del v
    t.cuda.empty_cache()
    floor = min(times)
    cost = np.zeros(R)
    for (a, b), tm in zip(slabs, times):
        ws = w[a:b].sum()
        share = (w[a:b] / ws) if ws > 0 else np.full(b - a, 1.0 / (b - a))
        cost[a:b] = max(tm - floor, 0.0) * share
    return cost


def probe_plane_work(length, resolution, sdf_trunc, origin, frames, intrinsic, probe_resolution=128, device=None):
    """Per-plane work estimate for `balanced_slab_ranges`: integrates `frames` -- an iterable of (depth_f32 CUDA
    tensor [H,W] in metres, extrinsic 4x4) -- into a coarse NoColor copy of the cube (probe_resolution^3, same
    length/origin) and counts the updates per z plane (weight = number of frames that updated a voxel).  Every
    rank runs the same probe on the same frames and gets the same answer: no collective."""
    from .pipelines import integration
    Rp = int(min(probe_resolution, resolution))
    vol = integration.UniformTSDFVolume(length, Rp, max(sdf_trunc, length / Rp), integration.TSDFVolumeColorType.NoColor,
                                        origin, device=device)
    for depth_f32, ext in frames:
        vol.integrate_device(depth_f32, None, intrinsic, ext)
    _, w, _ = vol.export_arrays()                    # [x, y, z]
    per_plane = np.asarray(w, np.float64).sum(axis=(1, 2))      # [x, y, z] -> work per x plane (the slab axis)
    rep = int(np.ceil(resolution / Rp))
    return np.repeat(per_plane, rep)[: int(resolution)] / rep


# --------------------------------------------------------------------------------------------------
# halo exchange
# --------------------------------------------------------------------------------------------------
def halo_plan(rank, world, n_owned, pad_lo=_lib.PAD_LO, pad_hi=_lib.PAD_HI):
    """Messages of one rank as (kind, peer, first_plane, n_planes) in units of planes of the padded slab
    array (plane p holds z = x_begin - pad_lo + p).  `send_down` = first pad_hi owned planes -> rank-1's top
    pad; `send_up` = last pad_lo owned planes -> rank+1's bottom pad."""
    ops = []
    if n_owned < max(pad_lo, pad_hi) and world > 1:
        raise RuntimeError("slab of %d planes is thinner than the halo (%d): use fewer ranks" %
                           (n_owned, max(pad_lo, pad_hi)))
    if rank > 0:
        ops.append(("send", rank - 1, pad_lo, pad_hi))                       # my first owned planes go down
        ops.append(("recv", rank - 1, 0, pad_lo))                            # bottom pad comes from below
    if rank < world - 1:
        ops.append(("send", rank + 1, pad_lo + n_owned - pad_lo, pad_lo))    # my last owned planes go up
        ops.append(("recv", rank + 1, pad_lo + n_owned, pad_hi))             # top pad comes from above
    return ops


def exchange_halos(arrays, n_owned, rank=None, world=None, group=None, pad_lo=_lib.PAD_LO, pad_hi=_lib.PAD_HI):
    """Fill the halo planes of every array in `arrays` (torch tensors shaped [pad_lo + n_owned + pad_hi, ...],
    plane-major, contiguous) from the neighbouring ranks.  One batch of point-to-point messages
    (ncclGroupStart/End under NCCL)."""
    dist = _dist()
    if rank is None or world is None:
        rank, world = world_info(group)
    if world == 1:
        return
    p2p = []
    for a in arrays:
        if a is None:
            continue
        for kind, peer, first, n in halo_plan(rank, world, n_owned, pad_lo, pad_hi):
            view = a[first:first + n]
            gpeer = peer if group is None else dist.get_global_rank(group, peer)
            p2p.append(dist.P2POp(dist.isend if kind == "send" else dist.irecv, view, gpeer, group))
    for w in dist.batch_isend_irecv(p2p):
        w.wait()


# --------------------------------------------------------------------------------------------------
# variable-length gathers
# --------------------------------------------------------------------------------------------------
def all_gather_counts(counts, device, group=None):
    """counts: sequence of python ints of this rank -> int64 array [world, len(counts)] (same on all ranks)."""
    import torch
    dist = _dist()
    rank, world = world_info(group)
    mine = torch.tensor([int(c) for c in counts], dtype=torch.int64, device=device)
    if world == 1:
        return mine.cpu().numpy()[None, :]
    out = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(out, mine, group=group)
    return torch.stack(out).cpu().numpy()


def exclusive_bases(per_rank_counts):
    """Exclusive prefix sum over ranks: base[r] = sum(counts[:r])."""
    c = np.asarray(per_rank_counts, np.int64)
    return np.concatenate([[0], np.cumsum(c)[:-1]]).astype(np.int64)


def gather_rows(tensor, counts, dst=0, group=None):
    """Concatenate row blocks of different lengths on rank `dst` (rank order).  `counts[r]` = rows of rank r
    (from all_gather_counts).  Returns the concatenated tensor on dst, None elsewhere."""
    import torch
    dist = _dist()
    rank, world = world_info(group)
    if world == 1:
        return tensor
    g = (lambda r: r) if group is None else (lambda r: dist.get_global_rank(group, r))
    if rank == dst:
        total = int(sum(int(c) for c in counts))
        out = torch.empty((total,) + tuple(tensor.shape[1:]), dtype=tensor.dtype, device=tensor.device)
        ops, off = [], 0
        for r in range(world):
            n = int(counts[r])
            if r == dst:
                out[off:off + n].copy_(tensor)
            elif n > 0:
                ops.append(dist.P2POp(dist.irecv, out[off:off + n], g(r), group))
            off += n
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()
        return out
    if int(counts[rank]) > 0:
        for w in dist.batch_isend_irecv([dist.P2POp(dist.isend, tensor.contiguous(), g(dst), group)]):
            w.wait()
    return None


# --------------------------------------------------------------------------------------------------
# the sharded volume
# --------------------------------------------------------------------------------------------------
class _RawCuda:
    """Exposes a raw device pointer as an int32 array to torch (`torch.as_tensor`) without copying."""

    def __init__(self, ptr, n_int32):
        self.__cuda_array_interface__ = {"shape": (int(n_int32),), "typestr": "<i4", "data": (int(ptr), False),
                                         "version": 2, "strides": None}


class SlabVolume:
    """`UniformTSDFVolume` sharded into slabs of x planes over the ranks of a process group (BASELINE configs 3 and 5).

        vol = SlabVolume(length, resolution, sdf_trunc, color_type, origin)   # every rank, same arguments
        vol.integrate(rgbd, intrinsic, extrinsic)                             # every rank, every frame
        mesh = vol.extract_triangle_mesh()                                    # TriangleMesh on rank 0, None elsewhere

    The mesh is the single-volume mesh: same vertex and triangle sets (vertex ids are global)."""

    def __init__(self, length, resolution, sdf_trunc, color_type=None, origin=(0.0, 0.0, 0.0), group=None, device=None,
                 x_ranges=None, local=None):
        """x_ranges: optional [(x_begin, x_end)] per rank (contiguous, covering [0, resolution)), e.g. from
        `balanced_slab_ranges`; default = equal plane counts.  local: this rank's slab if it already exists (a
        `UniformTSDFVolume(x_range=x_ranges[rank])`), else it is created here."""
        from .pipelines import integration
        self.group = group
        self.rank, self.world = world_info(group)
        self.resolution = int(resolution)
        self.timings = {}
        ct = integration.TSDFVolumeColorType.RGB8 if color_type is None else color_type
        if x_ranges is None:
            x_ranges = slab_ranges(resolution, self.world)
        x_ranges = [(int(a), int(b)) for a, b in x_ranges]
        if (len(x_ranges) != self.world or x_ranges[0][0] != 0 or x_ranges[-1][1] != self.resolution or
                any(x_ranges[i][1] != x_ranges[i + 1][0] for i in range(self.world - 1))):
            raise RuntimeError("SlabVolume: x_ranges must be %d contiguous slabs covering [0, %d)" % (self.world, self.resolution))
        self.x_begin, self.x_end = x_ranges[self.rank]
        if self.world > 1 and (self.x_end - self.x_begin) < max(_lib.PAD_LO, _lib.PAD_HI):
            raise RuntimeError("SlabVolume: %d planes over %d ranks leaves slabs thinner than the halo" %
                               (self.resolution, self.world))
        if local is not None:
            if (local.x_begin, local.x_end, local.resolution) != (self.x_begin, self.x_end, self.resolution):
                raise RuntimeError("SlabVolume: `local` is slab [%d,%d) of %d, this rank owns [%d,%d) of %d" % (
                    local.x_begin, local.x_end, local.resolution, self.x_begin, self.x_end, self.resolution))
            self.local = local
        else:
            self.local = integration.UniformTSDFVolume(length, resolution, sdf_trunc, ct, origin,
                                                       x_range=(self.x_begin, self.x_end), device=device)
        self._halo_valid = False

    def _timed(self, key, fn):
        """Runs fn(); when `self.timings` collection is on (`timings["on"]`), brackets it with CUDA events on the
        current stream + a synchronize and accumulates milliseconds under `key`."""
        if not self.timings.get("on"):
            return fn()
        import torch
        dev = self.local._ctx.tdev
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(dev)
        e0.record()
        out = fn()
        e1.record()
        torch.cuda.synchronize(dev)
        self.timings[key] = self.timings.get(key, 0.0) + e0.elapsed_time(e1)
        return out

    # ---- integration: no collective -------------------------------------------------------------
    def integrate(self, image, intrinsic, extrinsic):
        self.local.integrate(image, intrinsic, extrinsic)
        self._halo_valid = False

    def integrate_device(self, depth_f32, color, intrinsic, extrinsic):
        self.local.integrate_device(depth_f32, color, intrinsic, extrinsic)
        self._halo_valid = False

    def integrate_host(self, depth, color, intrinsic, extrinsic, depth_scale=1000.0, depth_trunc=3.0):
        self.local.integrate_host(depth, color, intrinsic, extrinsic, depth_scale, depth_trunc)
        self._halo_valid = False

    def reset(self):
        self.local.reset()
        self._halo_valid = False

    # ---- exchange -------------------------------------------------------------------------------
    def exchange_halos(self):
        """One-plane-below / two-planes-above halo of {tsdf, weight} and colour from the neighbour slabs."""
        if not self._halo_valid:
            self.local._ctx.join()   # the planes are sent straight from the volume's tensors
            exchange_halos([self.local._tsdfw, self.local._color, self.local._rows], self.x_end - self.x_begin, self.rank, self.world,
                           self.group)
            self._halo_valid = True

    # ---- extraction -----------------------------------------------------------------------------
    def extract_triangle_mesh(self, dst=0):
        """Halo exchange -> per-slab marching cubes with global vertex ids -> gather on rank `dst`."""
        import torch
        from . import geometry
        v = self.local
        dev = v._ctx.tdev
        self._timed("halo_ms", self.exchange_halos)
        nv, nt = self._timed("count_ms", v._mesh_count)
        counts = self._timed("seam_ms", lambda: all_gather_counts([nv, nt], dev, self.group))   # [world, 2]
        vbase = exclusive_bases(counts[:, 0])
        # fills the slab's vertex-id table
        V, Cc = self._timed("emit_ms", lambda: v._mesh_vertices(nv, vertex_base=int(vbase[self.rank])))
        if self.world > 1:
            self._timed("seam_ms", self._exchange_seam)
        T = self._timed("emit_ms", lambda: v._mesh_triangles(nt))

        def gather():
            Vg = gather_rows(V, counts[:, 0], dst, self.group)
            Cg = gather_rows(Cc, counts[:, 0], dst, self.group) if Cc is not None else None
            Tg = gather_rows(T, counts[:, 1], dst, self.group)
            return Vg, Cg, Tg
        Vg, Cg, Tg = self._timed("gather_ms", gather)
        self.last_counts = counts
        if self.rank != dst:
            return None
        m = geometry.TriangleMesh()
        m.vertices = Vg.cpu().numpy()
        m.triangles = Tg.cpu().numpy()
        if Cg is not None:
            m.vertex_colors = Cg.cpu().numpy()
        return m

    def _exchange_seam(self):
        """Triangles of the slab's top cube layer reference vertices owned by the first plane of the slab above:
        that plane of the vertex-id table (R*R*4 bytes) travels one rank down."""
        import ctypes as C
        import torch
        dist = _dist()
        v = self.local
        send, recv, nbytes = C.c_void_p(), C.c_void_p(), C.c_int64(0)
        _lib.check(_lib.load().b2r_mesh_seam_buffers(v._ctx.handle, C.byref(v._desc), C.byref(send), C.byref(recv),
                                                     C.byref(nbytes)))
        n = nbytes.value // 4
        g = (lambda r: r) if self.group is None else (lambda r: dist.get_global_rank(self.group, r))
        ops = []
        torch.cuda.current_stream(v._ctx.tdev).synchronize()   # the id table was written on this stream
        if self.rank > 0:
            ts = torch.as_tensor(_RawCuda(send.value, n), device=v._ctx.tdev)
            ops.append(dist.P2POp(dist.isend, ts, g(self.rank - 1), self.group))
        if self.rank < self.world - 1:
            tr = torch.as_tensor(_RawCuda(recv.value, n), device=v._ctx.tdev)
            ops.append(dist.P2POp(dist.irecv, tr, g(self.rank + 1), self.group))
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()

    def extract_point_cloud(self, dst=0):
        from . import geometry
        v = self.local
        self.exchange_halos()
        P, N, Cc = v.extract_point_cloud_device()
        counts = all_gather_counts([int(P.shape[0])], v._ctx.tdev, self.group)[:, 0]
        Pg = gather_rows(P, counts, dst, self.group)
        Ng = gather_rows(N, counts, dst, self.group)
        Cg = gather_rows(Cc, counts, dst, self.group) if Cc is not None else None
        if self.rank != dst:
            return None
        pcd = geometry.PointCloud()
        pcd._p.set_dev(Pg)
        pcd._n.set_dev(Ng)
        if Cg is not None:
            pcd._c.set_dev(Cg)
        return pcd


# --------------------------------------------------------------------------------------------------
# several slabs of one cube in ONE process (one GPU, or the GPUs of one process)
# --------------------------------------------------------------------------------------------------
def _raw_i32(ptr, n, tdev):
    import torch
    return torch.as_tensor(_RawCuda(ptr, n), device=tdev)


class LocalSlabs:
    """The slab decomposition of `SlabVolume` without a process group: every slab is a `UniformTSDFVolume` with its
    own library context (own scratch, side stream and extraction state) on one device of this process, and the
    exchange steps -- halo planes, the seam plane of the vertex-id table -- are device-to-device copies instead of
    NCCL messages.  Same kernels, same C-ABI calls (`vertex_base`, `b2r_mesh_seam_buffers`), same result: the
    single-volume mesh / point cloud.  Uses: cubes cut into slabs that are extracted one after the other with a
    bounded scratch footprint, and the seam logic under test on a single GPU (tests/test_slab_seam_gpu.py)."""

    def __init__(self, length, resolution, sdf_trunc, color_type=None, origin=(0.0, 0.0, 0.0), x_ranges=None,
                 n_slabs=2, device=None):
        from . import runtime
        from .pipelines import integration
        self.resolution = int(resolution)
        ct = integration.TSDFVolumeColorType.RGB8 if color_type is None else color_type
        if x_ranges is None:
            x_ranges = slab_ranges(resolution, n_slabs)
        x_ranges = [(int(a), int(b)) for a, b in x_ranges]
        n = len(x_ranges)
        if (x_ranges[0][0] != 0 or x_ranges[-1][1] != self.resolution or
                any(x_ranges[i][1] != x_ranges[i + 1][0] for i in range(n - 1))):
            raise RuntimeError("LocalSlabs: x_ranges must be contiguous slabs covering [0, %d)" % self.resolution)
        if n > 1 and min(b - a for a, b in x_ranges) < max(_lib.PAD_LO, _lib.PAD_HI):
            raise RuntimeError("LocalSlabs: a slab is thinner than the halo (%d planes)" % max(_lib.PAD_LO, _lib.PAD_HI))
        self.x_ranges = x_ranges
        self.slabs = [integration.UniformTSDFVolume(length, resolution, sdf_trunc, ct, origin, x_range=zr, device=device,
                                                    context=runtime.new_context(device)) for zr in x_ranges]
        self._halo_valid = False

    def __len__(self):
        return len(self.slabs)

    def integrate(self, image, intrinsic, extrinsic):
        for v in self.slabs:
            v.integrate(image, intrinsic, extrinsic)
        self._halo_valid = False

    def integrate_host(self, depth, color, intrinsic, extrinsic, depth_scale=1000.0, depth_trunc=3.0):
        for v in self.slabs:
            v.integrate_host(depth, color, intrinsic, extrinsic, depth_scale, depth_trunc)
        self._halo_valid = False

    def integrate_device(self, depth_f32, color, intrinsic, extrinsic):
        for v in self.slabs:
            v.integrate_device(depth_f32, color, intrinsic, extrinsic)
        self._halo_valid = False

    def reset(self):
        for v in self.slabs:
            v.reset()
        self._halo_valid = False

    def import_arrays(self, tsdf, weight, color=None):
        for v, (a, b) in zip(self.slabs, self.x_ranges):
            v.import_arrays(tsdf[a:b], weight[a:b], None if color is None else color[a:b])
        self._halo_valid = False

    def exchange_halos(self):
        """Same planes as `halo_plan` moves between ranks: first PAD_HI owned planes of slab k+1 -> top pad of slab k,
        last PAD_LO owned planes of slab k -> bottom pad of slab k+1."""
        if self._halo_valid:
            return
        lo, hi = _lib.PAD_LO, _lib.PAD_HI
        for v in self.slabs:
            v._ctx.join()
        for k in range(len(self.slabs) - 1):
            a, b = self.slabs[k], self.slabs[k + 1]
            na = a.x_end - a.x_begin
            for ta, tb in ((a._tsdfw, b._tsdfw), (a._color, b._color), (a._rows, b._rows)):
                if ta is None:
                    continue
                ta[lo + na:lo + na + hi].copy_(tb[lo:lo + hi])
                tb[0:lo].copy_(ta[lo + na - lo:lo + na])
        self._halo_valid = True

    def extract_triangle_mesh_device(self):
        """(vertices, triangles, colors|None) CUDA tensors of the whole cube; vertex ids are global."""
        import ctypes as C
        import torch
        self.exchange_halos()
        counts = [v._mesh_count() for v in self.slabs]
        # NOTE: every slab has its own context, so the count state of slab k survives the count of slab k+1
        bases = exclusive_bases([c[0] for c in counts])
        Vs, Cs = [], []
        for v, (nv, _), base in zip(self.slabs, counts, bases):
            V, Cc = v._mesh_vertices(nv, vertex_base=int(base))
            Vs.append(V)
            Cs.append(Cc)
        lib = _lib.load()
        bufs = []
        for v in self.slabs:
            send, recv, nbytes = C.c_void_p(), C.c_void_p(), C.c_int64(0)
            _lib.check(lib.b2r_mesh_seam_buffers(v._ctx.handle, C.byref(v._desc), C.byref(send), C.byref(recv),
                                                 C.byref(nbytes)))
            bufs.append((send.value, recv.value, nbytes.value // 4))
        for k in range(len(self.slabs) - 1):      # first owned plane of slab k+1's id table -> pad plane above slab k
            tdev = self.slabs[k]._ctx.tdev
            _raw_i32(bufs[k][1], bufs[k][2], tdev).copy_(_raw_i32(bufs[k + 1][0], bufs[k + 1][2], tdev))
        Ts = [v._mesh_triangles(nt) for v, (_, nt) in zip(self.slabs, counts)]
        V = torch.cat(Vs)
        T = torch.cat(Ts)
        Cc = torch.cat(Cs) if Cs[0] is not None else None
        return V, T, Cc

    def extract_triangle_mesh(self):
        from . import geometry
        V, T, Cc = self.extract_triangle_mesh_device()
        m = geometry.TriangleMesh()
        m.vertices = V.cpu().numpy()
        m.triangles = T.cpu().numpy()
        if Cc is not None:
            m.vertex_colors = Cc.cpu().numpy()
        return m

    def extract_point_cloud(self):
        import torch
        from . import geometry
        self.exchange_halos()
        parts = [v.extract_point_cloud_device() for v in self.slabs]
        pcd = geometry.PointCloud()
        pcd._p.set_dev(torch.cat([p[0] for p in parts]))
        pcd._n.set_dev(torch.cat([p[1] for p in parts]))
        if parts[0][2] is not None:
            pcd._c.set_dev(torch.cat([p[2] for p in parts]))
        return pcd


# --------------------------------------------------------------------------------------------------
# order-independent mesh digest (to compare the gathered mesh of N slabs with the single-volume mesh)
# --------------------------------------------------------------------------------------------------
def _mix64(x):
    x = x.astype(np.uint64, copy=True)
    x ^= x >> np.uint64(30)
    x *= np.uint64(0xBF58476D1CE4E5B9)
    x ^= x >> np.uint64(27)
    x *= np.uint64(0x94D049BB133111EB)
    x ^= x >> np.uint64(31)
    return x


def mesh_checksum(vertices, triangles, quantum):
    """Digest of a triangle mesh that does not depend on vertex order, triangle order, vertex numbering or on which
    corner a triangle starts with (its orientation does count): vertices are quantised to `quantum` (metres; e.g.
    1e-4 voxel, the parity tolerance) and hashed; a triangle is hashed from its corners' vertex hashes rotated so the
    smallest comes first.  Returns {"vertices": n, "triangles": n, "vertex_sum": hex, "triangle_sum": hex}: equal
    digests <=> equal quantised vertex sets and equal oriented triangle sets (up to 64-bit hash collisions)."""
    V = np.asarray(vertices, np.float64).reshape(-1, 3)
    T = np.asarray(triangles, np.int64).reshape(-1, 3)
    with np.errstate(over="ignore"):
        q = np.round(V / float(quantum)).astype(np.int64).view(np.uint64)
        vh = _mix64(q[:, 0] + _mix64(q[:, 1] + _mix64(q[:, 2])))
        vsum = int(np.add.reduce(vh, dtype=np.uint64)) if len(vh) else 0
        tsum = 0
        if len(T):
            h = vh[T]
            # canonical rotation = the lexicographically smallest of the three (ties between equal corner hashes --
            # degenerate triangles, which Open3D keeps -- are then resolved the same way whatever the start corner)
            r1, r2 = np.roll(h, -1, axis=1), np.roll(h, -2, axis=1)

            def less(a, b):
                return (a[:, 0] < b[:, 0]) | ((a[:, 0] == b[:, 0]) & ((a[:, 1] < b[:, 1]) | ((a[:, 1] == b[:, 1]) & (a[:, 2] < b[:, 2]))))
            best = np.where(less(r1, h)[:, None], r1, h)
            h = np.where(less(r2, best)[:, None], r2, best)
            th = _mix64(h[:, 0] + _mix64(h[:, 1] * np.uint64(3) + _mix64(h[:, 2] * np.uint64(5))))
            tsum = int(np.add.reduce(th, dtype=np.uint64))
    return {"vertices": int(len(V)), "triangles": int(len(T)), "vertex_sum": "%016x" % vsum,
            "triangle_sum": "%016x" % tsum}


# --------------------------------------------------------------------------------------------------
# batched ICP: pairs round-robin over the ranks, no collective on the data path
# --------------------------------------------------------------------------------------------------
def round_robin(n_items, rank, world):
    """Indices of the items rank `rank` owns: p with p % world == rank."""
    return list(range(rank, n_items, world))


def icp_round_robin(sources, targets, max_correspondence_distance, inits=None, estimation_method=None, criteria=None,
                    gather=True, group=None):
    """BASELINE config 4: pair p is registered on rank p % world (`registration_icp_batch` on that rank's GPU).
    Returns the full list of RegistrationResult in pair order on every rank when `gather` (one
    all_gather_object of the 4x4 + two scalars per pair -- results only, no point data), else {pair: result}
    for this rank's pairs."""
    from .pipelines import registration
    dist = _dist()
    rank, world = world_info(group)
    mine = round_robin(len(sources), rank, world)
    res = registration.registration_icp_batch([sources[p] for p in mine], [targets[p] for p in mine],
                                              max_correspondence_distance,
                                              None if inits is None else [inits[p] for p in mine],
                                              estimation_method, criteria, with_correspondences=not gather)
    local = dict(zip(mine, res))
    if not gather:
        return local
    if world == 1:
        return [local[p] for p in range(len(sources))]
    slim = {p: (r.transformation, r.fitness, r.inlier_rmse, getattr(r, "iterations", 0), getattr(r, "converged", False))
            for p, r in local.items()}
    parts = [None] * world
    dist.all_gather_object(parts, slim, group=group)
    out = [None] * len(sources)
    for part in parts:
        for p, (T, f, e, it, cv) in part.items():
            r = registration.RegistrationResult()
            r.transformation, r.fitness, r.inlier_rmse, r.iterations, r.converged = np.asarray(T), f, e, it, cv
            out[p] = r
    return out
