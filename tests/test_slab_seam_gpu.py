"""The slab seam logic of SURVEY.md 8(e) on ONE GPU: `parallel.LocalSlabs` cuts a cube into unequal slabs (x planes) that live in
one process, fills their halo planes with device copies (the planes `parallel.halo_plan` moves over NCCL between
ranks), runs b2r_extract_mesh_count / _vertices (vertex_base) / b2r_mesh_seam_buffers / _triangles per slab and
concatenates.  The union must be the oracle's single-volume mesh / point cloud -- the check tests/mgpu_worker.py does
on 2..8 GPUs, here without a process group so that it runs on the 1-GPU test box."""
import numpy as np
import pytest

import helpers as hp
from b200recon import parallel, synthetic as syn

pytestmark = pytest.mark.gpu


def _check_mesh(ref_mesh, mesh, vl, color_tol):
    Vr, Tr, Cr = ref_mesh
    V, T, Cc = np.asarray(mesh.vertices), np.asarray(mesh.triangles), np.asarray(mesh.vertex_colors)
    assert len(V) == len(Vr) and len(T) == len(Tr), (len(V), len(Vr), len(T), len(Tr))
    Vs_r, Ts_r, ord_r = hp.canon_mesh(Vr, Tr)
    Vs, Ts, ord_g = hp.canon_mesh(V, T)
    assert np.abs(Vs - Vs_r).max() <= 1e-4 * vl
    assert np.array_equal(Ts, Ts_r)
    if Cr is not None and len(Cr):
        assert np.abs(Cc[ord_g] - Cr[ord_r]).max() <= color_tol
    # the order-independent digest bench.py prints for the N-GPU mesh agrees with the canonical comparison
    q = 1e-4 * vl
    assert parallel.mesh_checksum(V, T, q) == parallel.mesh_checksum(Vr, Tr, q)


def _check_points(ref_pc, pcd, vl):
    P, N, _ = ref_pc
    Pg, Ng = np.asarray(pcd.points), np.asarray(pcd.normals)
    assert len(Pg) == len(P), (len(Pg), len(P))
    a, oa = hp.sort_rows(P)
    b, ob = hp.sort_rows(Pg)
    assert np.abs(a - b).max() <= 1e-4 * vl
    assert np.abs(N[oa] - Ng[ob]).max() < 1e-9


@pytest.mark.parametrize("ranges", [[(0, 40), (40, 44), (44, 96)], [(0, 46), (46, 48), (48, 51), (51, 96)]])
def test_three_unequal_slabs_frames_scene(o3d, orc, ranges):
    """Fused frames: every slab bit-exact against the oracle's planes, union mesh / points == single-volume oracle.
    Slabs are ranges of x planes; the thin middle slabs (4 and 2-3 planes) sit where the object is (x index ~27-69),
    so the seams cut the table and the object."""
    R, L, trunc = 96, 0.384, 0.016
    K = hp.small_intrinsic(320, 240)
    frames = hp.make_frames(5, K)
    origin = syn.volume_origin(L)
    RGB8 = o3d.pipelines.integration.TSDFVolumeColorType.RGB8
    intr = o3d.camera.PinholeCameraIntrinsic(K["width"], K["height"], *hp.kt(K))
    vol = parallel.LocalSlabs(L, R, trunc, RGB8, origin, x_ranges=ranges)
    ref = orc.UniformTSDFVolume(L, R, trunc, 1, origin)
    for d16, rgb, ext in frames:
        vol.integrate_host(d16, rgb, intr, ext)
        ref.integrate(orc.depth_to_float(d16), rgb, hp.kt(K), ext)
    for v, (zb, ze) in zip(vol.slabs, ranges):
        ts, w, _ = v.export_arrays()
        assert np.array_equal(w, ref.weight()[zb:ze]), "slab weights differ"
        assert np.array_equal(ts.view(np.uint32), ref.tsdf()[zb:ze].view(np.uint32)), "slab tsdf differs"
    ref_mesh = ref.extract_triangle_mesh()
    assert len(ref_mesh[0]) > 1000
    # the surface must really cross the seams: vertices on both sides of every cut
    zv = (ref_mesh[0][:, 0] - origin[0]) / (L / R)        # slab axis: x
    for (_, cut) in ranges[:-1]:
        assert (zv < cut).any() and (zv > cut).any()
    _check_mesh(ref_mesh, vol.extract_triangle_mesh(), L / R, 2.0 / 255.0)
    _check_points(ref.extract_point_cloud(), vol.extract_point_cloud(), L / R)


def test_white_noise_volume_across_seams(o3d, orc):
    """Every cube configuration, with zero-weight holes, across three seams (incl. a slab of PAD_HI planes)."""
    R = 27
    rng = np.random.default_rng(11)
    ts = rng.uniform(-1, 1, (R, R, R)).astype(np.float32)
    w = (rng.random((R, R, R)) > 0.08).astype(np.float32) * rng.integers(1, 5, (R, R, R)).astype(np.float32)
    col = (np.round(rng.uniform(0, 255, (R, R, R, 3)) * 4) / 4).astype(np.float32)
    RGB8 = o3d.pipelines.integration.TSDFVolumeColorType.RGB8
    vol = parallel.LocalSlabs(1.2, R, 0.1, RGB8, (0.1, -0.2, 0.3), x_ranges=[(0, 9), (9, 11), (11, 20), (20, 27)])
    vol.import_arrays(ts, w, col)
    ref = orc.UniformTSDFVolume(1.2, R, 0.1, 1, (0.1, -0.2, 0.3))
    ref.tsdf()[:] = ts
    ref.weight()[:] = w
    ref.color()[:] = col
    ref_mesh = ref.extract_triangle_mesh()
    assert len(ref_mesh[1]) > 5000
    _check_mesh(ref_mesh, vol.extract_triangle_mesh(), 1.2 / R, 1e-9)
    P, N, _ = ref.extract_point_cloud()
    pcd = vol.extract_point_cloud()
    Pg, Ng = np.asarray(pcd.points), np.asarray(pcd.normals)
    assert len(P) == len(Pg)
    a, oa = hp.sort_rows(P)
    b, ob = hp.sort_rows(Pg)
    assert np.abs(a - b).max() < 1e-12 and np.abs(N[oa] - Ng[ob]).max() < 1e-9


def test_single_slab_equals_plain_volume(o3d):
    """LocalSlabs with one slab is the plain volume (no exchange)."""
    R = 20
    rng = np.random.default_rng(3)
    ts = rng.uniform(-1, 1, (R, R, R)).astype(np.float32)
    w = np.ones((R, R, R), np.float32)
    NoColor = o3d.pipelines.integration.TSDFVolumeColorType.NoColor
    a = parallel.LocalSlabs(1.0, R, 0.1, NoColor, n_slabs=1)
    a.import_arrays(ts, w)
    b = o3d.pipelines.integration.UniformTSDFVolume(1.0, R, 0.1, NoColor)
    b.import_arrays(ts, w)
    ma, mb = a.extract_triangle_mesh(), b.extract_triangle_mesh()
    assert np.array_equal(np.asarray(ma.vertices), np.asarray(mb.vertices))
    assert np.array_equal(np.asarray(ma.triangles), np.asarray(mb.triangles))
