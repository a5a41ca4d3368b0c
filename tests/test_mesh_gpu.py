"""Marching cubes / point extraction parity with the oracle: identical counts, vertex positions within
1e-4 voxel (expected: identical doubles), identical triangle sets after order canonicalisation."""
import numpy as np
import pytest

import helpers as hp
from b200recon import synthetic as syn

pytestmark = pytest.mark.gpu


def _fused_pair(o3d, orc, R, L, trunc, n_frames, K):
    frames = hp.make_frames(n_frames, K)
    origin = syn.volume_origin(L)
    ref = orc.UniformTSDFVolume(L, R, trunc, 1, origin)
    vol = o3d.pipelines.integration.UniformTSDFVolume(L, R, trunc, o3d.pipelines.integration.TSDFVolumeColorType.RGB8,
                                                      origin)
    intr = o3d.camera.PinholeCameraIntrinsic(K["width"], K["height"], *hp.kt(K))
    for d16, rgb, ext in frames:
        ref.integrate(orc.depth_to_float(d16), rgb, hp.kt(K), ext)
        vol.integrate_host(d16, rgb, intr, ext)
    return ref, vol


def _check_mesh(ref_mesh, mesh, vl, color_tol=2.0 / 255.0):
    Vr, Tr, Cr = ref_mesh
    V, T, Cc = np.asarray(mesh.vertices), np.asarray(mesh.triangles), np.asarray(mesh.vertex_colors)
    assert len(V) == len(Vr) and len(T) == len(Tr), (len(V), len(Vr), len(T), len(Tr))
    Vs_r, Ts_r, ord_r = hp.canon_mesh(Vr, Tr)
    Vs, Ts, ord_g = hp.canon_mesh(V, T)
    assert np.abs(Vs - Vs_r).max() <= 1e-4 * vl
    assert np.array_equal(Ts, Ts_r)
    if Cr is not None and len(Cr):
        assert np.abs(Cc[ord_g] - Cr[ord_r]).max() <= color_tol


@pytest.mark.parametrize("R,L,trunc", [(64, 0.384, 0.024), (96, 0.384, 0.016)])
def test_mesh_matches_oracle(o3d, orc, R, L, trunc):
    K = hp.small_intrinsic(320, 240)
    ref, vol = _fused_pair(o3d, orc, R, L, trunc, 6, K)
    ref_mesh = ref.extract_triangle_mesh()
    assert len(ref_mesh[0]) > 1000
    _check_mesh(ref_mesh, vol.extract_triangle_mesh(), L / R)


def test_mesh_white_noise_volume(o3d, orc):
    """Every cube configuration incl. zero-weight holes: import the same random TSDF on both sides."""
    R = 24
    rng = np.random.default_rng(7)
    ts = rng.uniform(-1, 1, (R, R, R)).astype(np.float32)
    w = (rng.random((R, R, R)) > 0.08).astype(np.float32) * rng.integers(1, 5, (R, R, R)).astype(np.float32)
    col = np.round(rng.uniform(0, 255, (R, R, R, 3)) * 4) / 4
    ref = orc.UniformTSDFVolume(1.2, R, 0.1, 1, (0.1, -0.2, 0.3))
    ref.tsdf()[:] = ts
    ref.weight()[:] = w
    ref.color()[:] = col
    vol = o3d.pipelines.integration.UniformTSDFVolume(1.2, R, 0.1, o3d.pipelines.integration.TSDFVolumeColorType.RGB8,
                                                      (0.1, -0.2, 0.3))
    vol.import_arrays(ts, w, col.astype(np.float32))
    ts2, w2, c2 = vol.export_arrays()
    assert np.array_equal(ts2, ts) and np.array_equal(w2, w) and np.array_equal(c2, col.astype(np.float32))
    ref_mesh = ref.extract_triangle_mesh()
    assert len(ref_mesh[1]) > 5000
    _check_mesh(ref_mesh, vol.extract_triangle_mesh(), 1.2 / R, color_tol=1e-9)


def test_mesh_empty_volume(o3d):
    vol = o3d.pipelines.integration.UniformTSDFVolume(1.0, 16, 0.1, o3d.pipelines.integration.TSDFVolumeColorType.RGB8)
    m = vol.extract_triangle_mesh()
    assert len(m.vertices) == 0 and len(m.triangles) == 0
    p = vol.extract_point_cloud()
    assert len(p.points) == 0


def test_point_cloud_matches_oracle(o3d, orc):
    K = hp.small_intrinsic(320, 240)
    ref, vol = _fused_pair(o3d, orc, 64, 0.384, 0.024, 6, K)
    P, N, Cc = ref.extract_point_cloud()
    pcd = vol.extract_point_cloud()
    Pg, Ng, Cg = np.asarray(pcd.points), np.asarray(pcd.normals), np.asarray(pcd.colors)
    assert len(Pg) == len(P) and len(P) > 1000
    a, oa = hp.sort_rows(P)
    b, ob = hp.sort_rows(Pg)
    assert np.abs(a - b).max() <= 1e-4 * (0.384 / 64)
    assert np.abs(N[oa] - Ng[ob]).max() < 1e-9
    assert np.abs(Cc[oa] - Cg[ob]).max() <= 2.0 / 255.0


def test_point_cloud_white_noise(o3d, orc):
    R = 20
    rng = np.random.default_rng(3)
    ts = rng.uniform(-1, 1, (R, R, R)).astype(np.float32)
    w = (rng.random((R, R, R)) > 0.1).astype(np.float32)
    ref = orc.UniformTSDFVolume(1.0, R, 0.1, 0)
    ref.tsdf()[:] = ts
    ref.weight()[:] = w
    vol = o3d.pipelines.integration.UniformTSDFVolume(1.0, R, 0.1, o3d.pipelines.integration.TSDFVolumeColorType.NoColor)
    vol.import_arrays(ts, w)
    P, N, _ = ref.extract_point_cloud()
    pcd = vol.extract_point_cloud()
    Pg, Ng = np.asarray(pcd.points), np.asarray(pcd.normals)
    assert len(P) == len(Pg) and len(P) > 500
    a, oa = hp.sort_rows(P)
    b, ob = hp.sort_rows(Pg)
    assert np.abs(a - b).max() < 1e-12
    assert np.abs(N[oa] - Ng[ob]).max() < 1e-9


def test_row_marks_cover_negative_rows_and_stale_codes_are_not_read(o3d, orc):
    """b2r_volume::d_rows: every (x, z) row holding a negative tsdf value is marked (the only rows an active cube or a surface
    point can need), the free space in front of the surfaces is not, and cells that classification skips keep whatever codes an
    earlier extraction left in the context's scratch -- here a white-noise volume of the same size, active cubes everywhere --
    without anyone reading them."""
    from b200recon import _lib
    R, L, trunc = 96, 0.384, 0.016
    rng = np.random.default_rng(5)
    noise = o3d.pipelines.integration.UniformTSDFVolume(L, R, trunc, o3d.pipelines.integration.TSDFVolumeColorType.RGB8,
                                                        syn.volume_origin(L))
    noise.import_arrays(rng.uniform(-1, 1, (R, R, R)).astype(np.float32), np.ones((R, R, R), np.float32),
                        np.zeros((R, R, R, 3), np.float32))
    assert len(noise.extract_triangle_mesh().triangles) > 100000
    assert len(noise.extract_point_cloud().points) > 100000
    del noise
    K = hp.small_intrinsic(320, 240)
    ref, vol = _fused_pair(o3d, orc, R, L, trunc, 6, K)
    ts, w, _ = vol.export_arrays()
    marks = vol._rows.cpu().numpy()[_lib.PAD_LO:_lib.PAD_LO + R]          # [x][z]
    neg = (ts < 0).any(axis=1)                                            # ts is [x][y][z]
    assert neg.sum() > 100 and marks[neg].all()
    touched = (w != 0).any(axis=1)
    assert marks.astype(bool).sum() < 0.6 * touched.sum(), "marks are not output-sensitive"
    ref_mesh = ref.extract_triangle_mesh()
    _check_mesh(ref_mesh, vol.extract_triangle_mesh(), L / R)
    P, N, _ = ref.extract_point_cloud()
    pcd = vol.extract_point_cloud()
    Pg, Ng = np.asarray(pcd.points), np.asarray(pcd.normals)
    assert len(Pg) == len(P)
    a, oa = hp.sort_rows(P)
    b, ob = hp.sort_rows(Pg)
    assert np.abs(a - b).max() <= 1e-4 * (L / R) and np.abs(N[oa] - Ng[ob]).max() < 1e-9
