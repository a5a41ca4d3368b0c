"""BASELINE config[0] in miniature: the reference's stitch_sequence call sequence (tests/stitch_mirror.py) on
b200recon (GPU) against the same sequence on the CPU oracle (tests/oracle_o3d.py), identical synthetic frames.
Tolerance: every ICP pose within the north_star ICP tolerance (1e-5 rotation, 1e-6 m translation), fitness equal,
and the fused cloud equal point for point within 1e-6 m (same cells, same kept indices)."""
import numpy as np
import pytest

import helpers as hp
import oracle_o3d
from stitch_mirror import Stitcher

pytestmark = pytest.mark.gpu


def frames(n, W=320, H=240, step_deg=2.0):
    from b200recon import synthetic as syn
    from oracle import oracle as orc
    K = hp.small_intrinsic(W, H)
    poses = syn.orbit_poses(n, radius=0.45, elevation_deg=35.0, az_span=np.deg2rad(step_deg) * n)
    cols, deps = [], []
    for p in poses:
        d16, rgb = orc.raycast_scene(p, hp.kt(K), W, H, checker=0.03)
        cols.append(rgb)
        deps.append(d16)
    return K, cols, deps


def to_gpu(o3d, pc):
    g = o3d.geometry.PointCloud()
    g.points = o3d.utility.Vector3dVector(pc.points)
    if pc.has_colors():
        g.colors = o3d.utility.Vector3dVector(pc.colors)
    if pc.has_normals():
        g.normals = o3d.utility.Vector3dVector(pc.normals)
    return g


def same_step(la, lb):
    for (ka, Ta, fa, ra, na), (kb, Tb, fb, rb, nb) in zip(la, lb):
        assert ka == kb and na == nb
        assert np.abs(Ta[:3, :3] - Tb[:3, :3]).max() <= 1e-5
        assert np.abs(Ta[:3, 3] - Tb[:3, 3]).max() <= 1e-6
        assert fa == fb
        assert abs(ra - rb) <= 1e-9


@pytest.mark.parametrize("voxel,min_pts", [(0.02, 1000), (0.008, 500)])
def test_stitch_sequence_every_step_matches_oracle(o3d, voxel, min_pts):
    """Every step of the sequence on IDENTICAL inputs: the oracle runs the chain, and each of its steps (register =
    2x preprocess + colored ICP + point-to-point ICP; voxel_down_sample; remove_statistical_outlier) is repeated on
    the GPU from the oracle's own input clouds.  (A free-running comparison is chaotic, see the next test.)"""
    K, cols, deps = frames(5)
    kw = dict(voxel_size=voxel, distance_threshold=0.05, min_points_for_outlier_removal=min_pts)
    G = Stitcher(o3d, o3d.camera.PinholeCameraIntrinsic(K["width"], K["height"], *hp.kt(K)), **kw)
    A = Stitcher(oracle_o3d, oracle_o3d.camera.PinholeCameraIntrinsic(K["width"], K["height"], *hp.kt(K)), **kw)
    comb = A.cloud(cols[0], deps[0])
    g0 = G.cloud(cols[0], deps[0])
    assert np.array_equal(np.asarray(g0.points), comb.points) and np.array_equal(np.asarray(g0.colors), comb.colors)
    cleaned = 0
    for i in range(1, len(cols)):
        cur = A.cloud(cols[i], deps[i])
        na, ng = len(A.log), len(G.log)
        T, _ = A.register(cur, comb)
        G.register(to_gpu(o3d, cur), to_gpu(o3d, comb))
        same_step(G.log[ng:], A.log[na:])
        cur.transform(T)
        comb += cur
        if i % 2 == 0:
            gc = to_gpu(o3d, comb)
            comb = comb.voxel_down_sample(voxel)
            gd = gc.voxel_down_sample(voxel)
            assert np.array_equal(np.asarray(gd.points), comb.points)
            assert np.array_equal(np.asarray(gd.colors), comb.colors)
            if len(comb.points) > min_pts:
                comb, idx = comb.remove_statistical_outlier(nb_neighbors=20, std_ratio=2.0)
                go, gidx = gd.remove_statistical_outlier(nb_neighbors=20, std_ratio=2.0)
                assert gidx == idx
                assert np.array_equal(np.asarray(go.points), comb.points)
                cleaned += 1
    assert cleaned == 2 and len(A.log) == 8
    assert all(f > 0.5 for (_, _, f, _, _) in G.log)


@pytest.mark.parametrize("voxel,min_pts", [(0.02, 1000)])
def test_stitch_sequence_free_running(o3d, voxel, min_pts):
    """The whole sequence run independently on both sides.  The chain is chaotic at the 1e-16 level: the down-sampled
    clouds are near-regular grids, so a last-bit difference in a pose (device cos/sin vs glibc) can change which point is
    the 30th neighbour of another, hence a colour gradient, hence an ICP trajectory -- Open3D itself does not reproduce
    its own poses from run to run (OpenMP sum order).  First registration: tight (identical inputs);
    afterwards only agreement of the fused result at the scale of the voxel size is asserted."""
    K, cols, deps = frames(5)
    kw = dict(voxel_size=voxel, distance_threshold=0.05, min_points_for_outlier_removal=min_pts)
    a = Stitcher(o3d, o3d.camera.PinholeCameraIntrinsic(K["width"], K["height"], *hp.kt(K)), **kw)
    b = Stitcher(oracle_o3d, oracle_o3d.camera.PinholeCameraIntrinsic(K["width"], K["height"], *hp.kt(K)), **kw)
    ca, cb = a.stitch(cols, deps), b.stitch(cols, deps)
    assert len(a.log) == len(b.log) == 8
    same_step(a.log[:2], b.log[:2])
    for (ka, Ta, fa, ra, na), (kb, Tb, fb, rb, nb) in zip(a.log, b.log):
        assert np.abs(Ta - Tb).max() < 0.05 and abs(fa - fb) < 0.01
    Pa, Pb = np.asarray(ca.points), np.asarray(cb.points)
    assert abs(len(Pa) - len(Pb)) <= 0.10 * len(Pb)
    from scipy.spatial import cKDTree
    d, _ = cKDTree(Pb).query(Pa, k=1)
    assert np.median(d) < voxel
