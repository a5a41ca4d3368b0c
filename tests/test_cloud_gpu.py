"""GPU parity of the point-cloud preprocessing around ICP (cloud.cu through the
Open3D-named PointCloud methods) against the oracle's restatement of Open3D 0.18.

Tolerances (floating point, stated here because north_star names none for these rows):
  * neighbour SETS / kept-index lists / cell assignment: exact (index work);
  * down-sampled means: bit-exact (per-cell sums run in ascending point order like Open3D's sequential loop --
    anything less makes the near-regular down-sampled grids pick different 30th neighbours downstream);
  * normals: 1e-9 per component (device acos/cos differ from glibc's by an ulp; the eigenvector is well conditioned
    on these clouds), sign included;
  * colour gradient: 1e-7 relative (3x3 normal equations of up to 30 rows);
  * colored ICP: the north_star ICP tolerance, rotation 1e-5 / translation 1e-6 m."""
import numpy as np
import pytest

from test_icp_gpu import se3, surface_cloud

pytestmark = pytest.mark.gpu


def cloud(o3d, P, N=None, Cc=None):
    p = o3d.geometry.PointCloud()
    p.points = o3d.utility.Vector3dVector(P)
    if N is not None:
        p.normals = o3d.utility.Vector3dVector(N)
    if Cc is not None:
        p.colors = o3d.utility.Vector3dVector(Cc)
    return p


def texture(P):
    """smooth colour field in [0,1] with a usable gradient everywhere"""
    return np.stack([0.5 + 0.4 * np.sin(40.0 * P[:, 0] + 1.0), 0.5 + 0.4 * np.cos(35.0 * P[:, 1]),
                     0.5 + 0.4 * np.sin(30.0 * (P[:, 0] + P[:, 1]))], 1)


def sorted_rows(*arrs):
    A = np.concatenate([np.asarray(a, np.float64) for a in arrs], 1)
    key = np.round(A[:, :3], 9)
    return A[np.lexsort((key[:, 2], key[:, 1], key[:, 0]))]


@pytest.mark.parametrize("mode,radius,k", [("hybrid", 0.004, 30), ("hybrid", 0.02, 30), ("knn", 0.0, 20), ("hybrid", 0.05, 100),
                                           ("knn", 0.0, 1)])
def test_knn_search_is_exact(o3d, orc, mode, radius, k):
    from b200recon import cloud_ops
    P, _ = surface_cloud(20000, 5, noise=0.0003)
    P = np.concatenate([P, np.array([[0.5, 0.5, 0.5], [-0.4, 0.3, 0.2]])])   # two far outliers: long ring walks
    Q = P[::7]
    sp = o3d.geometry.KDTreeSearchParamHybrid(radius, k) if mode == "hybrid" else o3d.geometry.KDTreeSearchParamKNN(k)
    idx, d2 = cloud_ops.knn_search(P, Q, sp)
    ridx, rd2 = orc.knn(P, Q, 2 if mode == "hybrid" else 0, radius, k)
    assert np.array_equal(idx, ridx)
    assert np.array_equal(d2, rd2)      # same float64 operations in the same order
    # independent check of the oracle itself
    from scipy.spatial import cKDTree
    d, j = cKDTree(P).query(Q[:200], k=k, distance_upper_bound=radius if mode == "hybrid" else np.inf)
    d, j = d.reshape(200, -1), j.reshape(200, -1)
    for a in range(200):
        ok = np.isfinite(d[a]) & ((d[a] ** 2 < radius * radius) if mode == "hybrid" else True)
        assert set(j[a][ok]) == set(ridx[a][ridx[a] >= 0])


@pytest.mark.parametrize("vs", [0.02, 0.005, 0.3])
def test_voxel_down_sample_matches_oracle(o3d, orc, vs):
    P, N = surface_cloud(60000, 7, noise=0.0002)
    Cc = texture(P)
    out = cloud(o3d, P, N, Cc).voxel_down_sample(vs)
    rp, rn, rc = orc.voxel_down_sample(P, vs, N, Cc)
    assert len(out.points) == len(rp)
    # cells are emitted by first point index on both sides: same order, no canonicalisation needed
    assert np.array_equal(np.asarray(out.points), rp)
    assert np.array_equal(np.asarray(out.normals), rn)
    assert np.array_equal(np.asarray(out.colors), rc)
    # points only (the reference's own call: stitcher.py:151 on a cloud without normals)
    out2 = cloud(o3d, P).voxel_down_sample(vs)
    assert not out2.has_normals() and not out2.has_colors()
    assert np.array_equal(np.asarray(out2.points), rp)


def test_voxel_down_sample_edge_cases(o3d):
    with pytest.raises(RuntimeError):
        cloud(o3d, np.zeros((4, 3))).voxel_down_sample(0.0)
    assert len(o3d.geometry.PointCloud().voxel_down_sample(0.01).points) == 0
    one = cloud(o3d, np.tile([[0.1, 0.2, 0.3]], (5, 1))).voxel_down_sample(0.01)
    assert len(one.points) == 1 and np.allclose(np.asarray(one.points)[0], [0.1, 0.2, 0.3], atol=1e-15)
    with pytest.raises(RuntimeError):   # 21-bit cell coordinates (Open3D: "voxel_size is too small.")
        cloud(o3d, np.array([[0.0, 0, 0], [1.0, 1, 1]])).voxel_down_sample(1e-8)


@pytest.mark.parametrize("mode,radius,k", [("hybrid", 0.04, 30), ("hybrid", 0.003, 30), ("knn", 0.0, 30)])
def test_estimate_normals_matches_oracle(o3d, orc, mode, radius, k):
    P, Ntrue = surface_cloud(30000, 9, noise=0.0002)
    P = np.concatenate([P, np.array([[0.5, 0.5, 0.5]])])       # isolated point: < 3 neighbours -> (0,0,1) in hybrid mode
    sp = o3d.geometry.KDTreeSearchParamHybrid(radius, k) if mode == "hybrid" else o3d.geometry.KDTreeSearchParamKNN(k)
    pc = cloud(o3d, P)
    pc.estimate_normals(sp)
    ref = orc.estimate_normals(P, 2 if mode == "hybrid" else 0, radius, k)
    got = np.asarray(pc.normals)
    assert np.abs(got - ref).max() <= 1e-9
    if mode == "hybrid":
        assert np.array_equal(got[-1], [0.0, 0.0, 1.0])
    if radius == 0.04 or mode == "knn":   # enough neighbours: close to the analytic surface normal (up to sign)
        c = np.abs(np.sum(got[:-1] * Ntrue, 1))
        assert np.median(c) > 0.99
    # pre-existing normals: the new ones are flipped to agree with them
    prev = -ref
    pc2 = cloud(o3d, P, prev)
    pc2.estimate_normals(sp)
    ref2 = orc.estimate_normals(P, 2 if mode == "hybrid" else 0, radius, k, prev=prev)
    assert np.abs(np.asarray(pc2.normals) - ref2).max() <= 1e-9


def test_remove_statistical_outlier_matches_oracle(o3d, orc):
    rng = np.random.default_rng(3)
    P, _ = surface_cloud(20000, 13, noise=0.0003)
    P = np.concatenate([P, rng.uniform(-0.2, 0.2, size=(300, 3))])   # floating outliers
    Cc = texture(P)
    out, idx = cloud(o3d, P, None, Cc).remove_statistical_outlier(nb_neighbors=20, std_ratio=2.0)
    ridx, _ = orc.statistical_outlier(P, 20, 2.0)
    assert np.array_equal(np.asarray(idx, np.int64), ridx)
    assert np.array_equal(np.asarray(out.points), P[ridx])
    assert np.array_equal(np.asarray(out.colors), Cc[ridx])
    assert len(ridx) < len(P) - 100
    with pytest.raises(RuntimeError):
        cloud(o3d, P).remove_statistical_outlier(0, 2.0)


def test_color_gradient_matches_oracle(o3d, orc):
    from b200recon import cloud_ops
    P, N = surface_cloud(15000, 21, noise=0.0001)
    Cc = texture(P)
    g = cloud_ops.color_gradient(cloud(o3d, P, N, Cc), 0.02, 30).cpu().numpy()
    ref = orc.color_gradient(P, N, Cc, 0.02, 30)
    scale = np.abs(ref).max()
    assert np.abs(g - ref).max() <= 1e-7 * scale


def test_colored_icp_matches_oracle(o3d, orc):
    reg = o3d.pipelines.registration
    T_true = se3(0.006, -0.008, 0.01, [0.0015, -0.001, 0.0008])
    Ps, _ = surface_cloud(8000, 31, noise=0.0001)
    Pt, Nt = surface_cloud(9000, 32, noise=0.0001)
    Cs, Ct = texture(Ps), texture(Pt)                 # colours attached to the surface, then the source is moved
    Ti = np.linalg.inv(T_true)
    Ps = Ps @ Ti[:3, :3].T + Ti[:3, 3]
    s, t = cloud(o3d, Ps, None, Cs), cloud(o3d, Pt, Nt, Ct)
    crit = reg.ICPConvergenceCriteria(relative_fitness=1e-6, relative_rmse=1e-6, max_iteration=50)   # stitcher.py:99-102
    res = reg.registration_colored_icp(s, t, 0.01, np.eye(4), reg.TransformationEstimationForColoredICP(), crit)
    ref = orc.registration_colored_icp(Ps, Cs, Pt, Nt, Ct, 0.01, np.eye(4), 0.968, 1e-6, 1e-6, 50)
    T, Tr = res.transformation, ref["transformation"]
    assert np.abs(T[:3, :3] - Tr[:3, :3]).max() <= 1e-5
    assert np.abs(T[:3, 3] - Tr[:3, 3]).max() <= 1e-6
    assert res.iterations == ref["iterations"]
    assert res.fitness == ref["fitness"]
    assert abs(res.inlier_rmse - ref["inlier_rmse"]) <= 1e-9
    # and it actually registers: close to the known motion
    assert np.abs(T[:3, 3] - T_true[:3, 3]).max() < 5e-4
    # errors Open3D raises
    with pytest.raises(RuntimeError):
        reg.registration_colored_icp(cloud(o3d, Ps), t, 0.01)
    with pytest.raises(RuntimeError):
        reg.registration_colored_icp(s, cloud(o3d, Pt, None, Ct), 0.01)


def test_cloud_ops_small_and_degenerate_inputs(o3d, orc):
    """fewer points than neighbours asked for, duplicates, empty clouds, a single point"""
    g = o3d.geometry
    rng = np.random.default_rng(5)
    P = rng.uniform(-0.05, 0.05, size=(7, 3))
    pc = cloud(o3d, P)
    pc.estimate_normals(g.KDTreeSearchParamKNN(30))            # k > n: all 7 points are the neighbourhood
    assert np.abs(np.asarray(pc.normals) - orc.estimate_normals(P, 0, 0.0, 30)).max() <= 1e-9
    out, idx = pc.remove_statistical_outlier(20, 2.0)
    ridx, _ = orc.statistical_outlier(P, 20, 2.0)
    assert idx == ridx.tolist()
    D = np.concatenate([P, P, P[:3]])                           # exact duplicates: ties broken by index
    from b200recon import cloud_ops
    i1, d1 = cloud_ops.knn_search(D, D, g.KDTreeSearchParamKNN(5))
    i2, d2 = orc.knn(D, D, 0, 0.0, 5)
    assert np.array_equal(i1, i2) and np.array_equal(d1, d2)
    dd = cloud(o3d, D).voxel_down_sample(0.01)
    rp, _, _ = orc.voxel_down_sample(D, 0.01)
    assert np.array_equal(np.asarray(dd.points), rp)
    one = cloud(o3d, P[:1])
    one.estimate_normals(g.KDTreeSearchParamHybrid(0.1, 30))
    assert np.array_equal(np.asarray(one.normals), [[0.0, 0.0, 1.0]])
    e = g.PointCloud()
    e.estimate_normals(g.KDTreeSearchParamHybrid(0.1, 30))
    assert len(e.points) == 0
    o, i = e.remove_statistical_outlier(20, 2.0)
    assert len(o.points) == 0 and i == []
    with pytest.raises(RuntimeError):
        pc.estimate_normals(g.KDTreeSearchParamHybrid(0.1, 1000))   # max_nn beyond the 128-neighbour heap
