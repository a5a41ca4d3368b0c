"""GPU parity of batched ICP (b2r_icp_batch through registration_icp) against the oracle's restatement of
Open3D RegistrationICP.  north_star tolerance: rotation within 1e-5, translation within
1e-6 m; fitness must agree exactly (same correspondence COUNT), inlier_rmse to 1e-9."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROT_TOL = 1e-5
TRANS_TOL = 1e-6


def surface_cloud(n, seed, noise=0.0):
    """Points + analytic normals on an ellipsoid resting on a plane patch (YCB-banana scale)."""
    rng = np.random.default_rng(seed)
    r = np.array([0.084, 0.034, 0.020])
    c = np.array([0.0, 0.0, 0.020])
    ne = n // 2
    d = rng.normal(size=(ne, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d[:, 2] = np.abs(d[:, 2])
    pe = c + d * r
    nrm_e = d / r
    nrm_e /= np.linalg.norm(nrm_e, axis=1, keepdims=True)
    pp = np.zeros((n - ne, 3))
    pp[:, :2] = rng.uniform(-0.15, 0.15, size=(n - ne, 2))
    keep = (pp[:, 0] / r[0]) ** 2 + (pp[:, 1] / r[1]) ** 2 > 0.5
    pp = pp[keep]
    nrm_p = np.tile([0.0, 0.0, 1.0], (len(pp), 1))
    P = np.concatenate([pe, pp])
    N = np.concatenate([nrm_e, nrm_p])
    if noise > 0:
        P = P + rng.normal(scale=noise, size=P.shape)
    return P, N


def se3(rx, ry, rz, t):
    ca, sa, cb, sb, cg, sg = np.cos(rx), np.sin(rx), np.cos(ry), np.sin(ry), np.cos(rz), np.sin(rz)
    Rx = np.array([[1, 0, 0], [0, ca, -sa], [0, sa, ca]])
    Ry = np.array([[cb, 0, sb], [0, 1, 0], [-sb, 0, cb]])
    Rz = np.array([[cg, -sg, 0], [sg, cg, 0], [0, 0, 1]])
    T = np.eye(4)
    T[:3, :3] = Rz @ Ry @ Rx
    T[:3, 3] = t
    return T


def make_pair(o3d, n_src, n_tgt, seed, T_true, noise=0.0):
    Ps, _ = surface_cloud(n_src, seed, noise)
    Pt, Nt = surface_cloud(n_tgt, seed + 1000, noise)
    Ti = np.linalg.inv(T_true)
    Ps = Ps @ Ti[:3, :3].T + Ti[:3, 3]        # registering src -> tgt should recover T_true
    s = o3d.geometry.PointCloud()
    s.points = o3d.utility.Vector3dVector(Ps)
    t = o3d.geometry.PointCloud()
    t.points = o3d.utility.Vector3dVector(Pt)
    t.normals = o3d.utility.Vector3dVector(Nt)
    return s, t, Ps, Pt, Nt


def check(res, ref, exact_corr=True):
    T, Tr = res.transformation, ref["transformation"]
    assert np.abs(T[:3, :3] - Tr[:3, :3]).max() <= ROT_TOL, np.abs(T[:3, :3] - Tr[:3, :3]).max()
    assert np.abs(T[:3, 3] - Tr[:3, 3]).max() <= TRANS_TOL, np.abs(T[:3, 3] - Tr[:3, 3]).max()
    assert res.iterations == ref["iterations"], (res.iterations, ref["iterations"])
    assert res.fitness == ref["fitness"], (res.fitness, ref["fitness"])
    assert abs(res.inlier_rmse - ref["inlier_rmse"]) <= 1e-9
    if exact_corr:
        cs = np.asarray(res.correspondence_set)
        assert len(cs) == ref["n_corr"]
        dense = -np.ones(len(ref["corr"]), np.int32)
        dense[cs[:, 0]] = cs[:, 1]
        assert np.array_equal(dense, ref["corr"])


@pytest.mark.parametrize("kind", [1, 0])
@pytest.mark.parametrize("max_dist", [0.05, 0.004])
def test_icp_matches_oracle(o3d, orc, kind, max_dist):
    reg = o3d.pipelines.registration
    T_true = se3(0.01, -0.015, 0.02, [0.002, -0.001, 0.0015])
    s, t, Ps, Pt, Nt = make_pair(o3d, 6000, 7000, 3, T_true, noise=0.0002)
    est = reg.TransformationEstimationPointToPlane() if kind == 1 else reg.TransformationEstimationPointToPoint()
    res = reg.registration_icp(s, t, max_dist, np.eye(4), est, reg.ICPConvergenceCriteria(1e-6, 1e-6, 30))
    ref = orc.registration_icp(Ps, Pt, Nt, max_dist, np.eye(4), kind, 1e-6, 1e-6, 30)
    check(res, ref)
    assert res.fitness > 0.3


def test_icp_known_transform_recovered(o3d):
    """Known-answer test (SURVEY 8(c) fixture 3): same surface samples moved by a small SE(3)."""
    reg = o3d.pipelines.registration
    T_true = se3(0.004, 0.003, -0.005, [0.001, 0.0005, -0.0008])
    P, N = surface_cloud(20000, 11)
    Ti = np.linalg.inv(T_true)
    s = o3d.geometry.PointCloud()
    s.points = o3d.utility.Vector3dVector(P @ Ti[:3, :3].T + Ti[:3, 3])
    t = o3d.geometry.PointCloud()
    t.points = o3d.utility.Vector3dVector(P)
    t.normals = o3d.utility.Vector3dVector(N)
    res = reg.registration_icp(s, t, 0.01, np.eye(4), reg.TransformationEstimationPointToPlane(),
                               reg.ICPConvergenceCriteria(1e-12, 1e-12, 50))
    assert res.fitness == 1.0
    assert np.abs(res.transformation - T_true).max() < 1e-6
    assert res.inlier_rmse < 1e-6


def test_icp_nonidentity_init_and_zero_iterations(o3d, orc):
    reg = o3d.pipelines.registration
    T_true = se3(0.02, 0.01, -0.03, [0.004, 0.003, -0.002])
    s, t, Ps, Pt, Nt = make_pair(o3d, 3000, 3000, 5, T_true, noise=0.0001)
    init = se3(0.018, 0.012, -0.028, [0.0035, 0.0032, -0.0018])
    res = reg.registration_icp(s, t, 0.02, init, reg.TransformationEstimationPointToPlane(),
                               reg.ICPConvergenceCriteria(1e-6, 1e-6, 5))
    ref = orc.registration_icp(Ps, Pt, Nt, 0.02, init, 1, 1e-6, 1e-6, 5)
    check(res, ref)
    ev = reg.evaluate_registration(s, t, 0.02, init)
    ref0 = orc.registration_icp(Ps, Pt, None, 0.02, init, 0, 1e-6, 1e-6, 0)
    assert ev.fitness == ref0["fitness"] and abs(ev.inlier_rmse - ref0["inlier_rmse"]) < 1e-12
    assert np.array_equal(ev.transformation, init)


def test_icp_nearest_neighbour_is_exact(o3d):
    """Independent NN oracle (SURVEY 8(c) fixture 4): scipy cKDTree vs the voxel-hash search."""
    from scipy.spatial import cKDTree
    reg = o3d.pipelines.registration
    rng = np.random.default_rng(0)
    Pt = rng.uniform(-0.1, 0.1, size=(5000, 3))
    Ps = rng.uniform(-0.12, 0.12, size=(4000, 3))
    s = o3d.geometry.PointCloud()
    s.points = o3d.utility.Vector3dVector(Ps)
    t = o3d.geometry.PointCloud()
    t.points = o3d.utility.Vector3dVector(Pt)
    for r in (0.004, 0.01, 0.05):
        ev = reg.evaluate_registration(s, t, r)
        d, j = cKDTree(Pt).query(Ps, k=1, distance_upper_bound=r)
        ok = np.isfinite(d) & (d * d < r * r)
        cs = np.asarray(ev.correspondence_set)
        assert np.array_equal(cs[:, 0], np.nonzero(ok)[0])
        assert np.array_equal(cs[:, 1], j[ok])
        assert ev.fitness == ok.sum() / len(Ps)
        assert abs(ev.inlier_rmse - np.sqrt((d[ok] ** 2).sum() / ok.sum())) < 1e-12


def test_icp_batch_ragged_and_empty(o3d, orc):
    """Batched pairs of different sizes, including an empty source, an empty target and a pair with no
    correspondence in range (Open3D: fitness = 0, inlier_rmse = 0, identity update)."""
    reg = o3d.pipelines.registration
    pairs, refs = [], []
    specs = [(2000, 2500, 1), (700, 300, 2), (5000, 4000, 3)]
    for ns, nt, seed in specs:
        T_true = se3(0.01 * seed, -0.005, 0.004, [0.001, 0.002 * seed, -0.001])
        s, t, Ps, Pt, Nt = make_pair(o3d, ns, nt, seed, T_true, noise=0.0002)
        pairs.append((s, t))
        refs.append(orc.registration_icp(Ps, Pt, Nt, 0.02, np.eye(4), 1, 1e-6, 1e-6, 30))
    empty = o3d.geometry.PointCloud()
    s0, t0 = pairs[0]
    far = o3d.geometry.PointCloud()
    far.points = o3d.utility.Vector3dVector(np.asarray(s0.points) + 10.0)
    en = o3d.geometry.PointCloud()
    en.points = o3d.utility.Vector3dVector(np.zeros((0, 3)))
    srcs = [p[0] for p in pairs] + [empty, far]
    tgts = [p[1] for p in pairs] + [t0, t0]
    out = reg.registration_icp_batch(srcs, tgts, 0.02, None, reg.TransformationEstimationPointToPlane(),
                                     reg.ICPConvergenceCriteria(1e-6, 1e-6, 30))
    for res, ref in zip(out[:3], refs):
        check(res, ref)
    for res in out[3:]:
        assert res.fitness == 0.0 and res.inlier_rmse == 0.0 and len(res.correspondence_set) == 0
        assert np.array_equal(res.transformation, np.eye(4))
    # empty target with point-to-point (no normals needed)
    out2 = reg.registration_icp_batch([s0], [en], 0.02, None, reg.TransformationEstimationPointToPoint())
    assert out2[0].fitness == 0.0 and np.array_equal(out2[0].transformation, np.eye(4))


def test_icp_errors(o3d):
    reg = o3d.pipelines.registration
    s, t, *_ = make_pair(o3d, 100, 100, 1, np.eye(4))
    with pytest.raises(RuntimeError, match="Invalid max_correspondence_distance"):
        reg.registration_icp(s, t, 0.0)
    nonrm = o3d.geometry.PointCloud()
    nonrm.points = t.points
    with pytest.raises(RuntimeError, match="require pre-computed normal vectors"):
        reg.registration_icp(s, nonrm, 0.02, np.eye(4), reg.TransformationEstimationPointToPlane())
