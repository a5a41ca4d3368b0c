"""CPU checks of the ORACLE itself (no GPU): the C++ restatement (oracle/*.cpp) against
  * the independent numpy float32 twin (oracle/numpy_twin.py) -- bit-exact TSDF/weights,
  * the marching-cubes table invariants and sha1 digests recorded in SURVEY.md Appendix B,
  * closed-form geometry (sphere SDF: watertight, outward-facing, vertices on the sphere),
  * scipy.spatial.cKDTree for the exact nearest-neighbour search,
  * known-answer ICP (recover a known SE(3)),
  * the committed golden vectors in tests/golden/ (regression pins of the oracle; the reference has
    no tests or golden vectors of its own -- SURVEY.md section 4 -- so parity vs the real Open3D wheel
    stays "unpinned").
"""
import hashlib
import os
import re

import numpy as np
import pytest

import helpers as hp
from conftest import ROOT
from b200recon import synthetic as syn


def _tables():
    txt = open(os.path.join(ROOT, "oracle", "mc_tables.inc")).read()
    txt = re.sub(r"//.*", "", txt)
    arrays = {}
    for m in re.finditer(r"(\w+)\s*((?:\[\d*\])+)\s*=\s*\{(.*?)\};", txt, re.S):
        nums = [int(v, 0) for v in re.findall(r"-?0x[0-9a-fA-F]+|-?\d+", m.group(3))]
        arrays[m.group(1).lower()] = nums
    return arrays


def test_mc_tables_match_survey_digests():
    t = _tables()
    edge = next(v for k, v in t.items() if "edge_table" in k or k.endswith("edgetable"))
    tri = next(v for k, v in t.items() if "tri_table" in k or k.endswith("tritable"))
    assert len(edge) == 256 and len(tri) == 256 * 16
    assert edge[1] == 0x109 and edge[3] == 0x30a and edge[255] == 0 and sum(edge) == 524160
    assert hashlib.sha1(",".join(map(str, edge)).encode()).hexdigest() == "29366ad64889e0d89bddfe2371eeee6300c591ec"
    assert hashlib.sha1(",".join(str(v) for v in tri).encode()).hexdigest() == "ed05a8b5140ea5b6f72036640886e3f856cdd991"
    # edge_table == sign-change mask of edge_to_vert; every tri_table row uses exactly those edges
    e2v = [(0, 1), (1, 2), (3, 2), (0, 3), (4, 5), (5, 6), (7, 6), (4, 7), (0, 4), (1, 5), (2, 6), (3, 7)]
    hist = {}
    for c in range(256):
        mask = 0
        for e, (a, b) in enumerate(e2v):
            if ((c >> a) & 1) != ((c >> b) & 1):
                mask |= 1 << e
        assert mask == edge[c]
        row = [v for v in tri[c * 16:(c + 1) * 16] if v >= 0]
        assert len(row) % 3 == 0
        used = 0
        for v in row:
            used |= 1 << v
        assert used == edge[c]
        hist[len(row) // 3] = hist.get(len(row) // 3, 0) + 1
    assert hist == {0: 2, 1: 16, 2: 50, 3: 80, 4: 76, 5: 32}


def _small_frames(n=3, W=80, H=60):
    K = hp.small_intrinsic(W, H)
    return K, hp.make_frames(n, K)


def test_integrate_cpp_matches_numpy_twin(orc):
    from oracle import numpy_twin as tw
    K, frames = _small_frames()
    R, L, trunc = 28, 0.42, 0.03
    origin = syn.volume_origin(L)
    vol = orc.UniformTSDFVolume(L, R, trunc, 1, origin)
    ts = np.zeros((R, R, R), np.float32)
    w = np.zeros((R, R, R), np.float32)
    col = np.zeros((R, R, R, 3), np.float64)
    for d16, rgb, ext in frames:
        df = orc.depth_to_float(d16)
        n_c = vol.integrate(df, rgb, hp.kt(K), ext)
        n_t = tw.integrate(ts, w, col, df, rgb, hp.kt(K), ext, L, trunc, origin)
        assert n_c == n_t and n_c > 0
    assert np.array_equal(w, vol.weight())
    assert np.array_equal(ts.view(np.uint32), vol.tsdf().view(np.uint32))
    assert np.abs(col - vol.color()).max() < 1e-9


def test_depth_conversion_and_backprojection(orc):
    rng = np.random.default_rng(5)
    H, W = 12, 16
    d16 = rng.integers(0, 4000, size=(H, W)).astype(np.uint16)
    d16[0, :4] = 0
    df = orc.depth_to_float(d16)
    ref = d16.astype(np.float32) / np.float32(1000.0)
    ref[ref >= np.float32(3.0)] = 0
    assert np.array_equal(df, ref)
    d8 = rng.integers(0, 255, size=(H, W)).astype(np.uint8)
    assert np.array_equal(orc.depth_to_float(d8), (d8.astype(np.float32) / np.float32(255.0)) / np.float32(1000.0))
    rgb = rng.integers(0, 255, size=(H, W, 3)).astype(np.uint8)
    Kt = (20.0, 21.0, 7.5, 6.25)
    ext = np.eye(4)
    ext[:3, 3] = [0.1, -0.2, 0.3]
    P, Cc = orc.backproject(df, rgb, Kt, ext)
    v, u = np.nonzero(df > 0)          # row-major pixel order, invalid dropped
    z = df[v, u].astype(np.float64)
    cam = np.stack([(u - Kt[2]) * z / Kt[0], (v - Kt[3]) * z / Kt[1], z], 1)
    pose = np.linalg.inv(ext)
    assert np.allclose(P, cam @ pose[:3, :3].T + pose[:3, 3], atol=1e-12)
    assert np.array_equal(Cc, rgb[v, u].astype(np.float64) / 255.0)


def _sphere_volume(orc, R=40, L=1.0, r=0.3, color_type=0):
    vol = orc.UniformTSDFVolume(L, R, 0.1, color_type, (-L / 2, -L / 2, -L / 2))
    vl = L / R
    c = (np.arange(R) + 0.5) * vl - L / 2
    X, Y, Z = np.meshgrid(c, c, c, indexing="ij")
    sdf = np.sqrt(X * X + Y * Y + Z * Z) - r
    vol.tsdf()[...] = np.clip(sdf / 0.1, -1, 1).astype(np.float32)
    vol.weight()[...] = 1.0
    return vol, vl


def test_marching_cubes_sphere_is_closed_and_outward(orc):
    vol, vl = _sphere_volume(orc)
    V, T, _ = vol.extract_triangle_mesh()
    assert len(T) > 1000
    rad = np.linalg.norm(V, axis=1)
    assert np.abs(rad - 0.3).max() < 0.25 * vl           # linear interpolation of a near-linear field
    # watertight: every undirected edge is shared by exactly two triangles, with opposite directions
    e = np.concatenate([T[:, [0, 1]], T[:, [1, 2]], T[:, [2, 0]]]).astype(np.int64)
    key = e[:, 0] * (len(V) + 1) + e[:, 1]
    rkey = e[:, 1] * (len(V) + 1) + e[:, 0]
    assert len(np.unique(key)) == len(key)
    assert np.array_equal(np.sort(key), np.sort(rkey))
    # Open3D's swapped winding (i, i+2, i+1) faces free space (positive TSDF): outward for a sphere
    n = np.cross(V[T[:, 1]] - V[T[:, 0]], V[T[:, 2]] - V[T[:, 0]])
    cen = V[T].mean(1)
    assert ((n * cen).sum(1) > 0).mean() > 0.999
    # welded: no duplicated vertex
    assert len(np.unique(np.round(V, 12), axis=0)) == len(V)


def test_marching_cubes_skips_unobserved_cubes(orc):
    vol, _ = _sphere_volume(orc, R=24)
    nfull = len(vol.extract_triangle_mesh()[1])
    vol.weight()[:12] = 0.0                                 # half the cube never observed
    V, T, _ = vol.extract_triangle_mesh()
    assert 0 < len(T) < nfull
    assert V[:, 0].min() > -0.5 + 12 * (1.0 / 24)           # no vertex inside the unobserved half


def test_extract_point_cloud_sphere(orc):
    vol, vl = _sphere_volume(orc, R=32)
    P, N, _ = vol.extract_point_cloud()
    assert len(P) > 500
    rad = np.linalg.norm(P, axis=1)
    assert np.abs(rad - 0.3).max() < 0.25 * vl
    assert np.abs(np.linalg.norm(N, axis=1) - 1).max() < 1e-9
    assert ((N * P).sum(1) / rad).min() > 0.95              # gradient normals point outward


def test_nearest_matches_ckdtree(orc):
    from scipy.spatial import cKDTree
    rng = np.random.default_rng(1)
    tgt = rng.uniform(-0.1, 0.1, size=(3000, 3))
    q = rng.uniform(-0.12, 0.12, size=(2000, 3))
    for r in (0.004, 0.02, 0.05):
        idx, d2 = orc.nearest(tgt, q, r)
        d, j = cKDTree(tgt).query(q, k=1, distance_upper_bound=r)
        ok = np.isfinite(d) & (d * d < r * r)
        assert np.array_equal(idx >= 0, ok)
        assert np.array_equal(idx[ok], j[ok])
        assert np.allclose(d2[ok], d[ok] ** 2, rtol=0, atol=1e-15)


@pytest.mark.parametrize("kind", [0, 1])
def test_icp_recovers_known_transform(orc, kind):
    import test_icp_gpu as tig
    T_true = tig.se3(0.004, 0.003, -0.005, [0.001, 0.0005, -0.0008])
    P, N = tig.surface_cloud(6000, 11)
    Ti = np.linalg.inv(T_true)
    src = P @ Ti[:3, :3].T + Ti[:3, 3]
    res = orc.registration_icp(src, P, N, 0.01, np.eye(4), kind, 1e-12, 1e-12, 60)
    assert res["fitness"] == 1.0
    assert np.abs(res["transformation"] - T_true).max() < 1e-6
    assert res["inlier_rmse"] < 1e-6


def test_icp_degenerate_inputs(orc):
    P = np.random.default_rng(0).uniform(size=(50, 3))
    far = P + 10.0
    res = orc.registration_icp(far, P, None, 0.02, np.eye(4), 0, 1e-6, 1e-6, 30)
    assert res["fitness"] == 0.0 and res["inlier_rmse"] == 0.0 and res["n_corr"] == 0
    assert np.array_equal(res["transformation"], np.eye(4))
    with pytest.raises(RuntimeError):
        orc.registration_icp(P, P, None, 0.0, np.eye(4), 0)
    with pytest.raises(RuntimeError):
        orc.registration_icp(P, P, None, 0.02, np.eye(4), 1)    # point-to-plane without normals


def test_raycast_scene_is_deterministic_and_plausible(orc):
    K = hp.small_intrinsic(160, 120)
    pose = syn.orbit_poses(4, radius=0.5)[1]
    d1, c1 = orc.raycast_scene(pose, hp.kt(K), 160, 120)
    d2, c2 = orc.raycast_scene(pose, hp.kt(K), 160, 120)
    assert np.array_equal(d1, d2) and np.array_equal(c1, c2)
    assert d1.min() > 0 and 300 < np.median(d1) < 1500      # plane fills the frame, camera ~0.5 m away
    assert d1[60, 80] < np.median(d1)                        # the object is in front of the plane


# ------------------------------------------------------------------------------------- golden vectors
GOLD = os.path.join(ROOT, "tests", "golden", "oracle_golden.npz")


def test_golden_vectors(orc):
    """tests/golden/oracle_golden.npz is written by tests/golden/make_golden.py from the oracle; this pins the
    oracle against silent drift (compiler flags, refactors).  Inputs are regenerated from seeds."""
    import golden.make_golden as mg
    g = np.load(GOLD)
    cur = mg.compute(orc)
    for k in g.files:
        a, b = g[k], cur[k]
        assert a.shape == b.shape and a.dtype == b.dtype, k
        if a.dtype.kind == "f" and k.startswith(("icp_", "cicp_", "normals_", "grad_", "outlier_stats", "scal_vertex_sum", "scal_point_sum", "scal_normal_sum")):   # OpenMP / libm last bits
            assert np.allclose(a, b, rtol=0, atol=1e-12), k
        else:
            assert np.array_equal(a, b), k
