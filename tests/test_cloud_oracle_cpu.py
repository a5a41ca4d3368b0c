"""CPU checks of the oracle's point-cloud preprocessing (oracle/cloud_oracle.cpp, colored ICP in icp_oracle.cpp)
against independent numpy / scipy restatements -- the oracle is what the GPU tests (tests/test_cloud_gpu.py)
are compared with, so it is checked first.  No GPU needed."""
import numpy as np
import pytest
from scipy.spatial import cKDTree

from test_icp_gpu import se3, surface_cloud


def texture(P):
    return np.stack([0.5 + 0.4 * np.sin(40.0 * P[:, 0] + 1.0), 0.5 + 0.4 * np.cos(35.0 * P[:, 1]),
                     0.5 + 0.4 * np.sin(30.0 * (P[:, 0] + P[:, 1]))], 1)


def test_voxel_down_sample_vs_numpy(orc):
    P, N = surface_cloud(20000, 7, 0.0002)
    Cc = texture(P)
    vs = 0.02
    op, on, oc = orc.voxel_down_sample(P, vs, N, Cc)
    key = np.floor((P - (P.min(0) - 0.5 * vs)) / vs).astype(np.int64)
    _, first, inv = np.unique(key, axis=0, return_index=True, return_inverse=True)
    inv = inv.reshape(-1)
    order = np.argsort(first)                       # cells by first point index
    rank = np.empty(len(first), np.int64)
    rank[order] = np.arange(len(first))
    cell = rank[inv]
    assert len(op) == len(first)
    for src, out in ((P, op), (N, on), (Cc, oc)):
        ref = np.zeros((len(first), 3))
        np.add.at(ref, cell, src)
        ref /= np.bincount(cell)[:, None]
        assert np.abs(out - ref).max() < 1e-12
    with pytest.raises(RuntimeError):
        orc.voxel_down_sample(P, 0.0)
    assert len(orc.voxel_down_sample(np.zeros((0, 3)), 0.01)[0]) == 0


@pytest.mark.parametrize("mode,radius,k", [(2, 0.04, 30), (0, 0.0, 12)])
def test_estimate_normals_vs_eigh(orc, mode, radius, k):
    P, Ntrue = surface_cloud(6000, 9, 0.0002)
    Nn = orc.estimate_normals(P, mode, radius, k)
    assert np.abs(np.linalg.norm(Nn, axis=1) - 1.0).max() < 1e-12
    tree = cKDTree(P)
    for i in range(0, len(P), 37):
        d, j = tree.query(P[i], k=k, distance_upper_bound=radius if mode == 2 else np.inf)
        j = j[np.isfinite(d)]
        if len(j) < 3:
            continue
        Cm = np.cov(P[j].T, bias=True)
        w, v = np.linalg.eigh(Cm)
        if w[1] - w[0] < 1e-3 * w[2]:
            continue   # near-degenerate neighbourhoods: the eigenvector itself is ill-conditioned
        assert abs(abs(v[:, 0] @ Nn[i]) - 1.0) < 1e-8
    assert np.median(np.abs(np.sum(Nn * Ntrue, 1))) > 0.99
    # fewer than 3 neighbours -> (0,0,1); pre-existing normals fix the sign
    iso = np.concatenate([P, [[1.0, 1.0, 1.0]]])
    assert np.array_equal(orc.estimate_normals(iso, 2, 0.01, 30)[-1], [0.0, 0.0, 1.0])
    flipped = orc.estimate_normals(P, mode, radius, k, prev=-Nn)
    assert np.abs(flipped + Nn).max() < 1e-15


def test_statistical_outlier_vs_numpy(orc):
    rng = np.random.default_rng(3)
    P, _ = surface_cloud(8000, 13, 0.0003)
    P = np.concatenate([P, rng.uniform(-0.2, 0.2, size=(150, 3))])
    idx, avg = orc.statistical_outlier(P, 20, 2.0)
    d, _ = cKDTree(P).query(P, k=20)
    a = d.mean(1)
    assert np.abs(a - avg).max() < 1e-13
    thr = a.mean() + 2.0 * a.std(ddof=1)
    assert np.array_equal(idx, np.nonzero(a < thr)[0])
    assert len(idx) < len(P) - 50


def test_color_gradient_recovers_linear_field(orc):
    """intensity = g . p on a plane with normal n: the estimated gradient is the tangential part of g."""
    rng = np.random.default_rng(1)
    P = np.zeros((4000, 3))
    P[:, :2] = rng.uniform(-0.1, 0.1, size=(4000, 2))
    N = np.tile([0.0, 0.0, 1.0], (4000, 1))
    g = np.array([2.0, -1.5, 0.7])
    it = 0.5 + P @ g
    G = orc.color_gradient(P, N, np.stack([it, it, it], 1), 0.02, 30)
    assert np.abs(G - [2.0, -1.5, 0.0]).max() < 1e-8
    # fewer than 4 neighbours: gradient stays zero
    far = np.concatenate([P, [[5.0, 5.0, 0.0]]])
    G2 = orc.color_gradient(far, np.tile([0.0, 0.0, 1.0], (4001, 1)), np.stack([np.r_[it, 0.3]] * 3, 1), 0.02, 30)
    assert np.array_equal(G2[-1], [0.0, 0.0, 0.0])


def test_colored_icp_recovers_known_transform(orc):
    T_true = se3(0.006, -0.008, 0.01, [0.0015, -0.001, 0.0008])
    P, N = surface_cloud(6000, 31)
    Cc = texture(P)
    Ti = np.linalg.inv(T_true)
    Ps = P @ Ti[:3, :3].T + Ti[:3, 3]
    r = orc.registration_colored_icp(Ps, Cc, P, N, Cc, 0.01, np.eye(4), 0.968, 1e-9, 1e-9, 50)
    assert r["fitness"] == 1.0
    assert np.abs(r["transformation"] - T_true).max() < 1e-5
    # lambda_geometric = 1 switches the photometric term off: same result as point-to-plane ICP
    a = orc.registration_colored_icp(Ps, Cc, P, N, Cc, 0.01, np.eye(4), 1.0, 1e-6, 1e-6, 10)
    b = orc.registration_icp(Ps, P, N, 0.01, np.eye(4), 1, 1e-6, 1e-6, 10)
    assert np.abs(a["transformation"] - b["transformation"]).max() < 1e-12
    assert a["iterations"] == b["iterations"]
