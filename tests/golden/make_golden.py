"""Writes tests/golden/oracle_golden.npz: small outputs of the ORACLE on seeded inputs.

The reference (geconf/3d-reconstruction-planning) has no tests, golden vectors or saved outputs for this
path, and Open3D 0.18 -- where its arithmetic lives -- cannot be imported here, so these vectors pin the
oracle against itself (drift detection), not against the real wheel: parity stays "unpinned".

    python tests/golden/make_golden.py        # regenerate (only when the oracle changes on purpose)
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.join(ROOT, "3d-reconstruction-planning_b200"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def _digest(a):
    return np.frombuffer(hashlib.sha1(np.ascontiguousarray(a).tobytes()).digest(), np.uint8).copy()


def compute(orc):
    import helpers as hp
    import test_icp_gpu as tig
    from b200recon import synthetic as syn
    out = {}
    # --- TSDF integrate + extraction on a 32^3 cube, 3 frames of 96x72
    K = hp.small_intrinsic(96, 72)
    frames = hp.make_frames(3, K)
    R, L, trunc = 32, 0.384, 0.036
    vol = orc.UniformTSDFVolume(L, R, trunc, 1, syn.volume_origin(L))
    n = 0
    for d16, rgb, ext in frames:
        n += vol.integrate(orc.depth_to_float(d16), rgb, hp.kt(K), ext)
    out["frame0_depth_sha1"] = _digest(frames[0][0])
    out["frame0_color_sha1"] = _digest(frames[0][1])
    out["n_updates"] = np.array([n], np.int64)
    out["tsdf_sha1"] = _digest(vol.tsdf())
    out["weight_sha1"] = _digest(vol.weight())
    out["tsdf_slice"] = vol.tsdf()[16, 10:22, 4:12].copy()
    out["weight_slice"] = vol.weight()[16, 10:22, 4:12].copy()
    V, T, Cc = vol.extract_triangle_mesh()
    out["mesh_counts"] = np.array([len(V), len(T)], np.int64)
    out["mesh_vertices_sha1"] = _digest(V)
    out["mesh_triangles_sha1"] = _digest(T)
    out["mesh_first_vertices"] = V[:8].copy()
    P, N, _ = vol.extract_point_cloud()
    out["pcd_count"] = np.array([len(P)], np.int64)
    out["pcd_points_sha1"] = _digest(P)
    # --- ICP, both estimators
    T_true = tig.se3(0.01, -0.015, 0.02, [0.002, -0.001, 0.0015])
    Ps, _ = tig.surface_cloud(1500, 3, 0.0002)
    Pt, Nt = tig.surface_cloud(1800, 1003, 0.0002)
    Ti = np.linalg.inv(T_true)
    Ps = Ps @ Ti[:3, :3].T + Ti[:3, 3]
    for kind in (0, 1):
        r = orc.registration_icp(Ps, Pt, Nt, 0.01, np.eye(4), kind, 1e-6, 1e-6, 30)
        out["icp_T_%d" % kind] = r["transformation"]
        out["icp_stats_%d" % kind] = np.array([r["fitness"], r["inlier_rmse"]])
        out["corr_sha1_%d" % kind] = _digest(r["corr"])
        out["iters_%d" % kind] = np.array([r["iterations"], r["n_corr"]], np.int64)
    # --- point-cloud preprocessing (rows f1/f2) and colored ICP (f3)
    import test_cloud_oracle_cpu as tco
    P, N = tig.surface_cloud(3000, 7, 0.0002)
    Cc = tco.texture(P)
    dp, dn, dc = orc.voxel_down_sample(P, 0.01, N, Cc)
    out["vds_count"] = np.array([len(dp)], np.int64)
    out["vds_points_sha1"] = _digest(dp)
    out["vds_colors_sha1"] = _digest(dc)
    out["vds_first_points"] = dp[:6].copy()
    nn = orc.estimate_normals(dp, orc.SEARCH_HYBRID, 0.03, 30)
    out["normals_first"] = nn[:6].copy()
    out["normals_sum"] = nn.sum(0)
    idx, avg = orc.statistical_outlier(P, 20, 2.0)
    out["outlier_kept_sha1"] = _digest(idx)
    out["outlier_stats"] = np.array([len(idx), avg.mean()], np.float64)
    g = orc.color_gradient(dp, nn, dc, 0.03, 30)
    out["grad_first"] = g[:6].copy()
    Ps2 = P @ Ti[:3, :3].T + Ti[:3, 3]
    r = orc.registration_colored_icp(Ps2, Cc, P, N, Cc, 0.01, np.eye(4), 0.968, 1e-6, 1e-6, 50)
    out["cicp_T"] = r["transformation"]
    out["cicp_stats"] = np.array([r["fitness"], r["inlier_rmse"]])
    out["cicp_iters"] = np.array([r["iterations"], r["n_corr"]], np.int64)
    # --- ScalableTSDFVolume (row f4): 2 frames of 160x120 into 8^3 units of 6 mm voxels, stride 4
    K2 = hp.small_intrinsic(160, 120)
    sv = orc.ScalableTSDFVolume(0.006, 0.024, 1, 8, 4)
    ns = 0
    for d16, rgb, ext in hp.make_frames(2, K2, radius=0.35, elevation=65.0):
        ns += sv.integrate(orc.depth_to_float(d16), rgb, hp.kt(K2), ext)
    keys = sv.unit_keys()
    order = np.lexsort((keys[:, 2], keys[:, 1], keys[:, 0]))
    keys = keys[order]
    out["scal_updates_units"] = np.array([ns, len(keys), sv.touched], np.int64)
    out["scal_keys_sha1"] = _digest(keys)
    out["scal_tsdf_sha1"] = _digest(np.stack([sv.unit_arrays(k)[0] for k in keys]))
    out["scal_weight_sha1"] = _digest(np.stack([sv.unit_arrays(k)[1] for k in keys]))
    out["scal_color_sha1"] = _digest(np.stack([sv.unit_arrays(k)[2] for k in keys]))
    Vs, Ts, _ = sv.extract_triangle_mesh()
    Ps, Ns, _ = sv.extract_point_cloud()
    out["scal_counts"] = np.array([len(Vs), len(Ts), len(Ps)], np.int64)
    out["scal_vertex_sum"] = Vs.sum(0)
    out["scal_point_sum"] = Ps.sum(0)
    out["scal_normal_sum"] = Ns.sum(0)
    return out


if __name__ == "__main__":
    from oracle import oracle
    oracle.build()
    np.savez(os.path.join(HERE, "oracle_golden.npz"), **compute(oracle))
    print("wrote", os.path.join(HERE, "oracle_golden.npz"))
