import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests"))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "3d-reconstruction-planning_b200"))
import numpy as np
import b200recon as o3d
from oracle import oracle as orc
import test_icp_gpu as tig
orc.build()
reg = o3d.pipelines.registration
T_true = tig.se3(0.01, -0.015, 0.02, [0.002, -0.001, 0.0015])
s, t, Ps, Pt, Nt = tig.make_pair(o3d, 6000, 7000, 3, T_true, noise=0.0002)
for r in (0.05, 0.004):
    for it in (0, 1, 2, 3, 30):
        res = reg.registration_icp(s, t, r, np.eye(4), reg.TransformationEstimationPointToPlane(), reg.ICPConvergenceCriteria(1e-6, 1e-6, it))
        ref = orc.registration_icp(Ps, Pt, Nt, r, np.eye(4), 1, 1e-6, 1e-6, it)
        cs = np.asarray(res.correspondence_set)
        dense = -np.ones(len(ref["corr"]), np.int32); dense[cs[:, 0]] = cs[:, 1]
        print("r", r, "maxit", it, "gpu it", res.iterations, "ref it", ref["iterations"], "fit", res.fitness, ref["fitness"],
              "rmse", res.inlier_rmse, ref["inlier_rmse"], "dT", np.abs(res.transformation - ref["transformation"]).max(),
              "corr mismatch", int((dense != ref["corr"]).sum()))
# ---- first-iteration normal equations, numpy vs device
os.environ["B2R_ICP_DEBUG"] = "1"
from scipy.spatial import cKDTree
r = 0.05
d, j = cKDTree(Pt).query(Ps, k=1, distance_upper_bound=r)
ok = np.isfinite(d)
q = Ps[ok]; tt = Pt[j[ok]]; nn = Nt[j[ok]]
rr = ((q - tt) * nn).sum(1)
J = np.concatenate([np.cross(q, nn), nn], 1)
JTJ = J.T @ J; JTr = J.T @ rr
x = np.linalg.solve(JTJ, -JTr)
print("numpy K", ok.sum(), "sum d2", (d[ok] ** 2).sum())
print("numpy JTJ upper", [JTJ[a, b] for a in range(6) for b in range(a, 6)])
print("numpy JTr", JTr, "x", x)
sys.stdout.flush()
res = reg.registration_icp(s, t, r, np.eye(4), reg.TransformationEstimationPointToPlane(), reg.ICPConvergenceCriteria(1e-6, 1e-6, 1))
