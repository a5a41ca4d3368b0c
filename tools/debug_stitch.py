import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "3d-reconstruction-planning_b200")]
import numpy as np
import b200recon as o3d
import helpers as hp, oracle_o3d
from oracle import oracle as orc
from stitch_mirror import Stitcher
from test_stitch_gpu import frames
from b200recon import cloud_ops
K, cols, deps = frames(5)
for scene in ("default",):
    a = Stitcher(o3d, o3d.camera.PinholeCameraIntrinsic(K["width"], K["height"], *hp.kt(K)))
    b = Stitcher(oracle_o3d, oracle_o3d.camera.PinholeCameraIntrinsic(K["width"], K["height"], *hp.kt(K)))
    c0a, c1a = a.cloud(cols[0], deps[0]), a.cloud(cols[1], deps[1])
    c0b, c1b = b.cloud(cols[0], deps[0]), b.cloud(cols[1], deps[1])
    print("backproject diff", np.abs(np.asarray(c0a.points) - c0b.points).max(), np.abs(np.asarray(c0a.colors) - c0b.colors).max())
    sa, _ = a.preprocess(c1a); ta, _ = a.preprocess(c0a)
    sb, _ = b.preprocess(c1b); tb, _ = b.preprocess(c0b)
    print("down diff", np.abs(np.asarray(sa.points) - sb.points).max(), "normals diff", np.abs(np.asarray(ta.normals) - tb.normals).max(),
          "colors", np.abs(np.asarray(ta.colors) - tb.colors).max())
    ga = cloud_ops.color_gradient(ta, 0.1, 30).cpu().numpy()
    gb = orc.color_gradient(tb.points, tb.normals, tb.colors, 0.1, 30)
    print("grad diff", np.abs(ga - gb).max(), np.abs(gb).max())
    reg = o3d.pipelines.registration
    for it in (0, 1, 2, 3, 5, 10, 20, 50):
        ra = reg.registration_colored_icp(sa, ta, 0.05, np.eye(4), reg.TransformationEstimationForColoredICP(), reg.ICPConvergenceCriteria(1e-6, 1e-6, it))
        rb = orc.registration_colored_icp(sb.points, sb.colors, tb.points, tb.normals, tb.colors, 0.05, np.eye(4), 0.968, 1e-6, 1e-6, it, tgt_grad=gb)
        print(it, "T diff", np.abs(ra.transformation - rb["transformation"]).max(), ra.fitness, rb["fitness"], ra.inlier_rmse - rb["inlier_rmse"], ra.iterations, rb["iterations"])
