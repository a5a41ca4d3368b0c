"""BASELINE configs[0] / SURVEY.md 8(a) row a4 on the GPU: the 12 banana frames of the simulator's camera path
(main.py:68-136,213-232; frames ray-cast from the committed YCB mesh, tests/golden/README.md) through the fused path
`pipelines.fusion.fuse_sequence` -- ICP pose -> extrinsic = inverse pose -> vol.integrate -> extract_triangle_mesh.

  * every step against the oracle ON THE ORACLE'S OWN INPUTS (the way tests/test_stitch_gpu.py does it): back-projection
    and down-sampling bit-equal, every registration within north_star's ICP tolerance (1e-5 rotation, 1e-6 m), the
    integrated volume bit-exact, the model cloud handed to the next registration identical, the final mesh identical;
  * free-running on the GPU: poses recovered from a perturbed plan, mesh within one voxel of the source banana surface;
  * the reference's own `stitcher.py` (staged by build() into oracle/_ref/, never committed) EXECUTING on the GPU:
    `RGBDStitcher.stitch_sequence` and both loaders, against the mirror sequence and the oracle."""
import importlib.util
import os
import sys

import numpy as np
import pytest

import helpers as hp
import oracle_o3d
from b200recon.pipelines import fusion

pytestmark = pytest.mark.gpu

VOXEL, TRUNC, RES = 0.002, 0.01, 192          # ~2 mm voxels (configs[0]), 0.384 m cube around the banana
REG_VOXEL, MAX_DIST = 0.004, 0.01


@pytest.fixture(scope="module")
def banana(orc):
    return hp.banana_frames(12, 320, 240)


def _origin():
    L = RES * VOXEL
    return np.array([-L / 2, -L / 2, -0.1 * L])


def to_gpu(o3d, pc):
    g = o3d.geometry.PointCloud()
    g.points = o3d.utility.Vector3dVector(pc.points)
    if pc.has_colors():
        g.colors = o3d.utility.Vector3dVector(pc.colors)
    if pc.has_normals():
        g.normals = o3d.utility.Vector3dVector(pc.normals)
    return g


@pytest.mark.parametrize("estimation,target", [("colored", "model"), ("point_to_plane", "previous")])
def test_fused_path_every_step_matches_oracle(o3d, orc, banana, estimation, target):
    K, cols, deps, poses, tris = banana
    n = 6 if target == "previous" else 12
    cols, deps, poses = cols[:n], deps[:n], poses[:n]
    init = hp.perturb_poses(poses)
    L = RES * VOXEL
    A = oracle_o3d
    vol_a = A.pipelines.integration.UniformTSDFVolume(L, RES, TRUNC, 1, _origin())
    intr_a = A.camera.PinholeCameraIntrinsic(K["width"], K["height"], *hp.kt(K))
    steps = []

    def grab(i, rgbd, next_target, result, T):
        c = A.geometry.PointCloud()
        if next_target is not None:
            c._p, c._n, c._c = next_target._p.copy(), next_target._n.copy(), None if next_target._c is None else next_target._c.copy()
        steps.append((c, result, T.copy()))
    ra = fusion.fuse_sequence(cols, deps, intr_a, vol_a, init_poses=init, max_correspondence_distance=MAX_DIST,
                              registration_voxel_size=REG_VOXEL, estimation=estimation, target=target, api=A, on_frame=grab)

    # ---- the same steps on the GPU, each from the oracle's inputs ----
    G = o3d.geometry
    intr = o3d.camera.PinholeCameraIntrinsic(K["width"], K["height"], *hp.kt(K))
    RGB8 = o3d.pipelines.integration.TSDFVolumeColorType.RGB8
    vol = o3d.pipelines.integration.UniformTSDFVolume(L, RES, TRUNC, RGB8, _origin())
    for i in range(n):
        rgbd = G.RGBDImage.create_from_color_and_depth(G.Image(cols[i]), G.Image(deps[i]), depth_scale=1000.0,
                                                       depth_trunc=3.0, convert_rgb_to_intensity=False)
        cloud = G.PointCloud.create_from_rgbd_image(rgbd, intr)
        work = cloud.voxel_down_sample(REG_VOXEL)
        Pa, _ = orc.backproject(orc.depth_to_float(deps[i]), cols[i], hp.kt(K), np.eye(4))
        assert np.array_equal(np.asarray(cloud.points), Pa)
        wa, _, _ = orc.voxel_down_sample(Pa, REG_VOXEL)
        assert np.array_equal(np.asarray(work.points), wa)
        T_a = steps[i][2]
        if i > 0:
            tgt_a, res_a = steps[i - 1][0], steps[i][1]
            init_i = steps[i - 1][2] @ np.linalg.inv(init[i - 1]) @ init[i]
            rg = fusion.register_frame(work, to_gpu(o3d, tgt_a), init_i, MAX_DIST, estimation=estimation)
            Tg = np.asarray(rg.transformation)
            assert np.abs(Tg[:3, :3] - T_a[:3, :3]).max() <= 1e-5, (i, np.abs(Tg - T_a).max())
            assert np.abs(Tg[:3, 3] - T_a[:3, 3]).max() <= 1e-6, (i, np.abs(Tg - T_a).max())
            assert rg.fitness == res_a.fitness and abs(rg.inlier_rmse - res_a.inlier_rmse) <= 1e-9
        vol.integrate(rgbd, intr, np.linalg.inv(T_a))           # the ORACLE's pose: identical input to integrate
        if target == "model" and i + 1 < n and i % 4 == 0:
            # the cloud the next registration runs against: surface points + normals of the volume so far
            pg = vol.extract_point_cloud()
            ta = steps[i][0]
            a, oa = hp.sort_rows(ta._p)
            b, ob = hp.sort_rows(np.asarray(pg.points))
            assert a.shape == b.shape and np.abs(a - b).max() <= 1e-4 * VOXEL
            assert np.abs(ta._n[oa] - np.asarray(pg.normals)[ob]).max() < 1e-9
            assert np.abs(ta._c[oa] - np.asarray(pg.colors)[ob]).max() <= 2.0 / 255.0
    ts, w, _ = vol.export_arrays()
    assert np.array_equal(w, vol_a.v.weight()) and np.array_equal(ts.view(np.uint32), vol_a.v.tsdf().view(np.uint32))
    mesh = vol.extract_triangle_mesh()
    Va, Ta = np.asarray(ra.mesh.vertices), np.asarray(ra.mesh.triangles)
    assert len(Va) > 20000
    Vs_a, Ts_a, _ = hp.canon_mesh(Va, Ta)
    Vs, Ts, _ = hp.canon_mesh(np.asarray(mesh.vertices), np.asarray(mesh.triangles))
    assert Vs.shape == Vs_a.shape and np.abs(Vs - Vs_a).max() <= 1e-4 * VOXEL and np.array_equal(Ts, Ts_a)


def test_fuse_sequence_free_running_reconstructs_the_banana(o3d, banana):
    """The product call, nothing from the oracle: poses recovered from the perturbed plan (mean error 5 mm / 0.4 deg),
    mesh within one voxel of the surface the frames were ray-cast from; and agreement with the oracle's own free run
    at the level the chaotic chain allows (packed 10:10:10 model colours vs float64 ones enter the colour term)."""
    K, cols, deps, poses, tris = banana
    init = hp.perturb_poses(poses)
    L = RES * VOXEL
    intr = o3d.camera.PinholeCameraIntrinsic(K["width"], K["height"], *hp.kt(K))
    RGB8 = o3d.pipelines.integration.TSDFVolumeColorType.RGB8
    vol = o3d.pipelines.integration.UniformTSDFVolume(L, RES, TRUNC, RGB8, _origin())
    launches0 = vol._ctx.launch_count()
    res = fusion.fuse_sequence(cols, deps, intr, vol, init_poses=init, max_correspondence_distance=MAX_DIST,
                               registration_voxel_size=REG_VOXEL, estimation="colored", target="model")
    assert vol._ctx.launch_count() - launches0 > 12 * 10            # the whole chain ran as kernels of the library
    errs = [hp.pose_error(T, P) for T, P in zip(res.poses, poses)]
    assert max(e[0] for e in errs) < 0.004 and max(e[1] for e in errs) < 0.6, errs
    assert errs[-1][0] < 0.0015, errs[-1]
    V = np.asarray(res.mesh.vertices)
    from scipy.spatial import cKDTree
    obj = V[V[:, 2] > 2 * VOXEL]
    assert len(obj) > 3000
    d, _ = cKDTree(hp.sample_triangles(tris)).query(obj)
    assert np.percentile(d, 95) < VOXEL and np.median(d) < 0.5 * VOXEL, (np.median(d), np.percentile(d, 95))
    A = oracle_o3d
    vol_a = A.pipelines.integration.UniformTSDFVolume(L, RES, TRUNC, 1, _origin())
    ra = fusion.fuse_sequence(cols, deps, A.camera.PinholeCameraIntrinsic(K["width"], K["height"], *hp.kt(K)), vol_a,
                              init_poses=init, max_correspondence_distance=MAX_DIST, registration_voxel_size=REG_VOXEL,
                              estimation="colored", target="model", api=A)
    for Tg, Ta in zip(res.poses, ra.poses):
        et, er = hp.pose_error(Tg, Ta)
        assert et < 1e-3 and er < 0.1, (et, er)
    assert abs(len(V) - len(ra.mesh.vertices)) <= 0.02 * len(V)


# ------------------------------------------------------------------------------------------------------
# the reference's own stitcher.py, executing on the GPU
# ------------------------------------------------------------------------------------------------------
REF_STAGED = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "stitcher.py")


def _load_reference_stitcher(o3d):
    """`import open3d as o3d` inside the reference file resolves to b200recon."""
    saved = sys.modules.get("open3d")
    sys.modules["open3d"] = o3d
    try:
        spec = importlib.util.spec_from_file_location("ref_stitcher_gpu", REF_STAGED)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    finally:
        if saved is None:
            sys.modules.pop("open3d", None)
        else:
            sys.modules["open3d"] = saved
    return mod


@pytest.mark.skipif(not os.path.exists(REF_STAGED), reason="oracle/_ref/stitcher.py not staged (build() copies it from "
                    "/root/reference where that exists; the file is never committed)")
def test_reference_stitcher_executes_on_gpu(o3d, orc, tmp_path, capsys):
    from stitch_mirror import Stitcher
    import cv2
    mod = _load_reference_stitcher(o3d)
    from b200recon import synthetic as syn
    W, H = 320, 240
    K = hp.small_intrinsic(W, H)
    posesq = syn.orbit_poses(5, radius=0.45, elevation_deg=35.0, az_span=np.deg2rad(2.0) * 5)
    cols, deps = [], []
    for p in posesq:
        d16, rgb = orc.raycast_scene(p, hp.kt(K), W, H, checker=0.03)
        cols.append(rgb)
        deps.append(d16)
    intr = o3d.camera.PinholeCameraIntrinsic(W, H, *hp.kt(K))
    st = mod.RGBDStitcher(intr)
    assert (st.voxel_size, st.distance_threshold, st.optimization_modulus) == (0.02, 0.05, 2)     # stitcher.py:17-19

    # loaders: the simulator's folder layout (stitcher.py:202-233) and the recorder's (stitcher.py:236-258)
    cam = tmp_path / "camera"
    (cam / "rgb").mkdir(parents=True)
    (cam / "depth").mkdir()
    rs_rgb, rs_depth = tmp_path / "rs" / "rgb", tmp_path / "rs" / "depth"
    rs_rgb.mkdir(parents=True)
    rs_depth.mkdir()
    for i, (c, d) in enumerate(zip(cols, deps)):
        cv2.imwrite(str(cam / "rgb" / ("rgb_%04d.png" % i)), c)
        cv2.imwrite(str(cam / "depth" / ("depth_%04d.png" % i)), d)          # uint16 mm PNG
        cv2.imwrite(str(rs_rgb / ("%04d.png" % i)), c)
        np.save(str(rs_depth / ("%04d.npy" % i)), d)
    c2, d2 = st.load_dataset_two_folders(str(cam), "rgb", "depth")
    assert len(c2) == len(d2) == 4 and all(np.array_equal(a, b) for a, b in zip(c2, cols)) and \
        all(np.array_equal(a, b) for a, b in zip(d2, deps))                  # 5 frames -> 4 (stitcher.py:228-231)
    c3, d3 = st.load_dataset_realsense(str(rs_rgb), str(rs_depth))
    assert len(c3) == 4 and all(np.array_equal(a, b) for a, b in zip(d3, deps))
    cwd = os.getcwd()
    os.chdir(tmp_path)
    try:
        c4, d4 = st.load_default()                                           # './camera' (stitcher.py:202-206)
    finally:
        os.chdir(cwd)
    assert len(c4) == 4 and np.array_equal(d4[3], deps[3])

    # the hot loop itself: stitcher.py:114-166 verbatim, every o3d.* call a kernel of libb200recon.so
    ctx = __import__("b200recon").runtime.ctx()
    l0 = ctx.launch_count()
    combined = st.stitch_sequence(cols, deps)
    assert ctx.launch_count() - l0 > 100
    out = capsys.readouterr().out
    assert "Number of points:" in out and "Combined cloud type" in out       # the prints of stitcher.py:141-142 ran
    Pc = np.asarray(combined.points)
    assert len(Pc) > 500 and combined.has_colors()
    # its pieces
    pc0 = st.create_point_cloud_from_rgbd(cols[0], deps[0])
    Pa, Ca = orc.backproject(orc.depth_to_float(deps[0]), cols[0], hp.kt(K), np.eye(4))
    assert np.array_equal(np.asarray(pc0.points), Pa) and np.array_equal(np.asarray(pc0.colors), Ca)
    down, fpfh = st.preprocess_point_cloud(pc0)
    assert down.has_normals() and fpfh.data.shape == (33, len(down.points))
    T, fit = st.register_point_clouds(st.create_point_cloud_from_rgbd(cols[1], deps[1]), pc0)
    assert np.asarray(T).shape == (4, 4) and fit > 0.5
    # same sequence through the mirror on the same back end: identical (the GPU path is deterministic)
    m = Stitcher(o3d, intr).stitch(cols, deps)
    assert np.array_equal(np.asarray(m.points), Pc)
    # and against the oracle's free run: agreement at the scale of the voxel size (chaotic chain, see test_stitch_gpu.py)
    b = Stitcher(oracle_o3d, oracle_o3d.camera.PinholeCameraIntrinsic(W, H, *hp.kt(K))).stitch(cols, deps)
    Pb = np.asarray(b.points)
    assert abs(len(Pc) - len(Pb)) <= 0.10 * len(Pb)
    from scipy.spatial import cKDTree
    dd, _ = cKDTree(Pb).query(Pc, k=1)
    assert np.median(dd) < st.voxel_size
