"""Frame-side kernels vs the oracle: depth conversion (A.1), back-projection (A.2), synthetic ray caster."""
import numpy as np
import pytest

import helpers as hp
from b200recon import synthetic as syn

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype", [np.uint8, np.uint16, np.float32])
def test_depth_to_float(o3d, orc, dtype):
    rng = np.random.default_rng(1)
    if dtype == np.float32:
        d = (rng.random((48, 64)) * 5000).astype(np.float32)
    else:
        d = rng.integers(0, np.iinfo(dtype).max, (48, 64)).astype(dtype)
    d[0, :5] = 0
    c = np.zeros((48, 64, 3), np.uint8)
    rgbd = o3d.geometry.RGBDImage.create_from_color_and_depth(o3d.geometry.Image(c), o3d.geometry.Image(d),
                                                              depth_scale=1000.0, depth_trunc=3.0,
                                                              convert_rgb_to_intensity=False)
    got = np.asarray(rgbd.depth)
    want = orc.depth_to_float(d, 1000.0, 3.0)
    assert got.dtype == np.float32 and np.array_equal(got.view(np.uint32), want.view(np.uint32))


def test_backproject_matches_oracle(o3d, orc):
    (d16, rgb, ext), = hp.make_frames(1)
    d16[100:200, 300:400] = 0
    intr = o3d.camera.PinholeCameraIntrinsic(640, 480, *hp.KT640)
    rgbd = o3d.geometry.RGBDImage.create_from_color_and_depth(o3d.geometry.Image(rgb), o3d.geometry.Image(d16),
                                                              depth_scale=1000.0, depth_trunc=3.0,
                                                              convert_rgb_to_intensity=False)
    for E in (None, ext):
        pcd = o3d.geometry.PointCloud.create_from_rgbd_image(rgbd, intr) if E is None else \
            o3d.geometry.PointCloud.create_from_rgbd_image(rgbd, intr, E)
        P, Cc = orc.backproject(orc.depth_to_float(d16), rgb, hp.KT640, E)
        got = np.asarray(pcd.points)
        assert got.shape == P.shape and len(P) == int((orc.depth_to_float(d16) > 0).sum())
        assert np.abs(got - P).max() < 1e-12      # same order (row-major, invalid dropped)
        assert np.array_equal(np.asarray(pcd.colors), Cc)
        assert pcd.has_colors() and not pcd.has_normals()


def test_backproject_empty_and_ragged(o3d, orc):
    intr = o3d.camera.PinholeCameraIntrinsic(33, 7, 30.0, 30.0, 16.0, 3.0)
    d = np.zeros((7, 33), np.uint16)
    c = np.zeros((7, 33, 3), np.uint8)
    rgbd = o3d.geometry.RGBDImage.create_from_color_and_depth(o3d.geometry.Image(c), o3d.geometry.Image(d),
                                                              convert_rgb_to_intensity=False)
    pcd = o3d.geometry.PointCloud.create_from_rgbd_image(rgbd, intr)
    assert len(pcd.points) == 0 and pcd.is_empty()
    d[3, 5] = 1234
    d[6, 32] = 2999
    d[0, 0] = 3000   # == trunc -> dropped
    rgbd = o3d.geometry.RGBDImage.create_from_color_and_depth(o3d.geometry.Image(c), o3d.geometry.Image(d),
                                                              convert_rgb_to_intensity=False)
    pcd = o3d.geometry.PointCloud.create_from_rgbd_image(rgbd, intr)
    P, _ = orc.backproject(orc.depth_to_float(d), c, (30.0, 30.0, 16.0, 3.0))
    assert len(P) == 2 and np.abs(np.asarray(pcd.points) - P).max() < 1e-12


def test_raycast_twin_bit_identical(o3d, orc):
    K = hp.K640
    for k, pose in enumerate(syn.orbit_poses(3, radius=0.5)):
        d_ref, c_ref = orc.raycast_scene(pose, hp.KT640, 640, 480)
        d, c = syn.render_frame(pose, K)
        assert np.array_equal(d.cpu().numpy(), d_ref), k
        assert np.array_equal(c.cpu().numpy(), c_ref), k


def test_raycast_mesh_mode(o3d, orc):
    # a small tetrahedron standing on the plane
    v = np.array([[0.05, 0, 0], [-0.03, 0.04, 0], [-0.03, -0.04, 0], [0, 0, 0.06]], np.float32)
    f = np.array([[0, 1, 2], [0, 1, 3], [1, 2, 3], [2, 0, 3]])
    tris = v[f].reshape(-1, 9)
    K = hp.small_intrinsic(160, 120)
    pose = syn.orbit_poses(1, radius=0.4)[0]
    d_ref, c_ref = orc.raycast_scene(pose, hp.kt(K), 160, 120, tris=tris)
    d, c = syn.render_frame(pose, K, tris=tris)
    assert np.array_equal(d.cpu().numpy(), d_ref)
    assert np.array_equal(c.cpu().numpy(), c_ref)
    assert (c_ref[..., 0] == 230).any() or (c_ref[..., 0] == 120).any()   # object pixels exist
