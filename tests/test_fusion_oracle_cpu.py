"""BASELINE configs[0] on the CPU ORACLE alone (no GPU): the 12 banana frames of the simulator's camera path through
`pipelines.fusion.fuse_sequence` with the oracle wrapped in Open3D's names (tests/oracle_o3d.py).  Checks the host
logic of the fused path (pose chaining, extrinsic = inverse pose, target choice) and that the configuration the GPU
test and `bench.py --workload banana12` use actually reconstructs the banana: poses recovered from a perturbed planned
trajectory, mesh on the source surface."""
import numpy as np
import pytest

import helpers as hp
import oracle_o3d
from b200recon.pipelines import fusion


@pytest.fixture(scope="module")
def banana(orc):
    return hp.banana_frames(12, 160, 120)


def _volume(R=128, vl=0.003):
    L = R * vl
    return oracle_o3d.pipelines.integration.UniformTSDFVolume(L, R, 0.012, 1, (-L / 2, -L / 2, -0.1 * L)), vl


def test_fuse_sequence_banana_colored_frame_to_model(banana):
    K, cols, deps, poses, tris = banana
    intr = oracle_o3d.camera.PinholeCameraIntrinsic(K["width"], K["height"], *hp.kt(K))
    vol, vl = _volume()
    init = hp.perturb_poses(poses)
    seen = []
    res = fusion.fuse_sequence(cols, deps, intr, vol, init_poses=init, max_correspondence_distance=0.012,
                               registration_voxel_size=0.005, estimation="colored", target="model", api=oracle_o3d,
                               on_frame=lambda i, rgbd, cloud, r, T: seen.append(i))
    assert seen == list(range(12)) and len(res.poses) == 12 and res.registrations[0] is None
    assert np.array_equal(res.poses[0], poses[0])
    errs = [hp.pose_error(T, P) for T, P in zip(res.poses, poses)]
    errs0 = [hp.pose_error(I, P) for I, P in zip(init, poses)]
    # the registration improves on the perturbed plan (mean 5 mm / 0.4 degrees) and does not drift
    assert np.mean([e[0] for e in errs[1:]]) < 0.5 * np.mean([e[0] for e in errs0[1:]])
    assert max(e[0] for e in errs) < 0.004 and max(e[1] for e in errs) < 0.6
    V = np.asarray(res.mesh.vertices)
    assert len(V) > 5000 and len(res.mesh.triangles) > 10000
    from scipy.spatial import cKDTree
    obj = V[V[:, 2] > 2 * vl]                       # vertices above the table = the banana
    assert len(obj) > 1000
    d, _ = cKDTree(hp.sample_triangles(tris)).query(obj)
    assert np.percentile(d, 95) < vl and np.median(d) < 0.5 * vl        # within one voxel of the source surface


def test_fuse_sequence_host_logic(banana):
    K, cols, deps, poses, tris = banana
    intr = oracle_o3d.camera.PinholeCameraIntrinsic(K["width"], K["height"], *hp.kt(K))
    vol, _ = _volume(64, 0.006)
    with pytest.raises(ValueError):
        fusion.fuse_sequence(cols[:3], deps[:2], intr, vol, api=oracle_o3d)
    with pytest.raises(ValueError):
        fusion.fuse_sequence(cols[:3], deps[:3], intr, vol, init_poses=poses[:2], api=oracle_o3d)
    with pytest.raises(ValueError):
        fusion.fuse_sequence(cols[:2], deps[:2], intr, vol, estimation="fpfh", init_poses=poses[:2], api=oracle_o3d)
    # without init poses the first frame defines the world; point-to-plane against the previous frame
    r = fusion.fuse_sequence(cols[:2], deps[:2], intr, vol, max_correspondence_distance=0.02, registration_voxel_size=0.006,
                             api=oracle_o3d, extract_mesh=False)
    assert np.array_equal(r.poses[0], np.eye(4)) and r.mesh is None and r.registrations[1].fitness > 0.5
