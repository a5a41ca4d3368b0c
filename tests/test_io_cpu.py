"""b200recon/io.py: the on-disk formats either side of the path.  Host Python only -- no GPU:
PLY write -> read round trips (cloud and mesh, colours and normals), the two frame layouts the reference writes
(bullet_camera.py:71-85 `camera/rgb/rgb_%04d.png` + `camera/depth/depth_%04d.png`; data_recorder.py:236-243,305-315
`rgb/%04d.jpg` + `depth/%04d.npy` + `metadata.json`), the loaders' count rule (stitcher.py:228-231,253-256) and the
pose source `wtraj.txt` (main.py:156-165, xyzw quaternions per Expansion-GRR/grr/utils.py:101-105)."""
import json
import os

import numpy as np
import pytest

import helpers as hp
from b200recon import geometry, io


def _cloud(n=500, seed=0, colors=True, normals=True):
    rng = np.random.default_rng(seed)
    pc = geometry.PointCloud()
    pc.points = rng.uniform(-1, 1, (n, 3))
    if colors:
        pc.colors = rng.integers(0, 256, (n, 3)) / 255.0
    if normals:
        v = rng.normal(size=(n, 3))
        pc.normals = v / np.linalg.norm(v, axis=1, keepdims=True)
    return pc


@pytest.mark.parametrize("ascii_", [False, True])
@pytest.mark.parametrize("colors,normals", [(True, True), (False, False), (True, False)])
def test_ply_point_cloud_round_trip(tmp_path, ascii_, colors, normals):
    pc = _cloud(300, 1, colors, normals)
    f = str(tmp_path / "merged_cloud.ply")
    assert io.write_point_cloud(f, pc, write_ascii=ascii_)
    back = io.read_point_cloud(f)
    assert np.array_equal(np.asarray(back.points), np.asarray(pc.points))          # float64 properties: exact
    assert back.has_colors() == colors and back.has_normals() == normals
    if colors:      # colours are stored as uchar like Open3D does: exact for k/255 inputs
        assert np.abs(np.asarray(back.colors) - np.asarray(pc.colors)).max() < 1e-12
    if normals:
        assert np.array_equal(np.asarray(back.normals), np.asarray(pc.normals))


def test_ply_empty_cloud(tmp_path):
    f = str(tmp_path / "empty.ply")
    io.write_point_cloud(f, geometry.PointCloud())
    assert len(io.read_point_cloud(f).points) == 0


def test_ply_mesh_round_trip(tmp_path):
    rng = np.random.default_rng(2)
    m = geometry.TriangleMesh()
    m.vertices = rng.uniform(-1, 1, (200, 3))
    m.triangles = rng.integers(0, 200, (350, 3))
    m.vertex_colors = rng.integers(0, 256, (200, 3)) / 255.0
    f = str(tmp_path / "mesh.ply")
    assert io.write_triangle_mesh(f, m)
    b = io.read_triangle_mesh(f)
    assert np.array_equal(np.asarray(b.vertices), np.asarray(m.vertices))
    assert np.array_equal(np.asarray(b.triangles), np.asarray(m.triangles))
    assert np.abs(np.asarray(b.vertex_colors) - np.asarray(m.vertex_colors)).max() < 1e-12


def test_reads_the_reference_banana_mesh():
    """The committed YCB mesh (binary PLY, float vertices, `list uchar int` faces) -- the geometry of configs[0]."""
    m = io.read_triangle_mesh(hp.BANANA_PLY)
    V, T = np.asarray(m.vertices), np.asarray(m.triangles)
    assert V.shape == (7451, 3) and T.shape == (14898, 3)
    assert T.min() == 0 and T.max() == 7450
    ext = V.max(0) - V.min(0)
    assert np.allclose(ext, [0.161, 0.125, 0.047], atol=2e-3)


def _write_frames(cols, deps, folder, layout):
    import cv2
    if layout == "bullet":          # bullet_camera.py:71-85
        os.makedirs(os.path.join(folder, "rgb"))
        os.makedirs(os.path.join(folder, "depth"))
        for i, (c, d) in enumerate(zip(cols, deps)):
            cv2.imwrite(os.path.join(folder, "rgb", "rgb_%04d.png" % i), c)
            cv2.imwrite(os.path.join(folder, "depth", "depth_%04d.png" % i), d)
    else:                           # data_recorder.py:236-243
        os.makedirs(os.path.join(folder, "rgb"))
        os.makedirs(os.path.join(folder, "depth"))
        for i, (c, d) in enumerate(zip(cols, deps)):
            cv2.imwrite(os.path.join(folder, "rgb", "%04d.png" % i), c)
            np.save(os.path.join(folder, "depth", "%04d.npy" % i), d)
        json.dump({"collection_info": {"total_frames": len(cols)}, "camera_intrinsics": {"fx": 615.67},
                   "frames": [{"frame_id": i} for i in range(len(cols))]}, open(os.path.join(folder, "metadata.json"), "w"))


def test_read_frames_bullet_camera_layout(tmp_path):
    rng = np.random.default_rng(4)
    cols = [rng.integers(0, 256, (24, 32, 3), dtype=np.uint8) for _ in range(5)]
    deps8 = [rng.integers(0, 256, (24, 32), dtype=np.uint8) for _ in range(5)]       # the simulator saves uint8 depth
    folder = str(tmp_path / "camera")
    _write_frames(cols, deps8, folder, "bullet")
    c, d = io.read_frames_bullet_camera(folder)
    assert len(c) == len(d) == 4                     # 5 frames -> trailing one dropped (stitcher.py:228-231)
    for k in range(4):
        assert np.array_equal(c[k], cols[k]) and np.array_equal(d[k], deps8[k]) and d[k].dtype == np.uint8
    # uint16 millimetre PNGs survive IMREAD_UNCHANGED as well
    folder2 = str(tmp_path / "camera16")
    deps16 = [rng.integers(0, 3000, (24, 32)).astype(np.uint16) for _ in range(2)]
    _write_frames(cols[:2], deps16, folder2, "bullet")
    c, d = io.read_frames_bullet_camera(folder2)
    assert d[0].dtype == np.uint16 and np.array_equal(d[1], deps16[1])


def test_read_frames_realsense_layout(tmp_path):
    rng = np.random.default_rng(5)
    cols = [rng.integers(0, 256, (24, 32, 3), dtype=np.uint8) for _ in range(3)]
    deps = [rng.integers(0, 3000, (24, 32)).astype(np.uint16) for _ in range(3)]
    folder = str(tmp_path / "robot_data")
    _write_frames(cols, deps, folder, "realsense")
    c, d, meta = io.read_frames_realsense(folder)
    assert len(c) == len(d) == 2                     # count made even (stitcher.py:253-256)
    assert np.array_equal(c[0], cols[0]) and np.array_equal(d[1], deps[1]) and d[0].dtype == np.uint16
    assert meta["collection_info"]["total_frames"] == 3 and meta["camera_intrinsics"]["fx"] == 615.67


def test_read_trajectory_wtraj_formats(tmp_path):
    from scipy.spatial.transform import Rotation
    rng = np.random.default_rng(6)
    pos = rng.uniform(-1, 1, (7, 3))
    quat = Rotation.random(7, random_state=1).as_quat()             # scipy order: x, y, z, w
    t = np.linspace(0, 10, 7)
    f1 = str(tmp_path / "wtraj.txt")                                  # main.py:156-165: (t, [array(pos), array(quat)])
    with open(f1, "w") as f:
        for k in range(7):
            row = (t[k], [pos[k], quat[k]])
            f.write(",".join(map(str, row)) + "\n")
    times, poses = io.read_trajectory(f1)
    assert np.allclose(times, t) and poses.shape == (7, 4, 4)
    for k in range(7):
        assert np.allclose(poses[k][:3, :3], Rotation.from_quat(quat[k]).as_matrix(), atol=1e-7)
        assert np.allclose(poses[k][:3, 3], pos[k], atol=1e-7)
    f2 = str(tmp_path / "wtraj_input.txt")                            # main.py:138-140: [[x,y,z],[qx,qy,qz,qw]] rows
    with open(f2, "w") as f:
        for k in range(7):
            f.write(",".join(map(str, [list(pos[k]), list(quat[k])])) + "\n")
    times2, poses2 = io.read_trajectory(f2)
    assert times2 is None and np.allclose(poses2, poses, atol=1e-7)
    rel = io.relative_poses(poses)
    assert np.allclose(rel[0], np.eye(4)) and np.allclose(poses[0] @ rel[3], poses[3])
    # main.py:213-232: 12 pictures over 500 waypoints, first and last waypoint included
    idx = io.select_waypoints(500, 12)
    assert idx[0] == 0 and idx[-1] == 499 and len(idx) == 12 and all(b > a for a, b in zip(idx, idx[1:]))
    assert np.allclose(io.pose_from_position_quaternion([1, 2, 3], [0, 0, 0, 2.0]), [[1, 0, 0, 1], [0, 1, 0, 2], [0, 0, 1, 3], [0, 0, 0, 1]])


@pytest.mark.skipif(not os.path.exists("/root/reference/wtraj.txt"), reason="/root/reference is not present on this box")
def test_reads_the_reference_trajectories():
    times, poses = io.read_trajectory("/root/reference/wtraj.txt")
    assert poses.shape == (500, 4, 4) and abs(times[-1] - 10.0) < 1e-9
    assert np.allclose(poses[:, :3, :3] @ poses[:, :3, :3].transpose(0, 2, 1), np.eye(3), atol=1e-6)
    _, planned = io.read_trajectory("/root/reference/wtraj_input.txt")
    assert planned.shape == (500, 4, 4)
    # the executed end-effector path follows the planned semicircle (GRR's IK error: centimetres)
    assert np.abs(poses[:, :3, 3] - planned[:, :3, 3]).max() < 0.05
