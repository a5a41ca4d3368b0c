"""Frame and geometry I/O either side of the path.

- PLY writer/reader for clouds and meshes (`merged_cloud.ply` is what the reference's authors saved,
  .MISSING_LARGE_BLOBS:1).
- frame readers for the two on-disk layouts the reference produces: `camera/rgb/rgb_%04d.png` +
  `camera/depth/depth_%04d.png` (bullet_camera.py:71-85) and `rgb/%04d.jpg` + `depth/%04d.npy` +
  `metadata.json` (data_recorder.py:236-243,305-315).
"""
import json
import os
import re

import numpy as np

from . import geometry


def write_point_cloud(filename, pointcloud, write_ascii=False):
    P = np.asarray(pointcloud.points)
    has_c, has_n = pointcloud.has_colors(), pointcloud.has_normals()
    props = ["property double x", "property double y", "property double z"]
    cols = [P]
    if has_n:
        props += ["property double nx", "property double ny", "property double nz"]
        cols.append(np.asarray(pointcloud.normals))
    dt = [("x", "<f8"), ("y", "<f8"), ("z", "<f8")] + ([("nx", "<f8"), ("ny", "<f8"), ("nz", "<f8")] if has_n else [])
    if has_c:
        props += ["property uchar red", "property uchar green", "property uchar blue"]
        dt += [("r", "u1"), ("g", "u1"), ("b", "u1")]
    hdr = "ply\nformat %s 1.0\nelement vertex %d\n%s\nend_header\n" % (
        "ascii" if write_ascii else "binary_little_endian", len(P), "\n".join(props))
    c8 = (np.clip(np.asarray(pointcloud.colors), 0, 1) * 255.0).astype(np.uint8) if has_c else None
    with open(filename, "wb") as f:
        f.write(hdr.encode())
        if write_ascii:
            for i in range(len(P)):
                row = " ".join(repr(float(v)) for c in cols for v in c[i])
                if has_c:
                    row += " %d %d %d" % tuple(c8[i])
                f.write((row + "\n").encode())
        else:
            rec = np.zeros(len(P), dtype=dt)
            rec["x"], rec["y"], rec["z"] = P[:, 0], P[:, 1], P[:, 2]
            if has_n:
                N = cols[1]
                rec["nx"], rec["ny"], rec["nz"] = N[:, 0], N[:, 1], N[:, 2]
            if has_c:
                rec["r"], rec["g"], rec["b"] = c8[:, 0], c8[:, 1], c8[:, 2]
            f.write(rec.tobytes())
    return True


def write_triangle_mesh(filename, mesh, write_ascii=False):
    V = np.asarray(mesh.vertices)
    T = np.asarray(mesh.triangles)
    has_c = mesh.has_vertex_colors()
    props = ["property double x", "property double y", "property double z"]
    dt = [("x", "<f8"), ("y", "<f8"), ("z", "<f8")]
    if has_c:
        props += ["property uchar red", "property uchar green", "property uchar blue"]
        dt += [("r", "u1"), ("g", "u1"), ("b", "u1")]
    hdr = ("ply\nformat binary_little_endian 1.0\nelement vertex %d\n%s\nelement face %d\n"
           "property list uchar uint vertex_indices\nend_header\n") % (len(V), "\n".join(props), len(T))
    rec = np.zeros(len(V), dtype=dt)
    if len(V):
        rec["x"], rec["y"], rec["z"] = V[:, 0], V[:, 1], V[:, 2]
        if has_c:
            c8 = (np.clip(np.asarray(mesh.vertex_colors), 0, 1) * 255.0).astype(np.uint8)
            rec["r"], rec["g"], rec["b"] = c8[:, 0], c8[:, 1], c8[:, 2]
    frec = np.zeros(len(T), dtype=[("n", "u1"), ("i", "<u4", (3,))])
    frec["n"] = 3
    if len(T):
        frec["i"] = T.astype(np.uint32)
    with open(filename, "wb") as f:
        f.write(hdr.encode())
        f.write(rec.tobytes())
        f.write(frec.tobytes())
    return True


_PLY_T = {"char": "i1", "uchar": "u1", "short": "<i2", "ushort": "<u2", "int": "<i4", "uint": "<u4",
          "float": "<f4", "double": "<f8", "int8": "i1", "uint8": "u1", "int16": "<i2", "uint16": "<u2",
          "int32": "<i4", "uint32": "<u4", "float32": "<f4", "float64": "<f8"}


def _read_ply(filename):
    with open(filename, "rb") as f:
        assert f.readline().strip() == b"ply"
        fmt = None
        elements = []
        while True:
            line = f.readline().decode().strip()
            if line == "end_header":
                break
            tok = line.split()
            if not tok:
                continue
            if tok[0] == "format":
                fmt = tok[1]
            elif tok[0] == "element":
                elements.append([tok[1], int(tok[2]), []])
            elif tok[0] == "property":
                elements[-1][2].append(tok[1:])
        out = {}
        for name, count, props in elements:
            if fmt == "ascii":
                rows = [f.readline().decode().split() for _ in range(count)]
                if any(p[0] == "list" for p in props):
                    out[name] = {"list": np.array([[int(v) for v in r[1:1 + int(r[0])]] for r in rows], np.int64)}
                else:
                    arr = np.array(rows, dtype=np.float64) if count else np.zeros((0, len(props)))
                    out[name] = {p[1]: arr[:, i] for i, p in enumerate(props)}
            else:
                assert fmt == "binary_little_endian", fmt
                if any(p[0] == "list" for p in props):
                    p = props[0]
                    ct, it = _PLY_T[p[1]], _PLY_T[p[2]]
                    # assume a constant list length (triangles), read the first count to learn it
                    pos = f.tell()
                    n0 = int(np.frombuffer(f.read(np.dtype(ct).itemsize), ct)[0]) if count else 3
                    f.seek(pos)
                    rec = np.dtype([("n", ct), ("i", it, (n0,))])
                    data = np.frombuffer(f.read(rec.itemsize * count), rec)
                    out[name] = {"list": data["i"].astype(np.int64)}
                else:
                    rec = np.dtype([(p[1], _PLY_T[p[0]]) for p in props])
                    data = np.frombuffer(f.read(rec.itemsize * count), rec)
                    out[name] = {p[1]: data[p[1]] for p in props}
        return out


def read_point_cloud(filename):
    d = _read_ply(filename)["vertex"]
    pcd = geometry.PointCloud()
    pcd.points = np.stack([d["x"], d["y"], d["z"]], 1).astype(np.float64)
    if "nx" in d:
        pcd.normals = np.stack([d["nx"], d["ny"], d["nz"]], 1).astype(np.float64)
    if "red" in d:
        pcd.colors = np.stack([d["red"], d["green"], d["blue"]], 1).astype(np.float64) / 255.0
    return pcd


def read_triangle_mesh(filename):
    d = _read_ply(filename)
    m = geometry.TriangleMesh()
    v = d["vertex"]
    m.vertices = np.stack([v["x"], v["y"], v["z"]], 1).astype(np.float64)
    if "face" in d:
        m.triangles = d["face"]["list"].astype(np.int32)
    if "red" in v:
        m.vertex_colors = np.stack([v["red"], v["green"], v["blue"]], 1).astype(np.float64) / 255.0
    return m


def read_frames_realsense(data_folder):
    """data_recorder.py layout: rgb/%04d.jpg (or .png/.npy), depth/%04d.npy (z16 mm), metadata.json.
    Mirrors stitcher.py:236-258 (count made even)."""
    rgb_dir, depth_dir = os.path.join(data_folder, "rgb"), os.path.join(data_folder, "depth")
    names = sorted(os.listdir(depth_dir))
    colors, depths = [], []
    for n in names:
        stem = os.path.splitext(n)[0]
        depths.append(np.load(os.path.join(depth_dir, n)))
        for ext in (".npy", ".jpg", ".png"):
            p = os.path.join(rgb_dir, stem + ext)
            if os.path.exists(p):
                if ext == ".npy":
                    colors.append(np.load(p))
                else:
                    import cv2
                    colors.append(cv2.imread(p))
                break
    if len(colors) % 2:
        colors, depths = colors[:-1], depths[:-1]
    meta = None
    mp = os.path.join(data_folder, "metadata.json")
    if os.path.exists(mp):
        meta = json.load(open(mp))
    return colors, depths, meta


def read_frames_bullet_camera(folder_path, rgb_foldername="rgb", depth_foldername="depth", modulus=2):
    """The simulator's on-disk layout (bullet_camera.py:71-85): `<folder>/rgb/rgb_%04d.png` (BGR as cv2 writes it) and
    `<folder>/depth/depth_%04d.png` (the GL depth buffer x 255 as **uint8** -- Open3D then reads it as v/255/1000 m,
    SURVEY A.1: garbage in, but defined behaviour).  Same result as RGBDStitcher.load_dataset_two_folders
    (stitcher.py:208-233): sorted file order, cv2.imread / IMREAD_UNCHANGED, trailing frames dropped so that the count
    is a multiple of `modulus` (optimization_modulus = 2, stitcher.py:19,228-231)."""
    import cv2
    colors, depths = [], []
    rgb_folder = os.path.join(folder_path, rgb_foldername)
    for fn in sorted(os.listdir(rgb_folder)):
        colors.append(cv2.imread(os.path.join(rgb_folder, fn)))
    depth_folder = os.path.join(folder_path, depth_foldername)
    for fn in sorted(os.listdir(depth_folder)):
        depths.append(cv2.imread(os.path.join(depth_folder, fn), cv2.IMREAD_UNCHANGED))
    if modulus > 1 and len(colors) % modulus:
        for _ in range(len(colors) % modulus):
            colors.pop()
            depths.pop()
    return colors, depths


def pose_from_position_quaternion(position, quat_xyzw):
    """4x4 pose from a position and an (x, y, z, w) quaternion -- scipy's Rotation.from_quat order, which is what
    Expansion-GRR/grr/utils.py:101-105 (matrix_to_quat) writes into wtraj.txt."""
    x, y, z, w = [float(v) for v in quat_xyzw]
    n = np.sqrt(x * x + y * y + z * z + w * w)
    if n == 0.0:
        raise ValueError("zero quaternion")
    x, y, z, w = x / n, y / n, z / n, w / n
    T = np.eye(4)
    T[:3, :3] = [[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                 [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                 [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]]
    T[:3, 3] = [float(v) for v in position]
    return T


_CALL = re.compile(r"[A-Za-z_][A-Za-z_0-9.]*\(")
_NUM = re.compile(r"[-+]?(?:\d+\.?\d*|\.\d+)(?:[eE][-+]?\d+)?")


def read_trajectory(filename):
    """Pose source of the simulator run: `wtraj.txt` (main.py:156-165) / `wtraj_input.txt` (main.py:138-140), one
    waypoint per line as Python reprs -- `t,[array([x, y, z]), array([qx, qy, qz, qw])]` resp.
    `[x, y, z],[qx, qy, qz, qw]` (no time stamp).  Returns (times float64 [N] or None, poses float64 [N,4,4]); the
    poses are the end-effector / camera-link frames in the robot's world, usable as ICP `init` after expressing them
    relative to the first frame (`relative_poses`)."""
    times, poses = [], []
    with open(filename) as f:
        for line in f:
            # drop the constructor names of the reprs ("array(", numpy 2's "np.float64(") before picking the numbers
            v = [float(x) for x in _NUM.findall(_CALL.sub("(", line))]
            if not v:
                continue
            if len(v) == 8:
                times.append(v[0])
                v = v[1:]
            if len(v) != 7:
                raise ValueError("read_trajectory: cannot parse %r" % line[:80])
            poses.append(pose_from_position_quaternion(v[:3], v[3:]))
    return (np.asarray(times) if len(times) == len(poses) and times else None), np.asarray(poses).reshape(-1, 4, 4)


def relative_poses(poses, reference=0):
    """poses[i] expressed in the frame of poses[reference]: inv(poses[reference]) @ poses[i]."""
    P = np.asarray(poses, np.float64).reshape(-1, 4, 4)
    inv0 = np.linalg.inv(P[reference])
    return np.stack([inv0 @ p for p in P])


def select_waypoints(n_waypoints, n_images):
    """Indices of the waypoints at which main.py:213-232 takes its `img_total` pictures: image k is taken as soon as
    1 + floor((img_total - 1) / maneuver_time * t) exceeds the count, i.e. at trajectory fraction k / (img_total - 1)."""
    if n_images <= 1:
        return [0] * max(n_images, 0)
    return [int(round(k * (n_waypoints - 1) / (n_images - 1))) for k in range(n_images)]
