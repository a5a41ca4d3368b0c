"""Shared helpers for the parity tests (inputs are produced by the ORACLE's CPU twin so both sides see
bit-identical frames)."""
import numpy as np

from b200recon import synthetic as syn
from oracle import oracle as orc

K640 = syn.D435_640x480
KT640 = (K640["fx"], K640["fy"], K640["cx"], K640["cy"])


def small_intrinsic(W=160, H=120):
    s = W / 640.0
    return dict(width=W, height=H, fx=K640["fx"] * s, fy=K640["fy"] * s, cx=K640["cx"] * s, cy=K640["cy"] * s)


def kt(K):
    return (K["fx"], K["fy"], K["cx"], K["cy"])


def make_frames(n, K=None, radius=0.5, elevation=30.0, **scene):
    """n synthetic frames (depth uint16 mm, colour uint8) + world->camera extrinsics."""
    K = K or K640
    poses = syn.orbit_poses(n, radius=radius, elevation_deg=elevation)
    frames = []
    for p in poses:
        d16, rgb = orc.raycast_scene(p, kt(K), K["width"], K["height"], **scene)
        frames.append((d16, rgb, syn.extrinsic_from_pose(p)))
    return frames


def canon_mesh(V, T, decimals=9):
    """Order-independent canonical form: vertices sorted lexicographically (after rounding only for
    the sort key), triangles remapped, rotated so the smallest id is first (orientation kept), sorted."""
    V = np.asarray(V, np.float64)
    T = np.asarray(T, np.int64)
    key = np.round(V, decimals)
    order = np.lexsort((key[:, 2], key[:, 1], key[:, 0]))
    inv = np.empty(len(V), np.int64)
    inv[order] = np.arange(len(V))
    Vs = V[order]
    Tm = inv[T] if len(T) else T
    if len(Tm):
        k = np.argmin(Tm, axis=1)
        Tm = np.stack([np.roll(r, -s) for r, s in zip(Tm, k)]) if len(Tm) < 20000 else _roll_rows(Tm, k)
        Tm = Tm[np.lexsort((Tm[:, 2], Tm[:, 1], Tm[:, 0]))]
    return Vs, Tm, order


def _roll_rows(T, k):
    idx = (np.arange(3)[None, :] + k[:, None]) % 3
    return np.take_along_axis(T, idx, axis=1)


def sort_rows(A, decimals=9):
    A = np.asarray(A, np.float64)
    key = np.round(A[:, :3], decimals)
    order = np.lexsort((key[:, 2], key[:, 1], key[:, 0]))
    return A[order], order


# ---- BASELINE configs[0]: the 12 pictures main.py takes of the banana (tests/golden/README.md) ----
BANANA_PLY = __import__("os").path.join(__import__("os").path.dirname(__import__("os").path.abspath(__file__)), "golden",
                                        "banana_poisson_nontextured.ply")


def banana_frames(n_images=12, W=320, H=240, checker=0.04):
    """(K, colors, depths, camera->world poses, triangles): frames of the banana mesh on the ground plane, ray-cast by
    the oracle's CPU twin from the simulator's camera path (main.py:68-136, 213-232)."""
    K = small_intrinsic(W, H)
    tris = syn.load_banana_triangles(BANANA_PLY)
    poses = syn.main_orbit_poses(n_images)
    cols, deps = [], []
    for p in poses:
        d16, rgb = orc.raycast_scene(p, kt(K), W, H, tris=tris, checker=checker)
        cols.append(rgb)
        deps.append(d16)
    return K, cols, deps, poses, tris


def perturb_poses(poses, rot_sigma=0.004, trans_sigma=0.003, seed=3):
    """The planned trajectory as a robot would report it: every pose but the first off by a small random SE(3)
    (sigma 0.23 degrees per axis, 3 mm per axis)."""
    from scipy.spatial.transform import Rotation
    rng = np.random.default_rng(seed)
    out = []
    for p in poses:
        q = np.array(p, np.float64)
        q[:3, :3] = Rotation.from_rotvec(rng.normal(0, rot_sigma, 3)).as_matrix() @ q[:3, :3]
        q[:3, 3] += rng.normal(0, trans_sigma, 3)
        out.append(q)
    out[0] = np.array(poses[0], np.float64)
    return out


def sample_triangles(tris, per_triangle=20, seed=0):
    """Points spread over a triangle soup (float [T,9]) -- a dense stand-in for the surface."""
    rng = np.random.default_rng(seed)
    a = np.asarray(tris, np.float64).reshape(-1, 3, 3)
    r1, r2 = rng.random((len(a), per_triangle, 1)), rng.random((len(a), per_triangle, 1))
    m = r1 + r2 > 1
    r1, r2 = np.where(m, 1 - r1, r1), np.where(m, 1 - r2, r2)
    return (a[:, None, 0] + r1 * (a[:, None, 1] - a[:, None, 0]) + r2 * (a[:, None, 2] - a[:, None, 0])).reshape(-1, 3)


def pose_error(T, P):
    """(translation error in metres, rotation error in degrees) between two 4x4 poses."""
    T, P = np.asarray(T), np.asarray(P)
    c = (np.trace(T[:3, :3].T @ P[:3, :3]) - 1.0) / 2.0
    return float(np.linalg.norm(T[:3, 3] - P[:3, 3])), float(np.degrees(np.arccos(np.clip(c, -1.0, 1.0))))
