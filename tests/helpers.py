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
