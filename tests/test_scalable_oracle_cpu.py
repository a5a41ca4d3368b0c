"""The oracle's ScalableTSDFVolume restatement (oracle/scalable_oracle.cpp, SURVEY.md A.3 last bullet) checked against
the oracle's own UniformTSDFVolume and against properties Open3D's algorithm guarantees (CPU only)."""
import numpy as np

import helpers as hp

VL, TRUNC = 0.006, 0.024


def _frames(n=4):
    K = hp.small_intrinsic(320, 240)
    return hp.make_frames(n, K, radius=0.35, elevation=65.0), K


def test_unit_equals_uniform_volume_at_unit_origin(orc):
    """A unit IS a UniformTSDFVolume(unit_length, 16, trunc, origin = index * unit_length) integrated with the frames
    that touched it: with one frame, every opened unit must equal that uniform volume bit for bit."""
    frames, K = _frames(1)
    d16, rgb, ext = frames[0]
    df = orc.depth_to_float(d16)
    sv = orc.ScalableTSDFVolume(VL, TRUNC, 1, 16, 4)
    n = sv.integrate(df, rgb, hp.kt(K), ext)
    keys = sv.unit_keys()
    assert len(keys) == sv.touched and len(keys) > 30 and n > 0
    total = 0
    for k in keys[::5]:
        uv = orc.UniformTSDFVolume(VL * 16, 16, TRUNC, 1, k.astype(np.float64) * (VL * 16))
        total += uv.integrate(df, rgb, hp.kt(K), ext)
        t1, w1, c1 = sv.unit_arrays(k)
        assert np.array_equal(uv.tsdf(), t1) and np.array_equal(uv.weight(), w1) and np.array_equal(uv.color(), c1)
    assert total > 0


def test_touched_units_follow_the_depth_samples(orc):
    """Every opened unit lies within sdf_trunc (per axis) of a stride-4 back-projected depth sample, and every such
    sample's own unit is open."""
    frames, K = _frames(2)
    sv = orc.ScalableTSDFVolume(VL, TRUNC, 0, 16, 4)
    pts = []
    for d16, rgb, ext in frames:
        df = orc.depth_to_float(d16)
        sv.integrate(df, None, hp.kt(K), ext)
        sub = np.zeros_like(df)
        sub[::4, ::4] = df[::4, ::4]
        P, _ = orc.backproject(sub, None, hp.kt(K), ext)
        pts.append(P)
    P = np.concatenate(pts)
    UL = VL * 16
    keys = set(map(tuple, sv.unit_keys().tolist()))
    own = set(map(tuple, np.floor(P / UL).astype(np.int64).tolist()))
    assert own <= keys
    lo, hi = np.floor((P - TRUNC) / UL).astype(np.int64), np.floor((P + TRUNC) / UL).astype(np.int64)
    reach = set()
    for dx in (0, 1):
        for dy in (0, 1):
            for dz in (0, 1):
                sel = np.stack([np.where(dx, hi[:, 0], lo[:, 0]), np.where(dy, hi[:, 1], lo[:, 1]),
                                np.where(dz, hi[:, 2], lo[:, 2])], 1)
                reach |= set(map(tuple, sel.tolist()))
    assert keys == reach          # trunc < unit length: the range spans at most 2 units per axis


def test_scalable_agrees_with_uniform_near_the_surface(orc):
    """Where the scalable volume has integrated every frame the uniform one has (equal weights), the TSDF differs only
    by the z-chain's rounding (chain restarted per unit vs. marched through the cube): <= 1e-3."""
    frames, K = _frames(4)
    NU = 8
    off = np.array([-4, -4, -2])
    sv = orc.ScalableTSDFVolume(VL, TRUNC, 0, 16, 4)
    uv = orc.UniformTSDFVolume(VL * 16 * NU, 16 * NU, TRUNC, 0, off * VL * 16.0)
    for d16, rgb, ext in frames:
        df = orc.depth_to_float(d16)
        sv.integrate(df, None, hp.kt(K), ext)
        uv.integrate(df, None, hp.kt(K), ext)
    ts, w, _, outside = sv.dense_arrays(off, NU)
    assert outside == 0
    ut, uw = uv.tsdf(), uv.weight()
    assert (w <= uw).all()                       # a unit is integrated at most once per frame
    same = (w == uw) & (w > 0)
    assert same.sum() > 10000
    assert np.abs(ts - ut)[same].max() <= 1e-3
    # meshes: the scalable mesh is the uniform mesh restricted to opened units -- its vertices lie on uniform-mesh edges
    Vs, Ts, _ = sv.extract_triangle_mesh()
    Vu, Tu, _ = uv.extract_triangle_mesh()
    assert 0 < len(Vs) <= len(Vu) and len(Ts) <= len(Tu)
    from scipy.spatial import cKDTree
    d, _ = cKDTree(Vu).query(Vs)
    assert np.quantile(d, 0.99) < 0.05 * VL
    P, N, _ = sv.extract_point_cloud()
    assert len(P) > 1000 and np.abs(np.linalg.norm(N, axis=1) - 1).max() < 1e-12


def test_bounded_hash_is_the_unbounded_one_inside_the_cube(orc):
    """`set_bounds` (what bench.py uses to give the CPU arm the product's bounded backing cube): the units inside the
    cube are bit-identical to the unbounded hash's, units outside are not opened, and every touch outside is counted."""
    frames, K = _frames(2)
    a = orc.ScalableTSDFVolume(VL, TRUNC, 1, 16, 4)
    b = orc.ScalableTSDFVolume(VL, TRUNC, 1, 16, 4)
    for d16, rgb, ext in frames:
        a.integrate(orc.depth_to_float(d16), rgb, hp.kt(K), ext)
    ka = a.unit_keys()
    lo = ka.min(axis=0) + 1                      # a cube that cuts one layer of units off on the low side of each axis
    n = int((ka.max(axis=0) - lo).max()) + 1
    b.set_bounds(lo, n)
    for d16, rgb, ext in frames:
        b.integrate(orc.depth_to_float(d16), rgb, hp.kt(K), ext)
    kb = b.unit_keys()
    inside = np.all((ka >= lo) & (ka < lo + n), axis=1)
    assert 0 < inside.sum() < len(ka)
    assert {tuple(k) for k in kb} == {tuple(k) for k in ka[inside]}
    assert b.clipped() > 0 and a.clipped() == 0
    for k in kb[::7]:
        for x, y in zip(a.unit_arrays(k), b.unit_arrays(k)):
            assert np.array_equal(x, y)
