"""ScalableTSDFVolume (SURVEY.md section 8 row f4, A.3 last bullet) -- `b2r_integrate_units` + the dense extraction
kernels against the oracle's hash-of-units restatement of Open3D 0.18 (oracle/scalable_oracle.cpp).

Bars: set of opened units identical; TSDF and weights of every unit bit-exact (float32, Open3D's operation order with
the z-chain restarted per unit); colour within 1.5/255 (packed 10:10:10 running mean vs float64, as for the uniform
volume); mesh: identical counts and triangle sets, vertices within 1e-4 voxel (the oracle positions come from global
voxel indices, ours from cube-local indices + origin: last-bit differences); point cloud: same, normals 1e-9."""
import numpy as np
import pytest

import helpers as hp

pytestmark = pytest.mark.gpu

VL, TRUNC, RES = 0.006, 0.024, 256


def _pair(o3d, orc, frames, K, color_type=1, stride=4, unit=16, resolution=RES):
    integ = o3d.pipelines.integration
    ref = orc.ScalableTSDFVolume(VL, TRUNC, color_type, unit, stride)
    vol = integ.ScalableTSDFVolume(VL, TRUNC, integ.TSDFVolumeColorType(color_type), unit, stride, resolution=resolution)
    vol.count_updates = True
    intr = o3d.camera.PinholeCameraIntrinsic(K["width"], K["height"], *hp.kt(K))
    n_ref = 0
    touched = []
    for d16, rgb, ext in frames:
        df = orc.depth_to_float(d16)
        col = rgb if color_type == 1 else None
        n_ref += ref.integrate(df, col, hp.kt(K), ext)
        rgbd = o3d.geometry.RGBDImage.create_from_color_and_depth(o3d.geometry.Image(rgb), o3d.geometry.Image(d16),
                                                                  convert_rgb_to_intensity=False)
        vol.integrate(rgbd, intr, ext)
        touched.append((ref.touched, vol.units_touched_last_frame()))
    return ref, vol, n_ref, touched


def _frames(n=6):
    K = hp.small_intrinsic(320, 240)
    return hp.make_frames(n, K, radius=0.35, elevation=65.0), K


@pytest.mark.parametrize("color_type", [1, 0])
def test_scalable_units_match_oracle(o3d, orc, color_type):
    frames, K = _frames()
    ref, vol, n_ref, touched = _pair(o3d, orc, frames, K, color_type)
    assert vol.units_clipped() == 0            # the scene fits the backing cube: nothing Open3D would have opened is lost
    for a, b in touched:
        assert a == b and a > 50
    assert vol.updates() == n_ref
    keys_r = ref.unit_keys()
    keys_g = vol.unit_keys()
    assert vol.units_opened() == len(keys_r)
    assert sorted(map(tuple, keys_r.tolist())) == sorted(map(tuple, keys_g.tolist()))
    ts, w, col = vol.export_arrays()
    seen = np.zeros(ts.shape, bool)
    r = 16
    for k in keys_r:
        a, b, c = ((k - vol.unit_offset) * r).tolist()
        t1, w1, c1 = ref.unit_arrays(k)
        assert np.array_equal(w[a:a + r, b:b + r, c:c + r], w1)
        assert np.array_equal(ts[a:a + r, b:b + r, c:c + r], t1)
        if color_type == 1:
            m = w1 > 0
            assert not m.any() or np.abs(col[a:a + r, b:b + r, c:c + r] - c1)[m].max() <= 1.5
        seen[a:a + r, b:b + r, c:c + r] = True
    assert not w[~seen].any() and not ts[~seen].any()      # never-opened units stay zero


def test_scalable_mesh_and_points_match_oracle(o3d, orc):
    frames, K = _frames()
    ref, vol, _, _ = _pair(o3d, orc, frames, K, 1)
    Vr, Tr, Cr = ref.extract_triangle_mesh()
    mesh = vol.extract_triangle_mesh()
    V, T, Cc = np.asarray(mesh.vertices), np.asarray(mesh.triangles), np.asarray(mesh.vertex_colors)
    assert len(Vr) > 2000 and len(V) == len(Vr) and len(T) == len(Tr)
    Vs_r, Ts_r, ord_r = hp.canon_mesh(Vr, Tr)
    Vs, Ts, ord_g = hp.canon_mesh(V, T)
    assert np.abs(Vs - Vs_r).max() <= 1e-4 * VL
    assert np.array_equal(Ts, Ts_r)
    assert np.abs(Cc[ord_g] - Cr[ord_r]).max() <= 2.0 / 255.0
    P, N, Cp = ref.extract_point_cloud()
    pcd = vol.extract_point_cloud()
    Pg, Ng, Cg = np.asarray(pcd.points), np.asarray(pcd.normals), np.asarray(pcd.colors)
    assert len(P) > 2000 and len(Pg) == len(P)
    # match by nearest neighbour (a rounded sort key is not stable under the last-bit position differences)
    from scipy.spatial import cKDTree
    d, j = cKDTree(Pg).query(P)
    assert d.max() <= 1e-4 * VL and len(np.unique(j)) == len(P)
    assert np.abs(N - Ng[j]).max() < 1e-9
    assert np.abs(Cp - Cg[j]).max() <= 2.0 / 255.0


@pytest.mark.parametrize("stride,unit", [(1, 16), (8, 8), (3, 4)])
def test_scalable_stride_and_unit_size(o3d, orc, stride, unit):
    frames, K = _frames(3)
    ref, vol, n_ref, touched = _pair(o3d, orc, frames, K, 0, stride=stride, unit=unit)
    assert vol.units_clipped() == 0
    assert [a for a, _ in touched] == [b for _, b in touched]
    assert vol.updates() == n_ref
    ts, w, _ = vol.export_arrays()
    for k in ref.unit_keys()[::7]:
        a, b, c = ((k - vol.unit_offset) * unit).tolist()
        t1, w1, _ = ref.unit_arrays(k)
        assert np.array_equal(w[a:a + unit, b:b + unit, c:c + unit], w1)
        assert np.array_equal(ts[a:a + unit, b:b + unit, c:c + unit], t1)


def test_scalable_clipping_is_counted_and_reset_clears(o3d, orc):
    """A cube too small for the scene: touches outside are counted, the units inside still match the oracle."""
    frames, K = _frames(2)
    ref, vol, _, _ = _pair(o3d, orc, frames, K, 0, resolution=64)
    keys_r = ref.unit_keys()
    inside = np.all((keys_r >= vol.unit_offset) & (keys_r < vol.unit_offset + 4), axis=1)
    assert vol.units_clipped() > 0 and (~inside).any()
    assert vol.units_opened() == int(inside.sum())
    ts, w, _ = vol.export_arrays()
    for k in keys_r[inside]:
        a, b, c = ((k - vol.unit_offset) * 16).tolist()
        t1, w1, _ = ref.unit_arrays(k)
        assert np.array_equal(w[a:a + 16, b:b + 16, c:c + 16], w1) and np.array_equal(ts[a:a + 16, b:b + 16, c:c + 16], t1)
    vol.reset()
    assert vol.units_opened() == 0 and vol.units_clipped() == 0 and len(vol.unit_keys()) == 0
    ts, w, _ = vol.export_arrays()
    assert not w.any() and not ts.any()


def test_scalable_errors(o3d):
    integ = o3d.pipelines.integration
    with pytest.raises(RuntimeError):
        integ.ScalableTSDFVolume(0.004, 0.02, integ.TSDFVolumeColorType.RGB8, 32)           # unit > 16
    with pytest.raises(RuntimeError):
        integ.ScalableTSDFVolume(0.004, 0.02, integ.TSDFVolumeColorType.RGB8, 16, resolution=100)
    vol = integ.ScalableTSDFVolume(0.004, 0.02, integ.TSDFVolumeColorType.RGB8, resolution=64)
    intr = o3d.camera.PinholeCameraIntrinsic(160, 120, 150.0, 150.0, 80.0, 60.0)
    rgbd = o3d.geometry.RGBDImage.create_from_color_and_depth(o3d.geometry.Image(np.zeros((50, 160, 3), np.uint8)),
                                                              o3d.geometry.Image(np.zeros((50, 160), np.uint16)))
    with pytest.raises(RuntimeError):
        vol.integrate(rgbd, intr, np.eye(4))                                                  # size mismatch
    e = vol.extract_triangle_mesh()
    assert len(e.vertices) == 0
