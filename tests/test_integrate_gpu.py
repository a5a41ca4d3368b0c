"""Parity of the CUDA TSDF integration (through the C ABI) with the CPU oracle.
Bar: TSDF |delta| <= 1e-5 and weights bit-exact -- the kernel replays Open3D's
float32 operation order, so the test demands bit-exact TSDF as well."""
import numpy as np
import pytest

import helpers as hp
from b200recon import synthetic as syn

pytestmark = pytest.mark.gpu


def _run_pair(o3d, orc, R, L, trunc, frames, K, color_type=1, x_ranges=None, origin=None, via_host=False):
    origin = syn.volume_origin(L) if origin is None else origin
    kt = hp.kt(K)
    ref = orc.UniformTSDFVolume(L, R, trunc, color_type, origin)
    intr = o3d.camera.PinholeCameraIntrinsic(K["width"], K["height"], *kt)
    ct = o3d.pipelines.integration.TSDFVolumeColorType(color_type)
    x_ranges = x_ranges or [(0, R)]
    vols = [o3d.pipelines.integration.UniformTSDFVolume(L, R, trunc, ct, origin, x_range=zr) for zr in x_ranges]
    for v in vols:
        v.count_updates = True
    n_ref = 0
    for d16, rgb, ext in frames:
        df = orc.depth_to_float(d16)
        col = None
        if color_type == 1:
            col = rgb
        elif color_type == 2:
            col = (rgb.astype(np.float32).mean(2) / 255.0).astype(np.float32)
        n_ref += ref.integrate(df, col, kt, ext)
        for v in vols:
            if via_host:
                v.integrate_host(d16, rgb if color_type == 1 else None, intr, ext)
            else:
                rgbd = o3d.geometry.RGBDImage.create_from_color_and_depth(
                    o3d.geometry.Image(col if col is not None else rgb), o3d.geometry.Image(d16),
                    depth_scale=1000.0, depth_trunc=3.0, convert_rgb_to_intensity=False)
                v.integrate(rgbd, intr, ext)
    return ref, vols, n_ref


def _compare(ref, vols, n_ref, color_type, color_tol=1.5):
    t_ref, w_ref, c_ref = ref.tsdf(), ref.weight(), ref.color()
    n_gpu = 0
    for v in vols:
        ts, w, col = v.export_arrays()
        zb, ze = v.x_begin, v.x_end
        assert np.array_equal(w, w_ref[zb:ze]), "weights must be bit-exact"
        assert np.array_equal(ts.view(np.uint32), t_ref[zb:ze].view(np.uint32)), "tsdf must be bit-exact"
        assert np.abs(ts - t_ref[zb:ze]).max() <= 1e-5
        if color_type == 1:
            m = w > 0
            err = np.abs(col - c_ref[zb:ze])[m]
            assert err.size == 0 or err.max() <= color_tol, err.max()  # packed 10:10:10 mean vs float64
        elif color_type == 2:
            m = w > 0
            assert not m.any() or np.abs(col[..., 0] - c_ref[zb:ze, :, :, 0])[m].max() <= 1e-5
        n_gpu += v.updates()
    assert n_gpu == n_ref
    assert n_ref > 0


@pytest.mark.parametrize("R,L,trunc", [(64, 0.384, 0.024), (96, 0.384, 0.016), (128, 0.512, 0.012)])
def test_integrate_bit_exact(o3d, orc, R, L, trunc):
    K = hp.small_intrinsic(320, 240)
    frames = hp.make_frames(6, K)
    ref, vols, n = _run_pair(o3d, orc, R, L, trunc, frames, K)
    _compare(ref, vols, n, 1)


def test_integrate_full_res_frames_host_path(o3d, orc):
    """640x480 D435 frames through the end-to-end HOST entry point (uint16 depth, H2D inside)."""
    frames = hp.make_frames(4)
    ref, vols, n = _run_pair(o3d, orc, 128, 0.512, 0.012, frames, hp.K640, via_host=True)
    _compare(ref, vols, n, 1)


def test_integrate_odd_resolution_scalar_path(o3d, orc):
    K = hp.small_intrinsic(160, 120)
    frames = hp.make_frames(3, K)
    ref, vols, n = _run_pair(o3d, orc, 67, 0.402, 0.02, frames, K)
    _compare(ref, vols, n, 1)


@pytest.mark.parametrize("color_type", [0, 2])
def test_integrate_color_types(o3d, orc, color_type):
    K = hp.small_intrinsic(160, 120)
    frames = hp.make_frames(3, K)
    ref, vols, n = _run_pair(o3d, orc, 64, 0.384, 0.024, frames, K, color_type=color_type)
    _compare(ref, vols, n, color_type)


def test_integrate_slabs_match_whole(o3d, orc):
    """slab sharding: slabs are ranges of x planes and own whole columns, so the union of the
    slabs is bit-identical to the single volume."""
    K = hp.small_intrinsic(320, 240)
    frames = hp.make_frames(4, K)
    ref, vols, n = _run_pair(o3d, orc, 96, 0.384, 0.016, frames, K, x_ranges=[(0, 24), (24, 48), (48, 72), (72, 96)])
    _compare(ref, vols, n, 1)
    ref, vols, n = _run_pair(o3d, orc, 96, 0.384, 0.016, frames, K, x_ranges=[(0, 37), (37, 96)])
    _compare(ref, vols, n, 1)


def test_integrate_camera_inside_and_behind(o3d, orc):
    """Camera inside the volume with half of the voxels behind it, tilted view, holes in the depth."""
    K = hp.small_intrinsic(160, 120)
    frames = hp.make_frames(5, K, radius=0.12, elevation=55.0)
    for d16, _, _ in frames:
        d16[::7, ::5] = 0          # invalid pixels
        d16[10:20, 30:60] = 40000  # beyond depth_trunc -> 0
    ref, vols, n = _run_pair(o3d, orc, 80, 0.64, 0.03, frames, K)
    _compare(ref, vols, n, 1)


def test_integrate_many_frames_color_drift(o3d, orc):
    K = hp.small_intrinsic(160, 120)
    frames = hp.make_frames(40, K)
    ref, vols, n = _run_pair(o3d, orc, 64, 0.384, 0.024, frames, K)
    _compare(ref, vols, n, 1, color_tol=3.0)


def test_integrate_256_frames_color_bound(o3d, orc, capsys):
    """The packed 10:10:10 running mean rounds to a quarter level at every update (and stops moving once
    |x - c| * 4 / (w + 1) < 1/2), Open3D keeps a float64 mean: the deviation after a 256-frame orbit -- the length of the
    bench sequence -- is bounded here.  TSDF and weights stay bit-exact over the same 256 frames."""
    K = hp.small_intrinsic(160, 120)
    frames = hp.make_frames(256, K)
    ref, vols, n = _run_pair(o3d, orc, 64, 0.384, 0.024, frames, K, via_host=True)
    _compare(ref, vols, n, 1, color_tol=16.0)
    _, w, col = vols[0].export_arrays()
    err = np.abs(col - ref.color())[w > 0]
    with capsys.disabled():
        print("\n[color drift] 256 frames: max |dc| = %.3f / 255, mean %.4f, p99.9 %.3f, max weight %d"
              % (err.max(), err.mean(), np.percentile(err, 99.9), int(w.max())))
    assert err.mean() < 1.5        # measured: max 10.7 / 255 (voxels on colour edges, whose input alternates), mean 0.98 / 255


def test_integrate_errors(o3d):
    intr = o3d.camera.PinholeCameraIntrinsic(160, 120, 150.0, 150.0, 80.0, 60.0)
    vol = o3d.pipelines.integration.UniformTSDFVolume(0.4, 32, 0.02, o3d.pipelines.integration.TSDFVolumeColorType.RGB8)
    d = np.zeros((100, 160), np.uint16)
    c = np.zeros((100, 160, 3), np.uint8)
    rgbd = o3d.geometry.RGBDImage.create_from_color_and_depth(o3d.geometry.Image(c), o3d.geometry.Image(d),
                                                              convert_rgb_to_intensity=False)
    with pytest.raises(RuntimeError):
        vol.integrate(rgbd, intr, np.eye(4))          # size mismatch with the intrinsic
    with pytest.raises(RuntimeError):
        o3d.geometry.RGBDImage.create_from_color_and_depth(o3d.geometry.Image(np.zeros((50, 160, 3), np.uint8)),
                                                           o3d.geometry.Image(d))
    # empty frame: nothing updated, nothing crashes
    d2 = np.zeros((120, 160), np.uint16)
    c2 = np.zeros((120, 160, 3), np.uint8)
    rgbd2 = o3d.geometry.RGBDImage.create_from_color_and_depth(o3d.geometry.Image(c2), o3d.geometry.Image(d2),
                                                               convert_rgb_to_intensity=False)
    vol.count_updates = True
    vol.integrate(rgbd2, intr, np.eye(4))
    assert vol.updates() == 0
    ts, w, _ = vol.export_arrays()
    assert not w.any() and not ts.any()
    vol.reset()


def test_integrate_rejects_frames_of_2_23_pixels_and_recovers(o3d, orc):
    """The streaming kernel's pixel index lives in a float32 mantissa next to 2^23 (exact below 2^24): frames of 2^23 pixels or
    more are rejected by the host with a clear error, and the context's frame slot is usable afterwards -- the next, normal
    frame integrates bit-exactly."""
    import torch
    from b200recon import runtime
    NoColor = o3d.pipelines.integration.TSDFVolumeColorType.NoColor
    vol = o3d.pipelines.integration.UniformTSDFVolume(0.384, 64, 0.024, NoColor, syn.volume_origin(0.384))
    big = o3d.camera.PinholeCameraIntrinsic(4096, 2048, 3000.0, 3000.0, 2048.0, 1024.0)       # exactly 2^23 pixels
    depth = torch.ones((2048, 4096), dtype=torch.float32, device=runtime.ctx(0).tdev)
    with pytest.raises(RuntimeError, match="2\\^23"):
        vol.integrate_device(depth, None, big, np.eye(4))
    del depth
    K = hp.small_intrinsic(320, 240)
    frames = hp.make_frames(2, K)
    ref, vols, n_ref = _run_pair(o3d, orc, 64, 0.384, 0.024, frames, K, color_type=0)
    _compare(ref, vols, n_ref, 0)
    # ... and on the volume that saw the error
    kt = hp.kt(K)
    intr = o3d.camera.PinholeCameraIntrinsic(K["width"], K["height"], *kt)
    for d16, rgb, ext in frames:
        vol.integrate_host(d16, None, intr, ext)
    ts, w, _ = vol.export_arrays()
    assert np.array_equal(w, ref.weight()) and np.array_equal(ts.view(np.uint32), ref.tsdf().view(np.uint32))


def test_shared_reciprocal_division_is_ieee(o3d):
    """The kernel's float32 divisions (projection, running mean) share a refined reciprocal; on the device they
    must equal the `/` operator bit for bit (2^28 pseudo-random cases incl. near-half-way quotients)."""
    import ctypes as C
    from b200recon import _lib, runtime
    c = runtime.ctx()
    for seed in (1, 12345):
        bad = (C.c_int64 * 4)(-1, -1, -1, -1)
        _lib.check(_lib.load().b2r_selftest_div(c.handle, 1 << 28, seed, bad, c.stream()))
        assert list(bad) == [0, 0, 0, 0], list(bad)


def test_chain_jump_equals_stepped_chain(o3d):
    """chain_jump (a measured-and-shelved alternative to walking the planes below a z-slab, csrc/tsdf_integrate.cu)
    reaches a slab's first plane by mantissa arithmetic instead of walking Open3D's float32 adds; on the
    device the result must equal the stepped chain bit for bit (2^22 cases per seed: camera-like magnitudes, zero
    crossings, exact ties, binade edges, denormals / huge steps, increments below half an ulp)."""
    import ctypes as C
    from b200recon import _lib, runtime
    c = runtime.ctx()
    for seed in (3, 987654321):
        bad = C.c_int64(-1)
        _lib.check(_lib.load().b2r_selftest_chain(c.handle, 1 << 22, seed, C.byref(bad), c.stream()))
        assert bad.value == 0, bad.value
