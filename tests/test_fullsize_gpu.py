"""BASELINE.json's full sizes (640x480 frames into a dense 1024^3 TSDF+colour volume, config[2]) checked through
checks that do not need the whole cube on the CPU:

  * bit-exact parity with the ORACLE on x-ranges of the cube (Open3D parallelises Integrate over x, an x-range is
    a complete sub-problem: 8 x 1024 x 1024 voxels finish in seconds on the host);
  * slab consistency: three unequal slabs (x planes) integrated separately are bit-identical, plane for plane, to the single
    volume, and their update counts add up (chain replay across slabs);
  * the mesh of the full volume is welded and consistently oriented: every directed edge occurs exactly once, every
    index is in range, counts are stable across two extractions; surface points == sign-change edges of the same
    volume: point count equals the count of the extraction pass and every normal is a unit vector."""
import numpy as np
import pytest

import helpers as hp
from b200recon import synthetic as syn

pytestmark = pytest.mark.gpu

R, L, TRUNC = 1024, 2.048, 0.01


def _frames(n):
    poses = syn.orbit_poses(256, radius=0.9, elevation_deg=30.0)[:: 256 // n][:n]
    from oracle import oracle as orc
    out = []
    for p in poses:
        d16, rgb = orc.raycast_scene(p, hp.KT640, 640, 480)
        out.append((d16, rgb, syn.extrinsic_from_pose(p)))
    return out


def _volume(o3d, x_range=None):
    v = o3d.pipelines.integration.UniformTSDFVolume(L, R, TRUNC, o3d.pipelines.integration.TSDFVolumeColorType.RGB8,
                                                    syn.volume_origin(L), x_range=x_range)
    v.count_updates = True
    return v


def test_full_size_volume(o3d, orc):
    import torch
    frames = _frames(3)
    intr = o3d.camera.PinholeCameraIntrinsic(640, 480, *hp.KT640)
    whole = _volume(o3d)
    for d16, rgb, ext in frames:
        whole.integrate_host(d16, rgb, intr, ext)
    whole._ctx.sync()
    n_whole = whole.updates()
    assert n_whole > 100e6
    PL = 1   # B2R_PAD_LO: padded plane p holds x = x_begin - 1 + p; memory is [x plane][z][y]

    # ---- oracle parity on three x-ranges (left part of the frustum, centre, right part)
    for x0 in (300, 508, 700):
        ref = orc.UniformTSDFVolume(L, R, TRUNC, 1, syn.volume_origin(L), x_range=(x0, x0 + 8))
        n_ref = 0
        for d16, rgb, ext in frames:
            n_ref += ref.integrate(orc.depth_to_float(d16), rgb, hp.KT640, ext)
        g = whole._tsdfw[PL + x0:PL + x0 + 8].permute(0, 2, 1, 3).contiguous().cpu().numpy()   # [x, y, z, {tsdf, w}]
        assert np.array_equal(g[..., 1], ref.weight()), "weights must be bit-exact (x-range %d)" % x0
        assert np.array_equal(g[..., 0].view(np.uint32), ref.tsdf().view(np.uint32)), "tsdf must be bit-exact"
        assert int((g[..., 1] > 0).sum()) > 0 and n_ref > 0
        assert float(g[..., 1].sum()) == float(n_ref)       # weight = number of updates of a voxel
        del ref

    # ---- slab consistency
    n_slabs = 0
    for zr in ((0, 417), (417, 471), (471, 1024)):
        s = _volume(o3d, zr)
        for d16, rgb, ext in frames:
            s.integrate_host(d16, rgb, intr, ext)
        s._ctx.sync()
        n_slabs += s.updates()
        nz = zr[1] - zr[0]
        assert torch.equal(s._tsdfw[PL:PL + nz].view(torch.int32), whole._tsdfw[PL + zr[0]:PL + zr[1]].view(torch.int32))
        assert torch.equal(s._color[PL:PL + nz], whole._color[PL + zr[0]:PL + zr[1]])
        del s
        torch.cuda.empty_cache()
    assert n_slabs == n_whole

    # ---- mesh: welded, consistently oriented, reproducible
    V, T, Cc = whole.extract_triangle_mesh_device()
    nv, nt = int(V.shape[0]), int(T.shape[0])
    assert nv > 100000 and nt > 200000
    assert int(T.min()) >= 0 and int(T.max()) < nv
    Tl = T.to(torch.int64)
    de = torch.cat([Tl[:, 0] * nv + Tl[:, 1], Tl[:, 1] * nv + Tl[:, 2], Tl[:, 2] * nv + Tl[:, 0]])
    _, counts = torch.unique(de, return_counts=True)
    assert int(counts.max()) == 1, "a directed edge is used by two triangles: inconsistent orientation or unwelded vertices"
    assert int(torch.unique(Tl).numel()) == nv, "every vertex is referenced"
    V2, T2, _ = whole.extract_triangle_mesh_device()
    assert torch.equal(V, V2) and torch.equal(T, T2)
    lo = torch.tensor(syn.volume_origin(L), device=V.device)
    assert bool((V >= lo).all()) and bool((V <= lo + L).all())
    assert bool(((Cc >= 0) & (Cc <= 1.0)).all())
    # vertices lie on sign changes of a TSDF truncated at 5 voxels: within the truncation band of the analytic scene
    z = V[:, 2]
    e = ((V[:, 0] / 0.084) ** 2 + (V[:, 1] / 0.034) ** 2 + ((z - 0.02) / 0.020) ** 2).sqrt()
    near = (z.abs() < 2 * TRUNC) | ((e - 1).abs() * 0.020 < 2 * TRUNC)
    assert float(near.double().mean()) > 0.98

    # ---- surface points
    P, N, Pc = whole.extract_point_cloud_device()
    assert int(P.shape[0]) > 100000
    nn = torch.linalg.norm(N, dim=1)
    assert float((nn - 1).abs().max()) < 1e-12
    assert bool((P >= lo).all()) and bool((P <= lo + L).all())
