"""Multi-GPU parity worker, launched by tests/test_multi_gpu.py (or by hand) as
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tests/mgpu_worker.py
One rank per GPU over NCCL.  SlabVolume (slabs of x planes) vs the CPU oracle's single volume: every slab bit-exact,
the gathered mesh / point cloud identical to the oracle's, white-noise volume (all cube cases across the
seams), round-robin batched ICP identical to the single-rank batch."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (HERE, ROOT, os.path.join(ROOT, "3d-reconstruction-planning_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import b200recon as o3d  # noqa: E402
import helpers as hp  # noqa: E402
from b200recon import parallel, synthetic as syn  # noqa: E402
from oracle import oracle as orc  # noqa: E402


def check_mesh(ref_mesh, mesh, vl, color_tol):
    Vr, Tr, Cr = ref_mesh
    V, T, Cc = np.asarray(mesh.vertices), np.asarray(mesh.triangles), np.asarray(mesh.vertex_colors)
    assert len(V) == len(Vr) and len(T) == len(Tr), (len(V), len(Vr), len(T), len(Tr))
    Vs_r, Ts_r, ord_r = hp.canon_mesh(Vr, Tr)
    Vs, Ts, ord_g = hp.canon_mesh(V, T)
    assert np.abs(Vs - Vs_r).max() <= 1e-4 * vl
    assert np.array_equal(Ts, Ts_r)
    if Cr is not None and len(Cr):
        assert np.abs(Cc[ord_g] - Cr[ord_r]).max() <= color_tol


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if rank == 0:
        orc.build()
    dist.barrier()
    RGB8 = o3d.pipelines.integration.TSDFVolumeColorType.RGB8

    # ---- 1. fused frames: slabs bit-exact, gathered mesh / points == oracle --------------------
    R, L, trunc = 96, 0.384, 0.016
    K = hp.small_intrinsic(320, 240)
    frames = hp.make_frames(5, K)
    origin = syn.volume_origin(L)
    intr = o3d.camera.PinholeCameraIntrinsic(K["width"], K["height"], *hp.kt(K))
    # work-balanced (unequal) slabs from a coarse probe integration of the same frames -- what bench.py shards with
    from b200recon import runtime
    probe = []
    for d16, rgb, ext in frames:
        df = torch.from_numpy(orc.depth_to_float(d16)).to(runtime.ctx().tdev)
        probe.append((df, ext))
    work = parallel.probe_plane_work(L, R, trunc, origin, probe, intr, probe_resolution=32)
    ranges = parallel.balanced_slab_ranges(work, world)
    assert ranges[0][0] == 0 and ranges[-1][1] == R
    vol = parallel.SlabVolume(L, R, trunc, RGB8, origin, x_ranges=ranges)
    ref = orc.UniformTSDFVolume(L, R, trunc, 1, origin)
    for d16, rgb, ext in frames:
        vol.integrate_host(d16, rgb, intr, ext)
        ref.integrate(orc.depth_to_float(d16), rgb, hp.kt(K), ext)
    ts, w, _ = vol.local.export_arrays()
    zb, ze = vol.x_begin, vol.x_end
    assert np.array_equal(w, ref.weight()[zb:ze]), "slab weights differ"
    assert np.array_equal(ts.view(np.uint32), ref.tsdf()[zb:ze].view(np.uint32)), "slab tsdf differs"
    mesh = vol.extract_triangle_mesh()
    pcd = vol.extract_point_cloud()
    if rank == 0:
        check_mesh(ref.extract_triangle_mesh(), mesh, L / R, 2.0 / 255.0)
        P, N, Cc = ref.extract_point_cloud()
        Pg, Ng = np.asarray(pcd.points), np.asarray(pcd.normals)
        assert len(Pg) == len(P) and len(P) > 1000, (len(Pg), len(P))
        a, oa = hp.sort_rows(P)
        b, ob = hp.sort_rows(Pg)
        assert np.abs(a - b).max() <= 1e-4 * (L / R)
        assert np.abs(N[oa] - Ng[ob]).max() < 1e-9
    else:
        assert mesh is None and pcd is None

    # ---- 2. white-noise volume: every cube configuration across the seams --------------------------
    R2 = 8 * world + 3
    rng = np.random.default_rng(11)
    tsn = rng.uniform(-1, 1, (R2, R2, R2)).astype(np.float32)
    wn = (rng.random((R2, R2, R2)) > 0.08).astype(np.float32) * rng.integers(1, 5, (R2, R2, R2)).astype(np.float32)
    coln = (np.round(rng.uniform(0, 255, (R2, R2, R2, 3)) * 4) / 4).astype(np.float32)
    vol2 = parallel.SlabVolume(1.2, R2, 0.1, RGB8, (0.1, -0.2, 0.3))
    zb, ze = vol2.x_begin, vol2.x_end
    vol2.local.import_arrays(tsn[zb:ze], wn[zb:ze], coln[zb:ze])
    mesh2 = vol2.extract_triangle_mesh()
    pcd2 = vol2.extract_point_cloud()
    if rank == 0:
        ref2 = orc.UniformTSDFVolume(1.2, R2, 0.1, 1, (0.1, -0.2, 0.3))
        ref2.tsdf()[:] = tsn
        ref2.weight()[:] = wn
        ref2.color()[:] = coln
        check_mesh(ref2.extract_triangle_mesh(), mesh2, 1.2 / R2, 1e-9)
        P, N, _ = ref2.extract_point_cloud()
        Pg, Ng = np.asarray(pcd2.points), np.asarray(pcd2.normals)
        assert len(P) == len(Pg), (len(P), len(Pg))
        a, oa = hp.sort_rows(P)
        b, ob = hp.sort_rows(Pg)
        assert np.abs(a - b).max() < 1e-12 and np.abs(N[oa] - Ng[ob]).max() < 1e-9

    # ---- 3. batched ICP, pairs round-robin over the ranks ------------------------------------------
    rng = np.random.default_rng(5)
    srcs, tgts = [], []
    for p in range(2 * world + 1):
        pts = rng.uniform(-0.1, 0.1, (1500, 3))
        pts[:, 2] = 0.05 * np.sin(12 * pts[:, 0]) * np.cos(9 * pts[:, 1])
        nrm = np.stack([-0.6 * np.cos(12 * pts[:, 0]) * np.cos(9 * pts[:, 1]),
                        0.45 * np.sin(12 * pts[:, 0]) * np.sin(9 * pts[:, 1]), np.ones(len(pts))], 1)
        nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
        a = 0.01 * (p + 1)
        Rz = np.array([[np.cos(a), -np.sin(a), 0], [np.sin(a), np.cos(a), 0], [0, 0, 1]])
        s = o3d.geometry.PointCloud()
        s.points = o3d.utility.Vector3dVector((pts - [0.002, 0.001, 0.0]) @ Rz)
        t = o3d.geometry.PointCloud()
        t.points = o3d.utility.Vector3dVector(pts)
        t.normals = o3d.utility.Vector3dVector(nrm)
        srcs.append(s)
        tgts.append(t)
    est = o3d.pipelines.registration.TransformationEstimationPointToPlane()
    got = parallel.icp_round_robin(srcs, tgts, 0.02, estimation_method=est)
    assert len(got) == len(srcs) and all(g is not None for g in got)
    if rank == 0:
        want = o3d.pipelines.registration.registration_icp_batch(srcs, tgts, 0.02, estimation_method=est)
        for g, wv in zip(got, want):
            assert np.allclose(g.transformation, wv.transformation, atol=1e-12)
            assert abs(g.fitness - wv.fitness) < 1e-15 and abs(g.inlier_rmse - wv.inlier_rmse) < 1e-12
    dist.barrier()
    if rank == 0:
        print("MGPU OK world=%d" % world)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
