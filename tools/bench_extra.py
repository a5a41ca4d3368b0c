#!/usr/bin/env python3
"""Secondary measurements of the path: batched ICP pairs/s (BASELINE config 4), marching-cubes /
point extraction time, and the point-cloud preprocessing rows (f1/f2) at full-frame size.  CUDA-event timing.

    python tools/bench_extra.py [--pairs 64] [--res 512] [--frames 64]
Prints one JSON object.  bench.py imports `icp_bench` / `mesh_bench` for the extra keys of its JSON line."""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for _p in (ROOT, os.path.join(ROOT, "3d-reconstruction-planning_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402


def _ev():
    import torch
    return torch.cuda.Event(enable_timing=True)


def timed(fn, reps=1):
    import torch
    torch.cuda.synchronize()
    e0, e1 = _ev(), _ev()
    e0.record()
    out = None
    for _ in range(reps):
        out = None          # drop the previous result first: its buffers go back to torch's cache instead of a fresh cudaMalloc
        out = fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps, out


def make_clouds(n_frames, radius=0.5, K=None, with_normals=True, normal_radius=0.004):
    """Back-projected full-frame clouds (~300 k points at 640x480) of consecutive orbit frames, moved to the
    WORLD frame by the known pose so that consecutive clouds overlap; normals by estimate_normals."""
    import b200recon as o3d
    from b200recon import synthetic as syn
    K = K or syn.D435_640x480
    intr = o3d.camera.PinholeCameraIntrinsic(K["width"], K["height"], K["fx"], K["fy"], K["cx"], K["cy"])
    poses = syn.orbit_poses(64, radius=radius, elevation_deg=30.0)[:n_frames]
    clouds = []
    t_norm = 0.0
    for p in poses:
        d16, rgb = syn.render_frame(p, K)
        rgbd = o3d.geometry.RGBDImage.create_from_color_and_depth(o3d.geometry.Image(rgb.cpu().numpy()),
                                                                  o3d.geometry.Image(d16.cpu().numpy()),
                                                                  convert_rgb_to_intensity=False)
        pc = o3d.geometry.PointCloud.create_from_rgbd_image(rgbd, intr, syn.extrinsic_from_pose(p))
        if with_normals:
            ms, _ = timed(lambda: pc.estimate_normals(o3d.geometry.KDTreeSearchParamHybrid(normal_radius, 30)))
            t_norm += ms
        clouds.append(pc)
    return clouds, t_norm / max(1, len(clouds))


def icp_bench(n_pairs=64, max_dist=0.004, rank=0, world=1, reps=1):
    """BASELINE config 4: n_pairs frame pairs x ~300 k points, point-to-plane, criteria (1e-6, 1e-6, 30), pair p on
    GPU p % world.  The source of pair p is cloud p perturbed by a small known SE(3); the target is cloud p+1 of the orbit.
    Returns dict(pairs_per_s (this rank), ms, iterations, fitness...)."""
    import b200recon as o3d
    reg = o3d.pipelines.registration
    mine = [p for p in range(n_pairs) if p % world == rank]
    base, t_norm = make_clouds(min(9, len(mine) + 1))
    src, tgt, inits = [], [], []
    ca, sa = np.cos(0.004), np.sin(0.004)
    T0 = np.array([[ca, -sa, 0, 0.001], [sa, ca, 0, -0.0008], [0, 0, 1, 0.0005], [0, 0, 0, 1.0]])
    for i, p in enumerate(mine):
        a = base[i % (len(base) - 1)]
        b = base[i % (len(base) - 1) + 1]
        s = a.clone()
        s.transform(T0)
        src.append(s)
        tgt.append(b)
        inits.append(np.eye(4))
    est = reg.TransformationEstimationPointToPlane()
    crit = reg.ICPConvergenceCriteria(1e-6, 1e-6, 30)
    run = lambda: reg.registration_icp_batch(src, tgt, max_dist, inits, est, crit, with_correspondences=False)
    run()   # warm-up (scratch allocation)
    ms, res = timed(run, reps)
    npts = float(np.mean([len(s) for s in src]))
    return {"pairs": len(mine), "ms": ms, "pairs_per_s": len(mine) / (ms * 1e-3), "points_per_cloud": npts,
            "max_correspondence_distance": max_dist, "estimation": "point_to_plane",
            "iterations_mean": float(np.mean([r.iterations for r in res])),
            "fitness_mean": float(np.mean([r.fitness for r in res])),
            "rmse_mean": float(np.mean([r.inlier_rmse for r in res])),
            "estimate_normals_ms_per_cloud": t_norm}


def mesh_bench(vol, reps=2):
    """extract_triangle_mesh / extract_point_cloud of an integrated volume (device arrays, no host copy)."""
    vol.extract_triangle_mesh_device()      # warm-up calls: scratch and output allocation
    ms_m, m = timed(lambda: vol.extract_triangle_mesh_device(), reps)
    vol.extract_point_cloud_device()
    ms_p, p = timed(lambda: vol.extract_point_cloud_device(), reps)
    return {"mesh_ms": ms_m, "vertices": int(m[0].shape[0]), "triangles": int(m[1].shape[0]),
            "point_cloud_ms": ms_p, "points": int(p[0].shape[0])}


def sharded_mesh_bench(vol, ranges, voxel_length):
    """extract_triangle_mesh of the slab-sharded cube through parallel.SlabVolume (every rank calls this with its own
    slab `vol` = slab `ranges[rank]`): halo exchange of {tsdf,w}/colour planes -> per-slab marching cubes with global
    vertex ids -> seam plane of the id table -> gather on rank 0.  CUDA-event time per phase (max over ranks), the
    gathered counts and an order-independent digest of the mesh (parallel.mesh_checksum at 1e-4 voxel): the digest
    of the N-GPU mesh must equal the 1-GPU one.  With one rank the exchanges are no-ops."""
    import torch
    import torch.distributed as dist
    from b200recon import parallel
    rank, world = parallel.world_info()
    sv = parallel.SlabVolume(vol.length, vol.resolution, vol.sdf_trunc, vol.color_type, vol.origin, x_ranges=ranges,
                             local=vol)
    sv.extract_triangle_mesh()                       # warm-up: scratch allocation, NCCL channel set-up
    sv._halo_valid = False
    if world > 1:       # rank 0 has just copied the gathered mesh to the host: without this the other ranks' first
        dist.barrier()  # phase (the halo exchange) would be charged rank 0's ~40 ms of host work
    torch.cuda.synchronize(vol._ctx.tdev)
    sv.timings = {"on": True}
    t0 = time.perf_counter()
    mesh = sv.extract_triangle_mesh()
    wall = (time.perf_counter() - t0) * 1e3
    keys = ["halo_ms", "count_ms", "seam_ms", "emit_ms", "gather_ms"]
    tm = torch.tensor([sv.timings.get(k, 0.0) for k in keys] + [wall], dtype=torch.float64, device=vol._ctx.tdev)
    if world > 1:
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
    tm = tm.cpu().tolist()
    out = {k: v for k, v in zip(keys, tm[:-1])}
    out["sharded_mesh_ms"] = sum(tm[:-1])
    out["sharded_mesh_wall_ms"] = tm[-1]
    out["slabs"] = world
    out["per_slab_vertices"] = [int(v) for v in sv.last_counts[:, 0]]
    out["per_slab_triangles"] = [int(v) for v in sv.last_counts[:, 1]]
    out["exchange"] = ("none (single slab)" if world == 1 else
                       "NCCL send/recv: %d+%d halo planes of {tsdf,w}+colour per seam, 1 vertex-id plane per seam, "
                       "vertices/triangles to rank 0" % (2, 1))
    if rank == 0:
        V, T = np.asarray(mesh.vertices), np.asarray(mesh.triangles)
        out["checksum"] = parallel.mesh_checksum(V, T, 1e-4 * voxel_length)
        out["vertices"], out["triangles"] = int(len(V)), int(len(T))
    return out


def tsdf2048_bench(rank, world, device, steps=64, n_frames=32):
    """BASELINE configs[4] in the driver's own record (the driver only runs the default workload): 1280x720 frames into a
    2048^3 TSDF+colour cube (4.096 m @ 2 mm, 103 GB) sharded into x-slabs over the `world` ranks of this run.  Slab boundaries:
    the coarse work probe + two feedback steps on per-frame time, like the main workload.  Frames resident in HBM,
    CUDA-event time of `steps` integrate calls, max over ranks; every rank must call it."""
    import torch
    import torch.distributed as dist
    import b200recon as o3d
    from b200recon import parallel, runtime, synthetic as syn, _lib
    import bench
    wl = "tsdf2048"
    K = bench.camera(wl)
    R, L, trunc, poses = bench.orbit(wl)
    intr = o3d.camera.PinholeCameraIntrinsic(K["width"], K["height"], K["fx"], K["fy"], K["cx"], K["cy"])
    ctx = runtime.ctx(device)
    dev = ctx.tdev
    frames = []
    for p in poses[:: max(1, len(poses) // n_frames)][:n_frames]:
        d16, rgb = syn.render_frame(p, K)
        d = torch.empty((K["height"], K["width"]), dtype=torch.float32, device=dev)
        _lib.check(_lib.load().b2r_depth_to_float(ctx.handle, runtime.ptr(d16), 1, K["height"], K["width"], 1000.0, 3.0,
                                                  runtime.ptr(d), ctx.stream()))
        frames.append((d, rgb.contiguous(), syn.extrinsic_from_pose(p)))
    origin = syn.volume_origin(L)
    RGB8 = o3d.pipelines.integration.TSDFVolumeColorType.RGB8
    work = parallel.probe_plane_work(L, R, trunc, origin, [(d, e) for d, _, e in frames[::2]], intr, probe_resolution=128,
                                     device=device)
    ranges = parallel.balanced_slab_ranges(work, world)

    def per_frame_ms(v, n):
        for f in frames[:3]:
            v.integrate_device(f[0], f[1], intr, f[2])
        ctx.sync()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if world > 1:
            dist.barrier()
        e0.record()
        for i in range(n):
            f = frames[i % len(frames)]
            v.integrate_device(f[0], f[1], intr, f[2])
        ctx.join()
        e1.record()
        torch.cuda.synchronize(dev)
        return e0.elapsed_time(e1) / n

    for _ in range(2 if world > 1 else 0):
        v = o3d.pipelines.integration.UniformTSDFVolume(L, R, trunc, RGB8, origin, x_range=ranges[rank], device=device)
        tt = torch.zeros(world, dtype=torch.float64, device=dev)
        tt[rank] = per_frame_ms(v, len(frames))
        dist.all_reduce(tt)
        ranges = parallel.rebalance_slab_ranges(ranges, work, tt.cpu().numpy())
        del v
        torch.cuda.empty_cache()
    parallel.check_same_ranges(ranges, R, device=dev)
    v = o3d.pipelines.integration.UniformTSDFVolume(L, R, trunc, RGB8, origin, x_range=ranges[rank], device=device)
    ms = per_frame_ms(v, steps)
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    del v
    torch.cuda.empty_cache()
    return {"workload": "tsdf2048: 1280x720 frames into a 2048^3 TSDF+colour cube (4.096 m @ 2 mm, 103 GB), x-slab x%d" % world,
            "frames_per_s": 1000.0 / ms, "ms_per_step": ms, "steps": steps, "slab_boundaries": [a for a, _ in ranges] + [R],
            "frames": "%d resident frames of the 1000-frame orbit, cycled" % len(frames)}


BANANA_PLY = os.path.join(ROOT, "tests", "golden", "banana_poisson_nontextured.ply")


def banana12_frames(K=None, n_images=12):
    """BASELINE configs[0]: the 12 pictures main.py:213-232 takes on its semicircular camera path (main.py:68-136) of
    the YCB banana on the floor, ray-cast on the GPU from the committed mesh (the reference's own frames are not
    committed; its GL camera is unavailable on this box).  Host numpy frames, as stitcher.py's loaders return them."""
    from b200recon import synthetic as syn, runtime
    K = K or syn.D435_640x480
    tris = syn.load_banana_triangles(BANANA_PLY)
    tdev = runtime.to_device(tris)
    poses = syn.main_orbit_poses(n_images)
    cols, deps = [], []
    for p in poses:
        d16, rgb = syn.render_frame(p, K, tris_dev=tdev, tris=tris, checker=0.04)
        cols.append(rgb.cpu().numpy())
        deps.append(d16.cpu().numpy())
    return K, cols, deps, poses, tris


def _perturbed(poses, rot_sigma=0.004, trans_sigma=0.003, seed=3):
    """the planned trajectory as a robot reports it (wtraj.txt, main.py:156-165): off by a small random SE(3) per pose"""
    rng = np.random.default_rng(seed)
    out = []
    for p in poses:
        w = rng.normal(0, rot_sigma, 3)
        th = np.linalg.norm(w)
        Kx = np.array([[0, -w[2], w[1]], [w[2], 0, -w[0]], [-w[1], w[0], 0]])
        Rw = np.eye(3) + (np.sin(th) / th) * Kx + ((1 - np.cos(th)) / th ** 2) * (Kx @ Kx)
        q = np.array(p, np.float64)
        q[:3, :3] = Rw @ q[:3, :3]
        q[:3, 3] += rng.normal(0, trans_sigma, 3)
        out.append(q)
    out[0] = np.array(poses[0], np.float64)
    return out


BANANA_CFG = dict(resolution=256, voxel=0.002, trunc=0.01, reg_voxel=0.004, max_dist=0.01, estimation="colored",
                  target="model")


def _banana_quality(res, poses, tris, voxel):
    from scipy.spatial import cKDTree
    et, er = [], []
    for T, P in zip(res.poses, poses):
        et.append(float(np.linalg.norm(T[:3, 3] - P[:3, 3])))
        c = (np.trace(T[:3, :3].T @ P[:3, :3]) - 1.0) / 2.0
        er.append(float(np.degrees(np.arccos(np.clip(c, -1.0, 1.0)))))
    V = np.asarray(res.mesh.vertices)
    rng = np.random.default_rng(0)
    a = np.asarray(tris, np.float64).reshape(-1, 3, 3)
    r1, r2 = rng.random((len(a), 20, 1)), rng.random((len(a), 20, 1))
    m = r1 + r2 > 1
    r1, r2 = np.where(m, 1 - r1, r1), np.where(m, 1 - r2, r2)
    S = (a[:, None, 0] + r1 * (a[:, None, 1] - a[:, None, 0]) + r2 * (a[:, None, 2] - a[:, None, 0])).reshape(-1, 3)
    obj = V[V[:, 2] > 2 * voxel]
    d, _ = cKDTree(S).query(obj)
    return {"pose_err_mm_max": 1e3 * max(et), "pose_err_mm_last": 1e3 * et[-1], "pose_err_deg_max": max(er),
            "mesh_vertices": int(len(V)), "mesh_triangles": int(len(res.mesh.triangles)),
            "object_vertices": int(len(obj)), "surface_dist_mm_median": 1e3 * float(np.median(d)),
            "surface_dist_mm_p95": 1e3 * float(np.percentile(d, 95)), "voxel_mm": 1e3 * voxel}


def banana12_bench(reps=3, frames=None):
    """The fused path (pipelines.fusion.fuse_sequence: create_point_cloud_from_rgbd -> ICP -> integrate -> ... ->
    extract_triangle_mesh) on BASELINE configs[0], end to end from host frames, wall clock (every step blocks on its
    result, like stitcher.py's loop)."""
    import torch
    import b200recon as o3d
    from b200recon.pipelines import fusion
    K, cols, deps, poses, tris = frames or banana12_frames()
    cfg = BANANA_CFG
    R, vl = cfg["resolution"], cfg["voxel"]
    L = R * vl
    intr = o3d.camera.PinholeCameraIntrinsic(K["width"], K["height"], K["fx"], K["fy"], K["cx"], K["cy"])
    integ = o3d.pipelines.integration
    init = _perturbed(poses)
    vol = integ.UniformTSDFVolume(L, R, cfg["trunc"], integ.TSDFVolumeColorType.RGB8, (-L / 2, -L / 2, -0.1 * L))
    ctx = vol._ctx

    def run():
        vol.reset()
        return fusion.fuse_sequence(cols, deps, intr, vol, init_poses=init, max_correspondence_distance=cfg["max_dist"],
                                    registration_voxel_size=cfg["reg_voxel"], estimation=cfg["estimation"],
                                    target=cfg["target"])
    run()                                   # warm-up: scratch allocation
    torch.cuda.synchronize()
    l0 = ctx.launch_count()
    t0 = time.perf_counter()
    for _ in range(reps):
        res = run()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / reps
    out = {"workload": "banana12: %d frames %dx%d on the main.py camera path, fuse_sequence(%s ICP frame-to-%s, %g mm "
                       "registration clouds, r = %g m) into %d^3 @ %g mm + mesh" % (
                           len(cols), K["width"], K["height"], cfg["estimation"], cfg["target"], 1e3 * cfg["reg_voxel"],
                           cfg["max_dist"], R, 1e3 * vl),
           "sequence_ms": 1e3 * dt, "frames_per_s": len(cols) / dt, "gpu_launches_per_sequence": (ctx.launch_count() - l0) // reps,
           "iterations_mean": float(np.mean([r.iterations for r in res.registrations[1:]]))}
    out.update(_banana_quality(res, poses, tris, vl))
    return out, (K, cols, deps, poses, tris)


def banana12_cpu(frames):
    """The same sequence on the oracle port (tests/oracle_o3d.py wraps oracle/ in Open3D's names) on the host cores."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_o3d as A
    from b200recon.pipelines import fusion
    from oracle import oracle as orc
    K, cols, deps, poses, tris = frames
    cfg = BANANA_CFG
    R, vl = cfg["resolution"], cfg["voxel"]
    L = R * vl
    intr = A.camera.PinholeCameraIntrinsic(K["width"], K["height"], K["fx"], K["fy"], K["cx"], K["cy"])
    vol = A.pipelines.integration.UniformTSDFVolume(L, R, cfg["trunc"], 1, (-L / 2, -L / 2, -0.1 * L))
    t0 = time.perf_counter()
    res = fusion.fuse_sequence(cols, deps, intr, vol, init_poses=_perturbed(poses), max_correspondence_distance=cfg["max_dist"],
                               registration_voxel_size=cfg["reg_voxel"], estimation=cfg["estimation"], target=cfg["target"],
                               api=A)
    dt = time.perf_counter() - t0
    out = {"sequence_ms": 1e3 * dt, "frames_per_s": len(cols) / dt, "cores": orc.max_threads(), "kind": "port",
           "sample": "the whole 12-frame sequence once"}
    out.update(_banana_quality(res, poses, tris, vl))
    return out


def scalable_bench(resolution=1024, n_frames=64, radius=0.9, steps=None):
    """ScalableTSDFVolume.integrate (row f4): 640x480 frames of the bench orbit into 2 mm voxels, 16^3 units, stride 4,
    trunc 10 mm, frames resident in HBM.  Reports frames/s, touched units and updates per frame and the algorithmic
    bandwidth (24 B per voxel update) of the two kernels together."""
    import torch
    import b200recon as o3d
    from b200recon import synthetic as syn, _lib, runtime
    integ = o3d.pipelines.integration
    c = runtime.ctx()
    K = syn.D435_640x480
    intr = o3d.camera.PinholeCameraIntrinsic(K["width"], K["height"], K["fx"], K["fy"], K["cx"], K["cy"])
    vol = integ.ScalableTSDFVolume(0.002, 0.01, integ.TSDFVolumeColorType.RGB8, 16, 4, resolution=resolution,
                                   origin=syn.volume_origin(resolution * 0.002))
    frames = []
    for p in syn.orbit_poses(n_frames, radius=radius, elevation_deg=30.0):
        d16, rgb = syn.render_frame(p, K)
        d = torch.empty((K["height"], K["width"]), dtype=torch.float32, device=d16.device)
        _lib.check(_lib.load().b2r_depth_to_float(c.handle, runtime.ptr(d16), 1, K["height"], K["width"], 1000.0, 3.0,
                                                  runtime.ptr(d), c.stream()))
        frames.append((d, rgb.contiguous(), syn.extrinsic_from_pose(p)))
    steps = steps or 4 * n_frames
    for i in range(8):
        vol.integrate_device(*frames[i % n_frames][:2], intr, frames[i % n_frames][2])
    vol.reset()
    vol.count_updates = True

    def run():
        for i in range(steps):
            f = frames[i % n_frames]
            vol.integrate_device(f[0], f[1], intr, f[2])
    ms, _ = timed(run)
    upd = vol.updates()
    touched = vol.units_touched_last_frame()
    return {"frames_per_s": steps / (ms * 1e-3), "ms_per_frame": ms / steps, "steps": steps, "resolution": resolution,
            "voxel_mm": 2.0, "unit": 16, "stride": 4, "updates_per_frame": upd / steps,
            "units_touched_last_frame": touched, "units_opened": vol.units_opened(), "units_clipped": vol.units_clipped(),
            "alg_GBps": upd / steps * 24 / (ms / steps * 1e-3) / 1e9,
            "visited_voxels_per_frame": touched * 4096}


def preprocess_bench():
    import b200recon as o3d
    clouds, _ = make_clouds(2, with_normals=False)
    pc = clouds[0]
    out = {"points": len(pc)}
    out["voxel_down_sample_20mm_ms"], d = timed(lambda: pc.voxel_down_sample(0.02), 3)
    out["voxel_down_sample_2mm_ms"], d2 = timed(lambda: pc.voxel_down_sample(0.002), 3)
    out["cells_20mm"], out["cells_2mm"] = len(d), len(d2)
    out["estimate_normals_hybrid_4mm_30_ms"], _ = timed(lambda: pc.estimate_normals(o3d.geometry.KDTreeSearchParamHybrid(0.004, 30)), 2)
    out["remove_statistical_outlier_20_ms"], _ = timed(lambda: pc.remove_statistical_outlier(20, 2.0), 2)
    out["estimate_normals_down2mm_hybrid_40mm_30_ms"], _ = timed(lambda: d2.estimate_normals(o3d.geometry.KDTreeSearchParamHybrid(0.04, 30)), 2)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--pairs", type=int, default=64)
    ap.add_argument("--res", type=int, default=512)
    ap.add_argument("--frames", type=int, default=64)
    args = ap.parse_args()
    import b200recon as o3d
    from b200recon import synthetic as syn
    out = {}
    t0 = time.time()
    out["preprocess"] = preprocess_bench()
    out["icp_tight"] = icp_bench(args.pairs, 0.004)
    out["icp_wide"] = icp_bench(min(args.pairs, 8), 0.05)
    K = syn.D435_640x480
    L = args.res * 0.002
    intr = o3d.camera.PinholeCameraIntrinsic(K["width"], K["height"], K["fx"], K["fy"], K["cx"], K["cy"])
    vol = o3d.pipelines.integration.UniformTSDFVolume(L, args.res, 0.01, o3d.pipelines.integration.TSDFVolumeColorType.RGB8,
                                                      syn.volume_origin(L))
    import torch
    for p in syn.orbit_poses(args.frames, radius=0.5 if args.res <= 512 else 0.9):
        d16, rgb = syn.render_frame(p, K)
        vol.integrate_host(d16.cpu().numpy(), rgb.cpu().numpy(), intr, syn.extrinsic_from_pose(p))
    out["mesh_%d" % args.res] = mesh_bench(vol)
    del vol
    out["scalable"] = scalable_bench(1024 if args.res >= 1024 else 512, args.frames, 0.9 if args.res >= 1024 else 0.5)
    out["wall_s"] = time.time() - t0
    print(json.dumps(out))


if __name__ == "__main__":
    main()
