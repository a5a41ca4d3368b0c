"""Extraction timings on the bench volume (tsdf1024 orbit, NF frames): every repetition of extract_triangle_mesh_device /
extract_point_cloud_device timed on its own (CUDA events + wall), to separate first-call allocation from the steady state.
    python tools/extract_probe.py [workload] [NF]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tools"), os.path.join(ROOT, "3d-reconstruction-planning_b200")]
import torch
import b200recon as o3d
from b200recon import synthetic as syn
import bench

wl = sys.argv[1] if len(sys.argv) > 1 else "tsdf1024"
NF = int(sys.argv[2]) if len(sys.argv) > 2 else 64
K = bench.camera(wl)
R, L, trunc, poses = bench.orbit(wl)
intr = o3d.camera.PinholeCameraIntrinsic(K["width"], K["height"], K["fx"], K["fy"], K["cx"], K["cy"])
vol = o3d.pipelines.integration.UniformTSDFVolume(L, R, trunc, o3d.pipelines.integration.TSDFVolumeColorType.RGB8, syn.volume_origin(L))
for p in poses[:: max(1, len(poses) // NF)][:NF]:
    d16, rgb = syn.render_frame(p, K)
    vol.integrate_host(d16.cpu().numpy(), rgb.cpu().numpy(), intr, syn.extrinsic_from_pose(p))
torch.cuda.synchronize()
if vol._rows is not None:
    print("marked rows: %d of %d" % (int((vol._rows != 0).sum()), vol._rows.numel()))
for name, fn in (("mesh", vol.extract_triangle_mesh_device), ("points", vol.extract_point_cloud_device)):
    for i in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        e0.record()
        out = fn()
        e1.record()
        torch.cuda.synchronize()
        print("%s rep %d: %.3f ms (events) %.3f ms (wall), %d rows out" % (name, i, e0.elapsed_time(e1), 1e3 * (time.perf_counter() - t0), out[0].shape[0]))
        del out
