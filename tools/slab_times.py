"""Per-slab streaming-kernel time on ONE GPU for a given list of z boundaries (tuning aid for the slab balancer):
    python tools/slab_times.py 0,109,128,151,173,196,220,248,1024"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "3d-reconstruction-planning_b200")]
import torch
import b200recon as o3d
from b200recon import runtime, synthetic as syn
import bench

cuts = [int(v) for v in sys.argv[1].split(",")]
wl = sys.argv[2] if len(sys.argv) > 2 else "tsdf1024"
K = bench.camera(wl)
R, L, trunc, poses = bench.orbit(wl)
intr = o3d.camera.PinholeCameraIntrinsic(K["width"], K["height"], K["fx"], K["fy"], K["cx"], K["cy"])
ctx = runtime.ctx(0)
frames = []
for p in poses[:: max(1, len(poses) // 24)][:24]:
    d16, rgb = syn.render_frame(p, K)
    frames.append((d16.cpu().numpy(), rgb.cpu().numpy(), syn.extrinsic_from_pose(p)))
for a, b in zip(cuts[:-1], cuts[1:]):
    v = o3d.pipelines.integration.UniformTSDFVolume(L, R, trunc, o3d.pipelines.integration.TSDFVolumeColorType.RGB8,
                                                    syn.volume_origin(L), z_range=(a, b))
    v.count_updates = True
    for f in frames[:3]:
        v.integrate_host(f[0], f[1], intr, f[2])
    ctx.sync()
    u0 = v.updates()
    ctx.kernel_timing(True)
    for f in frames:
        v.integrate_host(f[0], f[1], intr, f[2])
    ctx.kernel_timing(False)
    ms, n = ctx.kernel_time()
    print("[%4d,%4d) kernel %.4f ms/launch  updates/launch %.2fM" % (a, b, ms / n, (v.updates() - u0) / n / 1e6))
    del v
    torch.cuda.empty_cache()
