"""Per-slab streaming-kernel time on ONE GPU for a given list of z boundaries (tuning aid for the slab balancer):
    python tools/slab_times.py 0,109,128,151,173,196,220,248,1024"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "3d-reconstruction-planning_b200")]
import torch
import b200recon as o3d
from b200recon import runtime, synthetic as syn
import bench

wl = sys.argv[2] if len(sys.argv) > 2 else "tsdf1024"
auto = sys.argv[1].startswith("auto:")
cuts = [] if auto else [int(v) for v in sys.argv[1].split(",")]
K = bench.camera(wl)
R, L, trunc, poses = bench.orbit(wl)
intr = o3d.camera.PinholeCameraIntrinsic(K["width"], K["height"], K["fx"], K["fy"], K["cx"], K["cy"])
ctx = runtime.ctx(0)
frames = []
for p in poses[:: max(1, len(poses) // 24)][:24]:
    d16, rgb = syn.render_frame(p, K)
    frames.append((d16.cpu().numpy(), rgb.cpu().numpy(), syn.extrinsic_from_pose(p)))
from b200recon import parallel, _lib


def slab_time(a, b):
    v = o3d.pipelines.integration.UniformTSDFVolume(L, R, trunc, o3d.pipelines.integration.TSDFVolumeColorType.RGB8,
                                                    syn.volume_origin(L), x_range=(a, b))
    for f in frames[:3]:
        v.integrate_host(f[0], f[1], intr, f[2])
    ctx.kernel_timing(True)
    for f in frames:
        v.integrate_host(f[0], f[1], intr, f[2])
    ctx.kernel_timing(False)
    ms, n = ctx.kernel_time()
    del v
    torch.cuda.empty_cache()
    return ms / n


if auto:      # emulate bench.py's setup for N ranks on one GPU: probe -> cost profile -> cut -> feedback steps
    n = int(sys.argv[1].split(":")[1])
    dev = ctx.tdev
    dfs = []
    for f in frames:
        d = torch.from_numpy(f[0].astype("float32") / 1000.0).to(dev)
        dfs.append((d, torch.from_numpy(f[1]).to(dev), f[2]))
    work = parallel.probe_plane_work(L, R, trunc, syn.volume_origin(L), [(d, e) for d, _, e in dfs[::2]], intr, 128)
    cost = parallel.profile_plane_cost(L, R, trunc, syn.volume_origin(L), dfs[::3], intr, work, chunk=32)
    print("cost profile per 32 planes (us):", [round(1000 * cost[a:a + 32].sum(), 1) for a in range(0, 384, 32)])
    ranges = parallel.balanced_slab_ranges(cost, n)
    for it in range(3):
        times = [slab_time(a, b) for a, b in ranges]
        print("it%d" % it, [a for a, _ in ranges] + [R], "times us", [round(1000 * t, 1) for t in times], "max", round(1000 * max(times), 1))
        ranges = parallel.rebalance_slab_ranges(ranges, cost, times)
    sys.exit(0)

for a, b in zip(cuts[:-1], cuts[1:]):
    v = o3d.pipelines.integration.UniformTSDFVolume(L, R, trunc, o3d.pipelines.integration.TSDFVolumeColorType.RGB8,
                                                    syn.volume_origin(L), x_range=(a, b))
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
