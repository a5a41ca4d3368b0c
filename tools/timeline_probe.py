"""Device-side timeline of consecutive integrate calls on ONE GPU (b2r_ctx_trace: the kernels stamp %globaltimer
themselves), for the whole cube or for the slabs (x planes) of an N-way split:
    python tools/timeline_probe.py 0,1024                 # the whole cube
    python tools/timeline_probe.py 0,109,131,153,176,199,224,259,1024 [tsdf1024] [host|device]
Prints, per slab, the median duration of frame_prepare / integrate_plan / integrate_items, the step (start of one
items kernel to the start of the next), the gap between consecutive items kernels, and the host time per call."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "3d-reconstruction-planning_b200")]
import numpy as np
import torch
import b200recon as o3d
from b200recon import runtime, synthetic as syn
import bench

cuts = [int(v) for v in sys.argv[1].split(",")]
wl = sys.argv[2] if len(sys.argv) > 2 else "tsdf1024"
mode = sys.argv[3] if len(sys.argv) > 3 else "host"
NF = int(os.environ.get("NF", "64"))
K = bench.camera(wl)
R, L, trunc, poses = bench.orbit(wl)
intr = o3d.camera.PinholeCameraIntrinsic(K["width"], K["height"], K["fx"], K["fy"], K["cx"], K["cy"])
ctx = runtime.ctx(0)
frames = []
for p in poses[:: max(1, len(poses) // NF)][:NF]:
    d16, rgb = syn.render_frame(p, K)
    if mode == "host":
        frames.append((d16.cpu().pin_memory(), rgb.cpu().pin_memory(), syn.extrinsic_from_pose(p)))
    else:
        frames.append(((d16.float() / 1000.0).contiguous(), rgb, syn.extrinsic_from_pose(p)))


def run(v):
    for f in frames:
        if mode == "host":
            v.integrate_host(f[0], f[1], intr, f[2])
        else:
            v.integrate_device(f[0], f[1], intr, f[2])


print("workload %s, %d frames, %s frames; times in us (median [p10 p90])" % (wl, len(frames), mode))
for a, b in zip(cuts[:-1], cuts[1:]):
    v = o3d.pipelines.integration.UniformTSDFVolume(L, R, trunc, o3d.pipelines.integration.TSDFVolumeColorType.RGB8,
                                                    syn.volume_origin(L), x_range=(a, b))
    run(v)
    ctx.sync()
    ctx.trace(len(frames))
    t0 = time.perf_counter()
    run(v)
    t_issue = time.perf_counter() - t0
    ctx.sync()
    t_all = time.perf_counter() - t0
    tr, n_live, n_items = ctx.trace_read()
    tr = tr.astype(np.float64) / 1000.0
    ctx.trace(0)
    dur = tr[:, :, 1] - tr[:, :, 0]
    st = tr[:, 2, 0]
    step = np.diff(st)
    gap = tr[1:, 2, 0] - tr[:-1, 2, 1]
    # how long before the items kernel of frame i could start was its plan finished?
    slack = tr[1:, 2, 0] - tr[1:, 1, 1]

    def q(x):
        return "%6.1f [%5.1f %5.1f]" % (np.median(x), np.percentile(x, 10), np.percentile(x, 90))
    print("[%4d,%4d) prepare %s  cull %s  plan %s  items %s  step %s  gap %s  plan_end->items_start %s  live tiles %d items %d  host issue %.1f us/call, wall %.1f us/call"
          % (a, b, q(dur[:, 0]), q(dur[:, 3]), q(dur[:, 1]), q(dur[:, 2]), q(step), q(gap), q(slack), np.median(n_live), np.median(n_items),
             1e6 * t_issue / len(frames), 1e6 * t_all / len(frames)))
    raw = ctx.trace_raw.astype(np.float64)
    nw = 148 * 5 * 4                      # resident streaming warps (5 blocks per SM, 4 warps per block)
    dur_i = (raw[:, 5] - raw[:, 4]) / 1000.0
    first_dry = (raw[:, 5] - raw[:, 9]) / 1000.0          # kernel end - first warp out of work
    mean_dry = dur_i - raw[:, 10] / nw / 1000.0           # mean over warps of (kernel end - own dry time)
    print("           streaming tail: first warp dry %s us before the end, mean warp idle %s us = %.1f %% of the kernel"
          % (q(first_dry), q(mean_dry), 100.0 * np.median(mean_dry / dur_i)))
    if os.environ.get("DUMP"):
        base = tr[0, 0, 0]
        for i in range(min(len(tr), int(os.environ["DUMP"]))):
            print("  frame %2d  prepare %7.1f-%7.1f  cull %7.1f-%7.1f  plan %7.1f-%7.1f  items %7.1f-%7.1f" % (
                i, tr[i, 0, 0] - base, tr[i, 0, 1] - base, tr[i, 3, 0] - base, tr[i, 3, 1] - base, tr[i, 1, 0] - base, tr[i, 1, 1] - base,
                tr[i, 2, 0] - base, tr[i, 2, 1] - base))
    del v
    torch.cuda.empty_cache()
