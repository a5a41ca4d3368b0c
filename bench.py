#!/usr/bin/env python3
"""bench.py -- headline benchmark of the reconstruction hot path (BASELINE.json metric:
"TSDF frames/s & voxel-updates/s (1-8 B200, %HBM peak); ICP pairs/s").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload tsdf1024|tsdf512] [--impl b200|reference]

A "step" is the integration of ONE synthetic 640x480 RGB-D frame of the orbit into the dense TSDF+colour
volume.  Default workload `tsdf1024` = BASELINE config[2] (1024^3 @ 2 mm, 256-frame orbit; the size the
north_star target is quoted on, and it fits one B200: 12.9 GB); `tsdf512` = config[1] (512^3, 64 frames).
For N > 1 (torchrun, one rank per GPU) the cube is sharded into slabs of x planes (x-slabs: whole columns per rank): every rank integrates every frame
into its own slab, no data-path collective -> total work fixed -> "scaling": "strong".

  value     frames/s with the frames already resident in HBM (ray-cast on the GPU before timing).
  e2e       frames/s through the public API with HOST (pinned) uint16 depth + uint8 colour frames:
            H2D copies + depth conversion inside the timed region, 8-byte D2H result per step.
  roofline  tsdf_integrate kernel alone: CUDA events on the launching stream around every launch;
            achieved = (voxel-updates * 24 B + frame bytes) / duration vs MEASURED_PEAKS.json hbm_gbs.
  cpu_baseline / --impl reference: the oracle (C++/OpenMP restatement of Open3D's integrate -- the
            reference's own CPU path, Open3D 0.18, cannot be installed here) on the box's host cores.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "3d-reconstruction-planning_b200")
for _p in (ROOT, PKG):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402

WORKLOADS = {
    # name: (resolution, length m, sdf_trunc m, orbit frames, orbit radius m)   SURVEY.md 8(d)
    "tsdf512": (512, 1.024, 0.01, 64, 0.5),
    "tsdf1024": (1024, 2.048, 0.01, 256, 0.9),
    "tsdf2048": (2048, 4.096, 0.01, 1000, 1.8),     # BASELINE config[4]: 1280x720 stream into 2048^3 (103 GB: 8 GPUs, or one)
}
CAMERAS = {"tsdf2048": "SYN_1280x720"}               # every other workload: the D435 640x480 intrinsics of stitcher.py:264-267


def camera(workload):
    from b200recon import synthetic as syn
    return getattr(syn, CAMERAS.get(workload, "D435_640x480"))
def ncu_traffic(workload):
    """dram__bytes_read.sum + dram__bytes_write.sum of ONE integrate_items launch, from the committed `ncu --set full`
    capture of this command (profiles/ncu_traffic.json: {workload: {bytes, profile, commit}}) -- per launch, like
    roofline.achieved.  It is a profiler figure, not measured in this run: the source is named next to it, and
    workloads without a capture report null."""
    p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    try:
        e = json.load(open(p)).get(workload)
        if e:
            return int(e["bytes"]), "%s (kernel source at commit %s)" % (e["profile"], e["commit"])
    except Exception:
        pass
    return None, None


B_UPD = 24  # algorithmic bytes per voxel update: float2 R+W (16) + packed colour R+W (8)  


def frame_bytes(K):
    n = K["width"] * K["height"]
    return n * (8 + 4)  # prepared {depth,multiplier} float2 + packed colour, read once per launch


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


_NVML_POLL = r"""
import sys, time
import pynvml as N
N.nvmlInit()
key = sys.argv[1]
try:
    h = N.nvmlDeviceGetHandleByUUID(key if key.startswith("GPU-") else "GPU-" + key) if len(key) > 4 else N.nvmlDeviceGetHandleByIndex(int(key))
except Exception:
    h = N.nvmlDeviceGetHandleByIndex(int(sys.argv[2]))
mx = N.nvmlDeviceGetMaxClockInfo(h, N.NVML_CLOCK_SM)
period = float(sys.argv[3]) * 1e-3
out = sys.stdout
while True:
    t = time.monotonic()
    sm = N.nvmlDeviceGetClockInfo(h, N.NVML_CLOCK_SM)
    r = N.nvmlDeviceGetCurrentClocksEventReasons(h)
    out.write("%.6f %d %d %d\n" % (t, sm, mx, r))
    out.flush()
    time.sleep(period)
"""


class ClockSampler:
    """SM clock and clock-event reasons sampled DURING the timed regions.  A helper process polls NVML every
    `period_ms` (nvidia-smi -lms cannot go below ~20 ms per row and takes longer to start than an 8-GPU timed region
    lasts); every row carries the system-wide monotonic time at which it was taken, and only rows that fall inside a
    timed window (begin()/end() around the event-bracketed loop) are reported."""

    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
               0x80: "hw_power_brake_slowdown"}

    def __init__(self, device, period_ms=1.0):
        self.device = device
        self.period_ms = period_ms
        self.rows = []          # (monotonic time, sm MHz, max MHz, reasons bitmask)
        self.windows = []       # [t0, t1]
        self.proc = None

    def start(self):
        key = str(self.device)
        try:
            import torch
            u = getattr(torch.cuda.get_device_properties(self.device), "uuid", None)
            if u is not None:
                key = str(u)
        except Exception:
            pass
        try:
            self.proc = subprocess.Popen([sys.executable, "-c", _NVML_POLL, key, str(self.device), str(self.period_ms)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None
            return
        self.t = threading.Thread(target=self._read, daemon=True)
        self.t.start()
        t_end = time.monotonic() + 10.0         # wait for the first row: the process is up and sampling
        while not self.rows and time.monotonic() < t_end and self.proc.poll() is None:
            time.sleep(0.01)

    def _read(self):
        for line in self.proc.stdout:
            f = line.split()
            if len(f) == 4:
                self.rows.append((float(f[0]), float(f[1]), float(f[2]), int(f[3])))

    def begin(self):
        self.windows.append([time.monotonic(), None])

    def end(self):
        self.windows[-1][1] = time.monotonic()

    def stop(self):
        if not self.proc or not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvml sampler unavailable"], "samples": 0}
        time.sleep(0.05)        # rows in flight in the pipe
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        ok = [r for r in self.rows if any(a <= r[0] <= (b or r[0]) for a, b in self.windows)]
        sm = [r[1] for r in ok]
        bits = 0
        for r in ok:
            bits |= r[3]
        reasons = sorted(n for b, n in self.REASONS.items() if bits & b)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(r[2] for r in self.rows),
                "reasons": reasons, "samples": len(sm), "timed_windows": len(self.windows),
                "period_ms": self.period_ms, "source": "NVML polled by a helper process"}


def orbit(workload):
    from b200recon import synthetic as syn
    R, L, trunc, nf, radius = WORKLOADS[workload]
    poses = syn.orbit_poses(nf, radius=radius, elevation_deg=30.0)
    return R, L, trunc, poses


# ------------------------------------------------------------------------------------------------
# CPU arm: the oracle on the host cores (cpu_baseline and --impl reference)
# ------------------------------------------------------------------------------------------------
def cpu_threads():
    """All host cores for the oracle's OpenMP loops, whatever OMP_NUM_THREADS the launcher exported (torchrun sets it
    to 1 for its workers); returns the team size actually in force -- the `cores` the JSON reports."""
    from oracle import oracle as orc
    n = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    orc.set_num_threads(n)
    return orc.max_threads()


def cpu_frames(workload, n, K):
    from b200recon import synthetic as syn
    from oracle import oracle as orc
    R, L, trunc, poses = orbit(workload)
    kt = (K["fx"], K["fy"], K["cx"], K["cy"])
    stride = max(1, len(poses) // n)
    out = []
    for p in poses[::stride][:n]:
        d16, rgb = orc.raycast_scene(p, kt, K["width"], K["height"])
        out.append((orc.depth_to_float(d16), rgb, syn.extrinsic_from_pose(p)))
    return out


def cpu_volumes(workload, n_ranges=8):
    """Oracle volumes over `n_ranges` x-ranges spread EVENLY over the cube (Open3D parallelises Integrate over x, so a
    set of x-ranges is the natural linear sample; a single centred range over-represents the dense middle of the
    cube -- the orbit looks at the cube centre).  Together they cover as many x planes as fit comfortably in host
    RAM; returns (volumes, scale = R / planes covered, ranges)."""
    from b200recon import synthetic as syn
    from oracle import oracle as orc
    R, L, trunc, _ = orbit(workload)
    try:
        avail = int([l for l in open("/proc/meminfo") if l.startswith("MemAvailable")][0].split()[1]) * 1024
    except Exception:
        avail = 8 << 30
    per_x = R * R * 32  # tsdf f32 + weight f32 + colour 3 x f64
    nx = int(min(R, max(n_ranges, (min(avail // 3, 6 << 30)) // per_x)))
    n_ranges = max(1, min(n_ranges, nx))
    w = max(1, nx // n_ranges)
    ranges = []
    for k in range(n_ranges):
        x0 = int((k + 0.5) * R / n_ranges - 0.5 * w)
        x0 = max(0, min(R - w, x0))
        ranges.append((x0, x0 + w))
    vols = [orc.UniformTSDFVolume(L, R, trunc, 1, syn.volume_origin(L), x_range=r) for r in ranges]
    return vols, R / float(w * n_ranges), ranges


def run_cpu(workload, steps, warmup, K, budget_s=25.0):
    """frames/s of the oracle; each step = one frame integrated into every sampled x-range, scaled linearly to R."""
    kt = (K["fx"], K["fy"], K["cx"], K["cy"])
    cpu_threads()
    vols, scale, xr = cpu_volumes(workload)
    frames = cpu_frames(workload, max(1, min(8, steps)), K)
    for i in range(max(1, warmup)):
        d, c, e = frames[i % len(frames)]
        for v in vols:
            v.integrate(d, c, kt, e)
    t0 = time.perf_counter()
    n = upd = 0
    while n < steps and (time.perf_counter() - t0) < budget_s:
        d, c, e = frames[n % len(frames)]
        for v in vols:
            upd += v.integrate(d, c, kt, e)
        n += 1
    dt = time.perf_counter() - t0
    R = WORKLOADS[workload][0]
    fps = n / (dt * scale)
    sample = "%d frames of %s, %d x-ranges of %d planes spread evenly over [0,%d) (first at x=%d; time scaled x%.2f to the full cube)" % (
        n, workload, len(xr), xr[0][1] - xr[0][0], R, xr[0][0], scale)
    return fps, upd * scale / dt, n, sample, dt


def run_cpu_icp(K, max_dist=0.004):
    """One pair of BASELINE config 4 on the oracle (exact k-d tree + OpenMP, Open3D's parallelisation): two
    consecutive orbit frames back-projected to ~300 k points each, analytic-free normals from the oracle's
    estimate_normals; returns pairs/s on the host cores."""
    from b200recon import synthetic as syn
    from oracle import oracle as orc
    kt = (K["fx"], K["fy"], K["cx"], K["cy"])
    cpu_threads()
    poses = syn.orbit_poses(64, radius=0.5, elevation_deg=30.0)[:2]
    clouds = []
    for p in poses:
        d16, rgb = orc.raycast_scene(p, kt, K["width"], K["height"])
        P, _ = orc.backproject(orc.depth_to_float(d16), rgb, kt, syn.extrinsic_from_pose(p))
        clouds.append(P)
    t0 = time.perf_counter()
    Nn = orc.estimate_normals(clouds[1], orc.SEARCH_HYBRID, 0.004, 30)
    t_n = time.perf_counter() - t0
    ca, sa = np.cos(0.004), np.sin(0.004)
    T0 = np.array([[ca, -sa, 0, 0.001], [sa, ca, 0, -0.0008], [0, 0, 1, 0.0005], [0, 0, 0, 1.0]])
    src = clouds[0] @ T0[:3, :3].T + T0[:3, 3]
    t0 = time.perf_counter()
    r = orc.registration_icp(src, clouds[1], Nn, max_dist, np.eye(4), 1, 1e-6, 1e-6, 30)
    dt = time.perf_counter() - t0
    return {"value": 1.0 / dt, "unit": "pairs/s", "cores": cpu_threads(), "kind": "port", "sample": "1 pair of config 4 "
            "(2 x %d points, point-to-plane, r = %g m)" % (len(src), max_dist), "iterations": r["iterations"],
            "estimate_normals_s": t_n}


def run_cpu_scalable(K, n_frames=4, resolution=1024):
    """ScalableTSDFVolume.integrate on the oracle (hash of 16^3 units, OpenMP over the frame's touched units):
    frames/s on the host cores for the configuration tools/bench_extra.scalable_bench measures on the GPU -- the SAME
    work: the oracle's hash is bounded to the product's backing cube (`set_bounds`), touches outside are counted on
    both sides instead of opened (Open3D's own hash is unbounded: 8156 instead of 3635 units per frame on this orbit)."""
    from b200recon import synthetic as syn
    from oracle import oracle as orc
    kt = (K["fx"], K["fy"], K["cx"], K["cy"])
    cpu_threads()
    poses = syn.orbit_poses(64, radius=0.9, elevation_deg=30.0)
    sv = orc.ScalableTSDFVolume(0.002, 0.01, 1, 16, 4)
    nu = resolution // 16
    origin = syn.volume_origin(resolution * 0.002)
    sv.set_bounds([int(np.floor(float(o) / (16 * 0.002) + 1e-9)) for o in origin], nu)
    frames = []
    for p in poses[:: max(1, len(poses) // n_frames)][:n_frames]:
        d16, rgb = orc.raycast_scene(p, kt, K["width"], K["height"])
        frames.append((orc.depth_to_float(d16), rgb, syn.extrinsic_from_pose(p)))
    sv.integrate(*frames[0][:2], kt, frames[0][2])          # warm-up (page faults of the first units)
    t0 = time.perf_counter()
    upd = 0
    for d, rgb, ext in frames:
        upd += sv.integrate(d, rgb, kt, ext)
    dt = time.perf_counter() - t0
    return {"value": len(frames) / dt, "unit": "frames/s", "cores": cpu_threads(), "kind": "port",
            "sample": "%d frames 640x480, 2 mm voxels, 16^3 units, stride 4, hash bounded to the GPU's %d^3 backing cube "
                      "(%d units touched in the last frame, %d touches outside the cube counted, not opened)"
                      % (len(frames), resolution, sv.touched, sv.clipped()), "voxel_updates_per_s": upd / dt,
            "units_touched_last_frame": int(sv.touched)}


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def run_gpu(args):
    import torch
    import torch.distributed as dist
    import b200recon as o3d
    from b200recon import runtime, synthetic as syn

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    K = camera(args.workload)
    kt = (K["fx"], K["fy"], K["cx"], K["cy"])
    intr = o3d.camera.PinholeCameraIntrinsic(K["width"], K["height"], *kt)
    R, L, trunc, poses = orbit(args.workload)
    ctx = runtime.ctx(local)
    dev = ctx.tdev

    # synthetic frames: ray-cast on this GPU (resident in HBM), raw uint16 depth + uint8 colour
    nf = min(len(poses), max(args.steps + args.warmup, 1))
    d16 = torch.empty((nf, K["height"], K["width"]), dtype=torch.uint16, device=dev)
    rgb = torch.empty((nf, K["height"], K["width"], 3), dtype=torch.uint8, device=dev)
    exts = []
    for i in range(nf):
        syn.render_frame(poses[i], K, out_depth=d16[i], out_color=rgb[i])
        exts.append(syn.extrinsic_from_pose(poses[i]))
    # device-resident arm input: float32 metres depth (what RGBDImage holds), converted outside the timed region
    depth_f = torch.empty((nf, K["height"], K["width"]), dtype=torch.float32, device=dev)
    from b200recon import _lib
    import ctypes as C
    for i in range(nf):
        _lib.check(_lib.load().b2r_depth_to_float(ctx.handle, runtime.ptr(d16[i]), 1, K["height"], K["width"], 1000.0,
                                                  3.0, runtime.ptr(depth_f[i]), ctx.stream()))
    # slab (x planes) owned by this rank.  Equal plane counts would leave most ranks idle (a table-top orbit only touches the
    # planes between the cameras and the table), so the slab boundaries come from a coarse probe integration of 16
    # frames of the sequence: contiguous slabs of equal WORK (parallel.balanced_slab_ranges); every rank computes the
    # same boundaries, no collective.
    from b200recon import parallel
    if world > 1 and not args.equal_slabs:
        probe = [(depth_f[k], exts[k]) for k in range(0, nf, max(1, nf // 16))]
        work = parallel.probe_plane_work(L, R, trunc, syn.volume_origin(L), probe, intr, probe_resolution=128, device=local)
        # measured cost per 32-plane slab (occluded planes under the table cost time without a single update)
        cprobe = [(depth_f[k], rgb[k], exts[k]) for k in range(0, nf, max(1, nf // 8))]
        work = parallel.profile_plane_cost(L, R, trunc, syn.volume_origin(L), cprobe, intr, work, chunk=32, device=local)
        # the profile is a TIMING, so it differs from rank to rank: every rank must cut the slabs from the same numbers
        # (mean over the ranks -- which also halves the noise), or the slabs would not tile the cube
        work = parallel.agree_plane_work(work, device=dev)
        ranges = parallel.balanced_slab_ranges(work, world)
        # three feedback steps on what the bench measures -- the time per frame of every rank on its slab (64 frames,
        # CUDA events around the loop; NOT the streaming kernel alone: for the first and last slab, which hold most of
        # the cube's planes, planning is a larger share of the frame than for the thin middle slabs): rescale the work
        # estimate of each slab to its measured time, cut again (setup, outside the timed region; one all_reduce of
        # world floats per step)
        for _ in range(3):
            zb, ze = ranges[rank]
            v = o3d.pipelines.integration.UniformTSDFVolume(L, R, trunc, o3d.pipelines.integration.TSDFVolumeColorType.RGB8,
                                                            syn.volume_origin(L), x_range=(zb, ze), device=local)
            v.count_updates = False
            for k in range(3):
                v.integrate_device(depth_f[k % nf], rgb[k % nf], intr, exts[k % nf])
            ctx.sync()
            f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            f0.record()
            ks = list(range(0, nf, max(1, nf // 64)))
            for k in ks:
                v.integrate_device(depth_f[k], rgb[k], intr, exts[k])
            ctx.join()
            f1.record()
            torch.cuda.synchronize()
            tt = torch.zeros(world, dtype=torch.float64, device=dev)
            tt[rank] = f0.elapsed_time(f1) / len(ks)
            dist.all_reduce(tt)
            ranges = parallel.rebalance_slab_ranges(ranges, work, tt.cpu().numpy())
            del v
            torch.cuda.empty_cache()
    else:
        ranges = parallel.slab_ranges(R, world)
    if world > 1:      # all ranks must hold the same tiling of [0, R)
        parallel.check_same_ranges(ranges, R, device=dev)
    zb, ze = ranges[rank]
    if args.emulate_rank:      # tuning aid: one rank's slab of an N-way split on a single GPU (not a bench line)
        r_, n_ = [int(v) for v in args.emulate_rank.split("/")]
        probe = [(depth_f[k], exts[k]) for k in range(0, nf, max(1, nf // 16))]
        work = parallel.probe_plane_work(L, R, trunc, syn.volume_origin(L), probe, intr, probe_resolution=128, device=local)
        ranges = parallel.balanced_slab_ranges(work, n_)
        zb, ze = ranges[r_]
    vol = o3d.pipelines.integration.UniformTSDFVolume(L, R, trunc, o3d.pipelines.integration.TSDFVolumeColorType.RGB8,
                                                      syn.volume_origin(L), x_range=(zb, ze), device=local)
    vol.count_updates = False   # switched on for the counting pass and the e2e pass only
    # pinned host copies for the e2e arm
    d16_h = d16.cpu().pin_memory()
    rgb_h = rgb.cpu().pin_memory()
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step_dev(i):
        k = i % nf
        vol.integrate_device(depth_f[k], rgb[k], intr, exts[k])

    def step_e2e(i):
        # pinned HOST frame -> H2D + conversion + plan + streaming kernel; the step's result (running update
        # count, 8 bytes) is written device -> pinned host by the streaming kernel's last warp
        k = i % nf
        vol.integrate_host(d16_h[k], rgb_h[k], intr, exts[k])

    # numpy (pageable) copies for the Open3D-named arm: what cv2.imread / np.load hand to stitcher.py
    d16_np = d16_h.numpy()
    rgb_np = rgb_h.numpy()
    G = o3d.geometry

    def step_o3d(i):
        # the call sequence of stitcher.py:34-42 followed by north_star's integrate(rgbd, intrinsic, extrinsic)
        k = i % nf
        rgbd = G.RGBDImage.create_from_color_and_depth(G.Image(rgb_np[k]), G.Image(d16_np[k]), depth_scale=1000.0,
                                                       depth_trunc=3.0, convert_rgb_to_intensity=False)
        vol.integrate(rgbd, intr, exts[k])

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()

    def timed(step, steps, warmup):
        vol.reset()
        for i in range(warmup):
            step(i)
        barrier()
        if rank == 0:
            sampler.begin()
        upd0 = vol.updates()
        l0 = ctx.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(warmup, warmup + steps):
            step(i)
        ctx.join()          # the streaming kernels run on the library's side stream: e1 goes behind the last one
        e1.record()
        barrier()
        if rank == 0:
            sampler.end()
        ms = e0.elapsed_time(e1)
        upd = vol.updates() - upd0
        timed.local_updates = upd          # this rank's own share (roofline of rank 0's kernel)
        launches = ctx.launch_count() - l0
        timed.rank_ms = [ms]
        if world > 1:
            tr_ = torch.zeros(world, dtype=torch.float64, device=dev)
            tr_[rank] = ms
            dist.all_reduce(tr_)
            timed.rank_ms = [float(x) for x in tr_.cpu()]
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
            u = torch.tensor([upd], dtype=torch.int64, device=dev)
            dist.all_reduce(u)
            upd = int(u.item())
        return ms, upd, launches

    ms, _, launches = timed(step_dev, args.steps, args.warmup)
    rank_ms_per_step = [round(x / args.steps, 5) for x in timed.rank_ms]
    # the end-to-end arms come right after the resident one: the boxes run power-capped and the SM clock sags over a long run
    # (1 965 -> ~1 840 MHz), so an arm timed four passes later is measured on a slower GPU than `value` was
    vol.count_updates = True        # the step's result (running update count) is part of the end-to-end definition
    e_ms, e_upd, _ = timed(step_e2e, args.steps, args.warmup)
    o_ms, o_upd, _ = timed(step_o3d, args.steps, args.warmup)
    vol.count_updates = False
    # dominant kernel alone: CUDA events (on the stream it is launched on) around every integrate_items launch
    # of the same steps, run again
    ctx.kernel_timing(True)
    timed(step_dev, args.steps, args.warmup)
    ctx.kernel_timing(False)
    k_ms, k_n = ctx.kernel_time()
    k_n_timed = args.steps
    # kernel_time covers warm-up launches too; time only the last `steps` launches' share
    k_ms_timed = k_ms * k_n_timed / max(k_n, 1)
    # voxel updates of exactly these steps: the same sequence replayed with the counting kernel variant (the
    # count is a deterministic function of the frame sequence since reset)
    vol.count_updates = True
    c_ms, upd, _ = timed(step_dev, args.steps, args.warmup)
    upd_k_local = timed.local_updates
    assert o_upd == e_upd == upd, "the three arms integrated different voxel sets: %r" % ((upd, e_upd, o_upd),)
    if os.environ.get("B2R_BENCH_DIAG"):        # where the e2e step spends its extra time (not part of the JSON contract)
        vol.count_updates = False
        n_ms, _, _ = timed(step_e2e, args.steps, args.warmup)
        vol.count_updates = True
        timed(step_e2e, 8, 0)
        sys.stderr.write("diag ms/step: resident %.4f  resident+count %.4f  e2e %.4f  e2e without count %.4f\n"
                         % (ms / args.steps, c_ms / args.steps, e_ms / args.steps, n_ms / args.steps))
    clocks = sampler.stop() if rank == 0 else None
    mirror = ctx.updates_mirror()
    assert mirror == vol.updates(), "host mirror of the step result disagrees with the device counter"

    # ---- the metric's other halves: mesh extraction of the volume just built (sharded: halo exchange -> per-slab
    #      marching cubes -> seam -> gather on rank 0) and ICP pairs/s (BASELINE config 4, both radii) ----
    extra = {}
    if not args.no_extras:
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import bench_extra
        if R <= 1024:
            try:
                extra["mesh"] = bench_extra.sharded_mesh_bench(vol, ranges, L / R)
                if world == 1:
                    extra["mesh"].update(bench_extra.mesh_bench(vol))
            except Exception as e:  # e.g. out of memory for the mesh of a very large volume: report, do not hide
                extra["mesh"] = {"error": repr(e)}
        vol.reset()
        for key, radius in (("icp", 0.004), ("icp_r0.05", 0.05)):
            icp = bench_extra.icp_bench(args.icp_pairs, radius, rank=rank, world=world)
            if world > 1:
                t = torch.tensor([icp["ms"]], dtype=torch.float64, device=dev)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                icp["ms"] = float(t.item())
            icp["pairs"] = args.icp_pairs
            icp["pairs_per_s"] = args.icp_pairs / (icp["ms"] * 1e-3)
            icp["distribution"] = "pair p -> GPU p %% %d, no collective" % world
            extra[key] = icp
        if world >= 8 and args.workload == "tsdf1024":      # BASELINE configs[4] on the same 8 GPUs (every rank calls it)
            del vol
            torch.cuda.empty_cache()
            try:
                extra["tsdf2048"] = bench_extra.tsdf2048_bench(rank, world, local)
            except Exception as e:
                extra["tsdf2048"] = {"error": repr(e)}
        if world == 1 and R <= 1024:
            try:
                extra["scalable"] = bench_extra.scalable_bench(1024, 64, 0.9)
            except Exception as e:
                extra["scalable"] = {"error": repr(e)}
        if world == 1:      # BASELINE configs[0]: the fused ICP -> integrate -> mesh path on the 12 banana frames
            try:
                extra["banana12"], run_gpu.banana_frames = bench_extra.banana12_bench()
            except Exception as e:
                extra["banana12"] = {"error": repr(e)}

    out = None
    if rank == 0:
        peak, peak_src = peaks()
        nz_frac = (ze - zb) / R
        upd_rank0_per_launch = upd_k_local / max(args.steps, 1)      # rank 0's own slab, its own kernel time
        alg_bytes = upd_rank0_per_launch * B_UPD + frame_bytes(K)
        achieved = alg_bytes / (k_ms_timed / max(args.steps, 1) * 1e-3) / 1e9
        traffic, traffic_src = ncu_traffic(args.workload) if world == 1 else (None, None)
        fps = args.steps / (ms * 1e-3)
        out = {
            "metric": "TSDF integration frames/s (%dx%d RGB-D into dense TSDF+colour volume)" % (K["width"], K["height"]),
            "value": fps, "unit": "frames/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "rank_ms_per_step": rank_ms_per_step,
            "config": {"workload": args.workload, "resolution": R, "voxel_mm": 1000.0 * L / R, "frames_in_orbit": len(poses),
                       "frame": "%dx%d uint16 depth + uint8 rgb" % (K["width"], K["height"]), "sharding": "x-slab x%d (slabs of x planes, whole columns per rank)" % world, "slab_boundaries": [a for a, _ in ranges] + [R],
                       "l2": "volume (%.1f GB) >> L2; consecutive frames touch overlapping voxels as in a real "
                             "sequence, no artificial flush" % (R ** 3 * 12 / 1e9)},
            "voxel_updates_per_s": upd / (ms * 1e-3),
            "voxel_visits_per_s": float(R) ** 3 * fps,
            "gpu_launches": launches,
            "clocks": clocks,
            "e2e": {"value": args.steps / (e_ms * 1e-3), "unit": "frames/s",
                    "h2d_bytes_per_step": int(d16_h[0].numel() * 2 + rgb_h[0].numel()), "d2h_bytes_per_step": 8,
                    "note": "integrate_host: host frames in pinned memory; H2D copies, conversion and planning of "
                            "frame i+1 overlap the streaming kernel of frame i (library side stream)",
                    "voxel_updates_per_s": e_upd / (e_ms * 1e-3)},
            "e2e_o3d": {"value": args.steps / (o_ms * 1e-3), "unit": "frames/s",
                        "h2d_bytes_per_step": int(d16_h[0].numel() * 2 + rgb_h[0].numel()), "d2h_bytes_per_step": 8,
                        "note": "the Open3D-named call surface of stitcher.py:34-42 + north_star: numpy (pageable) "
                                "uint16 depth + uint8 colour -> geometry.Image x2 -> RGBDImage.create_from_color_and_depth"
                                "(convert_rgb_to_intensity=False) -> vol.integrate(rgbd, intrinsic, extrinsic)"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "traffic_source": traffic_src,
                         "kernel": "integrate_tiles_kernel (rank 0's slab: plan + stream fused)" if (ze - zb) <= 96 else "integrate_items_kernel", "peak_source": peak_src,
                         "kernel_ms_per_launch": k_ms_timed / max(args.steps, 1),
                         "alg_bytes_per_launch": alg_bytes, "bytes_per_update": B_UPD,
                         "frac_of_nominal_8TBs": achieved / 8000.0,
                         # the same algorithmic bytes over the WHOLE step (prepare + plan + streaming kernel + hand-offs)
                         "step_achieved": alg_bytes / (ms / args.steps * 1e-3) / 1e9,
                         "step_frac": alg_bytes / (ms / args.steps * 1e-3) / 1e9 / peak,
                         "step_frac_of_nominal_8TBs": alg_bytes / (ms / args.steps * 1e-3) / 1e9 / 8000.0},
            "notes": {"voxel_updates": "counted in a separate pass of the same steps with the counting kernel variant "
                                       "(Open3D's integrate returns nothing); the timed `value` pass runs without the count",
                      "parity": "bit-exact against oracle/ (a restatement of Open3D 0.18; UNPINNED against the real "
                                "Open3D binary, which cannot be installed here)"},
        }
        out.update(extra)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return out


def main_banana12(args, rank):
    """BASELINE configs[0] as its own line: the fused path (ICP pose -> integrate -> mesh) over the 12 banana frames.
    A step = one frame through pipelines.fusion.fuse_sequence, host frames in, every copy inside (value == e2e: the
    path has no device-resident variant -- each registration blocks on its result like stitcher.py's loop).  One GPU:
    the sequence is strictly serial (frame i registers against the model of frames < i), ranks > 0 exit."""
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import bench_extra
    metric = "fused ICP -> TSDF integrate -> mesh frames/s (12 x 640x480 banana frames, BASELINE configs[0])"
    if args.impl == "reference":
        import torch  # the frames come from the GPU ray caster when there is a GPU, else from its CPU twin
        if torch.cuda.is_available():
            frames = bench_extra.banana12_frames()
        else:
            sys.path.insert(0, os.path.join(ROOT, "tests"))
            import helpers as hp
            frames = hp.banana_frames(12, 640, 480)
        cpu_threads()
        c = bench_extra.banana12_cpu(frames)
        emit({"impl": "reference", "metric": metric, "value": c["frames_per_s"], "unit": "frames/s", "n_gpus": args.gpus,
              "gpus_used": 0, "steps": 12, "warmup": 0, "ms_per_step": c["sequence_ms"] / 12, "higher_is_better": True,
              "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
              "config": {"workload": "banana12"}, "cpu_baseline": c,
              "e2e": {"value": c["frames_per_s"], "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})
        return
    import torch
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    sampler = ClockSampler(int(os.environ.get("LOCAL_RANK", "0")))
    sampler.start()
    sampler.begin()
    g, frames = bench_extra.banana12_bench(reps=max(1, min(args.steps, 5)))
    sampler.end()
    clocks = sampler.stop()
    K = frames[0]
    out = {"metric": metric, "value": g["frames_per_s"], "unit": "frames/s", "n_gpus": 1, "steps": 12, "warmup": 12,
           "ms_per_step": g["sequence_ms"] / 12, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
           "dtype": "f64", "data": "synthetic", "config": {"workload": "banana12", "detail": g["workload"]},
           "gpu_launches": g["gpu_launches_per_sequence"], "clocks": clocks,
           "e2e": {"value": g["frames_per_s"], "unit": "frames/s",
                   "h2d_bytes_per_step": int(K["width"] * K["height"] * 5), "d2h_bytes_per_step": 16 * 8 + 16,
                   "note": "wall clock over the whole sequence incl. mesh read-back; value == e2e"},
           "roofline": None, "quality": {k: v for k, v in g.items() if k not in ("workload", "sequence_ms", "frames_per_s")}}
    if not args.no_cpu_baseline:
        cpu_threads()
        out["cpu_baseline"] = bench_extra.banana12_cpu(frames)
        out["cpu_baseline"]["value"], out["cpu_baseline"]["unit"] = out["cpu_baseline"]["frames_per_s"], "frames/s"
    emit(out)


def emit(obj):
    """The one JSON line of the contract, written to the ORIGINAL stdout (libraries -- NCCL's version banner, torchrun
    -- print to fd 1 too; everything else of this process is sent to stderr, see main)."""
    os.write(_REAL_STDOUT, (json.dumps(obj) + "\n").encode())


_REAL_STDOUT = 1


def main():
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)           # fd 1 -> stderr for everything except emit()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=256)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="tsdf1024", choices=sorted(WORKLOADS) + ["banana12"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--equal-slabs", action="store_true", help="N > 1: equal plane counts instead of work-balanced slabs")
    ap.add_argument("--emulate-rank", default="", help="r/N: run rank r's slab of an N-way split on one GPU (tuning aid)")
    ap.add_argument("--no-extras", action="store_true", help="skip the ICP pairs/s and mesh-extraction measurements")
    ap.add_argument("--icp-pairs", type=int, default=64, help="BASELINE config 4: frame pairs of ~300 k points")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    if args.workload == "banana12":
        return main_banana12(args, rank)
    K = camera(args.workload)

    if args.impl == "reference":
        if rank != 0:
            return
        fps, ups, n, sample, dt = run_cpu(args.workload, args.steps, args.warmup, K, budget_s=120.0)
        cores = cpu_threads()
        R = WORKLOADS[args.workload][0]
        emit(({
            "impl": "reference", "metric": "TSDF integration frames/s (%dx%d RGB-D into dense TSDF+colour volume)" % (K["width"], K["height"]),
            "value": fps, "unit": "frames/s", "n_gpus": args.gpus, "gpus_used": 0, "steps": n, "warmup": args.warmup,
            "ms_per_step": 1000.0 / fps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": {"workload": args.workload, "resolution": R},
            "voxel_updates_per_s": ups,
            "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "note": "reference = Open3D 0.18 CPU wheel (not installable here); timed: the oracle's C++/OpenMP "
                    "restatement of UniformTSDFVolume::Integrate, parallelised over x like Open3D"}))
        return

    out = run_gpu(args)
    if out is not None:
        if not args.no_cpu_baseline and args.gpus == 1:
            fps, ups, n, sample, dt = run_cpu(args.workload, 8, 1, K, budget_s=25.0)
            out["cpu_baseline"] = {"value": fps, "unit": "frames/s", "cores": cpu_threads(), "kind": "port",
                                   "sample": sample, "voxel_updates_per_s": ups}
            if "icp" in out:
                out["icp"]["cpu_baseline"] = run_cpu_icp(K)
            if isinstance(out.get("scalable"), dict) and "error" not in out["scalable"]:
                out["scalable"]["cpu_baseline"] = run_cpu_scalable(K)
            if isinstance(out.get("banana12"), dict) and "error" not in out["banana12"]:
                sys.path.insert(0, os.path.join(ROOT, "tools"))
                import bench_extra
                out["banana12"]["cpu_baseline"] = bench_extra.banana12_cpu(run_gpu.banana_frames)
        emit(out)


if __name__ == "__main__":
    main()
