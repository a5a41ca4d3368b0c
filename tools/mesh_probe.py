import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tools"), os.path.join(ROOT, "3d-reconstruction-planning_b200")]
import b200recon as o3d
from b200recon import synthetic as syn
import bench_extra
R = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
nf = int(sys.argv[2]) if len(sys.argv) > 2 else 32
K = syn.D435_640x480
L = R * 0.002
intr = o3d.camera.PinholeCameraIntrinsic(K["width"], K["height"], K["fx"], K["fy"], K["cx"], K["cy"])
vol = o3d.pipelines.integration.UniformTSDFVolume(L, R, 0.01, o3d.pipelines.integration.TSDFVolumeColorType.RGB8, syn.volume_origin(L))
for p in syn.orbit_poses(nf, radius=0.5 if R <= 512 else 0.9):
    d16, rgb = syn.render_frame(p, K)
    vol.integrate_host(d16.cpu().numpy(), rgb.cpu().numpy(), intr, syn.extrinsic_from_pose(p))
print(bench_extra.mesh_bench(vol, reps=1))
