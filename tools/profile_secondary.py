"""Runs the secondary kernels once each at bench sizes so that ncu can capture them:
    ncu --set full -k regex:'mc_classify|pc_count2|icp_iterate|cl_normals|cl_avgdist|backproject' ... python tools/profile_secondary.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tools"), os.path.join(ROOT, "3d-reconstruction-planning_b200")]
import numpy as np
import b200recon as o3d
from b200recon import synthetic as syn
import bench_extra

K = syn.D435_640x480
R = 1024
L = R * 0.002
intr = o3d.camera.PinholeCameraIntrinsic(K["width"], K["height"], K["fx"], K["fy"], K["cx"], K["cy"])
vol = o3d.pipelines.integration.UniformTSDFVolume(L, R, 0.01, o3d.pipelines.integration.TSDFVolumeColorType.RGB8, syn.volume_origin(L))
for p in syn.orbit_poses(8, radius=0.9):
    d16, rgb = syn.render_frame(p, K)
    vol.integrate_host(d16.cpu().numpy(), rgb.cpu().numpy(), intr, syn.extrinsic_from_pose(p))
vol.extract_triangle_mesh_device()
vol.extract_point_cloud_device()
clouds, _ = bench_extra.make_clouds(3, with_normals=True)
clouds[0].remove_statistical_outlier(20, 2.0)
clouds[0].voxel_down_sample(0.02)
print(bench_extra.icp_bench(8, 0.004))
