"""TEST INFRASTRUCTURE: the call sequence of the reference's RGBDStitcher (stitcher.py:21-166) written against
a module `o3d` passed in, so the same sequence runs on b200recon (GPU) and on tests/oracle_o3d.py (CPU oracle).
The reference file itself cannot be read at test time on the GPU box (/root/reference is absent there); the
drop-in of the real file is checked by tests/test_dropin_cpu.py in the build container.

Sequence per stitcher.py: create_point_cloud_from_rgbd (:34-46) for every frame; register_point_clouds (:88-112)
= voxel_down_sample(0.02) + estimate_normals(Hybrid(0.04, 30)) + compute_fpfh_feature (dead) on both clouds,
registration_colored_icp(0.05, init, ForColoredICP, Criteria(1e-6, 1e-6, 50)) when both have colours, then
registration_icp(0.05, init, PointToPoint); transform; +=; every 2nd frame voxel_down_sample + (if > 1000 points)
remove_statistical_outlier(20, 2.0) (:139-159)."""
import numpy as np


class Stitcher:
    def __init__(self, o3d, intrinsic, voxel_size=0.02, distance_threshold=0.05, optimization_modulus=2,
                 min_points_for_outlier_removal=1000):
        self.o3d, self.intrinsic = o3d, intrinsic
        self.voxel_size, self.distance_threshold = voxel_size, distance_threshold
        self.optimization_modulus = optimization_modulus
        self.min_points = min_points_for_outlier_removal
        self.log = []

    def cloud(self, color, depth):
        g = self.o3d.geometry
        rgbd = g.RGBDImage.create_from_color_and_depth(g.Image(color), g.Image(depth), depth_scale=1000.0, depth_trunc=3.0,
                                                       convert_rgb_to_intensity=False)
        return g.PointCloud.create_from_rgbd_image(rgbd, self.intrinsic)

    def preprocess(self, pcd):
        g, reg = self.o3d.geometry, self.o3d.pipelines.registration
        down = pcd.voxel_down_sample(self.voxel_size)
        down.estimate_normals(g.KDTreeSearchParamHybrid(radius=self.voxel_size * 2, max_nn=30))
        fpfh = reg.compute_fpfh_feature(down, g.KDTreeSearchParamHybrid(radius=self.voxel_size * 5, max_nn=100))
        return down, fpfh

    def register(self, source, target, init=None):
        reg = self.o3d.pipelines.registration
        init = np.identity(4) if init is None else init
        sd, _ = self.preprocess(source)
        td, _ = self.preprocess(target)
        if source.has_colors() and target.has_colors():
            rc = reg.registration_colored_icp(sd, td, self.distance_threshold, init, reg.TransformationEstimationForColoredICP(),
                                              reg.ICPConvergenceCriteria(relative_fitness=1e-6, relative_rmse=1e-6, max_iteration=50))
            init = rc.transformation
            self.log.append(("colored", np.array(rc.transformation), rc.fitness, rc.inlier_rmse, len(np.asarray(sd.points))))
        r = reg.registration_icp(sd, td, self.distance_threshold, init, reg.TransformationEstimationPointToPoint())
        self.log.append(("p2p", np.array(r.transformation), r.fitness, r.inlier_rmse, len(np.asarray(td.points))))
        return r.transformation, r.fitness

    def stitch(self, colors, depths):
        combined = self.cloud(colors[0], depths[0])
        for i in range(1, len(colors)):
            cur = self.cloud(colors[i], depths[i])
            T, _ = self.register(cur, combined)
            cur.transform(T)
            combined += cur
            if i % self.optimization_modulus == 0:
                combined = combined.voxel_down_sample(self.voxel_size)
                if len(np.asarray(combined.points)) > self.min_points:
                    combined, _ = combined.remove_statistical_outlier(nb_neighbors=20, std_ratio=2.0)
        return combined
