"""The fused path of SURVEY.md 8(a) row a4 / BASELINE configs[0]: what north_star's "drops in under
main.py/stitcher.py" adds over `RGBDStitcher.stitch_sequence` (stitcher.py:114-166) -- the pose that ICP finds for a
frame is used at once as the frame's extrinsic for TSDF integration, and the result is a mesh, not a concatenated
cloud:

    for each frame i:   cloud_i = create_point_cloud_from_rgbd(color_i, depth_i)            stitcher.py:21-48
                        T_i     = registration_icp(cloud_i -> cloud_{i-1} in the world,      stitcher.py:73-112, with
                                  init, TransformationEstimationPointToPlane).transformation north_star's estimator
                        volume.integrate(rgbd_i, intrinsic, extrinsic = inv(T_i))            Open3D UniformTSDFVolume
    mesh = volume.extract_triangle_mesh()

`fuse_sequence` is written ONLY against the Open3D-named API (`api.geometry.*`, `api.pipelines.*`): `api` defaults to
this package (every step a CUDA kernel behind the C ABI); any module exposing the same names works -- Open3D itself,
or the CPU checker the parity tests hand in -- so that one and the same sequence runs on both sides.

Differences from stitch_sequence, on purpose: the registration target is the previous FRAME's cloud (not the growing,
re-down-sampled union), every registration starts from the previous pose (or from the planned trajectory, e.g.
`io.read_trajectory('wtraj.txt')`, main.py:156-165) instead of the identity (stitcher.py:132), and FPFH / colored ICP /
outlier removal are not needed because the volume does the averaging.
"""
import numpy as np


class FusionResult:
    """poses: camera->world 4x4 per frame (poses[0] = the first frame's pose = the world frame unless given);
    registrations: RegistrationResult per frame (None for frame 0); volume: the integrated TSDF volume;
    mesh: `volume.extract_triangle_mesh()` (None if extract_mesh=False)."""

    def __init__(self):
        self.poses = []
        self.registrations = []
        self.volume = None
        self.mesh = None


def _api(api):
    if api is not None:
        return api
    import b200recon
    return b200recon


def register_frame(source_cloud, target_cloud, init, max_correspondence_distance, criteria=None, estimation="point_to_plane",
                   api=None):
    """One frame-to-frame registration of the fused path: the new frame's camera-frame cloud against the previous
    frame's cloud in the world (which carries normals).

    estimation = "point_to_plane": `registration_icp(..., TransformationEstimationPointToPlane)`, north_star's
                 estimator.  On a table-top scene (95 % of the pixels on one plane) three degrees of freedom -- sliding
                 in the plane, turning about its normal -- are held by the object's few points only and drift.
    estimation = "colored": `registration_colored_icp(..., TransformationEstimationForColoredICP)` as
                 register_point_clouds does first (stitcher.py:94-103): the photometric residual pins the sliding
                 directions on textured surfaces.  Needs colours on both clouds."""
    reg = _api(api).pipelines.registration
    if estimation == "colored":
        criteria = criteria or reg.ICPConvergenceCriteria(relative_fitness=1e-6, relative_rmse=1e-6, max_iteration=50)
        return reg.registration_colored_icp(source_cloud, target_cloud, max_correspondence_distance, init,
                                            reg.TransformationEstimationForColoredICP(), criteria)
    if estimation != "point_to_plane":
        raise ValueError("estimation must be 'point_to_plane' or 'colored'")
    criteria = criteria or reg.ICPConvergenceCriteria(relative_fitness=1e-6, relative_rmse=1e-6, max_iteration=30)
    return reg.registration_icp(source_cloud, target_cloud, max_correspondence_distance, init,
                                reg.TransformationEstimationPointToPlane(), criteria)


def fuse_sequence(color_images, depth_images, intrinsic, volume, init_poses=None, max_correspondence_distance=0.01,
                  normal_radius=0.01, normal_max_nn=30, registration_voxel_size=None, criteria=None, depth_scale=1000.0,
                  depth_trunc=3.0, extract_mesh=True, estimation="point_to_plane", target="previous", api=None,
                  on_frame=None):
    """Register-and-integrate a sequence of RGB-D frames.

    color_images / depth_images : lists of ndarrays as stitcher.py's loaders return them (uint8 [H,W,3]; uint16 mm /
                                  uint8 / float32 depth), same length.
    volume        : a `pipelines.integration.UniformTSDFVolume` (or ScalableTSDFVolume / parallel.SlabVolume) to
                    integrate into; its world frame is the frame of poses[0].
    init_poses    : optional camera->world poses from another source (the planned / FK trajectory, `wtraj.txt`).  Only
                    their RELATIVE motion is used: init_i = T_{i-1} . inv(P_{i-1}) . P_i, with T_0 = P_0.  Without
                    them init_i = T_{i-1} (constant-position model) and T_0 = identity.
    target        : "previous" -- register against the previous frame's cloud (frame-to-frame; pose errors add up along
                    the sequence), or "model" -- against `volume.extract_point_cloud()` of everything integrated so far
                    (frame-to-model: the surface points, normals and colours of the TSDF; errors do not chain).
    estimation    : "point_to_plane" (north_star) or "colored" (stitcher.py:94-103), see `register_frame`.
    registration_voxel_size : if set, ICP runs on `voxel_down_sample`d clouds (faster; integration still uses the
                    full frame).
    on_frame(i, rgbd, cloud, result, T) : optional observer, called after the frame has been integrated; `cloud` is
                    the next frame's registration target (world frame, with normals).
    Returns a FusionResult."""
    o3 = _api(api)
    g = o3.geometry
    if len(color_images) != len(depth_images):
        raise ValueError("Number of color and depth images must match")          # stitcher.py:118-119
    out = FusionResult()
    out.volume = volume
    if target not in ("previous", "model"):
        raise ValueError("target must be 'previous' or 'model'")
    target_cloud = None
    T_prev = None
    P = None if init_poses is None else np.asarray(init_poses, np.float64).reshape(-1, 4, 4)
    if P is not None and len(P) != len(color_images):
        raise ValueError("init_poses must hold one pose per frame")
    for i, (color, depth) in enumerate(zip(color_images, depth_images)):
        rgbd = g.RGBDImage.create_from_color_and_depth(g.Image(color), g.Image(depth), depth_scale=depth_scale,
                                                       depth_trunc=depth_trunc, convert_rgb_to_intensity=False)
        cloud = g.PointCloud.create_from_rgbd_image(rgbd, intrinsic)             # camera frame, stitcher.py:45-46
        work = cloud.voxel_down_sample(registration_voxel_size) if registration_voxel_size else cloud
        result = None
        if i == 0:
            T = np.eye(4) if P is None else P[0].copy()
        else:
            init = T_prev if P is None else T_prev @ np.linalg.inv(P[i - 1]) @ P[i]
            result = register_frame(work, target_cloud, init, max_correspondence_distance, criteria, estimation, api=o3)
            T = np.asarray(result.transformation, np.float64).copy()
        volume.integrate(rgbd, intrinsic, np.linalg.inv(T))                      # extrinsic = world -> camera
        if target == "model":
            # the next target is the model itself: surface points of the TSDF with its gradient normals and colours
            if i + 1 < len(color_images):
                target_cloud = volume.extract_point_cloud()
        else:
            # this frame becomes the next target: normals in the camera frame, then moved to the world with the pose
            work.estimate_normals(g.KDTreeSearchParamHybrid(radius=normal_radius, max_nn=normal_max_nn))
            work.transform(T)
            target_cloud = work
        out.poses.append(T)
        out.registrations.append(result)
        T_prev = T
        if on_frame is not None:
            on_frame(i, rgbd, target_cloud, result, T)
    if extract_mesh:
        out.mesh = volume.extract_triangle_mesh()
    return out
