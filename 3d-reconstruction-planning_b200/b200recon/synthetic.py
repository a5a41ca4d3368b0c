"""Synthetic RGB-D sequences: camera orbits and scene constants.

Replaces the reference's frame sources, which need a GL renderer or camera hardware
("unavailable on this box"): bullet_camera.py:48-85 (PyBullet GL camera) and
data_recorder.py:104-121 (RealSense D435).  The orbit mirrors main.py:68-136 (camera circling
the object, looking at it).  Host-side float64 pose math only; pixels are produced by the CUDA
ray caster (b2r_raycast_scene).
"""
import numpy as np

# D435 colour intrinsics used everywhere in the reference (stitcher.py:264-267, main.py:241-244)
D435_640x480 = dict(width=640, height=480, fx=615.6707153320312, fy=615.962158203125,
                    cx=326.0557861328125, cy=240.55592346191406)
# synthetic 720p mode for BASELINE config 5 (the reference has no 720p mode)
SYN_1280x720 = dict(width=1280, height=720, fx=923.5, fy=923.5, cx=639.5, cy=359.5)

# YCB 011_banana bounding box / 2 as an ellipsoid resting on the plane z = 0
ELLIPSOID_R = (0.084, 0.034, 0.020)
ELLIPSOID_C = (0.0, 0.0, 0.020)


def look_at(eye, target, up=(0.0, 0.0, 1.0)):
    """camera->world 4x4 for a pinhole camera at `eye` looking at `target`
    (camera axes: +z forward, +x right, +y down, like Open3D / the D435)."""
    eye = np.asarray(eye, np.float64)
    target = np.asarray(target, np.float64)
    f = target - eye
    f /= np.linalg.norm(f)
    upv = np.asarray(up, np.float64)
    r = np.cross(f, upv)
    if np.linalg.norm(r) < 1e-9:
        r = np.cross(f, np.array([0.0, 1.0, 0.0]))
    r /= np.linalg.norm(r)
    d = np.cross(f, r)  # camera +y (down)
    c2w = np.eye(4)
    c2w[:3, 0], c2w[:3, 1], c2w[:3, 2], c2w[:3, 3] = r, d, f, eye
    return c2w


def orbit_poses(n_frames, radius=0.5, elevation_deg=30.0, target=(0.0, 0.0, 0.02), az0=0.0, az_span=2 * np.pi):
    """n_frames camera->world poses on a circle of `radius` around `target`."""
    el = np.deg2rad(elevation_deg)
    out = []
    for k in range(n_frames):
        az = az0 + az_span * k / n_frames
        eye = np.asarray(target, np.float64) + radius * np.array(
            [np.cos(el) * np.cos(az), np.cos(el) * np.sin(az), np.sin(el)])
        out.append(look_at(eye, target))
    return out


def main_orbit_poses(n_images=12, object_point=(0.0, 0.0, 0.0), circ_rad=0.3, circ_height=0.15, path_disc=500):
    """The simulator's camera path (main.py:68-136): a semicircle of radius 0.3 m in the vertical plane through the
    object at azimuth 3*pi/4, centre shifted 0.15*cos(pi/4) towards -x/-y and 0.15 m up, discretised into 500 waypoints,
    of which `img_total = 12` are photographed at equal time steps (main.py:213-232, `io.select_waypoints`).  The
    camera looks at the object with up = +z (bullet_camera.py:56-62 `computeViewMatrix(eye, target, [0,0,1])`).
    Returns camera->world 4x4 poses (camera axes +z forward, +x right, +y down)."""
    from .io import select_waypoints
    o = np.asarray(object_point, np.float64)
    a = np.linspace(0.0, np.pi, path_disc)
    x = o[0] - 0.15 * np.cos(np.pi / 4) + circ_rad * np.cos(a) * np.cos(3 * np.pi / 4)
    y = o[1] - 0.15 * np.cos(np.pi / 4) + circ_rad * np.cos(a) * np.sin(3 * np.pi / 4)
    z = circ_height + o[2] + circ_rad * np.sin(a)
    return [look_at((x[i], y[i], z[i]), o) for i in select_waypoints(path_disc, n_images)]


def extrinsic_from_pose(cam_to_world):
    """Open3D extrinsic = world->camera."""
    return np.linalg.inv(np.asarray(cam_to_world, np.float64))


def volume_origin(length):
    """cube of side L with the ground plane z=0 inside, object at the xy centre."""
    return np.array([-0.5 * length, -0.5 * length, -0.1 * length])


# ---------------------------------------------------------------------------------------------------
# device ray caster (b2r_raycast_scene)
# ---------------------------------------------------------------------------------------------------
def make_scene(cam_to_world, K, ell_c=ELLIPSOID_C, ell_r=ELLIPSOID_R, plane_z=0.0, has_plane=True, checker=0.04,
               depth_scale=1000.0, tris=None):
    """Fill the C scene struct (float32 parameters, identical to the oracle twin's inputs)."""
    from . import _lib
    sc = _lib.Scene()
    m = np.asarray(cam_to_world, np.float64)[:3, :4].astype(np.float32).reshape(-1)
    for i in range(12):
        sc.c2w[i] = float(m[i])
    sc.fx, sc.fy, sc.cx, sc.cy = [float(np.float32(K[k])) for k in ("fx", "fy", "cx", "cy")]
    sc.W, sc.H = int(K["width"]), int(K["height"])
    for i in range(3):
        sc.ell_c[i] = float(np.float32(ell_c[i]))
        sc.ell_r[i] = float(np.float32(ell_r[i]))
    sc.plane_z = float(np.float32(plane_z))
    sc.has_plane = int(bool(has_plane))
    sc.checker = float(np.float32(checker))
    sc.depth_scale = float(np.float32(depth_scale))
    sc.n_tris = 0
    if tris is not None:
        tr = np.ascontiguousarray(tris, np.float32).reshape(-1, 9)
        sc.n_tris = len(tr)
        mn, mx = tr.reshape(-1, 3).min(0), tr.reshape(-1, 3).max(0)
        for i in range(3):
            sc.mesh_min[i] = float(mn[i])
            sc.mesh_max[i] = float(mx[i])
    return sc


def render_frame(cam_to_world, K, tris_dev=None, tris=None, out_depth=None, out_color=None, **scene_kw):
    """Ray-cast one frame on the GPU -> (depth uint16 [H,W], color uint8 [H,W,3]) CUDA tensors."""
    import ctypes as C
    from . import _lib, runtime
    c = runtime.ctx()
    t = runtime.torch()
    sc = make_scene(cam_to_world, K, tris=tris, **scene_kw)
    if tris is not None and tris_dev is None:
        tris_dev = runtime.to_device(np.ascontiguousarray(tris, np.float32).reshape(-1, 9))
    H, W = sc.H, sc.W
    depth = out_depth if out_depth is not None else t.empty((H, W), dtype=t.uint16, device=c.tdev)
    color = out_color if out_color is not None else t.empty((H, W, 3), dtype=t.uint8, device=c.tdev)
    _lib.check(_lib.load().b2r_raycast_scene(c.handle, C.byref(sc), runtime.ptr(tris_dev), runtime.ptr(depth),
                                             runtime.ptr(color), c.stream()))
    return depth, color


def load_banana_triangles(path):
    """Triangles (float32 [T,9]) of a PLY mesh such as 011_banana/poisson/nontextured.ply, moved so that
    its bounding box rests on z = 0 and is centred in xy."""
    from . import io
    m = io.read_triangle_mesh(path)
    V = np.asarray(m.vertices).copy()
    T = np.asarray(m.triangles)
    mn, mx = V.min(0), V.max(0)
    V[:, 0] -= 0.5 * (mn[0] + mx[0])
    V[:, 1] -= 0.5 * (mn[1] + mx[1])
    V[:, 2] -= mn[2]
    return V[T].reshape(-1, 9).astype(np.float32)
