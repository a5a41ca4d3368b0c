"""ORACLE -- TEST INFRASTRUCTURE ONLY.  Slow numpy twin of the C++ oracle for tiny sizes.

Independent second restatement of SURVEY.md A.3 (UniformTSDFVolume::Integrate) used to cross-check
oracle/recon_oracle.cpp: numpy float32 arrays give IEEE single arithmetic with one rounding per
operation (no FMA), the z-chain is a Python loop so the incremental float32 adds are replayed.
"""
import numpy as np

f32 = np.float32


def integrate(tsdf, weight, color, depth, rgb, K, extrinsic, length, sdf_trunc, origin):
    """tsdf/weight float32 [R,R,R] ([x,y,z]), color float64 [R,R,R,3] or None -- updated in place."""
    R = tsdf.shape[0]
    H, W = depth.shape
    fx, fy, cx, cy = [f32(v) for v in K]
    E = np.asarray(extrinsic, np.float64).astype(f32)
    vl = f32(length / R)
    hvl = f32(vl * f32(0.5))
    trunc = f32(sdf_trunc)
    trunc_inv = f32(1.0) / trunc
    Es = (E * vl).astype(f32)
    safe_w, safe_h = f32(W) - f32(0.0001), f32(H) - f32(0.0001)
    ffl_inv0, ffl_inv1 = f32(1.0) / fx, f32(1.0) / fy
    xx = ((np.arange(W, dtype=f32) - cx) * ffl_inv0).astype(f32)
    yy = ((np.arange(H, dtype=f32) - cy) * ffl_inv1).astype(f32)
    mult = np.sqrt((xx * xx)[None, :] + (yy * yy)[:, None] + f32(1.0)).astype(f32)
    idx = np.arange(R, dtype=f32)
    px = ((hvl + vl * idx).astype(f32).astype(np.float64) + origin[0]).astype(f32)
    py = ((hvl + vl * idx).astype(f32).astype(np.float64) + origin[1]).astype(f32)
    pz = f32(np.float64(hvl) + origin[2])
    PX, PY = np.meshgrid(px, py, indexing="ij")
    pc = []
    for r in range(3):
        v = ((E[r, 0] * PX + E[r, 1] * PY).astype(f32) + E[r, 2] * pz).astype(f32) + E[r, 3] * f32(1.0)
        pc.append(v.astype(f32))
    updates = 0
    for z in range(R):
        X, Y, Z = pc
        with np.errstate(divide="ignore", invalid="ignore", over="ignore"):
            u_f = ((X * fx).astype(f32) / Z + cx).astype(f32) + f32(0.5)
            v_f = ((Y * fy).astype(f32) / Z + cy).astype(f32) + f32(0.5)
        ok = (Z > 0) & (u_f >= f32(0.0001)) & (u_f < safe_w) & (v_f >= f32(0.0001)) & (v_f < safe_h)
        u = np.where(ok, u_f, 0).astype(np.int32)
        v = np.where(ok, v_f, 0).astype(np.int32)
        d = depth[v, u]
        ok &= d > 0
        sdf = ((d - Z).astype(f32) * mult[v, u]).astype(f32)
        ok &= sdf > -trunc
        s = np.minimum(f32(1.0), (sdf * trunc_inv).astype(f32))
        t, w = tsdf[:, :, z], weight[:, :, z]
        newt = (((t * w).astype(f32) + s).astype(f32) / (w + f32(1.0)).astype(f32)).astype(f32)
        tsdf[:, :, z] = np.where(ok, newt, t)
        if color is not None:
            c = color[:, :, z, :]
            wd = w.astype(np.float64)[..., None]
            w1 = (w + f32(1.0)).astype(np.float64)[..., None]
            newc = (c * wd + rgb[v, u, :].astype(np.float64)) / w1
            color[:, :, z, :] = np.where(ok[..., None], newc, c)
        weight[:, :, z] = np.where(ok, (w + f32(1.0)).astype(f32), w)
        updates += int(ok.sum())
        pc = [(pc[r] + Es[r, 2]).astype(f32) for r in range(3)]
    return updates
