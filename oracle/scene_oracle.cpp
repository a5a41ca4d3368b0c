// ORACLE -- TEST INFRASTRUCTURE ONLY.
//
// CPU twin of the headless synthetic frame source: replaces
// the reference's GL camera (bullet_camera.py:48-85, "unavailable on this box") and the D435
// (data_recorder.py:104-121) with plain ray casting: analytic ellipsoid resting on the ground
// plane z=0, or a triangle mesh (the committed 011_banana YCB mesh) on that plane; colour is a
// procedural 4 cm checker of the world position.  All arithmetic is float32 with separately
// rounded + - * / sqrt (build with -ffp-contract=off) so that the CUDA kernel, compiled with
// -fmad=false, produces bit-identical uint16 depth and uint8 colour.
#include <cmath>
#include <cstdint>
#include <cstring>

extern "C" {

struct OrcScene {
    float c2w[12];      // camera->world, row-major 3x4 [R|t]
    float fx, fy, cx, cy;
    int W, H;
    float ell_c[3], ell_r[3];   // ellipsoid centre / semi-axes (ell_r[0] <= 0: no ellipsoid)
    float plane_z;              // ground plane height (world z)
    int has_plane;
    float checker;              // checker period (m)
    float depth_scale;          // 1000 -> millimetres
    int n_tris;                 // > 0: mesh mode
    float mesh_min[3], mesh_max[3];
};

static inline void shade(const float* p, int is_obj, float period, uint8_t* rgb) {
    const int ix = (int)floorf(p[0] / period), iy = (int)floorf(p[1] / period), iz = (int)floorf(p[2] / period);
    const int odd = (ix + iy + iz) & 1;
    if (is_obj) {
        if (odd) { rgb[0] = 120; rgb[1] = 160; rgb[2] = 30; } else { rgb[0] = 230; rgb[1] = 200; rgb[2] = 40; }
    } else {
        if (odd) { rgb[0] = 70; rgb[1] = 90; rgb[2] = 110; } else { rgb[0] = 210; rgb[1] = 180; rgb[2] = 140; }
    }
}

// tris: float32 [n_tris, 9] (v0,v1,v2 world coordinates).  depth: uint16 [H,W]; color: uint8 [H,W,3].
int orc_raycast_scene(const OrcScene* sc, const float* tris, uint16_t* depth, uint8_t* color) {
    const int W = sc->W, H = sc->H;
    const float* M = sc->c2w;
    const float ox = M[3], oy = M[7], oz = M[11];
#pragma omp parallel for schedule(dynamic, 4)
    for (int v = 0; v < H; ++v) {
        for (int u = 0; u < W; ++u) {
            const float xc = ((float)u - sc->cx) / sc->fx;
            const float yc = ((float)v - sc->cy) / sc->fy;
            // world direction of the camera-frame ray (xc, yc, 1): t is the z-depth directly
            const float dx = (M[0] * xc + M[1] * yc) + M[2];
            const float dy = (M[4] * xc + M[5] * yc) + M[6];
            const float dz = (M[8] * xc + M[9] * yc) + M[10];
            float tbest = 3.0e38f; int hit = 0, is_obj = 0;
            if (sc->has_plane && dz != 0.0f) {
                const float t = (sc->plane_z - oz) / dz;
                if (t > 0.0f && t < tbest) { tbest = t; hit = 1; is_obj = 0; }
            }
            if (sc->n_tris <= 0 && sc->ell_r[0] > 0.0f) {
                // |S(o + t d - c)|^2 = 1 with S = diag(1/r)
                const float ax = dx / sc->ell_r[0], ay = dy / sc->ell_r[1], az = dz / sc->ell_r[2];
                const float bx = (ox - sc->ell_c[0]) / sc->ell_r[0], by = (oy - sc->ell_c[1]) / sc->ell_r[1],
                            bz = (oz - sc->ell_c[2]) / sc->ell_r[2];
                const float A = (ax * ax + ay * ay) + az * az;
                const float B = (ax * bx + ay * by) + az * bz;
                const float C = ((bx * bx + by * by) + bz * bz) - 1.0f;
                const float disc = B * B - A * C;
                if (disc > 0.0f) {
                    const float t = (-B - sqrtf(disc)) / A;
                    if (t > 0.0f && t < tbest) { tbest = t; hit = 1; is_obj = 1; }
                }
            }
            if (sc->n_tris > 0) {
                // slab test against the mesh AABB first
                float t0 = 0.0f, t1 = tbest; int ok = 1;
                const float o[3] = {ox, oy, oz}, d[3] = {dx, dy, dz};
                for (int k = 0; k < 3 && ok; ++k) {
                    if (d[k] == 0.0f) { if (o[k] < sc->mesh_min[k] || o[k] > sc->mesh_max[k]) ok = 0; continue; }
                    float ta = (sc->mesh_min[k] - o[k]) / d[k], tb = (sc->mesh_max[k] - o[k]) / d[k];
                    if (ta > tb) { const float s = ta; ta = tb; tb = s; }
                    if (ta > t0) t0 = ta;
                    if (tb < t1) t1 = tb;
                    if (t0 > t1) ok = 0;
                }
                if (ok) {
                    for (int i = 0; i < sc->n_tris; ++i) {
                        const float* T = tris + (int64_t)i * 9;
                        // Moller-Trumbore, float32, fixed operation order
                        const float e1x = T[3] - T[0], e1y = T[4] - T[1], e1z = T[5] - T[2];
                        const float e2x = T[6] - T[0], e2y = T[7] - T[1], e2z = T[8] - T[2];
                        const float px = dy * e2z - dz * e2y, py = dz * e2x - dx * e2z, pz = dx * e2y - dy * e2x;
                        const float det = (e1x * px + e1y * py) + e1z * pz;
                        if (det > -1e-12f && det < 1e-12f) continue;
                        const float inv = 1.0f / det;
                        const float sx = ox - T[0], sy = oy - T[1], sz = oz - T[2];
                        const float uu = ((sx * px + sy * py) + sz * pz) * inv;
                        if (uu < 0.0f || uu > 1.0f) continue;
                        const float qx = sy * e1z - sz * e1y, qy = sz * e1x - sx * e1z, qz = sx * e1y - sy * e1x;
                        const float vv = ((dx * qx + dy * qy) + dz * qz) * inv;
                        if (vv < 0.0f || uu + vv > 1.0f) continue;
                        const float t = ((e2x * qx + e2y * qy) + e2z * qz) * inv;
                        if (t > 0.0f && t < tbest) { tbest = t; hit = 1; is_obj = 1; }
                    }
                }
            }
            const int64_t pix = (int64_t)v * W + u;
            uint16_t dq = 0; uint8_t rgb[3] = {0, 0, 0};
            if (hit) {
                const float mm = tbest * sc->depth_scale + 0.5f;
                if (mm < 65535.0f) dq = (uint16_t)(int)mm;
                const float p[3] = {ox + tbest * dx, oy + tbest * dy, oz + tbest * dz};
                shade(p, is_obj, sc->checker, rgb);
            }
            depth[pix] = dq;
            color[pix * 3 + 0] = rgb[0]; color[pix * 3 + 1] = rgb[1]; color[pix * 3 + 2] = rgb[2];
        }
    }
    return 0;
}

}  // extern "C"
