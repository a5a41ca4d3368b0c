// Headless synthetic frame source: one thread per pixel casts the camera
// ray against an analytic ellipsoid + ground plane, or brute-force against a triangle mesh staged
// through shared memory (the 14 898-triangle YCB banana of BASELINE config 1).  Stands in for the
// reference's GL camera (bullet_camera.py:48-85) and the D435 (data_recorder.py:104-121), which are
// "unavailable on this box".  float32 with separate roundings (-fmad=false) -> bit-identical to the
// CPU twin in oracle/scene_oracle.cpp.
#include "common.cuh"

namespace b2r {

__device__ __forceinline__ void shade(float px, float py, float pz, int is_obj, float period, uint8_t* rgb) {
    const int ix = (int)floorf(px / period), iy = (int)floorf(py / period), iz = (int)floorf(pz / period);
    const int odd = (ix + iy + iz) & 1;
    if (is_obj) {
        if (odd) { rgb[0] = 120; rgb[1] = 160; rgb[2] = 30; } else { rgb[0] = 230; rgb[1] = 200; rgb[2] = 40; }
    } else {
        if (odd) { rgb[0] = 70; rgb[1] = 90; rgb[2] = 110; } else { rgb[0] = 210; rgb[1] = 180; rgb[2] = 140; }
    }
}

constexpr int RC_TX = 32, RC_TY = 8;
constexpr int RC_TRI_CHUNK = 256;  // triangles staged per pass (256 * 36 B = 9 KB)

__global__ void __launch_bounds__(RC_TX* RC_TY) raycast_scene_kernel(const __grid_constant__ b2r_scene sc,
                                                                      const float* __restrict__ tris,
                                                                      uint16_t* __restrict__ depth,
                                                                      uint8_t* __restrict__ color) {
    __shared__ float s_tri[RC_TRI_CHUNK * 9];
    const int u = blockIdx.x * RC_TX + threadIdx.x;
    const int v = blockIdx.y * RC_TY + threadIdx.y;
    const bool inside = (u < sc.W) && (v < sc.H);
    const float* M = sc.c2w;
    const float ox = M[3], oy = M[7], oz = M[11];
    const float xc = ((float)u - sc.cx) / sc.fx;
    const float yc = ((float)v - sc.cy) / sc.fy;
    const float dx = (M[0] * xc + M[1] * yc) + M[2];
    const float dy = (M[4] * xc + M[5] * yc) + M[6];
    const float dz = (M[8] * xc + M[9] * yc) + M[10];
    float tbest = 3.0e38f;
    int hit = 0, is_obj = 0;
    if (sc.has_plane && dz != 0.0f) {
        const float t = (sc.plane_z - oz) / dz;
        if (t > 0.0f && t < tbest) { tbest = t; hit = 1; is_obj = 0; }
    }
    if (sc.n_tris <= 0 && sc.ell_r[0] > 0.0f) {
        const float ax = dx / sc.ell_r[0], ay = dy / sc.ell_r[1], az = dz / sc.ell_r[2];
        const float bx = (ox - sc.ell_c[0]) / sc.ell_r[0], by = (oy - sc.ell_c[1]) / sc.ell_r[1],
                    bz = (oz - sc.ell_c[2]) / sc.ell_r[2];
        const float A = (ax * ax + ay * ay) + az * az;
        const float B = (ax * bx + ay * by) + az * bz;
        const float C = ((bx * bx + by * by) + bz * bz) - 1.0f;
        const float disc = B * B - A * C;
        if (disc > 0.0f) {
            const float t = (-B - sqrtf(disc)) / A;
            if (t > 0.0f && t < tbest) { tbest = t; hit = 1; is_obj = 1; }
        }
    }
    if (sc.n_tris > 0) {
        float t0 = 0.0f, t1 = tbest;
        int ok = inside ? 1 : 0;
        const float o[3] = {ox, oy, oz}, d[3] = {dx, dy, dz};
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            if (!ok) break;
            if (d[k] == 0.0f) {
                if (o[k] < sc.mesh_min[k] || o[k] > sc.mesh_max[k]) ok = 0;
                continue;
            }
            float ta = (sc.mesh_min[k] - o[k]) / d[k], tb = (sc.mesh_max[k] - o[k]) / d[k];
            if (ta > tb) { const float s = ta; ta = tb; tb = s; }
            if (ta > t0) t0 = ta;
            if (tb < t1) t1 = tb;
            if (t0 > t1) ok = 0;
        }
        const int any_ok = __syncthreads_or(ok);
        if (any_ok) {
            const int tid = threadIdx.y * RC_TX + threadIdx.x;
            for (int base = 0; base < sc.n_tris; base += RC_TRI_CHUNK) {
                const int cnt = min(RC_TRI_CHUNK, sc.n_tris - base);
                __syncthreads();
                for (int i = tid; i < cnt * 9; i += RC_TX * RC_TY) s_tri[i] = tris[(int64_t)base * 9 + i];
                __syncthreads();
                if (!ok) continue;
                for (int i = 0; i < cnt; ++i) {
                    const float* T = s_tri + i * 9;
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
    }
    if (!inside) return;
    const int64_t pix = (int64_t)v * sc.W + u;
    uint16_t dq = 0;
    uint8_t rgb[3] = {0, 0, 0};
    if (hit) {
        const float mm = tbest * sc.depth_scale + 0.5f;
        if (mm < 65535.0f) dq = (uint16_t)(int)mm;
        shade(ox + tbest * dx, oy + tbest * dy, oz + tbest * dz, is_obj, sc.checker, rgb);
    }
    depth[pix] = dq;
    color[pix * 3 + 0] = rgb[0];
    color[pix * 3 + 1] = rgb[1];
    color[pix * 3 + 2] = rgb[2];
}

}  // namespace b2r

using namespace b2r;

extern "C" int b2r_raycast_scene(b2r_ctx* ctx, const b2r_scene* scene, const float* d_tris, uint16_t* d_depth,
                                 uint8_t* d_color, void* stream) {
    B2R_CHECK(ctx && scene && d_depth && d_color, -1, "b2r_raycast_scene: NULL argument");
    B2R_CHECK(scene->n_tris <= 0 || d_tris, -1, "b2r_raycast_scene: mesh mode without triangles");
    B2R_CHECK(scene->W > 0 && scene->H > 0, -5, "b2r_raycast_scene: bad image size %dx%d", scene->W, scene->H);
    B2R_CUDA(cudaSetDevice(ctx->device));
    dim3 block(RC_TX, RC_TY), grid(cdiv(scene->W, RC_TX), cdiv(scene->H, RC_TY));
    raycast_scene_kernel<<<grid, block, 0, S(stream)>>>(*scene, d_tris, d_depth, d_color);
    B2R_LAUNCH_CHECK(ctx);
    return 0;
}
