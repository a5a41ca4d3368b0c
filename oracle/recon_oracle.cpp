// ORACLE -- TEST INFRASTRUCTURE ONLY.  Not part of the product path.
//
// CPU restatement (C++17 + OpenMP) of the Open3D 0.18.0 legacy-CPU arithmetic that the
// reference's stitcher.py delegates to (reference call sites: stitcher.py:34-46 image
// conversion + back-projection; north_star API UniformTSDFVolume.integrate /
// extract_triangle_mesh / extract_point_cloud).  Open3D itself is a pinned third-party wheel
// (environment.yml:72-73, open3d==0.18.0) that is absent from /root/reference and cannot be
// installed here, and the reference holds no tests or golden vectors for this path, so:
//
//     *** PARITY UNPINNED against the real Open3D binary. ***
//
// What this file follows is SURVEY.md Appendix A (A.1 image conversion, A.2 back-projection,
// A.3 UniformTSDFVolume::Integrate, A.4 ExtractTriangleMesh / ExtractPointCloud) and
// Appendix B (tables, sha1-checked by tools/gen_mc_tables.py).
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
// may load this library.  Build: see oracle/Makefile (-O3 -fopenmp -ffp-contract=off, generic
// x86-64: float32 results must not depend on FMA contraction).
#include <omp.h>

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <unordered_map>
#include <vector>

#include "recon_internal.h"

extern "C" {

// OpenMP team size of the oracle (torchrun exports OMP_NUM_THREADS=1 to its workers; the timed CPU legs of bench.py
// ask for all host cores explicitly and report what they got).
void orc_set_num_threads(int n) { if (n > 0) omp_set_num_threads(n); }
int orc_max_threads(void) { return omp_get_max_threads(); }

// ------------------------------------------------------------------------------------------
// A.1  Image::ConvertDepthToFloatImage(depth_scale, depth_trunc)
// dtype: 0 = uint8 (v/255.0f), 1 = uint16 ((float)v), 2 = float32 (v).
// ------------------------------------------------------------------------------------------
int orc_depth_to_float(const void* depth, int dtype, int H, int W, double depth_scale,
                       double depth_trunc, float* out) {
    const float scale = (float)depth_scale, trunc = (float)depth_trunc;
    const int64_t n = (int64_t)H * W;
    for (int64_t i = 0; i < n; ++i) {
        float p;
        if (dtype == 0) p = (float)((const uint8_t*)depth)[i] / 255.0f;
        else if (dtype == 1) p = (float)((const uint16_t*)depth)[i];
        else if (dtype == 2) p = ((const float*)depth)[i];
        else return -1;
        p /= scale;
        if (p >= trunc) p = 0.0f;
        out[i] = p;
    }
    return 0;
}

// 4x4 inverse of a rigid/affine matrix in double (general Gauss-Jordan with partial pivoting;
// Open3D uses Eigen's inverse() -- results agree to rounding, points are compared with a
// tolerance in the tests, not bit-exactly).
static void inv4(const double* m, double* out) {
    double a[4][8];
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) { a[i][j] = m[i * 4 + j]; a[i][4 + j] = (i == j); }
    for (int c = 0; c < 4; ++c) {
        int p = c;
        for (int r = c + 1; r < 4; ++r) if (std::fabs(a[r][c]) > std::fabs(a[p][c])) p = r;
        if (p != c) for (int j = 0; j < 8; ++j) std::swap(a[p][j], a[c][j]);
        double d = a[c][c];
        for (int j = 0; j < 8; ++j) a[c][j] /= d;
        for (int r = 0; r < 4; ++r) if (r != c) {
            double f = a[r][c];
            if (f != 0.0) for (int j = 0; j < 8; ++j) a[r][j] -= f * a[c][j];
        }
    }
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) out[i * 4 + j] = a[i][4 + j];
}

// ------------------------------------------------------------------------------------------
// A.2  PointCloud::CreateFromRGBDImage(rgbd, intrinsic, extrinsic, project_valid_depth_only)
// depth: float32 metres.  color: uint8 with nc channels (3 or 4), or NULL.  Row-major scan,
// invalid (d <= 0) pixels dropped.  Returns the number of points written.
// ------------------------------------------------------------------------------------------
int64_t orc_backproject(const float* depth, const uint8_t* color, int nc, int H, int W, double fx,
                        double fy, double cx, double cy, const double* extrinsic, double* pts,
                        double* cols) {
    double pose[16];
    inv4(extrinsic, pose);
    int64_t n = 0;
    for (int i = 0; i < H; ++i) {
        for (int j = 0; j < W; ++j) {
            const float d = depth[(int64_t)i * W + j];
            if (!(d > 0)) continue;
            const double z = (double)d;
            const double x = (j - cx) * z / fx;
            const double y = (i - cy) * z / fy;
            for (int r = 0; r < 3; ++r)
                pts[n * 3 + r] = ((pose[r * 4 + 0] * x + pose[r * 4 + 1] * y) + pose[r * 4 + 2] * z) +
                                 pose[r * 4 + 3];
            if (color && cols) {
                const uint8_t* pc = color + ((int64_t)i * W + j) * nc;
                cols[n * 3 + 0] = pc[0] / 255.0;
                cols[n * 3 + 1] = pc[(nc - 1) / 2] / 255.0;
                cols[n * 3 + 2] = pc[nc - 1] / 255.0;
            }
            ++n;
        }
    }
    return n;
}

// ------------------------------------------------------------------------------------------
// A.3  UniformTSDFVolume.  Storage window: x in [x_begin, x_end) (the whole cube by default);
// the window exists only so that bench.py can time a bounded x-range of a 512^3/1024^3 volume
// without allocating all of it -- Open3D parallelises Integrate over x, so an x-range is the
// natural sample.  IndexOf(x,y,z) = x*R*R + y*R + z (z fastest), as Open3D.
// ------------------------------------------------------------------------------------------

void* orc_volume_create(double length, int resolution, double sdf_trunc, int color_type,
                        const double* origin, int x_begin, int x_end) {
    OrcVolume* v = new OrcVolume();
    v->length = length; v->R = resolution; v->sdf_trunc = sdf_trunc; v->color_type = color_type;
    v->voxel_length = length / (double)resolution;
    for (int i = 0; i < 3; ++i) v->origin[i] = origin ? origin[i] : 0.0;
    if (x_end <= x_begin) { x_begin = 0; x_end = resolution; }
    v->x_begin = x_begin; v->x_end = x_end;
    const int64_t n = (int64_t)(x_end - x_begin) * resolution * resolution;
    v->tsdf.assign(n, 0.0f); v->weight.assign(n, 0.0f);
    if (color_type != 0) v->color.assign(n * 3, 0.0);
    return v;
}
void orc_volume_destroy(void* h) { delete (OrcVolume*)h; }
void orc_volume_reset(void* h) {
    OrcVolume* v = (OrcVolume*)h;
    std::fill(v->tsdf.begin(), v->tsdf.end(), 0.0f);
    std::fill(v->weight.begin(), v->weight.end(), 0.0f);
    std::fill(v->color.begin(), v->color.end(), 0.0);
}
float* orc_volume_tsdf(void* h) { return ((OrcVolume*)h)->tsdf.data(); }
float* orc_volume_weight(void* h) { return ((OrcVolume*)h)->weight.data(); }
double* orc_volume_color(void* h) {
    OrcVolume* v = (OrcVolume*)h; return v->color.empty() ? nullptr : v->color.data();
}

// Image::CreateDepthToCameraDistanceMultiplierFloatImage (all float32).
void orc_distance_multiplier(int H, int W, double fx, double fy, double cx, double cy, float* out) {
    const float ffl_inv0 = 1.0f / (float)fx, ffl_inv1 = 1.0f / (float)fy;
    const float fpp0 = (float)cx, fpp1 = (float)cy;
    std::vector<float> xx(W), yy(H);
    for (int j = 0; j < W; ++j) xx[j] = ((float)j - fpp0) * ffl_inv0;
    for (int i = 0; i < H; ++i) yy[i] = ((float)i - fpp1) * ffl_inv1;
    for (int i = 0; i < H; ++i)
        for (int j = 0; j < W; ++j)
            out[(int64_t)i * W + j] = sqrtf(xx[j] * xx[j] + yy[i] * yy[i] + 1.0f);
}

// UniformTSDFVolume::IntegrateWithDepthToCameraDistanceMultiplier, the columns of ONE x index (the body of
// Open3D's `omp parallel for` over x).  Shared by the uniform volume and by the 16^3 units of the scalable one.

void make_frame(IntegrateFrame& F, const OrcVolume* vol, const float* depth, const void* color, const float* mult,
                       int H, int W, double fx_d, double fy_d, double cx_d, double cy_d, const double* extrinsic) {
    F.depth = depth; F.color = color; F.mult = mult; F.H = H; F.W = W;
    F.fx = (float)fx_d; F.fy = (float)fy_d; F.cx = (float)cx_d; F.cy = (float)cy_d;
    F.vl = (float)vol->voxel_length;
    F.hvl = F.vl * 0.5f;
    F.trunc = (float)vol->sdf_trunc;
    F.trunc_inv = 1.0f / F.trunc;
    for (int r = 0; r < 4; ++r)
        for (int c = 0; c < 4; ++c) { F.E[r][c] = (float)extrinsic[r * 4 + c]; F.Es[r][c] = F.E[r][c] * F.vl; }
    F.safe_w = (float)W - 0.0001f; F.safe_h = (float)H - 0.0001f;
}

int64_t integrate_x(OrcVolume* vol, const IntegrateFrame& F, int x) {
    const float fx = F.fx, fy = F.fy, cx = F.cx, cy = F.cy, vl = F.vl, hvl = F.hvl, trunc = F.trunc, trunc_inv = F.trunc_inv;
    const float safe_w = F.safe_w, safe_h = F.safe_h;
    const float* depth = F.depth; const float* mult = F.mult;
    const int W = F.W, R = vol->R;
    const uint8_t* rgb8 = (const uint8_t*)F.color;
    const float* gray = (const float*)F.color;
    int64_t updates = 0;
    for (int y = 0; y < R; ++y) {
        // pt_3d_homo(float(hvl + vl*x + origin_(0)), ...): the float sum is promoted to
        // double by origin_ (Vector3d) and rounded back to float.
        const float px = (float)((double)(hvl + vl * (float)x) + vol->origin[0]);
        const float py = (float)((double)(hvl + vl * (float)y) + vol->origin[1]);
        const float pz = (float)((double)hvl + vol->origin[2]);
        // Eigen 4x4 * 4x1 float, coefficient-based lazy product, column-major packets:
        // ((col0*x + col1*y) + col2*z) + col3*1, separate mul/add roundings (no FMA).
        float pc[3];
        for (int r = 0; r < 3; ++r)
            pc[r] = ((F.E[r][0] * px + F.E[r][1] * py) + F.E[r][2] * pz) + F.E[r][3] * 1.0f;
        for (int z = 0; z < R; ++z, pc[0] += F.Es[0][2], pc[1] += F.Es[1][2], pc[2] += F.Es[2][2]) {
            if (pc[2] <= 0) continue;
            const float u_f = pc[0] * fx / pc[2] + cx + 0.5f;
            const float v_f = pc[1] * fy / pc[2] + cy + 0.5f;
            if (!(u_f >= 0.0001f && u_f < safe_w && v_f >= 0.0001f && v_f < safe_h)) continue;
            const int u = (int)u_f, v = (int)v_f;
            const float d = depth[(int64_t)v * W + u];
            if (d <= 0.0f) continue;
            const float sdf = (d - pc[2]) * mult[(int64_t)v * W + u];
            if (sdf > -trunc) {
                const float tsdf = std::min(1.0f, sdf * trunc_inv);
                const int64_t i = vol->idx(x, y, z);
                const float w = vol->weight[i];
                vol->tsdf[i] = (vol->tsdf[i] * w + tsdf) / (w + 1.0f);
                if (vol->color_type == 1) {
                    const uint8_t* p = rgb8 + ((int64_t)v * W + u) * 3;
                    double* c = &vol->color[i * 3];
                    const double wd = (double)w, w1 = (double)(w + 1.0f);
                    c[0] = (c[0] * wd + (double)p[0]) / w1;
                    c[1] = (c[1] * wd + (double)p[1]) / w1;
                    c[2] = (c[2] * wd + (double)p[2]) / w1;
                } else if (vol->color_type == 2) {
                    double* c = &vol->color[i * 3];
                    const double g = (double)gray[(int64_t)v * W + u];
                    const double wd = (double)w, w1 = (double)(w + 1.0f);
                    c[0] = (c[0] * wd + g) / w1; c[1] = c[0]; c[2] = c[0];
                }
                vol->weight[i] = w + 1.0f;
                ++updates;
            }
        }
    }
    return updates;
}

// UniformTSDFVolume::Integrate.
// depth: float32 metres [H,W]; color: uint8 [H,W,3] (RGB8) or float32 [H,W] (Gray32) or NULL.
// Returns the number of voxel updates performed (for the voxel-updates/s metric), <0 on error.
int64_t orc_volume_integrate(void* h, const float* depth, const void* color, int H, int W,
                             double fx_d, double fy_d, double cx_d, double cy_d,
                             const double* extrinsic /*row-major world->camera*/) {
    OrcVolume* vol = (OrcVolume*)h;
    if (vol->color_type != 0 && !color) return -2;
    std::vector<float> mult((size_t)H * W);
    orc_distance_multiplier(H, W, fx_d, fy_d, cx_d, cy_d, mult.data());
    IntegrateFrame F;
    make_frame(F, vol, depth, color, mult.data(), H, W, fx_d, fy_d, cx_d, cy_d, extrinsic);
    int64_t updates = 0;
#pragma omp parallel for schedule(static) reduction(+ : updates)
    for (int x = vol->x_begin; x < vol->x_end; ++x) updates += integrate_x(vol, F, x);
    return updates;
}

// ------------------------------------------------------------------------------------------
// A.4  ExtractTriangleMesh: serial x,y,z scan, vertices welded through an edge hash, created
// lazily by the first ACTIVE cube that touches the edge; winding (i, i+2, i+1).
// Requires the full cube to be stored (x window == [0,R)).
// ------------------------------------------------------------------------------------------
struct Key4 {
    int x, y, z, a;
    bool operator==(const Key4& o) const { return x == o.x && y == o.y && z == o.z && a == o.a; }
};
struct Key4Hash {
    size_t operator()(const Key4& k) const {
        uint64_t h = (uint64_t)(uint32_t)k.x * 0x9E3779B97F4A7C15ull;
        h ^= (uint64_t)(uint32_t)k.y * 0xC2B2AE3D27D4EB4Full + (h << 6) + (h >> 2);
        h ^= (uint64_t)(uint32_t)k.z * 0x165667B19E3779F9ull + (h << 6) + (h >> 2);
        h ^= (uint64_t)k.a + (h << 6) + (h >> 2);
        return (size_t)h;
    }
};

int orc_volume_extract_mesh(void* h, int64_t* n_vertices, int64_t* n_triangles) {
    OrcVolume* vol = (OrcVolume*)h;
    if (vol->x_begin != 0 || vol->x_end != vol->R) return -1;
    const int R = vol->R;
    const double vl = vol->voxel_length, hvl = 0.5 * vl;
    vol->verts.clear(); vol->vcols.clear(); vol->tris.clear();
    std::unordered_map<Key4, int, Key4Hash> edge_to_index;
    int e2v[12];
    for (int x = 0; x < R - 1; ++x) {
        for (int y = 0; y < R - 1; ++y) {
            for (int z = 0; z < R - 1; ++z) {
                int cube_index = 0;
                float f[8];
                double c[8][3];
                for (int i = 0; i < 8; ++i) {
                    const int64_t vi = vol->idx(x + mc_shift[i][0], y + mc_shift[i][1], z + mc_shift[i][2]);
                    if (vol->weight[vi] == 0.0f) { cube_index = 0; break; }
                    f[i] = vol->tsdf[vi];
                    if (f[i] < 0.0f) cube_index |= (1 << i);
                    if (vol->color_type == 1) {
                        for (int k = 0; k < 3; ++k) c[i][k] = vol->color[vi * 3 + k] / 255.0;
                    } else if (vol->color_type == 2) {
                        for (int k = 0; k < 3; ++k) c[i][k] = vol->color[vi * 3 + k];
                    }
                }
                if (cube_index == 0 || cube_index == 255) continue;
                for (int i = 0; i < 12; ++i) {
                    if (!(mc_edge_table[cube_index] & (1 << i))) continue;
                    Key4 key{x + mc_edge_shift[i][0], y + mc_edge_shift[i][1], z + mc_edge_shift[i][2],
                             mc_edge_shift[i][3]};
                    auto it = edge_to_index.find(key);
                    if (it != edge_to_index.end()) { e2v[i] = it->second; continue; }
                    e2v[i] = (int)(vol->verts.size() / 3);
                    edge_to_index.emplace(key, e2v[i]);
                    double pt[3] = {hvl + vl * key.x, hvl + vl * key.y, hvl + vl * key.z};
                    const double f0 = std::fabs((double)f[mc_edge_to_vert[i][0]]);
                    const double f1 = std::fabs((double)f[mc_edge_to_vert[i][1]]);
                    pt[key.a] += f0 * vl / (f0 + f1);
                    for (int k = 0; k < 3; ++k) vol->verts.push_back(pt[k] + vol->origin[k]);
                    if (vol->color_type != 0) {
                        const double* c0 = c[mc_edge_to_vert[i][0]];
                        const double* c1 = c[mc_edge_to_vert[i][1]];
                        for (int k = 0; k < 3; ++k) vol->vcols.push_back((f1 * c0[k] + f0 * c1[k]) / (f0 + f1));
                    }
                }
                for (int i = 0; mc_tri_table[cube_index][i] != -1; i += 3) {
                    vol->tris.push_back(e2v[mc_tri_table[cube_index][i]]);
                    vol->tris.push_back(e2v[mc_tri_table[cube_index][i + 2]]);
                    vol->tris.push_back(e2v[mc_tri_table[cube_index][i + 1]]);
                }
            }
        }
    }
    *n_vertices = (int64_t)(vol->verts.size() / 3);
    *n_triangles = (int64_t)(vol->tris.size() / 3);
    return 0;
}

// GetTSDFAt / GetNormalAt (trilinear, no weight check) -- p is WITHOUT origin.
static double orc_tsdf_at(const OrcVolume* vol, const double* p) {
    const double vl = vol->voxel_length;
    int idx[3]; double r[3];
    for (int k = 0; k < 3; ++k) {
        const double g = p[k] / vl - 0.5;
        idx[k] = (int)std::floor(g);
        r[k] = g - idx[k];
    }
    auto T = [&](int dx, int dy, int dz) -> double {
        return (double)vol->tsdf[vol->idx(idx[0] + dx, idx[1] + dy, idx[2] + dz)];
    };
    return (1 - r[0]) * ((1 - r[1]) * ((1 - r[2]) * T(0, 0, 0) + r[2] * T(0, 0, 1)) +
                         r[1] * ((1 - r[2]) * T(0, 1, 0) + r[2] * T(0, 1, 1))) +
           r[0] * ((1 - r[1]) * ((1 - r[2]) * T(1, 0, 0) + r[2] * T(1, 0, 1)) +
                   r[1] * ((1 - r[2]) * T(1, 1, 0) + r[2] * T(1, 1, 1)));
}

int orc_volume_extract_points(void* h, int64_t* n_points) {
    OrcVolume* vol = (OrcVolume*)h;
    if (vol->x_begin != 0 || vol->x_end != vol->R) return -1;
    const int R = vol->R;
    const double vl = vol->voxel_length, hvl = 0.5 * vl;
    vol->verts.clear(); vol->vcols.clear(); vol->normals.clear();
    for (int x = 1; x < R - 1; ++x) {
        for (int y = 1; y < R - 1; ++y) {
            for (int z = 1; z < R - 1; ++z) {
                const int idx0[3] = {x, y, z};
                const int64_t i0 = vol->idx(x, y, z);
                const float w0 = vol->weight[i0], f0 = vol->tsdf[i0];
                if (!(w0 != 0.0f && f0 < 0.98f && f0 >= -0.98f)) continue;
                const double p0[3] = {hvl + vl * x, hvl + vl * y, hvl + vl * z};
                for (int a = 0; a < 3; ++a) {
                    int idx1[3] = {x, y, z};
                    idx1[a] += 1;
                    if (!(idx1[a] < R - 1)) continue;
                    const int64_t i1 = vol->idx(idx1[0], idx1[1], idx1[2]);
                    const float w1 = vol->weight[i1], f1 = vol->tsdf[i1];
                    if (!(w1 != 0.0f && f1 < 0.98f && f1 >= -0.98f && f0 * f1 < 0)) continue;
                    const float r0 = std::fabs(f0), r1 = std::fabs(f1);
                    double p1[3] = {p0[0], p0[1], p0[2]};
                    p1[a] += vl;
                    double p[3] = {p0[0], p0[1], p0[2]};
                    p[a] = (p0[a] * (double)r1 + p1[a] * (double)r0) / (double)(r0 + r1);
                    for (int k = 0; k < 3; ++k) vol->verts.push_back(p[k] + vol->origin[k]);
                    if (vol->color_type != 0) {
                        const double s = vol->color_type == 1 ? 255.0 : 1.0;
                        for (int k = 0; k < 3; ++k) {
                            const double c0 = vol->color[i0 * 3 + k], c1 = vol->color[i1 * 3 + k];
                            vol->vcols.push_back((c0 * (double)r1 + c1 * (double)r0) / (double)(r0 + r1) / s);
                        }
                    }
                    // GetNormalAt(p): central differences with half_gap = 0.99 * voxel_length
                    double n[3];
                    const double half_gap = 0.99 * vl;
                    for (int k = 0; k < 3; ++k) {
                        double pp[3] = {p[0], p[1], p[2]};
                        pp[k] -= half_gap;
                        const double t0 = orc_tsdf_at(vol, pp);
                        pp[k] += 2 * half_gap;
                        const double t1 = orc_tsdf_at(vol, pp);
                        n[k] = t1 - t0;
                    }
                    const double nn = std::sqrt(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
                    for (int k = 0; k < 3; ++k) vol->normals.push_back(nn > 0 ? n[k] / nn : n[k]);
                }
                (void)idx0;
            }
        }
    }
    *n_points = (int64_t)(vol->verts.size() / 3);
    return 0;
}

const double* orc_volume_vertices(void* h) { return ((OrcVolume*)h)->verts.data(); }
const double* orc_volume_vertex_colors(void* h) { return ((OrcVolume*)h)->vcols.data(); }
const double* orc_volume_normals(void* h) { return ((OrcVolume*)h)->normals.data(); }
const int32_t* orc_volume_triangles(void* h) { return ((OrcVolume*)h)->tris.data(); }

}  // extern "C"
