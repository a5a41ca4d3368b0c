// ORACLE -- TEST INFRASTRUCTURE ONLY.  Not part of the product path.
//
// CPU restatement of Open3D 0.18.0's ScalableTSDFVolume (SURVEY.md A.3 last bullet, row f4): a hash of
// volume_unit_resolution^3 UniformTSDFVolume units keyed by the integer unit index.  Restated from memory of
// upstream cpp/open3d/pipelines/integration/ScalableTSDFVolume.cpp -- *** PARITY UNPINNED *** like the rest of
// oracle/ (Open3D is an un-vendored wheel, environment.yml:72-73; the reference holds no tests for it).
//
//   Integrate:  points = CreateFromDepthImage(depth, intrinsic, extrinsic, stride = depth_sampling_stride);
//               for every point, every unit between LocateVolumeUnit(p - trunc) and LocateVolumeUnit(p + trunc)
//               that this frame has not touched yet is opened (if new) and integrated with the uniform
//               per-voxel routine (origin = index * unit_length; the float32 z-chain restarts in every unit).
//   ExtractTriangleMesh / ExtractPointCloud: per unit, per voxel, neighbours looked up across units, a missing
//               unit reads as weight 0 / tsdf 0; vertex positions from GLOBAL voxel indices (no origin).
//   GetTSDFAt:  trilinear over 8 voxels around p; returns 0 when the unit holding the base voxel is missing.
//
// This file is compiled into liborc.so together with recon_oracle.cpp (it reuses OrcVolume / integrate_x through
// the small internal header below).
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <memory>
#include <unordered_map>
#include <unordered_set>
#include <vector>

#include "recon_internal.h"

namespace {

struct UKey {
    int x, y, z;
    bool operator==(const UKey& o) const { return x == o.x && y == o.y && z == o.z; }
};
struct UKeyHash {
    size_t operator()(const UKey& k) const {
        uint64_t h = (uint64_t)(uint32_t)k.x * 0x9E3779B97F4A7C15ull;
        h ^= (uint64_t)(uint32_t)k.y * 0xC2B2AE3D27D4EB4Full + (h << 6) + (h >> 2);
        h ^= (uint64_t)(uint32_t)k.z * 0x165667B19E3779F9ull + (h << 6) + (h >> 2);
        return (size_t)h;
    }
};

struct OrcScalable {
    double voxel_length, sdf_trunc, unit_length;
    int color_type, res, stride;
    std::unordered_map<UKey, std::unique_ptr<OrcVolume>, UKeyHash> units;
    std::vector<UKey> order;   // opening order (only to make the export deterministic)
    std::vector<double> verts, vcols, normals;
    std::vector<int32_t> tris;
    // optional bounding cube in unit indices [blo, bhi): units outside are not opened (counted in `clipped`).  Open3D
    // has no such bound; it is what the product's bounded dense backing cube does (b2r_integrate_units counts the same
    // touches), so that a CPU / GPU throughput pair compares the same work.
    bool bounded = false;
    int blo[3] = {0, 0, 0}, bhi[3] = {0, 0, 0};
    int64_t clipped = 0;

    const OrcVolume* find(const UKey& k) const {
        auto it = units.find(k);
        return it == units.end() ? nullptr : it->second.get();
    }
    OrcVolume* open(const UKey& k) {
        auto it = units.find(k);
        if (it != units.end()) return it->second.get();
        const double origin[3] = {(double)k.x * unit_length, (double)k.y * unit_length, (double)k.z * unit_length};
        OrcVolume* v = (OrcVolume*)orc_volume_create(unit_length, res, sdf_trunc, color_type, origin, 0, res);
        units.emplace(k, std::unique_ptr<OrcVolume>(v));
        order.push_back(k);
        return v;
    }
    // voxel (idx may run one past the unit in any axis) -> weight / tsdf / colour, missing unit = zeros
    bool voxel(const UKey& u, int x, int y, int z, float& w, float& f, const double** c) const {
        UKey k = u;
        if (x >= res) { x -= res; k.x += 1; }
        if (y >= res) { y -= res; k.y += 1; }
        if (z >= res) { z -= res; k.z += 1; }
        const OrcVolume* v = find(k);
        if (!v) { w = 0.0f; f = 0.0f; if (c) *c = nullptr; return false; }
        const int64_t i = v->idx(x, y, z);
        w = v->weight[i]; f = v->tsdf[i];
        if (c) *c = v->color.empty() ? nullptr : &v->color[i * 3];
        return true;
    }
    int locate(double p) const { return (int)std::floor(p / unit_length); }

    double tsdf_at(const double* p) const {
        double pl[3]; UKey k0; int i0[3]; double r[3];
        for (int a = 0; a < 3; ++a) pl[a] = p[a] - 0.5 * voxel_length;
        k0 = UKey{locate(pl[0]), locate(pl[1]), locate(pl[2])};
        if (!find(k0)) return 0.0;
        const int kk[3] = {k0.x, k0.y, k0.z};
        for (int a = 0; a < 3; ++a) {
            const double g = (pl[a] - (double)kk[a] * unit_length) / voxel_length;
            i0[a] = (int)std::floor(g);
            if (i0[a] < 0) i0[a] = 0;
            if (i0[a] >= res) i0[a] = res - 1;
            r[a] = g - i0[a];
        }
        auto T = [&](int dx, int dy, int dz) -> double {
            float w, f;
            voxel(k0, i0[0] + dx, i0[1] + dy, i0[2] + dz, w, f, nullptr);
            return (double)f;
        };
        return (1 - r[0]) * ((1 - r[1]) * ((1 - r[2]) * T(0, 0, 0) + r[2] * T(0, 0, 1)) +
                             r[1] * ((1 - r[2]) * T(0, 1, 0) + r[2] * T(0, 1, 1))) +
               r[0] * ((1 - r[1]) * ((1 - r[2]) * T(1, 0, 0) + r[2] * T(1, 0, 1)) +
                       r[1] * ((1 - r[2]) * T(1, 1, 0) + r[2] * T(1, 1, 1)));
    }
};

struct EKey {
    int x, y, z, a;
    bool operator==(const EKey& o) const { return x == o.x && y == o.y && z == o.z && a == o.a; }
};
struct EKeyHash {
    size_t operator()(const EKey& k) const {
        return UKeyHash()(UKey{k.x, k.y, k.z}) * 4u + (size_t)k.a;
    }
};

}  // namespace

extern "C" {

void* orc_scalable_create(double voxel_length, double sdf_trunc, int color_type, int unit_resolution, int stride) {
    OrcScalable* s = new OrcScalable();
    s->voxel_length = voxel_length; s->sdf_trunc = sdf_trunc; s->color_type = color_type;
    s->res = unit_resolution; s->stride = stride;
    s->unit_length = voxel_length * unit_resolution;
    return s;
}
void orc_scalable_destroy(void* h) { delete (OrcScalable*)h; }
void orc_scalable_reset(void* h) { OrcScalable* s = (OrcScalable*)h; s->units.clear(); s->order.clear(); s->clipped = 0; }
void orc_scalable_set_bounds(void* h, const int32_t* lo, const int32_t* hi) {
    OrcScalable* s = (OrcScalable*)h;
    s->bounded = lo && hi;
    for (int a = 0; a < 3 && s->bounded; ++a) { s->blo[a] = lo[a]; s->bhi[a] = hi[a]; }
}
int64_t orc_scalable_clipped(void* h) { return ((OrcScalable*)h)->clipped; }
int64_t orc_scalable_num_units(void* h) { return (int64_t)((OrcScalable*)h)->order.size(); }
void orc_scalable_unit_keys(void* h, int32_t* out) {
    OrcScalable* s = (OrcScalable*)h;
    for (size_t i = 0; i < s->order.size(); ++i) { out[i * 3] = s->order[i].x; out[i * 3 + 1] = s->order[i].y; out[i * 3 + 2] = s->order[i].z; }
}
// unit contents in Open3D's order [x][y][z]; colour 3 doubles per voxel (NULL to skip)
int orc_scalable_unit_arrays(void* h, const int32_t* key, float* tsdf, float* weight, double* color) {
    OrcScalable* s = (OrcScalable*)h;
    const OrcVolume* v = s->find(UKey{key[0], key[1], key[2]});
    if (!v) return -1;
    std::memcpy(tsdf, v->tsdf.data(), v->tsdf.size() * sizeof(float));
    std::memcpy(weight, v->weight.data(), v->weight.size() * sizeof(float));
    if (color && !v->color.empty()) std::memcpy(color, v->color.data(), v->color.size() * sizeof(double));
    return 0;
}

// ScalableTSDFVolume::Integrate.  camera_pose = extrinsic^-1 is passed in (the caller inverts once, so that the
// oracle and the library under test see the same float64 matrix).  Returns voxel updates, <0 on error;
// *touched = units integrated by this frame.
int64_t orc_scalable_integrate(void* h, const float* depth, const void* color, int H, int W, double fx, double fy,
                               double cx, double cy, const double* extrinsic, const double* pose, int64_t* touched) {
    OrcScalable* s = (OrcScalable*)h;
    if (s->color_type != 0 && !color) return -2;
    std::vector<float> mult((size_t)H * W);
    orc_distance_multiplier(H, W, fx, fy, cx, cy, mult.data());
    std::unordered_set<UKey, UKeyHash> seen;
    std::vector<OrcVolume*> work;
    for (int i = 0; i < H; i += s->stride) {
        for (int j = 0; j < W; j += s->stride) {
            const float d = depth[(int64_t)i * W + j];
            if (!(d > 0)) continue;
            const double z = (double)d;
            const double x = (j - cx) * z / fx;
            const double y = (i - cy) * z / fy;
            double p[3];
            for (int r = 0; r < 3; ++r)
                p[r] = ((pose[r * 4 + 0] * x + pose[r * 4 + 1] * y) + pose[r * 4 + 2] * z) + pose[r * 4 + 3];
            int lo[3], hi[3];
            for (int a = 0; a < 3; ++a) { lo[a] = s->locate(p[a] - s->sdf_trunc); hi[a] = s->locate(p[a] + s->sdf_trunc); }
            for (int ux = lo[0]; ux <= hi[0]; ++ux)
                for (int uy = lo[1]; uy <= hi[1]; ++uy)
                    for (int uz = lo[2]; uz <= hi[2]; ++uz) {
                        const UKey k{ux, uy, uz};
                        if (s->bounded && (ux < s->blo[0] || ux >= s->bhi[0] || uy < s->blo[1] || uy >= s->bhi[1] ||
                                           uz < s->blo[2] || uz >= s->bhi[2])) {
                            ++s->clipped;
                            continue;
                        }
                        if (seen.insert(k).second) work.push_back(s->open(k));
                    }
        }
    }
    if (touched) *touched = (int64_t)work.size();
    if (work.empty()) return 0;
    IntegrateFrame F;
    make_frame(F, work[0], depth, color, mult.data(), H, W, fx, fy, cx, cy, extrinsic);
    int64_t updates = 0;
    const int64_t n = (int64_t)work.size();
#pragma omp parallel for schedule(dynamic, 8) reduction(+ : updates)
    for (int64_t u = 0; u < n; ++u)
        for (int x = 0; x < s->res; ++x) updates += integrate_x(work[u], F, x);
    return updates;
}

int orc_scalable_extract_mesh(void* h, int64_t* n_vertices, int64_t* n_triangles) {
    OrcScalable* s = (OrcScalable*)h;
    const int res = s->res;
    const double vl = s->voxel_length, hvl = 0.5 * vl;
    s->verts.clear(); s->vcols.clear(); s->tris.clear();
    std::unordered_map<EKey, int, EKeyHash> edge_to_index;
    int e2v[12];
    for (const UKey& u : s->order) {
        for (int x = 0; x < res; ++x)
            for (int y = 0; y < res; ++y)
                for (int z = 0; z < res; ++z) {
                    int cube_index = 0;
                    float f[8];
                    double c[8][3];
                    for (int i = 0; i < 8; ++i) {
                        float w; const double* col;
                        s->voxel(u, x + mc_shift[i][0], y + mc_shift[i][1], z + mc_shift[i][2], w, f[i], &col);
                        if (w == 0.0f) { cube_index = 0; break; }
                        if (f[i] < 0.0f) cube_index |= (1 << i);
                        if (s->color_type == 1) for (int k = 0; k < 3; ++k) c[i][k] = col[k] / 255.0;
                        else if (s->color_type == 2) for (int k = 0; k < 3; ++k) c[i][k] = col[k];
                    }
                    if (cube_index == 0 || cube_index == 255) continue;
                    for (int i = 0; i < 12; ++i) {
                        if (!(mc_edge_table[cube_index] & (1 << i))) continue;
                        EKey key{u.x * res + x + mc_edge_shift[i][0], u.y * res + y + mc_edge_shift[i][1],
                                 u.z * res + z + mc_edge_shift[i][2], mc_edge_shift[i][3]};
                        auto it = edge_to_index.find(key);
                        if (it != edge_to_index.end()) { e2v[i] = it->second; continue; }
                        e2v[i] = (int)(s->verts.size() / 3);
                        edge_to_index.emplace(key, e2v[i]);
                        double pt[3] = {hvl + vl * key.x, hvl + vl * key.y, hvl + vl * key.z};
                        const double f0 = std::fabs((double)f[mc_edge_to_vert[i][0]]);
                        const double f1 = std::fabs((double)f[mc_edge_to_vert[i][1]]);
                        pt[key.a] += f0 * vl / (f0 + f1);
                        for (int k = 0; k < 3; ++k) s->verts.push_back(pt[k]);
                        if (s->color_type != 0) {
                            const double* c0 = c[mc_edge_to_vert[i][0]];
                            const double* c1 = c[mc_edge_to_vert[i][1]];
                            for (int k = 0; k < 3; ++k) s->vcols.push_back((f1 * c0[k] + f0 * c1[k]) / (f0 + f1));
                        }
                    }
                    for (int i = 0; mc_tri_table[cube_index][i] != -1; i += 3) {
                        s->tris.push_back(e2v[mc_tri_table[cube_index][i]]);
                        s->tris.push_back(e2v[mc_tri_table[cube_index][i + 2]]);
                        s->tris.push_back(e2v[mc_tri_table[cube_index][i + 1]]);
                    }
                }
    }
    *n_vertices = (int64_t)(s->verts.size() / 3);
    *n_triangles = (int64_t)(s->tris.size() / 3);
    return 0;
}

int orc_scalable_extract_points(void* h, int64_t* n_points) {
    OrcScalable* s = (OrcScalable*)h;
    const int res = s->res;
    const double vl = s->voxel_length, hvl = 0.5 * vl;
    s->verts.clear(); s->vcols.clear(); s->normals.clear();
    for (const UKey& u : s->order) {
        const OrcVolume* v0 = s->find(u);
        const double uo[3] = {(double)u.x * s->unit_length, (double)u.y * s->unit_length, (double)u.z * s->unit_length};
        for (int x = 0; x < res; ++x)
            for (int y = 0; y < res; ++y)
                for (int z = 0; z < res; ++z) {
                    const int64_t i0 = v0->idx(x, y, z);
                    const float w0 = v0->weight[i0], f0 = v0->tsdf[i0];
                    if (!(w0 != 0.0f && f0 < 0.98f && f0 >= -0.98f)) continue;
                    const double* c0 = v0->color.empty() ? nullptr : &v0->color[i0 * 3];
                    const double p0[3] = {hvl + vl * x + uo[0], hvl + vl * y + uo[1], hvl + vl * z + uo[2]};
                    for (int a = 0; a < 3; ++a) {
                        int i1[3] = {x, y, z};
                        i1[a] += 1;
                        float w1, f1; const double* c1;
                        s->voxel(u, i1[0], i1[1], i1[2], w1, f1, &c1);
                        if (!(w1 != 0.0f && f1 < 0.98f && f1 >= -0.98f && f0 * f1 < 0)) continue;
                        const float r0 = std::fabs(f0), r1 = std::fabs(f1);
                        double p1[3] = {p0[0], p0[1], p0[2]};
                        p1[a] += vl;
                        double p[3] = {p0[0], p0[1], p0[2]};
                        p[a] = (p0[a] * (double)r1 + p1[a] * (double)r0) / (double)(r0 + r1);
                        for (int k = 0; k < 3; ++k) s->verts.push_back(p[k]);
                        if (s->color_type != 0) {
                            const double sc = s->color_type == 1 ? 255.0 : 1.0;
                            for (int k = 0; k < 3; ++k)
                                s->vcols.push_back((c0[k] * (double)r1 + c1[k] * (double)r0) / (double)(r0 + r1) / sc);
                        }
                        double n[3];
                        const double half_gap = 0.99 * vl;
                        for (int k = 0; k < 3; ++k) {
                            double pp[3] = {p[0], p[1], p[2]};
                            pp[k] -= half_gap;
                            const double t0 = s->tsdf_at(pp);
                            pp[k] += 2 * half_gap;
                            const double t1 = s->tsdf_at(pp);
                            n[k] = t1 - t0;
                        }
                        const double nn = std::sqrt(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
                        for (int k = 0; k < 3; ++k) s->normals.push_back(nn > 0 ? n[k] / nn : n[k]);
                    }
                }
    }
    *n_points = (int64_t)(s->verts.size() / 3);
    return 0;
}

const double* orc_scalable_vertices(void* h) { return ((OrcScalable*)h)->verts.data(); }
const double* orc_scalable_vertex_colors(void* h) { return ((OrcScalable*)h)->vcols.data(); }
const double* orc_scalable_normals(void* h) { return ((OrcScalable*)h)->normals.data(); }
const int32_t* orc_scalable_triangles(void* h) { return ((OrcScalable*)h)->tris.data(); }

}  // extern "C"
