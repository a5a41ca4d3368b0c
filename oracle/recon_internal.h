// ORACLE -- TEST INFRASTRUCTURE ONLY.  Internal declarations shared by recon_oracle.cpp (uniform volume) and
// scalable_oracle.cpp (hash of 16^3 units built from the same per-voxel routine).
#pragma once
#include <cstdint>
#include <vector>

#include "mc_tables.inc"

struct OrcVolume {
    double length, voxel_length, sdf_trunc, origin[3];
    int R, color_type;  // 0 NoColor, 1 RGB8, 2 Gray32
    int x_begin, x_end;
    std::vector<float> tsdf, weight;
    std::vector<double> color;  // 3 per voxel (RGB8) / 1 used (Gray32)
    // extraction results
    std::vector<double> verts, vcols, normals;
    std::vector<int32_t> tris;
    inline int64_t idx(int x, int y, int z) const {
        return ((int64_t)(x - x_begin) * R + y) * R + z;
    }
};

struct IntegrateFrame {
    const float* depth; const void* color; const float* mult; int H, W;
    float fx, fy, cx, cy, E[4][4], Es[4][4], vl, hvl, trunc, trunc_inv, safe_w, safe_h;
};

extern "C" {
void* orc_volume_create(double length, int resolution, double sdf_trunc, int color_type, const double* origin, int x_begin,
                        int x_end);
void orc_distance_multiplier(int H, int W, double fx, double fy, double cx, double cy, float* out);
void make_frame(IntegrateFrame& F, const OrcVolume* vol, const float* depth, const void* color, const float* mult, int H,
                int W, double fx_d, double fy_d, double cx_d, double cy_d, const double* extrinsic);
int64_t integrate_x(OrcVolume* vol, const IntegrateFrame& F, int x);
}
