#pragma once
#include "common.cuh"

namespace b2r {

// slots inside ctx->small (uint32 index for 32-bit values, uint64 index for 64-bit values; the two
// index spaces are kept apart: 32-bit slots live in bytes [0,1024), 64-bit slots in [1024,4096)).
constexpr int SMALL_DMAX = 0;          // uint32: bits of the frame's max depth
constexpr int SMALL_ITEMS = 8;         // uint32: integrate work items planned for the frame
constexpr int SMALL_NEXT = 9;          // uint32: next work item to hand out
constexpr int SMALL_LIVE_TILES = 11;   // uint32: tiles that survived integrate_cull_kernel
constexpr int SMALL_TICKET = 10;       // uint32: warps of the streaming kernel that have finished (result write-back)
constexpr int SMALL_TOTAL64 = 128;     // uint64 index (byte 1024): scan totals
constexpr int SMALL_TOTAL64_B = 129;
constexpr int SMALL_TOTAL64_C = 130;
constexpr int SMALL_MC_NLIVE = 131;     // uint64: live virtual blocks of the last b2r_extract_mesh_count
constexpr int SMALL_PC_NLIVE = 132;     // uint64: live blocks of the last b2r_extract_points_count
constexpr int SMALL_PC_NCAND = 137;     // uint64: count blocks next to a marked row (b2r_extract_points_count's work list)
constexpr int SMALL_MC_NCELLS = 136;    // uint64: cells of the last b2r_extract_mesh_count's classify list

constexpr int B2R_MIP_LEVELS = 5;   // max-depth pyramid cells: 8, 16, 32, 64, 128 pixels
struct MipDesc {
    int w[B2R_MIP_LEVELS], h[B2R_MIP_LEVELS], off[B2R_MIP_LEVELS], total;
};
MipDesc mip_desc(int H, int W);

// Fills ctx->frame_mip (max-depth pyramid) and ctx->frame_dm (float4 {depth_m, multiplier, colour bits, 0} per pixel) for one frame.
// color_mode: 0 none, 1 uint8 x3, 2 float32 gray.
int prepare_frame(b2r_ctx* ctx, b2r_frame_slot* slot, const void* d_depth, int dtype, int convert, double depth_scale, double depth_trunc,
                  const void* d_color, int color_mode, const b2r_intrinsic* K, cudaStream_t stream);

void invert4(const double* m, double* out);

}  // namespace b2r
