// Marching cubes with welded vertices -- replaces Open3D UniformTSDFVolume::ExtractTriangleMesh
// (north_star API extract_triangle_mesh(); arithmetic in SURVEY.md A.4, tables Appendix B).
//
// Open3D runs a single-threaded x,y,z scan with an unordered_map edge -> vertex.  Here:
//   mc_classify  : thread per (x,y) column chunk marching z, 4 corner loads per cube (the lower
//                  plane's sign/valid nibble is carried in registers); writes one byte per cube:
//                  the cube index if the cube is ACTIVE (all 8 weights != 0, index not 0/255) else 0.
//   mc_count     : thread per voxel; each voxel OWNS its +x,+y,+z edges.  An edge carries a vertex
//                  iff one of its <=4 adjacent cubes is active and has a sign change on it (Open3D
//                  creates vertices lazily, from active cubes only).  Per-block vertex / triangle
//                  totals come from a warp-ballot/popc + shared-memory reduction.
//   scan         : exclusive prefix sums of the block totals (scan.cu).
//   mc_emit_vertices / mc_emit_triangles : warp-ballot + block prefix-sum compaction; vertex ids
//                  are published in a per-voxel table (id | flags<<29) so that triangles -- which
//                  reference edges owned by up to 7 neighbouring voxels -- are welded by lookup.
// Vertex positions use Open3D's exact double expression, so they are bit-identical; the ORDER of
// vertices/triangles is memory order instead of first-touch scan order.
//
// Slabs: a slab owns voxels (and cubes) of the x planes [x_begin, x_end) (memory is x-slowest: plane, then z, then y);
// it reads halo planes x_begin-1 and x_end (B2R_PAD_*), and the vertex-id table of plane x_end comes from the slab above
// through b2r_mesh_seam_buffers (one extra R*R*4-byte message).
#include "frame.cuh"
#include "scan.cuh"
#include "mc_tables.inc"

namespace b2r {

#ifndef B2R_MC_ZCHUNK
#define B2R_MC_ZCHUNK 8         // measured at 1024^3 after the bench orbit: 32 -> 288 us, 16 -> 224 us, 8 -> 200 us (classify)
#endif
constexpr int MC_ZCHUNK = B2R_MC_ZCHUNK;
constexpr int MC_THREADS = 256;

struct McGeom {
    int R, x_begin, x_end;
    double vl, ox, oy, oz;
    int color_type;
};

__device__ __forceinline__ int64_t vidx(int R, int x_begin, int x, int y, int z) {
    return ((int64_t)(x - x_begin + B2R_PAD_LO) * R + z) * R + y;
}

// sign/valid nibble of the 4 corners (x,y),(x+1,y),(x+1,y+1),(x,y+1) of plane z
__device__ __forceinline__ unsigned plane_nibble(const float2* __restrict__ tsdfw, int R, int x_begin, int x, int y,
                                                 int z, bool& valid) {
    const int64_t i = vidx(R, x_begin, x, y, z);
    const int64_t sx = (int64_t)R * R;                      // x + 1
    const float2 a = __ldg(&tsdfw[i]), d = __ldg(&tsdfw[i + 1]);
    const float2 b = __ldg(&tsdfw[i + sx]), c = __ldg(&tsdfw[i + sx + 1]);
    valid = (a.y != 0.0f) && (b.y != 0.0f) && (c.y != 0.0f) && (d.y != 0.0f);
    return (a.x < 0.0f ? 1u : 0u) | (b.x < 0.0f ? 2u : 0u) | (c.x < 0.0f ? 4u : 0u) | (d.x < 0.0f ? 8u : 0u);
}

// Classification runs on CELLS of 4 cube x planes x MC_ZCHUNK cube z steps (all y).
// Output sensitivity (b2r_volume::d_rows): an active cube has a corner with tsdf < 0, so a cube of row (x, z) can only be
// active when one of the rows (x, z), (x+1, z), (x, z+1), (x+1, z+1) is marked "may hold a negative value".  A cell none of
// whose (x..x+4, z0..z1) rows is marked has no active cube: it is never classified and its codes are never written -- the
// later passes read a cube's code only when row_active says its row holds an active cube (code_at), and such a row's
// cell was classified.  mc_cells_kernel, a warp per cell, looks at the cell's <= 165 marks and appends the live cells to
// a list (order irrelevant: every output is addressed by position); without marks every cell is live.
__global__ void __launch_bounds__(256) mc_cells_kernel(const uint8_t* __restrict__ rows, int R, int x_begin, int xc_lo, int xc_hi,
                                                       int cells_x, int cells_z, uint32_t* __restrict__ cells,
                                                       unsigned long long* __restrict__ n_cells) {
    const int cell = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (cell >= cells_x * cells_z) return;
    const int cx = cell / cells_z, cz = cell - cx * cells_z;
    const int xb0 = xc_lo + cx * 4, z0 = cz * MC_ZCHUNK, z1 = min(z0 + MC_ZCHUNK, R - 1);
    bool m = rows == nullptr;
    if (rows)
        for (int t = lane; t < 5 * (MC_ZCHUNK + 1); t += 32) {
            const int xx = xb0 + t / (MC_ZCHUNK + 1), zz = z0 + t % (MC_ZCHUNK + 1);
            m |= xx <= xc_hi + 1 && zz <= z1 && __ldg(&rows[(int64_t)(xx - x_begin + B2R_PAD_LO) * R + zz]) != 0;
        }
    if (__any_sync(0xffffffffu, m) && lane == 0) cells[atomicAdd(n_cells, 1ull)] = (uint32_t)cell;
}

// Persistent blocks of (64, 4) threads over (live cell, 64-column y tile) pairs: threadIdx.x -> y, threadIdx.y -> x.  A
// thread marches its (x, y) column along z (the cube index's low / high nibble are the z and z+1 corners, so the carry
// runs along z whatever the memory order; consecutive z are R voxels apart).
__global__ void __launch_bounds__(256) mc_classify_kernel(const float2* __restrict__ tsdfw, uint8_t* __restrict__ code,
                                                          uint8_t* __restrict__ row_active, int R, int x_begin, int xc_lo,
                                                          int xc_hi, int cells_z, int tiles_y,
                                                          const uint32_t* __restrict__ cells,
                                                          const unsigned long long* __restrict__ n_cells) {
    const int64_t n_work = (int64_t)*n_cells * tiles_y;
    for (int64_t wk = blockIdx.x; wk < n_work; wk += gridDim.x) {
        const int cell = (int)cells[wk / tiles_y], ty = (int)(wk % tiles_y);
        const int cx = cell / cells_z, cz = cell - cx * cells_z;
        const int y = ty * 64 + threadIdx.x;
        const int x = xc_lo + cx * 4 + threadIdx.y;
        const int z0 = cz * MC_ZCHUNK;
        const int z1 = min(z0 + MC_ZCHUNK, R - 1);  // cubes z0 .. z1-1 (no cube starts at z = R-1)
        if (x > xc_hi || y >= R || z0 >= z1) continue;
        if (y >= R - 1) {  // no cube starts on the last column
            for (int z = z0; z < z1; ++z) code[vidx(R, x_begin, x, y, z)] = 0;
            continue;
        }
        bool v_lo;
        unsigned n_lo = plane_nibble(tsdfw, R, x_begin, x, y, z0, v_lo);
        for (int z = z0; z < z1; ++z) {   // (#pragma unroll 4 was measured: 3.18 ms instead of 2.47 ms at 1024^3)
            bool v_hi;
            const unsigned n_hi = plane_nibble(tsdfw, R, x_begin, x, y, z + 1, v_hi);
            const unsigned c = n_lo | (n_hi << 4);
            const bool active = v_lo && v_hi && c != 0u && c != 255u;
            code[vidx(R, x_begin, x, y, z)] = active ? (uint8_t)c : (uint8_t)0;
            if (active) row_active[(int64_t)(x - x_begin + B2R_PAD_LO) * R + z] = 1;   // (x, z) row holds an active cube
            n_lo = n_hi; v_lo = v_hi;
        }
    }
}

// Code of cube (x, y, z); `on` = row_active of the cube's (x, z) row.  Rows without an active cube may never have been
// classified (mc_classify_kernel leaves unmarked blocks alone), so their codes are not read: they are 0 by definition.
__device__ __forceinline__ unsigned code_at(const uint8_t* __restrict__ code, int R, int x_begin, bool on, int x, int y, int z) {
    if (!on || y < 0) return 0u;
    return (unsigned)__ldg(&code[vidx(R, x_begin, x, y, z)]);
}
// row_active of cube row (x, z), 0 outside the cubes this slab classified
__device__ __forceinline__ bool row_on(const uint8_t* __restrict__ row_active, int R, int x_begin, int xc_lo, int xc_hi, int x, int z) {
    if (z < 0 || z > R - 2 || x < xc_lo || x > xc_hi) return false;
    return __ldg(&row_active[(int64_t)(x - x_begin + B2R_PAD_LO) * R + z]) != 0;
}
__device__ __forceinline__ unsigned xbit(unsigned c, int a, int b) { return ((c >> a) ^ (c >> b)) & 1u; }

// Owned-edge vertex flags (bit0 +x, bit1 +y, bit2 +z) of voxel (x,y,z).
__device__ __forceinline__ unsigned edge_flags(const uint8_t* __restrict__ code, const uint8_t* __restrict__ row_active, int R,
                                               int x_begin, int xc_lo, int xc_hi, int x, int y, int z, unsigned& own_code) {
    const bool r00 = row_on(row_active, R, x_begin, xc_lo, xc_hi, x, z), r0m = row_on(row_active, R, x_begin, xc_lo, xc_hi, x, z - 1);
    const bool rm0 = row_on(row_active, R, x_begin, xc_lo, xc_hi, x - 1, z), rmm = row_on(row_active, R, x_begin, xc_lo, xc_hi, x - 1, z - 1);
    const unsigned c000 = code_at(code, R, x_begin, r00, x, y, z);
    const unsigned c0m0 = code_at(code, R, x_begin, r00, x, y - 1, z);
    const unsigned c00m = code_at(code, R, x_begin, r0m, x, y, z - 1);
    const unsigned c0mm = code_at(code, R, x_begin, r0m, x, y - 1, z - 1);
    const unsigned cm00 = code_at(code, R, x_begin, rm0, x - 1, y, z);
    const unsigned cm0m = code_at(code, R, x_begin, rmm, x - 1, y, z - 1);
    const unsigned cmm0 = code_at(code, R, x_begin, rm0, x - 1, y - 1, z);
    own_code = c000;
    const unsigned fx = xbit(c000, 0, 1) | xbit(c0m0, 3, 2) | xbit(c00m, 4, 5) | xbit(c0mm, 7, 6);
    const unsigned fy = xbit(c000, 0, 3) | xbit(cm00, 1, 2) | xbit(c00m, 4, 7) | xbit(cm0m, 5, 6);
    const unsigned fz = xbit(c000, 0, 4) | xbit(cm00, 1, 5) | xbit(cmm0, 2, 6) | xbit(c0m0, 3, 7);
    return fx | (fy << 1) | (fz << 2);
}

// Output sensitivity: the surface touches a tiny fraction of the cube, so the per-voxel passes below run only on
// the "virtual blocks" (256 consecutive voxels = part of one or several (x, z) rows) that can produce anything.
// A voxel's owned edges / triangles depend on cubes of rows (x, z), (x-1, z), (x, z-1), (x-1, z-1) only; a
// block none of whose rows has such an active cube is dead: its vert_tab entries stay unwritten (never read) and
// its totals stay zero.  mc_live marks the live blocks, a scan + compaction turns the marks into a list, and the
// passes are persistent kernels walking that list (launching a block per 256 voxels of a 1024^3 cube costs more
// than the work of the live ones).
__global__ void __launch_bounds__(256) mc_live_kernel(const uint8_t* __restrict__ row_active, int R, int64_t n, int nblocks,
                                                      uint32_t* __restrict__ live) {
    const int vb = blockIdx.x * blockDim.x + threadIdx.x;
    if (vb >= nblocks) return;
    const int64_t t0 = (int64_t)vb * MC_THREADS;
    const int64_t t1 = min(t0 + MC_THREADS, n) - 1;
    unsigned any = 0;
    for (int64_t r = t0 / R; r <= t1 / R; ++r) {
        const int64_t zr = r / R;
        const int x = (int)(r - zr * R);
        const int64_t a = (zr + B2R_PAD_LO) * R + x;     // padded plane index >= 1: the plane below always exists
        any |= row_active[a] | row_active[a - R];
        if (x > 0) any |= row_active[a - 1] | row_active[a - R - 1];
    }
    live[vb] = any ? 1u : 0u;
}
__global__ void __launch_bounds__(256) mc_live_list_kernel(const uint32_t* __restrict__ live, const uint32_t* __restrict__ pos,
                                                           int nblocks, uint32_t* __restrict__ list) {
    const int vb = blockIdx.x * blockDim.x + threadIdx.x;
    if (vb < nblocks && live[vb]) list[pos[vb]] = (uint32_t)vb;
}

// flat thread per owned voxel (memory order).  Writes vert_tab[voxel] = flags << 29 (ids are added by
// mc_emit_vertices) and per-block totals.
__global__ void __launch_bounds__(MC_THREADS) mc_count_kernel(const uint8_t* __restrict__ code,
                                                              const uint8_t* __restrict__ row_active, int R, int x_begin,
                                                              int x_end, int xc_lo, int xc_hi,
                                                              const uint32_t* __restrict__ live_list,
                                                              const unsigned long long* __restrict__ n_live,
                                                              uint32_t* __restrict__ vert_tab,
                                                              uint32_t* __restrict__ blk_verts,
                                                              uint32_t* __restrict__ blk_tris) {
    __shared__ uint32_t s_v[MC_THREADS / 32], s_t[MC_THREADS / 32];
    const int64_t n = (int64_t)(x_end - x_begin) * R * R;
    const unsigned nl = (unsigned)*n_live;
    for (unsigned li = blockIdx.x; li < nl; li += gridDim.x) {
        const unsigned vb = live_list[li];
        const int64_t t = (int64_t)vb * MC_THREADS + threadIdx.x;
        unsigned nv = 0, nt = 0;
        if (t < n) {
            const int y = (int)(t % R), z = (int)((t / R) % R), x = x_begin + (int)(t / ((int64_t)R * R));
            unsigned own;
            const unsigned f = edge_flags(code, row_active, R, x_begin, xc_lo, xc_hi, x, y, z, own);
            nv = __popc(f);
            nt = d_mc_ntri_table[own];
            vert_tab[vidx(R, x_begin, x, y, z)] = f << 29;
        }
        // warp shuffle reduction, then one shared-memory pass per block
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            nv += __shfl_xor_sync(0xffffffffu, nv, o);
            nt += __shfl_xor_sync(0xffffffffu, nt, o);
        }
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        if (lane == 0) { s_v[warp] = nv; s_t[warp] = nt; }
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned a = 0, b = 0;
            for (int w = 0; w < MC_THREADS / 32; ++w) { a += s_v[w]; b += s_t[w]; }
            blk_verts[vb] = a;
            blk_tris[vb] = b;
        }
        __syncthreads();
    }
}

__device__ __forceinline__ void decode_color(uint32_t c, int color_type, double* out) {
    if (color_type == 1) {
        out[0] = (double)((float)(c & 1023u) * 0.25f) / 255.0;
        out[1] = (double)((float)((c >> 10) & 1023u) * 0.25f) / 255.0;
        out[2] = (double)((float)((c >> 20) & 1023u) * 0.25f) / 255.0;
    } else {
        const double g = (double)__uint_as_float(c);
        out[0] = g; out[1] = g; out[2] = g;
    }
}

__global__ void __launch_bounds__(MC_THREADS) mc_emit_vertices_kernel(const float2* __restrict__ tsdfw,
                                                                      const uint32_t* __restrict__ color,
                                                                      const __grid_constant__ McGeom G,
                                                                      const uint32_t* __restrict__ live_list,
                                                                      const unsigned long long* __restrict__ n_live,
                                                                      uint32_t* __restrict__ vert_tab,
                                                                      const uint32_t* __restrict__ blk_off,
                                                                      uint32_t vertex_base, double* __restrict__ verts,
                                                                      double* __restrict__ vcols) {
    __shared__ uint32_t smem[33];
    const int R = G.R;
    const int64_t n = (int64_t)(G.x_end - G.x_begin) * R * R;
    const unsigned nl = (unsigned)*n_live;
    for (unsigned li = blockIdx.x; li < nl; li += gridDim.x) {
        const unsigned vb = live_list[li];
        const int64_t t = (int64_t)vb * MC_THREADS + threadIdx.x;
        unsigned f = 0;
        int x = 0, y = 0, z = 0;
        int64_t vi = 0;
        if (t < n) {
            y = (int)(t % R); z = (int)((t / R) % R); x = G.x_begin + (int)(t / ((int64_t)R * R));
            vi = vidx(R, G.x_begin, x, y, z);
            f = vert_tab[vi] >> 29;
        }
        uint32_t total;
        const uint32_t local = block_exclusive_scan(__popc(f), smem, &total);
        if (total == 0 || f == 0) continue;
        const uint32_t first_local = blk_off[vb] + local;  // id within this slab
        vert_tab[vi] = (f << 29) | (vertex_base + first_local);
        const double hvl = 0.5 * G.vl;
        const float f0s = tsdfw[vi].x;
        double c0[3] = {0, 0, 0};
        if (vcols && G.color_type) decode_color(color[vi], G.color_type, c0);
        uint32_t k = first_local;
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            if (!(f & (1u << a))) continue;
            const int64_t ni = vi + (a == 0 ? (int64_t)R * R : a == 1 ? (int64_t)1 : (int64_t)R);      // +x, +y, +z
            const float f1s = tsdfw[ni].x;
            const double f0 = fabs((double)f0s), f1 = fabs((double)f1s);
            double pt[3] = {hvl + G.vl * (double)x, hvl + G.vl * (double)y, hvl + G.vl * (double)z};
            pt[a] += f0 * G.vl / (f0 + f1);
            verts[(int64_t)k * 3 + 0] = pt[0] + G.ox;
            verts[(int64_t)k * 3 + 1] = pt[1] + G.oy;
            verts[(int64_t)k * 3 + 2] = pt[2] + G.oz;
            if (vcols && G.color_type) {
                double c1[3];
                decode_color(color[ni], G.color_type, c1);
#pragma unroll
                for (int q = 0; q < 3; ++q) vcols[(int64_t)k * 3 + q] = (f1 * c0[q] + f0 * c1[q]) / (f0 + f1);
            }
            ++k;
        }
    }
}

__global__ void __launch_bounds__(MC_THREADS) mc_emit_triangles_kernel(const uint8_t* __restrict__ code,
                                                                       const uint8_t* __restrict__ row_active,
                                                                       const uint32_t* __restrict__ vert_tab, int R,
                                                                       int x_begin, int x_end, int xc_hi,
                                                                       const uint32_t* __restrict__ live_list,
                                                                       const unsigned long long* __restrict__ n_live,
                                                                       const uint32_t* __restrict__ blk_off,
                                                                       int32_t* __restrict__ tris) {
    __shared__ uint32_t smem[33];
    const int64_t n = (int64_t)(x_end - x_begin) * R * R;
    const unsigned nl = (unsigned)*n_live;
    for (unsigned li = blockIdx.x; li < nl; li += gridDim.x) {
        const unsigned vb = live_list[li];
        const int64_t t = (int64_t)vb * MC_THREADS + threadIdx.x;
        unsigned c = 0;
        int x = 0, y = 0, z = 0;
        if (t < n) {
            y = (int)(t % R); z = (int)((t / R) % R); x = x_begin + (int)(t / ((int64_t)R * R));
            if (x <= xc_hi && z <= R - 2 && row_active[(int64_t)(x - x_begin + B2R_PAD_LO) * R + z]) c = code[vidx(R, x_begin, x, y, z)];
        }
        const unsigned nt = d_mc_ntri_table[c];
        uint32_t total;
        const uint32_t local = block_exclusive_scan(nt, smem, &total);
        if (total == 0 || nt == 0) continue;
        int64_t o = ((int64_t)blk_off[vb] + local) * 3;
        int ids[12];
#pragma unroll
        for (int e = 0; e < 12; ++e) {
            ids[e] = -1;
            if (d_mc_edge_table[c] & (1 << e)) {
                const int ex = x + d_mc_edge_shift[e][0], ey = y + d_mc_edge_shift[e][1], ez = z + d_mc_edge_shift[e][2];
                const int axis = d_mc_edge_shift[e][3];
                const uint32_t ent = __ldg(&vert_tab[vidx(R, x_begin, ex, ey, ez)]);
                const unsigned fl = ent >> 29;
                ids[e] = (int)((ent & 0x1fffffffu) + __popc(fl & ((1u << axis) - 1u)));
            }
        }
        for (int i = 0; i < 15 && d_mc_tri_table[c][i] != -1; i += 3) {
            tris[o + 0] = ids[d_mc_tri_table[c][i]];
            tris[o + 1] = ids[d_mc_tri_table[c][i + 2]];
            tris[o + 2] = ids[d_mc_tri_table[c][i + 1]];
            o += 3;
        }
    }
}

static void cube_x_range(const b2r_volume* vol, int& xc_lo, int& xc_hi) {
    // cubes this slab must classify: its own (x in [x_begin, x_end) and x <= R-2) plus x_begin-1
    xc_lo = vol->x_begin > 0 ? vol->x_begin - 1 : 0;
    xc_hi = (vol->x_end - 1 < vol->resolution - 2) ? vol->x_end - 1 : vol->resolution - 2;
}

}  // namespace b2r

using namespace b2r;

extern "C" {

int b2r_extract_mesh_count(b2r_ctx* ctx, const b2r_volume* vol, int64_t* n_vertices, int64_t* n_triangles,
                           void* stream) {
    B2R_CHECK(ctx && vol && vol->d_tsdfw && n_vertices && n_triangles, -1, "b2r_extract_mesh_count: NULL argument");
    B2R_CUDA(cudaSetDevice(ctx->device));
    B2R_CUDA(join_side(ctx, S(stream)));   // items kernels of earlier integrate calls run on the side stream
    cudaStream_t st = S(stream);
    const int R = vol->resolution, zb = vol->x_begin, ze = vol->x_end, nz = ze - zb;
    *n_vertices = 0; *n_triangles = 0;
    ctx->mc_nv = ctx->mc_nt = 0;
    ctx->mc_R = R; ctx->mc_zb = zb; ctx->mc_ze = ze;
    ctx->mc_vol = vol->d_tsdfw; ctx->mc_gen = ctx->write_gen;
    if (nz <= 0 || R < 2) return 0;
    const size_t planes = (size_t)nz + B2R_PAD_LO + B2R_PAD_HI;
    const size_t nvox_pad = planes * R * R;
    if (ctx->mc_code.ensure(nvox_pad)) return -102;
    if (ctx->mc_vert_id.ensure(nvox_pad * sizeof(uint32_t))) return -102;
    const int64_t n = (int64_t)nz * R * R;
    const int nblocks = cdiv(n, MC_THREADS);
    if (ctx->mc_block_counts.ensure((size_t)nblocks * 2 * sizeof(uint32_t))) return -102;
    uint8_t* code = (uint8_t*)ctx->mc_code.ptr;
    uint32_t* vert_tab = (uint32_t*)ctx->mc_vert_id.ptr;
    uint32_t* blk_v = (uint32_t*)ctx->mc_block_counts.ptr;
    uint32_t* blk_t = blk_v + nblocks;
    int xc_lo, xc_hi;
    cube_x_range(vol, xc_lo, xc_hi);
    if (ctx->mc_flags.ensure(planes * R)) return -102;
    uint8_t* row_active = (uint8_t*)ctx->mc_flags.ptr;
    B2R_CUDA(cudaMemsetAsync(row_active, 0, planes * R, st));
    unsigned long long* tot = (unsigned long long*)ctx->small.ptr + SMALL_TOTAL64;
    if (ctx->mc_live.ensure((size_t)nblocks * 3 * sizeof(uint32_t))) return -102;
    if (xc_hi >= xc_lo) {
        // (the code array is not cleared: codes of rows without an active cube are never read)
        const int cells_x = cdiv(xc_hi - xc_lo + 1, 4), cells_z = cdiv(R - 1, MC_ZCHUNK), tiles_y = cdiv(R, 64);
        if (ctx->mc_cells.ensure((size_t)cells_x * cells_z * sizeof(uint32_t))) return -102;
        unsigned long long* n_cells = (unsigned long long*)ctx->small.ptr + SMALL_MC_NCELLS;
        B2R_CUDA(cudaMemsetAsync(n_cells, 0, sizeof(unsigned long long), st));
        mc_cells_kernel<<<cdiv(cells_x * cells_z, 8), 256, 0, st>>>((const uint8_t*)vol->d_rows, R, zb, xc_lo, xc_hi, cells_x, cells_z,
                                                                    (uint32_t*)ctx->mc_cells.ptr, n_cells);
        B2R_LAUNCH_CHECK(ctx);
        const int64_t max_work = (int64_t)cells_x * cells_z * tiles_y;
        const int grid = (int)(max_work < (int64_t)ctx->sm_count * 8 ? max_work : (int64_t)ctx->sm_count * 8);
        mc_classify_kernel<<<grid, dim3(64, 4), 0, st>>>((const float2*)vol->d_tsdfw, code, row_active, R, zb, xc_lo, xc_hi, cells_z,
                                                         tiles_y, (const uint32_t*)ctx->mc_cells.ptr, n_cells);
        B2R_LAUNCH_CHECK(ctx);
    }
    // live virtual blocks -> list (device-side count in tot[2]); dead blocks keep zero totals
    uint32_t* live = (uint32_t*)ctx->mc_live.ptr;
    uint32_t* live_pos = live + nblocks;
    uint32_t* live_list = live + 2 * (size_t)nblocks;
    mc_live_kernel<<<cdiv(nblocks, 256), 256, 0, st>>>(row_active, R, n, nblocks, live);
    B2R_LAUNCH_CHECK(ctx);
    if (int rc = scan_exclusive_u32(ctx, live, live_pos, nblocks, (unsigned long long*)ctx->small.ptr + SMALL_MC_NLIVE, st)) return rc;
    mc_live_list_kernel<<<cdiv(nblocks, 256), 256, 0, st>>>(live, live_pos, nblocks, live_list);
    B2R_LAUNCH_CHECK(ctx);
    B2R_CUDA(cudaMemsetAsync(blk_v, 0, (size_t)nblocks * 2 * sizeof(uint32_t), st));
    mc_count_kernel<<<ctx->sm_count * 8, MC_THREADS, 0, st>>>(code, row_active, R, zb, ze, xc_lo, xc_hi, live_list, (unsigned long long*)ctx->small.ptr + SMALL_MC_NLIVE, vert_tab, blk_v,
                                                              blk_t);
    B2R_LAUNCH_CHECK(ctx);
    if (int rc = scan_exclusive_u32(ctx, blk_v, blk_v, nblocks, tot, st)) return rc;
    if (int rc = scan_exclusive_u32(ctx, blk_t, blk_t, nblocks, tot + 1, st)) return rc;
    B2R_CUDA(cudaMemcpyAsync(ctx->pinned, tot, 2 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
    B2R_CUDA(cudaStreamSynchronize(st));
    ctx->mc_nv = (int64_t)((unsigned long long*)ctx->pinned)[0];
    ctx->mc_nt = (int64_t)((unsigned long long*)ctx->pinned)[1];
    B2R_CHECK(ctx->mc_nv < (1ll << 29), -8, "b2r_extract_mesh_count: %lld vertices exceed the 2^29 id space",
              (long long)ctx->mc_nv);
    *n_vertices = ctx->mc_nv; *n_triangles = ctx->mc_nt;
    return 0;
}

static int mc_state_ok(b2r_ctx* ctx, const b2r_volume* vol, const char* who) {
    B2R_CHECK(ctx && vol && vol->d_tsdfw, -1, "%s: NULL argument", who);
    B2R_CHECK(ctx->mc_R == vol->resolution && ctx->mc_zb == vol->x_begin && ctx->mc_ze == vol->x_end, -9,
              "%s: call b2r_extract_mesh_count on this volume first", who);
    B2R_CHECK(ctx->mc_vol == vol->d_tsdfw && ctx->mc_gen == ctx->write_gen, -9,
              "%s: the volume was written (or another volume counted) since b2r_extract_mesh_count; count again", who);
    return 0;
}

int b2r_extract_mesh_vertices(b2r_ctx* ctx, const b2r_volume* vol, int64_t vertex_base, double* d_vertices,
                              double* d_colors, void* stream) {
    if (int rc = mc_state_ok(ctx, vol, "b2r_extract_mesh_vertices")) return rc;
    B2R_CUDA(cudaSetDevice(ctx->device));
    const int R = vol->resolution, nz = vol->x_end - vol->x_begin;
    if (nz <= 0 || R < 2) return 0;
    B2R_CHECK(vertex_base >= 0 && vertex_base + ctx->mc_nv < (1ll << 29), -8,
              "b2r_extract_mesh_vertices: vertex ids exceed the 2^29 id space");
    B2R_CHECK(ctx->mc_nv == 0 || d_vertices, -1, "b2r_extract_mesh_vertices: NULL output");
    const int64_t n = (int64_t)nz * R * R;
    const int nblocks = cdiv(n, MC_THREADS);
    McGeom G{R, vol->x_begin, vol->x_end, vol->voxel_length, vol->origin[0], vol->origin[1], vol->origin[2],
             vol->color_type};
    const uint32_t* live_list = (const uint32_t*)ctx->mc_live.ptr + 2 * (size_t)nblocks;
    const unsigned long long* n_live = (const unsigned long long*)ctx->small.ptr + SMALL_MC_NLIVE;
    mc_emit_vertices_kernel<<<ctx->sm_count * 8, MC_THREADS, 0, S(stream)>>>(
        (const float2*)vol->d_tsdfw, (const uint32_t*)vol->d_color, G, live_list, n_live,
        (uint32_t*)ctx->mc_vert_id.ptr,
        (const uint32_t*)ctx->mc_block_counts.ptr, (uint32_t)vertex_base, d_vertices,
        vol->color_type ? d_colors : nullptr);
    B2R_LAUNCH_CHECK(ctx);
    return 0;
}

int b2r_mesh_seam_buffers(b2r_ctx* ctx, const b2r_volume* vol, void** d_send_first_plane, void** d_recv_top_plane,
                          int64_t* bytes) {
    if (int rc = mc_state_ok(ctx, vol, "b2r_mesh_seam_buffers")) return rc;
    const int64_t plane = (int64_t)vol->resolution * vol->resolution;
    uint32_t* tab = (uint32_t*)ctx->mc_vert_id.ptr;
    if (d_send_first_plane) *d_send_first_plane = tab + plane * B2R_PAD_LO;
    if (d_recv_top_plane) *d_recv_top_plane = tab + plane * (B2R_PAD_LO + (vol->x_end - vol->x_begin));
    if (bytes) *bytes = plane * (int64_t)sizeof(uint32_t);
    return 0;
}

int b2r_extract_mesh_triangles(b2r_ctx* ctx, const b2r_volume* vol, int32_t* d_triangles, void* stream) {
    if (int rc = mc_state_ok(ctx, vol, "b2r_extract_mesh_triangles")) return rc;
    B2R_CUDA(cudaSetDevice(ctx->device));
    const int R = vol->resolution, nz = vol->x_end - vol->x_begin;
    if (nz <= 0 || R < 2 || ctx->mc_nt == 0) return 0;
    B2R_CHECK(d_triangles, -1, "b2r_extract_mesh_triangles: NULL output");
    const int64_t n = (int64_t)nz * R * R;
    const int nblocks = cdiv(n, MC_THREADS);
    int xc_lo, xc_hi;
    cube_x_range(vol, xc_lo, xc_hi);
    const uint32_t* blk_t = (const uint32_t*)ctx->mc_block_counts.ptr + nblocks;
    const uint32_t* live_list = (const uint32_t*)ctx->mc_live.ptr + 2 * (size_t)nblocks;
    const unsigned long long* n_live = (const unsigned long long*)ctx->small.ptr + SMALL_MC_NLIVE;
    mc_emit_triangles_kernel<<<ctx->sm_count * 8, MC_THREADS, 0, S(stream)>>>((const uint8_t*)ctx->mc_code.ptr,
                                                                              (const uint8_t*)ctx->mc_flags.ptr,
                                                                              (const uint32_t*)ctx->mc_vert_id.ptr, R,
                                                                              vol->x_begin, vol->x_end, xc_hi, live_list, n_live,
                                                                              blk_t, d_triangles);
    B2R_LAUNCH_CHECK(ctx);
    return 0;
}

}  // extern "C"
