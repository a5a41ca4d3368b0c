// Surface point extraction -- replaces Open3D UniformTSDFVolume::ExtractPointCloud (north_star API
// extract_point_cloud(); arithmetic in SURVEY.md A.4): sign-change edges between valid voxels
// (w != 0, -0.98 <= f < 0.98), interpolated position, colour, and the normalised central-difference
// gradient of the trilinearly interpolated TSDF (half gap 0.99 voxel).  Two passes over the owned
// voxels (count -> scan -> emit) with warp-ballot/prefix-sum compaction; output in memory order.
#include "frame.cuh"
#include "scan.cuh"

namespace b2r {

constexpr int PC_THREADS = 256;

struct PcGeom {
    int R, x_begin, x_end;
    double vl, ox, oy, oz;
    int color_type;
    // ScalableTSDFVolume mode (unit_res > 0): Open3D computes positions from the unit's own origin + in-unit index
    // and GetTSDFAt returns 0 when the unit holding the sample's base voxel was never opened
    int unit_res, nu, uoff[3];
    double unit_len;
    const int* unit_stamp;
};

__device__ __forceinline__ int64_t pidx(int R, int x_begin, int x, int y, int z) {
    return ((int64_t)(x - x_begin + B2R_PAD_LO) * R + z) * R + y;      // x-slowest, then z, y fastest
}
__device__ __forceinline__ bool pc_valid(const float2 q) { return q.y != 0.0f && q.x < 0.98f && q.x >= -0.98f; }

// bit a set when the edge voxel -> voxel + e_a produces a point
__device__ __forceinline__ unsigned pc_flags(const float2* __restrict__ tsdfw, const PcGeom& G, int x, int y, int z) {
    const int R = G.R;
    if (x < 1 || y < 1 || z < 1 || x > R - 2 || y > R - 2 || z > R - 2) return 0u;
    const int64_t i0 = pidx(R, G.x_begin, x, y, z);
    const float2 q0 = __ldg(&tsdfw[i0]);
    if (!pc_valid(q0)) return 0u;
    unsigned f = 0;
    if (x + 1 < R - 1) { const float2 q = __ldg(&tsdfw[i0 + (int64_t)R * R]); if (pc_valid(q) && q0.x * q.x < 0.0f) f |= 1u; }
    if (y + 1 < R - 1) { const float2 q = __ldg(&tsdfw[i0 + 1]); if (pc_valid(q) && q0.x * q.x < 0.0f) f |= 2u; }
    if (z + 1 < R - 1) { const float2 q = __ldg(&tsdfw[i0 + R]); if (pc_valid(q) && q0.x * q.x < 0.0f) f |= 4u; }
    return f;
}

__global__ void __launch_bounds__(PC_THREADS) pc_count_kernel(const float2* __restrict__ tsdfw,
                                                              const __grid_constant__ PcGeom G,
                                                              uint32_t* __restrict__ blk_counts) {
    __shared__ uint32_t s_c[PC_THREADS / 32];
    const int R = G.R;
    const int64_t n = (int64_t)(G.x_end - G.x_begin) * R * R;
    const int64_t t = (int64_t)blockIdx.x * PC_THREADS + threadIdx.x;
    unsigned c = 0;
    if (t < n) {
        const int y = (int)(t % R), z = (int)((t / R) % R), x = G.x_begin + (int)(t / ((int64_t)R * R));
        c = __popc(pc_flags(tsdfw, G, x, y, z));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0) s_c[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned a = 0;
        for (int w = 0; w < PC_THREADS / 32; ++w) a += s_c[w];
        blk_counts[blockIdx.x] = a;
    }
}

// Streaming form of the count pass for even R: a thread loads TWO adjacent voxels with one 16-byte load and only
// looks at neighbours when one of them is a valid surface candidate (rare); a block of 256 threads covers 512
// voxels = two count blocks of the emit pass (warps 0-3 / 4-7).
__global__ void __launch_bounds__(PC_THREADS) pc_count2_kernel(const float2* __restrict__ tsdfw,
                                                               const __grid_constant__ PcGeom G,
                                                               uint32_t* __restrict__ blk_counts, int n_count_blocks) {
    __shared__ uint32_t s_c[PC_THREADS / 32];
    const int R = G.R;
    const int64_t n = (int64_t)(G.x_end - G.x_begin) * R * R;
    const int64_t t = ((int64_t)blockIdx.x * PC_THREADS + threadIdx.x) * 2;
    unsigned c = 0;
    if (t < n) {
        // owned voxels are contiguous in memory behind the B2R_PAD_LO halo planes: the streaming load needs no
        // index arithmetic; (x, y, z) are only worked out for the rare surface candidates
        const float4 q = __ldcs(reinterpret_cast<const float4*>(tsdfw + (int64_t)B2R_PAD_LO * R * R + t));
        const bool va = pc_valid(make_float2(q.x, q.y)), vb = pc_valid(make_float2(q.z, q.w));
        if (va | vb) {
            const int y = (int)(t % R), z = (int)((t / R) % R), x = G.x_begin + (int)(t / ((int64_t)R * R));
            if (va) c += __popc(pc_flags(tsdfw, G, x, y, z));
            if (vb) c += __popc(pc_flags(tsdfw, G, x, y + 1, z));
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0) s_c[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x < 2) {
        const int b = blockIdx.x * 2 + threadIdx.x;
        if (b < n_count_blocks) {
            unsigned a = 0;
            for (int w = 0; w < PC_THREADS / 64; ++w) a += s_c[threadIdx.x * (PC_THREADS / 64) + w];
            blk_counts[b] = a;
        }
    }
}

// Output-sensitive form (b2r_volume::d_rows, even R): a surface point needs q0.x * q1.x < 0, i.e. a NEGATIVE tsdf value at the
// voxel or at its +x / +y / +z neighbour, so a count block (256 consecutive voxels) none of whose (x, z) rows r has a mark
// at r, r + 1 (z + 1) or r + R (x + 1) has nothing to count.  pc_cand_kernel, a thread per count block, looks at the marks
// and appends the candidates to a list (warp-aggregated; order irrelevant, counts are stored by block id);
// pc_count_list_kernel counts them, persistent warps, a warp per candidate block; the counts are zeroed beforehand.
// Tried: skipping inside pc_count2_kernel (a block per 512 voxels with a barrier around the look) -- SLOWER than streaming
// the whole cube, 2.66 vs 1.88 ms at 1024^3; a warp per count block that looks and counts -- 0.77 ms, of which 0.44 the 4.2 M
// looks; a lane per look and the warp counting its batch's candidates one after the other -- 1.69 ms (the candidates cluster).
__global__ void __launch_bounds__(256) pc_cand_kernel(const uint8_t* __restrict__ rows, int R, int64_t n, int n_count_blocks,
                                                      uint32_t* __restrict__ list, unsigned long long* __restrict__ n_cand) {
    const int b = blockIdx.x * 256 + threadIdx.x, lane = threadIdx.x & 31;
    bool m = false;
    if (b < n_count_blocks) {
        const int64_t t0 = (int64_t)b * PC_THREADS, t1 = min(t0 + PC_THREADS, n) - 1;
        for (int64_t r = t0 / R; r <= t1 / R; ++r) {     // (r + 1 and r + R stay inside the B2R_PAD_HI planes)
            const uint8_t* mr = rows + (int64_t)B2R_PAD_LO * R + r;
            m |= (__ldg(mr) | __ldg(mr + 1) | __ldg(mr + R)) != 0;
        }
    }
    const unsigned live = __ballot_sync(0xffffffffu, m);
    if (!live) return;
    unsigned long long base = 0;
    if (lane == 0) base = atomicAdd(n_cand, (unsigned long long)__popc(live));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (m) list[base + __popc(live & ((1u << lane) - 1u))] = (uint32_t)b;
}

// Points of the voxel pair (x, y, z), (x, y + 1, z) at owned offset t (even: the pair is one 16-byte word, q).  The tests
// are pc_flags'; the +x and +z neighbours of both voxels come as one 16-byte word each, a's +y neighbour is b.
__device__ __forceinline__ unsigned pc_pair_count(const float2* __restrict__ own, int R, int64_t t, const float4 q, bool va, bool vb,
                                                  int x, int y, int z) {
    if (x < 1 || z < 1 || x > R - 2 || z > R - 2) return 0u;
    const bool ia = va && y >= 1 && y <= R - 2, ib = vb && y + 1 <= R - 2;
    if (!(ia | ib)) return 0u;
    const float4 qx = __ldg(reinterpret_cast<const float4*>(own + t + (int64_t)R * R));
    const float4 qz = __ldg(reinterpret_cast<const float4*>(own + t + R));
    const float2 qy = __ldg(own + t + 2);
    const bool xn = x + 1 < R - 1, zn = z + 1 < R - 1;
    unsigned c = 0;
    if (ia) {
        c += (xn && pc_valid(make_float2(qx.x, qx.y)) && q.x * qx.x < 0.0f) ? 1u : 0u;
        c += (y + 1 < R - 1 && vb && q.x * q.z < 0.0f) ? 1u : 0u;
        c += (zn && pc_valid(make_float2(qz.x, qz.y)) && q.x * qz.x < 0.0f) ? 1u : 0u;
    }
    if (ib) {
        c += (xn && pc_valid(make_float2(qx.z, qx.w)) && q.z * qx.z < 0.0f) ? 1u : 0u;
        c += (y + 2 < R - 1 && pc_valid(qy) && q.z * qy.x < 0.0f) ? 1u : 0u;
        c += (zn && pc_valid(make_float2(qz.z, qz.w)) && q.z * qz.z < 0.0f) ? 1u : 0u;
    }
    return c;
}

__global__ void __launch_bounds__(256) pc_count_list_kernel(const float2* __restrict__ tsdfw, const __grid_constant__ PcGeom G,
                                                            uint32_t* __restrict__ blk_counts, const uint32_t* __restrict__ list,
                                                            const unsigned long long* __restrict__ n_cand) {
    const int R = G.R, lane = threadIdx.x & 31;
    const int64_t n = (int64_t)(G.x_end - G.x_begin) * R * R;
    const float2* own = tsdfw + (int64_t)B2R_PAD_LO * R * R;
    const unsigned nwarps = gridDim.x * 8, nc = (unsigned)*n_cand;
    for (unsigned i = blockIdx.x * 8 + (threadIdx.x >> 5); i < nc; i += nwarps) {
        const unsigned b = list[i];
        const int64_t t0 = (int64_t)b * PC_THREADS;
        // the block's first voxel, once per block; a lane's own (x, y, z) only needs divisions when the block spans rows
        const int64_t r0 = t0 / R;
        const int yb = (int)(t0 - r0 * R), zb = (int)(r0 % R), xb = G.x_begin + (int)(r0 / R);
        unsigned c = 0;
#pragma unroll
        for (int k = 0; k < PC_THREADS / 64; ++k) {
            const int off = (k * 32 + lane) * 2;
            const int64_t t = t0 + off;
            if (t < n) {
                const float4 q = __ldcs(reinterpret_cast<const float4*>(own + t));
                const bool va = pc_valid(make_float2(q.x, q.y)), vb = pc_valid(make_float2(q.z, q.w));
                if (va | vb) {
                    int x = xb, y = yb + off, z = zb;
                    if (y >= R) { y = (int)(t % R); z = (int)((t / R) % R); x = G.x_begin + (int)(t / ((int64_t)R * R)); }
                    c += pc_pair_count(own, R, t, q, va, vb, x, y, z);
                }
            }
        }
        c = __reduce_add_sync(0xffffffffu, c);
        if (lane == 0) blk_counts[b] = c;
    }
}

// raw per-block counts -> marks -> (scan) -> list of the blocks that produce points: the emit pass is a persistent
// kernel over that list
__global__ void __launch_bounds__(256) pc_mark_kernel(const uint32_t* __restrict__ blk_counts, int n, uint32_t* __restrict__ live) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) live[i] = blk_counts[i] != 0u ? 1u : 0u;
}
__global__ void __launch_bounds__(256) pc_live_list_kernel(const uint32_t* __restrict__ live, const uint32_t* __restrict__ pos, int n,
                                                           uint32_t* __restrict__ list) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && live[i]) list[pos[i]] = (uint32_t)i;
}

// Open3D GetTSDFAt(p): trilinear blend of the 8 surrounding tsdf values (no weight check).
__device__ __forceinline__ double tsdf_at(const float2* __restrict__ tsdfw, const PcGeom& G, const double* p) {
    int idx[3];
    double r[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const double g = p[k] / G.vl - 0.5;
        idx[k] = (int)floor(g);
        r[k] = g - (double)idx[k];
    }
    const int R = G.R;
    const int64_t b = pidx(R, G.x_begin, idx[0], idx[1], idx[2]);
    const int64_t sx = (int64_t)R * R, sy = 1, sz = R;
    const double t000 = (double)__ldg(&tsdfw[b]).x, t001 = (double)__ldg(&tsdfw[b + sz]).x;
    const double t010 = (double)__ldg(&tsdfw[b + sy]).x, t011 = (double)__ldg(&tsdfw[b + sy + sz]).x;
    const double t100 = (double)__ldg(&tsdfw[b + sx]).x, t101 = (double)__ldg(&tsdfw[b + sx + sz]).x;
    const double t110 = (double)__ldg(&tsdfw[b + sx + sy]).x, t111 = (double)__ldg(&tsdfw[b + sx + sy + sz]).x;
    return (1 - r[0]) * ((1 - r[1]) * ((1 - r[2]) * t000 + r[2] * t001) + r[1] * ((1 - r[2]) * t010 + r[2] * t011)) +
           r[0] * ((1 - r[1]) * ((1 - r[2]) * t100 + r[2] * t101) + r[1] * ((1 - r[2]) * t110 + r[2] * t111));
}

// Open3D ScalableTSDFVolume::GetTSDFAt(p), p in WORLD coordinates: the unit and the in-unit voxel of
// p - vl/2 are located separately (index clamped into the unit), a missing base unit reads 0, neighbours in
// missing units read 0 (the dense cube holds zeros there).
__device__ __forceinline__ double tsdf_at_units(const float2* __restrict__ tsdfw, const PcGeom& G, const double* p) {
    int idx[3];
    double r[3];
    int u[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const double pl = p[k] - 0.5 * G.vl;
        const int index0 = (int)floor(pl / G.unit_len);
        const double g = (pl - (double)index0 * G.unit_len) / G.vl;
        int i0 = (int)floor(g);
        i0 = max(0, min(G.unit_res - 1, i0));
        r[k] = g - (double)i0;
        u[k] = index0 - G.uoff[k];
        idx[k] = u[k] * G.unit_res + i0;
    }
    if ((unsigned)u[0] >= (unsigned)G.nu || (unsigned)u[1] >= (unsigned)G.nu || (unsigned)u[2] >= (unsigned)G.nu) return 0.0;
    if (__ldg(&G.unit_stamp[(u[0] * G.nu + u[1]) * G.nu + u[2]]) == 0) return 0.0;
    const int R = G.R;
    auto T = [&](int dx, int dy, int dz) -> double {
        const int x = idx[0] + dx, y = idx[1] + dy, z = idx[2] + dz;
        if (x >= R || y >= R || z >= R) return 0.0;      // the neighbouring unit lies outside the cube: missing
        return (double)__ldg(&tsdfw[pidx(R, G.x_begin, x, y, z)]).x;
    };
    return (1 - r[0]) * ((1 - r[1]) * ((1 - r[2]) * T(0, 0, 0) + r[2] * T(0, 0, 1)) + r[1] * ((1 - r[2]) * T(0, 1, 0) + r[2] * T(0, 1, 1))) +
           r[0] * ((1 - r[1]) * ((1 - r[2]) * T(1, 0, 0) + r[2] * T(1, 0, 1)) + r[1] * ((1 - r[2]) * T(1, 1, 0) + r[2] * T(1, 1, 1)));
}

template <bool UNITS>
__global__ void __launch_bounds__(PC_THREADS) pc_emit_kernel(const float2* __restrict__ tsdfw,
                                                             const uint32_t* __restrict__ color,
                                                             const __grid_constant__ PcGeom G,
                                                             const uint32_t* __restrict__ blk_off,
                                                             const uint32_t* __restrict__ live_list,
                                                             const unsigned long long* __restrict__ n_live,
                                                             double* __restrict__ pts, double* __restrict__ nrm,
                                                             double* __restrict__ cols) {
    __shared__ uint32_t smem[33];
    const int R = G.R;
    const int64_t n = (int64_t)(G.x_end - G.x_begin) * R * R;
    const unsigned nl = (unsigned)*n_live;
    for (unsigned li = blockIdx.x; li < nl; li += gridDim.x) {
    const unsigned vb = live_list[li];
    const int64_t t = (int64_t)vb * PC_THREADS + threadIdx.x;
    unsigned f = 0;
    int x = 0, y = 0, z = 0;
    if (t < n) {
        y = (int)(t % R); z = (int)((t / R) % R); x = G.x_begin + (int)(t / ((int64_t)R * R));
        f = pc_flags(tsdfw, G, x, y, z);
    }
    uint32_t total;
    const uint32_t local = block_exclusive_scan(__popc(f), smem, &total);
    if (total == 0 || f == 0) continue;
    int64_t k = (int64_t)blk_off[vb] + local;
    const double hvl = 0.5 * G.vl;
    const int64_t i0 = pidx(R, G.x_begin, x, y, z);
    const float f0 = tsdfw[i0].x;
    double p0[3] = {hvl + G.vl * (double)x, hvl + G.vl * (double)y, hvl + G.vl * (double)z};
    if (UNITS) {     // (hvl + vl * in-unit index) + index.cast<double>() * unit_length, world coordinates
        const int c[3] = {x, y, z};
#pragma unroll
        for (int q = 0; q < 3; ++q)
            p0[q] = (hvl + G.vl * (double)(c[q] % G.unit_res)) + (double)(G.uoff[q] + c[q] / G.unit_res) * G.unit_len;
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        if (!(f & (1u << a))) continue;
        const int64_t i1 = i0 + (a == 0 ? (int64_t)R * R : a == 1 ? (int64_t)1 : (int64_t)R);      // +x, +y, +z
        const float f1 = tsdfw[i1].x;
        const float r0 = fabsf(f0), r1 = fabsf(f1);
        double p[3] = {p0[0], p0[1], p0[2]};
        const double p1a = p0[a] + G.vl;
        p[a] = (p0[a] * (double)r1 + p1a * (double)r0) / (double)(r0 + r1);
        pts[k * 3 + 0] = UNITS ? p[0] : p[0] + G.ox;
        pts[k * 3 + 1] = UNITS ? p[1] : p[1] + G.oy;
        pts[k * 3 + 2] = UNITS ? p[2] : p[2] + G.oz;
        if (cols && G.color_type) {
            const uint32_t c0 = color[i0], c1 = color[i1];
#pragma unroll
            for (int q = 0; q < 3; ++q) {
                double a0, a1, s;
                if (G.color_type == 1) {
                    a0 = (double)((float)((c0 >> (10 * q)) & 1023u) * 0.25f);
                    a1 = (double)((float)((c1 >> (10 * q)) & 1023u) * 0.25f);
                    s = 255.0;
                } else {
                    a0 = (double)__uint_as_float(c0); a1 = (double)__uint_as_float(c1); s = 1.0;
                }
                cols[k * 3 + q] = (a0 * (double)r1 + a1 * (double)r0) / (double)(r0 + r1) / s;
            }
        }
        // GetNormalAt(p): n(i) = GetTSDFAt(p + gap e_i) - GetTSDFAt(p - gap e_i), normalised
        const double half_gap = 0.99 * G.vl;
        double nv[3];
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            double pp[3] = {p[0], p[1], p[2]};
            pp[q] -= half_gap;
            const double t0 = UNITS ? tsdf_at_units(tsdfw, G, pp) : tsdf_at(tsdfw, G, pp);
            pp[q] += 2 * half_gap;
            const double t1 = UNITS ? tsdf_at_units(tsdfw, G, pp) : tsdf_at(tsdfw, G, pp);
            nv[q] = t1 - t0;
        }
        const double nn = sqrt(nv[0] * nv[0] + nv[1] * nv[1] + nv[2] * nv[2]);
#pragma unroll
        for (int q = 0; q < 3; ++q) nrm[k * 3 + q] = nn > 0 ? nv[q] / nn : nv[q];
        ++k;
    }
    }
}


}  // namespace b2r

using namespace b2r;

extern "C" {

int b2r_extract_points_count(b2r_ctx* ctx, const b2r_volume* vol, int64_t* n_points, void* stream) {
    B2R_CHECK(ctx && vol && vol->d_tsdfw && n_points, -1, "b2r_extract_points_count: NULL argument");
    B2R_CUDA(cudaSetDevice(ctx->device));
    B2R_CUDA(join_side(ctx, S(stream)));   // items kernels of earlier integrate calls run on the side stream
    cudaStream_t st = S(stream);
    const int R = vol->resolution, nz = vol->x_end - vol->x_begin;
    *n_points = 0;
    ctx->pc_n = 0;
    ctx->pc_vol = vol->d_tsdfw; ctx->pc_gen = ctx->write_gen;
    if (nz <= 0 || R < 3) return 0;
    const int64_t n = (int64_t)nz * R * R;
    const int nblocks = cdiv(n, PC_THREADS);
    if (ctx->pc_block_counts.ensure((size_t)nblocks * 4 * sizeof(uint32_t))) return -102;
    uint32_t* blk = (uint32_t*)ctx->pc_block_counts.ptr;
    uint32_t* live = blk + nblocks;
    uint32_t* live_pos = blk + 2 * (size_t)nblocks;
    uint32_t* live_list = blk + 3 * (size_t)nblocks;
    PcGeom G{R, vol->x_begin, vol->x_end, vol->voxel_length, vol->origin[0], vol->origin[1], vol->origin[2],
             vol->color_type};
    if (vol->d_rows && R % 2 == 0 && (((uintptr_t)vol->d_tsdfw) & 15) == 0) {
        unsigned long long* n_cand = (unsigned long long*)ctx->small.ptr + SMALL_PC_NCAND;
        B2R_CUDA(cudaMemsetAsync(blk, 0, (size_t)nblocks * sizeof(uint32_t), st));
        B2R_CUDA(cudaMemsetAsync(n_cand, 0, sizeof(unsigned long long), st));
        // (the candidate list borrows live_list: pc_live_list_kernel overwrites it after the count kernel has read it)
        pc_cand_kernel<<<cdiv(nblocks, 256), 256, 0, st>>>((const uint8_t*)vol->d_rows, R, n, nblocks, live_list, n_cand);
        B2R_LAUNCH_CHECK(ctx);
        pc_count_list_kernel<<<ctx->sm_count * 8, 256, 0, st>>>((const float2*)vol->d_tsdfw, G, blk, live_list, n_cand);
    } else if (R % 2 == 0 && (((uintptr_t)vol->d_tsdfw) & 15) == 0)
        pc_count2_kernel<<<cdiv(nblocks, 2), PC_THREADS, 0, st>>>((const float2*)vol->d_tsdfw, G, blk, nblocks);
    else
        pc_count_kernel<<<nblocks, PC_THREADS, 0, st>>>((const float2*)vol->d_tsdfw, G, blk);
    B2R_LAUNCH_CHECK(ctx);
    pc_mark_kernel<<<cdiv(nblocks, 256), 256, 0, st>>>(blk, nblocks, live);
    B2R_LAUNCH_CHECK(ctx);
    if (int rc = scan_exclusive_u32(ctx, live, live_pos, nblocks, (unsigned long long*)ctx->small.ptr + SMALL_PC_NLIVE, st)) return rc;
    pc_live_list_kernel<<<cdiv(nblocks, 256), 256, 0, st>>>(live, live_pos, nblocks, live_list);
    B2R_LAUNCH_CHECK(ctx);
    unsigned long long* tot = (unsigned long long*)ctx->small.ptr + SMALL_TOTAL64_C;
    if (int rc = scan_exclusive_u32(ctx, blk, blk, nblocks, tot, st)) return rc;
    B2R_CUDA(cudaMemcpyAsync(ctx->pinned, tot, sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
    B2R_CUDA(cudaStreamSynchronize(st));
    ctx->pc_n = (int64_t) * (unsigned long long*)ctx->pinned;
    *n_points = ctx->pc_n;
    return 0;
}

static int points_emit(b2r_ctx* ctx, const b2r_volume* vol, int unit_res, const int32_t* unit_offset, const int32_t* d_unit_stamp,
                       double* d_points, double* d_normals, double* d_colors, void* stream) {
    B2R_CUDA(cudaSetDevice(ctx->device));
    B2R_CHECK(ctx->pc_vol == vol->d_tsdfw && ctx->pc_gen == ctx->write_gen, -9,
              "b2r_extract_points_emit: the volume was written (or another volume counted) since b2r_extract_points_count; count again");
    const int R = vol->resolution, nz = vol->x_end - vol->x_begin;
    if (nz <= 0 || R < 3 || ctx->pc_n == 0) return 0;
    B2R_CHECK(d_points && d_normals, -1, "b2r_extract_points_emit: NULL output");
    const int64_t n = (int64_t)nz * R * R;
    const int nblocks = cdiv(n, PC_THREADS);
    PcGeom G;
    G.R = R; G.x_begin = vol->x_begin; G.x_end = vol->x_end; G.vl = vol->voxel_length;
    G.ox = vol->origin[0]; G.oy = vol->origin[1]; G.oz = vol->origin[2]; G.color_type = vol->color_type;
    G.unit_res = unit_res; G.nu = unit_res ? R / unit_res : 0; G.unit_len = vol->voxel_length * unit_res;
    for (int a = 0; a < 3; ++a) G.uoff[a] = unit_offset ? unit_offset[a] : 0;
    G.unit_stamp = d_unit_stamp;
    const float2* tw = (const float2*)vol->d_tsdfw;
    const uint32_t* col = (const uint32_t*)vol->d_color;
    const uint32_t* blk = (const uint32_t*)ctx->pc_block_counts.ptr;
    const unsigned long long* nlive = (const unsigned long long*)ctx->small.ptr + SMALL_PC_NLIVE;
    double* dc = vol->color_type ? d_colors : nullptr;
    if (unit_res)
        pc_emit_kernel<true><<<ctx->sm_count * 8, PC_THREADS, 0, S(stream)>>>(tw, col, G, blk, blk + 3 * (size_t)nblocks, nlive, d_points, d_normals, dc);
    else
        pc_emit_kernel<false><<<ctx->sm_count * 8, PC_THREADS, 0, S(stream)>>>(tw, col, G, blk, blk + 3 * (size_t)nblocks, nlive, d_points, d_normals, dc);
    B2R_LAUNCH_CHECK(ctx);
    return 0;
}

int b2r_extract_points_emit(b2r_ctx* ctx, const b2r_volume* vol, double* d_points, double* d_normals,
                            double* d_colors, void* stream) {
    B2R_CHECK(ctx && vol && vol->d_tsdfw, -1, "b2r_extract_points_emit: NULL argument");
    return points_emit(ctx, vol, 0, nullptr, nullptr, d_points, d_normals, d_colors, stream);
}

int b2r_extract_points_emit_units(b2r_ctx* ctx, const b2r_volume* vol, int unit_resolution, const int32_t* unit_offset,
                                  const int32_t* d_unit_stamp, double* d_points, double* d_normals, double* d_colors,
                                  void* stream) {
    B2R_CHECK(ctx && vol && vol->d_tsdfw && unit_offset && d_unit_stamp, -1, "b2r_extract_points_emit_units: NULL argument");
    B2R_CHECK(unit_resolution >= 1 && vol->resolution % unit_resolution == 0 && vol->x_begin == 0 && vol->x_end == vol->resolution,
              -6, "b2r_extract_points_emit_units: whole cube of whole units expected");
    return points_emit(ctx, vol, unit_resolution, unit_offset, d_unit_stamp, d_points, d_normals, d_colors, stream);
}

}  // extern "C"
