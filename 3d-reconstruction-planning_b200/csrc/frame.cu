// Frame-side kernels: depth conversion, frame preparation for integration
// ({depth, distance-multiplier} float2 image + packed colour, A.3), ordered back-projection (A.2).
// Reference call sites: stitcher.py:34-46.
#include <mutex>
#include "frame.cuh"
#include "scan.cuh"

namespace b2r {

__device__ __forceinline__ float load_depth_raw(const void* depth, int dtype, int64_t i) {
    if (dtype == 0) return (float)((const uint8_t*)depth)[i] / 255.0f;
    if (dtype == 1) return (float)((const uint16_t*)depth)[i];
    return ((const float*)depth)[i];
}

// O3D Image::ConvertDepthToFloatImage: p /= (float)scale; if (p >= (float)trunc) p = 0.
__global__ void depth_to_float_kernel(const void* __restrict__ depth, int dtype, int64_t n, float scale, float trunc,
                                      float* __restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float p = load_depth_raw(depth, dtype, i);
    p = p / scale;
    if (p >= trunc) p = 0.0f;
    out[i] = p;
}

// One pass per frame: metres depth (optionally converting from the raw dtype), the
// depth->camera-distance multiplier of O3D CreateDepthToCameraDistanceMultiplierFloatImage
// (float32, sqrtf(xx^2 + yy^2 + 1), same operation order) as one 8-byte pixel record {depth, multiplier}, and
// the colour packed to one 32-bit word in the volume's own format.
//
// A block is a 128 x 8 pixel strip (a thread = 8 pixels of a column, all 8 loads in flight together), so the
// max-depth pyramid comes out of registers: level 0 (8 x 8 cells) is complete inside a warp and is STORED (no
// atomics, no zero-fill needed); the strip contributes 16 atomicMax to levels 1..4 and the frame maximum instead
// of the 320 a row-per-block kernel needs for the same pixels (those same-address atomics were most of the old
// kernel's 13 us).  Levels 1..4 and the frame maximum must be zero on entry (see b2r_frame_slot::clean).
__global__ void __launch_bounds__(128) frame_prepare_kernel(const void* __restrict__ depth, int dtype, int convert, float scale,
                                                            float trunc, const void* __restrict__ color, int color_mode, int H,
                                                            int W, float cx, float cy, float ffl_inv_x, float ffl_inv_y,
                                                            float2* __restrict__ px, uint32_t* __restrict__ pcol, const MipDesc M,
                                                            unsigned int* __restrict__ mip, unsigned int* __restrict__ dmax_bits,
                                                            unsigned long long* __restrict__ tr) {
    if (threadIdx.x == 0) trace_begin(tr);
    const int u = blockIdx.x * 128 + threadIdx.x;
    const int v0 = blockIdx.y * 8;
    float m = 0.0f;
    if (u < W) {
        float d[8];
#pragma unroll
        for (int r = 0; r < 8; ++r) d[r] = (v0 + r < H) ? load_depth_raw(depth, dtype, (int64_t)(v0 + r) * W + u) : 0.0f;
        uint32_t c[8];
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            c[r] = 0u;
            if (v0 + r < H) {
                const int64_t i = (int64_t)(v0 + r) * W + u;
                if (color_mode == 1) {
                    const uint8_t* p = (const uint8_t*)color + i * 3;
                    // the volume's own packed format (10:10:10 fixed point, value * 4) so that the integrate kernel
                    // unpacks frame and voxel colour with the same three instructions
                    c[r] = ((uint32_t)p[0] << 2) | ((uint32_t)p[1] << 12) | ((uint32_t)p[2] << 22);
                } else if (color_mode == 2) {
                    c[r] = __float_as_uint(((const float*)color)[i]);
                }
            }
        }
        const float xx = ((float)u - cx) * ffl_inv_x;
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            if (v0 + r < H) {
                float dd = d[r];
                if (convert) {
                    dd = dd / scale;
                    if (dd >= trunc) dd = 0.0f;
                }
                const float yy = ((float)(v0 + r) - cy) * ffl_inv_y;
                const float mult = sqrtf(xx * xx + yy * yy + 1.0f);
                const int64_t i = (int64_t)(v0 + r) * W + u;
                px[i] = make_float2(dd, mult);
                pcol[i] = c[r];
                m = fmaxf(m, dd);      // (dd <= 0 and NaN never raise the maximum)
            }
        }
    }
    // Max-depth pyramid (cells of 8, 16, 32, 64, 128 pixels; positive floats order like their bit patterns) and
    // the frame-wide maximum.  The integrate kernels bound the depth a run of voxels can read by the maximum over
    // the cells covering its image footprint and skip voxels more than sdf_trunc behind it (hierarchical z).
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int cv = blockIdx.y;
#pragma unroll
    for (int o = 1; o < 8; o <<= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((lane & 7) == 0 && (u >> 3) < M.w[0]) mip[M.off[0] + cv * M.w[0] + (u >> 3)] = __float_as_uint(m);
    m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 8));
    if ((lane & 15) == 0 && m > 0.0f) atomicMax(&mip[M.off[1] + (cv >> 1) * M.w[1] + (u >> 4)], __float_as_uint(m));
    m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 16));
    __shared__ float s_m[4];
    if (lane == 0) {
        s_m[warp] = m;
        if (m > 0.0f) atomicMax(&mip[M.off[2] + (cv >> 2) * M.w[2] + (u >> 5)], __float_as_uint(m));
    }
    __syncthreads();
    if (threadIdx.x < 2) {
        const float m64 = fmaxf(s_m[threadIdx.x * 2], s_m[threadIdx.x * 2 + 1]);
        if (m64 > 0.0f) atomicMax(&mip[M.off[3] + (cv >> 3) * M.w[3] + (blockIdx.x * 2 + threadIdx.x)], __float_as_uint(m64));
        if (threadIdx.x == 0) {
            const float m128 = fmaxf(fmaxf(s_m[0], s_m[1]), fmaxf(s_m[2], s_m[3]));
            if (m128 > 0.0f) {
                atomicMax(&mip[M.off[4] + (cv >> 4) * M.w[4] + blockIdx.x], __float_as_uint(m128));
                atomicMax(dmax_bits, __float_as_uint(m128));
            }
            trace_end(tr);
        }
    }
}

MipDesc mip_desc(int H, int W) {
    MipDesc M;
    int off = 0;
    for (int l = 0; l < B2R_MIP_LEVELS; ++l) {
        const int cell = 8 << l;
        M.w[l] = (W + cell - 1) / cell;
        M.h[l] = (H + cell - 1) / cell;
        M.off[l] = off;
        off += M.w[l] * M.h[l];
    }
    M.total = off;
    return M;
}

int prepare_frame(b2r_ctx* ctx, b2r_frame_slot* slot, const void* d_depth, int dtype, int convert, double depth_scale, double depth_trunc,
                  const void* d_color, int color_mode, const b2r_intrinsic* K, cudaStream_t stream) {
    const int H = K->height, W = K->width;
    const size_t n = (size_t)H * W;
    if (slot->dm.ensure(n * sizeof(float2))) return -102;
    if (slot->rgba.ensure(n * sizeof(uint32_t))) return -102;
    const MipDesc M = mip_desc(H, W);
    const size_t mip_bytes = (size_t)(M.total + 1) * sizeof(float);
    if (slot->mip.bytes < mip_bytes || slot->counters.bytes < 64) slot->clean = false;
    if (slot->mip.ensure(mip_bytes)) return -102;
    if (slot->counters.ensure(64)) return -102;
    if (!slot->clean) {
        B2R_CUDA(cudaMemsetAsync(slot->mip.ptr, 0, mip_bytes, stream));
        B2R_CUDA(cudaMemsetAsync(slot->counters.ptr, 0, 64, stream));
    }
    slot->clean = false;      // set again by integrate_prepared once an items kernel that re-zeroes them is queued
    unsigned int* mip = (unsigned int*)slot->mip.ptr;
    unsigned int* dmax = mip + M.total;   // frame-wide maximum depth sits behind the pyramid
    const float ffl_inv_x = 1.0f / (float)K->fx, ffl_inv_y = 1.0f / (float)K->fy;
    dim3 block(128), grid(cdiv(W, 128), cdiv(H, 8));
    frame_prepare_kernel<<<grid, block, 0, stream>>>(d_depth, dtype, convert, (float)depth_scale, (float)depth_trunc,
                                                     d_color, color_mode, H, W, (float)K->cx, (float)K->cy, ffl_inv_x,
                                                     ffl_inv_y, (float2*)slot->dm.ptr, (uint32_t*)slot->rgba.ptr, M,
                                                     mip, dmax, trace_row(ctx, 0));
    B2R_LAUNCH_CHECK(ctx);
    return 0;
}

// ---------------------------------------------------------------------------------------------
// Back-projection with order-preserving compaction (row-major pixel order, invalid dropped).
// ---------------------------------------------------------------------------------------------
constexpr int BP_THREADS = 256;

__global__ void __launch_bounds__(BP_THREADS) backproject_count_kernel(const float* __restrict__ depth, int64_t n,
                                                                       uint32_t* __restrict__ block_counts) {
    const int64_t i = (int64_t)blockIdx.x * BP_THREADS + threadIdx.x;
    const bool valid = (i < n) && (depth[i] > 0.0f);
    const int c = __syncthreads_count(valid);
    if (threadIdx.x == 0) block_counts[blockIdx.x] = (uint32_t)c;
}

struct BackprojectParams {
    double pose[12];  // camera->world rows 0..2
    double fx, fy, cx, cy;
    int W, nc;
};

__global__ void __launch_bounds__(BP_THREADS) backproject_emit_kernel(const float* __restrict__ depth,
                                                                      const uint8_t* __restrict__ color, int64_t n,
                                                                      const __grid_constant__ BackprojectParams P,
                                                                      const uint32_t* __restrict__ block_offsets,
                                                                      double* __restrict__ pts, double* __restrict__ cols) {
    __shared__ uint32_t warp_counts[BP_THREADS / 32];
    const int64_t i = (int64_t)blockIdx.x * BP_THREADS + threadIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float d = 0.0f;
    if (i < n) d = depth[i];
    const bool valid = d > 0.0f;
    const unsigned ballot = __ballot_sync(0xffffffffu, valid);
    if (lane == 0) warp_counts[warp] = __popc(ballot);
    __syncthreads();
    uint32_t off = block_offsets[blockIdx.x];
    for (int w = 0; w < warp; ++w) off += warp_counts[w];
    off += __popc(ballot & ((1u << lane) - 1u));
    if (!valid) return;
    const int row = (int)(i / P.W), col = (int)(i % P.W);
    const double z = (double)d;
    const double x = (col - P.cx) * z / P.fx;
    const double y = (row - P.cy) * z / P.fy;
#pragma unroll
    for (int r = 0; r < 3; ++r)
        pts[(int64_t)off * 3 + r] = ((P.pose[r * 4 + 0] * x + P.pose[r * 4 + 1] * y) + P.pose[r * 4 + 2] * z) + P.pose[r * 4 + 3];
    if (color && cols) {
        const uint8_t* pc = color + i * P.nc;
        cols[(int64_t)off * 3 + 0] = pc[0] / 255.0;
        cols[(int64_t)off * 3 + 1] = pc[(P.nc - 1) / 2] / 255.0;
        cols[(int64_t)off * 3 + 2] = pc[P.nc - 1] / 255.0;
    }
}

// ---------------------------------------------------------------------------------------------
// b2r_backproject_static: the same back-projection for callers that cannot allocate (CUDA graph capture, serving loops).
// Its scratch -- the per-block valid counts, turned in place into the blocks' output offsets -- lives in a file-scope
// device array of fixed size instead of the context's growable cudaMalloc scratch.  Every GPU that loads the library
// holds these 64 KB until exit; every call on a device shares them (calls that overlap on the GPU race).
// ---------------------------------------------------------------------------------------------
constexpr int64_t BP_STATIC_MAX_PIXELS = 4194304;                            // 2^22: 2048 x 2048, 1920 x 1080, 1280 x 720 ...
constexpr int BP_STATIC_MAX_BLOCKS = (int)(BP_STATIC_MAX_PIXELS / BP_THREADS);   // 16384 counts = 64 KB
static __device__ __align__(256) uint32_t g_bp_static_scratch[BP_STATIC_MAX_BLOCKS];

// One block turns the <= 16384 block counts into exclusive offsets in place and writes the total (the number of points).
constexpr int BP_SCAN_THREADS = 1024;
__global__ void __launch_bounds__(BP_SCAN_THREADS) backproject_static_scan_kernel(uint32_t* __restrict__ counts, int nblocks,
                                                                                 long long* __restrict__ n_points) {
    __shared__ uint32_t warp_sums[BP_SCAN_THREADS / 32];
    constexpr int PER = BP_STATIC_MAX_BLOCKS / BP_SCAN_THREADS;              // 16 consecutive counts per thread
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, base = threadIdx.x * PER;
    uint32_t v[PER], sum = 0;
#pragma unroll
    for (int k = 0; k < PER; ++k) { v[k] = (base + k < nblocks) ? counts[base + k] : 0u; sum += v[k]; }
    uint32_t incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
    if (lane == 31) warp_sums[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        uint32_t w = warp_sums[lane], wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, wi, o); if (lane >= o) wi += t; }
        warp_sums[lane] = wi - w;                                            // exclusive prefix of the warp totals
        if (lane == 31) *n_points = (long long)wi;
    }
    __syncthreads();
    uint32_t run = warp_sums[warp] + incl - sum;
#pragma unroll
    for (int k = 0; k < PER; ++k) { if (base + k < nblocks) counts[base + k] = run; run += v[k]; }
}

// 4x4 inverse on the host (double, Gauss-Jordan with partial pivoting).
void invert4(const double* m, double* out) {
    double a[4][8];
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) {
            a[i][j] = m[i * 4 + j];
            a[i][4 + j] = (i == j) ? 1.0 : 0.0;
        }
    for (int c = 0; c < 4; ++c) {
        int p = c;
        for (int r = c + 1; r < 4; ++r)
            if (fabs(a[r][c]) > fabs(a[p][c])) p = r;
        if (p != c)
            for (int j = 0; j < 8; ++j) { double t = a[p][j]; a[p][j] = a[c][j]; a[c][j] = t; }
        const double d = a[c][c];
        for (int j = 0; j < 8; ++j) a[c][j] /= d;
        for (int r = 0; r < 4; ++r) {
            if (r == c) continue;
            const double f = a[r][c];
            if (f != 0.0)
                for (int j = 0; j < 8; ++j) a[r][j] -= f * a[c][j];
        }
    }
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) out[i * 4 + j] = a[i][4 + j];
}

}  // namespace b2r

using namespace b2r;

extern "C" {

int b2r_depth_to_float(b2r_ctx* ctx, const void* d_depth, int dtype, int H, int W, double depth_scale,
                       double depth_trunc, float* d_out, void* stream) {
    B2R_CHECK(ctx && d_depth && d_out, -1, "b2r_depth_to_float: NULL argument");
    B2R_CHECK(dtype >= 0 && dtype <= 2, -5, "b2r_depth_to_float: unsupported depth dtype %d", dtype);
    B2R_CUDA(cudaSetDevice(ctx->device));
    const int64_t n = (int64_t)H * W;
    if (n == 0) return 0;
    depth_to_float_kernel<<<cdiv(n, 256), 256, 0, S(stream)>>>(d_depth, dtype, n, (float)depth_scale,
                                                               (float)depth_trunc, d_out);
    B2R_LAUNCH_CHECK(ctx);
    return 0;
}

int b2r_backproject(b2r_ctx* ctx, const float* d_depth, const uint8_t* d_color, int nc, const b2r_intrinsic* K,
                    const double* extrinsic, double* d_points, double* d_colors, int64_t* n_points, void* stream) {
    B2R_CHECK(ctx && d_depth && K && extrinsic && d_points && n_points, -1, "b2r_backproject: NULL argument");
    B2R_CHECK(!d_color || nc == 3 || nc == 4, -5, "b2r_backproject: colour must have 3 or 4 channels, got %d", nc);
    B2R_CUDA(cudaSetDevice(ctx->device));
    const int64_t n = (int64_t)K->width * K->height;
    *n_points = 0;
    if (n == 0) return 0;
    const int nblocks = cdiv(n, BP_THREADS);
    if (ctx->scan_a.ensure((size_t)nblocks * sizeof(uint32_t))) return -102;
    uint32_t* counts = (uint32_t*)ctx->scan_a.ptr;
    unsigned long long* total = (unsigned long long*)ctx->small.ptr + SMALL_TOTAL64;
    cudaStream_t st = S(stream);
    backproject_count_kernel<<<nblocks, BP_THREADS, 0, st>>>(d_depth, n, counts);
    B2R_LAUNCH_CHECK(ctx);
    if (int rc = scan_exclusive_u32(ctx, counts, counts, nblocks, total, st)) return rc;
    BackprojectParams P;
    double pose[16];
    invert4(extrinsic, pose);
    for (int i = 0; i < 12; ++i) P.pose[i] = pose[i];
    P.fx = K->fx; P.fy = K->fy; P.cx = K->cx; P.cy = K->cy; P.W = K->width; P.nc = nc;
    backproject_emit_kernel<<<nblocks, BP_THREADS, 0, st>>>(d_depth, d_color, n, P, counts, d_points, d_colors);
    B2R_LAUNCH_CHECK(ctx);
    B2R_CUDA(cudaMemcpyAsync(ctx->pinned, total, sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
    B2R_CUDA(cudaStreamSynchronize(st));
    *n_points = (int64_t) * (unsigned long long*)ctx->pinned;
    return 0;
}

// The device address of g_bp_static_scratch on `device` (which must be current): in host code the array's name is a
// host-side stand-in, not a device pointer -- a kernel given it would fault or, where the GPU can reach host memory,
// silently use the wrong memory.  Looked up once per device and cached; no allocation.
static int bp_static_scratch(int device, uint32_t** out) {
    constexpr int MAX_DEV = 64;
    static std::mutex mu;
    static uint32_t* ptrs[MAX_DEV] = {nullptr};
    B2R_CHECK(device >= 0 && device < MAX_DEV, -6, "b2r_backproject_static: device %d out of range [0, %d)", device, MAX_DEV);
    std::lock_guard<std::mutex> lock(mu);
    if (!ptrs[device]) {
        void* p = nullptr;
        B2R_CUDA(cudaGetSymbolAddress(&p, g_bp_static_scratch));
        B2R_CHECK(p != nullptr, -100, "b2r_backproject_static: cudaGetSymbolAddress returned a NULL address");
        ptrs[device] = (uint32_t*)p;
    }
    *out = ptrs[device];
    return 0;
}

int b2r_backproject_static_max_pixels(int64_t* out) {
    B2R_CHECK(out, -1, "b2r_backproject_static_max_pixels: NULL argument");
    *out = BP_STATIC_MAX_PIXELS;
    return 0;
}

int b2r_backproject_static(int device, const float* d_depth, const uint8_t* d_color, int nc, const b2r_intrinsic* K,
                           const double* extrinsic, double* d_points, double* d_colors, int64_t* d_n_points,
                           void* stream) {
    B2R_CHECK(d_depth && K && extrinsic && d_points && d_n_points, -1, "b2r_backproject_static: NULL argument");
    B2R_CHECK(!d_color || nc == 3 || nc == 4, -5, "b2r_backproject_static: colour must have 3 or 4 channels, got %d", nc);
    B2R_CHECK(K->width >= 0 && K->height >= 0, -6, "b2r_backproject_static: bad image size %d x %d", K->width, K->height);
    const int64_t n = (int64_t)K->width * K->height;
    B2R_CHECK(n <= BP_STATIC_MAX_PIXELS, -7,
              "b2r_backproject_static: %d x %d = %lld pixels exceed the static scratch's maximum of %lld pixels; "
              "use b2r_backproject (which allocates) for larger images",
              K->width, K->height, (long long)n, (long long)BP_STATIC_MAX_PIXELS);
    B2R_CUDA(cudaSetDevice(device));
    uint32_t* counts = nullptr;
    if (int rc = bp_static_scratch(device, &counts)) return rc;
    cudaStream_t st = S(stream);
    // (nothing in the scratch has to start at zero: the count kernel overwrites the nblocks entries the other two read;
    //  the previous call's contents beyond them are never looked at)
    const int nblocks = (int)cdiv(n, BP_THREADS);
    if (nblocks > 0) backproject_count_kernel<<<nblocks, BP_THREADS, 0, st>>>(d_depth, n, counts);
    backproject_static_scan_kernel<<<1, BP_SCAN_THREADS, 0, st>>>(counts, nblocks, (long long*)d_n_points);
    if (nblocks > 0) {
        BackprojectParams P;
        double pose[16];
        invert4(extrinsic, pose);
        for (int i = 0; i < 12; ++i) P.pose[i] = pose[i];
        P.fx = K->fx; P.fy = K->fy; P.cx = K->cx; P.cy = K->cy; P.W = K->width; P.nc = nc;
        backproject_emit_kernel<<<nblocks, BP_THREADS, 0, st>>>(d_depth, d_color, n, P, counts, d_points, d_colors);
    }
    const cudaError_t le = cudaGetLastError();
    B2R_CHECK(le == cudaSuccess, -101, "b2r_backproject_static: kernel launch failed: %s", cudaGetErrorString(le));
    return 0;
}

}  // extern "C"
