// Device-wide exclusive prefix sum over uint32 (hand-written; used for order-preserving
// compaction in back-projection, marching cubes and point extraction).
#pragma once
#include "common.cuh"

namespace b2r {

// warp inclusive scan (shuffle-up ladder)
__device__ __forceinline__ uint32_t warp_inclusive_scan(uint32_t v, int lane) {
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t n = __shfl_up_sync(0xffffffffu, v, d);
        if (lane >= d) v += n;
    }
    return v;
}

// Block-wide exclusive scan of one value per thread (blockDim.x <= 1024, multiple of 32).
// `smem` must hold 33 uint32.  Returns the exclusive prefix; *block_total receives the sum.
__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t* smem, uint32_t* block_total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = (blockDim.x + 31) >> 5;
    const uint32_t inc = warp_inclusive_scan(v, lane);
    if (lane == 31) smem[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        uint32_t w = lane < nwarps ? smem[lane] : 0u;
        uint32_t winc = warp_inclusive_scan(w, lane);
        smem[lane] = winc - w;          // exclusive warp offsets
        if (lane == 31) smem[32] = winc;  // total
    }
    __syncthreads();
    const uint32_t res = smem[warp] + inc - v;
    *block_total = smem[32];
    __syncthreads();
    return res;
}

// out[i] = sum_{j<i} in[j]; *d_total (uint64, optional) = sum of all.  in may alias out.
// Tiled: per-tile sums, scan of the tile sums, per-tile scan (three launches; scratch in ctx->scan_b).
int scan_exclusive_u32(b2r_ctx* ctx, const uint32_t* d_in, uint32_t* d_out, int64_t n,
                       unsigned long long* d_total, cudaStream_t stream);

}  // namespace b2r
