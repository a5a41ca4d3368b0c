#include "scan.cuh"

namespace b2r {

constexpr int SCAN_THREADS = 1024;
constexpr int SCAN_ITEMS = 8;
constexpr int64_t SCAN_TILE = (int64_t)SCAN_THREADS * SCAN_ITEMS;

// Three launches: per-tile sums (one block per tile of 8192 items), exclusive scan of the tile sums (one block;
// 2^31 items are 262144 tiles), per-tile exclusive scan offset by the tile's prefix.  `in` may alias `out`:
// every block reads its own tile completely before it writes it.
__global__ void __launch_bounds__(SCAN_THREADS) scan_tile_sums_kernel(const uint32_t* __restrict__ in, int64_t n,
                                                                      uint32_t* __restrict__ tile_sums) {
    __shared__ uint32_t smem[33];
    const int64_t i0 = (int64_t)blockIdx.x * SCAN_TILE + (int64_t)threadIdx.x * SCAN_ITEMS;
    uint32_t s = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) s += (i0 + k < n) ? in[i0 + k] : 0u;
    uint32_t total;
    block_exclusive_scan(s, smem, &total);
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}

__global__ void __launch_bounds__(SCAN_THREADS) scan_tile_offsets_kernel(uint32_t* tile_sums, int64_t n_tiles,
                                                                         unsigned long long* total) {
    __shared__ uint32_t smem[33];
    unsigned long long carry = 0;
    for (int64_t base = 0; base < n_tiles; base += SCAN_THREADS) {
        const int64_t i = base + threadIdx.x;
        const uint32_t v = i < n_tiles ? tile_sums[i] : 0u;
        uint32_t t;
        const uint32_t excl = block_exclusive_scan(v, smem, &t);
        if (i < n_tiles) tile_sums[i] = (uint32_t)carry + excl;
        carry += t;
    }
    if (threadIdx.x == 0 && total) *total = carry;
}

__global__ void __launch_bounds__(SCAN_THREADS) scan_tiles_kernel(const uint32_t* in, uint32_t* out, int64_t n,
                                                                  const uint32_t* __restrict__ tile_offsets) {
    __shared__ uint32_t smem[33];
    const int64_t i0 = (int64_t)blockIdx.x * SCAN_TILE + (int64_t)threadIdx.x * SCAN_ITEMS;
    uint32_t v[SCAN_ITEMS];
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) v[k] = (i0 + k < n) ? in[i0 + k] : 0u;
    uint32_t s = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) s += v[k];
    uint32_t total;
    const uint32_t excl = block_exclusive_scan(s, smem, &total);   // (also the barrier between this tile's loads and stores)
    uint32_t run = tile_offsets[blockIdx.x] + excl;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) {
        if (i0 + k < n) out[i0 + k] = run;
        run += v[k];
    }
}

int scan_exclusive_u32(b2r_ctx* ctx, const uint32_t* d_in, uint32_t* d_out, int64_t n,
                       unsigned long long* d_total, cudaStream_t stream) {
    const int64_t n_tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    if (n_tiles == 0) {
        if (d_total) B2R_CUDA(cudaMemsetAsync(d_total, 0, sizeof(unsigned long long), stream));
        return 0;
    }
    if (ctx->scan_b.ensure((size_t)n_tiles * sizeof(uint32_t))) return -102;
    uint32_t* tiles = (uint32_t*)ctx->scan_b.ptr;
    scan_tile_sums_kernel<<<(int)n_tiles, SCAN_THREADS, 0, stream>>>(d_in, n, tiles);
    B2R_LAUNCH_CHECK(ctx);
    scan_tile_offsets_kernel<<<1, SCAN_THREADS, 0, stream>>>(tiles, n_tiles, d_total);
    B2R_LAUNCH_CHECK(ctx);
    scan_tiles_kernel<<<(int)n_tiles, SCAN_THREADS, 0, stream>>>(d_in, d_out, n, tiles);
    B2R_LAUNCH_CHECK(ctx);
    return 0;
}

}  // namespace b2r
