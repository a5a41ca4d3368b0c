#include "scan.cuh"

namespace b2r {

constexpr int SCAN_THREADS = 1024;
constexpr int SCAN_ITEMS = 8;

__global__ void __launch_bounds__(SCAN_THREADS) scan_exclusive_u32_kernel(const uint32_t* __restrict__ in,
                                                                           uint32_t* out, int64_t n,
                                                                           unsigned long long* total) {
    __shared__ uint32_t smem[33];
    unsigned long long carry = 0;
    const int64_t tile = (int64_t)SCAN_THREADS * SCAN_ITEMS;
    for (int64_t base = 0; base < n; base += tile) {
        const int64_t i0 = base + (int64_t)threadIdx.x * SCAN_ITEMS;
        uint32_t v[SCAN_ITEMS];
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS; ++k) v[k] = (i0 + k < n) ? in[i0 + k] : 0u;
        uint32_t s = 0;
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS; ++k) s += v[k];
        uint32_t tile_total;
        uint32_t excl = block_exclusive_scan(s, smem, &tile_total);
        uint32_t run = (uint32_t)carry + excl;
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS; ++k) {
            if (i0 + k < n) out[i0 + k] = run;
            run += v[k];
        }
        carry += tile_total;
    }
    if (threadIdx.x == 0 && total) *total = carry;
}

int scan_exclusive_u32(b2r_ctx* ctx, const uint32_t* d_in, uint32_t* d_out, int64_t n,
                       unsigned long long* d_total, cudaStream_t stream) {
    scan_exclusive_u32_kernel<<<1, SCAN_THREADS, 0, stream>>>(d_in, d_out, n, d_total);
    B2R_LAUNCH_CHECK(ctx);
    return 0;
}

}  // namespace b2r
