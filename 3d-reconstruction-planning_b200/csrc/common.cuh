// Shared internals of libb200recon.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/b200recon.h"

namespace b2r {

void set_error(const char* fmt, ...);

#define B2R_CUDA(expr)                                                                          \
    do {                                                                                        \
        cudaError_t _e = (expr);                                                                \
        if (_e != cudaSuccess) {                                                                \
            b2r::set_error("%s:%d: %s failed: %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
            return -100;                                                                        \
        }                                                                                       \
    } while (0)

#define B2R_CHECK(cond, code, ...)                                                              \
    do {                                                                                        \
        if (!(cond)) { b2r::set_error(__VA_ARGS__); return (code); }                            \
    } while (0)

#define B2R_LAUNCH_CHECK(ctx)                                                                   \
    do {                                                                                        \
        (ctx)->launches++;                                                                      \
        cudaError_t _e = cudaGetLastError();                                                    \
        if (_e != cudaSuccess) {                                                                \
            b2r::set_error("%s:%d: kernel launch failed: %s", __FILE__, __LINE__, cudaGetErrorString(_e)); \
            return -101;                                                                        \
        }                                                                                       \
    } while (0)

// A grow-only device scratch buffer.
struct Scratch {
    void* ptr = nullptr;
    size_t bytes = 0;
    int ensure(size_t need) {
        if (need <= bytes) return 0;
        if (ptr) { cudaFree(ptr); ptr = nullptr; bytes = 0; }
        size_t want = need + (need >> 3) + 256;
        cudaError_t e = cudaMalloc(&ptr, want);
        if (e != cudaSuccess) { set_error("cudaMalloc(%zu) failed: %s", want, cudaGetErrorString(e)); return -102; }
        bytes = want;
        return 0;
    }
    void release() { if (ptr) cudaFree(ptr); ptr = nullptr; bytes = 0; }
};

}  // namespace b2r

// Everything one frame needs between its preparation and its items kernel.  Two slots alternate so that
// frame i+1 is prepared and planned on the caller's stream while frame i still streams on the side stream.
#ifndef B2R_FRAME_SLOTS
#define B2R_FRAME_SLOTS 3   // slots allocated; how many are cycled through is decided per call (integrate_prepared)
#endif
struct b2r_frame_slot {
    b2r::Scratch dm, rgba, mip;      // prepared frame: {depth, multiplier} pairs, colour words, max-depth pyramid
    b2r::Scratch hdr, rec, counters; // work items of the frame and their counters
    b2r::Scratch live;               // tiles that survived the coarse cull
    cudaEvent_t ev_plan = nullptr, ev_items = nullptr;
    bool busy = false;               // ev_items has been recorded
    // host frames (b2r_integrate_host): raw depth/colour land here from the context's copy stream, so the H2D copy of
    // frame i+1 runs while the caller's stream still prepares / plans frame i
    b2r::Scratch raw_depth, raw_color;
    cudaEvent_t ev_h2d = nullptr, ev_prep = nullptr;
    bool prep_recorded = false, h2d_recorded = false;
    void* h_stage = nullptr;         // pinned host staging of b2r_integrate_host_staged (pageable caller frames)
    size_t h_stage_bytes = 0;
    // The atomicMax levels of the pyramid, the frame maximum and the work-item counters must be zero before a frame is
    // prepared.  The items kernel leaves them zeroed when it is done with them (its last warp), so in steady state no
    // memset node sits between the kernels of consecutive frames; `clean` is false whenever that did not happen
    // (fresh buffers, a frame prepared for the unit path, an error between prepare and launch).
    bool clean = false;
};

struct b2r_ctx {
    int device = 0;
    int sm_count = 148;
    int64_t launches = 0;
    // frame staging (prepared frame: {depth_m, multiplier} float2 + packed RGBA)
    b2r::Scratch frame_raw_depth, frame_raw_color, frame_f32;
    b2r_frame_slot fslot[B2R_FRAME_SLOTS];
    int cur_slot = 0;
    cudaStream_t side = nullptr;     // items kernels run here (B2R_ASYNC=0: on the caller's stream)
    cudaStream_t copy = nullptr;     // H2D copies of b2r_integrate_host
    bool async_items = true;
    cudaEvent_t last_items = nullptr;   // completion of the newest items kernel on the side stream
    b2r::Scratch small;       // counters, per-frame scalars (zero-initialised on creation)
    b2r::Scratch scan_a, scan_b;
    // marching cubes
    b2r::Scratch mc_code, mc_flags, mc_block_counts, mc_vert_id, mc_live, mc_cells;
    int64_t mc_nv = 0, mc_nt = 0;
    int mc_R = 0, mc_zb = 0, mc_ze = 0;
    // Extraction is count -> emit against state kept here: the emit calls check that they see the SAME volume buffer
    // and that nothing of this context has written a volume since the count (write_gen is bumped by every integrate /
    // import / reset), else error -9 instead of silently stale ids and positions.
    const void* mc_vol = nullptr;
    const void* pc_vol = nullptr;
    uint64_t write_gen = 1, mc_gen = 0, pc_gen = 0;
    // point extraction
    b2r::Scratch pc_block_counts;
    int64_t pc_n = 0;
    // ICP
    b2r::Scratch icp_a, icp_b, icp_c, icp_d, icp_e, icp_f;
    // point-cloud preprocessing (cloud.cu): grid hash, sorted points, per-cell accumulators
    b2r::Scratch cl_a, cl_b, cl_c, cl_d;
    // optional CUDA-event timing of the dominant kernel
    bool timing = false;
    std::vector<cudaEvent_t> ev_pool;
    size_t ev_used = 0;
    void* pinned = nullptr;   // small pinned host buffer for result read-back
    size_t pinned_bytes = 0;
    // optional device-side timeline (b2r_ctx_trace): per integrate call 16 uint64 = {~start, end} of frame_prepare,
    // integrate_plan, integrate_items, integrate_cull, then (live tiles << 32 | work items) and spares, %globaltimer nanoseconds, written by the kernels themselves
    b2r::Scratch trace;
    int trace_cap = 0, trace_n = 0;
    bool trace_on = false;
};

namespace b2r {
// Orders `st` after every items kernel launched so far (they run on the context's side stream).
static inline cudaError_t join_side(b2r_ctx* ctx, cudaStream_t st) {
    return ctx->last_items ? cudaStreamWaitEvent(st, ctx->last_items, 0) : cudaSuccess;
}
// %globaltimer timeline hooks (no-ops when `tr` is NULL): start is kept as the maximum of ~t, i.e. the earliest block.
__device__ __forceinline__ unsigned long long gtimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ void trace_begin(unsigned long long* tr) { if (tr) atomicMax(tr, ~gtimer_ns()); }
__device__ __forceinline__ void trace_end(unsigned long long* tr) { if (tr) atomicMax(tr + 1, gtimer_ns()); }
static inline unsigned long long* trace_row(b2r_ctx* ctx, int kernel) {
    if (!ctx->trace_on || ctx->trace_n >= ctx->trace_cap) return nullptr;
    return (unsigned long long*)ctx->trace.ptr + (size_t)ctx->trace_n * 16 + kernel * 2;
}
static inline cudaStream_t S(void* s) { return (cudaStream_t)s; }
static inline int cdiv(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }
}  // namespace b2r
