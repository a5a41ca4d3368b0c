// Context management + error reporting for the C ABI (include/b200recon.h).
#include <cstdarg>
#include <cstdlib>

#include "common.cuh"

namespace b2r {
static thread_local char g_err[1024] = "";
void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
}  // namespace b2r

extern "C" {

const char* b2r_last_error(void) { return b2r::g_err; }
int b2r_version(void) { return 100; }

int b2r_ctx_create(int device, b2r_ctx** out) {
    B2R_CHECK(out != nullptr, -1, "b2r_ctx_create: out is NULL");
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        b2r::set_error("b2r_ctx_create: no CUDA device available (%s); this library has no CPU fallback",
                       cudaGetErrorString(e));
        return -2;
    }
    B2R_CHECK(device >= 0 && device < n, -3, "b2r_ctx_create: device %d out of range [0,%d)", device, n);
    B2R_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    B2R_CUDA(cudaGetDeviceProperties(&prop, device));
    B2R_CHECK(prop.major == 10, -4, "b2r_ctx_create: device %d is sm_%d%d; this library is built for sm_100a only",
              device, prop.major, prop.minor);
    b2r_ctx* c = new b2r_ctx();
    c->device = device;
    c->sm_count = prop.multiProcessorCount;
    if (c->small.ensure(4096)) { delete c; return -102; }
    B2R_CUDA(cudaMemset(c->small.ptr, 0, c->small.bytes));
    {
        const char* e2 = getenv("B2R_ASYNC");
        c->async_items = !(e2 && e2[0] == '0');
    }
    {
        // The streaming kernels' stream gets the highest priority: when a streaming kernel ends, the next one (same
        // stream) and the preparation of the frame after it (caller's stream, waiting for the same frame slot) become
        // ready together, and whichever the hardware picks first delays the other -- with equal priorities it was the
        // preparation every other frame: 8.5 us between consecutive streaming kernels instead of 3.8 (timeline probe).
        int lo = 0, hi = 0;
        B2R_CUDA(cudaDeviceGetStreamPriorityRange(&lo, &hi));
        B2R_CUDA(cudaStreamCreateWithPriority(&c->side, cudaStreamNonBlocking, hi));
    }
    B2R_CUDA(cudaStreamCreateWithFlags(&c->copy, cudaStreamNonBlocking));
    for (auto& f : c->fslot) {
        B2R_CUDA(cudaEventCreateWithFlags(&f.ev_plan, cudaEventDisableTiming));
        B2R_CUDA(cudaEventCreateWithFlags(&f.ev_items, cudaEventDisableTiming));
        B2R_CUDA(cudaEventCreateWithFlags(&f.ev_h2d, cudaEventDisableTiming));
        B2R_CUDA(cudaEventCreateWithFlags(&f.ev_prep, cudaEventDisableTiming));
    }
    c->pinned_bytes = 1 << 16;
    B2R_CUDA(cudaMallocHost(&c->pinned, c->pinned_bytes));
    *out = c;
    return 0;
}

int b2r_ctx_destroy(b2r_ctx* c) {
    if (!c) return 0;
    cudaSetDevice(c->device);
    if (c->side) cudaStreamSynchronize(c->side);
    if (c->copy) cudaStreamSynchronize(c->copy);
    for (auto& f : c->fslot) {
        b2r::Scratch* fs[] = {&f.dm, &f.rgba, &f.mip, &f.hdr, &f.rec, &f.counters, &f.live, &f.raw_depth, &f.raw_color};
        for (auto* s : fs) s->release();
        if (f.h_stage) cudaFreeHost(f.h_stage);
        if (f.ev_plan) cudaEventDestroy(f.ev_plan);
        if (f.ev_items) cudaEventDestroy(f.ev_items);
        if (f.ev_h2d) cudaEventDestroy(f.ev_h2d);
        if (f.ev_prep) cudaEventDestroy(f.ev_prep);
    }
    if (c->side) cudaStreamDestroy(c->side);
    if (c->copy) cudaStreamDestroy(c->copy);
    b2r::Scratch* all[] = {&c->frame_raw_depth, &c->frame_raw_color, &c->frame_f32,
                           &c->small, &c->scan_a, &c->scan_b, &c->mc_code, &c->mc_flags, &c->mc_block_counts,
                           &c->mc_vert_id, &c->pc_block_counts, &c->icp_a, &c->icp_b, &c->icp_c, &c->icp_d,
                           &c->icp_e, &c->icp_f, &c->cl_a, &c->cl_b, &c->cl_c, &c->cl_d, &c->mc_live, &c->mc_cells, &c->trace};
    for (auto* s : all) s->release();
    for (auto e : c->ev_pool) cudaEventDestroy(e);
    if (c->pinned) cudaFreeHost(c->pinned);
    delete c;
    return 0;
}

int b2r_ctx_launch_count(b2r_ctx* c, int64_t* out) {
    B2R_CHECK(c && out, -1, "b2r_ctx_launch_count: NULL argument");
    *out = c->launches;
    return 0;
}

int b2r_ctx_kernel_timing(b2r_ctx* c, int enable) {
    B2R_CHECK(c, -1, "b2r_ctx_kernel_timing: NULL ctx");
    c->timing = enable != 0;
    if (enable) c->ev_used = 0;
    return 0;
}

int b2r_ctx_kernel_time(b2r_ctx* c, double* ms_total, int64_t* n_launches) {
    B2R_CHECK(c && ms_total && n_launches, -1, "b2r_ctx_kernel_time: NULL argument");
    B2R_CUDA(cudaSetDevice(c->device));
    double tot = 0.0;
    for (size_t i = 0; i + 1 < c->ev_used; i += 2) {
        B2R_CUDA(cudaEventSynchronize(c->ev_pool[i + 1]));
        float ms = 0.f;
        B2R_CUDA(cudaEventElapsedTime(&ms, c->ev_pool[i], c->ev_pool[i + 1]));
        tot += ms;
    }
    *ms_total = tot;
    *n_launches = (int64_t)(c->ev_used / 2);
    return 0;
}

int b2r_ctx_trace(b2r_ctx* c, int max_calls) {
    B2R_CHECK(c, -1, "b2r_ctx_trace: NULL ctx");
    B2R_CUDA(cudaSetDevice(c->device));
    c->trace_on = false;
    c->trace_n = 0;
    if (max_calls <= 0) return 0;
    if (c->trace.ensure((size_t)max_calls * 16 * sizeof(unsigned long long))) return -102;
    B2R_CUDA(cudaDeviceSynchronize());
    B2R_CUDA(cudaMemset(c->trace.ptr, 0, (size_t)max_calls * 16 * sizeof(unsigned long long)));
    c->trace_cap = max_calls;
    c->trace_on = true;
    return 0;
}

int b2r_ctx_trace_read(b2r_ctx* c, unsigned long long* out, int cap_calls, int* n_calls) {
    B2R_CHECK(c && out && n_calls, -1, "b2r_ctx_trace_read: NULL argument");
    B2R_CUDA(cudaSetDevice(c->device));
    B2R_CUDA(cudaDeviceSynchronize());
    const int n = c->trace_n < cap_calls ? c->trace_n : cap_calls;
    if (n > 0) B2R_CUDA(cudaMemcpy(out, c->trace.ptr, (size_t)n * 16 * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    for (int i = 0; i < n; ++i)
        for (int k = 0; k < 8; k += 2) out[i * 16 + k] = out[i * 16 + k] ? ~out[i * 16 + k] : 0ull;   // stored as max(~start)
    for (int i = 0; i < n; ++i) out[i * 16 + 9] = out[i * 16 + 9] ? ~out[i * 16 + 9] : 0ull;           // first dry warp, same form
    *n_calls = n;
    return 0;
}

int b2r_ctx_join(b2r_ctx* c, void* stream) {
    B2R_CHECK(c, -1, "b2r_ctx_join: NULL ctx");
    B2R_CUDA(cudaSetDevice(c->device));
    B2R_CUDA(b2r::join_side(c, b2r::S(stream)));
    return 0;
}

int b2r_ctx_updates_mirror(b2r_ctx* c, unsigned long long* out) {
    B2R_CHECK(c && out, -1, "b2r_ctx_updates_mirror: NULL argument");
    *out = *(volatile unsigned long long*)((char*)c->pinned + 1024);
    return 0;
}

int b2r_ctx_sync(b2r_ctx* c, void* stream) {
    B2R_CHECK(c, -1, "b2r_ctx_sync: NULL ctx");
    B2R_CUDA(cudaSetDevice(c->device));
    B2R_CUDA(b2r::join_side(c, b2r::S(stream)));
    B2R_CUDA(cudaStreamSynchronize(b2r::S(stream)));
    return 0;
}

}  // extern "C"
