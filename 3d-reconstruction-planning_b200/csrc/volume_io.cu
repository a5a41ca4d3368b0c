// Volume export / import between the device slab layout and Open3D's [x][y][z] order
// (tests, persistence of volume slabs -- SURVEY.md section 5 "checkpoint / resume").
#include "common.cuh"

namespace b2r {

__global__ void volume_export_kernel(const float2* __restrict__ tsdfw, const uint32_t* __restrict__ color, int R,
                                     int nz, int color_type, float* __restrict__ out_tsdf,
                                     float* __restrict__ out_w, float* __restrict__ out_color) {
    const int64_t n = (int64_t)nz * R * R;
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    const int y = (int)(t % R);
    const int z = (int)((t / R) % R);
    const int xl = (int)(t / ((int64_t)R * R));                     // x plane of the slab (nz = planes owned)
    const int64_t in = ((int64_t)(xl + B2R_PAD_LO) * R + z) * R + y;
    const int64_t out = ((int64_t)xl * R + y) * R + z;              // Open3D's [x][y][z], x restricted to the slab
    const float2 q = tsdfw[in];
    if (out_tsdf) out_tsdf[out] = q.x;
    if (out_w) out_w[out] = q.y;
    if (out_color && color_type != 0) {
        const uint32_t c = color[in];
        if (color_type == 1) {
            out_color[out * 3 + 0] = (float)(c & 1023u) * 0.25f;
            out_color[out * 3 + 1] = (float)((c >> 10) & 1023u) * 0.25f;
            out_color[out * 3 + 2] = (float)((c >> 20) & 1023u) * 0.25f;
        } else {
            const float g = __uint_as_float(c);
            out_color[out * 3 + 0] = g; out_color[out * 3 + 1] = g; out_color[out * 3 + 2] = g;
        }
    }
}

__global__ void volume_import_kernel(float2* __restrict__ tsdfw, uint32_t* __restrict__ color, int R, int nz,
                                     int color_type, const float* __restrict__ in_tsdf,
                                     const float* __restrict__ in_w, const float* __restrict__ in_color) {
    const int64_t n = (int64_t)nz * R * R;
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    const int y = (int)(t % R);
    const int z = (int)((t / R) % R);
    const int xl = (int)(t / ((int64_t)R * R));
    const int64_t dst = ((int64_t)(xl + B2R_PAD_LO) * R + z) * R + y;
    const int64_t src = ((int64_t)xl * R + y) * R + z;
    tsdfw[dst] = make_float2(in_tsdf[src], in_w[src]);
    if (color && color_type != 0 && in_color) {
        if (color_type == 1) {
            uint32_t q[3];
            for (int k = 0; k < 3; ++k) {
                float v = in_color[src * 3 + k] * 4.0f + 0.5f;
                v = fminf(fmaxf(v, 0.0f), 1023.0f);
                q[k] = (uint32_t)v;
            }
            color[dst] = q[0] | (q[1] << 10) | (q[2] << 20);
        } else {
            color[dst] = __float_as_uint(in_color[src * 3]);
        }
    }
}

}  // namespace b2r

using namespace b2r;

extern "C" {

int b2r_volume_export(b2r_ctx* ctx, const b2r_volume* vol, float* d_tsdf_out, float* d_weight_out,
                      float* d_color_out, void* stream) {
    B2R_CHECK(ctx && vol && vol->d_tsdfw, -1, "b2r_volume_export: NULL argument");
    B2R_CUDA(cudaSetDevice(ctx->device));
    B2R_CUDA(join_side(ctx, S(stream)));   // items kernels of earlier integrate calls run on the side stream
    const int nz = vol->x_end - vol->x_begin, R = vol->resolution;
    const int64_t n = (int64_t)nz * R * R;
    if (n == 0) return 0;
    volume_export_kernel<<<cdiv(n, 256), 256, 0, S(stream)>>>((const float2*)vol->d_tsdfw,
                                                              (const uint32_t*)vol->d_color, R, nz, vol->color_type,
                                                              d_tsdf_out, d_weight_out, d_color_out);
    B2R_LAUNCH_CHECK(ctx);
    return 0;
}

int b2r_volume_import(b2r_ctx* ctx, const b2r_volume* vol, const float* d_tsdf_in, const float* d_weight_in,
                      const float* d_color_in, void* stream) {
    B2R_CHECK(ctx && vol && vol->d_tsdfw && d_tsdf_in && d_weight_in, -1, "b2r_volume_import: NULL argument");
    ctx->write_gen++;
    B2R_CUDA(cudaSetDevice(ctx->device));
    B2R_CUDA(join_side(ctx, S(stream)));   // items kernels of earlier integrate calls run on the side stream
    const int nz = vol->x_end - vol->x_begin, R = vol->resolution;
    const int64_t n = (int64_t)nz * R * R;
    if (n == 0) return 0;
    volume_import_kernel<<<cdiv(n, 256), 256, 0, S(stream)>>>((float2*)vol->d_tsdfw, (uint32_t*)vol->d_color, R, nz,
                                                              vol->color_type, d_tsdf_in, d_weight_in, d_color_in);
    B2R_LAUNCH_CHECK(ctx);
    if (vol->d_rows)     // imported content: every owned row may hold weights (halo rows keep what they had)
        B2R_CUDA(cudaMemsetAsync((char*)vol->d_rows + (size_t)B2R_PAD_LO * R, 1, (size_t)nz * R, S(stream)));
    return 0;
}

}  // extern "C"
