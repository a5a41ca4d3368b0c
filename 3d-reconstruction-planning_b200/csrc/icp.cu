#include "common.cuh"
extern "C" int b2r_icp_batch(b2r_ctx* ctx, int n_pairs, const double* d_src_points, const int64_t* h_src_off,
                  const double* d_tgt_points, const double* d_tgt_normals, const int64_t* h_tgt_off,
                  const double* d_src_colors, const double* d_tgt_colors, const double* d_tgt_color_grad,
                  const double* h_init, const b2r_icp_params* params,
                  b2r_icp_result* h_results, int32_t* d_corr, void* stream) {
    b2r::set_error("b2r_icp_batch: not implemented yet");
    return -99;
}
