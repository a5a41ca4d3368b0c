/*
 * b200recon.h -- C ABI of libb200recon.so: the B200 (sm_100a) implementation of the RGB-D
 * reconstruction hot path that the reference's stitcher.py hands to Open3D 0.18.0.
 *
 * Every entry point names the reference interface it replaces.  "stitcher.py:NN" is the call
 * site in /root/reference; "O3D" is the Open3D 0.18.0 legacy API the call resolves to (the
 * arithmetic is restated in SURVEY.md Appendix A; Open3D sources are not in /root/reference).
 *
 * Conventions
 *   - extern "C", plain pointers and sizes; no C++/torch types.
 *   - every function returns 0 on success, <0 on error; b2r_last_error() gives the message
 *     (thread-local).  No exceptions cross the ABI.
 *   - `stream` is a cudaStream_t passed as void* (NULL = the legacy default stream).
 *   - pointers named d_* are DEVICE pointers owned by the caller (e.g. tensor.data_ptr());
 *     h_* are HOST pointers.  The library owns only the scratch inside a b2r_ctx.
 *   - calls are asynchronous on `stream` unless documented as blocking.
 *   - transforms are row-major double[16]; `extrinsic` = world->camera as in Open3D.
 *
 * Volume memory layout (device, chosen for coalescing on B200 -- NOT Open3D's AoS):
 *   slab of planes x in [x_begin, x_end) of an R x R x R cube, stored x-slowest, y-fastest (Open3D's own index is
 *   x*R*R + y*R + z: x-slowest too) with B2R_PAD_LO halo planes below and B2R_PAD_HI above:
 *       plane p = x - x_begin + B2R_PAD_LO,   index = (p * R + z) * R + y
 *   A slab owns WHOLE (x, y) columns: Open3D's float32 z-chain of a column and the per-column planning of a frame run
 *   once per column whatever the number of slabs, and the z steps of a column are R*8 bytes apart.
 *   tsdfw : float2 {tsdf, weight} per voxel (8 B);
 *   color : uint32 packed 10:10:10 fixed point (value*4, R in bits 0-9, G 10-19, B 20-29).
 */
#ifndef B200RECON_H
#define B200RECON_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2R_PAD_LO 1
#define B2R_PAD_HI 2

typedef struct b2r_ctx b2r_ctx;

/* O3D camera.PinholeCameraIntrinsic(width,height,fx,fy,cx,cy) -- stitcher.py:264-267 */
typedef struct b2r_intrinsic {
    int32_t width, height;
    double fx, fy, cx, cy;
} b2r_intrinsic;

/* O3D pipelines.integration.UniformTSDFVolume(length, resolution, sdf_trunc, color_type, origin)
 * color_type: 0 NoColor, 1 RGB8, 2 Gray32 (TSDFVolumeColorType). */
typedef struct b2r_volume {
    int32_t resolution;        /* R */
    int32_t x_begin, x_end;    /* planes owned by this slab (whole cube: 0, R) */
    int32_t color_type;
    double voxel_length;       /* length / resolution */
    double sdf_trunc;
    double origin[3];
    void* d_tsdfw;             /* float2 [(x_end-x_begin+PAD_LO+PAD_HI) * R * R] */
    void* d_color;             /* uint32, same count; NULL when color_type == 0 */
    void* d_rows;              /* optional uint8 [(x_end-x_begin+PAD_LO+PAD_HI) * R]: 1 = row (plane x, z) may hold a voxel
                                * with tsdf < 0 (a surface needs one: an active cube and a surface point both have a
                                * negative corner).  Maintained by the library (integrate marks the rows of work items
                                * that integrated a negative sample, reset clears, import sets all); halo planes travel
                                * with the volume's halo planes.  Extraction reads the volume only next to marked rows;
                                * NULL = no information, every row is read. */
} b2r_volume;

/* ---- library / context ------------------------------------------------------------------ */
const char* b2r_last_error(void);
int b2r_version(void);
/* One context per (process, device, logical stream of work).  Owns scratch buffers. */
int b2r_ctx_create(int device, b2r_ctx** out);
int b2r_ctx_destroy(b2r_ctx* ctx);
/* Number of kernels this context has launched so far (bench.py "gpu_launches"). */
int b2r_ctx_launch_count(b2r_ctx* ctx, int64_t* out);
/* b2r_integrate* return after queueing the frame's preparation and planning on `stream`; the streaming kernel
 * itself runs on a stream owned by the context so that the NEXT frame can be prepared meanwhile.  Every entry
 * point of this library that reads or writes a volume orders itself after that work; code that touches the
 * volume's buffers directly (e.g. an NCCL send of a halo plane) calls b2r_ctx_join first: it makes `stream` wait
 * (on the device, the host does not block) for every integration queued so far.  B2R_ASYNC=0 in the
 * environment keeps everything on the caller's stream. */
int b2r_ctx_join(b2r_ctx* ctx, void* stream);
int b2r_ctx_sync(b2r_ctx* ctx, void* stream); /* join + blocking */
/* Host mirror of the update counter passed to the most recent b2r_integrate* call as d_updates: the library
 * writes it (8 bytes, device -> mapped pinned host memory, by the last warp of each frame's streaming kernel:
 * no copy-engine operation between consecutive kernels).  Not blocking: valid for a frame once that frame has
 * completed (b2r_ctx_sync). */
int b2r_ctx_updates_mirror(b2r_ctx* ctx, unsigned long long* out);
/* Per-kernel timing of the dominant kernel (tsdf_integrate) with CUDA events recorded on the launching
 * stream around each launch: enable=1 starts (and clears) collection, enable=0 stops it.
 * b2r_ctx_kernel_time is blocking: total milliseconds and number of launches collected. */
int b2r_ctx_kernel_timing(b2r_ctx* ctx, int enable);
int b2r_ctx_kernel_time(b2r_ctx* ctx, double* ms_total, int64_t* n_launches);
/* Device-side timeline of the integrate calls (tuning / profiles/ evidence; the kernels write %globaltimer
 * themselves, nothing is added to the streams).  b2r_ctx_trace(ctx, max_calls > 0) starts recording the next
 * max_calls b2r_integrate* calls (blocking: synchronises the device), max_calls = 0 stops.  b2r_ctx_trace_read
 * (blocking) copies, per recorded call, 16 uint64: nanosecond stamps {start, end} of frame_prepare_kernel,
 * integrate_plan_kernel, integrate_items_kernel, integrate_cull_kernel (0 = kernel did not run), then
 * (tiles that survived the cull << 32 | work items of the frame) and 7 spare zeros. */
int b2r_ctx_trace(b2r_ctx* ctx, int max_calls);
int b2r_ctx_trace_read(b2r_ctx* ctx, unsigned long long* out, int cap_calls, int* n_calls);

/* ---- images / back-projection (stitcher.py:34-46) --------------------------------------- */
/* O3D Image.ConvertDepthToFloatImage(depth_scale, depth_trunc) inside
 * RGBDImage.create_from_color_and_depth (stitcher.py:38-42).
 * dtype: 0 uint8 (v/255), 1 uint16, 2 float32.  d_out: float32 [H*W] metres. */
int b2r_depth_to_float(b2r_ctx* ctx, const void* d_depth, int dtype, int H, int W, double depth_scale,
                       double depth_trunc, float* d_out, void* stream);
/* O3D PointCloud.create_from_rgbd_image(rgbd, intrinsic, extrinsic) (stitcher.py:45-46).
 * d_depth: float32 metres; d_color: uint8 [H,W,nc] (nc 3|4) or NULL.  Output in row-major pixel
 * order with invalid (d<=0) pixels dropped: d_points/d_colors double [<=H*W,3].
 * Blocking: returns the point count in *n_points. */
int b2r_backproject(b2r_ctx* ctx, const float* d_depth, const uint8_t* d_color, int nc, const b2r_intrinsic* K,
                    const double* extrinsic, double* d_points, double* d_colors, int64_t* n_points, void* stream);

/* b2r_backproject without any allocation -- NO counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same inputs, same points/colours in the same order as b2r_backproject; differences:
 *  - no context: `device` is the CUDA device ordinal (made current by the call);
 *  - the scratch (per-block counts -> offsets) is a file-scope static __device__ array of 64 KB that every GPU loading
 *    the library holds until exit; images of more than b2r_backproject_static_max_pixels() = 4 194 304 pixels (2^22) are
 *    rejected with -7 and a message in b2r_last_error();
 *  - non-blocking: the point count is written to *d_n_points (int64 in DEVICE memory, caller's) on `stream`;
 *  - calls no cudaMalloc / cudaFree / allocating routine; the first call on a device looks the array's address up
 *    (cudaGetSymbolAddress), later calls only launch three kernels.
 * All calls on a device share the scratch, which keeps the last call's contents (nothing in it needs clearing: each call
 * overwrites what it reads): calls that OVERLAP on the GPU (different streams, no ordering between them) race. */
int b2r_backproject_static(int device, const float* d_depth, const uint8_t* d_color, int nc, const b2r_intrinsic* K,
                           const double* extrinsic, double* d_points, double* d_colors, int64_t* d_n_points,
                           void* stream);
int b2r_backproject_static_max_pixels(int64_t* out);

/* ---- TSDF integration (O3D UniformTSDFVolume.integrate(rgbd, intrinsic, extrinsic)) ------ */
/* d_depth: float32 metres [H,W] (RGBDImage.depth); d_color: uint8 [H,W,3] (RGB8) or float32 [H,W]
 * (Gray32) or NULL (NoColor).  d_updates (optional, uint64 on device) is incremented by the number
 * of voxels updated. */
int b2r_integrate(b2r_ctx* ctx, const b2r_volume* vol, const float* d_depth, const void* d_color,
                  const b2r_intrinsic* K, const double* extrinsic, unsigned long long* d_updates, void* stream);
/* Same, end to end from HOST frame buffers: H2D copy + depth conversion (depth_scale/depth_trunc as
 * in stitcher.py:40-41) + integrate.  h_depth dtype as in b2r_depth_to_float.  Host buffers must be PINNED
 * (cudaMallocHost / torch pin_memory) and must stay valid and unmodified until the copy has completed: the copy
 * runs on a stream owned by the context and `stream` is ordered after it, so any later synchronisation of `stream`
 * (an event recorded on it after this call, b2r_ctx_sync) covers it.  Pageable memory: use the _staged variant. */
int b2r_integrate_host(b2r_ctx* ctx, const b2r_volume* vol, const void* h_depth, int depth_dtype,
                       const uint8_t* h_color, const b2r_intrinsic* K, const double* extrinsic, double depth_scale,
                       double depth_trunc, unsigned long long* d_updates, void* stream);
/* Same for PAGEABLE host frames (numpy arrays from cv2.imread / np.load as stitcher.py:219-220,250-251 produce them,
 * or geometry.Image buffers): the frame is first copied into a pinned staging buffer owned by the context (one per
 * frame slot), so the H2D transfer is an asynchronous DMA overlapping the previous frame's kernels, and the caller's
 * buffers may be freed or overwritten as soon as the call returns. */
int b2r_integrate_host_staged(b2r_ctx* ctx, const b2r_volume* vol, const void* h_depth, int depth_dtype,
                              const uint8_t* h_color, const b2r_intrinsic* K, const double* extrinsic, double depth_scale,
                              double depth_trunc, unsigned long long* d_updates, void* stream);
/* O3D pipelines.integration.ScalableTSDFVolume.integrate(rgbd, intrinsic, extrinsic) -- SURVEY.md section 8 row f4
 * / A.3 last bullet (Open3D 0.18 ScalableTSDFVolume::Integrate; no call site in the reference, named by
 * north_star next to UniformTSDFVolume).  Open3D's hash of unit_resolution^3-voxel units is laid out at its place
 * in a DENSE backing cube (`vol`, whole cube, origin = unit_offset * unit_length): a unit that was never opened
 * reads weight 0 like a missing hash entry, so extraction is b2r_extract_mesh_* / b2r_extract_points_* unchanged.
 * Per call: depth samples at `stride` are back-projected with camera_pose (= extrinsic^-1, float64, inverted by
 * the caller), every unit between LocateVolumeUnit(p - sdf_trunc) and LocateVolumeUnit(p + sdf_trunc) is stamped
 * with frame_id (>= 1, increasing) and integrated once with Open3D's per-unit arithmetic (column start from the
 * unit's own origin, float32 z-chain restarted per unit).
 *   d_unit_stamp  int32 [(R/unit_resolution)^3], zero before the first frame / after reset: 0 = unit not opened
 *   d_unit_list   int32 scratch of the same size (the frame's touched units)
 *   d_counters    uint64 [4] device: [0] units touched by this call, [1] unit touches that fell outside the
 *                 cube (running total -- Open3D would have opened those units), [2] units opened so far
 * No host synchronisation; everything runs on `stream` behind any b2r_integrate* still streaming. */
int b2r_integrate_units(b2r_ctx* ctx, const b2r_volume* vol, const float* d_depth, const void* d_color,
                        const b2r_intrinsic* K, const double* extrinsic, const double* camera_pose, int unit_resolution,
                        int stride, const int32_t* unit_offset, int32_t* d_unit_stamp, int32_t frame_id,
                        int32_t* d_unit_list, unsigned long long* d_counters, unsigned long long* d_updates,
                        void* stream);
/* Device self-test (blocking): the integrate kernel computes Open3D's IEEE float32 divisions with a shared
 * refined reciprocal; this runs n pseudo-random cases (projection-like, full-range, running-mean-like and
 * near-half-way quotients) through that code and through the `/` operator; mismatches[4] = count per family. */
int b2r_selftest_div(b2r_ctx* ctx, int64_t n, uint32_t seed, int64_t* mismatches, void* stream);
/* Device self-test (blocking): chain_jump (csrc/tsdf_integrate.cu, compiled out by default) advances Open3D's float32
 * chains `pt_camera += col(2)` n steps without walking them (mantissa arithmetic inside a binade, real steps
 * across binade edges / ties);
 * this runs n pseudo-random and adversarial (start, increment, steps) cases through that code and through the
 * stepped chain; *mismatches = cases whose bits differ. */
int b2r_selftest_chain(b2r_ctx* ctx, int64_t n, uint32_t seed, int64_t* mismatches, void* stream);
/* O3D UniformTSDFVolume.reset() */
int b2r_volume_reset(b2r_ctx* ctx, const b2r_volume* vol, void* stream);
/* Export owned planes in Open3D order [x][y][z] as separate float arrays (tests / persistence);
 * d_color_out: float32 [R,R,nz,3] in 0..255 scale or NULL. */
int b2r_volume_export(b2r_ctx* ctx, const b2r_volume* vol, float* d_tsdf_out, float* d_weight_out,
                      float* d_color_out, void* stream);
/* Inverse of b2r_volume_export (load a persisted slab / build test volumes); colour in 0..255 scale. */
int b2r_volume_import(b2r_ctx* ctx, const b2r_volume* vol, const float* d_tsdf_in, const float* d_weight_in,
                      const float* d_color_in, void* stream);

/* ---- marching cubes (O3D UniformTSDFVolume.extract_triangle_mesh()) ---------------------- */
/* Pass 1 (blocking): classify + count + scan.  Halo planes [x_begin-1, x_end] must be valid when the
 * slab is not the whole cube.  Returns this slab's vertex and triangle counts. */
int b2r_extract_mesh_count(b2r_ctx* ctx, const b2r_volume* vol, int64_t* n_vertices, int64_t* n_triangles,
                           void* stream);
/* Pass 2a: write vertex positions/colours (double [n_vertices,3]) and this slab's global vertex ids
 * (vertex_base = number of vertices on lower slabs).  d_vert_id: uint32 [(nz+PAD_LO+PAD_HI)*R*R*... ] is
 * owned by the context; b2r_mesh_seam_* give access to the seam plane for the NCCL exchange. */
int b2r_extract_mesh_vertices(b2r_ctx* ctx, const b2r_volume* vol, int64_t vertex_base, double* d_vertices,
                              double* d_colors, void* stream);
/* Seam exchange: device pointer + byte size of (a) the first owned plane of the vertex-id table (to
 * send to the slab below) and (b) the pad plane above the slab (to receive from the slab above). */
int b2r_mesh_seam_buffers(b2r_ctx* ctx, const b2r_volume* vol, void** d_send_first_plane, void** d_recv_top_plane,
                          int64_t* bytes);
/* Pass 2b: write triangles (int32 [n_triangles,3], global vertex ids, winding as Open3D). */
int b2r_extract_mesh_triangles(b2r_ctx* ctx, const b2r_volume* vol, int32_t* d_triangles, void* stream);

/* ---- O3D UniformTSDFVolume.extract_point_cloud() ------------------------------------------ */
int b2r_extract_points_count(b2r_ctx* ctx, const b2r_volume* vol, int64_t* n_points, void* stream); /* blocking */
int b2r_extract_points_emit(b2r_ctx* ctx, const b2r_volume* vol, double* d_points, double* d_normals,
                            double* d_colors, void* stream);
/* ScalableTSDFVolume.extract_point_cloud() on the dense backing cube of b2r_integrate_units (after
 * b2r_extract_points_count on the same volume): positions are interpolated in WORLD coordinates from the unit's own
 * origin + in-unit index, as Open3D's ScalableTSDFVolume::ExtractPointCloud does (its formula is not translation
 * invariant: (p0*r1 + p1*r0) / float(r0 + r1)), and the normals use ScalableTSDFVolume::GetTSDFAt (0 when the unit
 * of the sample's base voxel was never opened: d_unit_stamp == 0). */
int b2r_extract_points_emit_units(b2r_ctx* ctx, const b2r_volume* vol, int unit_resolution, const int32_t* unit_offset,
                                  const int32_t* d_unit_stamp, double* d_points, double* d_normals, double* d_colors,
                                  void* stream);

/* ---- ICP (O3D pipelines.registration.registration_icp, stitcher.py:94-110) ---------------- */
/* Batched: pair p uses source points [src_off[p], src_off[p+1]) and target points
 * [tgt_off[p], tgt_off[p+1]) of the concatenated arrays.
 * estimation: 0 TransformationEstimationPointToPoint, 1 ...PointToPlane, 2 ...ForColoredICP.
 * criteria = ICPConvergenceCriteria(relative_fitness, relative_rmse, max_iteration). */
typedef struct b2r_icp_params {
    double max_correspondence_distance;
    double relative_fitness, relative_rmse;
    int32_t max_iteration;
    int32_t estimation;
    double lambda_geometric;  /* colored ICP only (0.968) */
} b2r_icp_params;

typedef struct b2r_icp_result {  /* O3D RegistrationResult */
    double transformation[16];
    double fitness;
    double inlier_rmse;
    int64_t n_correspondences;
    int32_t iterations;
    int32_t converged;
} b2r_icp_result;

/* Blocking (returns results on the host).  d_corr (optional): int32 [total source points], target
 * index local to the pair or -1 (RegistrationResult.correspondence_set).
 * d_src_colors / d_tgt_colors / d_tgt_color_grad are used by estimation 2 only. */
int b2r_icp_batch(b2r_ctx* ctx, int n_pairs, const double* d_src_points, const int64_t* h_src_off,
                  const double* d_tgt_points, const double* d_tgt_normals, const int64_t* h_tgt_off,
                  const double* d_src_colors, const double* d_tgt_colors, const double* d_tgt_color_grad,
                  const double* h_init /* [n_pairs*16] */, const b2r_icp_params* params,
                  b2r_icp_result* h_results, int32_t* d_corr, void* stream);

/* ---- point-cloud preprocessing around ICP ---------------------- */
/* Neighbour search modes: O3D geometry.KDTreeSearchParamKNN(knn) / KDTreeSearchParamHybrid(radius, max_nn)
 * (stitcher.py:63,69).  Searches are exact (uniform-grid hash, ring walk with distance pruning); neighbours are
 * ordered by (squared distance, index) like a k-d tree result, the query point included; max_nn <= 128. */
#define B2R_SEARCH_KNN 0
#define B2R_SEARCH_HYBRID 2

/* O3D PointCloud.voxel_down_sample(voxel_size) -- stitcher.py:58,151.  Cell = floor((p - (min_bound -
 * voxel_size/2)) / voxel_size); outputs the per-cell means of points / normals / colours (d_normals, d_colors
 * and their outputs may be NULL), cells ordered by their first point index (Open3D: unordered_map order).  The sums
 * run in ascending point order per cell, like Open3D's sequential loop: the means are bit-identical to it.
 * Output buffers hold up to n points.  Blocking: *h_n_out = number of cells.
 * Errors: voxel_size <= 0 ("[VoxelDownSample] voxel_size <= 0."), voxel_size too small for the extent. */
int b2r_voxel_down_sample(b2r_ctx* ctx, const double* d_points, const double* d_normals, const double* d_colors,
                          int64_t n, double voxel_size, double* d_out_points, double* d_out_normals,
                          double* d_out_colors, int64_t* h_n_out, void* stream);

/* O3D PointCloud.estimate_normals(search_param, fast_normal_computation=True) -- stitcher.py:61-63.
 * Covariance of the neighbours (cumulants in search order) -> eigenvector of the smallest eigenvalue
 * (closed-form FastEigen3x3); fewer than 3 neighbours or a zero vector -> (0,0,1); d_prev_normals (optional):
 * pre-existing normals the result is flipped to agree with.  d_out may alias d_prev_normals. */
int b2r_estimate_normals(b2r_ctx* ctx, const double* d_points, int64_t n, int search_mode, double radius,
                         int max_nn, const double* d_prev_normals, double* d_out_normals, void* stream);

/* O3D PointCloud.remove_statistical_outlier(nb_neighbors, std_ratio) -- stitcher.py:158-159.  Mean distance to
 * the nb_neighbors nearest points (query included) per point; keeps points with 0 < mean < cloud mean +
 * std_ratio * sample std.  d_indices_out: int64 [n], kept indices ascending.  Blocking: *h_n_out = kept. */
int b2r_statistical_outlier(b2r_ctx* ctx, const double* d_points, int64_t n, int nb_neighbors, double std_ratio,
                            int64_t* d_indices_out, int64_t* h_n_out, void* stream);

/* O3D InitializePointCloudForColoredICP(target, KDTreeSearchParamHybrid(radius, max_nn)) -- the per-point colour
 * gradient registration_colored_icp (stitcher.py:94-102) precomputes on the target.  d_grad_out: double [n,3]. */
int b2r_color_gradient(b2r_ctx* ctx, const double* d_points, const double* d_normals, const double* d_colors,
                       int64_t n, double radius, int max_nn, double* d_grad_out, void* stream);

/* Neighbour lists of the search above, for tests: d_idx_out int32 [n_query, k] (-1 padded), d_d2_out double
 * [n_query, k].  Queries must lie inside the cloud's bounding box (they are cloud points in every caller). */
int b2r_knn_search(b2r_ctx* ctx, const double* d_points, int64_t n, const double* d_queries, int64_t n_query,
                   int search_mode, double radius, int k, int32_t* d_idx_out, double* d_d2_out, void* stream);

/* ---- synthetic frame source (replaces bullet_camera.py:48-85 / data_recorder.py:104-121) -- */
typedef struct b2r_scene {
    float c2w[12];            /* camera->world 3x4 row-major */
    float fx, fy, cx, cy;
    int32_t W, H;
    float ell_c[3], ell_r[3]; /* ellipsoid (ell_r[0] <= 0: none) */
    float plane_z;
    int32_t has_plane;
    float checker;
    float depth_scale;
    int32_t n_tris;           /* > 0: mesh mode (d_tris float32 [n_tris,9]) */
    float mesh_min[3], mesh_max[3];
} b2r_scene;
int b2r_raycast_scene(b2r_ctx* ctx, const b2r_scene* scene, const float* d_tris, uint16_t* d_depth,
                      uint8_t* d_color, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200RECON_H */
