// Batched ICP -- replaces Open3D pipelines::registration::RegistrationICP (stitcher.py:94-110; the
// north_star variant is TransformationEstimationPointToPlane; arithmetic in SURVEY.md A.5).
//
// Per pair, once:   a DENSE uniform grid over the target's bounding box (cell ~2.5 point spacings, at least r/16; one
//                   uint32 per cell = the end of the cell's run in the cell-sorted point array, so cell c is
//                   [E[c-1], E[c]) and a row of consecutive x cells is ONE contiguous run): count -> scan -> scatter.
//                   With 180 GB of HBM a few million cells per pair cost nothing, and a cell lookup is two adjacent
//                   loads instead of a 64-bit hash and a probe chain (round 1's open-addressing table).
// Per iteration:    TWO kernels.
//   icp_search_kernel      a thread per source point, small register footprint (occupancy hides the L2 latency): transform
//                          by the pair's current T (float64), walk the grid ring by ring around the query's cell -- every
//                          (dy, dz) row of a ring is one run of candidates, rows / end cells farther than the best
//                          distance so far are skipped from per-query face distances, the walk stops when the next ring
//                          cannot hold a closer point -- exact nearest target point within the radius (strict d^2 < r^2,
//                          ties -> lowest target index); writes the hit position and d^2.
//   icp_accumulate_kernel  a thread per 8 source points with room for the 29 float64 sums in registers (round 1's single
//                          kernel was capped at 40 registers for the search and spilled 4.9 KB per thread here): count,
//                          sum d^2 and the 27 normal-equation terms (21 upper J^T J + 6 J^T r for point-to-plane; centroid
//                          / cross-covariance sums for point-to-point; two residuals per correspondence for colored ICP),
//                          warp shuffles -> one shared-memory pass -> ONE atomic ticket per block; the last block of the
//                          pair sums the block partials in a fixed order (deterministic), evaluates Open3D's convergence
//                          test, solves the 6x6 system (LDL^T with diagonal pivoting) or the 3x3 SVD (Umeyama) and composes
//                          T <- update * T -- on the device, so the 2 (max_iteration + 1) launches run back to back with
//                          no host sync.
// Pairs are independent (blockIdx.y), so a batch fills the GPU; batches are distributed round-robin
// over GPUs by the host with no collective.
#include "frame.cuh"
#include "scan.cuh"

namespace b2r {

constexpr int ICP_THREADS = 256;
constexpr int ICP_ACC_PTS = 8;     // source points per thread of the accumulate kernel
constexpr int ICP_NACC = 29;       // [0] count, [1] sum d^2, [2..28] estimation sums
constexpr int ICP_MAX_RINGS = 16;  // cell size h >= r / ICP_MAX_RINGS (empty cells of the dense grid are cheap to skip)
constexpr long long ICP_MAX_CELLS = 1ll << 25;   // per pair (128 MB of uint32); coarser cells beyond that

struct IcpPair {
    long long src_off, n_src, tgt_off, n_tgt;
    long long cell_off;        // first cell of this pair in the batch's cell array
    int nx, ny, nz;            // grid dimensions (x fastest)
    int n_blocks;              // blocks of the accumulate kernel that hold points of this pair
    int rings;                 // K = ceil(r / h)
    double gmin[3];            // grid origin
    double h, inv_h;
    double bmin[3], bmax[3];   // target bounding box
    float feps;                // bound on |float32 (t - q) - exact (t - q)| for grid-relative float32 coordinates
};

struct IcpState {
    double T[16];
    double fitness, rmse, prev_fitness, prev_rmse;
    long long ncorr;
    int iter;        // number of updates applied so far
    int done;
    int converged;
    unsigned int ticket;
    double dbg[ICP_NACC + 6];
};

// ---- orderable encoding of doubles for atomicMin/atomicMax ---------------------------------------
__device__ __forceinline__ long long enc_double(double v) {
    long long b = __double_as_longlong(v);
    return b < 0 ? (b ^ 0x7fffffffffffffffll) : b;
}
__device__ __forceinline__ double dec_double(long long b) {
    return __longlong_as_double(b < 0 ? (b ^ 0x7fffffffffffffffll) : b);
}

__global__ void icp_bbox_init_kernel(long long* __restrict__ bb, int n_pairs) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_pairs * 6) bb[i] = ((i % 6) < 3) ? 0x7fffffffffffffffll : (long long)0x8000000000000000ull;
}

__global__ void __launch_bounds__(256) icp_bbox_kernel(const double* __restrict__ tgt, const IcpPair* __restrict__ pairs,
                                                       long long* __restrict__ bb) {
    const int p = blockIdx.y;
    const IcpPair P = pairs[p];
    double mn[3] = {1e300, 1e300, 1e300}, mx[3] = {-1e300, -1e300, -1e300};
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < P.n_tgt; i += (long long)gridDim.x * blockDim.x) {
        const double* q = tgt + (P.tgt_off + i) * 3;
#pragma unroll
        for (int k = 0; k < 3; ++k) { mn[k] = fmin(mn[k], q[k]); mx[k] = fmax(mx[k], q[k]); }
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            mn[k] = fmin(mn[k], __shfl_xor_sync(0xffffffffu, mn[k], o));
            mx[k] = fmax(mx[k], __shfl_xor_sync(0xffffffffu, mx[k], o));
        }
    }
    if ((threadIdx.x & 31) == 0 && mn[0] <= mx[0]) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            atomicMin(&bb[p * 6 + k], enc_double(mn[k]));
            atomicMax(&bb[p * 6 + 3 + k], enc_double(mx[k]));
        }
    }
}

// Host side of the grid setup: cell size from the radius and the point density, dimensions from the bounding box.
#ifndef B2R_ICP_WARM
#define B2R_ICP_WARM 1
#endif
#ifndef B2R_ICP_CELL_SPACINGS
#define B2R_ICP_CELL_SPACINGS 2.5
#endif
static void icp_grid_setup(IcpPair& P, const double* bmin, const double* bmax, double r) {
    double e[3];
    for (int k = 0; k < 3; ++k) {
        P.bmin[k] = P.n_tgt > 0 ? bmin[k] : 0.0;
        P.bmax[k] = P.n_tgt > 0 ? bmax[k] : 0.0;
        e[k] = P.bmax[k] - P.bmin[k];
    }
    // surface-like clouds: mean spacing ~ sqrt(largest bbox face / N)
    const double a = fmax(e[0] * e[1], fmax(e[1] * e[2], e[0] * e[2]));
    const double spacing = sqrt(fmax(a, 1e-300) / (double)(P.n_tgt > 0 ? P.n_tgt : 1));
    double h = fmax(B2R_ICP_CELL_SPACINGS * spacing, r / (double)ICP_MAX_RINGS);
    if (h > r) h = r;
    if (!(h > 0.0)) h = r;
    for (;;) {      // the grid covers the bounding box plus one cell on every side
        double cells = 1.0;
        for (int k = 0; k < 3; ++k) cells *= floor(e[k] / h) + 3.0;
        if (cells <= (double)ICP_MAX_CELLS) break;
        h *= 1.25;
    }
    P.h = h;
    P.inv_h = 1.0 / h;
    P.rings = (int)ceil(r / h - 1e-12);
    if (P.rings < 1) P.rings = 1;
    for (int k = 0; k < 3; ++k) P.gmin[k] = P.bmin[k] - h;
    P.nx = (int)floor(e[0] / h) + 3; P.ny = (int)floor(e[1] / h) + 3; P.nz = (int)floor(e[2] / h) + 3;
    // float32 pre-filter of the candidates: target and query coordinates relative to gmin are at most ext in magnitude,
    // each rounds to float32 with error <= ext * 2^-24, so a difference vector is off by <= sqrt(3) * 2 * ext * 2^-24
    // (~2.1e-7 ext); x2 for the float32 arithmetic on top
    const double ext = fmax(e[0], fmax(e[1], e[2])) + 2.0 * h + 2.0 * r;
    P.feps = (float)(4.2e-7 * ext);
}

// cell of a target point (inside the grid by construction; clamped against rounding at the box faces)
__device__ __forceinline__ uint32_t icp_cell_of(const IcpPair& P, const double* q) {
    int ix = (int)floor((q[0] - P.gmin[0]) * P.inv_h), iy = (int)floor((q[1] - P.gmin[1]) * P.inv_h),
        iz = (int)floor((q[2] - P.gmin[2]) * P.inv_h);
    ix = min(max(ix, 0), P.nx - 1); iy = min(max(iy, 0), P.ny - 1); iz = min(max(iz, 0), P.nz - 1);
    return (uint32_t)((iz * P.ny + iy) * P.nx + ix);
}

// count the target points per cell; remember each point's cell
__global__ void __launch_bounds__(256) icp_cell_count_kernel(const double* __restrict__ tgt, const IcpPair* __restrict__ pairs,
                                                             uint32_t* __restrict__ cells, uint32_t* __restrict__ cell_of) {
    const IcpPair P = pairs[blockIdx.y];
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P.n_tgt) return;
    const uint32_t c = icp_cell_of(P, tgt + (P.tgt_off + i) * 3);
    atomicAdd(&cells[P.cell_off + c], 1u);
    cell_of[P.tgt_off + i] = c;
}

// counting-sort scatter.  `cells` holds the exclusive scan of the counts (= every cell's start in the batch-wide sorted
// array) on entry; each point takes the next position of its cell, which leaves the cell's END in `cells` -- the form
// the search reads: cell c = [cells[c - 1], cells[c]), a run of consecutive cells = [cells[first - 1], cells[last]).
__global__ void __launch_bounds__(256) icp_cell_scatter_kernel(const double* __restrict__ tgt, const double* __restrict__ nrm,
                                                               const IcpPair* __restrict__ pairs, uint32_t* __restrict__ cells,
                                                               const uint32_t* __restrict__ cell_of,
                                                               double* __restrict__ s_pts, double* __restrict__ s_nrm,
                                                               int32_t* __restrict__ s_idx, float4* __restrict__ s_ptsf) {
    const IcpPair P = pairs[blockIdx.y];
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P.n_tgt) return;
    const long long g = P.tgt_off + i;
    const uint32_t pos = atomicAdd(&cells[P.cell_off + cell_of[g]], 1u);      // batch-wide sorted index
#pragma unroll
    for (int k = 0; k < 3; ++k) s_pts[(long long)pos * 4 + k] = tgt[g * 3 + k];
    s_pts[(long long)pos * 4 + 3] = __longlong_as_double((long long)i);      // the point's index in its cloud rides along
    if (nrm && s_nrm) {
#pragma unroll
        for (int k = 0; k < 3; ++k) s_nrm[(long long)pos * 3 + k] = nrm[g * 3 + k];
    }
    s_idx[pos] = (int32_t)i;
    s_ptsf[pos] = make_float4((float)(tgt[g * 3] - P.gmin[0]), (float)(tgt[g * 3 + 1] - P.gmin[1]), (float)(tgt[g * 3 + 2] - P.gmin[2]),
                              __int_as_float((int)i));
}

// ---- small dense solvers (one thread) ---------------------------------------------------------------
// Symmetric 6x6 solve, LDL^T with diagonal pivoting (largest remaining diagonal first, like Eigen's
// LDLT that Open3D's SolveJacobianSystemAndObtainExtrinsicMatrix calls).  Straight-line code: every
// array index is a compile-time constant after unrolling (the pivot swap is a chain of predicated
// exchanges), so the whole system lives in registers.
template <int K, int I>
__device__ __forceinline__ void ldlt6_swap(double (&A)[6][6], double (&b)[6], int (&perm)[6]) {
#pragma unroll
    for (int j = 0; j < 6; ++j) { const double t = A[K][j]; A[K][j] = A[I][j]; A[I][j] = t; }
#pragma unroll
    for (int j = 0; j < 6; ++j) { const double t = A[j][K]; A[j][K] = A[j][I]; A[j][I] = t; }
    { const double t = b[K]; b[K] = b[I]; b[I] = t; }
    { const int t = perm[K]; perm[K] = perm[I]; perm[I] = t; }
}

template <int K>
__device__ __forceinline__ void ldlt6_step(double (&A)[6][6], double (&b)[6], int (&perm)[6]) {
    int p = K;
    double best = fabs(A[K][K]);
#pragma unroll
    for (int i = K + 1; i < 6; ++i)
        if (fabs(A[i][i]) > best) { best = fabs(A[i][i]); p = i; }
    if (K + 1 < 6 && p == K + 1) ldlt6_swap<K, (K + 1 < 6 ? K + 1 : K)>(A, b, perm);
    if (K + 2 < 6 && p == K + 2) ldlt6_swap<K, (K + 2 < 6 ? K + 2 : K)>(A, b, perm);
    if (K + 3 < 6 && p == K + 3) ldlt6_swap<K, (K + 3 < 6 ? K + 3 : K)>(A, b, perm);
    if (K + 4 < 6 && p == K + 4) ldlt6_swap<K, (K + 4 < 6 ? K + 4 : K)>(A, b, perm);
    if (K + 5 < 6 && p == K + 5) ldlt6_swap<K, (K + 5 < 6 ? K + 5 : K)>(A, b, perm);
    const double d = A[K][K];
    if (d != 0.0) {
#pragma unroll
        for (int i = K + 1; i < 6; ++i) {
            const double l = A[i][K] / d;
#pragma unroll
            for (int j = K + 1; j < 6; ++j) A[i][j] -= l * A[K][j];
            A[i][K] = l;
        }
    } else {
#pragma unroll
        for (int i = K + 1; i < 6; ++i) A[i][K] = 0.0;
    }
}

__device__ void ldlt6_solve(const double* Ain, const double* bin, double* x) {
    double A[6][6], b[6];
    int perm[6];
#pragma unroll
    for (int i = 0; i < 6; ++i) {
        perm[i] = i;
        b[i] = bin[i];
#pragma unroll
        for (int j = 0; j < 6; ++j) A[i][j] = Ain[i * 6 + j];
    }
    ldlt6_step<0>(A, b, perm);
    ldlt6_step<1>(A, b, perm);
    ldlt6_step<2>(A, b, perm);
    ldlt6_step<3>(A, b, perm);
    ldlt6_step<4>(A, b, perm);
    ldlt6_step<5>(A, b, perm);
    double y[6], z[6];
#pragma unroll
    for (int i = 0; i < 6; ++i) {   // L y = P b
        double s = b[i];
#pragma unroll
        for (int j = 0; j < i; ++j) s -= A[i][j] * y[j];
        y[i] = s;
    }
#pragma unroll
    for (int i = 0; i < 6; ++i) y[i] = (A[i][i] != 0.0) ? y[i] / A[i][i] : 0.0;   // D
#pragma unroll
    for (int i = 5; i >= 0; --i) {  // L^T z = y
        double s = y[i];
#pragma unroll
        for (int j = i + 1; j < 6; ++j) s -= A[j][i] * z[j];
        z[i] = s;
    }
#pragma unroll
    for (int i = 0; i < 6; ++i) {   // x = P^T z
#pragma unroll
        for (int j = 0; j < 6; ++j)
            if (perm[i] == j) x[j] = z[i];
    }
}

// TransformVector6dToMatrix4d: R = Rz(x2) Ry(x1) Rx(x0), t = x[3:6]
__device__ void vec6_to_mat4(const double* x, double* T) {
    const double ca = cos(x[0]), sa = sin(x[0]), cb = cos(x[1]), sb = sin(x[1]), cg = cos(x[2]), sg = sin(x[2]);
    T[0] = cg * cb; T[1] = cg * sb * sa - sg * ca; T[2] = cg * sb * ca + sg * sa; T[3] = x[3];
    T[4] = sg * cb; T[5] = sg * sb * sa + cg * ca; T[6] = sg * sb * ca - cg * sa; T[7] = x[4];
    T[8] = -sb;     T[9] = cb * sa;                T[10] = cb * ca;               T[11] = x[5];
    T[12] = 0; T[13] = 0; T[14] = 0; T[15] = 1;
}

__device__ double det3(const double* M) {
    return M[0] * (M[4] * M[8] - M[5] * M[7]) - M[1] * (M[3] * M[8] - M[5] * M[6]) + M[2] * (M[3] * M[7] - M[4] * M[6]);
}

// Rotation of Umeyama's method from the 3x3 cross-covariance S (dst * src^T):  R = U diag(1,1,s) V^T.
// One-sided Jacobi SVD.
__device__ void umeyama_rotation(const double* Sin, double* Rm) {
    double A[9], V[9];
    for (int i = 0; i < 9; ++i) { A[i] = Sin[i]; V[i] = (i % 4 == 0) ? 1.0 : 0.0; }
    for (int sweep = 0; sweep < 60; ++sweep) {
        double off = 0.0;
        for (int p = 0; p < 2; ++p)
            for (int q = p + 1; q < 3; ++q) {
                double al = 0, be = 0, ga = 0;
                for (int i = 0; i < 3; ++i) { al += A[i * 3 + p] * A[i * 3 + p]; be += A[i * 3 + q] * A[i * 3 + q]; ga += A[i * 3 + p] * A[i * 3 + q]; }
                off = fmax(off, fabs(ga) / sqrt(fmax(al * be, 1e-300)));
                if (ga == 0.0) continue;
                const double zeta = (be - al) / (2.0 * ga);
                const double t = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
                const double c = 1.0 / sqrt(1.0 + t * t), s = c * t;
                for (int i = 0; i < 3; ++i) {
                    const double ap = A[i * 3 + p], aq = A[i * 3 + q];
                    A[i * 3 + p] = c * ap - s * aq; A[i * 3 + q] = s * ap + c * aq;
                    const double vp = V[i * 3 + p], vq = V[i * 3 + q];
                    V[i * 3 + p] = c * vp - s * vq; V[i * 3 + q] = s * vp + c * vq;
                }
            }
        if (off < 1e-15) break;
    }
    double S[3];
    for (int j = 0; j < 3; ++j) { double n = 0; for (int i = 0; i < 3; ++i) n += A[i * 3 + j] * A[i * 3 + j]; S[j] = sqrt(n); }
    int o0 = 0, o1 = 1, o2 = 2;  // sort singular values descending
    if (S[o0] < S[o1]) { const int t = o0; o0 = o1; o1 = t; }
    if (S[o1] < S[o2]) { const int t = o1; o1 = o2; o2 = t; }
    if (S[o0] < S[o1]) { const int t = o0; o0 = o1; o1 = t; }
    const int ord[3] = {o0, o1, o2};
    double U[9], Vs[9], Ss[3];
    for (int j = 0; j < 3; ++j) {
        Ss[j] = S[ord[j]];
        for (int i = 0; i < 3; ++i) { U[i * 3 + j] = Ss[j] > 1e-300 ? A[i * 3 + ord[j]] / Ss[j] : 0.0; Vs[i * 3 + j] = V[i * 3 + ord[j]]; }
    }
    if (Ss[0] <= 1e-300) { for (int i = 0; i < 9; ++i) Rm[i] = (i % 4 == 0) ? 1.0 : 0.0; return; }
    const double tiny = 1e-12 * Ss[0];
    if (Ss[1] <= tiny) {
        const double u0[3] = {U[0], U[3], U[6]};
        int k = 0;
        for (int i = 1; i < 3; ++i) if (fabs(u0[i]) < fabs(u0[k])) k = i;
        double e[3] = {0, 0, 0};
        e[k] = 1;
        double u1[3] = {u0[1] * e[2] - u0[2] * e[1], u0[2] * e[0] - u0[0] * e[2], u0[0] * e[1] - u0[1] * e[0]};
        const double n = sqrt(u1[0] * u1[0] + u1[1] * u1[1] + u1[2] * u1[2]);
        U[1] = u1[0] / n; U[4] = u1[1] / n; U[7] = u1[2] / n;
    }
    if (Ss[2] <= tiny) {
        const double u0[3] = {U[0], U[3], U[6]}, u1[3] = {U[1], U[4], U[7]};
        U[2] = u0[1] * u1[2] - u0[2] * u1[1]; U[5] = u0[2] * u1[0] - u0[0] * u1[2]; U[8] = u0[0] * u1[1] - u0[1] * u1[0];
    }
    const double sg = (det3(U) * det3(Vs) < 0) ? -1.0 : 1.0;
    for (int a = 0; a < 3; ++a)
        for (int b = 0; b < 3; ++b)
            Rm[a * 3 + b] = U[a * 3 + 0] * Vs[b * 3 + 0] + U[a * 3 + 1] * Vs[b * 3 + 1] + sg * U[a * 3 + 2] * Vs[b * 3 + 2];
}

__device__ void matmul4(const double* A, const double* B, double* C) {
    double t[16];
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) {
            double s = 0;
            for (int k = 0; k < 4; ++k) s += A[i * 4 + k] * B[k * 4 + j];
            t[i * 4 + j] = s;
        }
    for (int i = 0; i < 16; ++i) C[i] = t[i];
}

struct IcpArgs {
    const double* src;
    const double* s_pts;       // cell-sorted target points, 32 bytes each: x, y, z, original index (integer bits)
    const float4* s_ptsf;      // the same as float32 relative to the grid origin (pre-filter), .w = index bits
    const double* s_nrm;
    const int32_t* s_idx;
    const uint32_t* cells;     // per cell: END of its run in the sorted arrays (cells[-1] = 0 exists)
    int32_t* hit;              // per source point: sorted position of its correspondence, -1 = none
    double* hit_d2;            // and the squared distance
    const IcpPair* pairs;
    IcpState* states;
    double* partials;      // [pair][block][ICP_NACC]
    long long partial_stride;  // blocks reserved per pair
    int32_t* corr;         // optional [total source points]
    double r2;
    double rel_fitness, rel_rmse;
    int max_iter;
    int estimation;
    // colored ICP (estimation 2): source colours (original order), target intensity / colour gradient sorted by cell
    const double* src_col;
    const double* s_it;
    const double* s_grad;
    double sqrt_lg, sqrt_lp;   // sqrt(lambda_geometric), sqrt(1 - lambda_geometric)
};

// target intensity (r+g+b)/3 and colour gradient in cell-sorted order (s_idx maps sorted position -> point of the pair)
__global__ void __launch_bounds__(256) icp_gather_color_kernel(const double* __restrict__ tgt_col, const double* __restrict__ tgt_grad,
                                                               const IcpPair* __restrict__ pairs, const int32_t* __restrict__ s_idx,
                                                               double* __restrict__ s_it, double* __restrict__ s_grad) {
    const IcpPair P = pairs[blockIdx.y];
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P.n_tgt) return;
    const long long pos = P.tgt_off + i;                       // sorted positions of a pair occupy its own range
    const long long g = P.tgt_off + (long long)s_idx[pos];
    s_it[pos] = (tgt_col[g * 3] + tgt_col[g * 3 + 1] + tgt_col[g * 3 + 2]) / 3.0;
#pragma unroll
    for (int k = 0; k < 3; ++k) s_grad[pos * 3 + k] = tgt_grad[g * 3 + k];
}

// distance along one axis from the query to the cells `o` cells away from its own (em / ep: distance to the lower /
// upper face of its own cell)
__device__ __forceinline__ float icp_axis_dist(int o, float em, float ep, float h) {
    return o == 0 ? 0.0f : (o < 0 ? (float)(-o - 1) * h + em : (float)(o - 1) * h + ep);
}

// Candidates are pre-filtered in float32 (one 16-byte load, six float operations): grid-relative float32 coordinates
// give the squared distance to within the bound `feps` on the difference vector, so a candidate whose float32 distance
// exceeds sqrt(best) + feps (with a relative margin for the float32 arithmetic) cannot be the nearest neighbour nor tie
// with it.  Everything that passes is decided in float64 exactly as before: strict <, ties -> lowest target index.
struct IcpBest {
    double best;      // exact squared distance of the best candidate so far (r^2 before the first)
    float thr;        // float32 pass threshold derived from it
    int pos, j;
};
__device__ __forceinline__ float icp_thr(double best, float feps) {
    float sq;       // (MUFU approximation, 2 ulp: far inside the 1e-5 margins; the IEEE sqrtf sequence was 17 % of the kernel)
    asm("sqrt.approx.f32 %0, %1;" : "=f"(sq) : "f"((float)best));
    const float s = sq * 1.00001f + feps;
    return s * s * 1.00001f;
}
__device__ __forceinline__ void icp_exact(const IcpArgs& A, uint32_t m, double q0, double q1, double q2, float feps, IcpBest& B) {
    const double2 a = __ldg(reinterpret_cast<const double2*>(A.s_pts + (long long)m * 4));
    const double2 b = __ldg(reinterpret_cast<const double2*>(A.s_pts + (long long)m * 4 + 2));
    const int j = (int)__double_as_longlong(b.y);
    const double tx = a.x - q0, ty = a.y - q1, tz = b.x - q2;
    const double d2 = tx * tx + ty * ty + tz * tz;
    if (d2 < B.best || (d2 == B.best && B.pos >= 0 && j < B.j)) {
        B.best = d2; B.pos = (int)m; B.j = j;
        B.thr = icp_thr(d2, feps);
    }
}
// A run of candidates given as t -> m (IDX), four per turn with their loads in flight together (the serial load ->
// compare -> branch -> next load chain was what the kernel waited on).  The last turn repeats the final candidate,
// which changes nothing.
#define B2R_ICP_SCAN(TOT, IDX)                                                                   \
    for (uint32_t t = 0; t < (TOT); t += 4) {                                                    \
        uint32_t mm[4];                                                                          \
        float4 cc[4];                                                                            \
        _Pragma("unroll") for (int u = 0; u < 4; ++u) {                                          \
            const uint32_t tt = min(t + (uint32_t)u, (TOT) - 1u);                                \
            mm[u] = IDX(tt);                                                                     \
            cc[u] = __ldg(A.s_ptsf + mm[u]);                                                     \
        }                                                                                        \
        _Pragma("unroll") for (int u = 0; u < 4; ++u) {                                          \
            const float fx = cc[u].x - qf0, fy = cc[u].y - qf1, fz = cc[u].z - qf2;              \
            const float d2f = __fmaf_rn(fz, fz, __fmaf_rn(fy, fy, fx * fx));                     \
            if (d2f <= B.thr) icp_exact(A, mm[u], q0, q1, q2, feps, B);                          \
        }                                                                                        \
    }

// The search, a thread per source point.  What the first version of this kernel cost was not arithmetic but divergence:
// nested ring / row / trim / candidate loops with per-lane trip counts left 12.5 of 32 lanes active per instruction
// (profiles/r02_icp_search_4mm_ncu.txt: 578 M warp instructions per launch of 16 pairs).  Now rings 0 and 1 -- where
// almost every query ends -- are straight-line code: the own cell, then the three z planes of the 3x3x3 neighbourhood,
// each plane's (up to four) candidate runs walked by ONE flattened loop; rows and end cells that cannot hold a closer
// point are dropped from per-query face distances (float32 with a 1e-5 margin: conservative), the runs' ends are loaded
// together.  Rings >= 2 (queries without a neighbour within a cell size) keep the generic loop.
#ifndef B2R_ICP_SEARCH_MINB
#define B2R_ICP_SEARCH_MINB 5      // measured 3 / 4 / 5 / 6 blocks per SM (78 / 64 / 48 / 40 registers): 811 / 911 / 920 / 829 pairs per s at 4 mm, 301 / 365 / 377 / 341 at 5 cm
#endif
__global__ void __launch_bounds__(ICP_THREADS, B2R_ICP_SEARCH_MINB) icp_search_kernel(const __grid_constant__ IcpArgs A) {
    const int p = blockIdx.y;
    const IcpState& St = A.states[p];
    if (St.done) return;
    const IcpPair& P = A.pairs[p];
    const long long i = (long long)blockIdx.x * ICP_THREADS + threadIdx.x;
    if (i >= P.n_src) return;
    const double* s = A.src + (P.src_off + i) * 3;
    const double sx = s[0], sy = s[1], sz = s[2];
    const double q0 = St.T[0] * sx + St.T[1] * sy + St.T[2] * sz + St.T[3];
    const double q1 = St.T[4] * sx + St.T[5] * sy + St.T[6] * sz + St.T[7];
    const double q2 = St.T[8] * sx + St.T[9] * sy + St.T[10] * sz + St.T[11];
    const float feps = P.feps;
    IcpBest B{A.r2, icp_thr(A.r2, feps), -1, 0x7fffffff};
    const float qf0 = (float)(q0 - P.gmin[0]), qf1 = (float)(q1 - P.gmin[1]), qf2 = (float)(q2 - P.gmin[2]);
    // Warm start: after the first update the pose moves little, so the previous iteration's correspondence is (almost
    // always) still the nearest neighbour or close to it -- its distance bounds the search from the start, and every row /
    // cell / ring farther than that is dropped before anything is loaded.  The search stays exact: whatever is closer, or as
    // close with a lower index, still replaces it.
    if (B2R_ICP_WARM && St.iter > 0) {
        const int prev = A.hit[P.src_off + i];
        if (prev >= 0) icp_exact(A, (uint32_t)prev, q0, q1, q2, feps, B);
    }
    // reject queries farther than r from the target's bounding box
    double bd = 0.0;
    {
        const double qq[3] = {q0, q1, q2};
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const double d = qq[k] < P.bmin[k] ? P.bmin[k] - qq[k] : (qq[k] > P.bmax[k] ? qq[k] - P.bmax[k] : 0.0);
            bd += d * d;
        }
    }
    if (bd < A.r2 && P.n_tgt > 0) {
        const double h = P.h;
        const double g0 = (q0 - P.gmin[0]) * P.inv_h, g1 = (q1 - P.gmin[1]) * P.inv_h, g2 = (q2 - P.gmin[2]) * P.inv_h;
        const double fl0 = floor(g0), fl1 = floor(g1), fl2 = floor(g2);
        const int c0 = (int)fl0, c1 = (int)fl1, c2 = (int)fl2;       // may lie outside the grid (within r of the box)
        const int nx = P.nx, ny = P.ny, nz = P.nz;
        const uint32_t* E = A.cells + P.cell_off;
        // distances to the faces of the own cell, float32 (pruning only; every prune leaves a 1e-5 relative margin)
        const float hf = (float)h;
        const float em0 = (float)(g0 - fl0) * hf, ep0 = hf - em0, em1 = (float)(g1 - fl1) * hf, ep1 = hf - em1;
        const float em2 = (float)(g2 - fl2) * hf, ep2 = hf - em2;
        const float face = fminf(fminf(fminf(em0, ep0), fminf(em1, ep1)), fminf(em2, ep2));
        // ---- ring 0: the own cell ----
        if ((unsigned)c0 < (unsigned)nx && (unsigned)c1 < (unsigned)ny && (unsigned)c2 < (unsigned)nz) {
            const uint32_t* row = E + ((long long)c2 * ny + c1) * nx + c0;
            const uint32_t m0 = __ldg(row - 1), n0 = __ldg(row) - m0;
#define B2R_IDX0(T) (m0 + (T))
            B2R_ICP_SCAN(n0, B2R_IDX0)
#undef B2R_IDX0
        }
        // ---- ring 1: three planes, straight-line; skipped when nothing outside the own cell can be closer ----
        if (!(face * face * 0.99999f >= (float)B.best)) {
            const float dx2m = em0 * em0 * 0.99999f, dx2p = ep0 * ep0 * 0.99999f;
#pragma unroll
            for (int pz = 0; pz < 3; ++pz) {
                const int dz = pz == 0 ? 0 : (pz == 1 ? -1 : 1);           // the query's own plane first
                const int cz = c2 + dz;
                const float dz2 = dz == 0 ? 0.0f : (dz < 0 ? em2 * em2 : ep2 * ep2) * 0.99999f;
                const float thr = (float)B.best * 1.00001f + 1e-30f;
                const bool okz = (unsigned)cz < (unsigned)nz && dz2 <= thr;
                uint32_t ra[4], rn[4];          // up to four runs: start, length
#pragma unroll
                for (int py = 0; py < 3; ++py) {
                    const int dy = py - 1;
                    const int cy = c1 + dy;
                    const float row2 = dz2 + (dy == 0 ? 0.0f : (dy < 0 ? em1 * em1 : ep1 * ep1) * 0.99999f);
                    const bool ok = okz && (unsigned)cy < (unsigned)ny && row2 <= thr;
                    const bool centre = (dz == 0 && dy == 0);
                    const bool inl = ok && row2 + dx2m <= thr, inr = ok && row2 + dx2p <= thr;
                    const uint32_t* row = E + ((long long)(ok ? cz : 0) * ny + (ok ? cy : 0)) * nx;
                    if (!centre) {
                        int xa = inl ? c0 - 1 : c0, xb = inr ? c0 + 1 : c0;
                        xa = max(xa, 0); xb = min(xb, nx - 1);
                        const bool run = ok && xa <= xb;
                        const uint32_t a0 = run ? __ldg(row + xa - 1) : 0u, a1 = run ? __ldg(row + xb) : 0u;
                        ra[py] = a0; rn[py] = a1 - a0;
                        if (dz != 0 && py == 2) { ra[3] = 0u; rn[3] = 0u; }
                    } else {                                                // own cell is done: its two x neighbours
                        const bool rl = inl && c0 - 1 >= 0 && c0 - 1 < nx, rr = inr && c0 + 1 >= 0 && c0 + 1 < nx;
                        const uint32_t l0 = rl ? __ldg(row + c0 - 2) : 0u, l1 = rl ? __ldg(row + c0 - 1) : 0u;
                        const uint32_t r0 = rr ? __ldg(row + c0) : 0u, r1 = rr ? __ldg(row + c0 + 1) : 0u;
                        ra[1] = l0; rn[1] = l1 - l0;
                        ra[3] = r0; rn[3] = r1 - r0;
                    }
                }
                // one flattened loop over the plane's runs
                const uint32_t s1 = rn[0], s2 = s1 + rn[1], s3 = s2 + rn[2], tot = s3 + rn[3];
                const uint32_t o0 = ra[0], o1 = ra[1] - s1, o2 = ra[2] - s2, o3 = ra[3] - s3;
#define B2R_IDX4(T) ((T) + ((T) < s1 ? o0 : ((T) < s2 ? o1 : ((T) < s3 ? o2 : o3))))
                B2R_ICP_SCAN(tot, B2R_IDX4)
#undef B2R_IDX4
            }
        }
        // ---- rings >= 2 (rare): generic walk, a shell row is one run, an inner row contributes its two end cells ----
        for (int ring = 2; ring <= P.rings + 1; ++ring) {
            {
                const float lo = (float)(ring - 1) * hf + face;   // everything in ring >= `ring` is at least this far
                if (lo * lo * 0.99999f >= (float)B.best) break;
            }
            for (int dz = -ring; dz <= ring; ++dz) {
                const int cz = c2 + dz;
                if ((unsigned)cz >= (unsigned)nz) continue;
                const float ez = (float)icp_axis_dist(dz, em2, ep2, hf);
                for (int dy = -ring; dy <= ring; ++dy) {
                    const int cy = c1 + dy;
                    if ((unsigned)cy >= (unsigned)ny) continue;
                    const float ey = (float)icp_axis_dist(dy, em1, ep1, hf);
                    const float row2 = (ey * ey + ez * ez) * 0.99999f;
                    if (row2 > (float)B.best * 1.00001f + 1e-30f) continue;
                    const uint32_t* row = E + ((long long)cz * ny + cy) * nx;
                    const bool shell = (dz == -ring || dz == ring || dy == -ring || dy == ring);
                    const int nseg = shell ? 1 : 2;
                    for (int sg = 0; sg < nseg; ++sg) {
                        int xa, xb;
                        if (nseg == 1) { xa = -ring; xb = ring; }
                        else { xa = xb = (sg == 0 ? -ring : ring); }
                        // trim cells whose box is farther than the best distance so far from both ends of the run
                        const float thr = (float)B.best * 1.00001f + 1e-30f - row2;
                        while (xa <= xb) {
                            const float ex = (float)icp_axis_dist(xa, em0, ep0, hf);
                            if (ex * ex * 0.99999f <= thr) break;
                            ++xa;
                        }
                        while (xb >= xa) {
                            const float ex = (float)icp_axis_dist(xb, em0, ep0, hf);
                            if (ex * ex * 0.99999f <= thr) break;
                            --xb;
                        }
                        xa = max(xa + c0, 0);
                        xb = min(xb + c0, nx - 1);
                        if (xa > xb) continue;
                        const uint32_t mg = __ldg(row + xa - 1), ng = __ldg(row + xb) - mg;
#define B2R_IDXG(T) (mg + (T))
                        B2R_ICP_SCAN(ng, B2R_IDXG)
#undef B2R_IDXG
                    }
                }
            }
        }
    }
    if (A.corr) A.corr[P.src_off + i] = B.pos >= 0 ? B.j : -1;
    A.hit[P.src_off + i] = B.pos;
    A.hit_d2[P.src_off + i] = B.best;
}

template <int EST>
__global__ void __launch_bounds__(ICP_THREADS, 2) icp_accumulate_kernel(const __grid_constant__ IcpArgs A) {
    const int p = blockIdx.y;
    IcpState& St = A.states[p];
    if (St.done) return;
    const IcpPair P = A.pairs[p];
    if ((int)blockIdx.x >= P.n_blocks) return;
    __shared__ double s_acc[ICP_THREADS / 32][ICP_NACC];
    __shared__ int s_last;
    double T[12];
#pragma unroll
    for (int i = 0; i < 12; ++i) T[i] = St.T[i];
    const long long base = (long long)blockIdx.x * (ICP_THREADS * ICP_ACC_PTS);

    double acc[ICP_NACC];
#pragma unroll
    for (int i = 0; i < ICP_NACC; ++i) acc[i] = 0.0;
#pragma unroll 2
    for (int it = 0; it < ICP_ACC_PTS; ++it) {
        const long long i = base + (long long)it * ICP_THREADS + threadIdx.x;
        if (i >= P.n_src) break;
        const long long bestpos = (long long)A.hit[P.src_off + i];
        if (bestpos < 0) continue;
        const double best = A.hit_d2[P.src_off + i];
        const double* s = A.src + (P.src_off + i) * 3;
        const double sx = s[0], sy = s[1], sz = s[2];
        const double q[3] = {T[0] * sx + T[1] * sy + T[2] * sz + T[3], T[4] * sx + T[5] * sy + T[6] * sz + T[7],
                             T[8] * sx + T[9] * sy + T[10] * sz + T[11]};
        const double tx = A.s_pts[bestpos * 4], ty = A.s_pts[bestpos * 4 + 1], tz = A.s_pts[bestpos * 4 + 2];
        acc[0] += 1.0;
        acc[1] += best;
        if (EST == 2) {
            // TransformationEstimationForColoredICP: geometric + photometric residual per correspondence
            const double nt[3] = {A.s_nrm[bestpos * 3], A.s_nrm[bestpos * 3 + 1], A.s_nrm[bestpos * 3 + 2]};
            const double dit[3] = {A.s_grad[bestpos * 3], A.s_grad[bestpos * 3 + 1], A.s_grad[bestpos * 3 + 2]};
            const double dn = (q[0] - tx) * nt[0] + (q[1] - ty) * nt[1] + (q[2] - tz) * nt[2];
            const double* sc = A.src_col + (P.src_off + i) * 3;
            const double is = (sc[0] + sc[1] + sc[2]) / 3.0;
            const double vp[3] = {q[0] - dn * nt[0], q[1] - dn * nt[1], q[2] - dn * nt[2]};
            const double is0 = (dit[0] * (vp[0] - tx) + dit[1] * (vp[1] - ty) + dit[2] * (vp[2] - tz)) + A.s_it[bestpos];
            const double M[9] = {1.0 - nt[0] * nt[0], -nt[0] * nt[1], -nt[0] * nt[2], -nt[0] * nt[1], 1.0 - nt[1] * nt[1],
                                 -nt[1] * nt[2], -nt[0] * nt[2], -nt[1] * nt[2], 1.0 - nt[2] * nt[2]};
            double dm[3];
#pragma unroll
            for (int c = 0; c < 3; ++c) dm[c] = (-dit[0]) * M[c] + (-dit[1]) * M[3 + c] + (-dit[2]) * M[6 + c];
            double J[2][6], r[2];
            const double cg[3] = {q[1] * nt[2] - q[2] * nt[1], q[2] * nt[0] - q[0] * nt[2], q[0] * nt[1] - q[1] * nt[0]};
            const double cm[3] = {q[1] * dm[2] - q[2] * dm[1], q[2] * dm[0] - q[0] * dm[2], q[0] * dm[1] - q[1] * dm[0]};
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                J[0][k] = A.sqrt_lg * cg[k]; J[0][3 + k] = A.sqrt_lg * nt[k];
                J[1][k] = A.sqrt_lp * cm[k]; J[1][3 + k] = A.sqrt_lp * dm[k];
            }
            r[0] = A.sqrt_lg * dn;
            r[1] = A.sqrt_lp * (is - is0);
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                int k = 2;
#pragma unroll
                for (int a = 0; a < 6; ++a)
#pragma unroll
                    for (int b = a; b < 6; ++b) acc[k++] += J[e][a] * J[e][b];
#pragma unroll
                for (int a = 0; a < 6; ++a) acc[23 + a] += J[e][a] * r[e];
            }
        } else if (EST == 1) {
            const double nx = A.s_nrm[bestpos * 3], ny = A.s_nrm[bestpos * 3 + 1], nz = A.s_nrm[bestpos * 3 + 2];
            const double r = (q[0] - tx) * nx + (q[1] - ty) * ny + (q[2] - tz) * nz;
            const double J[6] = {q[1] * nz - q[2] * ny, q[2] * nx - q[0] * nz, q[0] * ny - q[1] * nx, nx, ny, nz};
            int k = 2;
#pragma unroll
            for (int a = 0; a < 6; ++a)
#pragma unroll
                for (int b = a; b < 6; ++b) acc[k++] += J[a] * J[b];
#pragma unroll
            for (int a = 0; a < 6; ++a) acc[23 + a] += J[a] * r;
        } else {
            acc[2] += q[0]; acc[3] += q[1]; acc[4] += q[2];
            acc[5] += tx; acc[6] += ty; acc[7] += tz;
            const double t3[3] = {tx, ty, tz};
#pragma unroll
            for (int a = 0; a < 3; ++a)
#pragma unroll
                for (int b = 0; b < 3; ++b) acc[8 + a * 3 + b] += t3[a] * q[b];
        }
    }

    // warp shuffle reduction -> one shared-memory pass -> block partial
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int i = 0; i < ICP_NACC; ++i) {
        double v = acc[i];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
        if (lane == 0) s_acc[warp][i] = v;
    }
    __syncthreads();
    double* mypart = A.partials + ((long long)p * A.partial_stride + blockIdx.x) * ICP_NACC;
    if (threadIdx.x < ICP_NACC) {
        double v = 0.0;
        for (int w = 0; w < ICP_THREADS / 32; ++w) v += s_acc[w][threadIdx.x];
        mypart[threadIdx.x] = v;
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned int t = atomicAdd(&St.ticket, 1u);   // the single atomic of this block
        s_last = (t == (unsigned int)P.n_blocks - 1u);
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();

    // ---- last block of the pair: fixed-order sum, convergence test, solve, compose ------------------
    __shared__ double s_sum[ICP_NACC];
    if (threadIdx.x < ICP_NACC) {
        double v = 0.0;
        const double* base_p = A.partials + (long long)p * A.partial_stride * ICP_NACC + threadIdx.x;
        for (int b = 0; b < P.n_blocks; ++b) v += __ldcg(base_p + (long long)b * ICP_NACC);
        s_sum[threadIdx.x] = v;
    }
    __syncthreads();
    if (threadIdx.x != 0) return;
    St.ticket = 0;
    const double K = s_sum[0];
    const double fitness = (K > 0 && P.n_src > 0) ? K / (double)P.n_src : 0.0;
    const double rmse = K > 0 ? sqrt(s_sum[1] / K) : 0.0;
    const int n = St.iter;
    St.prev_fitness = St.fitness; St.prev_rmse = St.rmse;
    St.fitness = fitness; St.rmse = rmse; St.ncorr = (long long)K;
    if (n >= 1 && fabs(St.prev_fitness - fitness) < A.rel_fitness && fabs(St.prev_rmse - rmse) < A.rel_rmse) {
        St.done = 1; St.converged = 1;
        return;
    }
    if (n == 0) for (int i = 0; i < ICP_NACC; ++i) St.dbg[i] = s_sum[i];
    if (n >= A.max_iter) { St.done = 1; return; }
    double upd[16];
    for (int i = 0; i < 16; ++i) upd[i] = (i % 5 == 0) ? 1.0 : 0.0;
    if (K > 0) {
        if (EST >= 1) {
            double JTJ[36], nb[6], x[6];
            int k = 2;
            for (int a = 0; a < 6; ++a)
                for (int b = a; b < 6; ++b) { JTJ[a * 6 + b] = s_sum[k]; JTJ[b * 6 + a] = s_sum[k]; ++k; }
            for (int a = 0; a < 6; ++a) nb[a] = -s_sum[23 + a];
            ldlt6_solve(JTJ, nb, x);
            if (n == 0) for (int i = 0; i < 6; ++i) St.dbg[ICP_NACC + i] = x[i];
            vec6_to_mat4(x, upd);
        } else {
            double ms[3], mt[3], Sg[9], Rm[9];
            for (int a = 0; a < 3; ++a) { ms[a] = s_sum[2 + a] / K; mt[a] = s_sum[5 + a] / K; }
            for (int a = 0; a < 3; ++a)
                for (int b = 0; b < 3; ++b) Sg[a * 3 + b] = s_sum[8 + a * 3 + b] / K - mt[a] * ms[b];
            umeyama_rotation(Sg, Rm);
            for (int a = 0; a < 3; ++a) {
                for (int b = 0; b < 3; ++b) upd[a * 4 + b] = Rm[a * 3 + b];
                upd[a * 4 + 3] = mt[a] - (Rm[a * 3 + 0] * ms[0] + Rm[a * 3 + 1] * ms[1] + Rm[a * 3 + 2] * ms[2]);
            }
        }
    }
    double Tn[16];
    matmul4(upd, St.T, Tn);
    for (int i = 0; i < 16; ++i) St.T[i] = Tn[i];
    St.iter = n + 1;
    __threadfence();
}

}  // namespace b2r

using namespace b2r;

static unsigned int next_pow2(unsigned long long v) {
    unsigned int p = 16;
    while ((unsigned long long)p < v && p < (1u << 30)) p <<= 1;
    return p;
}

extern "C" int b2r_icp_batch(b2r_ctx* ctx, int n_pairs, const double* d_src_points, const int64_t* h_src_off,
                             const double* d_tgt_points, const double* d_tgt_normals, const int64_t* h_tgt_off,
                             const double* d_src_colors, const double* d_tgt_colors, const double* d_tgt_color_grad,
                             const double* h_init, const b2r_icp_params* params, b2r_icp_result* h_results,
                             int32_t* d_corr, void* stream) {
    B2R_CHECK(ctx && h_src_off && h_tgt_off && params && h_results, -1, "b2r_icp_batch: NULL argument");
    B2R_CHECK(n_pairs >= 0, -5, "b2r_icp_batch: negative pair count");
    // Open3D RegistrationICP: LogError("Invalid max_correspondence_distance.")
    B2R_CHECK(params->max_correspondence_distance > 0.0, -10, "Invalid max_correspondence_distance.");
    B2R_CHECK(params->estimation >= 0 && params->estimation <= 2, -11,
              "b2r_icp_batch: unknown estimation %d (0 point-to-point, 1 point-to-plane, 2 colored)", params->estimation);
    // Open3D: "TransformationEstimationPointToPlane ... require pre-computed normal vectors for target PointCloud."
    B2R_CHECK(params->estimation == 0 || d_tgt_normals, -12,
              "TransformationEstimationPointToPlane and TransformationEstimationColoredICP require pre-computed "
              "normal vectors for target PointCloud.");
    // Open3D RegistrationColoredICP: LogError when either cloud has no colours; the gradient comes from b2r_color_gradient
    B2R_CHECK(params->estimation != 2 || (d_src_colors && d_tgt_colors && d_tgt_color_grad), -13,
              "ColoredICP requires source and target PointClouds with colors (and the target's colour gradient).");
    B2R_CHECK(params->estimation != 2 || (params->lambda_geometric >= 0.0 && params->lambda_geometric <= 1.0), -13,
              "ColoredICP: lambda_geometric must be in [0, 1]");
    if (n_pairs == 0) return 0;
    B2R_CUDA(cudaSetDevice(ctx->device));
    cudaStream_t st = S(stream);
    const long long Ns = h_src_off[n_pairs], Nt = h_tgt_off[n_pairs];
    B2R_CHECK((Ns == 0 || d_src_points) && (Nt == 0 || d_tgt_points), -1, "b2r_icp_batch: NULL point arrays");
    B2R_CHECK(Nt < (1ll << 31) && Ns < (1ll << 40), -8, "b2r_icp_batch: too many points");

    // ---- host-side pair table ----
    std::vector<IcpPair> hp(n_pairs);
    std::vector<IcpState> hs(n_pairs);
    int max_blocks = 1;
    long long max_nt = 0, max_ns = 0;
    for (int p = 0; p < n_pairs; ++p) {
        IcpPair& P = hp[p];
        memset(&P, 0, sizeof(P));
        P.src_off = h_src_off[p]; P.n_src = h_src_off[p + 1] - h_src_off[p];
        P.tgt_off = h_tgt_off[p]; P.n_tgt = h_tgt_off[p + 1] - h_tgt_off[p];
        B2R_CHECK(P.n_src >= 0 && P.n_tgt >= 0, -5, "b2r_icp_batch: offsets must be non-decreasing");
        P.n_blocks = (int)((P.n_src + ICP_THREADS * ICP_ACC_PTS - 1) / (ICP_THREADS * ICP_ACC_PTS));
        if (P.n_blocks < 1) P.n_blocks = 1;
        if (P.n_blocks > max_blocks) max_blocks = P.n_blocks;
        if (P.n_tgt > max_nt) max_nt = P.n_tgt;
        if (P.n_src > max_ns) max_ns = P.n_src;
        IcpState& Sx = hs[p];
        memset(&Sx, 0, sizeof(Sx));
        for (int i = 0; i < 16; ++i) Sx.T[i] = h_init ? h_init[p * 16 + i] : ((i % 5 == 0) ? 1.0 : 0.0);
    }

    // ---- scratch carving ----
    auto al = [](size_t v) { return (v + 255) & ~(size_t)255; };
    const size_t sz_pairs = al(sizeof(IcpPair) * n_pairs), sz_states = al(sizeof(IcpState) * n_pairs);
    const size_t sz_bb = al(sizeof(long long) * 6 * n_pairs);
    if (ctx->icp_a.ensure(sz_pairs + sz_states + sz_bb)) return -102;
    IcpPair* d_pairs = (IcpPair*)ctx->icp_a.ptr;
    IcpState* d_states = (IcpState*)((char*)ctx->icp_a.ptr + sz_pairs);
    long long* d_bb = (long long*)((char*)ctx->icp_a.ptr + sz_pairs + sz_states);
    const size_t sz_pts = al(sizeof(double) * 3 * (size_t)Nt), sz_pts4 = al(sizeof(double) * 4 * (size_t)Nt);
    const size_t sz_ptsf = al(sizeof(float4) * (size_t)Nt);
    if (ctx->icp_c.ensure(sz_pts4 + sz_pts + al(sizeof(int32_t) * (size_t)Nt) + al(sizeof(uint32_t) * (size_t)Nt) + sz_ptsf)) return -102;
    double* d_spts = (double*)ctx->icp_c.ptr;
    double* d_snrm = (double*)((char*)ctx->icp_c.ptr + sz_pts4);
    int32_t* d_sidx = (int32_t*)((char*)ctx->icp_c.ptr + sz_pts4 + sz_pts);
    uint32_t* d_cellof = (uint32_t*)((char*)ctx->icp_c.ptr + sz_pts4 + sz_pts + al(sizeof(int32_t) * (size_t)Nt));
    float4* d_sptsf = (float4*)((char*)ctx->icp_c.ptr + sz_pts4 + sz_pts + al(sizeof(int32_t) * (size_t)Nt) + al(sizeof(uint32_t) * (size_t)Nt));
    if (ctx->icp_d.ensure(sizeof(double) * ICP_NACC * (size_t)max_blocks * n_pairs)) return -102;
    double* d_partials = (double*)ctx->icp_d.ptr;
    if (ctx->icp_f.ensure(al(sizeof(int32_t) * (size_t)Ns) + al(sizeof(double) * (size_t)Ns) + 256)) return -102;
    int32_t* d_hit = (int32_t*)ctx->icp_f.ptr;
    double* d_hitd2 = (double*)((char*)ctx->icp_f.ptr + al(sizeof(int32_t) * (size_t)Ns));

    // ---- bounding boxes -> grids (one host round trip: the grid sizes decide the allocation) ----
    B2R_CUDA(cudaMemcpyAsync(d_pairs, hp.data(), sizeof(IcpPair) * n_pairs, cudaMemcpyHostToDevice, st));
    icp_bbox_init_kernel<<<cdiv(n_pairs * 6, 128), 128, 0, st>>>(d_bb, n_pairs);
    B2R_LAUNCH_CHECK(ctx);
    if (max_nt > 0) {
        const int bx = (int)std::min<long long>(cdiv(max_nt, 256), 256);
        icp_bbox_kernel<<<dim3(bx, n_pairs), 256, 0, st>>>(d_tgt_points, d_pairs, d_bb);
        B2R_LAUNCH_CHECK(ctx);
    }
    std::vector<long long> hbb((size_t)n_pairs * 6);
    B2R_CUDA(cudaMemcpyAsync(hbb.data(), d_bb, sizeof(long long) * 6 * n_pairs, cudaMemcpyDeviceToHost, st));
    B2R_CUDA(cudaStreamSynchronize(st));
    long long cell_total = 0;
    for (int p = 0; p < n_pairs; ++p) {
        double bmin[3], bmax[3];
        for (int k = 0; k < 3; ++k) {
            const long long lo = hbb[(size_t)p * 6 + k], hi = hbb[(size_t)p * 6 + 3 + k];
            const long long dl = lo < 0 ? (lo ^ 0x7fffffffffffffffll) : lo, dh = hi < 0 ? (hi ^ 0x7fffffffffffffffll) : hi;
            memcpy(&bmin[k], &dl, 8);
            memcpy(&bmax[k], &dh, 8);
        }
        icp_grid_setup(hp[p], bmin, bmax, params->max_correspondence_distance);
        hp[p].cell_off = cell_total;
        cell_total += (long long)hp[p].nx * hp[p].ny * hp[p].nz;
    }
    B2R_CHECK(cell_total < (1ll << 33), -8, "b2r_icp_batch: grids too large (%lld cells)", cell_total);
    if (ctx->icp_b.ensure(sizeof(uint32_t) * (size_t)(cell_total + 64))) return -102;
    uint32_t* d_cells = (uint32_t*)ctx->icp_b.ptr + 32;       // d_cells[-1] = 0: the start of the very first cell
    B2R_CUDA(cudaMemcpyAsync(d_pairs, hp.data(), sizeof(IcpPair) * n_pairs, cudaMemcpyHostToDevice, st));
    B2R_CUDA(cudaMemcpyAsync(d_states, hs.data(), sizeof(IcpState) * n_pairs, cudaMemcpyHostToDevice, st));
    // the host vectors die at return: make sure the copies are done before that (pageable memory
    // copies are staged synchronously by the runtime, this sync is for clarity)
    B2R_CUDA(cudaStreamSynchronize(st));
    B2R_CUDA(cudaMemsetAsync(ctx->icp_b.ptr, 0, sizeof(uint32_t) * (size_t)(cell_total + 64), st));

    // ---- build the target grids ----
    if (max_nt > 0) {
        dim3 grid(cdiv(max_nt, 256), n_pairs);
        icp_cell_count_kernel<<<grid, 256, 0, st>>>(d_tgt_points, d_pairs, d_cells, d_cellof);
        B2R_LAUNCH_CHECK(ctx);
        if (int rc = scan_exclusive_u32(ctx, d_cells, d_cells, cell_total, nullptr, st)) return rc;
        icp_cell_scatter_kernel<<<grid, 256, 0, st>>>(d_tgt_points, d_tgt_normals, d_pairs, d_cells, d_cellof, d_spts, d_snrm,
                                                      d_sidx, d_sptsf);
        B2R_LAUNCH_CHECK(ctx);
    }

    double *d_sit = nullptr, *d_sgrad = nullptr;
    if (params->estimation == 2 && Nt > 0) {
        if (ctx->icp_e.ensure(al(sizeof(double) * (size_t)Nt) + al(sizeof(double) * 3 * (size_t)Nt))) return -102;
        d_sit = (double*)ctx->icp_e.ptr;
        d_sgrad = (double*)((char*)ctx->icp_e.ptr + al(sizeof(double) * (size_t)Nt));
        if (max_nt > 0) {
            icp_gather_color_kernel<<<dim3(cdiv(max_nt, 256), n_pairs), 256, 0, st>>>(d_tgt_colors, d_tgt_color_grad, d_pairs, d_sidx,
                                                                                      d_sit, d_sgrad);
            B2R_LAUNCH_CHECK(ctx);
        }
    }

    // ---- iterate: max_iteration + 1 evaluations, no host sync in between ----
    IcpArgs A;
    A.src_col = d_src_colors; A.s_it = d_sit; A.s_grad = d_sgrad;
    A.sqrt_lg = sqrt(params->lambda_geometric); A.sqrt_lp = sqrt(1.0 - params->lambda_geometric);
    A.src = d_src_points; A.s_pts = d_spts; A.s_nrm = d_snrm; A.s_idx = d_sidx; A.s_ptsf = d_sptsf; A.cells = d_cells; A.hit = d_hit;
    A.hit_d2 = d_hitd2; A.pairs = d_pairs; A.states = d_states; A.partials = d_partials; A.partial_stride = max_blocks;
    A.corr = d_corr; A.r2 = params->max_correspondence_distance * params->max_correspondence_distance;
    A.rel_fitness = params->relative_fitness; A.rel_rmse = params->relative_rmse;
    A.max_iter = params->max_iteration < 0 ? 0 : params->max_iteration; A.estimation = params->estimation;
    const dim3 sgrid(cdiv(std::max<long long>(max_ns, 1), ICP_THREADS), n_pairs), agrid(max_blocks, n_pairs);
    for (int it = 0; it <= A.max_iter; ++it) {
        icp_search_kernel<<<sgrid, ICP_THREADS, 0, st>>>(A);
        B2R_LAUNCH_CHECK(ctx);
        if (A.estimation == 2) icp_accumulate_kernel<2><<<agrid, ICP_THREADS, 0, st>>>(A);
        else if (A.estimation == 1) icp_accumulate_kernel<1><<<agrid, ICP_THREADS, 0, st>>>(A);
        else icp_accumulate_kernel<0><<<agrid, ICP_THREADS, 0, st>>>(A);
        B2R_LAUNCH_CHECK(ctx);
    }
    B2R_CUDA(cudaMemcpyAsync(hs.data(), d_states, sizeof(IcpState) * n_pairs, cudaMemcpyDeviceToHost, st));
    B2R_CUDA(cudaStreamSynchronize(st));
    if (getenv("B2R_ICP_DEBUG")) for (int i = 0; i < ICP_NACC + 6; ++i) fprintf(stderr, "dbg[%d]=%.17g\n", i, hs[0].dbg[i]);
    for (int p = 0; p < n_pairs; ++p) {
        b2r_icp_result& Rr = h_results[p];
        for (int i = 0; i < 16; ++i) Rr.transformation[i] = hs[p].T[i];
        Rr.fitness = hs[p].fitness; Rr.inlier_rmse = hs[p].rmse; Rr.n_correspondences = hs[p].ncorr;
        Rr.iterations = hs[p].iter; Rr.converged = hs[p].converged;
    }
    return 0;
}
