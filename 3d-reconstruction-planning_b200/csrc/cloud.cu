// Point-cloud preprocessing around ICP -- SURVEY.md 8(f) rows f1/f2/f3:
//   b2r_voxel_down_sample      O3D PointCloud::VoxelDownSample            stitcher.py:58,151      (A.7)
//   b2r_estimate_normals       O3D PointCloud::EstimateNormals            stitcher.py:61-63       (A.8)
//   b2r_statistical_outlier    O3D PointCloud::RemoveStatisticalOutliers  stitcher.py:158-159     (A.9)
//   b2r_color_gradient         O3D InitializePointCloudForColoredICP      stitcher.py:94-102      (A.6)
//
// One spatial structure serves all of them: a uniform grid held in an open-addressing hash table keyed by the
// 3x21-bit cell coordinate, the points counting-sorted by cell (count -> scan -> scatter), as in icp.cu.
// Neighbour searches are EXACT: a thread walks the cells around its query ring by ring, keeps the k best
// candidates in a max-heap ordered by (squared distance, index), and stops once no unvisited cell can hold a
// closer point -- the result equals a k-d tree's SearchKNN / SearchHybrid (strict d^2 < r^2), ties broken by the
// lower index.  All arithmetic is float64 like Open3D's; the library is built with -fmad=false.
#include <cfloat>

#include "frame.cuh"
#include "scan.cuh"

namespace b2r {

constexpr int KCAP = 128;                      // largest supported max_nn / knn
constexpr unsigned long long CL_EMPTY = 0xffffffffffffffffull;

struct CloudGrid {            // device-resident description of one grid (written by grid_setup_kernel)
    double gmin[3];           // grid origin
    double h, inv_h;          // cell size
    double bmin[3], bmax[3];  // bounding box of the points
    int cmax[3];              // largest cell coordinate that holds a point
    int bad;                  // cell coordinates would not fit 21 bits
};

struct GridView {
    const CloudGrid* g;
    const unsigned long long* keys;
    const uint32_t* starts;
    const uint32_t* counts;
    const double* s_pts;      // points sorted by cell
    const int32_t* s_idx;     // their original indices
    unsigned int tab_mask;
};

__device__ __forceinline__ long long cl_enc(double v) {
    long long b = __double_as_longlong(v);
    return b < 0 ? (b ^ 0x7fffffffffffffffll) : b;
}
__device__ __forceinline__ double cl_dec(long long b) { return __longlong_as_double(b < 0 ? (b ^ 0x7fffffffffffffffll) : b); }
__device__ __forceinline__ unsigned long long cl_key(int ix, int iy, int iz) {
    return ((unsigned long long)(unsigned)ix << 42) | ((unsigned long long)(unsigned)iy << 21) | (unsigned long long)(unsigned)iz;
}
__device__ __forceinline__ unsigned int cl_hash(unsigned long long k) {
    k ^= k >> 33; k *= 0xff51afd7ed558ccdull; k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ull; k ^= k >> 33;
    return (unsigned int)k;
}

__global__ void cl_bbox_init_kernel(long long* __restrict__ bb) {
    if (threadIdx.x < 6) bb[threadIdx.x] = threadIdx.x < 3 ? 0x7fffffffffffffffll : (long long)0x8000000000000000ull;
}

__global__ void __launch_bounds__(256) cl_bbox_kernel(const double* __restrict__ pts, long long n, long long* __restrict__ bb) {
    double mn[3] = {DBL_MAX, DBL_MAX, DBL_MAX}, mx[3] = {-DBL_MAX, -DBL_MAX, -DBL_MAX};
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
#pragma unroll
        for (int k = 0; k < 3; ++k) { const double v = pts[i * 3 + k]; mn[k] = fmin(mn[k], v); mx[k] = fmax(mx[k], v); }
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
        for (int k = 0; k < 3; ++k) { atomicMin(&bb[k], cl_enc(mn[k])); atomicMax(&bb[3 + k], cl_enc(mx[k])); }
    }
}

// One thread.  voxel_size > 0: the down-sampling grid of Open3D (origin = min_bound - voxel_size/2, cell =
// voxel_size).  Otherwise a search grid: cell size from the point density (surface-like clouds: mean spacing
// ~ sqrt(largest bbox face / n)) so that the k nearest neighbours usually sit within one ring, capped by the radius.
__global__ void cl_grid_setup_kernel(CloudGrid* __restrict__ G, const long long* __restrict__ bb, long long n,
                                     double voxel_size, double radius, int k) {
    double e[3];
    for (int a = 0; a < 3; ++a) { G->bmin[a] = cl_dec(bb[a]); G->bmax[a] = cl_dec(bb[3 + a]); e[a] = G->bmax[a] - G->bmin[a]; }
    double h;
    if (voxel_size > 0.0) {
        h = voxel_size;
        for (int a = 0; a < 3; ++a) G->gmin[a] = G->bmin[a] - voxel_size * 0.5;
    } else {
        const double area = fmax(e[0] * e[1], fmax(e[1] * e[2], e[0] * e[2]));
        const double spacing = sqrt(fmax(area, 1e-300) / (double)(n > 0 ? n : 1));
        h = 0.6 * sqrt((double)(k > 0 ? k : 1)) * spacing;
        if (radius > 0.0 && h > radius) h = radius;
        const double emax = fmax(e[0], fmax(e[1], e[2]));
        if (h < emax * 1e-6) h = emax * 1e-6;      // keeps cell coordinates far inside 21 bits
        if (!(h > 0.0)) h = 1.0;                   // all points identical
        for (int a = 0; a < 3; ++a) G->gmin[a] = G->bmin[a] - 0.5 * h;
    }
    G->h = h;
    G->inv_h = 1.0 / h;
    int bad = 0;
    for (int a = 0; a < 3; ++a) {
        const double c = floor((G->bmax[a] - G->gmin[a]) / h);
        if (!(c < 2097151.0)) bad = 1;
        G->cmax[a] = bad ? 0 : (int)c;
    }
    G->bad = bad;
}

// Open3D VoxelDownSample: voxel_index = floor((p - voxel_min_bound) / voxel_size) -- a true division, kept for
// both grid kinds so that a point and a query with the same coordinates always land in the same cell.
__device__ __forceinline__ void cl_cell_of(const CloudGrid& G, const double* p, int c[3]) {
#pragma unroll
    for (int a = 0; a < 3; ++a) c[a] = (int)floor((p[a] - G.gmin[a]) / G.h);
}

// insert every point's cell, remember the slot, count points per slot; first[slot] = lowest point index
__global__ void __launch_bounds__(256) cl_insert_kernel(const double* __restrict__ pts, long long n, const CloudGrid* __restrict__ Gp,
                                                        unsigned long long* __restrict__ keys, uint32_t* __restrict__ counts,
                                                        uint32_t* __restrict__ first, uint32_t* __restrict__ slot_of,
                                                        unsigned int tab_mask) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const CloudGrid& G = *Gp;
    int c[3] = {0, 0, 0};
    if (!G.bad) cl_cell_of(G, pts + i * 3, c);   // bad grid: everything goes to one cell, the host reports the error
    const unsigned long long key = cl_key(c[0], c[1], c[2]);
    unsigned int s = cl_hash(key) & tab_mask;
    while (true) {
        const unsigned long long prev = atomicCAS(&keys[s], CL_EMPTY, key);
        if (prev == CL_EMPTY || prev == key) break;
        s = (s + 1) & tab_mask;
    }
    atomicAdd(&counts[s], 1u);
    if (first) atomicMin(&first[s], (uint32_t)i);
    slot_of[i] = s;
}

__global__ void __launch_bounds__(256) cl_scatter_kernel(const double* __restrict__ pts, long long n,
                                                         const uint32_t* __restrict__ starts, uint32_t* __restrict__ cursor,
                                                         const uint32_t* __restrict__ slot_of, double* __restrict__ s_pts,
                                                         int32_t* __restrict__ s_idx) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const unsigned int s = slot_of[i];
    const uint32_t pos = starts[s] + atomicAdd(&cursor[s], 1u);
#pragma unroll
    for (int k = 0; k < 3; ++k) s_pts[(long long)pos * 3 + k] = pts[i * 3 + k];
    s_idx[pos] = (int32_t)i;
}

// ---- exact k-nearest / hybrid search ---------------------------------------------------------------------------
// Max-heap over (d2, index): the root is the WORST of the kept candidates.
struct NbHeap {
    double d[KCAP];
    int j[KCAP];
    int n;
};
__device__ __forceinline__ bool nb_less(double da, int ja, double db, int jb) { return da < db || (da == db && ja < jb); }

__device__ __forceinline__ void heap_sift_down(NbHeap& H, int root, int end) {
    const double d0 = H.d[root];
    const int j0 = H.j[root];
    int i = root;
    while (true) {
        int c = 2 * i + 1;
        if (c >= end) break;
        if (c + 1 < end && nb_less(H.d[c], H.j[c], H.d[c + 1], H.j[c + 1])) ++c;
        if (!nb_less(d0, j0, H.d[c], H.j[c])) break;
        H.d[i] = H.d[c]; H.j[i] = H.j[c];
        i = c;
    }
    H.d[i] = d0; H.j[i] = j0;
}

__device__ __forceinline__ void heap_offer(NbHeap& H, int k, double d2, int j) {
    if (H.n < k) {   // sift up
        int i = H.n++;
        while (i > 0) {
            const int p = (i - 1) >> 1;
            if (!nb_less(H.d[p], H.j[p], d2, j)) break;
            H.d[i] = H.d[p]; H.j[i] = H.j[p];
            i = p;
        }
        H.d[i] = d2; H.j[i] = j;
    } else if (nb_less(d2, j, H.d[0], H.j[0])) {
        H.d[0] = d2; H.j[0] = j;
        heap_sift_down(H, 0, k);
    }
}

// The k nearest points to q with d^2 < r2 (r2 = +inf: plain kNN), left in H sorted ASCENDING by (d2, index).
// q must lie inside the cloud's bounding box (every caller queries cloud points).
__device__ void grid_search(const GridView& V, const double q[3], int k, double r2, NbHeap& H) {
    const CloudGrid& G = *V.g;
    H.n = 0;
    int c[3];
    cl_cell_of(G, q, c);
    double face = DBL_MAX;     // distance from q to the nearest face of its own cell
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const double lo = G.gmin[a] + (double)c[a] * G.h;
        face = fmin(face, fmin(q[a] - lo, lo + G.h - q[a]));
    }
    face = fmax(face, 0.0) * (1.0 - 1e-9);   // cells come from rounded arithmetic: keep the bound conservative
    const int max_ring = max(max(max(c[0], G.cmax[0] - c[0]), max(c[1], G.cmax[1] - c[1])), max(c[2], G.cmax[2] - c[2]));
    for (int ring = 0; ring <= max_ring; ++ring) {
        if (ring > 0) {
            // every point outside the rings visited so far is at least this far from q
            const double lo = (double)(ring - 1) * G.h + face;
            const double bound = H.n < k ? r2 : fmin(r2, H.d[0]);
            if (lo * lo > bound || lo * lo >= r2) break;
        }
        const int z0 = max(c[2] - ring, 0), z1 = min(c[2] + ring, G.cmax[2]);
        const int y0 = max(c[1] - ring, 0), y1 = min(c[1] + ring, G.cmax[1]);
        for (int cz = z0; cz <= z1; ++cz)
            for (int cy = y0; cy <= y1; ++cy) {
                const bool shell = (cz == c[2] - ring || cz == c[2] + ring || cy == c[1] - ring || cy == c[1] + ring);
                const int step = (shell || ring == 0) ? 1 : 2 * ring;
                for (int cx = c[0] - ring; cx <= c[0] + ring; cx += step) {
                    if (cx < 0 || cx > G.cmax[0]) continue;
                    const double lx = G.gmin[0] + (double)cx * G.h, ly = G.gmin[1] + (double)cy * G.h, lz = G.gmin[2] + (double)cz * G.h;
                    const double ex = q[0] < lx ? lx - q[0] : (q[0] > lx + G.h ? q[0] - (lx + G.h) : 0.0);
                    const double ey = q[1] < ly ? ly - q[1] : (q[1] > ly + G.h ? q[1] - (ly + G.h) : 0.0);
                    const double ez = q[2] < lz ? lz - q[2] : (q[2] > lz + G.h ? q[2] - (lz + G.h) : 0.0);
                    const double bound = H.n < k ? r2 : fmin(r2, H.d[0]);
                    if ((ex * ex + ey * ey + ez * ez) * (1.0 - 1e-9) > bound) continue;
                    const unsigned long long key = cl_key(cx, cy, cz);
                    unsigned int sl = cl_hash(key) & V.tab_mask;
                    unsigned long long kk;
                    while ((kk = __ldg(&V.keys[sl])) != key && kk != CL_EMPTY) sl = (sl + 1) & V.tab_mask;
                    if (kk != key) continue;
                    const uint32_t st = __ldg(&V.starts[sl]), cn = __ldg(&V.counts[sl]);
                    for (uint32_t m = st; m < st + cn; ++m) {
                        const double tx = __ldg(&V.s_pts[(long long)m * 3]) - q[0];
                        const double ty = __ldg(&V.s_pts[(long long)m * 3 + 1]) - q[1];
                        const double tz = __ldg(&V.s_pts[(long long)m * 3 + 2]) - q[2];
                        const double d2 = tx * tx + ty * ty + tz * tz;
                        if (!(d2 < r2)) continue;
                        if (H.n == k && d2 > H.d[0]) continue;
                        heap_offer(H, k, d2, __ldg(&V.s_idx[m]));
                    }
                }
            }
    }
    // heap sort: ascending (d2, index)
    for (int m = H.n - 1; m > 0; --m) {
        const double td = H.d[0]; const int tj = H.j[0];
        H.d[0] = H.d[m]; H.j[0] = H.j[m];
        H.d[m] = td; H.j[m] = tj;
        heap_sift_down(H, 0, m);
    }
}

// ---- estimate_normals -------------------------------------------------------------------------------------------
struct D3 { double x, y, z; };
__device__ __forceinline__ D3 d3_cross(const D3& a, const D3& b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
__device__ __forceinline__ double d3_dot(const D3& a, const D3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ D3 d3_scale(const D3& a, double s) { return {a.x * s, a.y * s, a.z * s}; }

// Open3D ComputeEigenvector0 / ComputeEigenvector1 / FastEigen3x3 (Geometric Tools' robust symmetric 3x3
// eigensolver), operation for operation.  A = {a00, a01, a02, a11, a12, a22}.
__device__ D3 eigvec0(const double A[6], double ev) {
    const D3 r0{A[0] - ev, A[1], A[2]}, r1{A[1], A[3] - ev, A[4]}, r2{A[2], A[4], A[5] - ev};
    const D3 r0xr1 = d3_cross(r0, r1), r0xr2 = d3_cross(r0, r2), r1xr2 = d3_cross(r1, r2);
    const double d0 = d3_dot(r0xr1, r0xr1), d1 = d3_dot(r0xr2, r0xr2), d2 = d3_dot(r1xr2, r1xr2);
    double dmax = d0;
    int imax = 0;
    if (d1 > dmax) { dmax = d1; imax = 1; }
    if (d2 > dmax) imax = 2;
    if (imax == 0) return d3_scale(r0xr1, 1.0 / sqrt(d0));
    if (imax == 1) return d3_scale(r0xr2, 1.0 / sqrt(d1));
    return d3_scale(r1xr2, 1.0 / sqrt(d2));
}

__device__ D3 eigvec1(const double A[6], const D3& e0, double ev1) {
    D3 U;
    if (fabs(e0.x) > fabs(e0.y)) {
        const double il = 1.0 / sqrt(e0.x * e0.x + e0.z * e0.z);
        U = {-e0.z * il, 0.0, e0.x * il};
    } else {
        const double il = 1.0 / sqrt(e0.y * e0.y + e0.z * e0.z);
        U = {0.0, e0.z * il, -e0.y * il};
    }
    const D3 V = d3_cross(e0, U);
    const D3 AU{A[0] * U.x + A[1] * U.y + A[2] * U.z, A[1] * U.x + A[3] * U.y + A[4] * U.z, A[2] * U.x + A[4] * U.y + A[5] * U.z};
    const D3 AV{A[0] * V.x + A[1] * V.y + A[2] * V.z, A[1] * V.x + A[3] * V.y + A[4] * V.z, A[2] * V.x + A[4] * V.y + A[5] * V.z};
    double m00 = d3_dot(U, AU) - ev1, m01 = d3_dot(U, AV), m11 = d3_dot(V, AV) - ev1;
    const double a00 = fabs(m00), a01 = fabs(m01), a11 = fabs(m11);
    if (a00 >= a11) {
        if (fmax(a00, a01) > 0.0) {
            if (a00 >= a01) { m01 /= m00; m00 = 1.0 / sqrt(1.0 + m01 * m01); m01 *= m00; }
            else { m00 /= m01; m01 = 1.0 / sqrt(1.0 + m00 * m00); m00 *= m01; }
            return {m01 * U.x - m00 * V.x, m01 * U.y - m00 * V.y, m01 * U.z - m00 * V.z};
        }
        return U;
    }
    if (fmax(a11, a01) > 0.0) {
        if (a11 >= a01) { m01 /= m11; m11 = 1.0 / sqrt(1.0 + m01 * m01); m01 *= m11; }
        else { m11 /= m01; m01 = 1.0 / sqrt(1.0 + m11 * m11); m11 *= m01; }
        return {m11 * U.x - m01 * V.x, m11 * U.y - m01 * V.y, m11 * U.z - m01 * V.z};
    }
    return U;
}

__device__ D3 fast_eigen3x3(const double C[6]) {
    double mx = C[0];
#pragma unroll
    for (int i = 1; i < 6; ++i) mx = fmax(mx, C[i]);
    if (mx == 0.0) return {0.0, 0.0, 0.0};
    double A[6];
#pragma unroll
    for (int i = 0; i < 6; ++i) A[i] = C[i] / mx;
    const double norm = A[1] * A[1] + A[2] * A[2] + A[4] * A[4];
    if (norm > 0.0) {
        const double q = (A[0] + A[3] + A[5]) / 3.0;
        const double b00 = A[0] - q, b11 = A[3] - q, b22 = A[5] - q;
        const double p = sqrt((b00 * b00 + b11 * b11 + b22 * b22 + norm * 2.0) / 6.0);
        const double c00 = b11 * b22 - A[4] * A[4];
        const double c01 = A[1] * b22 - A[4] * A[2];
        const double c02 = A[1] * A[4] - b11 * A[2];
        const double det = (b00 * c00 - A[1] * c01 + A[2] * c02) / (p * p * p);
        double half_det = det * 0.5;
        half_det = fmin(fmax(half_det, -1.0), 1.0);
        const double angle = acos(half_det) / 3.0;
        const double two_thirds_pi = 2.09439510239319549;
        const double beta2 = cos(angle) * 2.0;
        const double beta0 = cos(angle + two_thirds_pi) * 2.0;
        const double beta1 = -(beta0 + beta2);
        const double e0 = q + p * beta0, e1 = q + p * beta1, e2 = q + p * beta2;
        if (half_det >= 0.0) {
            const D3 v2 = eigvec0(A, e2);
            if (e2 < e0 && e2 < e1) return v2;
            const D3 v1 = eigvec1(A, v2, e1);
            if (e1 < e0 && e1 < e2) return v1;
            return d3_cross(v1, v2);
        }
        const D3 v0 = eigvec0(A, e0);
        if (e0 < e1 && e0 < e2) return v0;
        const D3 v1 = eigvec1(A, v0, e1);
        if (e1 < e0 && e1 < e2) return v1;
        return d3_cross(v0, v1);
    }
    if (C[0] < C[3] && C[0] < C[5]) return {1.0, 0.0, 0.0};
    if (C[3] < C[0] && C[3] < C[5]) return {0.0, 1.0, 0.0};
    return {0.0, 0.0, 1.0};
}

__global__ void __launch_bounds__(128) cl_normals_kernel(const GridView V, const double* __restrict__ pts, long long n, int k,
                                                         double r2, const double* prev, double* out) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    NbHeap H;
    const double q[3] = {pts[i * 3], pts[i * 3 + 1], pts[i * 3 + 2]};
    grid_search(V, q, k, r2, H);
    D3 nrm{0.0, 0.0, 1.0};
    if (H.n >= 3) {
        // utility::ComputeCovariance: cumulants over the neighbours in search order
        double c[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
        for (int m = 0; m < H.n; ++m) {
            const double* p = pts + (long long)H.j[m] * 3;
            const double x = p[0], y = p[1], z = p[2];
            c[0] += x; c[1] += y; c[2] += z;
            c[3] += x * x; c[4] += x * y; c[5] += x * z; c[6] += y * y; c[7] += y * z; c[8] += z * z;
        }
        const double cnt = (double)H.n;
#pragma unroll
        for (int a = 0; a < 9; ++a) c[a] /= cnt;
        const double C[6] = {c[3] - c[0] * c[0], c[4] - c[0] * c[1], c[5] - c[0] * c[2],
                             c[6] - c[1] * c[1], c[7] - c[1] * c[2], c[8] - c[2] * c[2]};
        nrm = fast_eigen3x3(C);
    }
    const bool hp = prev != nullptr;
    const D3 pv = hp ? D3{prev[i * 3], prev[i * 3 + 1], prev[i * 3 + 2]} : D3{0.0, 0.0, 1.0};
    if (sqrt(d3_dot(nrm, nrm)) == 0.0) nrm = pv;
    if (hp && d3_dot(nrm, pv) < 0.0) nrm = d3_scale(nrm, -1.0);
    out[i * 3] = nrm.x; out[i * 3 + 1] = nrm.y; out[i * 3 + 2] = nrm.z;
}

// ---- remove_statistical_outlier ---------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) cl_avgdist_kernel(const GridView V, const double* __restrict__ pts, long long n, int k,
                                                         double* __restrict__ avg) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    NbHeap H;
    const double q[3] = {pts[i * 3], pts[i * 3 + 1], pts[i * 3 + 2]};
    grid_search(V, q, k, __longlong_as_double(0x7ff0000000000000ll), H);
    double s = 0.0;
    for (int m = 0; m < H.n; ++m) s += sqrt(H.d[m]);
    avg[i] = H.n > 0 ? s / (double)H.n : -1.0;
}

// Single block: cloud mean and sample standard deviation of the positive averages -> threshold.  Fixed
// reduction tree (deterministic); Open3D sums sequentially, the two differ by rounding only.
__global__ void __launch_bounds__(1024) cl_outlier_threshold_kernel(const double* __restrict__ avg, long long n, double std_ratio,
                                                                    double* __restrict__ out /* thr, mean, std, valid */) {
    __shared__ double s_a[32], s_b[32];
    __shared__ double s_mean, s_valid;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int pass = 0; pass < 2; ++pass) {
        double a = 0.0, b = 0.0;
        const double mean = pass ? s_mean : 0.0;
        for (long long i = threadIdx.x; i < n; i += blockDim.x) {
            const double v = avg[i];
            if (v > 0.0) {
                if (pass == 0) { a += v; b += 1.0; }
                else a += (v - mean) * (v - mean);
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { a += __shfl_down_sync(0xffffffffu, a, o); b += __shfl_down_sync(0xffffffffu, b, o); }
        if (lane == 0) { s_a[warp] = a; s_b[warp] = b; }
        __syncthreads();
        if (threadIdx.x == 0) {
            double ta = 0.0, tb = 0.0;
            for (int w = 0; w < (int)(blockDim.x >> 5); ++w) { ta += s_a[w]; tb += s_b[w]; }
            if (pass == 0) { s_valid = tb; s_mean = tb > 0.0 ? ta / tb : 0.0; }
            else {
                const double sd = sqrt(ta / (s_valid - 1.0));
                out[0] = s_mean + std_ratio * sd; out[1] = s_mean; out[2] = sd; out[3] = s_valid;
            }
        }
        __syncthreads();
    }
}

__global__ void __launch_bounds__(256) cl_outlier_flag_kernel(const double* __restrict__ avg, long long n,
                                                              const double* __restrict__ thr, uint32_t* __restrict__ flag) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double v = avg[i];
    flag[i] = (thr[3] > 0.0 && v > 0.0 && v < thr[0]) ? 1u : 0u;
}

__global__ void __launch_bounds__(256) cl_compact_index_kernel(const uint32_t* __restrict__ flag, const uint32_t* __restrict__ pos,
                                                               long long n, int64_t* __restrict__ out) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && flag[i]) out[pos[i]] = i;
}

// ---- colour gradient for colored ICP ----------------------------------------------------------------------------
// 3x3 symmetric solve, LDL^T with diagonal pivoting (Eigen ldlt(): largest remaining diagonal first).
__device__ void ldlt3_solve(const double* Ain, const double* b, double* x) {
    double A[9];
    int perm[3] = {0, 1, 2};
    for (int i = 0; i < 9; ++i) A[i] = Ain[i];
    for (int k = 0; k < 3; ++k) {
        int p = k;
        double best = fabs(A[k * 3 + k]);
        for (int i = k + 1; i < 3; ++i) if (fabs(A[i * 3 + i]) > best) { best = fabs(A[i * 3 + i]); p = i; }
        if (p != k) {
            for (int j = 0; j < 3; ++j) { const double t = A[k * 3 + j]; A[k * 3 + j] = A[p * 3 + j]; A[p * 3 + j] = t; }
            for (int j = 0; j < 3; ++j) { const double t = A[j * 3 + k]; A[j * 3 + k] = A[j * 3 + p]; A[j * 3 + p] = t; }
            const int t = perm[k]; perm[k] = perm[p]; perm[p] = t;
        }
        const double d = A[k * 3 + k];
        if (d == 0.0) continue;
        for (int i = k + 1; i < 3; ++i) {
            const double l = A[i * 3 + k] / d;
            for (int j = k + 1; j < 3; ++j) A[i * 3 + j] -= l * A[k * 3 + j];
            A[i * 3 + k] = l;
        }
    }
    double y[3], z[3];
    for (int i = 0; i < 3; ++i) { double s = b[perm[i]]; for (int j = 0; j < i; ++j) s -= A[i * 3 + j] * y[j]; y[i] = s; }
    for (int i = 0; i < 3; ++i) y[i] = (A[i * 3 + i] != 0.0) ? y[i] / A[i * 3 + i] : 0.0;
    for (int i = 2; i >= 0; --i) { double s = y[i]; for (int j = i + 1; j < 3; ++j) s -= A[j * 3 + i] * z[j]; z[i] = s; }
    for (int i = 0; i < 3; ++i) x[perm[i]] = z[i];
}

__global__ void __launch_bounds__(128) cl_color_gradient_kernel(const GridView V, const double* __restrict__ pts,
                                                                const double* __restrict__ nrm, const double* __restrict__ col,
                                                                long long n, int k, double r2, double* __restrict__ grad) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    NbHeap H;
    const double vt[3] = {pts[i * 3], pts[i * 3 + 1], pts[i * 3 + 2]};
    grid_search(V, vt, k, r2, H);
    double g[3] = {0.0, 0.0, 0.0};
    if (H.n >= 4) {
        const double nt[3] = {nrm[i * 3], nrm[i * 3 + 1], nrm[i * 3 + 2]};
        const double it = (col[i * 3] + col[i * 3 + 1] + col[i * 3 + 2]) / 3.0;
        double AtA[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, Atb[3] = {0, 0, 0};
        for (int m = 1; m < H.n; ++m) {
            const long long a = H.j[m];
            const double va[3] = {pts[a * 3], pts[a * 3 + 1], pts[a * 3 + 2]};
            const double d = (va[0] - vt[0]) * nt[0] + (va[1] - vt[1]) * nt[1] + (va[2] - vt[2]) * nt[2];
            const double row[3] = {(va[0] - d * nt[0]) - vt[0], (va[1] - d * nt[1]) - vt[1], (va[2] - d * nt[2]) - vt[2]};
            const double b = (col[a * 3] + col[a * 3 + 1] + col[a * 3 + 2]) / 3.0 - it;
#pragma unroll
            for (int r = 0; r < 3; ++r) {
#pragma unroll
                for (int c = 0; c < 3; ++c) AtA[r * 3 + c] += row[r] * row[c];
                Atb[r] += row[r] * b;
            }
        }
        const double w = (double)(H.n - 1);
        const double row[3] = {w * nt[0], w * nt[1], w * nt[2]};
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
            for (int c = 0; c < 3; ++c) AtA[r * 3 + c] += row[r] * row[c];
        ldlt3_solve(AtA, Atb, g);
    }
    grad[i * 3] = g[0]; grad[i * 3 + 1] = g[1]; grad[i * 3 + 2] = g[2];
}

__global__ void __launch_bounds__(128) cl_knn_kernel(const GridView V, const double* __restrict__ qs, long long nq, int k, double r2,
                                                     int32_t* __restrict__ idx, double* __restrict__ d2) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nq) return;
    NbHeap H;
    const double q[3] = {qs[i * 3], qs[i * 3 + 1], qs[i * 3 + 2]};
    grid_search(V, q, k, r2, H);
    for (int m = 0; m < k; ++m) {
        idx[i * k + m] = m < H.n ? H.j[m] : -1;
        d2[i * k + m] = m < H.n ? H.d[m] : 0.0;
    }
}

// ---- voxel_down_sample ------------------------------------------------------------------------------------------
// flag[i] = 1 when point i is the first (lowest index) point of its cell: the exclusive scan of the flags is the
// cell's output position, so cells come out ordered by their first point.
__global__ void __launch_bounds__(256) cl_first_flag_kernel(const uint32_t* __restrict__ slot_of, const uint32_t* __restrict__ first,
                                                            long long n, uint32_t* __restrict__ flag) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) flag[i] = (first[slot_of[i]] == (uint32_t)i) ? 1u : 0u;
}
// The scatter leaves every cell's index list in arbitrary order; Open3D adds the points of a cell in ascending
// index order (one sequential loop over the cloud), and float64 addition is not associative -- so each list is
// sorted first (one warp per hash slot, ascending-only bitonic network over the slot's segment, virtual +inf
// padding up to the next power of two) and then summed sequentially by one thread: bit-identical means.
__global__ void __launch_bounds__(128) cl_sort_cells_kernel(const uint32_t* __restrict__ starts, const uint32_t* __restrict__ counts,
                                                            unsigned int n_slots, int32_t* s_idx) {
    const unsigned int slot = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (slot >= n_slots) return;
    const uint32_t m = counts[slot];
    if (m < 2) return;
    int32_t* a = s_idx + starts[slot];
    const int lane = threadIdx.x & 31;
    uint32_t p2 = 2;
    while (p2 < m) p2 <<= 1;
    for (uint32_t k = 2; k <= p2; k <<= 1) {
        for (uint32_t j = k >> 1; j > 0; j >>= 1) {
            const bool flip = (j == (k >> 1));           // first stage of a merge mirrors, the rest are half-cleaners
            for (uint32_t t = lane; t < (p2 >> 1); t += 32) {
                // t-th compare-exchange of this stage: lo has bit j clear
                const uint32_t lo = ((t & ~(j - 1)) << 1) | (t & (j - 1));
                const uint32_t hi = flip ? (lo ^ (k - 1)) : (lo | j);
                if (hi < m) {                            // (lo < hi always; an out-of-range hi is +inf: no exchange)
                    const int32_t x = a[lo], y = a[hi];
                    if (x > y) { a[lo] = y; a[hi] = x; }
                }
            }
            __syncwarp();
        }
    }
}

// AccumulatedPoint::AddPoint / GetAverage*: sequential sums in ascending point order (NaN normals skipped),
// divided by the cell's point count.  One thread per cell (the thread of the cell's first point).
__global__ void __launch_bounds__(128) cl_cell_mean_kernel(const double* __restrict__ pts, const double* __restrict__ nrm,
                                                           const double* __restrict__ col, long long n,
                                                           const uint32_t* __restrict__ slot_of, const uint32_t* __restrict__ flag,
                                                           const uint32_t* __restrict__ pos, const uint32_t* __restrict__ starts,
                                                           const uint32_t* __restrict__ counts, const int32_t* __restrict__ s_idx,
                                                           double* __restrict__ out_p, double* __restrict__ out_n,
                                                           double* __restrict__ out_c) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || !flag[i]) return;
    const uint32_t slot = slot_of[i], st = starts[slot], m = counts[slot];
    double sp[3] = {0.0, 0.0, 0.0}, sn[3] = {0.0, 0.0, 0.0}, sc[3] = {0.0, 0.0, 0.0};
    for (uint32_t e = 0; e < m; ++e) {
        const long long j = s_idx[st + e];
#pragma unroll
        for (int k = 0; k < 3; ++k) sp[k] += pts[j * 3 + k];
        if (nrm) {
            const double x = nrm[j * 3], y = nrm[j * 3 + 1], z = nrm[j * 3 + 2];
            if (!(isnan(x) || isnan(y) || isnan(z))) { sn[0] += x; sn[1] += y; sn[2] += z; }
        }
        if (col) {
#pragma unroll
            for (int k = 0; k < 3; ++k) sc[k] += col[j * 3 + k];
        }
    }
    const long long c = pos[i];
    const double d = (double)m;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        out_p[c * 3 + k] = sp[k] / d;
        if (out_n) out_n[c * 3 + k] = sn[k] / d;
        if (out_c) out_c[c * 3 + k] = sc[k] / d;
    }
}

// ---- host side ------------------------------------------------------------------------------------------------------
static unsigned int cl_next_pow2(unsigned long long v) {
    unsigned int p = 16;
    while ((unsigned long long)p < v && p < (1u << 31)) p <<= 1;
    return p;
}
static size_t al256(size_t v) { return (v + 255) & ~(size_t)255; }

struct GridBuf {
    CloudGrid* d_grid;
    long long* d_bb;
    unsigned long long* keys;
    uint32_t *counts, *starts, *cursor, *first, *slot_of;
    double* s_pts;
    int32_t* s_idx;
    unsigned int tab_mask;
    GridView view() const { return GridView{d_grid, keys, starts, counts, s_pts, s_idx, tab_mask}; }
};

// Builds the grid of `n` points on `st` (no host sync).  sorted: also counting-sort the points by cell.
static int build_grid(b2r_ctx* ctx, const double* d_pts, int64_t n, double voxel_size, double radius, int k, bool sorted,
                      GridBuf& B, cudaStream_t st) {
    const unsigned int tsz = cl_next_pow2((unsigned long long)n * 2 + 16);
    B.tab_mask = tsz - 1;
    const size_t sz_hdr = al256(sizeof(CloudGrid)) + al256(6 * sizeof(long long));
    const size_t sz_keys = al256(sizeof(unsigned long long) * tsz), sz_u32 = al256(sizeof(uint32_t) * tsz);
    if (ctx->cl_a.ensure(sz_hdr + sz_keys + 4 * sz_u32 + al256(sizeof(uint32_t) * (size_t)n))) return -102;
    char* p = (char*)ctx->cl_a.ptr;
    B.d_grid = (CloudGrid*)p; p += al256(sizeof(CloudGrid));
    B.d_bb = (long long*)p; p += al256(6 * sizeof(long long));
    B.keys = (unsigned long long*)p; p += sz_keys;
    B.counts = (uint32_t*)p; p += sz_u32;
    B.starts = (uint32_t*)p; p += sz_u32;
    B.cursor = (uint32_t*)p; p += sz_u32;
    B.first = (uint32_t*)p; p += sz_u32;
    B.slot_of = (uint32_t*)p;
    B.s_pts = nullptr; B.s_idx = nullptr;
    if (sorted) {
        if (ctx->cl_b.ensure(al256(sizeof(double) * 3 * (size_t)n) + al256(sizeof(int32_t) * (size_t)n))) return -102;
        B.s_pts = (double*)ctx->cl_b.ptr;
        B.s_idx = (int32_t*)((char*)ctx->cl_b.ptr + al256(sizeof(double) * 3 * (size_t)n));
    }
    B2R_CUDA(cudaMemsetAsync(B.keys, 0xff, sizeof(unsigned long long) * tsz, st));
    B2R_CUDA(cudaMemsetAsync(B.counts, 0, sizeof(uint32_t) * tsz, st));
    B2R_CUDA(cudaMemsetAsync(B.cursor, 0, sizeof(uint32_t) * tsz, st));
    B2R_CUDA(cudaMemsetAsync(B.first, 0xff, sizeof(uint32_t) * tsz, st));
    cl_bbox_init_kernel<<<1, 32, 0, st>>>(B.d_bb);
    B2R_LAUNCH_CHECK(ctx);
    cl_bbox_kernel<<<(int)std::min<int64_t>(cdiv(n, 256), 1024), 256, 0, st>>>(d_pts, n, B.d_bb);
    B2R_LAUNCH_CHECK(ctx);
    cl_grid_setup_kernel<<<1, 1, 0, st>>>(B.d_grid, B.d_bb, n, voxel_size, radius, k);
    B2R_LAUNCH_CHECK(ctx);
    cl_insert_kernel<<<cdiv(n, 256), 256, 0, st>>>(d_pts, n, B.d_grid, B.keys, B.counts, B.first, B.slot_of,
                                                   B.tab_mask);
    B2R_LAUNCH_CHECK(ctx);
    if (sorted) {
        if (int rc = scan_exclusive_u32(ctx, B.counts, B.starts, tsz, nullptr, st)) return rc;
        cl_scatter_kernel<<<cdiv(n, 256), 256, 0, st>>>(d_pts, n, B.starts, B.cursor, B.slot_of, B.s_pts, B.s_idx);
        B2R_LAUNCH_CHECK(ctx);
    }
    return 0;
}

static int check_search(const char* who, int mode, double radius, int k) {
    B2R_CHECK(mode == B2R_SEARCH_KNN || mode == B2R_SEARCH_HYBRID, -5,
              "%s: search mode %d is not supported (KDTreeSearchParamKNN / KDTreeSearchParamHybrid only)", who, mode);
    B2R_CHECK(k >= 1 && k <= KCAP, -5, "%s: max_nn / knn must be in [1, %d], got %d", who, KCAP, k);
    B2R_CHECK(mode == B2R_SEARCH_KNN || radius > 0.0, -5, "%s: radius must be > 0", who);
    return 0;
}
static double search_r2(int mode, double radius) {
    return mode == B2R_SEARCH_KNN ? (double)INFINITY : radius * radius;
}

}  // namespace b2r

using namespace b2r;

extern "C" {

int b2r_voxel_down_sample(b2r_ctx* ctx, const double* d_points, const double* d_normals, const double* d_colors,
                          int64_t n, double voxel_size, double* d_out_points, double* d_out_normals,
                          double* d_out_colors, int64_t* h_n_out, void* stream) {
    B2R_CHECK(ctx && h_n_out, -1, "b2r_voxel_down_sample: NULL argument");
    // Open3D: LogError("[VoxelDownSample] voxel_size <= 0.")
    B2R_CHECK(voxel_size > 0.0, -10, "[VoxelDownSample] voxel_size <= 0.");
    B2R_CHECK(n >= 0 && n < (1ll << 31), -8, "b2r_voxel_down_sample: bad point count");
    *h_n_out = 0;
    if (n == 0) return 0;
    B2R_CHECK(d_points && d_out_points, -1, "b2r_voxel_down_sample: NULL point arrays");
    B2R_CUDA(cudaSetDevice(ctx->device));
    cudaStream_t st = S(stream);
    GridBuf B;
    if (int rc = build_grid(ctx, d_points, n, voxel_size, 0.0, 0, true, B, st)) return rc;
    const size_t sz_n = al256(sizeof(uint32_t) * (size_t)n);
    if (ctx->cl_c.ensure(2 * sz_n)) return -102;
    uint32_t* flag = (uint32_t*)ctx->cl_c.ptr;
    uint32_t* pos = (uint32_t*)((char*)ctx->cl_c.ptr + sz_n);
    unsigned long long* d_total = (unsigned long long*)ctx->small.ptr + SMALL_TOTAL64;
    const unsigned int n_slots = B.tab_mask + 1;
    cl_sort_cells_kernel<<<cdiv(n_slots, 4), 128, 0, st>>>(B.starts, B.counts, n_slots, B.s_idx);
    B2R_LAUNCH_CHECK(ctx);
    cl_first_flag_kernel<<<cdiv(n, 256), 256, 0, st>>>(B.slot_of, B.first, n, flag);
    B2R_LAUNCH_CHECK(ctx);
    if (int rc = scan_exclusive_u32(ctx, flag, pos, n, d_total, st)) return rc;
    cl_cell_mean_kernel<<<cdiv(n, 128), 128, 0, st>>>(d_points, d_normals, d_colors, n, B.slot_of, flag, pos, B.starts, B.counts,
                                                      B.s_idx, d_out_points, d_normals ? d_out_normals : nullptr,
                                                      d_colors ? d_out_colors : nullptr);
    B2R_LAUNCH_CHECK(ctx);
    CloudGrid hg;
    B2R_CUDA(cudaMemcpyAsync(ctx->pinned, d_total, 8, cudaMemcpyDeviceToHost, st));
    B2R_CUDA(cudaMemcpyAsync(&hg, B.d_grid, sizeof(hg), cudaMemcpyDeviceToHost, st));
    B2R_CUDA(cudaStreamSynchronize(st));
    // Open3D: LogError("[VoxelDownSample] voxel_size is too small.") when the index would overflow an int; here the
    // limit is the 21-bit cell coordinate of the hash key.
    B2R_CHECK(!hg.bad, -8, "[VoxelDownSample] voxel_size is too small.");
    *h_n_out = (int64_t)*(unsigned long long*)ctx->pinned;
    return 0;
}

int b2r_estimate_normals(b2r_ctx* ctx, const double* d_points, int64_t n, int search_mode, double radius, int max_nn,
                         const double* d_prev_normals, double* d_out_normals, void* stream) {
    B2R_CHECK(ctx, -1, "b2r_estimate_normals: NULL ctx");
    if (int rc = check_search("b2r_estimate_normals", search_mode, radius, max_nn)) return rc;
    B2R_CHECK(n >= 0 && n < (1ll << 31), -8, "b2r_estimate_normals: bad point count");
    if (n == 0) return 0;
    B2R_CHECK(d_points && d_out_normals, -1, "b2r_estimate_normals: NULL arrays");
    B2R_CUDA(cudaSetDevice(ctx->device));
    cudaStream_t st = S(stream);
    GridBuf B;
    if (int rc = build_grid(ctx, d_points, n, 0.0, search_mode == B2R_SEARCH_HYBRID ? radius : 0.0, max_nn, true, B, st)) return rc;
    cl_normals_kernel<<<cdiv(n, 128), 128, 0, st>>>(B.view(), d_points, n, max_nn, search_r2(search_mode, radius), d_prev_normals,
                                                    d_out_normals);
    B2R_LAUNCH_CHECK(ctx);
    return 0;
}

int b2r_statistical_outlier(b2r_ctx* ctx, const double* d_points, int64_t n, int nb_neighbors, double std_ratio,
                            int64_t* d_indices_out, int64_t* h_n_out, void* stream) {
    B2R_CHECK(ctx && h_n_out, -1, "b2r_statistical_outlier: NULL argument");
    // Open3D: LogError("[RemoveStatisticalOutliers] Illegal input parameters, the number of neighbors and standard
    // deviation ratio must be positive.")
    B2R_CHECK(nb_neighbors >= 1 && std_ratio > 0.0, -10,
              "[RemoveStatisticalOutliers] Illegal input parameters, the number of neighbors and standard deviation ratio must be positive.");
    B2R_CHECK(nb_neighbors <= KCAP, -5, "b2r_statistical_outlier: nb_neighbors must be <= %d", KCAP);
    B2R_CHECK(n >= 0 && n < (1ll << 31), -8, "b2r_statistical_outlier: bad point count");
    *h_n_out = 0;
    if (n == 0) return 0;
    B2R_CHECK(d_points && d_indices_out, -1, "b2r_statistical_outlier: NULL arrays");
    B2R_CUDA(cudaSetDevice(ctx->device));
    cudaStream_t st = S(stream);
    GridBuf B;
    if (int rc = build_grid(ctx, d_points, n, 0.0, 0.0, nb_neighbors, true, B, st)) return rc;
    const size_t sz_avg = al256(sizeof(double) * (size_t)n), sz_n = al256(sizeof(uint32_t) * (size_t)n);
    if (ctx->cl_c.ensure(sz_avg + 2 * sz_n + 256)) return -102;
    double* avg = (double*)ctx->cl_c.ptr;
    uint32_t* flag = (uint32_t*)((char*)ctx->cl_c.ptr + sz_avg);
    uint32_t* pos = (uint32_t*)((char*)ctx->cl_c.ptr + sz_avg + sz_n);
    double* thr = (double*)((char*)ctx->cl_c.ptr + sz_avg + 2 * sz_n);
    unsigned long long* d_total = (unsigned long long*)ctx->small.ptr + SMALL_TOTAL64;
    cl_avgdist_kernel<<<cdiv(n, 128), 128, 0, st>>>(B.view(), d_points, n, nb_neighbors, avg);
    B2R_LAUNCH_CHECK(ctx);
    cl_outlier_threshold_kernel<<<1, 1024, 0, st>>>(avg, n, std_ratio, thr);
    B2R_LAUNCH_CHECK(ctx);
    cl_outlier_flag_kernel<<<cdiv(n, 256), 256, 0, st>>>(avg, n, thr, flag);
    B2R_LAUNCH_CHECK(ctx);
    if (int rc = scan_exclusive_u32(ctx, flag, pos, n, d_total, st)) return rc;
    cl_compact_index_kernel<<<cdiv(n, 256), 256, 0, st>>>(flag, pos, n, d_indices_out);
    B2R_LAUNCH_CHECK(ctx);
    B2R_CUDA(cudaMemcpyAsync(ctx->pinned, d_total, 8, cudaMemcpyDeviceToHost, st));
    B2R_CUDA(cudaStreamSynchronize(st));
    *h_n_out = (int64_t)*(unsigned long long*)ctx->pinned;
    return 0;
}

int b2r_color_gradient(b2r_ctx* ctx, const double* d_points, const double* d_normals, const double* d_colors, int64_t n,
                       double radius, int max_nn, double* d_grad_out, void* stream) {
    B2R_CHECK(ctx, -1, "b2r_color_gradient: NULL ctx");
    if (int rc = check_search("b2r_color_gradient", B2R_SEARCH_HYBRID, radius, max_nn)) return rc;
    B2R_CHECK(n >= 0 && n < (1ll << 31), -8, "b2r_color_gradient: bad point count");
    if (n == 0) return 0;
    B2R_CHECK(d_points && d_normals && d_colors && d_grad_out, -1, "b2r_color_gradient: NULL arrays");
    B2R_CUDA(cudaSetDevice(ctx->device));
    cudaStream_t st = S(stream);
    GridBuf B;
    if (int rc = build_grid(ctx, d_points, n, 0.0, radius, max_nn, true, B, st)) return rc;
    cl_color_gradient_kernel<<<cdiv(n, 128), 128, 0, st>>>(B.view(), d_points, d_normals, d_colors, n, max_nn, radius * radius,
                                                           d_grad_out);
    B2R_LAUNCH_CHECK(ctx);
    return 0;
}

int b2r_knn_search(b2r_ctx* ctx, const double* d_points, int64_t n, const double* d_queries, int64_t n_query, int search_mode,
                   double radius, int k, int32_t* d_idx_out, double* d_d2_out, void* stream) {
    B2R_CHECK(ctx, -1, "b2r_knn_search: NULL ctx");
    if (int rc = check_search("b2r_knn_search", search_mode, radius, k)) return rc;
    B2R_CHECK(n > 0 && n < (1ll << 31) && n_query >= 0, -8, "b2r_knn_search: bad point count");
    if (n_query == 0) return 0;
    B2R_CHECK(d_points && d_queries && d_idx_out && d_d2_out, -1, "b2r_knn_search: NULL arrays");
    B2R_CUDA(cudaSetDevice(ctx->device));
    cudaStream_t st = S(stream);
    GridBuf B;
    if (int rc = build_grid(ctx, d_points, n, 0.0, search_mode == B2R_SEARCH_HYBRID ? radius : 0.0, k, true, B, st)) return rc;
    cl_knn_kernel<<<cdiv(n_query, 128), 128, 0, st>>>(B.view(), d_queries, n_query, k, search_r2(search_mode, radius), d_idx_out,
                                                      d_d2_out);
    B2R_LAUNCH_CHECK(ctx);
    return 0;
}

}  // extern "C"
