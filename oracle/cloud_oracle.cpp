// ORACLE -- TEST INFRASTRUCTURE ONLY.  Not part of the product path.
//
// CPU restatement of the Open3D 0.18.0 point-cloud preprocessing that stitcher.py calls around ICP
// (SURVEY.md 8(f) rows f1/f2/f3, Appendix A.6-A.9):
//   voxel_down_sample            stitcher.py:58,151      PointCloud::VoxelDownSample
//   estimate_normals             stitcher.py:61-63       PointCloud::EstimateNormals (FastEigen3x3)
//   remove_statistical_outlier   stitcher.py:158-159     PointCloud::RemoveStatisticalOutliers
//   colour gradient              stitcher.py:94-102      InitializePointCloudForColoredICP
// Neighbour searches are exact (oracle_kdtree.h stands in for nanoflann; ties -> lower index).
// PARITY UNPINNED against the real Open3D binary (open3d==0.18.0 is an un-vendored pip dependency,
// environment.yml:72-73; restated from the published algorithm, see recon_oracle.cpp's header).
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <unordered_map>
#include <vector>

#include "oracle_kdtree.h"

namespace {

struct V3 { double x, y, z; };
inline V3 cross(const V3& a, const V3& b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline double dot(const V3& a, const V3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline V3 scale(const V3& a, double s) { return {a.x * s, a.y * s, a.z * s}; }

// Open3D ComputeEigenvector0 (Geometric Tools, "A Robust Eigensolver for 3x3 Symmetric Matrices").
V3 eigenvector0(const double A[6], double ev) {  // A = {a00,a01,a02,a11,a12,a22}
    const V3 r0{A[0] - ev, A[1], A[2]}, r1{A[1], A[3] - ev, A[4]}, r2{A[2], A[4], A[5] - ev};
    const V3 r0xr1 = cross(r0, r1), r0xr2 = cross(r0, r2), r1xr2 = cross(r1, r2);
    const double d0 = dot(r0xr1, r0xr1), d1 = dot(r0xr2, r0xr2), d2 = dot(r1xr2, r1xr2);
    double dmax = d0; int imax = 0;
    if (d1 > dmax) { dmax = d1; imax = 1; }
    if (d2 > dmax) imax = 2;
    if (imax == 0) return scale(r0xr1, 1.0 / std::sqrt(d0));
    if (imax == 1) return scale(r0xr2, 1.0 / std::sqrt(d1));
    return scale(r1xr2, 1.0 / std::sqrt(d2));
}

// Open3D ComputeEigenvector1.
V3 eigenvector1(const double A[6], const V3& e0, double ev1) {
    V3 U;
    if (std::fabs(e0.x) > std::fabs(e0.y)) {
        const double il = 1.0 / std::sqrt(e0.x * e0.x + e0.z * e0.z);
        U = {-e0.z * il, 0.0, e0.x * il};
    } else {
        const double il = 1.0 / std::sqrt(e0.y * e0.y + e0.z * e0.z);
        U = {0.0, e0.z * il, -e0.y * il};
    }
    const V3 V = cross(e0, U);
    const V3 AU{A[0] * U.x + A[1] * U.y + A[2] * U.z, A[1] * U.x + A[3] * U.y + A[4] * U.z, A[2] * U.x + A[4] * U.y + A[5] * U.z};
    const V3 AV{A[0] * V.x + A[1] * V.y + A[2] * V.z, A[1] * V.x + A[3] * V.y + A[4] * V.z, A[2] * V.x + A[4] * V.y + A[5] * V.z};
    double m00 = dot(U, AU) - ev1, m01 = dot(U, AV), m11 = dot(V, AV) - ev1;
    const double a00 = std::fabs(m00), a01 = std::fabs(m01), a11 = std::fabs(m11);
    if (a00 >= a11) {
        if (std::max(a00, a01) > 0) {
            if (a00 >= a01) { m01 /= m00; m00 = 1 / std::sqrt(1 + m01 * m01); m01 *= m00; }
            else { m00 /= m01; m01 = 1 / std::sqrt(1 + m00 * m00); m00 *= m01; }
            return {m01 * U.x - m00 * V.x, m01 * U.y - m00 * V.y, m01 * U.z - m00 * V.z};
        }
        return U;
    }
    if (std::max(a11, a01) > 0) {
        if (a11 >= a01) { m01 /= m11; m11 = 1 / std::sqrt(1 + m01 * m01); m01 *= m11; }
        else { m11 /= m01; m01 = 1 / std::sqrt(1 + m11 * m11); m11 *= m01; }
        return {m11 * U.x - m01 * V.x, m11 * U.y - m01 * V.y, m11 * U.z - m01 * V.z};
    }
    return U;
}

// Open3D FastEigen3x3: eigenvector of the smallest eigenvalue of a symmetric 3x3.
V3 fast_eigen3x3(const double C[6]) {
    double A[6];
    double mx = C[0];
    for (int i = 1; i < 6; ++i) mx = std::max(mx, C[i]);   // Matrix3d::maxCoeff (signed, all 9 entries)
    if (mx == 0) return {0, 0, 0};
    for (int i = 0; i < 6; ++i) A[i] = C[i] / mx;
    const double norm = A[1] * A[1] + A[2] * A[2] + A[4] * A[4];
    if (norm > 0) {
        const double q = (A[0] + A[3] + A[5]) / 3;
        const double b00 = A[0] - q, b11 = A[3] - q, b22 = A[5] - q;
        const double p = std::sqrt((b00 * b00 + b11 * b11 + b22 * b22 + norm * 2) / 6);
        const double c00 = b11 * b22 - A[4] * A[4];
        const double c01 = A[1] * b22 - A[4] * A[2];
        const double c02 = A[1] * A[4] - b11 * A[2];
        const double det = (b00 * c00 - A[1] * c01 + A[2] * c02) / (p * p * p);
        double half_det = det * 0.5;
        half_det = std::min(std::max(half_det, -1.0), 1.0);
        const double angle = std::acos(half_det) / 3.0;
        const double two_thirds_pi = 2.09439510239319549;
        const double beta2 = std::cos(angle) * 2;
        const double beta0 = std::cos(angle + two_thirds_pi) * 2;
        const double beta1 = -(beta0 + beta2);
        const double e0 = q + p * beta0, e1 = q + p * beta1, e2 = q + p * beta2;
        if (half_det >= 0) {
            const V3 v2 = eigenvector0(A, e2);
            if (e2 < e0 && e2 < e1) return v2;
            const V3 v1 = eigenvector1(A, v2, e1);
            if (e1 < e0 && e1 < e2) return v1;
            return cross(v1, v2);
        }
        const V3 v0 = eigenvector0(A, e0);
        if (e0 < e1 && e0 < e2) return v0;
        const V3 v1 = eigenvector1(A, v0, e1);
        if (e1 < e0 && e1 < e2) return v1;
        return cross(v0, v1);
    }
    if (C[0] < C[3] && C[0] < C[5]) return {1, 0, 0};
    if (C[3] < C[0] && C[3] < C[5]) return {0, 1, 0};
    return {0, 0, 1};
}

// utility::ComputeCovariance: cumulants over the neighbours in search order.
void covariance(const double* pts, const std::vector<std::pair<double, int>>& nb, double C[6]) {
    double c[9] = {0};
    for (const auto& e : nb) {
        const double* p = pts + (int64_t)e.second * 3;
        c[0] += p[0]; c[1] += p[1]; c[2] += p[2];
        c[3] += p[0] * p[0]; c[4] += p[0] * p[1]; c[5] += p[0] * p[2];
        c[6] += p[1] * p[1]; c[7] += p[1] * p[2]; c[8] += p[2] * p[2];
    }
    const double n = (double)nb.size();
    for (int i = 0; i < 9; ++i) c[i] /= n;
    C[0] = c[3] - c[0] * c[0]; C[3] = c[6] - c[1] * c[1]; C[5] = c[8] - c[2] * c[2];
    C[1] = c[4] - c[0] * c[1]; C[2] = c[5] - c[0] * c[2]; C[4] = c[7] - c[1] * c[2];
}

// 3x3 symmetric solve, LDL^T with diagonal pivoting (Eigen ldlt(), as in icp_oracle.cpp).
void ldlt3(const double* Ain, const double* b, double* x) {
    const int n = 3;
    double A[9]; int perm[3] = {0, 1, 2};
    std::memcpy(A, Ain, sizeof(A));
    for (int k = 0; k < n; ++k) {
        int p = k; double best = std::fabs(A[k * n + k]);
        for (int i = k + 1; i < n; ++i) if (std::fabs(A[i * n + i]) > best) { best = std::fabs(A[i * n + i]); p = i; }
        if (p != k) {
            for (int j = 0; j < n; ++j) std::swap(A[k * n + j], A[p * n + j]);
            for (int j = 0; j < n; ++j) std::swap(A[j * n + k], A[j * n + p]);
            std::swap(perm[k], perm[p]);
        }
        const double d = A[k * n + k];
        if (d == 0.0) continue;
        for (int i = k + 1; i < n; ++i) {
            const double l = A[i * n + k] / d;
            for (int j = k + 1; j < n; ++j) A[i * n + j] -= l * A[k * n + j];
            A[i * n + k] = l;
        }
    }
    double y[3], z[3];
    for (int i = 0; i < n; ++i) { double s = b[perm[i]]; for (int j = 0; j < i; ++j) s -= A[i * n + j] * y[j]; y[i] = s; }
    for (int i = 0; i < n; ++i) y[i] = (A[i * n + i] != 0.0) ? y[i] / A[i * n + i] : 0.0;
    for (int i = n - 1; i >= 0; --i) { double s = y[i]; for (int j = i + 1; j < n; ++j) s -= A[j * n + i] * z[j]; z[i] = s; }
    for (int i = 0; i < n; ++i) x[perm[i]] = z[i];
}

struct Key { int x, y, z; bool operator==(const Key& o) const { return x == o.x && y == o.y && z == o.z; } };
struct KeyHash { size_t operator()(const Key& k) const {
    size_t h = 1469598103934665603ull;
    for (int v : {k.x, k.y, k.z}) { h ^= (size_t)(uint32_t)v; h *= 1099511628211ull; }
    return h; } };
struct Acc { double p[3] = {0, 0, 0}, n[3] = {0, 0, 0}, c[3] = {0, 0, 0}; int64_t num = 0, first = 0; };

}  // namespace

extern "C" {

// PointCloud::VoxelDownSample (A.7).  normals / colors may be NULL.  Output order: Open3D walks a
// std::unordered_map (unreproducible) -> cells are emitted by FIRST POINT INDEX here; tests compare sets.
// Returns the number of output points, or -1 for voxel_size <= 0 (Open3D: LogError).
int64_t orc_voxel_down_sample(const double* pts, const double* nrm, const double* col, int64_t n, double voxel_size,
                              double* out_pts, double* out_nrm, double* out_col) {
    if (voxel_size <= 0.0) return -1;
    if (n == 0) return 0;
    double mn[3] = {pts[0], pts[1], pts[2]};
    for (int64_t i = 1; i < n; ++i) for (int k = 0; k < 3; ++k) mn[k] = std::min(mn[k], pts[i * 3 + k]);
    for (int k = 0; k < 3; ++k) mn[k] -= voxel_size * 0.5;
    std::unordered_map<Key, Acc, KeyHash> cells;
    for (int64_t i = 0; i < n; ++i) {
        Key key{(int)std::floor((pts[i * 3] - mn[0]) / voxel_size), (int)std::floor((pts[i * 3 + 1] - mn[1]) / voxel_size),
                (int)std::floor((pts[i * 3 + 2] - mn[2]) / voxel_size)};
        auto it = cells.find(key);
        if (it == cells.end()) { it = cells.emplace(key, Acc()).first; it->second.first = i; }
        Acc& a = it->second;
        for (int k = 0; k < 3; ++k) a.p[k] += pts[i * 3 + k];
        if (nrm && !std::isnan(nrm[i * 3]) && !std::isnan(nrm[i * 3 + 1]) && !std::isnan(nrm[i * 3 + 2]))
            for (int k = 0; k < 3; ++k) a.n[k] += nrm[i * 3 + k];
        if (col) for (int k = 0; k < 3; ++k) a.c[k] += col[i * 3 + k];
        a.num++;
    }
    std::vector<const Acc*> order;
    order.reserve(cells.size());
    for (auto& kv : cells) order.push_back(&kv.second);
    std::sort(order.begin(), order.end(), [](const Acc* a, const Acc* b) { return a->first < b->first; });
    int64_t m = 0;
    for (const Acc* a : order) {
        const double d = (double)a->num;
        for (int k = 0; k < 3; ++k) {
            out_pts[m * 3 + k] = a->p[k] / d;
            if (nrm && out_nrm) out_nrm[m * 3 + k] = a->n[k] / d;   // 0.18: "call NormalizeNormals() afterwards if necessary"
            if (col && out_col) out_col[m * 3 + k] = a->c[k] / d;
        }
        ++m;
    }
    return m;
}

// mode: 0 = KNN(k), 1 = Radius(r), 2 = Hybrid(r, k).  Neighbours sorted by (d2, index), query included.
static int search(const orc::KDTree& tree, const double* q, int mode, double radius, int k, int64_t n,
                  std::vector<std::pair<double, int>>& nb) {
    const double inf = std::numeric_limits<double>::infinity();
    if (mode == 0) return tree.knn(q, k, inf, nb);
    if (mode == 1) return tree.knn(q, (int)std::min<int64_t>(n, 1 << 30), radius * radius, nb);
    return tree.knn(q, k, radius * radius, nb);
}

// PointCloud::EstimateNormals(search_param, fast_normal_computation=true) (A.8).  prev may be NULL
// (no pre-existing normals); otherwise the new normal is flipped to agree with it.
void orc_estimate_normals(const double* pts, int64_t n, int mode, double radius, int max_nn, const double* prev, double* out) {
    orc::KDTree tree; tree.build(pts, n);
#pragma omp parallel
    {
        std::vector<std::pair<double, int>> nb;
#pragma omp for schedule(static)
        for (int64_t i = 0; i < n; ++i) {
            V3 nrm{0, 0, 1};
            if (search(tree, pts + i * 3, mode, radius, max_nn, n, nb) >= 3) {
                double C[6];
                covariance(pts, nb, C);
                nrm = fast_eigen3x3(C);
            }   // else covariance = Identity -> FastEigen3x3 returns (0,0,1)
            if (std::sqrt(dot(nrm, nrm)) == 0.0) nrm = prev ? V3{prev[i * 3], prev[i * 3 + 1], prev[i * 3 + 2]} : V3{0, 0, 1};
            if (prev && nrm.x * prev[i * 3] + nrm.y * prev[i * 3 + 1] + nrm.z * prev[i * 3 + 2] < 0.0) nrm = scale(nrm, -1.0);
            out[i * 3] = nrm.x; out[i * 3 + 1] = nrm.y; out[i * 3 + 2] = nrm.z;
        }
    }
}

// PointCloud::RemoveStatisticalOutliers(nb_neighbors, std_ratio) (A.9): kept indices (ascending) -> idx_out,
// returns their number; avg_out (optional) receives the per-point mean neighbour distance.
int64_t orc_statistical_outlier(const double* pts, int64_t n, int nb_neighbors, double std_ratio, int64_t* idx_out, double* avg_out) {
    if (n == 0) return 0;
    orc::KDTree tree; tree.build(pts, n);
    std::vector<double> avg((size_t)n, -1.0);
    int64_t valid = 0;
#pragma omp parallel
    {
        std::vector<std::pair<double, int>> nb;
#pragma omp for schedule(static) reduction(+ : valid)
        for (int64_t i = 0; i < n; ++i) {
            const int m = tree.knn(pts + i * 3, nb_neighbors, std::numeric_limits<double>::infinity(), nb);
            if (m > 0) {
                ++valid;
                double s = 0.0;
                for (const auto& e : nb) s += std::sqrt(e.first);
                avg[i] = s / (double)m;
            }
        }
    }
    if (avg_out) std::memcpy(avg_out, avg.data(), sizeof(double) * (size_t)n);
    if (valid == 0) return 0;
    double mean = 0.0;
    for (int64_t i = 0; i < n; ++i) mean += avg[i] > 0 ? avg[i] : 0.0;
    mean /= (double)valid;
    double sq = 0.0;
    for (int64_t i = 0; i < n; ++i) sq += avg[i] > 0 ? (avg[i] - mean) * (avg[i] - mean) : 0.0;
    const double sd = std::sqrt(sq / (double)(valid - 1));
    const double thr = mean + std_ratio * sd;
    int64_t m = 0;
    for (int64_t i = 0; i < n; ++i) if (avg[i] > 0 && avg[i] < thr) idx_out[m++] = i;
    return m;
}

// InitializePointCloudForColoredICP(target, KDTreeSearchParamHybrid(radius, max_nn)) (A.6): per-point colour gradient.
void orc_color_gradient(const double* pts, const double* nrm, const double* col, int64_t n, double radius, int max_nn, double* grad) {
    orc::KDTree tree; tree.build(pts, n);
#pragma omp parallel
    {
        std::vector<std::pair<double, int>> nb;
#pragma omp for schedule(static)
        for (int64_t k = 0; k < n; ++k) {
            grad[k * 3] = grad[k * 3 + 1] = grad[k * 3 + 2] = 0.0;
            const double* vt = pts + k * 3; const double* nt = nrm + k * 3;
            const double it = (col[k * 3] + col[k * 3 + 1] + col[k * 3 + 2]) / 3.0;
            const int nn = tree.knn(vt, max_nn, radius * radius, nb);
            if (nn < 4) continue;
            double AtA[9] = {0}, Atb[3] = {0};
            for (int i = 1; i < nn; ++i) {
                const int a = nb[i].second;
                const double* va = pts + (int64_t)a * 3;
                const double d = (va[0] - vt[0]) * nt[0] + (va[1] - vt[1]) * nt[1] + (va[2] - vt[2]) * nt[2];
                const double row[3] = {(va[0] - d * nt[0]) - vt[0], (va[1] - d * nt[1]) - vt[1], (va[2] - d * nt[2]) - vt[2]};
                const double b = (col[(int64_t)a * 3] + col[(int64_t)a * 3 + 1] + col[(int64_t)a * 3 + 2]) / 3.0 - it;
                for (int r = 0; r < 3; ++r) { for (int c = 0; c < 3; ++c) AtA[r * 3 + c] += row[r] * row[c]; Atb[r] += row[r] * b; }
            }
            const double w = (double)(nn - 1);
            const double row[3] = {w * nt[0], w * nt[1], w * nt[2]};
            for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) AtA[r * 3 + c] += row[r] * row[c];
            ldlt3(AtA, Atb, grad + k * 3);
        }
    }
}

// Neighbour lists for cross-checks (vs scipy.spatial.cKDTree): idx_out [n, k] padded with -1.
void orc_knn(const double* pts, int64_t n, const double* q, int64_t nq, int mode, double radius, int k, int32_t* idx_out, double* d2_out) {
    orc::KDTree tree; tree.build(pts, n);
    std::vector<std::pair<double, int>> nb;
    for (int64_t i = 0; i < nq; ++i) {
        const int m = search(tree, q + i * 3, mode, radius, k, n, nb);
        for (int j = 0; j < k; ++j) {
            idx_out[i * k + j] = j < m ? nb[j].second : -1;
            d2_out[i * k + j] = j < m ? nb[j].first : 0.0;
        }
    }
}

}  // extern "C"
