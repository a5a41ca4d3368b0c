// ORACLE -- TEST INFRASTRUCTURE ONLY.  Not part of the product path.
//
// CPU restatement of Open3D 0.18.0 pipelines::registration::RegistrationICP with
// TransformationEstimationPointToPlane / PointToPoint (reference call sites stitcher.py:94-110;
// north_star names the point-to-plane variant), following SURVEY.md Appendix A.5.
// The nearest-neighbour search is an exact KD-tree (Open3D: nanoflann SearchHybrid(max_nn=1)),
// strict d^2 < r^2.  PARITY UNPINNED against the real Open3D binary (see recon_oracle.cpp).
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <numeric>
#include <vector>

#include "oracle_kdtree.h"

namespace {

// y = M(4x4 row-major) * [p,1]
inline void xform(const double* M, const double* p, double* o) {
    for (int r = 0; r < 3; ++r)
        o[r] = M[r * 4 + 0] * p[0] + M[r * 4 + 1] * p[1] + M[r * 4 + 2] * p[2] + M[r * 4 + 3];
}
void matmul4(const double* A, const double* B, double* C) {
    double t[16];
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) {
            double s = 0;
            for (int k = 0; k < 4; ++k) s += A[i * 4 + k] * B[k * 4 + j];
            t[i * 4 + j] = s;
        }
    std::memcpy(C, t, sizeof(t));
}

// Dense symmetric solve A x = b (n<=6) by LDL^T with diagonal pivoting (what Eigen's ldlt()
// does: pick the largest remaining diagonal entry as pivot).  No definiteness check, like
// utility::SolveJacobianSystemAndObtainExtrinsicMatrix.
void ldlt_solve(int n, const double* Ain, const double* b, double* x) {
    double A[36]; int perm[6];
    for (int i = 0; i < n * n; ++i) A[i] = Ain[i];
    for (int i = 0; i < n; ++i) perm[i] = i;
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
    double y[6];
    for (int i = 0; i < n; ++i) {  // L y = P b
        double s = b[perm[i]];
        for (int j = 0; j < i; ++j) s -= A[i * n + j] * y[j];
        y[i] = s;
    }
    for (int i = 0; i < n; ++i) y[i] = (A[i * n + i] != 0.0) ? y[i] / A[i * n + i] : 0.0;
    double z[6];
    for (int i = n - 1; i >= 0; --i) {  // L^T z = y
        double s = y[i];
        for (int j = i + 1; j < n; ++j) s -= A[j * n + i] * z[j];
        z[i] = s;
    }
    for (int i = 0; i < n; ++i) x[perm[i]] = z[i];
}

// utility::TransformVector6dToMatrix4d: R = Rz(x[2]) * Ry(x[1]) * Rx(x[0]), t = x[3:6].
void vec6_to_mat4(const double* x, double* T) {
    const double ca = std::cos(x[0]), sa = std::sin(x[0]);
    const double cb = std::cos(x[1]), sb = std::sin(x[1]);
    const double cg = std::cos(x[2]), sg = std::sin(x[2]);
    const double Rx[9] = {1, 0, 0, 0, ca, -sa, 0, sa, ca};
    const double Ry[9] = {cb, 0, sb, 0, 1, 0, -sb, 0, cb};
    const double Rz[9] = {cg, -sg, 0, sg, cg, 0, 0, 0, 1};
    double t[9], Rm[9];
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) {
        double s = 0; for (int k = 0; k < 3; ++k) s += Rz[i * 3 + k] * Ry[k * 3 + j]; t[i * 3 + j] = s; }
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) {
        double s = 0; for (int k = 0; k < 3; ++k) s += t[i * 3 + k] * Rx[k * 3 + j]; Rm[i * 3 + j] = s; }
    for (int i = 0; i < 16; ++i) T[i] = 0;
    for (int i = 0; i < 3; ++i) { for (int j = 0; j < 3; ++j) T[i * 4 + j] = Rm[i * 3 + j]; T[i * 4 + 3] = x[3 + i]; }
    T[15] = 1;
}

// One-sided Jacobi SVD of a 3x3 (A = U diag(s) V^T), double.  Used for Umeyama.
void svd3(const double* Ain, double* U, double* S, double* V) {
    double A[9]; std::memcpy(A, Ain, sizeof(A));
    for (int i = 0; i < 9; ++i) V[i] = (i % 4 == 0);
    for (int sweep = 0; sweep < 60; ++sweep) {
        double off = 0;
        for (int p = 0; p < 2; ++p) for (int q = p + 1; q < 3; ++q) {
            double alpha = 0, beta = 0, gamma = 0;
            for (int i = 0; i < 3; ++i) { alpha += A[i * 3 + p] * A[i * 3 + p]; beta += A[i * 3 + q] * A[i * 3 + q]; gamma += A[i * 3 + p] * A[i * 3 + q]; }
            off = std::max(off, std::fabs(gamma) / std::sqrt(std::max(alpha * beta, 1e-300)));
            if (gamma == 0.0) continue;
            const double zeta = (beta - alpha) / (2.0 * gamma);
            const double t = (zeta >= 0 ? 1.0 : -1.0) / (std::fabs(zeta) + std::sqrt(1.0 + zeta * zeta));
            const double c = 1.0 / std::sqrt(1.0 + t * t), s = c * t;
            for (int i = 0; i < 3; ++i) {
                const double ap = A[i * 3 + p], aq = A[i * 3 + q];
                A[i * 3 + p] = c * ap - s * aq; A[i * 3 + q] = s * ap + c * aq;
                const double vp = V[i * 3 + p], vq = V[i * 3 + q];
                V[i * 3 + p] = c * vp - s * vq; V[i * 3 + q] = s * vp + c * vq;
            }
        }
        if (off < 1e-15) break;
    }
    for (int j = 0; j < 3; ++j) {
        double n = 0; for (int i = 0; i < 3; ++i) n += A[i * 3 + j] * A[i * 3 + j];
        S[j] = std::sqrt(n);
    }
    // sort descending
    int ord[3] = {0, 1, 2};
    std::sort(ord, ord + 3, [&](int a, int b) { return S[a] > S[b]; });
    double A2[9], V2[9], S2[3];
    for (int j = 0; j < 3; ++j) { S2[j] = S[ord[j]]; for (int i = 0; i < 3; ++i) { A2[i * 3 + j] = A[i * 3 + ord[j]]; V2[i * 3 + j] = V[i * 3 + ord[j]]; } }
    std::memcpy(V, V2, sizeof(V2)); std::memcpy(S, S2, sizeof(S2));
    // U columns = A columns / s; complete degenerate columns by cross products
    for (int j = 0; j < 3; ++j) for (int i = 0; i < 3; ++i) U[i * 3 + j] = S[j] > 1e-300 ? A2[i * 3 + j] / S[j] : 0.0;
    auto col = [&](double* M, int j, double* o) { o[0] = M[j]; o[1] = M[3 + j]; o[2] = M[6 + j]; };
    auto setcol = [&](double* M, int j, const double* v) { M[j] = v[0]; M[3 + j] = v[1]; M[6 + j] = v[2]; };
    const double tiny = 1e-12 * std::max(S[0], 1e-300);
    if (S[0] <= 1e-300) { for (int i = 0; i < 9; ++i) U[i] = (i % 4 == 0); return; }
    if (S[1] <= tiny) {  // rank 1: pick any unit vector orthogonal to u0
        double u0[3]; col(U, 0, u0);
        double e[3] = {0, 0, 0}; int k = 0; for (int i = 1; i < 3; ++i) if (std::fabs(u0[i]) < std::fabs(u0[k])) k = i; e[k] = 1;
        double u1[3] = {u0[1] * e[2] - u0[2] * e[1], u0[2] * e[0] - u0[0] * e[2], u0[0] * e[1] - u0[1] * e[0]};
        const double n = std::sqrt(u1[0] * u1[0] + u1[1] * u1[1] + u1[2] * u1[2]);
        for (int i = 0; i < 3; ++i) u1[i] /= n;
        setcol(U, 1, u1);
    }
    if (S[2] <= tiny) {
        double u0[3], u1[3]; col(U, 0, u0); col(U, 1, u1);
        double u2[3] = {u0[1] * u1[2] - u0[2] * u1[1], u0[2] * u1[0] - u0[0] * u1[2], u0[0] * u1[1] - u0[1] * u1[0]};
        setcol(U, 2, u2);
    }
}
double det3(const double* M) {
    return M[0] * (M[4] * M[8] - M[5] * M[7]) - M[1] * (M[3] * M[8] - M[5] * M[6]) + M[2] * (M[3] * M[7] - M[4] * M[6]);
}

struct Result { double fitness = 0, rmse = 0; std::vector<int32_t> corr; /* per source: target idx or -1 */ int64_t K = 0; };

// GetRegistrationResultAndCorrespondences: 1-NN within max_dist (strict), fitness, inlier_rmse.
void evaluate(const std::vector<double>& src, const orc::KDTree& tree, double max_dist, Result& res) {
    const int64_t Ns = (int64_t)src.size() / 3;
    res.corr.assign(Ns, -1);
    double err2 = 0; int64_t K = 0;
    const double r2 = max_dist * max_dist;
#pragma omp parallel for schedule(static) reduction(+ : err2, K)
    for (int64_t i = 0; i < Ns; ++i) {
        double d2; int j = tree.nearest(&src[i * 3], r2, &d2);
        if (j >= 0) { res.corr[i] = j; err2 += d2; ++K; }
    }
    res.K = K;
    if (K == 0) { res.fitness = 0; res.rmse = 0; }
    else { res.fitness = (double)K / (double)Ns; res.rmse = std::sqrt(err2 / (double)K); }
}

}  // namespace

static int run_icp(const double* src_pts, int64_t Ns, const double* tgt_pts, const double* tgt_normals,
                   int64_t Nt, double max_dist, const double* init, int kind, double rel_fitness,
                   double rel_rmse, int max_iter, double* T_out, double* fitness, double* rmse,
                   int32_t* corr_out, int64_t* n_corr, int* iterations, const double* src_col,
                   const double* tgt_col, const double* tgt_grad, double lambda_geometric) {
    if (max_dist <= 0) return -1;
    if (kind >= 1 && !tgt_normals) return -2;
    if (kind == 2 && !(src_col && tgt_col && tgt_grad)) return -3;
    orc::KDTree tree; tree.build(tgt_pts, Nt);
    double T[16];
    std::memcpy(T, init, sizeof(T));
    std::vector<double> pcd((size_t)Ns * 3);
    for (int64_t i = 0; i < Ns; ++i) xform(T, &src_pts[i * 3], &pcd[i * 3]);
    Result res;
    evaluate(pcd, tree, max_dist, res);
    int it = 0;
    for (; it < max_iter; ++it) {
        double upd[16];
        for (int i = 0; i < 16; ++i) upd[i] = (i % 5 == 0);
        if (res.K > 0) {
            if (kind == 2) {
                // TransformationEstimationForColoredICP::ComputeTransformation (A.6): two residuals per
                // correspondence, geometric sqrt(lambda) (vs-vt).nt and photometric sqrt(1-lambda) (is - is0_proj).
                double JTJ[36] = {0}, JTr[6] = {0};
                const double sg = std::sqrt(lambda_geometric), sp = std::sqrt(1.0 - lambda_geometric);
                for (int64_t i = 0; i < Ns; ++i) {
                    const int j = res.corr[i]; if (j < 0) continue;
                    const double* vs = &pcd[i * 3]; const double* vt = &tgt_pts[(int64_t)j * 3];
                    const double* nt = &tgt_normals[(int64_t)j * 3];
                    const double* dit = &tgt_grad[(int64_t)j * 3];
                    const double dn = (vs[0] - vt[0]) * nt[0] + (vs[1] - vt[1]) * nt[1] + (vs[2] - vt[2]) * nt[2];
                    double J[2][6], r[2];
                    const double cr[3] = {vs[1] * nt[2] - vs[2] * nt[1], vs[2] * nt[0] - vs[0] * nt[2], vs[0] * nt[1] - vs[1] * nt[0]};
                    for (int k = 0; k < 3; ++k) { J[0][k] = sg * cr[k]; J[0][3 + k] = sg * nt[k]; }
                    r[0] = sg * dn;
                    const double vp[3] = {vs[0] - dn * nt[0], vs[1] - dn * nt[1], vs[2] - dn * nt[2]};
                    const double is = (src_col[i * 3] + src_col[i * 3 + 1] + src_col[i * 3 + 2]) / 3.0;
                    const double it = (tgt_col[(int64_t)j * 3] + tgt_col[(int64_t)j * 3 + 1] + tgt_col[(int64_t)j * 3 + 2]) / 3.0;
                    const double is0 = (dit[0] * (vp[0] - vt[0]) + dit[1] * (vp[1] - vt[1]) + dit[2] * (vp[2] - vt[2])) + it;
                    const double M[9] = {1.0 - nt[0] * nt[0], -nt[0] * nt[1], -nt[0] * nt[2], -nt[0] * nt[1], 1.0 - nt[1] * nt[1],
                                         -nt[1] * nt[2], -nt[0] * nt[2], -nt[1] * nt[2], 1.0 - nt[2] * nt[2]};
                    double dm[3];
                    for (int c = 0; c < 3; ++c) dm[c] = (-dit[0]) * M[c] + (-dit[1]) * M[3 + c] + (-dit[2]) * M[6 + c];
                    const double cm[3] = {vs[1] * dm[2] - vs[2] * dm[1], vs[2] * dm[0] - vs[0] * dm[2], vs[0] * dm[1] - vs[1] * dm[0]};
                    for (int k = 0; k < 3; ++k) { J[1][k] = sp * cm[k]; J[1][3 + k] = sp * dm[k]; }
                    r[1] = sp * (is - is0);
                    for (int e = 0; e < 2; ++e)
                        for (int a = 0; a < 6; ++a) { for (int b = 0; b < 6; ++b) JTJ[a * 6 + b] += J[e][a] * J[e][b]; JTr[a] += J[e][a] * r[e]; }
                }
                double nb[6], x[6];
                for (int a = 0; a < 6; ++a) nb[a] = -JTr[a];
                ldlt_solve(6, JTJ, nb, x);
                vec6_to_mat4(x, upd);
            } else if (kind == 1) {
                double JTJ[36] = {0}, JTr[6] = {0};
                // (Open3D: OpenMP with private partials; sum order is unpinned -> serial here.)
                for (int64_t i = 0; i < Ns; ++i) {
                    const int j = res.corr[i]; if (j < 0) continue;
                    const double* vs = &pcd[i * 3]; const double* vt = &tgt_pts[(int64_t)j * 3];
                    const double* nt = &tgt_normals[(int64_t)j * 3];
                    const double r = (vs[0] - vt[0]) * nt[0] + (vs[1] - vt[1]) * nt[1] + (vs[2] - vt[2]) * nt[2];
                    const double J[6] = {vs[1] * nt[2] - vs[2] * nt[1], vs[2] * nt[0] - vs[0] * nt[2],
                                         vs[0] * nt[1] - vs[1] * nt[0], nt[0], nt[1], nt[2]};
                    for (int a = 0; a < 6; ++a) { for (int b = 0; b < 6; ++b) JTJ[a * 6 + b] += J[a] * J[b]; JTr[a] += J[a] * r; }
                }
                double nb[6], x[6];
                for (int a = 0; a < 6; ++a) nb[a] = -JTr[a];
                ldlt_solve(6, JTJ, nb, x);
                vec6_to_mat4(x, upd);
            } else {
                // Eigen::umeyama(src, dst, false)
                double ms[3] = {0, 0, 0}, mt[3] = {0, 0, 0};
                for (int64_t i = 0; i < Ns; ++i) { const int j = res.corr[i]; if (j < 0) continue;
                    for (int k = 0; k < 3; ++k) { ms[k] += pcd[i * 3 + k]; mt[k] += tgt_pts[(int64_t)j * 3 + k]; } }
                for (int k = 0; k < 3; ++k) { ms[k] /= (double)res.K; mt[k] /= (double)res.K; }
                double sigma[9] = {0};
                for (int64_t i = 0; i < Ns; ++i) { const int j = res.corr[i]; if (j < 0) continue;
                    for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b)
                        sigma[a * 3 + b] += (tgt_pts[(int64_t)j * 3 + a] - mt[a]) * (pcd[i * 3 + b] - ms[b]); }
                for (int a = 0; a < 9; ++a) sigma[a] /= (double)res.K;
                double U[9], S[3], V[9];
                svd3(sigma, U, S, V);
                double sgn[3] = {1, 1, 1};
                if (det3(U) * det3(V) < 0) sgn[2] = -1;
                double Rm[9];
                for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b) {
                    double s = 0; for (int k = 0; k < 3; ++k) s += U[a * 3 + k] * sgn[k] * V[b * 3 + k]; Rm[a * 3 + b] = s; }
                for (int a = 0; a < 3; ++a) {
                    for (int b = 0; b < 3; ++b) upd[a * 4 + b] = Rm[a * 3 + b];
                    upd[a * 4 + 3] = mt[a] - (Rm[a * 3 + 0] * ms[0] + Rm[a * 3 + 1] * ms[1] + Rm[a * 3 + 2] * ms[2]);
                }
            }
        }
        matmul4(upd, T, T);
        for (int64_t i = 0; i < Ns; ++i) { double o[3]; xform(upd, &pcd[i * 3], o); pcd[i * 3] = o[0]; pcd[i * 3 + 1] = o[1]; pcd[i * 3 + 2] = o[2]; }
        const double bf = res.fitness, br = res.rmse;
        evaluate(pcd, tree, max_dist, res);
        if (std::fabs(bf - res.fitness) < rel_fitness && std::fabs(br - res.rmse) < rel_rmse) { ++it; break; }
    }
    std::memcpy(T_out, T, sizeof(T));
    *fitness = res.fitness; *rmse = res.rmse;
    if (n_corr) *n_corr = res.K;
    if (corr_out) std::memcpy(corr_out, res.corr.data(), sizeof(int32_t) * (size_t)Ns);
    if (iterations) *iterations = it;
    return 0;
}

extern "C" {

// kind: 0 = PointToPoint (Umeyama, no scaling), 1 = PointToPlane.
// corr_out: int32[Ns] (target index per source point or -1) may be NULL.
// Returns 0, or <0 on the errors Open3D raises (max_dist<=0 -> -1, p2plane without normals -> -2).
int orc_registration_icp(const double* src_pts, int64_t Ns, const double* tgt_pts, const double* tgt_normals,
                         int64_t Nt, double max_dist, const double* init, int kind, double rel_fitness,
                         double rel_rmse, int max_iter, double* T_out, double* fitness, double* rmse,
                         int32_t* corr_out, int64_t* n_corr, int* iterations) {
    return run_icp(src_pts, Ns, tgt_pts, tgt_normals, Nt, max_dist, init, kind, rel_fitness, rel_rmse, max_iter, T_out,
                   fitness, rmse, corr_out, n_corr, iterations, nullptr, nullptr, nullptr, 0.968);
}

// RegistrationColoredICP (stitcher.py:94-102; A.6) once the target's colour gradient is known
// (orc_color_gradient with radius = 2 * max_dist, max_nn = 30).  -3: colours / gradient missing.
int orc_registration_colored_icp(const double* src_pts, const double* src_col, int64_t Ns, const double* tgt_pts,
                                 const double* tgt_normals, const double* tgt_col, const double* tgt_grad, int64_t Nt,
                                 double max_dist, const double* init, double lambda_geometric, double rel_fitness,
                                 double rel_rmse, int max_iter, double* T_out, double* fitness, double* rmse,
                                 int32_t* corr_out, int64_t* n_corr, int* iterations) {
    return run_icp(src_pts, Ns, tgt_pts, tgt_normals, Nt, max_dist, init, 2, rel_fitness, rel_rmse, max_iter, T_out,
                   fitness, rmse, corr_out, n_corr, iterations, src_col, tgt_col, tgt_grad, lambda_geometric);
}

// Exposed for cross-checks in tests (vs scipy.spatial.cKDTree).
void orc_nearest(const double* tgt, int64_t Nt, const double* q, int64_t Nq, double max_dist, int32_t* idx, double* d2) {
    orc::KDTree tree; tree.build(tgt, Nt);
    const double r2 = max_dist * max_dist;
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < Nq; ++i) idx[i] = tree.nearest(&q[i * 3], r2, &d2[i]);
}

}  // extern "C"
