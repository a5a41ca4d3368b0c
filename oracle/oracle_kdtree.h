// ORACLE -- TEST INFRASTRUCTURE ONLY.
// Exact KD-tree over 3-D double points (stands in for nanoflann inside Open3D's
// geometry::KDTreeFlann: SearchKNN / SearchRadius / SearchHybrid).  Median split on the widest
// axis, leaves of <= 10 points, results exact; among equal distances the lower index wins.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <limits>
#include <numeric>
#include <utility>
#include <vector>

namespace orc {

class KDTree {
public:
    void build(const double* pts, int64_t n) {
        pts_ = pts; n_ = n;
        order_.resize((size_t)n);
        std::iota(order_.begin(), order_.end(), 0);
        nodes_.clear();
        if (n > 0) { nodes_.reserve((size_t)(2 * n / kLeaf + 8)); build_rec(0, n); }
    }

    // nearest neighbour with squared distance strictly below r2; -1 if none.
    int nearest(const double* q, double r2, double* d2_out) const {
        if (n_ == 0) { *d2_out = 0; return -1; }
        int best = -1; double bd = r2;
        nn_rec(0, q, best, bd);
        *d2_out = best >= 0 ? bd : 0.0;
        return best;
    }

    // up to k nearest neighbours with squared distance < r2 (r2 = +inf for pure kNN), sorted
    // ascending by (distance, index).  Returns the count.
    int knn(const double* q, int k, double r2, std::vector<std::pair<double, int>>& out) const {
        out.clear();
        if (n_ == 0 || k <= 0) return 0;
        knn_rec(0, q, k, r2, out);
        std::sort_heap(out.begin(), out.end());
        return (int)out.size();
    }

private:
    static constexpr int kLeaf = 10;
    struct Node { int64_t lo, hi; int axis; double split; int left, right; double bmin[3], bmax[3]; };
    const double* pts_ = nullptr; int64_t n_ = 0;
    std::vector<int32_t> order_;
    std::vector<Node> nodes_;

    int build_rec(int64_t lo, int64_t hi) {
        Node nd; nd.lo = lo; nd.hi = hi; nd.left = nd.right = -1; nd.axis = 0; nd.split = 0;
        for (int k = 0; k < 3; ++k) { nd.bmin[k] = std::numeric_limits<double>::infinity(); nd.bmax[k] = -nd.bmin[k]; }
        for (int64_t i = lo; i < hi; ++i)
            for (int k = 0; k < 3; ++k) {
                const double v = pts_[(int64_t)order_[i] * 3 + k];
                nd.bmin[k] = std::min(nd.bmin[k], v); nd.bmax[k] = std::max(nd.bmax[k], v);
            }
        const int id = (int)nodes_.size();
        nodes_.push_back(nd);
        if (hi - lo > kLeaf) {
            int ax = 0; double w = nd.bmax[0] - nd.bmin[0];
            for (int k = 1; k < 3; ++k) if (nd.bmax[k] - nd.bmin[k] > w) { w = nd.bmax[k] - nd.bmin[k]; ax = k; }
            if (w > 0) {
                const int64_t mid = (lo + hi) / 2;
                std::nth_element(order_.begin() + lo, order_.begin() + mid, order_.begin() + hi,
                                 [&](int32_t a, int32_t b) { return pts_[(int64_t)a * 3 + ax] < pts_[(int64_t)b * 3 + ax]; });
                nodes_[id].axis = ax;
                nodes_[id].split = pts_[(int64_t)order_[mid] * 3 + ax];
                const int l = build_rec(lo, mid);
                const int r = build_rec(mid, hi);
                nodes_[id].left = l; nodes_[id].right = r;
            }
        }
        return id;
    }
    inline double box_d2(const Node& nd, const double* q) const {
        double s = 0;
        for (int k = 0; k < 3; ++k) {
            const double d = q[k] < nd.bmin[k] ? nd.bmin[k] - q[k] : (q[k] > nd.bmax[k] ? q[k] - nd.bmax[k] : 0.0);
            s += d * d;
        }
        return s;
    }
    inline double dist2(const double* q, int32_t j) const {
        const double dx = q[0] - pts_[(int64_t)j * 3], dy = q[1] - pts_[(int64_t)j * 3 + 1], dz = q[2] - pts_[(int64_t)j * 3 + 2];
        return dx * dx + dy * dy + dz * dz;
    }
    void nn_rec(int id, const double* q, int& best, double& bd) const {
        const Node& nd = nodes_[id];
        if (box_d2(nd, q) > bd) return;
        if (nd.left < 0) {
            for (int64_t i = nd.lo; i < nd.hi; ++i) {
                const int32_t j = order_[i];
                const double d = dist2(q, j);
                if (d < bd || (d == bd && best >= 0 && j < best)) { bd = d; best = j; }
            }
            return;
        }
        const bool left_first = q[nd.axis] < nd.split;
        nn_rec(left_first ? nd.left : nd.right, q, best, bd);
        nn_rec(left_first ? nd.right : nd.left, q, best, bd);
    }
    void knn_rec(int id, const double* q, int k, double r2, std::vector<std::pair<double, int>>& heap) const {
        const Node& nd = nodes_[id];
        const double bound = (int)heap.size() < k ? r2 : std::min(r2, heap.front().first);
        const double bd = box_d2(nd, q);
        if (bd > bound || (bd >= r2)) return;
        if (nd.left < 0) {
            for (int64_t i = nd.lo; i < nd.hi; ++i) {
                const int32_t j = order_[i];
                const double d = dist2(q, j);
                if (!(d < r2)) continue;
                std::pair<double, int> cand(d, j);
                if ((int)heap.size() < k) { heap.push_back(cand); std::push_heap(heap.begin(), heap.end()); }
                else if (cand < heap.front()) { std::pop_heap(heap.begin(), heap.end()); heap.back() = cand; std::push_heap(heap.begin(), heap.end()); }
            }
            return;
        }
        const bool left_first = q[nd.axis] < nd.split;
        knn_rec(left_first ? nd.left : nd.right, q, k, r2, heap);
        knn_rec(left_first ? nd.right : nd.left, q, k, r2, heap);
    }
};

}  // namespace orc
