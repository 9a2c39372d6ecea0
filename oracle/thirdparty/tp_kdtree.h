// TEST INFRASTRUCTURE ONLY (oracle).
//
// Exact nearest-neighbour search standing in for pcl::KdTreeFLANN<PointXYZI> (FLANN KDTreeSingleIndex,
// leaf size 15, eps 0, sorted results).  FLANN is NOT vendored in /root/reference; call sites are
// featureAssociation.cpp:1054,1165,1615-1616,1786-1787 and mapOptmization.cpp:1005-1006,1099,1181,
// 1333-1334.  Parity UNPINNED by reference fixtures; exact search has a unique answer except for
// equal-distance ties, which this stand-in resolves by lowest point index.
//
// A real kd-tree (median split on the widest axis, leaves of <= 15 points) so that CPU-baseline timings
// are not sandbagged by a brute-force stand-in.  Squared distances are float32, summed x,y,z in order.
#pragma once
#include <algorithm>
#include <cfloat>
#include <cstdint>
#include <vector>

namespace tp {

class KdTree3 {
public:
    void build(const float* pts, size_t n, size_t stride_floats) {
        pts_ = pts; n_ = n; stride_ = stride_floats;
        order_.resize(n);
        for (size_t i = 0; i < n; ++i) order_[i] = static_cast<int>(i);
        nodes_.clear();
        if (n) { nodes_.reserve(2 * n / 8 + 4); build_rec(0, static_cast<int>(n)); }
    }
    size_t size() const { return n_; }

    // k nearest, ascending by (distance, index). Returns number found (min(k, n)).
    int knn(const float q[3], int k, std::vector<int>& idx, std::vector<float>& d2) const {
        idx.clear(); d2.clear();
        if (n_ == 0 || k <= 0) return 0;
        Res r; r.k = k;
        search(0, q, r, false, 0.0f);
        finish(r, idx, d2);
        return static_cast<int>(idx.size());
    }
    // all points with squared distance <= radius^2 (FLANN radiusSearch, sorted ascending).
    int radius(const float q[3], float radius, std::vector<int>& idx, std::vector<float>& d2) const {
        idx.clear(); d2.clear();
        if (n_ == 0) return 0;
        Res r; r.k = 0;
        search(0, q, r, true, radius * radius);
        finish(r, idx, d2);
        return static_cast<int>(idx.size());
    }

private:
    struct Node { int lo, hi, axis, left, right; float split; };
    struct Hit { float d; int i; bool operator<(const Hit& o) const { return d < o.d || (d == o.d && i < o.i); } };
    struct Res { int k; std::vector<Hit> heap; };   // max-heap on (d,i) for knn; plain list for radius

    const float* pts_ = nullptr; size_t n_ = 0, stride_ = 4;
    std::vector<int> order_;
    std::vector<Node> nodes_;

    float coord(int p, int a) const { return pts_[static_cast<size_t>(p) * stride_ + a]; }

    int build_rec(int lo, int hi) {
        int id = static_cast<int>(nodes_.size());
        nodes_.push_back(Node{lo, hi, -1, -1, -1, 0.0f});
        if (hi - lo <= 15) return id;
        float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
        for (int i = lo; i < hi; ++i)
            for (int a = 0; a < 3; ++a) { float v = coord(order_[i], a); mn[a] = std::min(mn[a], v); mx[a] = std::max(mx[a], v); }
        int ax = 0;
        for (int a = 1; a < 3; ++a) if (mx[a] - mn[a] > mx[ax] - mn[ax]) ax = a;
        if (!(mx[ax] > mn[ax])) return id;   // all identical: keep as a leaf
        int mid = (lo + hi) / 2;
        std::nth_element(order_.begin() + lo, order_.begin() + mid, order_.begin() + hi,
                         [&](int a, int b) { return coord(a, ax) < coord(b, ax); });
        float split = coord(order_[mid], ax);
        int l = build_rec(lo, mid);
        int r = build_rec(mid, hi);
        nodes_[id].axis = ax; nodes_[id].split = split; nodes_[id].left = l; nodes_[id].right = r;
        return id;
    }

    float worst(const Res& r, bool is_radius, float r2) const {
        if (is_radius) return r2;
        return (static_cast<int>(r.heap.size()) < r.k) ? FLT_MAX : r.heap.front().d;
    }

    void offer(Res& r, bool is_radius, float r2, float d, int i) const {
        Hit h{d, i};
        if (is_radius) { if (d <= r2) r.heap.push_back(h); return; }
        if (static_cast<int>(r.heap.size()) < r.k) { r.heap.push_back(h); std::push_heap(r.heap.begin(), r.heap.end()); }
        else if (h < r.heap.front()) { std::pop_heap(r.heap.begin(), r.heap.end()); r.heap.back() = h; std::push_heap(r.heap.begin(), r.heap.end()); }
    }

    void search(int id, const float q[3], Res& r, bool is_radius, float r2) const {
        const Node& nd = nodes_[id];
        if (nd.axis < 0) {
            for (int i = nd.lo; i < nd.hi; ++i) {
                int p = order_[i];
                float dx = coord(p, 0) - q[0], dy = coord(p, 1) - q[1], dz = coord(p, 2) - q[2];
                float d = dx * dx + dy * dy + dz * dz;
                offer(r, is_radius, r2, d, p);
            }
            return;
        }
        float diff = q[nd.axis] - nd.split;
        int near = diff < 0 ? nd.left : nd.right, far = diff < 0 ? nd.right : nd.left;
        search(near, q, r, is_radius, r2);
        // visit the far side unless it is strictly farther than the current worst (ties must be explored)
        if (diff * diff <= worst(r, is_radius, r2)) search(far, q, r, is_radius, r2);
    }

    void finish(Res& r, std::vector<int>& idx, std::vector<float>& d2) const {
        std::sort(r.heap.begin(), r.heap.end());
        for (const Hit& h : r.heap) { idx.push_back(h.i); d2.push_back(h.d); }
    }
};

}  // namespace tp
