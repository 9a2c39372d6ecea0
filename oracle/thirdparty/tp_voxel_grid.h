// TEST INFRASTRUCTURE ONLY (oracle).
//
// Restatement of the published algorithm of pcl::VoxelGrid<pcl::PointXYZI>::applyFilter (PCL 1.7/1.8,
// pcl/filters/impl/voxel_grid.hpp) with the settings the reference uses: setLeafSize(l,l,l) only, all
// fields down-sampled, no field filter, no minimum point count.  PCL is NOT vendored in /root/reference;
// the reference's call sites are featureAssociation.cpp:225,778-782 and mapOptmization.cpp:249-257,
// 1009-1010,1058-1063,1070-1089.  Parity for this module is therefore UNPINNED by any reference fixture.
//
// Points are 4 floats (x,y,z,intensity).  Output order = ascending linear voxel index; each output is the
// float32 mean of the members taken in sorted order (PCL sorts with std::sort, so the order inside a voxel
// is unspecified there; here members keep ascending input order, which makes the oracle deterministic).
#pragma once
#include <algorithm>
#include <cfloat>
#include <climits>
#include <cmath>
#include <cstdint>
#include <vector>

namespace tp {

struct VoxelKey { int idx; int pt; };

// in: n points (4 floats each). out: resized to 4*m floats. Returns m.
inline size_t voxel_grid_filter(const float* in, size_t n, float leaf, std::vector<float>& out) {
    out.clear();
    if (n == 0) return 0;
    const float inv = 1.0f / leaf;   // Eigen::Array4f::Ones() / leaf_size_
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (size_t i = 0; i < n; ++i)
        for (int a = 0; a < 3; ++a) {
            float v = in[4 * i + a];
            if (v < mn[a]) mn[a] = v;
            if (v > mx[a]) mx[a] = v;
        }
    int64_t d[3];
    for (int a = 0; a < 3; ++a) d[a] = static_cast<int64_t>((mx[a] - mn[a]) * inv) + 1;
    if (d[0] * d[1] * d[2] > static_cast<int64_t>(INT_MAX)) {   // PCL warns and returns the input unchanged
        out.assign(in, in + 4 * n);
        return n;
    }
    int minb[3], maxb[3], divb[3];
    for (int a = 0; a < 3; ++a) {
        minb[a] = static_cast<int>(std::floor(mn[a] * inv));
        maxb[a] = static_cast<int>(std::floor(mx[a] * inv));
        divb[a] = maxb[a] - minb[a] + 1;
    }
    const int mul1 = divb[0], mul2 = divb[0] * divb[1];
    std::vector<VoxelKey> keys(n);
    for (size_t i = 0; i < n; ++i) {
        int i0 = static_cast<int>(std::floor(in[4 * i + 0] * inv) - static_cast<float>(minb[0]));
        int i1 = static_cast<int>(std::floor(in[4 * i + 1] * inv) - static_cast<float>(minb[1]));
        int i2 = static_cast<int>(std::floor(in[4 * i + 2] * inv) - static_cast<float>(minb[2]));
        keys[i].idx = i0 + i1 * mul1 + i2 * mul2;
        keys[i].pt = static_cast<int>(i);
    }
    std::stable_sort(keys.begin(), keys.end(), [](const VoxelKey& a, const VoxelKey& b) { return a.idx < b.idx; });
    size_t first = 0;
    while (first < n) {
        size_t last = first + 1;
        while (last < n && keys[last].idx == keys[first].idx) ++last;
        float s[4] = {0, 0, 0, 0};
        for (size_t k = first; k < last; ++k)
            for (int a = 0; a < 4; ++a) s[a] += in[4 * keys[k].pt + a];
        const float cnt = static_cast<float>(last - first);
        for (int a = 0; a < 4; ++a) out.push_back(s[a] / cnt);
        first = last;
    }
    return out.size() / 4;
}

}  // namespace tp
