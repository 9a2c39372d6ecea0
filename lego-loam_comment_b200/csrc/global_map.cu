// The map products of mapOptimization that are not on the per-scan path:
//   which 0  publishGlobalMap (mapOptmization.cpp:757-800): key poses within globalMapVisualizationSearchRadius of the
//            robot, thinned by VoxelGrid(1.0) over the POSES (the averaged intensity picks the frame, quirk Q16), their
//            corner + surf + outlier clouds moved to the map frame and filtered with VoxelGrid(0.4)
//            -> /laser_cloud_surround, and finalCloud.pcd at shutdown (731)
//   which 1  cornerMap.pcd  (734-749): corner clouds of ALL key frames, VoxelGrid(0.2)
//   which 2  surfaceMap.pcd (734-751): surf + outlier clouds of all key frames, VoxelGrid(0.4)
//   which 3  trajectory.pcd (754): cloudKeyPoses3D (intensity = key-frame index)
// The key-frame clouds already live in the lane's device pool (sensor frame); one kernel moves the selected ones by
// their key poses into a scratch cloud and the batched VoxelGrid filter (voxel_grid.cu) reduces it.  Scratch memory is
// allocated per call: this runs at 0.2 Hz in the reference (724-728), not per scan.
#include <algorithm>
#include <climits>
#include <cmath>
#include <cstring>
#include <fstream>
#include <limits>
#include <locale>
#include <sstream>
#include <stdexcept>
#include "engine.h"
#include "map_kernels.cuh"

namespace lego {

namespace {
struct GmEntry { int src, n; long long dst; float cR, sR, cP, sP, cY, sY, x, y, z; };
}

// transformPointCloud(cloud, &cloudKeyPoses6D->points[i]) (mapOptmization.cpp:577-606) for a list of key-frame clouds
__global__ void __launch_bounds__(256) gm_assemble_kernel(const float4* __restrict__ pool, const GmEntry* __restrict__ entries, int nE,
                                                          float4* __restrict__ dst) {
    for (int e = blockIdx.x; e < nE; e += gridDim.x) {
        const GmEntry en = entries[e];
        MapSinCos t;
        t.cRoll = en.cR; t.sRoll = en.sR; t.cPitch = en.cP; t.sPitch = en.sP; t.cYaw = en.cY; t.sYaw = en.sY;
        t.tX = en.x; t.tY = en.y; t.tZ = en.z;
        for (int i = threadIdx.x; i < en.n; i += blockDim.x) dst[en.dst + i] = map_transform(t, pool[en.src + i]);
    }
}

void Engine::globalMap(int lane, int which, std::vector<float4>& out) {
    out.clear();
    consumeMapRecords();   // the key frame of the last mapping call may still be in flight
    MoLaneHost& H = moHost[lane];
    const int K = (int)H.keyFrames.size();
    if (K == 0) return;
    if (which == 3) {
        for (int i = 0; i < K; ++i) out.push_back(make_float4(H.keyFrames[i].x, H.keyFrames[i].y, H.keyFrames[i].z, (float)i));
        return;
    }
    // ---- which key frames, in the reference's order ----
    std::vector<int> ids;
    if (which == 0) {
        const float r = 500.0f, r2 = r * r;   // globalMapVisualizationSearchRadius, utility.h:136
        std::vector<std::pair<float, int> > hits;
        for (int i = 0; i < K; ++i) {
            const float dx = H.keyFrames[i].x - H.currentRobotPos[0], dy = H.keyFrames[i].y - H.currentRobotPos[1], dz = H.keyFrames[i].z - H.currentRobotPos[2];
            const float d = dx * dx + dy * dy + dz * dz;
            if (d <= r2) hits.push_back(std::make_pair(d, i));
        }
        std::sort(hits.begin(), hits.end());   // radiusSearch returns ascending distance
        std::vector<float4> poses, ds;
        for (auto& h : hits) { const MoKeyFrame& k = H.keyFrames[h.second]; poses.push_back(make_float4(k.x, k.y, k.z, (float)h.second)); }
        host_voxel_grid(poses, 1.0f, ds);      // downSizeFilterGlobalMapKeyPoses, mapOptmization.cpp:256
        for (const float4& p : ds) ids.push_back((int)p.w);
    } else {
        for (int i = 0; i < K; ++i) ids.push_back(i);
    }
    std::vector<GmEntry> entries;
    long long total = 0;
    auto push = [&](const MoKeyFrame& k, int src, int n) {
        if (n <= 0) return;
        GmEntry e;
        e.src = src; e.n = n; e.dst = total;
        e.cR = cosf(k.roll); e.sR = sinf(k.roll); e.cP = cosf(k.pitch); e.sP = sinf(k.pitch); e.cY = cosf(k.yaw); e.sY = sinf(k.yaw);
        e.x = k.x; e.y = k.y; e.z = k.z;
        entries.push_back(e); total += n;
    };
    for (int id : ids) {
        if (id < 0 || id >= K) continue;
        const MoKeyFrame& k = H.keyFrames[id];
        if (which == 0 || which == 1) push(k, k.poolOff, k.nC);
        if (which == 0 || which == 2) { push(k, k.poolOff + k.nC, k.nS); push(k, k.poolOff + k.nC + k.nS, k.nO); }
    }
    if (total == 0) return;
    if (total > (long long)INT_MAX / 2) throw std::runtime_error("global map: too many points for one filter call");
    const float leaf = which == 1 ? 0.2f : 0.4f;   // downSizeFilterCorner / downSizeFilterSurf / downSizeFilterGlobalMapKeyFrames
    // ---- scratch for this call ----
    DevArena a; a.stream = stream;
    const int cap = (int)total;
    float4* raw = a.alloc<float4>(cap, false);
    float4* ds = a.alloc<float4>(cap, false);
    int* counts = a.alloc<int>(2);
    GmEntry* dEntries = a.alloc<GmEntry>(entries.size(), false);
    VgSeg g; std::memset(&g, 0, sizeof(g));
    g.in = raw; g.nIn = counts; g.out = ds; g.nOut = counts + 1; g.active = nullptr; g.leaf = leaf; g.cap = cap; g.outCap = cap;
    VgBatch b; b.nSeg = 1; b.maxBlocks = (cap + VG_TILE - 1) / VG_TILE;
    g.keys[0] = a.alloc<unsigned>(cap, false); g.keys[1] = a.alloc<unsigned>(cap, false);
    g.vals[0] = a.alloc<unsigned>(cap, false); g.vals[1] = a.alloc<unsigned>(cap, false);
    g.blockHist = a.alloc<int>((size_t)256 * b.maxBlocks); g.blockHeads = a.alloc<int>(b.maxBlocks); g.state = a.alloc<int>(16);
    b.segs = a.alloc<VgSeg>(1, false);
    const int hCounts[2] = {cap, 0};
    cuda_check(cudaMemcpyAsync(b.segs, &g, sizeof(VgSeg), cudaMemcpyHostToDevice, stream), "gm table");
    cuda_check(cudaMemcpyAsync(counts, hCounts, sizeof(hCounts), cudaMemcpyHostToDevice, stream), "gm counts");
    cuda_check(cudaMemcpyAsync(dEntries, entries.data(), sizeof(GmEntry) * entries.size(), cudaMemcpyHostToDevice, stream), "gm entries");
    {
        ProfScope ps(this, "gm_assemble_kernel");
        const int grid = (int)std::min<size_t>(entries.size(), 148 * 8);
        gm_assemble_kernel<<<grid, 256, 0, stream>>>(mo.pool + (size_t)lane * mo.poolCap, dEntries, (int)entries.size(), raw);
        launches += 1;
    }
    { ProfScope ps(this, "vg_filter(globalMap)"); launches += vg_filter(b, stream); }
    int nOut = 0;
    cuda_check(cudaMemcpyAsync(&nOut, counts + 1, sizeof(int), cudaMemcpyDeviceToHost, stream), "gm count back");
    cuda_check(cudaStreamSynchronize(stream), "gm sync");
    out.resize((size_t)nOut);
    if (nOut > 0) cuda_check(cudaMemcpyAsync(out.data(), ds, sizeof(float4) * (size_t)nOut, cudaMemcpyDeviceToHost, stream), "gm cloud back");
    cuda_check(cudaStreamSynchronize(stream), "gm sync");
}

// pcl::io::savePCDFileASCII for a PointXYZI cloud (pcl/io/impl/pcd_io.hpp, PCDWriter::writeASCII with precision 8;
// PCL is not vendored in the reference: layout restated from the published writer, see oracle/thirdparty/tp_pcd.h).
int write_pcd_ascii(const char* path, const float* xyzi, long n) {
    std::ofstream fs(path, std::ios::out | std::ios::trunc);
    if (!fs.is_open()) return -1;
    fs.precision(8);
    fs.imbue(std::locale::classic());
    fs << "# .PCD v0.7 - Point Cloud Data file format\nVERSION 0.7\nFIELDS x y z intensity\nSIZE 4 4 4 4\nTYPE F F F F\nCOUNT 1 1 1 1\n"
       << "WIDTH " << n << "\nHEIGHT 1\nVIEWPOINT 0 0 0 1 0 0 0\nPOINTS " << n << "\nDATA ascii\n";
    std::ostringstream line;
    line.precision(8);
    line.imbue(std::locale::classic());
    for (long i = 0; i < n; ++i) {
        line.str(std::string());
        for (int k = 0; k < 4; ++k) {
            const float v = xyzi[4 * i + k];
            if (std::isnan(v)) line << "nan"; else line << v;
            if (k < 3) line << " ";
        }
        fs << line.str() << "\n";
    }
    fs.close();
    return fs.fail() ? -1 : 0;
}

}  // namespace lego
