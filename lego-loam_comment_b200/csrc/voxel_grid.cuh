// Batched pcl::VoxelGrid<PointXYZI>::filter on the device: the "sort-by-voxel-key reduction" replacing the calls at
// mapOptmization.cpp:1009-1010 (key poses, done on the host), 1058-1063 (local map) and 1070-1089 (current scan).
// One call processes a whole table of independent SEGMENTS (lanes x clouds) with the same launches:
//   bounding box -> voxel index per point (PCL's float operations, no FMA) -> stable LSD radix sort of
//   (index, point id) with 8-bit digits -> one output per distinct index = float32 mean of its members in input order.
// Output order = ascending voxel index, exactly PCL's.  The radix sort is hand-written (no CUB): segments have very
// different sizes that only the device knows, so every kernel sizes its grid for the capacity and blocks whose tile
// lies beyond the segment's live size exit at once; passes whose digit is constant over the segment are skipped.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace lego {

constexpr int VG_TILE = 4096;      // keys per CTA per pass
constexpr int VG_THREADS = 256;

struct VgSeg {
    const float4* in;        // input cloud
    const int* nIn;          // its size (device counter)
    float4* out;             // filtered cloud
    int* nOut;               // its size
    const int* active;       // nullptr = always; else *active == 0 leaves the segment (and nOut) untouched
    float leaf;              // setLeafSize(leaf, leaf, leaf)
    int cap, outCap;         // capacities of in / scratch and of out (points)
    unsigned* keys[2];       // ping-pong scratch [cap]
    unsigned* vals[2];
    int* blockHist;          // [256][maxBlocks] digit-major block histograms of the current pass, then block offsets
    int* blockHeads;         // [maxBlocks]
    int* state;              // [16]: 0..2 ordered-int min xyz, 3..5 ordered-int max xyz, 6 current buffer, 7 pending flip,
                             //       8 skip this pass, 9 live size n, 10 error (output truncated),
                             //       12 the grid would exceed INT_MAX voxels (PCL returns the input unchanged: identity keys)
};

struct VgBatch {
    VgSeg* segs;             // device table
    int nSeg;
    int maxBlocks;           // ceil(max cap / VG_TILE)
};

// Enqueues the launches of one batched filter; returns the number of launches.
int vg_filter(const VgBatch& b, cudaStream_t stream);
// Only the first half: bounding box, voxel keys and the stable radix sort.  Afterwards vg_sorted_vals(seg)[j] is the input
// index of the j-th point in (voxel index, input order) order; out / nOut are not touched.  Used by the incremental local
// map (incremental_map.cu), which sorts just the key frames that entered the neighbourhood.
int vg_sort_keys(const VgBatch& b, cudaStream_t stream);
__device__ __forceinline__ const unsigned* vg_sorted_vals(const VgSeg& s) { return s.vals[s.state[6] ^ s.state[7]]; }

}  // namespace lego
