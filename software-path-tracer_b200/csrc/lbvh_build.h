// Device-side BVH8 build (lbvh_build.cu).
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <string>
#include <vector>

namespace cudapt {

struct LbvhResult {
  uint32_t n_nodes = 0;
  uint32_t max_depth = 0;
  std::vector<uint32_t> level_start;  // breadth-first: level d = [level_start[d], level_start[d + 1])
};

// The builder's temporaries (keys, ids, binary tree, work queues, sort space): one allocation, grown on demand and kept
// by the caller between builds; release with cudaFree(p).
struct LbvhScratch {
  void* p = nullptr;
  size_t bytes = 0;
};

// Builds the tree over the nt triangles of (d_vertices, d_indices) into d_nodes (capacity node_capacity >= nt / 2 + 8
// nodes) and d_tris (nt records, leaf order); d_tri_box (nt) and d_node_box (node_capacity) are the refit's scratch
// arrays and stay valid for later refits.  Synchronises the stream.
bool build_lbvh_device(const float* d_vertices, const uint32_t* d_indices, const uint32_t* d_material_id, uint32_t nt,
                       const float world_lo[3], const float world_hi[3], void* d_nodes, uint32_t node_capacity, void* d_tris,
                       void* d_tri_box, void* d_node_box, LbvhResult* out, LbvhScratch* scratch, uint64_t* launches,
                       cudaStream_t stream, std::string* err);

}  // namespace cudapt
