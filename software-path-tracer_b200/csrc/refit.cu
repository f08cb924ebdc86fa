// K7: BVH8 refit on the device (dirty-bit -> refit instead of the reference's full rebuild,
// CPT.cpp:122-131,156-160,362-403).  After a vertex update only the 12 B/vertex positions cross PCIe; the packed
// triangles and every node's child boxes + quantisation grid are recomputed where they live:
//   repack_triangles_kernel : one thread per leaf-order triangle -> PackedTri + padded box
//   refit_level_kernel      : one launch per tree level, deepest first (nodes are stored breadth-first, so a level
//                             is a contiguous index range and all its children are already done); one thread per node
// Both run the SAME __host__ __device__ code as the host builder (bvh8_refit.h), in double where the builder uses
// double, so the bytes equal a host refit's.  Launch count = 1 + depth (<= 41).
#include "bvh8_refit.h"
#include "kernels.h"

namespace cudapt {
namespace {

__global__ void __launch_bounds__(256) repack_triangles_kernel(const float* __restrict__ vertices, const uint32_t* __restrict__ indices,
                                                               uint32_t nt, PackedTri* __restrict__ tris, Box3* __restrict__ tri_box) {
  const uint32_t k = blockIdx.x * 256u + threadIdx.x;
  if (k >= nt) return;
  PackedTri T = tris[k];
  Box3 b;
  pack_triangle_hd(vertices, indices, T.prim, T.material, &T, &b);
  tris[k] = T;
  tri_box[k] = b;
}

__global__ void __launch_bounds__(128) refit_level_kernel(Bvh8Node* __restrict__ nodes, uint32_t begin, uint32_t end,
                                                          Box3* __restrict__ node_box, const Box3* __restrict__ tri_box) {
  const uint32_t i = begin + blockIdx.x * 128u + threadIdx.x;
  if (i >= end) return;
  refit_node_hd(nodes, i, node_box, tri_box);
}

}  // namespace

cudaError_t launch_refit(const float* d_vertices, const uint32_t* d_indices, uint32_t nt, void* d_nodes, void* d_tris,
                         void* d_tri_box, void* d_node_box, const uint32_t* level_start, uint32_t n_levels, uint32_t n_nodes,
                         uint64_t* launches, cudaStream_t stream) {
  repack_triangles_kernel<<<(nt + 255u) / 256u, 256, 0, stream>>>(d_vertices, d_indices, nt, static_cast<PackedTri*>(d_tris),
                                                                  static_cast<Box3*>(d_tri_box));
  ++*launches;
  for (uint32_t l = n_levels; l-- > 0;) {
    const uint32_t begin = level_start[l], end = l + 1 < n_levels ? level_start[l + 1] : n_nodes;
    if (end <= begin) continue;
    refit_level_kernel<<<(end - begin + 127u) / 128u, 128, 0, stream>>>(static_cast<Bvh8Node*>(d_nodes), begin, end,
                                                                        static_cast<Box3*>(d_node_box), static_cast<const Box3*>(d_tri_box));
    ++*launches;
  }
  return cudaGetLastError();
}

}  // namespace cudapt
