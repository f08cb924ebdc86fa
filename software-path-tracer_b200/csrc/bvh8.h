// 8-wide quantised BVH ("compressed wide BVH", after Ylitie, Karras & Laine, HPG 2017) -- memory format
// and host builder interface.  Replaces Embree's rtcCommitScene (CPT.cpp:177,403) for triangle scenes.
//
// Node = 80 bytes = five 128-bit words, read with five LDG.128:
//   w0: float px, py, pz (node box minimum)        | u8 ex, ey, ez (biased exponents of the per-axis
//                                                     quantisation step 2^(e-127)) | u8 imask (bit s set <=>
//                                                     child slot s is an internal node)
//   w1: u32 child_base (index of the first internal child; internal children are contiguous, in slot order)
//       u32 tri_base   (index of the first triangle of this node's leaf children; contiguous)
//       u32 tri_valid  (bit 8*j + s set <=> the leaf child in slot s owns a j-th triangle, j < 3; the node's
//                       triangles are stored contiguously in (j, slot) order, so the triangle behind bit b is
//                       tri_base + popc(tri_valid & ((1 << b) - 1)), and the candidates of an 8-bit box-hit
//                       mask h are (h * 0x010101) & tri_valid: one IMAD instead of a bit spread)
//       u32 reserved
//   w2: u8 qlo_x[8], qlo_y[8]      w3: u8 qlo_z[8], qhi_x[8]      w4: u8 qhi_y[8], qhi_z[8]
// Child box = p + q * 2^(e-127), rounded outwards by the builder (exact-arithmetic conservative).
//
// Child slots are assigned so that slot s holds the child lying furthest along -sign_s, sign_s =
// ((s&1)?-1:+1, (s&2)?-1:+1, (s&4)?-1:+1); a ray with direction-sign octant o then visits slots in
// decreasing (s ^ o ^ 7), an approximate front-to-back order that needs no per-ray sorting.
//
// Triangle = 48 bytes = three 128-bit words, stored in leaf order:
//   {v0.xyz, original index}, {e1.xyz = v1-v0, material id}, {e2.xyz = v2-v0, 0}
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace cudapt {

struct Bvh8Node {
  float p[3];
  uint8_t e[3];
  uint8_t imask;
  uint32_t child_base;
  uint32_t tri_base;
  uint32_t tri_valid;
  uint32_t reserved;
  uint8_t qlo_x[8], qlo_y[8], qlo_z[8];
  uint8_t qhi_x[8], qhi_y[8], qhi_z[8];
};
static_assert(sizeof(Bvh8Node) == 80, "Bvh8Node must be 80 bytes");

struct PackedTri {
  float v0[3];
  uint32_t prim;  // index in the caller's triangle array
  float e1[3];
  uint32_t material;
  float e2[3];
  uint32_t pad;
};
static_assert(sizeof(PackedTri) == 48, "PackedTri must be 48 bytes");

struct Bvh8 {
  std::vector<Bvh8Node> nodes;       // nodes[0] = root
  std::vector<PackedTri> tris;       // leaf order
  std::vector<uint32_t> tri_of_prim; // caller index -> leaf-order index
  uint32_t max_depth = 0;            // depth of the 8-wide tree (root = 1)
  std::vector<uint32_t> level_start; // nodes are emitted breadth-first: level d (0 = root) is [level_start[d], level_start[d+1])
  float world_lo[3] = {0, 0, 0}, world_hi[3] = {0, 0, 0};
  double sah_cost = 0.0;
};

constexpr uint32_t kBvh8MaxDepth = 40;  // traversal stack capacity (8 shared + 32 local entries)

// Builds a binned-SAH binary BVH (leaves <= 3 triangles) with `threads` host threads, collapses it to
// 8-wide nodes, orders children by octant and quantises.  Returns false with *err set on failure.
bool build_bvh8(const float* vertices, uint32_t nv, const uint32_t* indices, uint32_t nt, const uint32_t* material_id,
                int threads, Bvh8* out, std::string* err);

// Incremental update: same indices, moved vertices.  Keeps the tree, refits all boxes.
bool refit_bvh8(const float* vertices, uint32_t nv, const uint32_t* indices, uint32_t nt, Bvh8* bvh, std::string* err);

}  // namespace cudapt
