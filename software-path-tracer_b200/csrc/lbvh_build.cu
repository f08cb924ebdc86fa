// K8: BVH8 BUILD ON THE DEVICE (linear BVH) -- the "rebuild" half of SURVEY.md 8f-3 and the on-GPU alternative the
// north star names next to the host SAH builder (bvh_build.cpp).  For scenes whose TOPOLOGY changes between frames:
// a 1 M-triangle tree in milliseconds instead of ~0.6 s on 16 host threads, at the price of a worse tree (no SAH).
//
//   1. morton_kernel        63-bit Morton code of every triangle's box centre inside the scene's bounding cube
//   2. cub::DeviceRadixSort (keys, triangle ids)                       [library call: the sort is not our hot path]
//   3. radix_tree_kernel    binary radix tree over the sorted codes, one thread per internal node (Karras 2012:
//                           "Maximizing Parallelism in the Construction of BVHs, Octrees, and k-d Trees"; ties between
//                           equal codes are broken by the index, so the tree stays balanced on duplicates)
//   4. fit_binary_kernel    boxes and triangle counts bottom-up (one thread per leaf, the second arrival at a node
//                           continues upwards)
//   5. collapse_level_kernel (one launch per level of the 8-wide tree) opens binary nodes, largest box first, until a
//                           node has 8 children; subtrees of <= 3 triangles become the triangles of one leaf slot;
//                           greedy octant slot assignment as in the host builder; nodes are allocated level by level,
//                           so the array is breadth-first and the refit kernels apply unchanged
//   6. launch_refit (refit.cu) packs the triangles and computes every child box + quantisation grid with the code the
//                           host builder uses (bvh8_refit.h)
// Closest-hit results do not depend on the tree, so images equal the host-built tree's bit for bit.
#include <cub/device/device_radix_sort.cuh>

#include <string>
#include <vector>

#include "bvh8_refit.h"
#include "kernels.h"
#include "lbvh_build.h"

namespace cudapt {
namespace {

constexpr uint32_t kLeafFlag = 0x80000000u;
constexpr uint32_t kLeafMax = 3;  // triangles per leaf slot (three tri_valid planes)

struct Bin2 {  // binary radix tree: internal nodes 0 .. n-2; leaf k = sorted position k
  uint32_t* left;    // child reference: index | kLeafFlag for a leaf
  uint32_t* right;
  uint32_t* parent;  // of internal nodes [0, n-1) and of leaves [n-1, 2n-1)
  uint32_t* first;   // sorted range of the subtree
  uint32_t* count;   // triangles below (internal nodes)
  Box3* box;         // internal nodes
  uint32_t* arrivals;
};

__device__ __forceinline__ uint64_t spread21(uint32_t v) {  // 21 bits -> every third bit
  uint64_t x = v & 0x1FFFFFu;
  x = (x | (x << 32)) & 0x1F00000000FFFFull;
  x = (x | (x << 16)) & 0x1F0000FF0000FFull;
  x = (x | (x << 8)) & 0x100F00F00F00F00Full;
  x = (x | (x << 4)) & 0x10C30C30C30C30C3ull;
  x = (x | (x << 2)) & 0x1249249249249249ull;
  return x;
}

// also writes the caller-order triangle boxes the bottom-up pass needs
__global__ void __launch_bounds__(256) morton_kernel(const float* __restrict__ vertices, const uint32_t* __restrict__ indices,
                                                     uint32_t nt, float3 lo, float3 inv_extent, uint64_t* __restrict__ keys,
                                                     uint32_t* __restrict__ ids, Box3* __restrict__ prim_box) {
  const uint32_t i = blockIdx.x * 256u + threadIdx.x;
  if (i >= nt) return;
  PackedTri T;
  Box3 b;
  pack_triangle_hd(vertices, indices, i, 0u, &T, &b);
  prim_box[i] = b;
  const float cx = (0.5f * (b.lo[0] + b.hi[0]) - lo.x) * inv_extent.x;
  const float cy = (0.5f * (b.lo[1] + b.hi[1]) - lo.y) * inv_extent.y;
  const float cz = (0.5f * (b.lo[2] + b.hi[2]) - lo.z) * inv_extent.z;
  const float s = 2097151.0f;  // 2^21 - 1
  const uint32_t qx = (uint32_t)fminf(fmaxf(cx * s, 0.0f), s), qy = (uint32_t)fminf(fmaxf(cy * s, 0.0f), s),
                 qz = (uint32_t)fminf(fmaxf(cz * s, 0.0f), s);
  keys[i] = (spread21(qx) << 2) | (spread21(qy) << 1) | spread21(qz);
  ids[i] = i;
}

// length of the common prefix of the (key, index) pairs at sorted positions i and j; -1 outside the array
__device__ __forceinline__ int prefix_len(const uint64_t* __restrict__ keys, int n, int i, int j) {
  if (j < 0 || j >= n) return -1;
  const uint64_t a = keys[i], b = keys[j];
  if (a != b) return __clzll((long long)(a ^ b));
  return 64 + __clz(i ^ j);
}

__global__ void __launch_bounds__(256) radix_tree_kernel(const uint64_t* __restrict__ keys, int n, Bin2 t) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= n - 1) return;
  const int d = prefix_len(keys, n, i, i + 1) - prefix_len(keys, n, i, i - 1) >= 0 ? 1 : -1;
  const int dmin = prefix_len(keys, n, i, i - d);
  int lmax = 2;
  while (prefix_len(keys, n, i, i + lmax * d) > dmin) lmax *= 2;
  int l = 0;
  for (int s = lmax / 2; s >= 1; s /= 2)
    if (prefix_len(keys, n, i, i + (l + s) * d) > dmin) l += s;
  const int j = i + l * d;
  const int dnode = prefix_len(keys, n, i, j);
  int split = 0;
  for (int s = (l + 1) / 2;; s = (s + 1) / 2) {
    if (prefix_len(keys, n, i, i + (split + s) * d) > dnode) split += s;
    if (s <= 1) break;
  }
  const int gamma = i + split * d + min(d, 0);
  const int lo = min(i, j), hi = max(i, j);
  const uint32_t lref = lo == gamma ? (uint32_t)gamma | kLeafFlag : (uint32_t)gamma;
  const uint32_t rref = hi == gamma + 1 ? (uint32_t)(gamma + 1) | kLeafFlag : (uint32_t)(gamma + 1);
  t.left[i] = lref;
  t.right[i] = rref;
  t.first[i] = (uint32_t)lo;
  t.count[i] = (uint32_t)(hi - lo + 1);
  t.parent[(lref & kLeafFlag) ? (uint32_t)(n - 1) + (lref & ~kLeafFlag) : lref] = (uint32_t)i;
  t.parent[(rref & kLeafFlag) ? (uint32_t)(n - 1) + (rref & ~kLeafFlag) : rref] = (uint32_t)i;
  if (i == 0) t.parent[0] = 0xFFFFFFFFu;
}

__device__ __forceinline__ Box3 ref_box(const Bin2& t, const Box3* __restrict__ prim_box, const uint32_t* __restrict__ ids,
                                        uint32_t ref) {
  return (ref & kLeafFlag) ? prim_box[ids[ref & ~kLeafFlag]] : t.box[ref];
}

// boxes written by other threads during this launch: read around the (incoherent, line-granular) L1
__device__ __forceinline__ Box3 ref_box_cg(const Bin2& t, const Box3* __restrict__ prim_box, const uint32_t* __restrict__ ids,
                                           uint32_t ref) {
  if (ref & kLeafFlag) return prim_box[ids[ref & ~kLeafFlag]];
  Box3 b;
  const float* p = reinterpret_cast<const float*>(t.box + ref);
  for (int k = 0; k < 3; ++k) b.lo[k] = __ldcg(p + k), b.hi[k] = __ldcg(p + 3 + k);
  return b;
}

__global__ void __launch_bounds__(256) fit_binary_kernel(Bin2 t, const Box3* __restrict__ prim_box,
                                                         const uint32_t* __restrict__ ids, uint32_t n) {
  const uint32_t k = blockIdx.x * 256u + threadIdx.x;
  if (k >= n) return;
  uint32_t node = t.parent[(n - 1u) + k];
  while (node != 0xFFFFFFFFu) {
    __threadfence();
    if (atomicAdd(&t.arrivals[node], 1u) == 0u) return;  // the sibling subtree is not finished yet
    __threadfence();
    Box3 b = ref_box_cg(t, prim_box, ids, t.left[node]);
    box_grow(b, ref_box_cg(t, prim_box, ids, t.right[node]));
    t.box[node] = b;
    node = t.parent[node];
  }
}

__device__ __forceinline__ float box_area(const Box3& b) {
  const float dx = b.hi[0] - b.lo[0], dy = b.hi[1] - b.lo[1], dz = b.hi[2] - b.lo[2];
  return 2.0f * (dx * dy + dy * dz + dz * dx);
}

struct Work {
  uint32_t node2, node8;
};

// counters: [0] next free 8-wide node, [1] next free triangle slot, [2] entries in the output queue
__global__ void __launch_bounds__(128) collapse_level_kernel(Bin2 t, const Box3* __restrict__ prim_box,
                                                             const uint32_t* __restrict__ ids,
                                                             const uint32_t* __restrict__ material_id, const Work* __restrict__ qin,
                                                             uint32_t n_in, Work* __restrict__ qout, uint32_t* __restrict__ counters,
                                                             Bvh8Node* __restrict__ nodes, PackedTri* __restrict__ tris) {
  const uint32_t w = blockIdx.x * 128u + threadIdx.x;
  if (w >= n_in) return;
  const Work me = qin[w];
  uint32_t ref[8];
  float area[8];
  int n = 2;
  ref[0] = t.left[me.node2];
  ref[1] = t.right[me.node2];
  auto count_of = [&](uint32_t r) { return (r & kLeafFlag) ? 1u : t.count[r]; };
  auto openable = [&](uint32_t r) { return !(r & kLeafFlag) && t.count[r] > kLeafMax; };
  for (int i = 0; i < 2; ++i) area[i] = openable(ref[i]) ? box_area(t.box[ref[i]]) : -1.0f;
  while (n < 8) {
    int pick = -1;
    float pa = -1.0f;
    for (int i = 0; i < n; ++i)
      if (area[i] > pa) pa = area[i], pick = i;
    if (pick < 0) break;
    const uint32_t o = ref[pick];
    ref[pick] = t.left[o];
    ref[n] = t.right[o];
    area[pick] = openable(ref[pick]) ? box_area(t.box[ref[pick]]) : -1.0f;
    area[n] = openable(ref[n]) ? box_area(t.box[ref[n]]) : -1.0f;
    ++n;
  }
  // greedy octant assignment (bvh_build.cpp step 3): slot s takes the child lying furthest along -sign_s
  Box3 cb[8], nb;
  box_clear(nb);
  for (int i = 0; i < n; ++i) {
    cb[i] = ref_box(t, prim_box, ids, ref[i]);
    box_grow(nb, cb[i]);
  }
  int slot_of[8], child_in_slot[8];
  for (int i = 0; i < 8; ++i) slot_of[i] = -1, child_in_slot[i] = -1;
  for (int round = 0; round < n; ++round) {
    float best = INFINITY;
    int bs = -1, bi = -1;
    for (int s = 0; s < 8; ++s) {
      if (child_in_slot[s] >= 0) continue;
      const float sx = (s & 1) ? -1.f : 1.f, sy = (s & 2) ? -1.f : 1.f, sz = (s & 4) ? -1.f : 1.f;
      for (int i = 0; i < n; ++i) {
        if (slot_of[i] >= 0) continue;
        const float cx = 0.5f * (cb[i].lo[0] + cb[i].hi[0]) - 0.5f * (nb.lo[0] + nb.hi[0]);
        const float cy = 0.5f * (cb[i].lo[1] + cb[i].hi[1]) - 0.5f * (nb.lo[1] + nb.hi[1]);
        const float cz = 0.5f * (cb[i].lo[2] + cb[i].hi[2]) - 0.5f * (nb.lo[2] + nb.hi[2]);
        const float c = cx * sx + cy * sy + cz * sz;
        if (c < best) best = c, bs = s, bi = i;
      }
    }
    slot_of[bi] = bs;
    child_in_slot[bs] = bi;
  }
  uint32_t n_internal = 0, n_tris = 0;
  for (int i = 0; i < n; ++i) {
    if (openable(ref[i]))
      ++n_internal;
    else
      n_tris += count_of(ref[i]);
  }
  const uint32_t child_base = n_internal ? atomicAdd(&counters[0], n_internal) : 0u;
  const uint32_t tri_base = n_tris ? atomicAdd(&counters[1], n_tris) : 0u;
  const uint32_t q_base = n_internal ? atomicAdd(&counters[2], n_internal) : 0u;
  Bvh8Node N;
  memset(&N, 0, sizeof(N));
  N.child_base = child_base;
  N.tri_base = tri_base;
  uint32_t next_internal = 0;
  for (int s = 0; s < 8; ++s) {
    const int i = child_in_slot[s];
    if (i < 0) continue;
    if (openable(ref[i])) {
      N.imask |= (uint8_t)(1u << s);
      qout[q_base + next_internal] = Work{ref[i], child_base + next_internal};
      ++next_internal;
    } else {
      const uint32_t cnt = count_of(ref[i]);
      for (uint32_t j = 0; j < cnt; ++j) N.tri_valid |= 1u << (8u * j + (uint32_t)s);
    }
  }
  // triangles of the node in (j, slot) order = ascending bit order of tri_valid; geometry is filled in by the refit
  uint32_t k = tri_base;
  for (uint32_t j = 0; j < kLeafMax; ++j)
    for (int s = 0; s < 8; ++s) {
      if (!(N.tri_valid & (1u << (8u * j + (uint32_t)s)))) continue;
      const uint32_t r = ref[child_in_slot[s]];
      const uint32_t first = (r & kLeafFlag) ? (r & ~kLeafFlag) : t.first[r];
      const uint32_t prim = ids[first + j];
      PackedTri T;
      memset(&T, 0, sizeof(T));
      T.prim = prim;
      T.material = material_id ? material_id[prim] : 0u;
      tris[k++] = T;
    }
  nodes[me.node8] = N;
}

#define LB_TRY(expr)                                              \
  do {                                                            \
    cudaError_t e__ = (expr);                                     \
    if (e__ != cudaSuccess) {                                     \
      *err = std::string(#expr) + ": " + cudaGetErrorString(e__); \
      ok = false;                                                 \
      goto done;                                                  \
    }                                                             \
  } while (0)

}  // namespace

bool build_lbvh_device(const float* d_vertices, const uint32_t* d_indices, const uint32_t* d_material_id, uint32_t nt,
                       const float world_lo[3], const float world_hi[3], void* d_nodes, uint32_t node_capacity, void* d_tris,
                       void* d_tri_box, void* d_node_box, LbvhResult* out, LbvhScratch* scratch, uint64_t* launches,
                       cudaStream_t stream, std::string* err) {
  if (nt < 2) {
    *err = "the device builder needs at least two triangles";
    return false;
  }
  bool ok = true;
  const uint32_t n = nt;
  uint64_t *d_keys = nullptr, *d_keys2 = nullptr;
  uint32_t *d_ids = nullptr, *d_ids2 = nullptr, *d_u32 = nullptr, *d_counters = nullptr;
  Box3 *d_prim_box = nullptr, *d_bin_box = nullptr;
  Work *d_q0 = nullptr, *d_q1 = nullptr;
  void* d_sort_tmp = nullptr;
  size_t sort_bytes = 0;
  Bin2 t{};
  std::vector<uint32_t> level_start;
  uint32_t n_level = 1, n_nodes = 1;
  {
    // one arena for all the temporaries, kept by the caller across rebuilds (a dozen cudaMalloc/cudaFree pairs cost
    // more than the kernels at 1 M triangles)
    LB_TRY(cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, d_keys, d_keys2, d_ids, d_ids2, (int)n, 0, 63, stream));
    size_t off = 0;
    auto take = [&off](size_t bytes) {
      const size_t at = off;
      off += (bytes + 255u) & ~(size_t)255u;
      return at;
    };
    const size_t o_keys = take(sizeof(uint64_t) * n), o_keys2 = take(sizeof(uint64_t) * n), o_ids = take(sizeof(uint32_t) * n);
    const size_t o_ids2 = take(sizeof(uint32_t) * n), o_prim = take(sizeof(Box3) * n), o_bin = take(sizeof(Box3) * n);
    const size_t o_u32 = take(sizeof(uint32_t) * (size_t)n * 7);  // left, right, first, count, arrivals: n each; parent: 2n
    const size_t o_cnt = take(sizeof(uint32_t) * 4), o_q0 = take(sizeof(Work) * (size_t)node_capacity);
    const size_t o_q1 = take(sizeof(Work) * (size_t)node_capacity), o_sort = take(sort_bytes);
    if (scratch->bytes < off) {
      cudaFree(scratch->p);
      scratch->p = nullptr, scratch->bytes = 0;
      LB_TRY(cudaMalloc(&scratch->p, off));
      scratch->bytes = off;
    }
    char* base = static_cast<char*>(scratch->p);
    d_keys = reinterpret_cast<uint64_t*>(base + o_keys), d_keys2 = reinterpret_cast<uint64_t*>(base + o_keys2);
    d_ids = reinterpret_cast<uint32_t*>(base + o_ids), d_ids2 = reinterpret_cast<uint32_t*>(base + o_ids2);
    d_prim_box = reinterpret_cast<Box3*>(base + o_prim), d_bin_box = reinterpret_cast<Box3*>(base + o_bin);
    d_u32 = reinterpret_cast<uint32_t*>(base + o_u32), d_counters = reinterpret_cast<uint32_t*>(base + o_cnt);
    d_q0 = reinterpret_cast<Work*>(base + o_q0), d_q1 = reinterpret_cast<Work*>(base + o_q1);
    d_sort_tmp = base + o_sort;
    t.left = d_u32, t.right = d_u32 + n, t.first = d_u32 + 2 * (size_t)n, t.count = d_u32 + 3 * (size_t)n;
    t.arrivals = d_u32 + 4 * (size_t)n, t.parent = d_u32 + 5 * (size_t)n;
    t.box = d_bin_box;
    const float ex = world_hi[0] - world_lo[0], ey = world_hi[1] - world_lo[1], ez = world_hi[2] - world_lo[2];
    const float3 lo = make_float3(world_lo[0], world_lo[1], world_lo[2]);
    // ONE cell size for the three axes (the bounding cube, not the bounding box): with per-axis normalisation a flat
    // scene (C4: 120 x 0.3 x 120) is cut along its thin axis at every third level, into halves that both span it
    const float emax = fmaxf(ex, fmaxf(ey, ez));
    const float ie = emax > 0.f ? 1.0f / emax : 0.f;
    const float3 inv = make_float3(ie, ie, ie);
    morton_kernel<<<(n + 255u) / 256u, 256, 0, stream>>>(d_vertices, d_indices, n, lo, inv, d_keys, d_ids, d_prim_box);
    ++*launches;
    LB_TRY(cub::DeviceRadixSort::SortPairs(d_sort_tmp, sort_bytes, d_keys, d_keys2, d_ids, d_ids2, (int)n, 0, 63, stream));
    LB_TRY(cudaMemsetAsync(t.arrivals, 0, sizeof(uint32_t) * n, stream));
    radix_tree_kernel<<<(n + 254u) / 256u, 256, 0, stream>>>(d_keys2, (int)n, t);
    fit_binary_kernel<<<(n + 255u) / 256u, 256, 0, stream>>>(t, d_prim_box, d_ids2, n);
    *launches += 2;
    // collapse, one launch per level of the 8-wide tree
    const uint32_t init[4] = {1u, 0u, 0u, 0u};
    const Work root{0u, 0u};
    LB_TRY(cudaMemcpyAsync(d_counters, init, sizeof(init), cudaMemcpyHostToDevice, stream));
    LB_TRY(cudaMemcpyAsync(d_q0, &root, sizeof(root), cudaMemcpyHostToDevice, stream));
    Work *qin = d_q0, *qout = d_q1;
    while (n_level) {
      if (level_start.size() >= kBvh8MaxDepth) {
        *err = "BVH too deep for the traversal stack";
        ok = false;
        goto done;
      }
      level_start.push_back(n_nodes - n_level);
      collapse_level_kernel<<<(n_level + 127u) / 128u, 128, 0, stream>>>(t, d_prim_box, d_ids2, d_material_id, qin, n_level, qout,
                                                                        d_counters, static_cast<Bvh8Node*>(d_nodes),
                                                                        static_cast<PackedTri*>(d_tris));
      ++*launches;
      uint32_t c[3];
      LB_TRY(cudaMemcpyAsync(c, d_counters, sizeof(c), cudaMemcpyDeviceToHost, stream));
      LB_TRY(cudaStreamSynchronize(stream));
      if (c[0] > node_capacity) {
        *err = "internal error: node capacity exceeded";
        ok = false;
        goto done;
      }
      n_level = c[2];
      n_nodes = c[0];
      const uint32_t zero = 0;
      LB_TRY(cudaMemcpyAsync(d_counters + 2, &zero, sizeof(zero), cudaMemcpyHostToDevice, stream));
      std::swap(qin, qout);
      if (n_level == 0 && c[1] != nt) {
        *err = "internal error: triangle count mismatch after the collapse";
        ok = false;
        goto done;
      }
    }
    LB_TRY(launch_refit(d_vertices, d_indices, nt, d_nodes, d_tris, d_tri_box, d_node_box, level_start.data(),
                        (uint32_t)level_start.size(), n_nodes, launches, stream));
    LB_TRY(cudaStreamSynchronize(stream));
    out->n_nodes = n_nodes;
    out->level_start = level_start;
    out->max_depth = (uint32_t)level_start.size();
  }
done:
  return ok;
}

}  // namespace cudapt
