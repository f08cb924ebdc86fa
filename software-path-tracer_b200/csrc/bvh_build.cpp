// Host builder of the 8-wide quantised BVH (format: bvh8.h).  Not on the per-frame path: the reference
// builds its Embree scene once per backend lifetime (CPT.cpp:156-160, :403); build time is reported
// separately in cudapt_stats.accel_build_ms.
//
//   1. binned-SAH binary BVH (16 bins x 3 axes, leaves <= 3 triangles), subtrees built by std::threads;
//   2. collapse to <= 8 children per node by repeatedly opening the child with the largest surface area;
//   3. octant-ordered slot assignment (greedy minimum of dot(child centre - node centre, sign_slot));
//   4. outward-rounded 8-bit quantisation of the child boxes against a per-node power-of-two grid;
//   5. breadth-first emission: internal children of a node are contiguous (indexable by popcount of
//      imask), leaf triangles of a node are contiguous (indexable by a 5-bit offset).
#include "bvh8.h"
#include "bvh8_refit.h"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <limits>
#include <thread>

namespace cudapt {
namespace {

constexpr int kBins = 16;
constexpr float kInf = std::numeric_limits<float>::infinity();

struct Box {
  float lo[3] = {kInf, kInf, kInf};
  float hi[3] = {-kInf, -kInf, -kInf};
  void grow(const float* l, const float* h) {
    for (int k = 0; k < 3; ++k) {
      lo[k] = std::min(lo[k], l[k]);
      hi[k] = std::max(hi[k], h[k]);
    }
  }
  void grow(const Box& b) { grow(b.lo, b.hi); }
  float area() const {
    const float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
    if (!(dx >= 0.f)) return 0.f;
    return 2.0f * (dx * dy + dy * dz + dz * dx);
  }
};

struct Node2 {
  Box box;
  uint32_t left = 0, right = 0;   // children (internal)
  uint32_t first = 0, count = 0;  // leaf: range in the permutation
  uint32_t total = 0;             // every node: triangles below it = perm[first .. first + total)
};

struct Builder {
  const float* blo;
  const float* bhi;
  const float* cen;
  std::vector<uint32_t> perm;
  std::vector<Node2> nodes;
  std::atomic<uint32_t> next{1};
  std::atomic<int> spare{0};
  uint32_t leaf_max = 3;   // triangles per leaf (<= 3: the node format has three tri_valid planes)
  float trav_cost = 0.8f;  // SAH cost of one more box test relative to one triangle test

  void build(uint32_t node, uint32_t begin, uint32_t end) {
    Box box, cbox;
    for (uint32_t i = begin; i < end; ++i) {
      const size_t p = perm[i];
      box.grow(blo + 3 * p, bhi + 3 * p);
      cbox.grow(cen + 3 * p, cen + 3 * p);
    }
    nodes[node].box = box;
    nodes[node].first = begin;
    nodes[node].total = end - begin;
    const uint32_t count = end - begin;
    auto make_leaf = [&]() {
      nodes[node].first = begin;
      nodes[node].count = count;
    };
    if (count == 1) return make_leaf();

    // binned SAH over the three axes
    float best_cost = kInf;
    int best_axis = -1, best_bin = -1;
    for (int axis = 0; axis < 3; ++axis) {
      const float ext = cbox.hi[axis] - cbox.lo[axis];
      if (!(ext > 0.f)) continue;
      const float scale = (float)kBins / ext;
      Box bb[kBins];
      uint32_t bc[kBins] = {0};
      for (uint32_t i = begin; i < end; ++i) {
        const size_t p = perm[i];
        int k = (int)((cen[3 * p + axis] - cbox.lo[axis]) * scale);
        k = std::min(std::max(k, 0), kBins - 1);
        bb[k].grow(blo + 3 * p, bhi + 3 * p);
        bc[k]++;
      }
      float right_area[kBins];
      uint32_t right_cnt[kBins];
      Box acc;
      uint32_t n = 0;
      for (int k = kBins - 1; k > 0; --k) {
        acc.grow(bb[k]);
        n += bc[k];
        right_area[k] = acc.area();
        right_cnt[k] = n;
      }
      acc = Box();
      n = 0;
      for (int k = 0; k < kBins - 1; ++k) {
        acc.grow(bb[k]);
        n += bc[k];
        if (n == 0 || right_cnt[k + 1] == 0) continue;
        const float c = acc.area() * (float)n + right_area[k + 1] * (float)right_cnt[k + 1];
        if (c < best_cost) best_cost = c, best_axis = axis, best_bin = k;
      }
    }
    uint32_t mid;
    if (best_axis < 0) {
      if (count <= leaf_max) return make_leaf();
      mid = begin + count / 2;  // coincident centroids: split the index range
    } else {
      const float split_cost = trav_cost + best_cost / std::max(box.area(), 1e-30f);
      if (count <= leaf_max && split_cost >= (float)count) return make_leaf();
      const float ext = cbox.hi[best_axis] - cbox.lo[best_axis];
      const float scale = (float)kBins / ext;
      const float lo = cbox.lo[best_axis];
      auto it = std::partition(perm.begin() + begin, perm.begin() + end, [&](uint32_t p) {
        int k = (int)((cen[3 * (size_t)p + best_axis] - lo) * scale);
        k = std::min(std::max(k, 0), kBins - 1);
        return k <= best_bin;
      });
      mid = (uint32_t)(it - perm.begin());
      if (mid == begin || mid == end) mid = begin + count / 2;
    }
    const uint32_t l = next.fetch_add(2);
    nodes[node].left = l;
    nodes[node].right = l + 1;
    nodes[node].count = 0;
    if (count > 32768 && spare.fetch_sub(1) > 0) {
      std::thread t([this, l, begin, mid]() { build(l, begin, mid); });
      build(l + 1, mid, end);
      t.join();
      spare.fetch_add(1);
    } else {
      if (count > 32768) spare.fetch_add(1);  // undo the failed reservation
      build(l, begin, mid);
      build(l + 1, mid, end);
    }
  }
};

struct Child {
  uint32_t node2;
  Box box;
  bool leaf;
};

// Optimal collapse of the binary tree into 8-wide nodes (the dynamic programme of Ylitie, Karras & Laine, HPG 2017,
// section 3.1): c[i] = cheapest SAH cost of representing the subtree of a binary node as a forest of at most i roots,
// each a leaf (<= leaf_max triangles) or an 8-wide node.  The greedy "open the largest child" collapse leaves the nodes
// 4.5/8 full (78 % of them leaf-only with ~3.5 leaves); this one 6.2/8.  Opt-in, see build_bvh8().
struct Collapse {
  float c[8];       // [1..7]
  uint8_t k[9];     // [2..7]: triangles^W roots given to the left child when the forest of i roots is split (0: same as i-1); [8]: the split of the 8-wide node
  uint8_t leaf1;    // the single-root form is a leaf (else an 8-wide node)
};

void collapse_costs(const std::vector<Node2>& n2, uint32_t n_used, uint32_t leaf_max, float c_node, float c_prim,
                    std::vector<Collapse>* out) {
  out->assign(n_used, Collapse{});
  for (uint32_t n = n_used; n-- > 0;) {  // children have larger indices than their parent
    const Node2& N = n2[n];
    Collapse& R = (*out)[n];
    const float A = N.box.area();
    if (N.count) {
      for (int i = 1; i <= 7; ++i) R.c[i] = A * (float)N.count * c_prim;
      R.leaf1 = 1;
      continue;
    }
    const Collapse &L = (*out)[N.left], &Rt = (*out)[N.right];
    float D[9];
    uint8_t kD[9];
    for (int j = 2; j <= 8; ++j) {
      D[j] = kInf;
      kD[j] = 0;
      for (int k = std::max(1, j - 7); k <= std::min(7, j - 1); ++k) {
        const float v = L.c[k] + Rt.c[j - k];
        if (v < D[j]) D[j] = v, kD[j] = (uint8_t)k;
      }
    }
    const float c_leaf = N.total <= leaf_max ? A * (float)N.total * c_prim : kInf;
    const float c_int = D[8] + A * c_node;
    R.leaf1 = c_leaf <= c_int;
    R.c[1] = R.leaf1 ? c_leaf : c_int;
    R.k[8] = kD[8];
    for (int i = 2; i <= 7; ++i) {
      if (D[i] < R.c[i - 1]) {
        R.c[i] = D[i];
        R.k[i] = kD[i];
      } else {
        R.c[i] = R.c[i - 1];
        R.k[i] = 0;
      }
    }
  }
}

void collect_roots(const std::vector<Node2>& n2, const std::vector<Collapse>& dp, uint32_t m, int i, std::vector<Child>* ch) {
  const Node2& M = n2[m];
  if (M.count) {
    ch->push_back({m, M.box, true});
    return;
  }
  while (i > 1 && dp[m].k[i] == 0) --i;
  if (i == 1) {
    ch->push_back({m, M.box, dp[m].leaf1 != 0});
    return;
  }
  collect_roots(n2, dp, M.left, dp[m].k[i], ch);
  collect_roots(n2, dp, M.right, i - dp[m].k[i], ch);
}

// Step 4 of the build and the triangle packing live in bvh8_refit.h (shared with the GPU refit kernels).
void quantise_node(const Box& nb, const Box* slot_box, const bool* slot_used, Bvh8Node* N) {
  Box3 b3, sb[8];
  std::memcpy(&b3, &nb, sizeof(b3));
  for (int s = 0; s < 8; ++s) std::memcpy(&sb[s], &slot_box[s], sizeof(Box3));
  quantise_node_hd(b3, sb, slot_used, N);
}
void pack_triangle(const float* vertices, const uint32_t* indices, uint32_t i, uint32_t material, PackedTri* T, float* lo3,
                   float* hi3) {
  Box3 b;
  pack_triangle_hd(vertices, indices, i, material, T, &b);
  for (int k = 0; k < 3; ++k) lo3[k] = b.lo[k], hi3[k] = b.hi[k];
}
static_assert(sizeof(Box) == sizeof(Box3), "Box layout");

}  // namespace

bool build_bvh8(const float* vertices, uint32_t nv, const uint32_t* indices, uint32_t nt, const uint32_t* material_id,
                int threads, Bvh8* out, std::string* err) {
  (void)nv;
  if (nt == 0) {
    *err = "empty mesh";
    return false;
  }
  // ---- packed triangles (caller order) and their boxes ------------------------------------------------
  std::vector<PackedTri> tri(nt);
  std::vector<float> blo(3 * (size_t)nt), bhi(3 * (size_t)nt), cen(3 * (size_t)nt);
  for (uint32_t i = 0; i < nt; ++i) {
    pack_triangle(vertices, indices, i, material_id ? material_id[i] : 0u, &tri[i], &blo[3 * (size_t)i], &bhi[3 * (size_t)i]);
    for (int k = 0; k < 3; ++k) cen[3 * (size_t)i + k] = 0.5f * blo[3 * (size_t)i + k] + 0.5f * bhi[3 * (size_t)i + k];
  }
  // ---- 1. binary SAH BVH -------------------------------------------------------------------------------
  Builder B;
  B.blo = blo.data();
  B.bhi = bhi.data();
  B.cen = cen.data();
  B.perm.resize(nt);
  for (uint32_t i = 0; i < nt; ++i) B.perm[i] = i;
  B.nodes.resize(2 * (size_t)nt + 1);
  B.spare.store(std::max(threads, 1) - 1);
  if (const char* e = std::getenv("CUDAPT_BVH_LEAF_MAX")) B.leaf_max = (uint32_t)std::min(std::max(std::atoi(e), 1), 3);
  if (const char* e = std::getenv("CUDAPT_BVH_TRAV_COST")) B.trav_cost = (float)std::atof(e);
  B.build(0, 0, nt);
  const std::vector<Node2>& n2 = B.nodes;

  // ---- 2-5. collapse, order, quantise, emit breadth-first ---------------------------------------------
  out->nodes.clear();
  out->tris.clear();
  out->tris.reserve(nt);
  out->tri_of_prim.assign(nt, 0u);
  out->max_depth = 0;
  out->level_start.clear();
  out->sah_cost = 0.0;
  for (int k = 0; k < 3; ++k) out->world_lo[k] = n2[0].box.lo[k], out->world_hi[k] = n2[0].box.hi[k];
  const double root_area = std::max((double)n2[0].box.area(), 1e-30);

  // CUDAPT_BVH_COLLAPSE=optimal selects the dynamic-programming collapse: 33 % fewer nodes and 4.5 % fewer node steps
  // per ray, but 2 % more triangle tests, which cost this traversal more (profiles/r01_notes.md: C3 step 19.94 vs
  // 19.95 ms, C4 88.5 vs 89.2 ms) -- the greedy collapse stays the default
  const char* ce = std::getenv("CUDAPT_BVH_COLLAPSE");
  const bool optimal = ce && std::string(ce) == "optimal";
  std::vector<Collapse> dp;
  if (optimal) {
    float c_node = 1.0f, c_prim = 1.0f;
    if (const char* e = std::getenv("CUDAPT_BVH_NODE_COST")) c_node = (float)std::atof(e);
    if (const char* e = std::getenv("CUDAPT_BVH_PRIM_COST")) c_prim = (float)std::atof(e);
    collapse_costs(n2, B.next.load(), B.leaf_max, c_node, c_prim, &dp);
  }
  struct Work {
    uint32_t node2, node8, depth;
  };
  std::deque<Work> queue;
  out->nodes.emplace_back();
  queue.push_back({0u, 0u, 1u});
  while (!queue.empty()) {
    const Work w = queue.front();
    queue.pop_front();
    out->max_depth = std::max(out->max_depth, w.depth);
    if (out->level_start.size() < w.depth) out->level_start.push_back(w.node8);  // breadth-first: first node of the level
    std::vector<Child> ch;
    if (n2[w.node2].count || (optimal && dp[w.node2].leaf1)) {
      ch.push_back({w.node2, n2[w.node2].box, true});  // whole scene is one leaf
    } else if (optimal) {
      collect_roots(n2, dp, n2[w.node2].left, dp[w.node2].k[8], &ch);
      collect_roots(n2, dp, n2[w.node2].right, 8 - dp[w.node2].k[8], &ch);
    } else {
      ch.push_back({n2[w.node2].left, n2[n2[w.node2].left].box, n2[n2[w.node2].left].count != 0});
      ch.push_back({n2[w.node2].right, n2[n2[w.node2].right].box, n2[n2[w.node2].right].count != 0});
      while (ch.size() < 8) {
        int pick = -1;
        float pa = -1.f;
        for (int i = 0; i < (int)ch.size(); ++i)
          if (!ch[i].leaf && ch[i].box.area() > pa) pa = ch[i].box.area(), pick = i;
        if (pick < 0) break;
        const Node2& o = n2[ch[pick].node2];
        ch[pick] = {o.left, n2[o.left].box, n2[o.left].count != 0};
        ch.push_back({o.right, n2[o.right].box, n2[o.right].count != 0});
      }
    }
    const int nc = (int)ch.size();
    Box nb;
    for (const Child& c : ch) nb.grow(c.box);
    out->sah_cost += (double)nb.area() / root_area;

    // 3. greedy octant assignment
    int slot_of[8], child_in_slot[8];
    std::fill(slot_of, slot_of + 8, -1);
    std::fill(child_in_slot, child_in_slot + 8, -1);
    float cost[8][8];
    for (int s = 0; s < 8; ++s) {
      const float sx = (s & 1) ? -1.f : 1.f, sy = (s & 2) ? -1.f : 1.f, sz = (s & 4) ? -1.f : 1.f;
      for (int i = 0; i < nc; ++i) {
        const float cx = 0.5f * (ch[i].box.lo[0] + ch[i].box.hi[0]) - 0.5f * (nb.lo[0] + nb.hi[0]);
        const float cy = 0.5f * (ch[i].box.lo[1] + ch[i].box.hi[1]) - 0.5f * (nb.lo[1] + nb.hi[1]);
        const float cz = 0.5f * (ch[i].box.lo[2] + ch[i].box.hi[2]) - 0.5f * (nb.lo[2] + nb.hi[2]);
        cost[s][i] = cx * sx + cy * sy + cz * sz;
      }
    }
    for (int round = 0; round < nc; ++round) {
      float best = kInf;
      int bs = -1, bi = -1;
      for (int s = 0; s < 8; ++s) {
        if (child_in_slot[s] >= 0) continue;
        for (int i = 0; i < nc; ++i) {
          if (slot_of[i] >= 0) continue;
          if (cost[s][i] < best) best = cost[s][i], bs = s, bi = i;
        }
      }
      slot_of[bi] = bs;
      child_in_slot[bs] = bi;
    }

    // 4. quantisation grid
    Bvh8Node N;
    std::memset(&N, 0, sizeof(N));
    {
      Box slot_box[8];
      bool slot_used[8];
      for (int s8 = 0; s8 < 8; ++s8) {
        slot_used[s8] = child_in_slot[s8] >= 0;
        if (slot_used[s8]) slot_box[s8] = ch[child_in_slot[s8]].box;
      }
      quantise_node(nb, slot_box, slot_used, &N);
    }

    // 5. emit children in slot order
    uint32_t n_internal = 0;
    for (int s = 0; s < 8; ++s)
      if (child_in_slot[s] >= 0 && !ch[child_in_slot[s]].leaf) ++n_internal;
    N.child_base = (uint32_t)out->nodes.size();
    N.tri_base = (uint32_t)out->tris.size();
    out->nodes.resize(out->nodes.size() + n_internal);
    uint32_t next_internal = N.child_base;
    for (int s = 0; s < 8; ++s) {
      const int i = child_in_slot[s];
      if (i < 0) continue;
      if (!ch[i].leaf) {
        N.imask |= (uint8_t)(1u << s);
        queue.push_back({ch[i].node2, next_internal++, w.depth + 1});
      } else {
        const uint32_t cnt = n2[ch[i].node2].total;  // 1..3 (a leaf of the binary tree, or a subtree the collapse merged)
        for (uint32_t j = 0; j < cnt; ++j) N.tri_valid |= 1u << (8 * j + s);
      }
    }
    // triangles of the node in (j, slot) order = ascending bit order of tri_valid
    for (uint32_t j = 0; j < 3; ++j)
      for (int s = 0; s < 8; ++s) {
        if (!(N.tri_valid & (1u << (8 * j + s)))) continue;
        const Node2& leaf = n2[ch[child_in_slot[s]].node2];
        const uint32_t prim = B.perm[leaf.first + j];
        out->tri_of_prim[prim] = (uint32_t)out->tris.size();
        out->tris.push_back(tri[prim]);
      }
    out->nodes[w.node8] = N;
  }
  if (out->max_depth > kBvh8MaxDepth) {
    *err = "BVH too deep for the traversal stack (" + std::to_string(out->max_depth) + " levels)";
    return false;
  }
  if (out->tris.size() != nt) {
    *err = "internal error: triangle count mismatch after collapse";
    return false;
  }
  return true;
}

// Refit after a vertex update (same topology): recomputes the packed triangles and, bottom-up, every node's child
// boxes and quantisation grid; tree structure, slot assignment and triangle order are kept.  Nodes were emitted
// breadth-first, so every child has a larger index than its parent and one reverse sweep suffices.
bool refit_bvh8(const float* vertices, uint32_t nv, const uint32_t* indices, uint32_t nt, Bvh8* bvh, std::string* err) {
  (void)nv;
  if (bvh->tris.size() != nt || bvh->nodes.empty()) {
    *err = "refit: triangle count differs from the built scene";
    return false;
  }
  std::vector<Box3> tri_box(nt);  // by leaf-order index
  for (uint32_t k = 0; k < nt; ++k) {
    PackedTri& T = bvh->tris[k];
    const uint32_t prim = T.prim, mat = T.material;
    if (prim >= nt) {
      *err = "refit: corrupt triangle record";
      return false;
    }
    pack_triangle_hd(vertices, indices, prim, mat, &T, &tri_box[k]);
  }
  const size_t nn = bvh->nodes.size();
  std::vector<Box3> node_box(nn);
  for (size_t i = nn; i-- > 0;) refit_node_hd(bvh->nodes.data(), (uint32_t)i, node_box.data(), tri_box.data());
  for (int k = 0; k < 3; ++k) bvh->world_lo[k] = node_box[0].lo[k], bvh->world_hi[k] = node_box[0].hi[k];
  return true;
}

}  // namespace cudapt
