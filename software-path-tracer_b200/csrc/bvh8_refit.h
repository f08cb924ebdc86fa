// Pieces of the BVH8 builder that the host (bvh_build.cpp: build + host refit) and the device (refit.cu: GPU refit)
// must execute IDENTICALLY: triangle packing, child-box quantisation, and the refit of one node from its children.
// Only IEEE binary32/binary64 + - * /, floor/ceil, frexp/ldexp and comparisons, and both sides are compiled without
// FMA contraction, so a GPU refit leaves exactly the bytes a host refit would (tests/test_gpu_refit.py).
#pragma once
#include <math.h>
#include <stdint.h>

#include "bvh8.h"

#ifdef __CUDACC__
#define BVH8_HD __host__ __device__ __forceinline__
#else
#define BVH8_HD inline
#endif

namespace cudapt {

struct Box3 {
  float lo[3], hi[3];
};
BVH8_HD void box_clear(Box3& b) {
  for (int k = 0; k < 3; ++k) b.lo[k] = INFINITY, b.hi[k] = -INFINITY;
}
BVH8_HD void box_grow(Box3& b, const Box3& o) {
  for (int k = 0; k < 3; ++k) {
    b.lo[k] = o.lo[k] < b.lo[k] ? o.lo[k] : b.lo[k];
    b.hi[k] = o.hi[k] > b.hi[k] ? o.hi[k] : b.hi[k];
  }
}
BVH8_HD float bvh8_min(float a, float b) { return b < a ? b : a; }  // std::min / std::max semantics
BVH8_HD float bvh8_max(float a, float b) { return a < b ? b : a; }

// Packed record + padded box of triangle `prim` (caller index): the record stores (v0, e1 = v1 - v0, e2 = v2 - v0),
// one rounding each; the box encloses v0, v0+e1, v0+e2 AND the caller's vertices, padded by 4e-7 relative.
BVH8_HD void pack_triangle_hd(const float* vertices, const uint32_t* indices, uint32_t prim, uint32_t material, PackedTri* T,
                              Box3* box) {
  const float* a = vertices + 3 * (size_t)indices[3 * (size_t)prim + 0];
  const float* b = vertices + 3 * (size_t)indices[3 * (size_t)prim + 1];
  const float* c = vertices + 3 * (size_t)indices[3 * (size_t)prim + 2];
  for (int k = 0; k < 3; ++k) {
    T->v0[k] = a[k];
    T->e1[k] = b[k] - a[k];
    T->e2[k] = c[k] - a[k];
    const float b2 = a[k] + T->e1[k], c2 = a[k] + T->e2[k];
    const float lo = bvh8_min(bvh8_min(a[k], b[k]), bvh8_min(c[k], bvh8_min(b2, c2)));
    const float hi = bvh8_max(bvh8_max(a[k], b[k]), bvh8_max(c[k], bvh8_max(b2, c2)));
    const float pad = 4e-7f * bvh8_max(fabsf(lo), fabsf(hi)) + 1e-30f;
    box->lo[k] = lo - pad;
    box->hi[k] = hi + pad;
  }
  T->prim = prim;
  T->material = material;
  T->pad = 0;
}

// Per-axis power-of-two grid anchored at the node box minimum and outward-rounded 8-bit child bounds (exact
// arithmetic in double, so p + q * 2^e always encloses the child box).  Writes p, e and the 48 bounds bytes.
BVH8_HD void quantise_node_hd(const Box3& nb, const Box3* slot_box, const bool* slot_used, Bvh8Node* Np) {
  Bvh8Node& N = *Np;
  int ex[3];
  for (int k = 0; k < 3; ++k) {
    N.p[k] = nb.lo[k];
    const double ext = (double)nb.hi[k] - (double)nb.lo[k];
    int e = -60;
    if (ext > 0.0) {
      int kk;
      (void)frexp(ext / 255.0, &kk);
      e = kk;
    }
    ex[k] = e < -125 ? -125 : (e > 126 ? 126 : e);
  }
  for (int k = 0; k < 3; ++k) {
    uint8_t* qlo = k == 0 ? N.qlo_x : (k == 1 ? N.qlo_y : N.qlo_z);
    uint8_t* qhi = k == 0 ? N.qhi_x : (k == 1 ? N.qhi_y : N.qhi_z);
    for (;;) {
      const double scale = ldexp(1.0, ex[k]);
      bool ok = true;
      for (int s = 0; s < 8 && ok; ++s) {
        if (!slot_used[s]) {
          qlo[s] = 255;  // empty slot: inverted box, never hit
          qhi[s] = 0;
          continue;
        }
        double l = floor(((double)slot_box[s].lo[k] - (double)N.p[k]) / scale);
        double h = ceil(((double)slot_box[s].hi[k] - (double)N.p[k]) / scale);
        while ((double)N.p[k] + l * scale > (double)slot_box[s].lo[k]) l -= 1.0;
        while ((double)N.p[k] + h * scale < (double)slot_box[s].hi[k]) h += 1.0;
        l = l < 0.0 ? 0.0 : l;
        if (h > 255.0) {
          ok = false;
          break;
        }
        qlo[s] = (uint8_t)l;
        qhi[s] = (uint8_t)(h < 0.0 ? 0.0 : h);
      }
      if (ok) break;
      ++ex[k];
    }
    N.e[k] = (uint8_t)(ex[k] + 127);
  }
}

BVH8_HD uint32_t bvh8_popc(uint32_t x) {
#ifdef __CUDA_ARCH__
  return (uint32_t)__popc(x);
#else
  return (uint32_t)__builtin_popcount(x);
#endif
}

// Refit of one node whose children are up to date: child boxes from node_box[] (internal children) or tri_box[]
// (leaf triangles, leaf order), new quantisation grid, and the node's own box into node_box[i].
BVH8_HD void refit_node_hd(Bvh8Node* nodes, uint32_t i, Box3* node_box, const Box3* tri_box) {
  Bvh8Node N = nodes[i];
  Box3 slot_box[8], nb;
  bool slot_used[8];
  box_clear(nb);
  for (int s = 0; s < 8; ++s) {
    slot_used[s] = false;
    box_clear(slot_box[s]);
    if (N.imask & (1u << s)) {
      const uint32_t child = N.child_base + bvh8_popc(N.imask & ((1u << s) - 1u));
      slot_box[s] = node_box[child];
      slot_used[s] = true;
    } else {
      for (uint32_t j = 0; j < 3; ++j) {
        const uint32_t bit = 8 * j + (uint32_t)s;
        if (!(N.tri_valid & (1u << bit))) continue;
        const uint32_t t = N.tri_base + bvh8_popc(N.tri_valid & ((1u << bit) - 1u));
        box_grow(slot_box[s], tri_box[t]);
        slot_used[s] = true;
      }
    }
    if (slot_used[s]) box_grow(nb, slot_box[s]);
  }
  quantise_node_hd(nb, slot_box, slot_used, &N);
  nodes[i] = N;
  node_box[i] = nb;
}

}  // namespace cudapt
