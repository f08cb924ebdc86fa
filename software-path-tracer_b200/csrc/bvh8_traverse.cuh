// Stack-based traversal of the 8-wide quantised BVH (bvh8.h) + the triangle test.  Written
// __host__ __device__ so that tests/cpp/bvh8_check.cu can run the very same code on the CPU against a
// brute-force search (that host build is test scaffolding only; the product calls it from CUDA kernels).
//
// Exactness: the triangle test is Moller-Trumbore on the stored (v0, e1, e2) record in a FIXED operation
// order (no FMA: the TU is compiled with -fmad=false), so t is bit-identical to the CPU oracle's.  The
// closest hit is "smallest t >= tnear, ties to the smallest caller triangle index", which does not depend
// on traversal order.  Box tests only have to be conservative; they use explicit fmaf and a widened exit
// distance.
#pragma once
#include "common.cuh"

namespace cudapt {

#ifdef __CUDA_ARCH__
#define PT_LDG4(ptr) __ldg(reinterpret_cast<const uint4*>(ptr))
#define PT_LDGF4(ptr) __ldg(reinterpret_cast<const float4*>(ptr))
__device__ __forceinline__ float pt_as_float(uint32_t u) { return __uint_as_float(u); }
__device__ __forceinline__ uint32_t pt_as_uint(float f) { return __float_as_uint(f); }
__device__ __forceinline__ int pt_bfind(uint32_t x) { return 31 - __clz(x); }
__device__ __forceinline__ int pt_popc(uint32_t x) { return __popc(x); }
#else
#define PT_LDG4(ptr) (*reinterpret_cast<const uint4*>(ptr))
#define PT_LDGF4(ptr) (*reinterpret_cast<const float4*>(ptr))
inline float pt_as_float(uint32_t u) {
  float f;
  memcpy(&f, &u, 4);
  return f;
}
inline uint32_t pt_as_uint(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  return u;
}
inline int pt_bfind(uint32_t x) { return 31 - __builtin_clz(x); }
inline int pt_popc(uint32_t x) { return __builtin_popcount(x); }
#endif

// 2^23 + (byte J of `word`) as a float: PRMT splices the byte into the mantissa of 2^23 (bit pattern `magic` =
// 0x4B000000).  ONE full-rate instruction per conversion; the 2^23 bias is folded into the slab constants of
// intersect_node().  (History, profiles/r01_notes.md: shift/mask + quarter-rate I2F -> PRMT + FADD -> PRMT.)
// `magic` is carried in a register (RayPre::magic, from a kernel argument) so that ptxas keeps the SELECTOR
// as the PRMT immediate; with a literal it kept 0x4B000000 as the immediate and re-materialised a selector
// register for every PRMT.
template <int J>
__host__ __device__ __forceinline__ float byte_as_biased_float(uint32_t word, uint32_t magic) {
#ifdef __CUDA_ARCH__
  uint32_t v;
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(v) : "r"(word), "r"(magic), "n"(0x7540 + J));
  return __uint_as_float(v);
#else
  (void)magic;
  return 8388608.0f + (float)((word >> (8 * J)) & 0xFFu);
#endif
}

struct TraceHit {
  float t;        // +inf = miss
  uint32_t tri;   // leaf-order triangle index, 0xFFFFFFFF = miss
  uint32_t prim;  // caller's triangle index (tie-break key)
};

struct RayPre {
  f3 o, d, idir;
  uint32_t octinv;  // 7 ^ (direction sign octant)
  uint32_t magic;   // 0x4B000000, see byte_as_biased_float()
  uint32_t perm_lut;  // device: shared-space byte address of the 256-entry xor_permute8(., octinv) table (0 = none)
};

__host__ __device__ __forceinline__ RayPre prepare_ray(f3 o, f3 d, uint32_t magic = 0x4B000000u) {
  RayPre r;
  r.magic = magic;
  r.o = o;
  r.d = d;
  const float kTiny = 1e-20f;
  const float dx = fabsf(d.x) < kTiny ? copysignf(kTiny, d.x) : d.x;
  const float dy = fabsf(d.y) < kTiny ? copysignf(kTiny, d.y) : d.y;
  const float dz = fabsf(d.z) < kTiny ? copysignf(kTiny, d.z) : d.z;
  r.idir = mk3(1.0f / dx, 1.0f / dy, 1.0f / dz);
  const uint32_t oct = (dx < 0.f ? 1u : 0u) | (dy < 0.f ? 2u : 0u) | (dz < 0.f ? 4u : 0u);
  r.octinv = 7u ^ oct;
  r.perm_lut = 0u;
  return r;
}

// Entry distance used by the box tests: slabs are entered slightly early so that hits at exactly tnear are never
// culled by box rounding; strictly non-zero (see the sign-bit test in intersect_node()).
__host__ __device__ __forceinline__ float box_tmin(float tnear) { return tnear > 0.0f ? tnear * 0.5f : -1e-30f; }

// Moller-Trumbore, two-sided, fixed operation order with FUSED products (explicit fmaf: one rounding each, the
// same nesting as oracle/twin/twin.cpp; DESIGN.md "Mesh integrator numerics").  Inside test on
// the det-scaled barycentrics with the sign of det folded in by an exact sign-bit flip; the one IEEE division
// (t = (e2 . q) / det) is only reached by candidates that are inside, so the common reject paths cost ~25 and
// ~45 instructions instead of ~90 (profiles/r01_notes.md: triangle tests ran at 4/32 lanes, 27 % of issue).
__host__ __device__ __forceinline__ float dot3f(f3 a, f3 b) { return fmaf(a.z, b.z, fmaf(a.y, b.y, a.x * b.x)); }
__host__ __device__ __forceinline__ f3 cross3f(f3 a, f3 b) {
  return mk3(fmaf(a.y, b.z, -(b.y * a.z)), fmaf(a.z, b.x, -(b.z * a.x)), fmaf(a.x, b.y, -(b.x * a.y)));
}
__host__ __device__ __forceinline__ bool tri_test(f3 v0, f3 e1, f3 e2, f3 o, f3 d, float& t) {
  const f3 pv = cross3f(d, e2);
  const float det = dot3f(e1, pv);
  if (det == 0.0f) return false;
  const uint32_t sb = pt_as_uint(det) & 0x80000000u;
  const float adet = pt_as_float(pt_as_uint(det) ^ sb);
  const f3 tv = o - v0;
  const float u = pt_as_float(pt_as_uint(dot3f(tv, pv)) ^ sb);
  if (!(u >= 0.0f && u <= adet)) return false;
  const f3 qv = cross3f(tv, e1);
  const float v = pt_as_float(pt_as_uint(dot3f(d, qv)) ^ sb);
  if (!(v >= 0.0f && (u + v) <= adet)) return false;
  t = dot3f(e2, qv) / det;
  return true;
}

// Tests the 8 quantised child boxes of one node; returns an 8-bit mask, bit s = the box in slot s is hit.
__host__ __device__ __forceinline__ uint32_t intersect_node(const uint4 w0, const uint4 w2, const uint4 w3, const uint4 w4,
                                                            const RayPre& r, float tmin, float tmax) {
  const float px = pt_as_float(w0.x), py = pt_as_float(w0.y), pz = pt_as_float(w0.z);
  const uint32_t ew = w0.w;
  const float adjx = pt_as_float((ew & 0xFFu) << 23) * r.idir.x;
  const float adjy = pt_as_float(((ew >> 8) & 0xFFu) << 23) * r.idir.y;
  const float adjz = pt_as_float(((ew >> 16) & 0xFFu) << 23) * r.idir.z;
  // t = q * adj + (p - o) * idir with q = (2^23 + q) - 2^23: the bias goes into the additive constant, and the
  // rounding error that constant then carries (<= |org2| * 2^-24, i.e. about one quantisation step) is given
  // back as slack, entry side down, exit side up, so the test stays conservative.
  const float kBias = 8388608.0f;
  const float orgx = (px - r.o.x) * r.idir.x, orgy = (py - r.o.y) * r.idir.y, orgz = (pz - r.o.z) * r.idir.z;
#ifdef __CUDA_ARCH__
  // directed rounding: the entry constant is rounded down, the exit constant up (FFMA.RM / FFMA.RP, full
  // rate), so the boxes grow by at most one ulp of 2^23 * adj = one quantisation step in total
  const float o0x = __fmaf_rd(-kBias, adjx, orgx), o1x = __fmaf_ru(-kBias, adjx, orgx);
  const float o0y = __fmaf_rd(-kBias, adjy, orgy), o1y = __fmaf_ru(-kBias, adjy, orgy);
  const float o0z = __fmaf_rd(-kBias, adjz, orgz), o1z = __fmaf_ru(-kBias, adjz, orgz);
#else
  const float org2x = fmaf(-kBias, adjx, orgx), org2y = fmaf(-kBias, adjy, orgy), org2z = fmaf(-kBias, adjz, orgz);
  const float slx = fabsf(org2x) * 1.3e-7f, sly = fabsf(org2y) * 1.3e-7f, slz = fabsf(org2z) * 1.3e-7f;
  const float o0x = org2x - slx, o1x = org2x + slx;
  const float o0y = org2y - sly, o1y = org2y + sly;
  const float o0z = org2z - slz, o1z = org2z + slz;
#endif
  const bool nx = r.idir.x < 0.f, ny = r.idir.y < 0.f, nz = r.idir.z < 0.f;
  // widened exit so rounding in the slab arithmetic can never cull a box the exact test would enter
  const float tmax_w = tmax * 1.000002f;
  const uint32_t magic = r.magic;
  uint32_t hits = 0;
  // Hit <=> cmin <= cmax.  On the device the test is the SIGN BIT of (cmax - cmin) (FADD, fma pipe), shifted into
  // the mask by one funnel shift per child: 9 alu-pipe instructions per node instead of FSETP + SEL + IADD3 = 20
  // (the node test is alu-pipe bound, profiles/r01_notes.md).  cmin > 0 always (tmin > 0 or tiny negative, see
  // the callers), so cmax - cmin is never -0 for cmin <= cmax, and no operand is NaN.
#ifdef __CUDA_ARCH__
#define PT_RECORD(BIT) hits = __funnelshift_l(__float_as_uint(cmax - cmin), hits, 1);
#else
#define PT_RECORD(BIT) if (cmin <= cmax) hits |= 1u << (BIT);
#endif
#pragma unroll
  for (int half = 1; half >= 0; --half) {  // children 7..0, so that the funnel leaves child k in bit k
    const uint32_t lox = half ? w2.y : w2.x, loy = half ? w2.w : w2.z, loz = half ? w3.y : w3.x;
    const uint32_t hix = half ? w3.w : w3.z, hiy = half ? w4.y : w4.x, hiz = half ? w4.w : w4.z;
    const uint32_t nearx = nx ? hix : lox, farx = nx ? lox : hix;
    const uint32_t neary = ny ? hiy : loy, fary = ny ? loy : hiy;
    const uint32_t nearz = nz ? hiz : loz, farz = nz ? loz : hiz;
#define PT_CHILD(J)                                                                \
    {                                                                              \
      const float t0x = fmaf(byte_as_biased_float<J>(nearx, magic), adjx, o0x);    \
      const float t0y = fmaf(byte_as_biased_float<J>(neary, magic), adjy, o0y);    \
      const float t0z = fmaf(byte_as_biased_float<J>(nearz, magic), adjz, o0z);    \
      const float t1x = fmaf(byte_as_biased_float<J>(farx, magic), adjx, o1x);     \
      const float t1y = fmaf(byte_as_biased_float<J>(fary, magic), adjy, o1y);     \
      const float t1z = fmaf(byte_as_biased_float<J>(farz, magic), adjz, o1z);     \
      const float cmin = fmaxf(fmaxf(t0x, t0y), fmaxf(t0z, tmin));                 \
      const float cmax = fminf(fminf(t1x, t1y), fminf(t1z, tmax_w));               \
      /* empty slots (inverted boxes) may pass; imask / tri_valid keep them out */ \
      PT_RECORD(4 * half + J)                                                      \
    }
    PT_CHILD(3) PT_CHILD(2) PT_CHILD(1) PT_CHILD(0)
#undef PT_CHILD
  }
#undef PT_RECORD
#ifdef __CUDA_ARCH__
  return ~hits & 0xFFu;  // the funnel collected sign bits: set = miss
#else
  return hits;
#endif
}

// Bit s of an 8-bit mask -> bit (s ^ k): three conditional swaps (neighbours, pairs, nibbles).
__host__ __device__ __forceinline__ uint32_t xor_permute8(uint32_t x, uint32_t k) {
  const uint32_t a = ((x & 0x55u) << 1) | ((x >> 1) & 0x55u);
  x = (k & 1u) ? a : x;
  const uint32_t b = ((x & 0x33u) << 2) | ((x >> 2) & 0x33u);
  x = (k & 2u) ? b : x;
  const uint32_t c = ((x & 0x0Fu) << 4) | ((x >> 4) & 0x0Fu);
  return (k & 4u) ? c : x;
}
// Stack concept: void push(uint2), uint2 pop(), bool empty().

// One traversal step.  Precondition: ngroup has at least one internal child left (ngroup.y > 0x00FFFFFF).
// Takes the highest-priority child, pushes what is left of the group, fetches that node (5 x LDG.128) and
// tests its 8 children.  On return ngroup describes the fetched node's internal hits (bits 24..31, by
// traversal priority slot ^ octinv) and (tri_base, tri_valid, tmask) its triangle hits (tmask bit b = 8*j + slot = candidate
// triangle tri_base + popc(tri_valid & ((1 << b) - 1))).
template <class Stack>
__host__ __device__ __forceinline__ void node_step(const uint4* __restrict__ nodes, const RayPre& r, float tmin_box,
                                                   float tmax, uint2& ngroup, Stack& stack, uint32_t& tri_base,
                                                   uint32_t& tri_valid, uint32_t& tmask) {
  const int bit = pt_bfind(ngroup.y);
  ngroup.y &= ~(1u << bit);
  const uint32_t base = ngroup.x, imask = ngroup.y & 0xFFu;
  if (ngroup.y > 0x00FFFFFFu) stack.push(ngroup);
  const uint32_t slot = ((uint32_t)bit - 24u) ^ r.octinv;
  const uint32_t rel = (uint32_t)pt_popc(imask & ~(0xFFFFFFFFu << slot));
  const uint4* np = nodes + 5 * (size_t)(base + rel);
  const uint4 w0 = PT_LDG4(np + 0), w1 = PT_LDG4(np + 1), w2 = PT_LDG4(np + 2), w3 = PT_LDG4(np + 3), w4 = PT_LDG4(np + 4);
  const uint32_t hits = intersect_node(w0, w2, w3, w4, r, tmin_box, tmax);
  const uint32_t node_imask = w0.w >> 24;
#ifdef __CUDA_ARCH__
  // priority permutation from a shared-memory table (one LDS.U8 on the lsu pipe instead of 12 alu-pipe
  // instructions; the kernels fill the table with xor_permute8 itself, see fill_perm_lut())
  uint32_t prio;
  asm("ld.shared.u8 %0, [%1];" : "=r"(prio) : "r"(r.perm_lut + (hits & node_imask)));
#else
  const uint32_t prio = xor_permute8(hits & node_imask, r.octinv);
#endif
  ngroup = make_uint2(w1.x, (prio << 24) | node_imask);
  tri_base = w1.y;
  tri_valid = w1.z;
  tmask = (hits * 0x010101u) & tri_valid;  // bit 8*j + s: slot s was hit and owns a j-th triangle
}

// Tests triangle `tri`; returns true when an any-hit query is satisfied.  Closest-hit rule: smallest
// t >= tnear, ties to the smallest caller index.
template <bool kAnyHit>
__host__ __device__ __forceinline__ bool tri_step(const float4* __restrict__ tris, uint32_t tri, const RayPre& r, float tnear,
                                                  float tmax, TraceHit& hit) {
  const float4* tp = tris + 3 * (size_t)tri;
  const float4 a = PT_LDGF4(tp + 0), b = PT_LDGF4(tp + 1), c = PT_LDGF4(tp + 2);
  float t;
  if (!tri_test(mk3(a.x, a.y, a.z), mk3(b.x, b.y, b.z), mk3(c.x, c.y, c.z), r.o, r.d, t)) return false;
  if (!(t >= tnear)) return false;
  const uint32_t prim = pt_as_uint(a.w);
  if (kAnyHit) {
    if (t < tmax) {
      hit.t = t;
      hit.tri = tri;
      hit.prim = prim;
      return true;
    }
  } else if (t < hit.t || (t == hit.t && prim < hit.prim)) {
    hit.t = t;
    hit.tri = tri;
    hit.prim = prim;
  }
  return false;
}

// Runs one ray to completion (host tests; the CUDA kernels drive node_step/tri_step themselves so they can
// refill lanes between steps).  kAnyHit = false: closest hit in [tnear, tmax).  kAnyHit = true: returns at the
// first triangle with tnear <= t < tmax (hit.tri != ~0u means occluded).
template <bool kAnyHit, bool kCount, class Stack>
__host__ __device__ __forceinline__ void bvh8_trace(const uint4* __restrict__ nodes, const float4* __restrict__ tris,
                                                    const RayPre& r, float tnear, float tmax, Stack& stack, TraceHit& hit,
                                                    uint32_t& n_nodes, uint32_t& n_tris) {
  hit.t = tmax;
  hit.tri = 0xFFFFFFFFu;
  hit.prim = 0xFFFFFFFFu;
  // slabs are entered slightly early so hits at exactly tnear are never culled by box rounding
  const float tmin_box = box_tmin(tnear);
  uint2 ngroup = make_uint2(0u, 0x80000000u);  // the root, as "slot 7^octinv of a virtual parent"
  for (;;) {
    uint32_t tri_base, tri_valid, tmask;
    node_step(nodes, r, tmin_box, kAnyHit ? tmax : hit.t, ngroup, stack, tri_base, tri_valid, tmask);
    if (kCount) ++n_nodes;
    while (tmask) {
      const int ti = pt_bfind(tmask);
      tmask &= ~(1u << ti);
      if (kCount) ++n_tris;
      const uint32_t tri = tri_base + (uint32_t)pt_popc(tri_valid & ((1u << ti) - 1u));
      if (tri_step<kAnyHit>(tris, tri, r, tnear, tmax, hit)) return;
    }
    if (ngroup.y <= 0x00FFFFFFu) {
      if (stack.empty()) return;
      ngroup = stack.pop();
    }
  }
}

}  // namespace cudapt
