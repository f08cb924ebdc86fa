// K1+K2a: camera rays as PACKETS.  One warp = 32 samples of ONE pixel (a pinhole camera: common origin, directions
// inside one pixel's footprint).  The warp generates the 32 rays (fused raygen: CPT.cpp:53-73 + jitter), walks the
// BVH8 ONCE for all of them with a warp-uniform stack, and tests every candidate triangle with all 32 lanes.
//
// Why: the per-ray kernel (trace_bvh.cu) is alu-pipe bound on the 8-box node test (~210 instructions per lane and
// node) and runs its triangle tests at 3-4 of 32 lanes (profiles/r01_notes.md).  Here the node test is SPLIT over
// the warp -- lane = (child slot c, frustum plane pl): 3 byte conversions + 1 plane evaluation per lane, one ballot
// -- and a triangle test occupies all 32 lanes.
//
// Exactness: the per-ray result is "smallest t >= tnear over all triangles, ties to the smallest caller index",
// evaluated by the SAME tri_test() as the per-ray kernel and the CPU twin; it does not depend on which superset of
// the ray's own candidates was tested.  The packet only has to be conservative: the four bounding planes of the
// 32 directions are fattened by 1e-5 (relative), children are culled in depth only against the farthest current hit
// of the packet (x 1.00001).  Rays of packets that stop paying off (budget of node + triangle steps, e.g. distant
// geometry much smaller than a pixel) are handed to the per-ray kernel through a fall-back queue.
#include <cstdlib>

#include "bvh8_traverse.cuh"
#include "kernels.h"

namespace cudapt {
namespace {

constexpr int kPacketBlock = 128;
constexpr int kPacketWarps = kPacketBlock / 32;
constexpr unsigned kFull = 0xffffffffu;

__device__ __forceinline__ float sel3(f3 v, int ax) { return ax == 0 ? v.x : (ax == 1 ? v.y : v.z); }

// order-preserving float <-> uint key (any sign), for the integer warp reductions
__device__ __forceinline__ uint32_t fkey(float f) {
  const uint32_t b = __float_as_uint(f);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float funkey(uint32_t k) { return __uint_as_float((k & 0x80000000u) ? (k & 0x7FFFFFFFu) : ~k); }

__device__ __forceinline__ uint32_t lds_u8(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t lds_u32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ uint4 lds_u128(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void sts_u32(uint32_t addr, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory"); }
__device__ __forceinline__ void sts_u128(uint32_t addr, uint4 v) {
  asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// hides how a value was computed, so that the compiler keeps it in a register instead of recomputing it
__device__ __forceinline__ uint32_t opaque(uint32_t v) {
  uint32_t r;
  asm volatile("mov.u32 %0, %1;" : "=r"(r) : "r"(v));
  return r;
}

__device__ __forceinline__ const uint32_t* opaque_ptr(const uint32_t* v) {
  const uint32_t* r;
  asm volatile("mov.u64 %0, %1;" : "=l"(r) : "l"(v));
  return r;
}

struct PacketArgs {
  PassParams p;
  uint32_t* fb_queue;   // rays handed to the per-ray kernel
  uint32_t* fb_count;
  int axis_k;           // dominant axis of the camera's forward vector
  float sigma;          // its sign
  int budget;           // 3 per node step + 4 per triangle step
  // hit list (kList): tiles of 64 packets are claimed dynamically; packets with a hit (or handed to the per-ray kernel)
  // leave one descriptor (path id >> 5) per 32-ray chunk in the tile's region, all-sky packets are finished here
  uint32_t* block_counter;
  uint32_t* block_hits;   // descriptors written per tile
  uint32_t* desc;         // tile b owns [b * 64 * kRays, (b + 1) * 64 * kRays)
};

constexpr uint32_t kNodeBytes = 96;            // staged node: 80 B + 3 scale words, padded
constexpr uint32_t kTriSlots = 24;             // a node has at most 24 triangles
constexpr uint32_t kWarpSmem = kNodeBytes + kTriSlots * 48u;

// kRays = rays per lane: a packet is 32 * kRays samples of one pixel (kRays consecutive 32-sample chunks), so the
// node steps -- the larger part of a packet's cost -- are shared by more rays when the pass is deep enough.
//
// kList (every pixel owned): the shade kernel's bounce-0 queue holds only the chunks of packets that hit something.  A
// packet whose rays all miss -- most of an open scene's camera rays -- is FINISHED here: radiance = sky, nothing else
// stored, never seen by the shade kernel (whose per-round bookkeeping was a quarter of the bounce-0 shade for those
// warps, profiles/r01_notes.md).  To keep the queue in image order without a global atomic per packet, a CTA claims
// TILES of 64 consecutive packets (one 8x8 pixel tile of one chunk group) and appends descriptors to the tile's own
// region through a shared-memory counter; hitlist_scan / hitlist_expand then lay the regions out in tile order.
//
// kMini (kRays == 1; passes of 1, 2, 4, 8 or 16 samples, slot-major path ids, kernels.h): a packet is 32 CONSECUTIVE IDS =
// 32 / S neighbouring pixels x S samples, so every lane has its own pixel; the bounding planes come from the rays' own
// directions as before, they are just wider.  Lanes on padding slots (outside the image) duplicate a valid lane's ray
// and store no result.  Precondition: every slot is owned (no partition, or the compacted one).
template <bool kCount, int kRays, bool kList, bool kMini>
__global__ void __launch_bounds__(kPacketBlock, kRays == 1 ? 8 : (kRays == 2 ? 6 : 4)) primary_packet_kernel(const PacketArgs a) {
  static_assert(!kMini || kRays == 1, "multi-pixel packets carry one ray per lane");
  __shared__ uint32_t s_blk, s_cnt;
  __shared__ __align__(16) unsigned char s_warp[kPacketWarps * kWarpSmem];
  __shared__ uint8_t s_lut[8 * 256];
  for (int i = threadIdx.x; i < 8 * 256; i += kPacketBlock) s_lut[i] = (uint8_t)xor_permute8((uint32_t)i & 0xFFu, (uint32_t)i >> 8);
  __syncthreads();
  const PassParams& p = a.p;
  const unsigned lane = threadIdx.x & 31u;
  const uint32_t n_warps = gridDim.x * kPacketWarps;
  const uint32_t n_tasks = kMini ? (p.n_slots << p.s_shift) >> 5 : p.n_slots * (p.samples / (32u * kRays));
  const int ax_k = a.axis_k, ax_a = (ax_k + 1) % 3, ax_b = (ax_k + 2) % 3;
  const float sigma = a.sigma;
  const f3 O = p.cam.origin;
  const float O_k = sel3(O, ax_k);
  // this warp's staging area: one node (20 words) + the candidate triangles of that node
  uint32_t sm_node = (uint32_t)__cvta_generic_to_shared(s_warp) + (threadIdx.x >> 5) * kWarpSmem;
  const uint32_t sm_lut = (uint32_t)__cvta_generic_to_shared(s_lut);
  // lane roles in the node test: child slot c, plane pl (0: u >= umin, 1: u <= umax, 2: v >= vmin, 3: v <= vmax)
  const uint32_t c = lane & 7u, pl = lane >> 3;
  const int ax_sel = pl < 2u ? ax_a : ax_b;
  const float ca = (pl & 1u) ? -1.0f : 1.0f;
  const float O_sel = sel3(O, ax_sel);
  // Staged node layout: words 0..19 = the node, words 20..22 = the three grid scales 2^(e-127) as floats (written by
  // lanes 20..22 from their own copy of word 3), so that p and scale of an axis share one address register.
  // The addresses go through opaque() once: ptxas otherwise re-derives them from %tid in every node step.
  const uint32_t a_ax_sel = opaque(sm_node + 4u * (uint32_t)ax_sel);  // p at +0, scale at +80
  const uint32_t a_ax_k = opaque(sm_node + 4u * (uint32_t)ax_k);
  const uint32_t a_q_sel = opaque(sm_node + ((pl & 1u) ? 32u : 56u) + 8u * (uint32_t)ax_sel + c);  // +1: far corner = hi; -1: lo
  const uint32_t a_q_k = opaque(sm_node + 32u + 8u * (uint32_t)ax_k + c);  // lo; hi is 24 bytes on
  const uint32_t a_stage = opaque(sm_node + 4u * lane);
  // staged value = prmt(word, 0, stage_sel) >> stage_shr: the word itself (lanes 0..19) or byte j of word 3 moved to
  // the exponent field (lanes 20 + j)
  const bool scale_lane = lane >= 20u && lane < 23u;
  const uint32_t stage_sel = opaque(scale_lane ? ((lane - 20u) << 12) | 0x444u : 0x3210u);
  const uint32_t stage_shr = opaque(scale_lane ? 1u : 0u);
  const uint32_t* node_words = opaque_ptr(reinterpret_cast<const uint32_t*>(p.scene.nodes) + (lane < 20u ? lane : 3u));
  sm_node = opaque(sm_node);

  uint32_t done_rays = 0, n_nodes = 0, n_tris = 0, n_bail = 0, sky_rays = 0;

  auto process = [&](const uint32_t task) {
    uint32_t chunk = 0, slot, x, y, id0, s_mini = 0;
    bool valid = true;  // kMini: this lane's slot is a pixel of the image
    if constexpr (kMini) {
      id0 = task * 32u + lane;  // slot-major path id (kernels.h slot_sample_id())
      slot = id0 >> p.s_shift;
      s_mini = id0 & ((1u << p.s_shift) - 1u);
      valid = slot_pixel(p, slot, x, y);
      const unsigned vm = __ballot_sync(kFull, valid);
      if (vm == 0u) return;
      const int src = __ffs(vm) - 1;  // padding lanes trace a copy of a valid lane's ray
      const uint32_t sx = __shfl_sync(kFull, x, src), sy = __shfl_sync(kFull, y, src), ss = __shfl_sync(kFull, s_mini, src);
      if (!valid) x = sx, y = sy, s_mini = ss;
    } else {
      chunk = fd_div(p.fd_slots, task), slot = task - chunk * p.n_slots;  // slots go tile by tile (kernels.h)
      if (!slot_pixel(p, slot, x, y) || (!p.dense && !owns_pixel(p.im, x, y))) return;  // warp-uniform
      id0 = ((chunk * kRays) * p.n_slots + slot) * 32u + lane;  // chunked path id (kernels.h path_id())
    }
    const uint32_t pixel = y * p.im.width + x;
    // ---- the 32 * kRays rays (K1); ray j of a lane is sample 32 * (chunk * kRays + j) + lane ----
    const uint32_t id_step = p.n_slots * 32u;                                // to the same lane of the next chunk
    f3 dir[kRays];
    float w[kRays];  // depth per unit t
    uint32_t ku_min = 0xFFFFFFFFu, ku_max = 0u, kv_min = 0xFFFFFFFFu, kv_max = 0u;
    bool bail = false;
#pragma unroll
    for (int j = 0; j < kRays; ++j) {
      const uint32_t s = kMini ? s_mini : (chunk * kRays + (uint32_t)j) * 32u + lane;
      const uint32_t id = id0 + (uint32_t)j * id_step;
      const uint32_t frame = p.first_frame + s * p.frame_stride;
      float jx = 0.0f, jy = 0.0f;
      if (p.ip.jitter) {
        const rand4 jit = philox_block(pixel, frame, 0u, 2u, p.ip.seed);
        jx = jit.u0, jy = jit.u1;
      }
      dir[j] = primary_dir(p.cam, p.im, x, y, jx, jy);
      // the origin {camera, tnear} of a camera ray is implied and queue 0 is the identity when every pixel is owned
      // (shade_kernel, wavefront.cu): neither is stored; the direction is stored at the end, if anything still needs it
      if (!kMini && !kList && !all_slots_owned(p)) {
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(p.st.counters + slot_queue_count(0), 32u);
        base = __shfl_sync(kFull, base, 0);
        p.st.queue[0][base + lane] = id;
      }
      w[j] = sigma * sel3(dir[j], ax_k);
      bail = bail || !(w[j] > 0.02f);
      const float inv_w = 1.0f / w[j];
      const uint32_t ku = fkey(sel3(dir[j], ax_a) * inv_w), kv = fkey(sel3(dir[j], ax_b) * inv_w);
      ku_min = min(ku_min, ku), ku_max = max(ku_max, ku), kv_min = min(kv_min, kv), kv_max = max(kv_max, kv);
    }
    bail = __any_sync(kFull, bail);
    // ---- bounding planes of the directions: lanes of planes 0/1 need a u limit, lanes of planes 2/3 a v limit ----
    const uint32_t kumin = __reduce_min_sync(kFull, ku_min), kumax = __reduce_max_sync(kFull, ku_max);
    const uint32_t kvmin = __reduce_min_sync(kFull, kv_min), kvmax = __reduce_max_sync(kFull, kv_max);
    const float e0 = funkey(pl < 2u ? kumin : kvmin), e1 = funkey(pl < 2u ? kumax : kvmax);
    const float dl = 1e-5f * (1.0f + fmaxf(fabsf(e0), fabsf(e1)));
    // plane of this lane: ca * D_sel + ck * D_k >= 0, D = far corner of the child box along (ca, ck) minus the origin
    const float lim = (pl & 1u) ? e1 + dl : e0 - dl;
    const float ck = (ca > 0.0f ? -lim : lim) * sigma;
    const uint32_t a_q_kp = opaque(a_q_k + (ck > 0.0f ? 24u : 0u));  // the k bound this lane's plane uses
    const uint32_t octinv = 7u ^ ((dir[0].x < 0.f ? 1u : 0u) | (dir[0].y < 0.f ? 2u : 0u) | (dir[0].z < 0.f ? 4u : 0u));
    const uint32_t oct = __shfl_sync(kFull, octinv, 0);
    const uint32_t a_lut = opaque(sm_lut + oct * 256u);

    TraceHit hit[kRays];
#pragma unroll
    for (int j = 0; j < kRays; ++j) {
      hit[j].t = __int_as_float(0x7f800000);
      hit[j].tri = 0xFFFFFFFFu;
      hit[j].prim = 0xFFFFFFFFu;
    }
    float kmax = __int_as_float(0x7f800000);
    uint2 ngroup = make_uint2(0u, 0x80000000u), sreg = make_uint2(0u, 0u);
    int sp = 0, cost = 0;

    while (!bail) {
      // ---- node step (warp-uniform control) ----
      const int bit = pt_bfind(ngroup.y);
      ngroup.y &= ~(1u << bit);
      if (ngroup.y > 0x00FFFFFFu) {
        if (sp == 32) {
          bail = true;
          break;
        }
        if ((int)lane == sp) sreg = ngroup;
        ++sp;
      }
      const uint32_t slot = ((uint32_t)bit - 24u) ^ oct;
      const uint32_t rel = (uint32_t)__popc((ngroup.y & 0xFFu) & ~(0xFFFFFFFFu << slot));
      const uint32_t node = ngroup.x + rel;
      // stage the 80-byte node through shared memory: ONE coalesced load, then every lane picks its bytes
      const uint32_t word = __ldg(node_words + 20u * (size_t)node);
      __syncwarp();
      if (lane < 23u) sts_u32(a_stage, __byte_perm(word, 0u, stage_sel) >> stage_shr);
      __syncwarp();
      const uint4 w1 = lds_u128(sm_node + 16u);
      const uint32_t node_imask = lds_u8(sm_node + 15u);
      const float p_sel = __uint_as_float(lds_u32(a_ax_sel)), s_sel = __uint_as_float(lds_u32(a_ax_sel + 80u));
      const float p_k = __uint_as_float(lds_u32(a_ax_k)), s_k = __uint_as_float(lds_u32(a_ax_k + 80u));
      const uint32_t q_sel = lds_u8(a_q_sel), q_klo = lds_u8(a_q_k), q_khi = lds_u8(a_q_k + 24u), q_kp = lds_u8(a_q_kp);
      const float D_sel = fmaf((float)q_sel, s_sel, p_sel - O_sel);
      const float bk = p_k - O_k;
      const float D_lo = fmaf((float)q_klo, s_k, bk), D_hi = fmaf((float)q_khi, s_k, bk), D_kp = fmaf((float)q_kp, s_k, bk);
      const float val = fmaf(ck, D_kp, ca * D_sel);
      const float n1 = sigma * D_lo, n2 = sigma * D_hi;
      const bool ok = val >= 0.0f && fminf(n1, n2) <= kmax && fmaxf(n1, n2) >= 0.0f;
      const unsigned bal = __ballot_sync(kFull, ok);
      const uint32_t hits = bal & (bal >> 8) & (bal >> 16) & (bal >> 24) & 0xFFu;
      const uint32_t prio = lds_u8(a_lut + (hits & node_imask));
      ngroup = make_uint2(w1.x, (prio << 24) | node_imask);
      const uint32_t tri_base = w1.y, tri_valid = w1.z;
      uint32_t tmask = (hits * 0x010101u) & tri_valid;
      cost += 3;
      if (kCount) ++n_nodes;
      // ---- every candidate triangle against all 32 rays ----
      if (tmask) {
        // lane b fetches candidate b (if any) into slot popc(lower candidates): all fetches of the node in flight together
        if ((tmask >> lane) & 1u) {
          const uint32_t below = (1u << lane) - 1u;
          const float4* tp = p.scene.tris + 3 * (size_t)(tri_base + (uint32_t)__popc(tri_valid & below));
          const float4 ta = __ldg(tp), tb = __ldg(tp + 1), tc = __ldg(tp + 2);
          const uint32_t dst = sm_node + kNodeBytes + 48u * (uint32_t)__popc(tmask & below);
          sts_u128(dst, *reinterpret_cast<const uint4*>(&ta));
          sts_u128(dst + 16u, *reinterpret_cast<const uint4*>(&tb));
          sts_u128(dst + 32u, *reinterpret_cast<const uint4*>(&tc));
        }
        __syncwarp();
        uint32_t src = sm_node + kNodeBytes;
        do {
          const uint32_t low = tmask & (0u - tmask);
          tmask ^= low;
          const uint32_t tri = tri_base + (uint32_t)__popc(tri_valid & (low - 1u));
          const uint4 ua = lds_u128(src), ub = lds_u128(src + 16u), uc = lds_u128(src + 32u);
          src += 48u;
          const f3 tv0 = mk3(__uint_as_float(ua.x), __uint_as_float(ua.y), __uint_as_float(ua.z));
          const f3 te1 = mk3(__uint_as_float(ub.x), __uint_as_float(ub.y), __uint_as_float(ub.z));
          const f3 te2 = mk3(__uint_as_float(uc.x), __uint_as_float(uc.y), __uint_as_float(uc.z));
#pragma unroll
          for (int j = 0; j < kRays; ++j) {
            float t;
            if (tri_test(tv0, te1, te2, O, dir[j], t) && t >= p.ip.tnear &&
                (t < hit[j].t || (t == hit[j].t && ua.w < hit[j].prim))) {  // tri_step<false>'s closest-hit rule
              hit[j].t = t;
              hit[j].tri = tri;
              hit[j].prim = ua.w;
            }
          }
          cost += 4;
          if (kCount) ++n_tris;
        } while (tmask);
        float far = hit[0].t * w[0];
#pragma unroll
        for (int j = 1; j < kRays; ++j) far = fmaxf(far, hit[j].t * w[j]);
        kmax = __uint_as_float(__reduce_max_sync(kFull, __float_as_uint(far))) * 1.00001f;
      }
      if (ngroup.y <= 0x00FFFFFFu) {
        if (sp == 0) break;
        --sp;
        ngroup.x = __shfl_sync(kFull, sreg.x, sp);
        ngroup.y = __shfl_sync(kFull, sreg.y, sp);
      }
      if (cost > a.budget * kRays) bail = true;
    }
    bool any_hit = true;
    if (kList && !bail) {
      bool mine = false;
#pragma unroll
      for (int j = 0; j < kRays; ++j) mine = mine || hit[j].tri != 0xFFFFFFFFu;
      any_hit = __any_sync(kFull, mine);
    }
    if (!any_hit) {
      // every ray of the packet leaves the scene: L = 0 + 1 * (sky * sky_scale), the shade kernel's miss branch for a
      // camera ray (CPT.cpp:231-235) written out, so the bits are the same
#pragma unroll
      for (int j = 0; j < kRays; ++j) {
        const f3 L = mk3(0.f, 0.f, 0.f) + mk3(1.0f, 1.0f, 1.0f) * (sky_radiance(p.ip, dir[j]) * p.ip.sky_scale);
        if (!kMini || valid) p.st.L[id0 + (uint32_t)j * id_step] = make_float4(L.x, L.y, L.z, 0.f);
      }
      // sample 0 of the pixel: ray 0 of lane 0 of chunk 0, or (kMini) the lane whose sample index is 0
      if (p.aov_prim != nullptr && (kMini ? (valid && s_mini == 0u) : (chunk == 0u && lane == 0u))) {
        p.aov_prim[pixel] = 0xFFFFFFFFu;
        p.aov_t[pixel] = __int_as_float(0x7f800000);
      }
      if (!kMini || valid) {
        sky_rays += (uint32_t)kRays;
        done_rays += (uint32_t)kRays;
      }
      return;
    }
#pragma unroll
    for (int j = 0; j < kRays; ++j)
      p.st.ray_d[id0 + (uint32_t)j * id_step] = make_float4(dir[j].x, dir[j].y, dir[j].z, __int_as_float(0x7f800000));
    if (kList) {  // one descriptor per chunk, in the tile's region
      uint32_t pos = 0;
      if (lane == 0) pos = atomicAdd(&s_cnt, (uint32_t)kRays);
      pos = __shfl_sync(kFull, pos, 0);
      if (lane < (uint32_t)kRays) a.desc[(size_t)s_blk * (64u * kRays) + pos + lane] = kMini ? task : (id0 + lane * id_step) >> 5;
    }
    if (!bail) {
#pragma unroll
      for (int j = 0; j < kRays; ++j)
        p.st.hit[id0 + (uint32_t)j * id_step] =
            make_uint2(__float_as_uint(hit[j].tri == 0xFFFFFFFFu ? __int_as_float(0x7f800000) : hit[j].t), hit[j].tri);
      if (!kMini || valid) done_rays += (uint32_t)kRays;
    } else {
      if constexpr (kMini) {  // only the lanes that are pixels go to the per-ray kernel, so that its ray count stays exact
        const unsigned vm = __ballot_sync(kFull, valid);
        const uint32_t rank = (uint32_t)__popc(vm & ((1u << lane) - 1u));
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(a.fb_count, (uint32_t)__popc(vm));
        base = __shfl_sync(kFull, base, 0);
        if (valid) {
          a.fb_queue[base + rank] = id0;
          p.st.ray_o[id0] = make_float4(O.x, O.y, O.z, p.ip.tnear);
        }
      } else {
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(a.fb_count, 32u * kRays);
        base = __shfl_sync(kFull, base, 0);
#pragma unroll
        for (int j = 0; j < kRays; ++j) {
          const uint32_t id = id0 + (uint32_t)j * id_step;
          a.fb_queue[base + 32u * (uint32_t)j + lane] = id;
          p.st.ray_o[id] = make_float4(O.x, O.y, O.z, p.ip.tnear);  // the per-ray kernel reads its rays from memory
        }
      }
      if (kCount) ++n_bail;
    }
  };

  if constexpr (kList) {
    const uint32_t n_blocks = kMini ? (n_tasks + 63u) >> 6 : n_tasks >> 6;  // n_slots is a multiple of 64: whole tiles unless kMini
    for (;;) {
      __syncthreads();  // the previous tile's count has been written
      if (threadIdx.x == 0) {
        s_blk = atomicAdd(a.block_counter, 1u);
        s_cnt = 0u;
      }
      __syncthreads();
      const uint32_t blk = s_blk;
      if (blk >= n_blocks) break;
      for (uint32_t t = threadIdx.x >> 5; t < 64u; t += kPacketWarps)
        if (!kMini || (blk << 6) + t < n_tasks) process((blk << 6) + t);
      __syncthreads();
      if (threadIdx.x == 0) a.block_hits[blk] = s_cnt;
    }
  } else {
    for (uint32_t task = blockIdx.x * kPacketWarps + (threadIdx.x >> 5); task < n_tasks; task += n_warps) process(task);
  }
  if (!kList && all_slots_owned(p) && blockIdx.x == 0 && threadIdx.x == 0) p.st.counters[slot_queue_count(0)] = p.n_slots * p.samples;  // (also the padding ids: the shade kernel skips them)
  const uint64_t wr = warp_sum_u32(done_rays);
  if (lane == 0 && wr) {
    atomicAdd(&p.counters[kCntRays], (unsigned long long)wr);
    atomicAdd(&p.counters[kCntPacketRays], (unsigned long long)wr);
    atomicAdd(&p.counters[kCntPackets], (unsigned long long)(wr / (32u * kRays)));
  }
  if (kList) {
    const uint64_t ws = warp_sum_u32(sky_rays);
    if (lane == 0 && ws) atomicAdd(&p.counters[kCntPaths], (unsigned long long)ws);
  }
  if (kCount && lane == 0) {
    atomicAdd(&p.counters[kCntPacketNodes], (unsigned long long)n_nodes);
    atomicAdd(&p.counters[kCntPacketTris], (unsigned long long)n_tris);
    atomicAdd(&p.counters[kCntPacketBail], (unsigned long long)n_bail);
  }
}

// Hit list, step 2: exclusive scan of the descriptors per tile (one CTA; a pass has at most a few 10^5 tiles) and the
// size of the shade kernel's bounce-0 queue.
__global__ void __launch_bounds__(1024) hitlist_scan_kernel(const uint32_t* __restrict__ block_hits, uint32_t n_blocks,
                                                            uint32_t* __restrict__ offsets, uint32_t* __restrict__ queue_count) {
  __shared__ uint32_t s_sum[1024];
  const uint32_t per = (n_blocks + 1023u) / 1024u, begin = threadIdx.x * per, end = min(begin + per, n_blocks);
  uint32_t sum = 0;
  for (uint32_t i = begin; i < end; ++i) sum += block_hits[i];
  s_sum[threadIdx.x] = sum;
  __syncthreads();
  for (uint32_t d = 1; d < 1024u; d <<= 1) {  // Hillis-Steele inclusive scan
    const uint32_t v = threadIdx.x >= d ? s_sum[threadIdx.x - d] : 0u;
    __syncthreads();
    s_sum[threadIdx.x] += v;
    __syncthreads();
  }
  uint32_t run = s_sum[threadIdx.x] - sum;
  for (uint32_t i = begin; i < end; ++i) {
    offsets[i] = run;
    run += block_hits[i];
  }
  if (threadIdx.x == 1023u) *queue_count = s_sum[1023] * 32u;
}

// Hit list, step 3: the descriptors of every tile, in tile order, expanded to the 32 path ids of their chunk.
__global__ void __launch_bounds__(128) hitlist_expand_kernel(const uint32_t* __restrict__ desc, const uint32_t* __restrict__ block_hits,
                                                             const uint32_t* __restrict__ offsets, uint32_t n_blocks, uint32_t region,
                                                             uint32_t* __restrict__ queue) {
  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  for (uint32_t b = blockIdx.x; b < n_blocks; b += gridDim.x) {
    const uint32_t n = block_hits[b], base = offsets[b];
    for (uint32_t k = warp; k < n; k += 4u) queue[(size_t)(base + k) * 32u + lane] = desc[(size_t)b * region + k] * 32u + lane;
  }
}

}  // namespace

bool primary_packet_shape(uint32_t samples) {
  return (samples >= 32u && samples % 32u == 0u) || samples == 1u || samples == 2u || samples == 4u || samples == 8u || samples == 16u;
}

// The bounce-0 hit list applies when every slot of the pass is owned: no partition, or the compacted slots of a tile
// partition (kernels.h); a sparse partition keeps an explicit queue built with one atomic per packet.
bool primary_packet_hitlist(const PassParams& p) {
  const char* e = getenv("CUDAPT_PACKET_HITLIST");  // read per pass: tests toggle it
  if ((e ? atoi(e) : 1) == 0 || !all_slots_owned(p) || !primary_packets_enabled(p)) return false;
  if (p.samples % 32u != 0u) {
    // multi-pixel packets: descriptors (one per packet, in blocks of 64) + per-block counts and offsets must fit the shadow
    // queue of the pass (one word per path) -- they do not for images of a tile or two
    const uint64_t n_tasks = ((uint64_t)p.n_slots * p.samples) >> 5, n_blocks = (n_tasks + 63u) >> 6;
    if (n_blocks * 64u + 2u * n_blocks > (uint64_t)p.n_slots * p.samples) return false;
  }
  return true;
}

// Packets apply to a pass of 32 k samples (one warp = 32 samples of one pixel, k = 1, 2 or 4 rays per lane) and to a pass
// of 1, 2, 4, 8 or 16 samples (one warp = 32 / S neighbouring pixels x S samples) when every slot of the pass is owned;
// CUDAPT_PRIMARY_PACKETS=0 switches them off; CUDAPT_PACKET_MINI: see below.
bool primary_packets_enabled(const PassParams& p) {
  const char* e = getenv("CUDAPT_PRIMARY_PACKETS");  // read per pass: tests toggle it
  const int on = e ? atoi(e) : 1;
  // a thin lens gives every ray its own origin: no common apex, no packets
  if (on == 0 || p.cam.lens_radius > 0.0f || !primary_packet_shape(p.samples)) return false;
  if (p.samples % 32u == 0u) return true;
  // measured on C3 at 1920x1080 (profiles/r02m_mini_probe_c3.log): 16 samples per pass -24 %, 8 -15 %, 4 neutral, 1 +54 %
  // (an 8x4-pixel frustum tests 11 triangles per packet) -> on from 8 samples; CUDAPT_PACKET_MINI=1 forces 1..16, =0 off
  const char* m = getenv("CUDAPT_PACKET_MINI");
  const bool wanted = m ? atoi(m) != 0 : p.samples >= 8u;
  return wanted && all_slots_owned(p);
}

cudaError_t launch_primary_packets(const PassParams& p, uint32_t* fb_queue, uint32_t* fb_count, bool counting, int sm_count,
                                   cudaStream_t s) {
  const char* be = getenv("CUDAPT_PACKET_BUDGET");
  const int budget = be ? (atoi(be) < 0 ? 0 : atoi(be)) : 400;
  PacketArgs a;
  a.p = p;
  a.fb_queue = fb_queue;
  a.fb_count = fb_count;
  const float f[3] = {p.cam.forward.x, p.cam.forward.y, p.cam.forward.z};
  int k = 0;
  for (int i = 1; i < 3; ++i)
    if (fabsf(f[i]) > fabsf(f[k])) k = i;
  a.axis_k = k;
  a.sigma = f[k] < 0.0f ? -1.0f : 1.0f;
  a.budget = budget;
  const char* re = getenv("CUDAPT_PACKET_RAYS");  // rays per lane: 1, 2 or 4 (default: the largest that divides the pass)
  const bool mini = p.chunked == 2u;
  const uint32_t chunks = p.samples / 32u;
  int rays = re ? atoi(re) : 4;
  rays = mini ? 1 : (rays >= 4 && chunks % 4u == 0u ? 4 : (rays >= 2 && chunks % 2u == 0u ? 2 : 1));
  const uint32_t n_tasks = mini ? (uint32_t)(((uint64_t)p.n_slots * p.samples) >> 5) : p.n_slots * (chunks / (uint32_t)rays);
  const uint32_t warps_per_cta = kPacketWarps;
  uint32_t grid = (uint32_t)sm_count * (rays == 1 ? 8u : (rays == 2 ? 6u : 4u));
  const uint32_t need = (n_tasks + warps_per_cta - 1) / warps_per_cta;
  if (need < grid) grid = need ? need : 1u;
  // hit list: descriptors, per-tile counts and offsets live in the (still unused) shadow queue of the pass
  const bool list = p.explicit_q0 != 0u;
  const uint32_t n_blocks = (n_tasks + 63u) >> 6, n_desc = mini ? n_blocks * 64u : p.n_slots * chunks;
  a.block_counter = fb_count + 2;  // free word of the fall-back counter block, zeroed with the pass's counters
  a.desc = p.st.shadow_queue;
  a.block_hits = p.st.shadow_queue + n_desc;
  uint32_t* offsets = a.block_hits + n_blocks;
#define PK_LAUNCH3(R, L, M)                                                    \
  if (counting)                                                                \
    primary_packet_kernel<true, R, L, M><<<grid, kPacketBlock, 0, s>>>(a);     \
  else                                                                         \
    primary_packet_kernel<false, R, L, M><<<grid, kPacketBlock, 0, s>>>(a)
#define PK_LAUNCH(R)             \
  if (list) {                    \
    PK_LAUNCH3(R, true, false);  \
  } else {                       \
    PK_LAUNCH3(R, false, false); \
  }
  if (mini) {
    if (list) {
      PK_LAUNCH3(1, true, true);
    } else {
      PK_LAUNCH3(1, false, true);
    }
  } else if (rays == 4) {
    PK_LAUNCH(4);
  } else if (rays == 2) {
    PK_LAUNCH(2);
  } else {
    PK_LAUNCH(1);
  }
#undef PK_LAUNCH
#undef PK_LAUNCH3
  if (list) {
    hitlist_scan_kernel<<<1, 1024, 0, s>>>(a.block_hits, n_blocks, offsets, p.st.counters + slot_queue_count(0));
    hitlist_expand_kernel<<<(uint32_t)sm_count * 8u, 128, 0, s>>>(a.desc, a.block_hits, offsets, n_blocks, 64u * (uint32_t)rays, p.st.queue[0]);
  }
  return cudaGetLastError();
}

}  // namespace cudapt
