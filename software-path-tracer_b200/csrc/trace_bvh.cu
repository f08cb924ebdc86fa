// K2b / K2c: persistent-thread traversal of the 8-wide quantised BVH (closest hit and any-hit).
// Replaces rtcIntersect1 (CPT.cpp:227) for triangle scenes.
//
// Launch shape: grid = SMs x kTraceBlocksPerSM CTAs of 128 threads, resident for the whole launch.  Each
// warp pulls rays from the bounce's queue through a global atomic counter (one atomic per refill, issued by
// lane 0 and broadcast with shfl); lanes whose ray finished are refilled as soon as the number of lanes
// still traversing together drops below kRefillThreshold (ballot/popc), so a warp keeps > 50 % of its
// lanes busy until the queue drains.  Traversal stack: 8 entries per thread in shared memory (interleaved
// by thread id: conflict-free 64-bit accesses), the rest in local memory.  Node and triangle fetches are
// 128-bit read-only loads (5 per node, 3 per triangle).
#include <cstdlib>

#include "bvh8_traverse.cuh"
#include "kernels.h"

namespace cudapt {
namespace {

constexpr int kSharedStack = 8;
constexpr int kLocalStack = 32;
constexpr int kRefillThreshold = 26;
constexpr int kPoolRays = 64;

__device__ __forceinline__ void sts64(uint32_t addr, uint2 v) {
  asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(addr), "r"(v.x), "r"(v.y) : "memory");
}
__device__ __forceinline__ uint2 lds64(uint32_t addr) {
  uint2 v;
  asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void sts32(uint32_t addr, uint32_t v) {
  asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t lds32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
  return v;
}

// s_lut[k * 256 + x] = xor_permute8(x, k): child-hit mask -> traversal priority order for octant key k.
__device__ __forceinline__ void fill_perm_lut(uint8_t* lut) {
  for (int i = threadIdx.x; i < 8 * 256; i += kTraceBlock) lut[i] = (uint8_t)xor_permute8((uint32_t)i & 0xFFu, (uint32_t)i >> 8);
  __syncthreads();
}

struct RegStack {
  uint32_t sbase;  // shared-space byte address of this thread's entry 0; entries kTraceBlock * 8 bytes apart
  uint2* local;
  int& sp;
  __device__ __forceinline__ void push(uint2 v) {
    if (sp < kSharedStack)
      sts64(sbase + (uint32_t)sp * (kTraceBlock * 8u), v);
    else if (sp < kSharedStack + kLocalStack)
      local[sp - kSharedStack] = v;
    ++sp;
  }
  __device__ __forceinline__ uint2 pop() {
    --sp;
    return sp < kSharedStack ? lds64(sbase + (uint32_t)sp * (kTraceBlock * 8u)) : local[sp - kSharedStack];
  }
  __device__ __forceinline__ bool empty() const { return sp == 0; }
};

// One resumable traversal state per lane.  The generic bvh8_trace() in bvh8_traverse.cuh runs a ray to
// completion; here the same steps are driven by an outer loop that can interleave refills.
template <bool kAnyHit, bool kCount, int kBlocksPerSM>
__global__ void __launch_bounds__(kTraceBlock, kBlocksPerSM)
    trace_kernel(const uint4* __restrict__ nodes, const float4* __restrict__ tris, const float4* __restrict__ ray_o,
                 const float4* __restrict__ ray_d, const uint32_t* __restrict__ queue, const uint32_t* __restrict__ count_ptr,
                 uint32_t* __restrict__ work_counter, uint2* __restrict__ hit_out, float4* __restrict__ L,
                 const float4* __restrict__ contrib, unsigned long long* __restrict__ stats, const int refill_threshold, const uint32_t magic) {
  __shared__ uint2 s_stack[kSharedStack * kTraceBlock];
  __shared__ uint8_t s_lut[8 * 256];
  fill_perm_lut(s_lut);
  const uint32_t lut_base = (uint32_t)__cvta_generic_to_shared(s_lut);
  const uint32_t n = *count_ptr;
  const unsigned lane = threadIdx.x & 31u;
  uint2 lstack[kLocalStack];
  int sp = 0;
  RegStack stack{(uint32_t)__cvta_generic_to_shared(s_stack) + threadIdx.x * 8u, lstack, sp};

  bool active = false, drained = false;
  uint32_t pool_next = 0, pool_left = 0;  // warp-uniform
  uint32_t path = 0;
  RayPre r;
  float tnear = 0.f, tmax = 0.f, tmin_box = 0.f;
  TraceHit hit;
  uint2 ngroup = make_uint2(0u, 0u);
  uint32_t n_nodes = 0, n_tris = 0;

  for (;;) {
    // ---- refill idle lanes (warp-synchronous) ----
    const unsigned idle = __ballot_sync(0xffffffffu, !active);
    if (idle && !drained) {
      // The warp owns a pool of kPoolRays consecutive queue entries, claimed with ONE atomic; idle lanes are
      // served from the pool, so the number of global atomics is n / kPoolRays however often lanes refill.
      const uint32_t want = (uint32_t)__popc(idle);
      if (pool_left < want) {  // warp-uniform
        // the few entries left in the old pool are served first, the new pool next time round
        if (pool_left == 0u) {
          uint32_t base = 0;
          if (lane == 0) base = atomicAdd(work_counter, (uint32_t)kPoolRays);
          pool_next = __shfl_sync(0xffffffffu, base, 0);
          pool_left = pool_next < n ? min((uint32_t)kPoolRays, n - pool_next) : 0u;
          if (pool_left == 0u) drained = true;
        }
      }
      if (!active) {
        const uint32_t rank = (uint32_t)__popc(idle & ((1u << lane) - 1u));
        const uint32_t i = pool_next + rank;
        if (rank < pool_left) {
          path = queue[i];
          const float4 o4 = ray_o[path], d4 = ray_d[path];
          r = prepare_ray(mk3(o4.x, o4.y, o4.z), mk3(d4.x, d4.y, d4.z), magic);
          r.perm_lut = lut_base + r.octinv * 256u;
          if (kAnyHit) {
            tnear = d4.w;  // shadow rays: {o, tmax}, {d, tnear}
            tmax = o4.w;
          } else {
            tnear = o4.w;
            tmax = d4.w;
          }
          tmin_box = box_tmin(tnear);
          hit.t = tmax;
          hit.tri = 0xFFFFFFFFu;
          hit.prim = 0xFFFFFFFFu;
          ngroup = make_uint2(0u, 0x80000000u);
          sp = 0;
          active = true;
        }
      }
      const uint32_t served = min(want, pool_left);
      pool_next += served;
      pool_left -= served;
    }
    if (__ballot_sync(0xffffffffu, active) == 0u) break;

    // ---- warp-synchronous traversal: every iteration each active lane does ONE node step (fetch + 8 box
    // tests), its triangle tests, and a pop if the node had no internal hits; the ballot at the bottom
    // reconverges the warp, so node steps always execute with all active lanes together ----
    for (;;) {
      if (active) {
        uint32_t tri_base, tri_valid, tmask;
        node_step(nodes, r, tmin_box, kAnyHit ? tmax : hit.t, ngroup, stack, tri_base, tri_valid, tmask);
        if (kCount) ++n_nodes;
        bool finished = false;
        while (tmask) {
          // lowest candidate first (the order of the tests does not matter): isolating the lowest set bit is
          // two alu instructions, against FLO (xu pipe) + shift + mask for the highest
          const uint32_t low = tmask & (0u - tmask);
          tmask ^= low;
          if (kCount) ++n_tris;
          const uint32_t tri = tri_base + (uint32_t)__popc(tri_valid & (low - 1u));
          if (tri_step<kAnyHit>(tris, tri, r, tnear, tmax, hit)) {
            finished = true;
            break;
          }
        }
        if (!finished && ngroup.y <= 0x00FFFFFFu) {
          if (stack.empty())
            finished = true;
          else
            ngroup = stack.pop();
        }
        if (finished) {
          if (kAnyHit) {
            if (hit.tri == 0xFFFFFFFFu) {  // unoccluded: add the light sample (path-private slot, no atomics)
              float4 l = L[path];
              const float4 c = contrib[path];
              l.x += c.x;
              l.y += c.y;
              l.z += c.z;
              L[path] = l;
            }
          } else {
            hit_out[path] = make_uint2(__float_as_uint(hit.tri == 0xFFFFFFFFu ? __int_as_float(0x7f800000) : hit.t), hit.tri);
          }
          active = false;
        }
      }
      const unsigned act = __ballot_sync(0xffffffffu, active);
      if (act == 0u) break;
      if (!drained && __popc(act) < refill_threshold) break;
    }
  }
  if (kCount) {
    const uint64_t wn = warp_sum_u32(n_nodes), wt = warp_sum_u32(n_tris);
    if (lane == 0) {
      atomicAdd(&stats[kCntNodes], (unsigned long long)wn);
      atomicAdd(&stats[kCntTris], (unsigned long long)wt);
    }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    atomicAdd(&stats[kCntRays], (unsigned long long)n);
    if (kAnyHit) atomicAdd(&stats[kCntShadowRays], (unsigned long long)n);
  }
}


int env_int(const char* name, int dflt, int lo, int hi) {
  const char* e = getenv(name);
  const int x = e ? atoi(e) : dflt;
  return x < lo ? lo : (x > hi ? hi : x);
}

// Tuning knob (profiles/r01_notes.md): lanes are refilled when fewer than this many of a warp's 32 are busy.
int refill_threshold() {
  static const int v = [] {
    const char* e = getenv("CUDAPT_REFILL_THRESHOLD");
    const int x = e ? atoi(e) : kRefillThreshold;
    return x < 1 ? 1 : (x > 32 ? 32 : x);
  }();
  return v;
}

// Register budget knob: CTAs per SM the kernel is compiled for (8 -> 64 registers, 6 -> 80, 5 -> 96).
int trace_blocks_per_sm() {
  static const int v = [] {
    const char* e = getenv("CUDAPT_TRACE_BLOCKS");
    const int x = e ? atoi(e) : kTraceBlocksPerSM;
    return (x == 6 || x == 8 || x == 9) ? x : kTraceBlocksPerSM;
  }();
  return v;
}

template <bool kAnyHit>
cudaError_t launch_trace(const PassParams& p, const float4* ro, const float4* rd, const uint32_t* queue, uint32_t* count,
                         uint32_t* work, uint2* hit, float4* L, const float4* contrib, bool counting, int sm_count,
                         cudaStream_t s) {
  const int bps = trace_blocks_per_sm();
  const dim3 grid(sm_count * bps), block(kTraceBlock);
  const int thr = refill_threshold();
#define PT_LAUNCH(B, C)                                                                                                 \
  trace_kernel<kAnyHit, C, B><<<grid, block, 0, s>>>(p.scene.nodes, p.scene.tris, ro, rd, queue, count, work, hit, L, contrib, \
                                                     p.counters, thr, 0x4B000000u)
#define PT_PICK(B)          \
  if (counting)             \
    PT_LAUNCH(B, true);     \
  else                      \
    PT_LAUNCH(B, false)
  switch (bps) {
    case 6: PT_PICK(6); break;
    case 9: PT_PICK(9); break;
    default: PT_PICK(8); break;
  }
#undef PT_PICK
#undef PT_LAUNCH
  return cudaGetLastError();
}

}  // namespace

cudaError_t launch_trace_closest(const PassParams& p, uint32_t bounce, int queue_in, bool counting, int sm_count,
                                 cudaStream_t s) {
  const WavefrontState& st = p.st;
  return launch_trace<false>(p, st.ray_o, st.ray_d, st.queue[queue_in], st.counters + slot_queue_count(bounce),
                             st.counters + slot_trace_work(bounce), st.hit, nullptr, nullptr, counting, sm_count, s);
}

cudaError_t launch_trace_queue(const PassParams& p, const uint32_t* queue, uint32_t* count, uint32_t* work, bool counting, int sm_count,
                               cudaStream_t s) {
  const WavefrontState& st = p.st;
  return launch_trace<false>(p, st.ray_o, st.ray_d, queue, count, work, st.hit, nullptr, nullptr, counting, sm_count, s);
}

cudaError_t launch_trace_shadow(const PassParams& p, uint32_t bounce, bool counting, int sm_count, cudaStream_t s) {
  const WavefrontState& st = p.st;
  return launch_trace<true>(p, st.sh_o, st.sh_d, st.shadow_queue, st.counters + slot_shadow_count(bounce),
                            st.counters + slot_shadow_work(bounce), nullptr, st.L, st.sh_c, counting, sm_count, s);
}

}  // namespace cudapt
