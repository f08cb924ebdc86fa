// K1 raygen, K3 shade (+K4 compaction fused as warp-aggregated queue appends), K6 accumulate of the mesh
// wavefront integrator.  One pass traces `samples` samples of every owned pixel together:
//
//   raygen -> [ trace_closest -> shade (-> trace_shadow) ] x max_bounces -> accumulate
//
// The per-bounce loop body follows the reference's trace_ray (CPT.cpp:211-281): hit point o += t*d,
// throughput scaled BEFORE Russian roulette, roulette once bounce_count > rr_start with p = max(throughput),
// strict '>' test, new origin offset by eps along the normal, sky added on miss (CPT.cpp:231-235).  What the
// reference does not have (materials, triangles, area-light NEE, Philox) follows DESIGN.md "Mesh integrator
// numerics"; every expression is written in that spec's operation order (TU compiled with -fmad=false) so
// the result is bit-comparable with the CPU oracle.
#include "kernels.h"

namespace cudapt {
namespace {

constexpr float kInvPi = 0.318309886f;
constexpr float kF0 = 0.04f;
constexpr uint32_t kFlagSpecular = 1u << 8;

// Warp-aggregated append: one atomic per warp, returns this lane's slot (valid only where pred).
__device__ __forceinline__ uint32_t warp_append(uint32_t* counter, bool pred) {
  const unsigned m = __ballot_sync(0xffffffffu, pred);
  if (m == 0u) return 0u;
  const unsigned lane = threadIdx.x & 31u;
  const int leader = __ffs(m) - 1;
  uint32_t base = 0;
  if ((int)lane == leader) base = atomicAdd(counter, (uint32_t)__popc(m));
  base = __shfl_sync(0xffffffffu, base, leader);
  return base + (uint32_t)__popc(m & ((1u << lane) - 1u));
}

// Block-aggregated append to up to two queues at once: warp ballots -> per-warp counts in shared memory ->
// ONE atomic per queue per CTA round.  (Per-warp atomics on a single counter serialise in the L2 atomic unit:
// 260 k same-address atomics per 8.3 M-path pass cost ~0.3 ms; profiles/r01_notes.md.)  Must be called by
// every thread of the CTA.
struct BlockAppend {
  uint32_t slot_a, slot_b;
};
__device__ __forceinline__ BlockAppend block_append2(uint32_t* counter_a, bool pred_a, uint32_t* counter_b, bool pred_b) {
  __shared__ uint32_t s_cnt[2][kShadeBlock / 32];
  __shared__ uint32_t s_base[2];
  const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  const unsigned ma = __ballot_sync(0xffffffffu, pred_a);
  const unsigned mb = counter_b ? __ballot_sync(0xffffffffu, pred_b) : 0u;
  if (lane == 0) {
    s_cnt[0][warp] = (uint32_t)__popc(ma);
    s_cnt[1][warp] = (uint32_t)__popc(mb);
  }
  __syncthreads();
  if (threadIdx.x < 2) {
    uint32_t total = 0;
#pragma unroll
    for (int w = 0; w < kShadeBlock / 32; ++w) {
      const uint32_t c = s_cnt[threadIdx.x][w];
      s_cnt[threadIdx.x][w] = total;  // exclusive prefix
      total += c;
    }
    uint32_t* ctr = threadIdx.x == 0 ? counter_a : counter_b;
    s_base[threadIdx.x] = (total && ctr) ? atomicAdd(ctr, total) : 0u;
  }
  __syncthreads();
  BlockAppend r;
  r.slot_a = s_base[0] + s_cnt[0][warp] + (uint32_t)__popc(ma & ((1u << lane) - 1u));
  r.slot_b = s_base[1] + s_cnt[1][warp] + (uint32_t)__popc(mb & ((1u << lane) - 1u));
  __syncthreads();  // s_cnt / s_base are reused by the next round
  return r;
}

// Orthonormal basis around unit n (Duff et al. 2017), operation order as in the spec.
__device__ __forceinline__ void onb(f3 n, f3& t, f3& b) {
  const float sign = copysignf(1.0f, n.z);
  const float a = -1.0f / (sign + n.z);
  const float bb = (n.x * n.y) * a;
  t = mk3(1.0f + (sign * (n.x * n.x)) * a, sign * bb, -sign * n.x);
  b = mk3(bb, sign + (n.y * n.y) * a, -n.y);
}
__device__ __forceinline__ f3 to_world(f3 t, f3 b, f3 n, float x, float y, float z) { return (x * t + y * b) + z * n; }
__device__ __forceinline__ float pow5(float x) {
  const float x2 = x * x;
  return (x2 * x2) * x;
}
__device__ __forceinline__ float ggx_g1(float a2, float x) { return (2.0f * x) / (x + sqrtf(a2 + (1.0f - a2) * (x * x))); }

// ---- K1 ----------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kShadeBlock) raygen_kernel(const PassParams p) {
  const uint32_t total = p.n_pixels * p.samples;
  const uint32_t stride = gridDim.x * blockDim.x;
  const uint32_t rounds = (total + stride - 1) / stride;
  uint32_t id = blockIdx.x * blockDim.x + threadIdx.x;
  for (uint32_t k = 0; k < rounds; ++k, id += stride) {
    bool live = id < total;
    uint32_t pixel = 0, x = 0, y = 0, s = 0;
    if (live) {
      s = fd_div(p.fd_pixels, id);
      pixel = id - s * p.n_pixels;
      y = fd_div(p.fd_width, pixel);
      x = pixel - y * p.im.width;
      live = owns_pixel(p.im, x, y);
    }
    if (live) {
      const uint32_t frame = p.first_frame + s * p.frame_stride;
      f3 o, d;
      primary_ray_philox(p.cam, p.im, p.ip.jitter != 0, p.ip.seed, pixel, frame, x, y, &o, &d);
      p.st.ray_o[id] = make_float4(o.x, o.y, o.z, p.ip.tnear);
      p.st.ray_d[id] = make_float4(d.x, d.y, d.z, __int_as_float(0x7f800000));
      // throughput (1,1,1 | specular) and radiance 0 of a camera ray are implied at bounce 0: not stored
    }
    if (p.im.part_world <= 1u) {
      if (live) p.st.queue[0][id] = id;  // every pixel is owned: the queue is the identity, no atomics
    } else {
      const uint32_t slot = warp_append(p.st.counters + slot_queue_count(0), live);
      if (live) p.st.queue[0][slot] = id;
    }
  }
  if (p.im.part_world <= 1u && blockIdx.x == 0 && threadIdx.x == 0) p.st.counters[slot_queue_count(0)] = total;
}

// ---- K3 ----------------------------------------------------------------------------------------------
struct ShadeLaunch {
  PassParams p;
  uint32_t bounce;   // bounces already taken by the paths in the input queue (0 = camera rays)
  int queue_in;
};

// 4 CTAs/SM (64 registers): the kernel is bound by the latency of its dependent gathers (queue -> hit ->
// triangle -> material), so resident warps matter more than registers (profiles/r01_notes.md).
__global__ void __launch_bounds__(kShadeBlock, kShadeBlocksPerSM) shade_kernel(const ShadeLaunch a) {
  const PassParams& p = a.p;
  const WavefrontState& st = p.st;
  const uint32_t n = st.counters[slot_queue_count(a.bounce)];
  uint32_t* next_count = st.counters + slot_queue_count(a.bounce + 1);
  uint32_t* shadow_count = st.counters + slot_shadow_count(a.bounce);
  const uint32_t* qin = st.queue[a.queue_in];
  uint32_t* qout = st.queue[a.queue_in ^ 1];
  const bool use_nee = p.ip.nee && p.scene.n_lights > 0;
  const uint32_t stride = gridDim.x * blockDim.x;
  const uint32_t rounds = (n + stride - 1) / stride;
  uint32_t n_shades = 0, n_paths = 0;
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  // Two-deep software pipeline over the dependent gathers queue -> hit -> triangle -> material: the queue entry is
  // fetched two rounds ahead and the hit record one round ahead, so a round starts with both in registers and its
  // critical path is (ray state | triangle) -> material instead of four dependent loads (profiles/r01_notes.md:
  // the kernel was latency bound, long_scoreboard 7.2 per issue at 50 % issue utilisation).
  // Camera rays traced as packets (p.chunked): queue 0 is the identity when every pixel is owned and is not stored,
  // and the ray origin {camera, tnear} is implied (primary_packet.cu stores it only for its fall-back rays).
  const bool camera_implied = a.bounce == 0 && p.chunked != 0u;
  const bool ident = camera_implied && all_slots_owned(p) && p.explicit_q0 == 0u;
  uint32_t path_1 = i < n ? (ident ? i : qin[i]) : 0u;                             // entry of this round
  uint32_t path_2 = i + stride < n ? (ident ? i + stride : qin[i + stride]) : 0u;  // entry of the next round
  uint2 hit_1 = st.hit[path_1];
  for (uint32_t k = 0; k < rounds; ++k, i += stride) {
    bool cont = false, shadow = false;
    const uint32_t path = path_1;
    const uint2 h = hit_1;
    path_1 = path_2;
    path_2 = (k + 2 < rounds && i + 2 * stride < n) ? (ident ? i + 2 * stride : qin[i + 2 * stride]) : 0u;
    if (k + 1 < rounds) hit_1 = st.hit[path_1];  // path_1 was loaded a round ago: no stall here
    uint32_t s = 0, pixel = 0;
    if (i < n && path_decode(p, path, pixel, s)) {
      const uint32_t frame = p.first_frame + s * p.frame_stride;
      const float t = __uint_as_float(h.x);
      const float4 o4 = camera_implied ? make_float4(p.cam.origin.x, p.cam.origin.y, p.cam.origin.z, p.ip.tnear) : st.ray_o[path];
      const float4 d4 = st.ray_d[path];
      const float4 thr4 = a.bounce == 0 ? make_float4(1.0f, 1.0f, 1.0f, __uint_as_float(kFlagSpecular)) : st.thr[path];
      f3 o = mk3(o4.x, o4.y, o4.z);
      const f3 d = mk3(d4.x, d4.y, d4.z);
      f3 thr = mk3(thr4.x, thr4.y, thr4.z);
      const uint32_t flags = __float_as_uint(thr4.w);
      const bool specular_in = (flags & kFlagSpecular) != 0u;
      const float4 L4 = a.bounce == 0 ? make_float4(0.f, 0.f, 0.f, 0.f) : st.L[path];
      f3 L = mk3(L4.x, L4.y, L4.z);
      bool L_dirty = a.bounce == 0;  // the first shade initialises the path's radiance slot

      if (a.bounce == 0 && s == 0 && p.aov_prim != nullptr) {
        p.aov_prim[pixel] = h.y == 0xFFFFFFFFu ? 0xFFFFFFFFu : __float_as_uint(__ldg(p.scene.tris + 3 * (size_t)h.y).w);
        p.aov_t[pixel] = t;
      }
      if (h.y == 0xFFFFFFFFu) {  // miss: CPT.cpp:231-235
        const f3 sky = sky_radiance(p.ip, d);
        L = L + thr * (sky * p.ip.sky_scale);
        L_dirty = true;
        ++n_paths;
      } else {
        ++n_shades;
        const float4 ta = __ldg(p.scene.tris + 3 * (size_t)h.y + 0);
        const float4 tb = __ldg(p.scene.tris + 3 * (size_t)h.y + 1);
        const float4 tc = __ldg(p.scene.tris + 3 * (size_t)h.y + 2);
        uint32_t mat_id = __float_as_uint(tb.w);
        if (mat_id >= p.scene.n_materials) mat_id = 0;
        const float4 m0 = __ldg(p.scene.materials + 2 * (size_t)mat_id + 0);
        const float4 m1 = __ldg(p.scene.materials + 2 * (size_t)mat_id + 1);
        const f3 albedo = mk3(m0.x, m0.y, m0.z);
        const uint32_t mtype = __float_as_uint(m0.w);
        const f3 Le = mk3(m1.x, m1.y, m1.z);
        const float mparam = m1.w;
        const f3 e1 = mk3(tb.x, tb.y, tb.z), e2 = mk3(tc.x, tc.y, tc.z);

        o = mk3(o.x + t * d.x, o.y + t * d.y, o.z + t * d.z);  // CPT.cpp:238-241
        const f3 Ng = normalize3(cross3(e1, e2));
        const bool front = dot3(Ng, d) < 0.0f;
        const f3 nrm = front ? Ng : -Ng;
        if (front && (Le.x > 0.0f || Le.y > 0.0f || Le.z > 0.0f) && (specular_in || !use_nee)) {
          L = L + thr * Le;
          L_dirty = true;
        }
        const uint32_t bounce = a.bounce + 1;  // CPT.cpp:263
        const rand4 r = philox_block(pixel, frame, bounce, 0u, p.ip.seed);  // {rr, lobe, u1, u2}
        f3 wi = mk3(0.f, 0.f, 1.f), weight = mk3(0.f, 0.f, 0.f);
        float side = 1.0f;
        bool specular_out = true, absorbed = false;
        f3 tt, bb;
        if (mtype == 1u) {  // mirror
          const float cosi = -dot3(d, nrm);
          wi = normalize3(d + nrm * (2.0f * cosi));
          weight = albedo;
        } else if (mtype == 2u) {  // dielectric
          const float ior = mparam;
          const float eta = front ? 1.0f / ior : ior;
          const float cosi = -dot3(d, nrm);
          const float sin2t = (eta * eta) * (1.0f - cosi * cosi);
          const float r0s = (1.0f - ior) / (1.0f + ior);
          const float r0 = r0s * r0s;
          bool reflect = true;
          float cost = 0.0f;
          if (sin2t < 1.0f) {
            cost = sqrtf(1.0f - sin2t);
            const float cx = front ? cosi : cost;
            const float F = r0 + (1.0f - r0) * pow5(1.0f - cx);
            reflect = r.u1 < F;
          }
          if (reflect) {
            wi = normalize3(d + nrm * (2.0f * cosi));
          } else {
            wi = normalize3(d * eta + nrm * (eta * cosi - cost));
            side = -1.0f;
          }
          weight = albedo;
        } else if (mtype == 3u) {  // GGX-coated diffuse, one-sample lobe selection (p = 1/2 each)
          const float a2 = mparam * mparam;
          const f3 wo = -d;
          const float nov = fmaxf(dot3(nrm, wo), 1e-6f);
          onb(nrm, tt, bb);
          float c, sn;
          sincos2pi(r.u3, c, sn);
          if (r.u1 < 0.5f) {
            const float cos2h = (1.0f - r.u2) / (1.0f + (a2 - 1.0f) * r.u2);
            const float ch = sqrtf(cos2h), sh = sqrtf(fmaxf(0.0f, 1.0f - cos2h));
            const f3 hh = normalize3(to_world(tt, bb, nrm, sh * c, sh * sn, ch));
            const float voh = dot3(wo, hh);
            if (!(voh > 0.0f)) {
              absorbed = true;
            } else {
              wi = normalize3(hh * (2.0f * voh) - wo);
              const float noi = dot3(nrm, wi);
              if (!(noi > 0.0f)) {
                absorbed = true;
              } else {
                const float noh = fmaxf(dot3(nrm, hh), 1e-6f);
                const float F = kF0 + (1.0f - kF0) * pow5(1.0f - voh);
                const float G = ggx_g1(a2, nov) * ggx_g1(a2, noi);
                const float w = (((F * G) * voh) / (nov * noh)) * 2.0f;
                weight = mk3(w, w, w);
              }
            }
          } else {
            const float ct = sqrtf(r.u2), stheta = sqrtf(1.0f - r.u2);
            wi = normalize3(to_world(tt, bb, nrm, stheta * c, stheta * sn, ct));
            const float Fd = kF0 + (1.0f - kF0) * pow5(1.0f - nov);
            weight = albedo * ((1.0f - Fd) * 2.0f);
          }
        } else {  // Lambert
          if (use_nee) {
            const rand4 q = philox_block(pixel, frame, bounce, 1u, p.ip.seed);  // {light, v1, v2, -}
            const uint32_t nl = p.scene.n_lights;
            uint32_t kl = (uint32_t)(q.u0 * (float)nl);
            if (kl >= nl) kl = nl - 1;
            const uint32_t lt = __ldg(p.scene.lights + kl);
            const float4 la = __ldg(p.scene.tris + 3 * (size_t)lt + 0);
            const float4 lb = __ldg(p.scene.tris + 3 * (size_t)lt + 1);
            const float4 lc4 = __ldg(p.scene.tris + 3 * (size_t)lt + 2);
            uint32_t lmat = __float_as_uint(lb.w);
            if (lmat >= p.scene.n_materials) lmat = 0;
            const float4 lm1 = __ldg(p.scene.materials + 2 * (size_t)lmat + 1);
            const f3 lv0 = mk3(la.x, la.y, la.z), le1 = mk3(lb.x, lb.y, lb.z), le2 = mk3(lc4.x, lc4.y, lc4.z);
            const f3 lc = cross3(le1, le2);
            const float ll = sqrtf(dot3(lc, lc));
            const f3 Nl = lc * (1.0f / ll);
            const float area = 0.5f * ll;
            const float su = sqrtf(q.u1);
            const float b1 = su * (1.0f - q.u2), b2 = su * q.u2;
            const f3 pl = (lv0 + le1 * b1) + le2 * b2;
            const f3 ps = mk3(o.x + nrm.x * p.ip.eps, o.y + nrm.y * p.ip.eps, o.z + nrm.z * p.ip.eps);
            f3 wl = pl - ps;
            const float dist2 = dot3(wl, wl);
            const float dist = sqrtf(dist2);
            wl = wl * (1.0f / dist);
            const float cos_s = dot3(nrm, wl);
            const float cos_l = -dot3(Nl, wl);
            if (cos_s > 0.0f && cos_l > 0.0f) {
              const float g = ((cos_s * cos_l) / dist2) * (area * (float)nl);
              const f3 le = mk3(lm1.x, lm1.y, lm1.z);
              const f3 cshadow = ((thr * albedo) * le) * (g * kInvPi);
              st.sh_o[path] = make_float4(ps.x, ps.y, ps.z, dist * 0.999f);
              st.sh_d[path] = make_float4(wl.x, wl.y, wl.z, p.ip.tnear);
              st.sh_c[path] = make_float4(cshadow.x, cshadow.y, cshadow.z, 0.f);
              shadow = true;
            }
          }
          float c, sn;
          sincos2pi(r.u3, c, sn);
          const float ct = sqrtf(r.u2), stheta = sqrtf(1.0f - r.u2);
          onb(nrm, tt, bb);
          wi = normalize3(to_world(tt, bb, nrm, stheta * c, stheta * sn, ct));
          weight = albedo;
          specular_out = false;
        }
        if (!absorbed) {
          thr = thr * weight;  // CPT.cpp:260: before Russian roulette
          bool alive = true;
          if ((int)bounce > p.ip.rr_start) {  // CPT.cpp:264-270
            const float q = fminf(max3(thr), 1.0f);
            if (!(q > 0.0f) || r.u0 > q) alive = false;
            else thr = thr / q;
          }
          if (alive && (int)bounce < p.ip.max_bounces) {  // CPT.cpp:211
            const float off = side * p.ip.eps;
            st.ray_o[path] = make_float4(o.x + nrm.x * off, o.y + nrm.y * off, o.z + nrm.z * off, p.ip.tnear);  // CPT.cpp:277-280
            st.ray_d[path] = make_float4(wi.x, wi.y, wi.z, __int_as_float(0x7f800000));
            st.thr[path] = make_float4(thr.x, thr.y, thr.z, __uint_as_float(bounce | (specular_out ? kFlagSpecular : 0u)));
            cont = true;
          }
        }
        if (!cont) ++n_paths;
      }
      if (L_dirty) st.L[path] = make_float4(L.x, L.y, L.z, 0.f);
    }
    const BlockAppend ap = block_append2(next_count, cont, use_nee ? shadow_count : nullptr, shadow);
    if (cont) qout[ap.slot_a] = path;
    if (shadow) st.shadow_queue[ap.slot_b] = path;
  }
  const uint64_t ws = warp_sum_u32(n_shades), wp = warp_sum_u32(n_paths);
  if ((threadIdx.x & 31u) == 0u && (ws | wp)) {
    atomicAdd(&p.counters[kCntShades], (unsigned long long)ws);
    atomicAdd(&p.counters[kCntPaths], (unsigned long long)wp);
  }
}

// ---- K6: add the pass's per-path radiance to the accumulation buffer, in sample order (CPT.cpp:77-80) --
__global__ void __launch_bounds__(kShadeBlock) accumulate_kernel(const PassParams p) {
  const uint32_t stride = gridDim.x * blockDim.x;
  if (p.chunked) {
    // chunked layout: slot by slot (8 neighbouring pixels of a row per 8 threads); the ids of a slot are 32 * n_slots apart
    for (uint32_t g = blockIdx.x * blockDim.x + threadIdx.x; g < p.n_slots; g += stride) {
      uint32_t x, y;
      if (!slot_pixel(p, g, x, y) || (!p.dense && !owns_pixel(p.im, x, y))) continue;
      const uint32_t pixel = y * p.im.width + x;
      float4 acc = p.accum[pixel];
      for (uint32_t s = 0; s < p.samples; ++s) {
        const float4 l = p.st.L[slot_sample_id(p, g, s)];
        acc.x += l.x;
        acc.y += l.y;
        acc.z += l.z;
        acc.w += 1.0f;
      }
      p.accum[pixel] = acc;
    }
    return;
  }
  for (uint32_t pixel = blockIdx.x * blockDim.x + threadIdx.x; pixel < p.n_pixels; pixel += stride) {
    const uint32_t y = fd_div(p.fd_width, pixel), x = pixel - y * p.im.width;
    if (!owns_pixel(p.im, x, y)) continue;
    float4 acc = p.accum[pixel];
    for (uint32_t s = 0; s < p.samples; ++s) {
      const float4 l = p.st.L[s * p.n_pixels + pixel];
      acc.x += l.x;
      acc.y += l.y;
      acc.z += l.z;
      acc.w += 1.0f;
    }
    p.accum[pixel] = acc;
  }
}

inline int grid_for(uint32_t work, int sm_count, int block) {
  const long long want = ((long long)work + block - 1) / block;
  const long long cap = (long long)sm_count * (2048 / block);
  return (int)(want < 1 ? 1 : (want < cap ? want : cap));
}

}  // namespace

cudaError_t launch_raygen(const PassParams& p, int sm_count, cudaStream_t s) {
  raygen_kernel<<<grid_for(p.n_pixels * p.samples, sm_count, kShadeBlock), kShadeBlock, 0, s>>>(p);
  return cudaGetLastError();
}

cudaError_t launch_shade(const PassParams& p, uint32_t bounce, int queue_in, int sm_count, cudaStream_t s) {
  ShadeLaunch a{p, bounce, queue_in};
  shade_kernel<<<sm_count * kShadeBlocksPerSM, kShadeBlock, 0, s>>>(a);
  return cudaGetLastError();
}

cudaError_t launch_accumulate(const PassParams& p, int sm_count, cudaStream_t s) {
  accumulate_kernel<<<grid_for(p.chunked ? p.n_slots : p.n_pixels, sm_count, kShadeBlock), kShadeBlock, 0, s>>>(p);
  return cudaGetLastError();
}

}  // namespace cudapt
