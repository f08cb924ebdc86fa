// Sphere integrator: the reference hot path (CPT.cpp:43-85 render loop, :197-284 trace_ray, :286-326
// sky / RNG / bounce) as ONE kernel.  "CPT.cpp" = libs/render/src/engines/pathtracer/backends/cpu/
// CPUPathTracer.cpp of imisumi/software-path-tracer.
//
// Design (B200): the whole scene (16 B per sphere) lives in shared memory and is read with broadcast
// LDS.128, so the only HBM traffic is the 32 B/pixel accumulation read-modify-write -- once per LAUNCH,
// not per frame: each thread owns one pixel, loops over the launch's frames and keeps the running sum in
// registers, adding frame after frame in the reference's order (CPT.cpp:77-80), which keeps the float sum
// bit-identical to the serial CPU loop.  The frame loop and the bounce loop are flattened into one
// "one ray per iteration" loop so a lane whose path ended starts its next frame immediately instead of
// idling until the slowest path of the warp finishes (path regeneration).
//
// Bit parity with the CPU path: sphere test in the exact operation order of the Embree stand-in
// (oracle/ref/embree_shim.cpp), camera/sky/accumulate as CPT.cpp, and the cosine
// bounce evaluated the way g++/glibc compile CPT.cpp:303-326 (double-precision sin/cos/sqrt, int abs()).
#include "common.cuh"
#include "kernels.h"

namespace cudapt {

namespace {

struct SphereHit {
  int idx;
  float t;
  f3 Ng;
};

// Closest hit over all spheres; candidate replaces the current hit only when strictly closer, so ties
// keep the lowest index (upload order = ascending render::NodeID).
__device__ __forceinline__ SphereHit intersect_spheres(const float4* __restrict__ sph, int n, f3 o, f3 d, float tnear) {
  SphereHit h;
  h.idx = -1;
  h.t = __int_as_float(0x7f800000);
  h.Ng = mk3(0.f, 0.f, 0.f);
  const float rd2 = 1.0f / ((d.x * d.x + d.y * d.y) + d.z * d.z);
#pragma unroll 2
  for (int i = 0; i < n; ++i) {
    const float4 c = sph[i];
    const float c0x = c.x - o.x, c0y = c.y - o.y, c0z = c.z - o.z;
    const float proj = ((c0x * d.x + c0y * d.y) + c0z * d.z) * rd2;
    const float px = c0x - proj * d.x, py = c0y - proj * d.y, pz = c0z - proj * d.z;
    const float l2 = (px * px + py * py) + pz * pz;
    const float r2 = c.w * c.w;
    if (l2 > r2) continue;
    const float td = sqrtf((r2 - l2) * rd2);
    const float tf = proj - td, tb = proj + td;
    const bool vf = tnear <= tf;  // tfar = +inf: the upper bound always holds for finite t
    const bool vb = tnear <= tb;
    if (!vf && !vb) continue;
    const float t = vf ? tf : tb;
    if (!(t < h.t)) continue;     // strictly closer only (ties keep the earlier sphere)
    const float sgn = vf ? -td : td;
    h.idx = i;
    h.t = t;
    h.Ng = mk3(sgn * d.x - px, sgn * d.y - py, sgn * d.z - pz);
  }
  return h;
}

// CPT.cpp:303-326 as compiled by g++ against glibc: unqualified sqrt/cos/sin bind
// to the double overloads and abs() to int abs(int).  sqrt in double then rounded to float equals sqrtf.
__device__ __forceinline__ f3 reference_bounce(f3 n, float u1, float u2) {
  const float cosTheta = sqrtf(u1);
  const float sinTheta = sqrtf(1.0f - u1);
  const float phi = (2.0f * 3.14159274f) * u2;
  double sd, cd;
  sincos((double)phi, &sd, &cd);
  const float x = (float)((double)sinTheta * cd);
  const float y = (float)((double)sinTheta * sd);
  const float z = cosTheta;
  const f3 up = ((int)n.z == 0) ? mk3(0.f, 0.f, 1.f) : mk3(1.f, 0.f, 0.f);
  const f3 tangent = normalize3(cross3(up, n));
  const f3 bitangent = cross3(n, tangent);
  return (x * tangent + y * bitangent) + z * n;
}

template <bool kPhilox>
__global__ void __launch_bounds__(kSphereBlockX* kSphereBlockY)
    render_spheres_kernel(const SphereLaunch a) {
  extern __shared__ float4 s_sph[];
  const int tid = threadIdx.y * kSphereBlockX + threadIdx.x;
  for (int i = tid; i < a.n_spheres; i += kSphereBlockX * kSphereBlockY) s_sph[i] = a.spheres[i];
  __syncthreads();

  const uint32_t x = blockIdx.x * kSphereBlockX + threadIdx.x;
  const uint32_t y = blockIdx.y * kSphereBlockY + threadIdx.y;
  const bool live = x < a.im.width && y < a.im.height && owns_pixel(a.im, x, y);
  uint32_t n_rays = 0, n_shades = 0, n_paths = 0;
  if (live) {
    const uint32_t pixel = y * a.im.width + x;
    float4 acc = a.accum[pixel];
    const f3 d0 = primary_dir(a.cam, a.im, x, y, 0.0f, 0.0f);

    uint32_t k = 0;
    uint32_t frame = a.first_frame;
    uint32_t rng = pcg_seed(x, y, a.im.width, frame);
    f3 o = a.cam.origin, d = d0;
    if (kPhilox && (a.ip.jitter || a.cam.lens_radius > 0.0f))
      primary_ray_philox(a.cam, a.im, a.ip.jitter != 0, a.ip.seed, pixel, frame, x, y, &o, &d);
    f3 thr = mk3(1.f, 1.f, 1.f);
    f3 L = mk3(0.f, 0.f, 0.f);
    int bounce = 0;

    while (k < a.n_frames) {
      bool done = false;
      const SphereHit h = intersect_spheres(s_sph, a.n_spheres, o, d, a.ip.tnear);
      ++n_rays;
      if (bounce == 0 && k == 0 && a.aov_prim != nullptr) {
        a.aov_prim[pixel] = h.idx < 0 ? 0xFFFFFFFFu : a.node_id[h.idx];
        a.aov_t[pixel] = h.t;
      }
      if (h.idx < 0) {  // CPT.cpp:231-235
        const f3 sky = sky_radiance(a.ip, d);
        L = L + thr * (a.ip.sky_scale == 1.0f ? sky : sky * a.ip.sky_scale);
        done = true;
      } else {
        ++n_shades;
        o.x += h.t * d.x;  // CPT.cpp:238-241
        o.y += h.t * d.y;
        o.z += h.t * d.z;
        const float inv_len = 1.0f / sqrtf(h.Ng.x * h.Ng.x + h.Ng.y * h.Ng.y + h.Ng.z * h.Ng.z);
        const f3 nrm = mk3(h.Ng.x * inv_len, h.Ng.y * inv_len, h.Ng.z * inv_len);
        thr = thr * a.ip.albedo;  // CPT.cpp:260, before Russian roulette
        ++bounce;                 // CPT.cpp:263
        rand4 r;
        if (kPhilox) r = philox_block(pixel, frame, (uint32_t)bounce, 0u, a.ip.seed);
        if (bounce > a.ip.rr_start) {  // CPT.cpp:264-270
          const float q = max3(thr);
          const float u = kPhilox ? r.u0 : pcg_next(rng);
          if (u > q) done = true;
          else thr = thr / q;
        }
        // CPT.cpp:211: after the max_bounces-th hit the loop exits; the direction the reference still draws
        // there (CPT.cpp:274) is never used, so it is not computed here.
        if (!done && bounce >= a.ip.max_bounces) done = true;
        if (!done) {
          float u1, u2;
          if (kPhilox) {
            u1 = r.u2, u2 = r.u3;
          } else {
            u1 = pcg_next(rng);
            u2 = pcg_next(rng);
          }
          d = reference_bounce(nrm, u1, u2);  // CPT.cpp:274
          o.x += nrm.x * a.ip.eps;            // CPT.cpp:277-280
          o.y += nrm.y * a.ip.eps;
          o.z += nrm.z * a.ip.eps;
        }
      }
      if (done) {
        acc.x += L.x;  // CPT.cpp:77-80
        acc.y += L.y;
        acc.z += L.z;
        acc.w += 1.0f;
        ++n_paths;
        ++k;
        frame += a.frame_stride;
        rng = pcg_seed(x, y, a.im.width, frame);
        o = a.cam.origin;
        d = d0;
        if (kPhilox && (a.ip.jitter || a.cam.lens_radius > 0.0f))
          primary_ray_philox(a.cam, a.im, a.ip.jitter != 0, a.ip.seed, pixel, frame, x, y, &o, &d);
        thr = mk3(1.f, 1.f, 1.f);
        L = mk3(0.f, 0.f, 0.f);
        bounce = 0;
      }
    }
    a.accum[pixel] = acc;
  }
  // one atomic per warp per counter
  const uint64_t wr = warp_sum_u32(n_rays), ws = warp_sum_u32(n_shades), wp = warp_sum_u32(n_paths);
  if ((tid & 31) == 0 && wp) {
    atomicAdd(&a.counters[kCntRays], (unsigned long long)wr);
    atomicAdd(&a.counters[kCntShades], (unsigned long long)ws);
    atomicAdd(&a.counters[kCntPaths], (unsigned long long)wp);
  }
}

// get_render_result(): CPT.cpp:87-117 + Color.h:7-10
__global__ void __launch_bounds__(256) resolve_rgba8_kernel(const float4* __restrict__ accum, uint32_t* __restrict__ out,
                                                            uint32_t n_pixels, float frame_count) {
  const uint32_t i = blockIdx.x * 256u + threadIdx.x;
  if (i >= n_pixels) return;
  const float4 a = accum[i];
  float c[4] = {a.x / frame_count, a.y / frame_count, a.z / frame_count, a.w / frame_count};
  uint32_t b[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    float v = c[k];
    v = (v < 0.0f) ? 0.0f : ((1.0f < v) ? 1.0f : v);  // std::clamp
    b[k] = __float2uint_rz(v * 255.0f) & 0xFFu;       // (uint8_t)(v * 255.0f)
  }
  out[i] = (b[0] << 24) | (b[1] << 16) | (b[2] << 8) | b[3];
}

}  // namespace

cudaError_t launch_render_spheres(const SphereLaunch& a, bool philox, cudaStream_t stream) {
  const dim3 block(kSphereBlockX, kSphereBlockY);
  const dim3 grid((a.im.width + kSphereBlockX - 1) / kSphereBlockX, (a.im.height + kSphereBlockY - 1) / kSphereBlockY);
  const size_t smem = sizeof(float4) * (size_t)a.n_spheres;
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(philox ? render_spheres_kernel<true> : render_spheres_kernel<false>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  if (philox)
    render_spheres_kernel<true><<<grid, block, smem, stream>>>(a);
  else
    render_spheres_kernel<false><<<grid, block, smem, stream>>>(a);
  return cudaGetLastError();
}

cudaError_t launch_resolve_rgba8(const float4* accum, uint32_t* out, uint32_t n_pixels, uint32_t frame_count,
                                 cudaStream_t stream) {
  resolve_rgba8_kernel<<<(n_pixels + 255u) / 256u, 256, 0, stream>>>(accum, out, n_pixels, (float)frame_count);
  return cudaGetLastError();
}

}  // namespace cudapt
