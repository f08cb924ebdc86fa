// Device-side building blocks shared by the sphere megakernel and the mesh wavefront kernels.
//
// Numerics contract: this translation unit set is compiled with -fmad=false -prec-div=true
// -prec-sqrt=true -ftz=false, so every + - * / sqrtf below is one correctly rounded IEEE-754 binary32
// operation, in the order written.  That order follows the reference
// (libs/render/src/engines/pathtracer/backends/cpu/CPUPathTracer.cpp, "CPT.cpp") and glm's scalar
// definitions, which is what makes the accumulation buffer bit-comparable with the CPU path.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace cudapt {

struct f3 {
  float x, y, z;
};
__host__ __device__ __forceinline__ f3 mk3(float x, float y, float z) { return f3{x, y, z}; }
__host__ __device__ __forceinline__ f3 operator+(f3 a, f3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
__host__ __device__ __forceinline__ f3 operator-(f3 a, f3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
__host__ __device__ __forceinline__ f3 operator*(f3 a, f3 b) { return mk3(a.x * b.x, a.y * b.y, a.z * b.z); }
__host__ __device__ __forceinline__ f3 operator*(f3 a, float s) { return mk3(a.x * s, a.y * s, a.z * s); }
__host__ __device__ __forceinline__ f3 operator*(float s, f3 a) { return mk3(s * a.x, s * a.y, s * a.z); }
__host__ __device__ __forceinline__ f3 operator/(f3 a, float s) { return mk3(a.x / s, a.y / s, a.z / s); }
__host__ __device__ __forceinline__ f3 operator-(f3 a) { return mk3(-a.x, -a.y, -a.z); }
// glm::dot(vec3) = (x*x' + y*y') + z*z'
__host__ __device__ __forceinline__ float dot3(f3 a, f3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
// glm::cross
__host__ __device__ __forceinline__ f3 cross3(f3 a, f3 b) {
  return mk3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y);
}
// glm::normalize = v * (1 / sqrt(dot(v, v)))
__host__ __device__ __forceinline__ f3 normalize3(f3 a) { return a * (1.0f / sqrtf(dot3(a, a))); }
__host__ __device__ __forceinline__ float max3(f3 a) { return fmaxf(fmaxf(a.x, a.y), a.z); }

// ---- RNG -------------------------------------------------------------------------------------------
// CPT.cpp:192-195
__device__ __forceinline__ uint32_t pcg_seed(uint32_t x, uint32_t y, uint32_t width, uint32_t frame) {
  return x + y * width + frame * 982451653u;
}
// CPT.cpp:294-301.  float(w)/4294967295.0f: the divisor is 2^32 in binary32, so the quotient equals
// RN(float(w)) * 2^-32 exactly (a power-of-two scale), range [0, 1] inclusive.
__device__ __forceinline__ float pcg_next(uint32_t& state) {
  state = state * 747796405u + 2891336453u;
  uint32_t w = ((state >> ((state >> 28) + 4u)) ^ state) * 277803737u;
  w = (w >> 22) ^ w;
  return __uint2float_rn(w) * 0x1p-32f;
}

// Philox4x32-10 keyed (0xB200CAFE, seed), counter (pixel, sample, bounce, block).
struct rand4 {
  float u0, u1, u2, u3;
};
__device__ __forceinline__ float u01(uint32_t w) { return __uint2float_rn(w >> 8) * 0x1p-24f; }
__device__ __forceinline__ rand4 philox_block(uint32_t pixel, uint32_t sample, uint32_t bounce, uint32_t block,
                                              uint32_t seed) {
  uint32_t c0 = pixel, c1 = sample, c2 = bounce, c3 = block;
  uint32_t k0 = 0xB200CAFEu, k1 = seed;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
    c0 = n0, c1 = lo1, c2 = n2, c3 = lo0;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  return rand4{u01(c0), u01(c1), u01(c2), u01(c3)};
}

// cos/sin(2*pi*u), u in [0,1]: quadrant split + Taylor/Horner with separate mul and add (DESIGN.md,
// "Mesh integrator numerics"); only + - * so CPU and GPU agree bit-for-bit.
__device__ __forceinline__ void sincos2pi(float u, float& c, float& s) {
  const float x = u * 4.0f;
  int q = (int)x;
  if (q > 3) q = 3;
  const float f = x - (float)q;
  const float a = f * 1.57079637f;
  const float a2 = a * a;
  float ps = -2.50521084e-08f;
  ps = ps * a2 + 2.75573192e-06f;
  ps = ps * a2 + -1.98412701e-04f;
  ps = ps * a2 + 8.33333377e-03f;
  ps = ps * a2 + -1.66666672e-01f;
  ps = ps * a2 + 1.0f;
  const float sn = ps * a;
  float pc = 2.08767570e-09f;
  pc = pc * a2 + -2.75573192e-07f;
  pc = pc * a2 + 2.48015876e-05f;
  pc = pc * a2 + -1.38888892e-03f;
  pc = pc * a2 + 4.16666679e-02f;
  pc = pc * a2 + -0.5f;
  const float cs = pc * a2 + 1.0f;
  c = (q == 0) ? cs : (q == 1) ? -sn : (q == 2) ? -cs : sn;
  s = (q == 0) ? sn : (q == 1) ? cs : (q == 2) ? -sn : -cs;
}

// ---- camera / sky ----------------------------------------------------------------------------------
struct CameraParams {
  f3 origin, right, up, forward;
  float tan_half_fov;
  float lens_radius, focus_distance;  // thin lens (Philox mode only); lens_radius 0 = pinhole
};
struct ImageParams {
  uint32_t width, height;
  float inv_width, inv_height, aspect;  // 1.0f/W, 1.0f/H, (float)W/(float)H  (CPT.cpp:53-54,65)
  // interleaved-tile partition (cudapt_set_tile_partition); world == 1 disables
  uint32_t part_rank, part_world, tile_w, tile_h, tiles_x;
};
// CPT.cpp:62-73; (jx, jy) = sub-pixel offset, 0 = the reference's pixel-corner sampling.
__device__ __forceinline__ f3 primary_dir(const CameraParams& cam, const ImageParams& im, uint32_t x, uint32_t y,
                                          float jx, float jy) {
  const float u = ((float)x + jx) * im.inv_width;
  const float v = 1.0f - ((float)y + jy) * im.inv_height;
  const float uv_x = ((u * 2.0f - 1.0f) * im.aspect) * cam.tan_half_fov;
  const float uv_y = (v * 2.0f - 1.0f) * cam.tan_half_fov;
  const float len = sqrtf(uv_x * uv_x + uv_y * uv_y + 1.0f);
  const float cx = uv_x / len, cy = uv_y / len, cz = 1.0f / len;
  return mk3((cx * cam.right.x + cy * cam.up.x) + cz * cam.forward.x,
             (cx * cam.right.y + cy * cam.up.y) + cz * cam.forward.y,
             (cx * cam.right.z + cy * cam.up.z) + cz * cam.forward.z);
}
// Camera ray of (pixel, frame) in Philox mode: sub-pixel jitter from block 2 words 0,1 and -- thin lens, SURVEY.md
// 8f-1 -- a lens point by REJECTION from the same stream (block 2 words 2,3, then blocks 3 and 4, two candidates each;
// the lens centre if all five fall outside the unit disk), so that only + - * / sqrt and comparisons are involved and
// the CPU twin (oracle/twin/twin.cpp primary_ray) reproduces origin and direction bit for bit.
__device__ __forceinline__ void primary_ray_philox(const CameraParams& cam, const ImageParams& im, bool jitter, uint32_t seed,
                                                   uint32_t pixel, uint32_t frame, uint32_t x, uint32_t y, f3* o, f3* d) {
  float jx = 0.0f, jy = 0.0f;
  const bool lens = cam.lens_radius > 0.0f;
  rand4 j;
  j.u0 = j.u1 = j.u2 = j.u3 = 0.0f;
  if (jitter || lens) j = philox_block(pixel, frame, 0u, 2u, seed);
  if (jitter) jx = j.u0, jy = j.u1;
  if (!lens) {
    *o = cam.origin;
    *d = primary_dir(cam, im, x, y, jx, jy);
    return;
  }
  float a = 0.0f, b = 0.0f;
  bool found = false;
  for (int c = 0; c < 5 && !found; ++c) {
    float ua, ub;
    if (c == 0) {
      ua = j.u2, ub = j.u3;
    } else {
      const rand4 q = philox_block(pixel, frame, 0u, 3u + (uint32_t)(c - 1) / 2u, seed);
      ua = ((c - 1) & 1) ? q.u2 : q.u0, ub = ((c - 1) & 1) ? q.u3 : q.u1;
    }
    const float ca = ua * 2.0f - 1.0f, cb = ub * 2.0f - 1.0f;
    if (ca * ca + cb * cb <= 1.0f) a = ca, b = cb, found = true;
  }
  const float lx = a * cam.lens_radius, ly = b * cam.lens_radius;
  const float u = ((float)x + jx) * im.inv_width;
  const float v = 1.0f - ((float)y + jy) * im.inv_height;
  const float uv_x = ((u * 2.0f - 1.0f) * im.aspect) * cam.tan_half_fov;
  const float uv_y = (v * 2.0f - 1.0f) * cam.tan_half_fov;
  const float fx = uv_x * cam.focus_distance - lx, fy = uv_y * cam.focus_distance - ly, fz = cam.focus_distance;
  const float len = sqrtf((fx * fx + fy * fy) + fz * fz);
  const float cx = fx / len, cy = fy / len, cz = fz / len;
  *d = mk3((cx * cam.right.x + cy * cam.up.x) + cz * cam.forward.x, (cx * cam.right.y + cy * cam.up.y) + cz * cam.forward.y,
           (cx * cam.right.z + cy * cam.up.z) + cz * cam.forward.z);
  *o = mk3((cam.origin.x + lx * cam.right.x) + ly * cam.up.x, (cam.origin.y + lx * cam.right.y) + ly * cam.up.y,
           (cam.origin.z + lx * cam.right.z) + ly * cam.up.z);
}
// CPT.cpp:286-292; glm::mix(x, y, a) = x*(1-a) + y*a
__device__ __forceinline__ f3 sample_sky(f3 d) {
  const float t = 0.5f * (d.y + 1.0f);
  const float omt = 1.0f - t;
  return mk3(1.0f * omt + 0.5f * t, 1.0f * omt + 0.7f * t, 1.0f * omt + 1.0f * t);
}
__device__ __forceinline__ bool owns_pixel(const ImageParams& im, uint32_t x, uint32_t y) {
  if (im.part_world <= 1u) return true;
  const uint32_t tx = x / im.tile_w, ty = y / im.tile_h;
  return (ty * im.tiles_x + tx + ty) % im.part_world == im.part_rank;
}

struct IntegratorParams {
  int32_t rng_mode, max_bounces, rr_start, jitter, nee;
  float tnear, eps, albedo, sky_scale;
  uint32_t seed;
  const float4* env;  // octahedral environment map (env_size x env_size texels, rgb + pad) or nullptr = procedural sky
  uint32_t env_size;
};

// Radiance arriving from direction d when nothing is hit: the reference's procedural sky (CPT.cpp:286-292), or --
// SURVEY.md 8f-4, after the commented-out Scene::setEnvironmentMap of libs/render/src/Scene.cpp:124-169 -- a
// bilinear lookup in an OCTAHEDRAL map (pole axis +Y).  The octahedral mapping needs only + - * / and
// comparisons (no atan2/acos), so the CPU oracle reproduces every texel address and weight bit for bit.
__device__ __forceinline__ f3 sky_radiance(const IntegratorParams& ip, f3 d) {
  if (ip.env == nullptr) return sample_sky(d);
  const float s = (fabsf(d.x) + fabsf(d.y)) + fabsf(d.z);
  float px = d.x / s, pz = d.z / s;
  if (d.y < 0.0f) {  // lower hemisphere: fold the corners outwards
    const float fx = 1.0f - fabsf(pz), fz = 1.0f - fabsf(px);
    px = d.x < 0.0f ? -fx : fx;
    pz = d.z < 0.0f ? -fz : fz;
  }
  const float n = (float)ip.env_size, hi = (float)(ip.env_size - 1u);
  float x = (px * 0.5f + 0.5f) * n - 0.5f, y = (pz * 0.5f + 0.5f) * n - 0.5f;
  x = x < 0.0f ? 0.0f : (x > hi ? hi : x);
  y = y < 0.0f ? 0.0f : (y > hi ? hi : y);
  const uint32_t x0 = (uint32_t)x, y0 = (uint32_t)y;
  const uint32_t x1 = x0 + 1u < ip.env_size ? x0 + 1u : x0, y1 = y0 + 1u < ip.env_size ? y0 + 1u : y0;
  const float fx = x - (float)x0, fy = y - (float)y0, gx = 1.0f - fx, gy = 1.0f - fy;
  const float4 c00 = __ldg(ip.env + (size_t)y0 * ip.env_size + x0), c10 = __ldg(ip.env + (size_t)y0 * ip.env_size + x1);
  const float4 c01 = __ldg(ip.env + (size_t)y1 * ip.env_size + x0), c11 = __ldg(ip.env + (size_t)y1 * ip.env_size + x1);
  return mk3((c00.x * gx + c10.x * fx) * gy + (c01.x * gx + c11.x * fx) * fy,
             (c00.y * gx + c10.y * fx) * gy + (c01.y * gx + c11.y * fx) * fy,
             (c00.z * gx + c10.z * fx) * gy + (c01.z * gx + c11.z * fx) * fy);
}

__device__ __forceinline__ uint64_t warp_sum_u32(uint32_t v) {
  // values are small per thread; sum in 32 bit across the warp
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

}  // namespace cudapt
