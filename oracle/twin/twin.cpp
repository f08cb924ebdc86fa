// TEST INFRASTRUCTURE (oracle, Tier 2: "CPU twin") -- never linked into or called by the product.
// See twin.h for what pins it.  "CPT.cpp" = libs/render/src/engines/pathtracer/backends/cpu/CPUPathTracer.cpp
// of the reference.  Compile with -ffp-contract=off: every expression below is
// written in the exact operation order that is part of the parity contract.
#include "twin.h"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <numeric>
#include <vector>
#include <atomic>
#include <thread>

namespace {

struct V3 {
  float x, y, z;
};
inline V3 v3(float x, float y, float z) { return V3{x, y, z}; }
inline V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline V3 operator*(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
inline V3 operator*(float s, V3 a) { return v3(s * a.x, s * a.y, s * a.z); }
inline V3 operator/(V3 a, float s) { return v3(a.x / s, a.y / s, a.z / s); }
inline V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }
// glm::dot for vec3: tmp = a*b; tmp.x + tmp.y + tmp.z  (glm/detail/func_geometric.inl compute_dot)
inline float dot(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
// glm::cross (glm/detail/func_geometric.inl compute_cross)
inline V3 cross(V3 a, V3 b) { return v3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y); }
// glm::normalize = v * inversesqrt(dot(v,v)), inversesqrt(x) = 1/sqrt(x)  (func_geometric.inl:98-105)
inline V3 normalize(V3 a) { return a * (1.0f / sqrtf(dot(a, a))); }
inline float max3(V3 a) { return std::max(std::max(a.x, a.y), a.z); }

// ------------------------------------------------------------------------------------------------------
// RNG
// ------------------------------------------------------------------------------------------------------
// CPT.cpp:192-195
inline uint32_t pcg_seed(uint32_t x, uint32_t y, uint32_t width, uint32_t frame) {
  return x + y * width + frame * 982451653U;
}
// CPT.cpp:294-301.  4294967295.0f is 2^32 in float, so the quotient is RN(float(w)) * 2^-32 in [0, 1].
inline float pcg_next(uint32_t& state, uint32_t* raw = nullptr) {
  state = state * 747796405u + 2891336453u;
  uint32_t result = ((state >> ((state >> 28) + 4)) ^ state) * 277803737u;
  result = (result >> 22) ^ result;
  if (raw) *raw = result;
  return (float)result / 4294967295.0f;
}

// Philox4x32-10 (Salmon et al., SC'11), the published round function and Weyl key schedule.
inline void philox4x32_10(const uint32_t c_in[4], const uint32_t k_in[2], uint32_t out[4]) {
  uint32_t c0 = c_in[0], c1 = c_in[1], c2 = c_in[2], c3 = c_in[3];
  uint32_t k0 = k_in[0], k1 = k_in[1];
  for (int r = 0; r < 10; ++r) {
    const uint64_t p0 = (uint64_t)0xD2511F53u * c0;
    const uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
    const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
    const uint32_t n1 = (uint32_t)p1;
    const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
    const uint32_t n3 = (uint32_t)p0;
    c0 = n0, c1 = n1, c2 = n2, c3 = n3;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  out[0] = c0, out[1] = c1, out[2] = c2, out[3] = c3;
}
constexpr uint32_t kPhiloxKey0 = 0xB200CAFEu;
inline float u01(uint32_t w) { return (float)(w >> 8) * 0x1p-24f; }  // [0,1), exact in float

struct Rand4 {
  float u[4];
};
inline Rand4 philox_block(uint32_t pixel, uint32_t sample, uint32_t bounce, uint32_t block, uint32_t seed) {
  const uint32_t ctr[4] = {pixel, sample, bounce, block};
  const uint32_t key[2] = {kPhiloxKey0, seed};
  uint32_t w[4];
  philox4x32_10(ctr, key, w);
  return Rand4{{u01(w[0]), u01(w[1]), u01(w[2]), u01(w[3])}};
}

// cos/sin of 2*pi*u, u in [0,1]: quadrant split + Taylor polynomials in Horner form with separate
// multiplies and adds.  Uses only + - * so that any IEEE-754 binary32 machine reproduces the bits.
inline void sincos2pi(float u, float& c, float& s) {
  const float x = u * 4.0f;
  int q = (int)x;
  if (q > 3) q = 3;  // u == 1.0 (possible in PCG mode) -> angle 2*pi handled as q=3, f=1
  const float f = x - (float)q;
  const float a = f * 1.57079637f;
  const float a2 = a * a;
  float ps = -2.50521084e-08f;             // -1/11!
  ps = ps * a2 + 2.75573192e-06f;          //  1/9!
  ps = ps * a2 + -1.98412701e-04f;         // -1/7!
  ps = ps * a2 + 8.33333377e-03f;          //  1/5!
  ps = ps * a2 + -1.66666672e-01f;         // -1/3!
  ps = ps * a2 + 1.0f;
  const float sn = ps * a;
  float pc = 2.08767570e-09f;              //  1/12!
  pc = pc * a2 + -2.75573192e-07f;         // -1/10!
  pc = pc * a2 + 2.48015876e-05f;          //  1/8!
  pc = pc * a2 + -1.38888892e-03f;         // -1/6!
  pc = pc * a2 + 4.16666679e-02f;          //  1/4!
  pc = pc * a2 + -0.5f;
  const float cs = pc * a2 + 1.0f;
  switch (q) {
    case 0: c = cs, s = sn; break;
    case 1: c = -sn, s = cs; break;
    case 2: c = -cs, s = -sn; break;
    default: c = sn, s = -cs; break;
  }
}

// ------------------------------------------------------------------------------------------------------
// Shared pieces: camera, sky, resolve
// ------------------------------------------------------------------------------------------------------
struct Camera {
  V3 origin, right, up, forward;
  float tan_half_fov;
  float lens_radius, focus_distance;
};
inline Camera camera_of(const twin_params& p) {
  return Camera{v3(p.cam_origin[0], p.cam_origin[1], p.cam_origin[2]), v3(p.cam_right[0], p.cam_right[1], p.cam_right[2]),
                v3(p.cam_up[0], p.cam_up[1], p.cam_up[2]), v3(p.cam_forward[0], p.cam_forward[1], p.cam_forward[2]),
                p.cam_tan_half_fov, p.rng_mode == 1 ? p.cam_lens_radius : 0.0f, p.cam_focus_distance};
}
// CPT.cpp:53-54,62-73 generalised: (jx,jy) is the sub-pixel offset (0 = reference corner sampling); with the
// default basis/tan_half_fov the extra multiplies are by exactly 1 or 0 and the reference bits are kept.
inline V3 primary_dir(const Camera& cam, uint32_t x, uint32_t y, uint32_t W, uint32_t H, float jx, float jy) {
  const float inv_height = 1.0f / H;
  const float inv_width = 1.0f / W;
  const float aspect_ratio = (float)W / (float)H;
  const float u = ((float)x + jx) * inv_width;
  const float v = 1.0f - ((float)y + jy) * inv_height;
  const float uv_x = ((u * 2.0f - 1.0f) * aspect_ratio) * cam.tan_half_fov;
  const float uv_y = (v * 2.0f - 1.0f) * cam.tan_half_fov;
  const float len = sqrtf(uv_x * uv_x + uv_y * uv_y + 1.0f);
  const float cx = uv_x / len, cy = uv_y / len, cz = 1.0f / len;
  return v3((cx * cam.right.x + cy * cam.up.x) + cz * cam.forward.x, (cx * cam.right.y + cy * cam.up.y) + cz * cam.forward.y,
            (cx * cam.right.z + cy * cam.up.z) + cz * cam.forward.z);
}
// Thin lens (DESIGN.md "camera"): a point on the lens disk by REJECTION from the Philox stream -- candidates (2u-1, 2v-1)
// from block 2 words 2,3, then blocks 3 and 4 (two candidates each), the first inside the unit disk wins, the lens
// centre if none does (p = 4.6e-4) -- so that only + - * / sqrt and comparisons are needed and CPU and GPU agree bit
// for bit.  The ray leaves origin + lx*right + ly*up towards the point where the pinhole ray of the same pixel sample
// crosses the plane of focus.
inline void primary_ray(const Camera& cam, uint32_t x, uint32_t y, uint32_t W, uint32_t H, float jx, float jy, const Rand4& j2,
                        uint32_t pixel, uint32_t frame, uint32_t seed, V3* o, V3* d) {
  if (!(cam.lens_radius > 0.0f)) {
    *o = cam.origin;
    *d = primary_dir(cam, x, y, W, H, jx, jy);
    return;
  }
  float a = 0.0f, b = 0.0f;
  bool found = false;
  for (int c = 0; c < 5 && !found; ++c) {
    float ua, ub;
    if (c == 0) {
      ua = j2.u[2], ub = j2.u[3];
    } else {
      const Rand4 q = philox_block(pixel, frame, 0, 3 + (uint32_t)(c - 1) / 2, seed);
      ua = q.u[((c - 1) & 1) * 2], ub = q.u[((c - 1) & 1) * 2 + 1];
    }
    const float ca = ua * 2.0f - 1.0f, cb = ub * 2.0f - 1.0f;
    if (ca * ca + cb * cb <= 1.0f) a = ca, b = cb, found = true;
  }
  const float lx = a * cam.lens_radius, ly = b * cam.lens_radius;
  const float inv_height = 1.0f / H;
  const float inv_width = 1.0f / W;
  const float aspect_ratio = (float)W / (float)H;
  const float u = ((float)x + jx) * inv_width;
  const float v = 1.0f - ((float)y + jy) * inv_height;
  const float uv_x = ((u * 2.0f - 1.0f) * aspect_ratio) * cam.tan_half_fov;
  const float uv_y = (v * 2.0f - 1.0f) * cam.tan_half_fov;
  const float fx = uv_x * cam.focus_distance - lx, fy = uv_y * cam.focus_distance - ly, fz = cam.focus_distance;
  const float len = sqrtf((fx * fx + fy * fy) + fz * fz);
  const float cx = fx / len, cy = fy / len, cz = fz / len;
  *d = v3((cx * cam.right.x + cy * cam.up.x) + cz * cam.forward.x, (cx * cam.right.y + cy * cam.up.y) + cz * cam.forward.y,
          (cx * cam.right.z + cy * cam.up.z) + cz * cam.forward.z);
  *o = v3((cam.origin.x + lx * cam.right.x) + ly * cam.up.x, (cam.origin.y + lx * cam.right.y) + ly * cam.up.y,
          (cam.origin.z + lx * cam.right.z) + ly * cam.up.z);
}
// CPT.cpp:286-292 with glm::mix(x, y, a) = x*(1-a) + y*a  (glm/detail/func_common.inl:140)
inline V3 sample_sky(V3 d) {
  const float t = 0.5f * (d.y + 1.0f);
  const float omt = 1.0f - t;
  return v3(1.0f * omt + 0.5f * t, 1.0f * omt + 0.7f * t, 1.0f * omt + 1.0f * t);
}

// Miss radiance: procedural sky, or the bilinear octahedral environment lookup specified in include/cudapt.h
// (cudapt_upload_environment) -- only + - * / and comparisons, same order as csrc/common.cuh sky_radiance().
inline V3 sky_radiance(const twin_params& p, V3 d) {
  if (p.env_rgb == nullptr || p.env_size <= 0) return sample_sky(d);
  const float s = (fabsf(d.x) + fabsf(d.y)) + fabsf(d.z);
  float px = d.x / s, pz = d.z / s;
  if (d.y < 0.0f) {
    const float fx = 1.0f - fabsf(pz), fz = 1.0f - fabsf(px);
    px = d.x < 0.0f ? -fx : fx;
    pz = d.z < 0.0f ? -fz : fz;
  }
  const uint32_t N = (uint32_t)p.env_size;
  const float n = (float)N, hi = (float)(N - 1u);
  float x = (px * 0.5f + 0.5f) * n - 0.5f, y = (pz * 0.5f + 0.5f) * n - 0.5f;
  x = x < 0.0f ? 0.0f : (x > hi ? hi : x);
  y = y < 0.0f ? 0.0f : (y > hi ? hi : y);
  const uint32_t x0 = (uint32_t)x, y0 = (uint32_t)y;
  const uint32_t x1 = x0 + 1u < N ? x0 + 1u : x0, y1 = y0 + 1u < N ? y0 + 1u : y0;
  const float fx = x - (float)x0, fy = y - (float)y0, gx = 1.0f - fx, gy = 1.0f - fy;
  const float* c00 = p.env_rgb + 3 * ((size_t)y0 * N + x0);
  const float* c10 = p.env_rgb + 3 * ((size_t)y0 * N + x1);
  const float* c01 = p.env_rgb + 3 * ((size_t)y1 * N + x0);
  const float* c11 = p.env_rgb + 3 * ((size_t)y1 * N + x1);
  return v3((c00[0] * gx + c10[0] * fx) * gy + (c01[0] * gx + c11[0] * fx) * fy,
            (c00[1] * gx + c10[1] * fx) * gy + (c01[1] * gx + c11[1] * fx) * fy,
            (c00[2] * gx + c10[2] * fx) * gy + (c01[2] * gx + c11[2] * fx) * fy);
}

inline bool owns_pixel(const twin_params& p, uint32_t x, uint32_t y) {
  if (p.part_world <= 1) return true;
  const uint32_t tx = x / (uint32_t)p.tile_w, ty = y / (uint32_t)p.tile_h;
  const uint32_t tiles_x = ((uint32_t)p.width + (uint32_t)p.tile_w - 1) / (uint32_t)p.tile_w;
  const uint32_t t = ty * tiles_x + tx;
  return (t + ty) % (uint32_t)p.part_world == (uint32_t)p.part_rank;
}

// ------------------------------------------------------------------------------------------------------
// Sphere integrator == the reference (CPT.cpp:197-326) with the Embree stand-in's sphere test
// (oracle/ref/embree_shim.cpp).
// ------------------------------------------------------------------------------------------------------
struct SphereHit {
  int idx;
  float t;
  V3 Ng;
};
inline SphereHit intersect_spheres(const float* sph, int n, V3 o, V3 d, float tnear) {
  SphereHit h{-1, std::numeric_limits<float>::infinity(), v3(0, 0, 0)};
  const float tfar = std::numeric_limits<float>::infinity();
  const float rd2 = 1.0f / ((d.x * d.x + d.y * d.y) + d.z * d.z);
  for (int i = 0; i < n; ++i) {
    const float* c = sph + 4 * i;
    const float c0x = c[0] - o.x, c0y = c[1] - o.y, c0z = c[2] - o.z;
    const float proj = ((c0x * d.x + c0y * d.y) + c0z * d.z) * rd2;
    const float px = c0x - proj * d.x, py = c0y - proj * d.y, pz = c0z - proj * d.z;
    const float l2 = (px * px + py * py) + pz * pz;
    const float r2 = c[3] * c[3];
    if (l2 > r2) continue;
    const float td = sqrtf((r2 - l2) * rd2);
    const float tf = proj - td, tb = proj + td;
    const bool vf = (tnear <= tf) && (tf <= tfar);
    const bool vb = (tnear <= tb) && (tb <= tfar);
    if (!vf && !vb) continue;
    const float t = vf ? tf : tb;
    if (h.idx >= 0 && !(t < h.t)) continue;
    const float sgn = vf ? -td : td;
    h.idx = i;
    h.t = t;
    h.Ng = v3(sgn * d.x - px, sgn * d.y - py, sgn * d.z - pz);
  }
  return h;
}

// CPT.cpp:303-326 as g++/glibc compile it: sqrt/cos/sin resolve to the double
// overloads, abs() to int abs(int).
inline V3 reference_bounce(V3 n, float u1, float u2) {
  const float cosTheta = (float)sqrt((double)u1);
  const float sinTheta = (float)sqrt((double)(1.0f - u1));
  const float phi = 2.0f * 3.14159265358979323846264338327950288f * u2;
  const float x = (float)((double)sinTheta * cos((double)phi));
  const float y = (float)((double)sinTheta * sin((double)phi));
  const float z = cosTheta;
  const V3 up = (std::abs((int)n.z) < 0.999f) ? v3(0, 0, 1) : v3(1, 0, 0);
  const V3 tangent = normalize(cross(up, n));
  const V3 bitangent = cross(n, tangent);
  return (x * tangent + y * bitangent) + z * n;
}

struct Counters {
  uint64_t paths = 0, rays = 0, shades = 0, nodes = 0, tris = 0;
};

// One reference path (CPT.cpp:197-284).  In Philox mode the same loop draws from
// block 0 of (pixel, sample, bounce): {rr, -, u1, u2}.
V3 trace_spheres(const twin_params& p, const float* sph, int n, V3 o, V3 d, uint32_t& rng, uint32_t pixel,
                 uint32_t sample, Counters& cnt, int* first_idx, float* first_t) {
  V3 acc = v3(0, 0, 0), thr = v3(1, 1, 1);
  int bounce = 0;
  while (bounce < p.max_bounces) {
    const SphereHit h = intersect_spheres(sph, n, o, d, p.tnear);
    ++cnt.rays;
    if (bounce == 0 && first_idx) {
      *first_idx = h.idx;
      *first_t = h.t;
    }
    if (h.idx < 0) {
      const V3 sky = sky_radiance(p, d);
      acc = acc + thr * (p.sky_scale == 1.0f ? sky : sky * p.sky_scale);
      break;
    }
    ++cnt.shades;
    o.x += h.t * d.x;
    o.y += h.t * d.y;
    o.z += h.t * d.z;
    const float inv_len = 1.0f / sqrtf(h.Ng.x * h.Ng.x + h.Ng.y * h.Ng.y + h.Ng.z * h.Ng.z);
    const V3 nrm = v3(h.Ng.x * inv_len, h.Ng.y * inv_len, h.Ng.z * inv_len);
    thr = thr * p.albedo;
    ++bounce;
    Rand4 r{};
    if (p.rng_mode == 1) r = philox_block(pixel, sample, (uint32_t)bounce, 0, p.seed);
    if (bounce > p.rr_start) {
      const float q = max3(thr);
      const float u = (p.rng_mode == 1) ? r.u[0] : pcg_next(rng);
      if (u > q) break;
      thr = thr / q;
    }
    float u1, u2;
    if (p.rng_mode == 1) {
      u1 = r.u[2], u2 = r.u[3];
    } else {
      u1 = pcg_next(rng);
      u2 = pcg_next(rng);
    }
    d = reference_bounce(nrm, u1, u2);
    o.x += nrm.x * p.eps;
    o.y += nrm.y * p.eps;
    o.z += nrm.z * p.eps;
  }
  return acc;
}

// ------------------------------------------------------------------------------------------------------
// Mesh integrator (no reference counterpart; spec in DESIGN.md "Mesh integrator numerics")
// ------------------------------------------------------------------------------------------------------
struct Tri {
  V3 v0, e1, e2;
};
struct Node {
  float lo[3], hi[3];
  uint32_t left;   // internal: index of left child (right = left+1); leaf: first index into prim list
  uint32_t count;  // 0 = internal, else number of prims
};
struct Mesh {
  std::vector<Tri> tris;
  std::vector<uint32_t> mat_id;
  std::vector<twin_material> mats;
  std::vector<Node> nodes;
  std::vector<uint32_t> prims;
  std::vector<uint32_t> lights;  // emissive triangle indices, ascending
};

// Moller-Trumbore on the stored (v0, e1, e2) record; two-sided; fixed operation order.  The inside test is
// done on the det-scaled barycentrics (sign of det folded in by flipping sign bits, which is exact), so the
// only division is t = (e2 . q) / det for candidates that are inside.  (DESIGN.md "Mesh integrator numerics")
inline float flip_sign(float x, uint32_t sign_bit) {
  uint32_t u;
  std::memcpy(&u, &x, 4);
  u ^= sign_bit;
  std::memcpy(&x, &u, 4);
  return x;
}
// The products of the triangle test are fused (one rounding per fma), in exactly this nesting; std::fmaf is exact
// whether or not the host has an FMA unit (the Makefile passes -mfma so that it is one instruction).
inline float dot_f(V3 a, V3 b) { return std::fmaf(a.z, b.z, std::fmaf(a.y, b.y, a.x * b.x)); }
inline V3 cross_f(V3 a, V3 b) {
  return v3(std::fmaf(a.y, b.z, -(b.y * a.z)), std::fmaf(a.z, b.x, -(b.z * a.x)), std::fmaf(a.x, b.y, -(b.x * a.y)));
}
inline bool tri_hit(const Tri& T, V3 o, V3 d, float& t, float& bu, float& bv) {
  const V3 pv = cross_f(d, T.e2);
  const float det = dot_f(T.e1, pv);
  if (det == 0.0f) return false;
  uint32_t sb;
  std::memcpy(&sb, &det, 4);
  sb &= 0x80000000u;
  const float adet = flip_sign(det, sb);  // |det|
  const V3 tv = o - T.v0;
  const float u = flip_sign(dot_f(tv, pv), sb);
  if (!(u >= 0.0f && u <= adet)) return false;
  const V3 qv = cross_f(tv, T.e1);
  const float v = flip_sign(dot_f(d, qv), sb);
  if (!(v >= 0.0f && (u + v) <= adet)) return false;
  t = dot_f(T.e2, qv) / det;
  bu = u;
  bv = v;
  return true;
}

struct MeshHit {
  int32_t prim;
  float t;
};
// closest: smallest t in [tnear, inf); ties -> smallest triangle index (traversal-order independent)
inline void consider(const Mesh& m, uint32_t prim, V3 o, V3 d, float tnear, MeshHit& best) {
  float t, u, v;
  if (!tri_hit(m.tris[prim], o, d, t, u, v)) return;
  if (!(t >= tnear)) return;
  if (t < best.t || (t == best.t && (int32_t)prim < best.prim)) {
    best.t = t;
    best.prim = (int32_t)prim;
  }
}

inline bool slab(const Node& nd, V3 o, V3 inv, float tnear, float tmax, float* entry = nullptr) {
  float t0 = (nd.lo[0] - o.x) * inv.x, t1 = (nd.hi[0] - o.x) * inv.x;
  float tmin = std::min(t0, t1), tmx = std::max(t0, t1);
  t0 = (nd.lo[1] - o.y) * inv.y, t1 = (nd.hi[1] - o.y) * inv.y;
  tmin = std::max(tmin, std::min(t0, t1)), tmx = std::min(tmx, std::max(t0, t1));
  t0 = (nd.lo[2] - o.z) * inv.z, t1 = (nd.hi[2] - o.z) * inv.z;
  tmin = std::max(tmin, std::min(t0, t1)), tmx = std::min(tmx, std::max(t0, t1));
  tmin = std::max(tmin, tnear * 0.5f);
  if (entry) *entry = tmin;
  // conservative: widen the exit distance (Ize 2013) so grazing rays are not culled before the exact test
  return tmin <= std::min(tmx * 1.0000005f + 1e-6f, tmax);
}
inline V3 safe_inv(V3 d) {
  auto f = [](float x) {
    if (fabsf(x) < 1e-30f) x = (std::signbit(x) ? -1e-30f : 1e-30f);
    return 1.0f / x;
  };
  return v3(f(d.x), f(d.y), f(d.z));
}

// Front-to-back: both children are tested when their parent is opened and the nearer one is visited first, so that
// the shrinking best.t culls the farther one (the result -- smallest t, ties to the smallest index -- does not depend
// on the order; the work does).
MeshHit closest_bvh(const Mesh& m, V3 o, V3 d, float tnear, Counters* cnt) {
  MeshHit best{-1, std::numeric_limits<float>::infinity()};
  if (m.nodes.empty()) return best;
  const V3 inv = safe_inv(d);
  struct Entry {
    uint32_t node;
    float t;
  };
  Entry stack[128];
  int sp = 0;
  if (cnt) ++cnt->nodes;
  float troot;
  if (!slab(m.nodes[0], o, inv, tnear, best.t, &troot)) return best;
  stack[sp++] = {0, troot};
  while (sp) {
    const Entry e = stack[--sp];
    if (e.t > best.t) continue;  // culled by a hit found since it was pushed
    const Node& nd = m.nodes[e.node];
    if (nd.count) {
      for (uint32_t i = 0; i < nd.count; ++i) {
        if (cnt) ++cnt->tris;
        consider(m, m.prims[nd.left + i], o, d, tnear, best);
      }
      continue;
    }
    float ta, tb;
    if (cnt) cnt->nodes += 2;
    const bool ha = slab(m.nodes[nd.left], o, inv, tnear, best.t, &ta), hb = slab(m.nodes[nd.left + 1], o, inv, tnear, best.t, &tb);
    if (ha && hb) {
      if (ta <= tb) {
        stack[sp++] = {nd.left + 1, tb};
        stack[sp++] = {nd.left, ta};
      } else {
        stack[sp++] = {nd.left, ta};
        stack[sp++] = {nd.left + 1, tb};
      }
    } else if (ha) {
      stack[sp++] = {nd.left, ta};
    } else if (hb) {
      stack[sp++] = {nd.left + 1, tb};
    }
  }
  return best;
}
// any hit with tnear <= t < tmax
bool occluded_bvh(const Mesh& m, V3 o, V3 d, float tnear, float tmax, Counters* cnt) {
  if (m.nodes.empty()) return false;
  const V3 inv = safe_inv(d);
  uint32_t stack[96];
  int sp = 0;
  stack[sp++] = 0;
  while (sp) {
    const Node& nd = m.nodes[stack[--sp]];
    if (cnt) ++cnt->nodes;
    if (!slab(nd, o, inv, tnear, tmax)) continue;
    if (nd.count) {
      for (uint32_t i = 0; i < nd.count; ++i) {
        if (cnt) ++cnt->tris;
        float t, u, v;
        if (tri_hit(m.tris[m.prims[nd.left + i]], o, d, t, u, v) && t >= tnear && t < tmax) return true;
      }
    } else {
      stack[sp++] = nd.left;
      stack[sp++] = nd.left + 1;
    }
  }
  return false;
}

void build_bvh(Mesh& m) {
  const uint32_t nt = (uint32_t)m.tris.size();
  m.prims.resize(nt);
  std::iota(m.prims.begin(), m.prims.end(), 0u);
  std::vector<float> cen(3 * (size_t)nt), blo(3 * (size_t)nt), bhi(3 * (size_t)nt);
  for (uint32_t i = 0; i < nt; ++i) {
    const Tri& T = m.tris[i];
    const V3 a = T.v0, b = T.v0 + T.e1, c = T.v0 + T.e2;
    const float ax[3] = {a.x, a.y, a.z}, bx[3] = {b.x, b.y, b.z}, cx[3] = {c.x, c.y, c.z};
    for (int k = 0; k < 3; ++k) {
      const float lo = std::min(ax[k], std::min(bx[k], cx[k])), hi = std::max(ax[k], std::max(bx[k], cx[k]));
      // v0+e1 is a rounded reconstruction of the vertex: pad by a few ulp so the box stays conservative
      const float pad = 4e-7f * std::max(fabsf(lo), fabsf(hi)) + 1e-30f;
      blo[3 * (size_t)i + k] = lo - pad;
      bhi[3 * (size_t)i + k] = hi + pad;
      cen[3 * (size_t)i + k] = 0.5f * (lo + hi);
    }
  }
  m.nodes.clear();
  m.nodes.reserve(2 * (size_t)nt / 2 + 16);
  struct Job {
    uint32_t node, begin, end;
  };
  std::vector<Job> jobs;
  m.nodes.push_back(Node{});
  jobs.push_back({0, 0, nt});
  while (!jobs.empty()) {
    const Job j = jobs.back();
    jobs.pop_back();
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    float clo[3] = {INFINITY, INFINITY, INFINITY}, chi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (uint32_t i = j.begin; i < j.end; ++i) {
      const size_t p = m.prims[i];
      for (int k = 0; k < 3; ++k) {
        lo[k] = std::min(lo[k], blo[3 * p + k]);
        hi[k] = std::max(hi[k], bhi[3 * p + k]);
        clo[k] = std::min(clo[k], cen[3 * p + k]);
        chi[k] = std::max(chi[k], cen[3 * p + k]);
      }
    }
    Node nd;
    for (int k = 0; k < 3; ++k) nd.lo[k] = lo[k], nd.hi[k] = hi[k];
    const uint32_t cnt = j.end - j.begin;
    if (cnt <= 4) {
      nd.left = j.begin;
      nd.count = cnt;
      m.nodes[j.node] = nd;
      continue;
    }
    // Binned surface-area heuristic (16 bins per axis on the centroid bounds, all three axes); the median of the
    // widest axis when no split plane separates the centroids.  The tree only shapes the WORK of a query -- any
    // conservative tree returns the same closest hit -- so this is about giving the CPU arm a fair structure.
    constexpr int kBins = 16;
    auto half_area = [](const float* l, const float* h) {
      const float dx = h[0] - l[0], dy = h[1] - l[1], dz = h[2] - l[2];
      return dx * dy + dy * dz + dz * dx;
    };
    float best_cost = INFINITY, best_pos = 0.f;
    int best_axis = -1;
    for (int axis = 0; axis < 3; ++axis) {
      const float ext = chi[axis] - clo[axis];
      if (!(ext > 0.f)) continue;
      struct Bin {
        float lo[3], hi[3];
        uint32_t n;
      } bins[kBins];
      for (Bin& b : bins) {
        for (int k = 0; k < 3; ++k) b.lo[k] = INFINITY, b.hi[k] = -INFINITY;
        b.n = 0;
      }
      const float scale = (float)kBins / ext;
      for (uint32_t i = j.begin; i < j.end; ++i) {
        const size_t p = m.prims[i];
        int b = (int)((cen[3 * p + axis] - clo[axis]) * scale);
        b = b < 0 ? 0 : (b >= kBins ? kBins - 1 : b);
        for (int k = 0; k < 3; ++k) {
          bins[b].lo[k] = std::min(bins[b].lo[k], blo[3 * p + k]);
          bins[b].hi[k] = std::max(bins[b].hi[k], bhi[3 * p + k]);
        }
        ++bins[b].n;
      }
      float right_area[kBins];
      uint32_t right_n[kBins];
      float rl[3] = {INFINITY, INFINITY, INFINITY}, rh[3] = {-INFINITY, -INFINITY, -INFINITY};
      uint32_t rn = 0;
      for (int b = kBins - 1; b > 0; --b) {
        for (int k = 0; k < 3; ++k) rl[k] = std::min(rl[k], bins[b].lo[k]), rh[k] = std::max(rh[k], bins[b].hi[k]);
        rn += bins[b].n;
        right_area[b] = rn ? half_area(rl, rh) : 0.f;
        right_n[b] = rn;
      }
      float ll[3] = {INFINITY, INFINITY, INFINITY}, lh[3] = {-INFINITY, -INFINITY, -INFINITY};
      uint32_t ln = 0;
      for (int b = 0; b + 1 < kBins; ++b) {
        for (int k = 0; k < 3; ++k) ll[k] = std::min(ll[k], bins[b].lo[k]), lh[k] = std::max(lh[k], bins[b].hi[k]);
        ln += bins[b].n;
        if (ln == 0 || right_n[b + 1] == 0) continue;
        const float cost = half_area(ll, lh) * (float)ln + right_area[b + 1] * (float)right_n[b + 1];
        if (cost < best_cost) best_cost = cost, best_axis = axis, best_pos = clo[axis] + (float)(b + 1) / scale;
      }
    }
    uint32_t mid = j.begin + cnt / 2;
    if (best_axis >= 0) {
      const int axis = best_axis;
      const float scale = (float)kBins / (chi[axis] - clo[axis]);
      const int split_bin = (int)((best_pos - clo[axis]) * scale + 0.5f);
      auto it = std::partition(m.prims.begin() + j.begin, m.prims.begin() + j.end, [&](uint32_t a) {
        int b = (int)((cen[3 * (size_t)a + axis] - clo[axis]) * scale);
        b = b < 0 ? 0 : (b >= kBins ? kBins - 1 : b);
        return b < split_bin;
      });
      mid = (uint32_t)(it - m.prims.begin());
    }
    if (best_axis < 0 || mid == j.begin || mid == j.end) {  // coincident centroids (or a degenerate partition): median
      int axis = 0;
      if (chi[1] - clo[1] > chi[axis] - clo[axis]) axis = 1;
      if (chi[2] - clo[2] > chi[axis] - clo[axis]) axis = 2;
      mid = j.begin + cnt / 2;
      std::nth_element(m.prims.begin() + j.begin, m.prims.begin() + mid, m.prims.begin() + j.end,
                       [&](uint32_t a, uint32_t b) {
                         const float ca = cen[3 * (size_t)a + axis], cb = cen[3 * (size_t)b + axis];
                         return ca < cb || (ca == cb && a < b);
                       });
    }
    nd.left = (uint32_t)m.nodes.size();
    nd.count = 0;
    m.nodes[j.node] = nd;
    m.nodes.push_back(Node{});
    m.nodes.push_back(Node{});
    jobs.push_back({nd.left, j.begin, mid});
    jobs.push_back({nd.left + 1, mid, j.end});
  }
}

// Orthonormal basis around unit n (Duff et al. 2017, "Building an Orthonormal Basis, Revisited").
inline void onb(V3 n, V3& t, V3& b) {
  const float sign = copysignf(1.0f, n.z);
  const float a = -1.0f / (sign + n.z);
  const float bb = (n.x * n.y) * a;
  t = v3(1.0f + (sign * (n.x * n.x)) * a, sign * bb, -sign * n.x);
  b = v3(bb, sign + (n.y * n.y) * a, -n.y);
}
inline V3 to_world(V3 t, V3 b, V3 n, float x, float y, float z) { return (x * t + y * b) + z * n; }
inline float pow5(float x) {
  const float x2 = x * x;
  return (x2 * x2) * x;
}
inline float ggx_g1(float a2, float x) { return (2.0f * x) / (x + sqrtf(a2 + (1.0f - a2) * (x * x))); }

constexpr float kInvPi = 0.318309886f;
constexpr float kF0 = 0.04f;

V3 trace_mesh(const twin_params& p, const Mesh& m, V3 o, V3 d, uint32_t pixel, uint32_t sample, Counters& cnt,
              int32_t* first_prim, float* first_t) {
  V3 L = v3(0, 0, 0), thr = v3(1, 1, 1);
  bool specular = true;
  const bool use_nee = p.nee && !m.lights.empty();
  int bounce = 0;
  while (bounce < p.max_bounces) {
    const MeshHit h = closest_bvh(m, o, d, p.tnear, &cnt);
    ++cnt.rays;
    if (bounce == 0 && first_prim) {
      *first_prim = h.prim;
      *first_t = h.t;
    }
    if (h.prim < 0) {
      const V3 sky = sky_radiance(p, d);
      L = L + thr * (sky * p.sky_scale);
      break;
    }
    ++cnt.shades;
    const Tri& T = m.tris[h.prim];
    const twin_material& mat = m.mats[m.mat_id[h.prim]];
    const V3 albedo = v3(mat.albedo[0], mat.albedo[1], mat.albedo[2]);
    o = v3(o.x + h.t * d.x, o.y + h.t * d.y, o.z + h.t * d.z);
    const V3 Ng = normalize(cross(T.e1, T.e2));
    const bool front = dot(Ng, d) < 0.0f;
    const V3 n = front ? Ng : -Ng;
    const V3 Le = v3(mat.emission[0], mat.emission[1], mat.emission[2]);
    if (front && (Le.x > 0.0f || Le.y > 0.0f || Le.z > 0.0f) && (specular || !use_nee)) L = L + thr * Le;
    ++bounce;
    const Rand4 r = philox_block(pixel, sample, (uint32_t)bounce, 0, p.seed);  // {rr, lobe, u1, u2}
    V3 wi;
    V3 weight;
    float side = 1.0f;  // offset along +n or -n
    V3 tt, bb;
    switch (mat.type) {
      default:
      case 0: {  // Lambert: cosine-weighted sample, weight = albedo
        if (use_nee) {
          const Rand4 q = philox_block(pixel, sample, (uint32_t)bounce, 1, p.seed);  // {light, v1, v2, -}
          const uint32_t nl = (uint32_t)m.lights.size();
          uint32_t k = (uint32_t)(q.u[0] * (float)nl);
          if (k >= nl) k = nl - 1;
          const uint32_t lp = m.lights[k];
          const Tri& Lt = m.tris[lp];
          const twin_material& lm = m.mats[m.mat_id[lp]];
          const V3 lc = cross(Lt.e1, Lt.e2);
          const float ll = sqrtf(dot(lc, lc));
          const V3 Nl = lc * (1.0f / ll);
          const float area = 0.5f * ll;
          const float su = sqrtf(q.u[1]);
          const float b1 = su * (1.0f - q.u[2]), b2 = su * q.u[2];
          const V3 pl = (Lt.v0 + Lt.e1 * b1) + Lt.e2 * b2;
          const V3 ps = v3(o.x + n.x * p.eps, o.y + n.y * p.eps, o.z + n.z * p.eps);
          V3 wl = pl - ps;
          const float dist2 = dot(wl, wl);
          const float dist = sqrtf(dist2);
          wl = wl * (1.0f / dist);
          const float cos_s = dot(n, wl);
          const float cos_l = -dot(Nl, wl);
          if (cos_s > 0.0f && cos_l > 0.0f) {
            ++cnt.rays;
            if (!occluded_bvh(m, ps, wl, p.tnear, dist * 0.999f, &cnt)) {
              const float g = ((cos_s * cos_l) / dist2) * (area * (float)nl);
              const V3 le = v3(lm.emission[0], lm.emission[1], lm.emission[2]);
              L = L + ((thr * albedo) * le) * (g * kInvPi);
            }
          }
        }
        float c, s;
        sincos2pi(r.u[3], c, s);
        const float ct = sqrtf(r.u[2]), st = sqrtf(1.0f - r.u[2]);
        onb(n, tt, bb);
        wi = normalize(to_world(tt, bb, n, st * c, st * s, ct));
        weight = albedo;
        specular = false;
        break;
      }
      case 1: {  // mirror
        const float cosi = -dot(d, n);
        wi = normalize(d + n * (2.0f * cosi));
        weight = albedo;
        specular = true;
        break;
      }
      case 2: {  // dielectric, Schlick Fresnel, stochastic reflect/refract
        const float ior = mat.param;
        const float eta = front ? 1.0f / ior : ior;
        const float cosi = -dot(d, n);
        const float sin2t = (eta * eta) * (1.0f - cosi * cosi);
        const float r0s = (1.0f - ior) / (1.0f + ior);
        const float r0 = r0s * r0s;
        bool reflect = true;
        float cost = 0.0f;
        if (sin2t < 1.0f) {
          cost = sqrtf(1.0f - sin2t);
          const float cx = front ? cosi : cost;
          const float F = r0 + (1.0f - r0) * pow5(1.0f - cx);
          reflect = r.u[1] < F;
        }
        if (reflect) {
          wi = normalize(d + n * (2.0f * cosi));
        } else {
          wi = normalize(d * eta + n * (eta * cosi - cost));
          side = -1.0f;
        }
        weight = albedo;
        specular = true;
        break;
      }
      case 3: {  // GGX-coated diffuse: one-sample lobe selection with probability 1/2 each
        const float alpha = mat.param;
        const float a2 = alpha * alpha;
        const V3 wo = -d;
        const float nov = std::max(dot(n, wo), 1e-6f);
        onb(n, tt, bb);
        float c, s;
        sincos2pi(r.u[3], c, s);
        if (r.u[1] < 0.5f) {
          const float cos2h = (1.0f - r.u[2]) / (1.0f + (a2 - 1.0f) * r.u[2]);
          const float ch = sqrtf(cos2h), sh = sqrtf(std::max(0.0f, 1.0f - cos2h));
          const V3 hh = normalize(to_world(tt, bb, n, sh * c, sh * s, ch));
          const float voh = dot(wo, hh);
          if (!(voh > 0.0f)) return L;
          wi = normalize(hh * (2.0f * voh) - wo);
          const float noi = dot(n, wi);
          if (!(noi > 0.0f)) return L;
          const float noh = std::max(dot(n, hh), 1e-6f);
          const float F = kF0 + (1.0f - kF0) * pow5(1.0f - voh);
          const float G = ggx_g1(a2, nov) * ggx_g1(a2, noi);
          const float w = (((F * G) * voh) / (nov * noh)) * 2.0f;
          weight = v3(w, w, w);
        } else {
          const float ct = sqrtf(r.u[2]), st = sqrtf(1.0f - r.u[2]);
          wi = normalize(to_world(tt, bb, n, st * c, st * s, ct));
          const float Fd = kF0 + (1.0f - kF0) * pow5(1.0f - nov);
          weight = albedo * ((1.0f - Fd) * 2.0f);
        }
        specular = true;  // no NEE on this lobe mix: emission is picked up by BSDF sampling
        break;
      }
    }
    thr = thr * weight;
    if (bounce > p.rr_start) {
      const float q = std::min(max3(thr), 1.0f);
      if (!(q > 0.0f) || r.u[0] > q) break;
      thr = thr / q;
    }
    d = wi;
    const float off = side * p.eps;
    o = v3(o.x + n.x * off, o.y + n.y * off, o.z + n.z * off);
  }
  return L;
}

// Rows are independent (every pixel owns its RNG stream, CPT.cpp:61), so they are handed out to
// std::threads in chunks of 4 from an atomic counter.  (No OpenMP: libgomp is absent from this image.)
template <class F>
void for_rows(const twin_params& p, F&& fn) {
  const int H = p.height;
  int nthreads = p.threads > 0 ? p.threads : (int)std::thread::hardware_concurrency();
  if (nthreads < 1) nthreads = 1;
  if (nthreads == 1) {
    for (int y = 0; y < H; ++y) fn((uint32_t)y);
    return;
  }
  std::atomic<int> next{0};
  auto work = [&]() {
    for (;;) {
      const int y0 = next.fetch_add(4);
      if (y0 >= H) return;
      for (int y = y0; y < std::min(y0 + 4, H); ++y) fn((uint32_t)y);
    }
  };
  std::vector<std::thread> pool;
  for (int i = 1; i < nthreads; ++i) pool.emplace_back(work);
  work();
  for (auto& t : pool) t.join();
}

}  // namespace

extern "C" {

void twin_default_params(twin_params* p) {
  std::memset(p, 0, sizeof(*p));
  p->width = 512;
  p->height = 512;  // Types.h:77-78
  p->rng_mode = 0;
  p->max_bounces = 4;
  p->rr_start = 2;
  p->tnear = 0.001f;
  p->eps = 1e-4f;
  p->albedo = 0.7f;
  p->sky_scale = 1.0f;
  p->cam_right[0] = 1.0f;
  p->cam_up[1] = 1.0f;
  p->cam_forward[2] = 1.0f;
  p->cam_tan_half_fov = 1.0f;
  p->cam_focus_distance = 1.0f;
  p->part_world = 1;
  p->tile_w = 32;
  p->tile_h = 32;
}

int twin_render_spheres(const twin_params* pp, const float* sph, int32_t n, uint32_t first_frame, uint32_t n_frames,
                        uint32_t frame_stride, float* accum, int32_t* primary_idx, float* primary_t,
                        twin_counters* counters) {
  const twin_params p = *pp;
  const uint32_t W = (uint32_t)p.width, H = (uint32_t)p.height;
  const Camera cam = camera_of(p);
  std::vector<Counters> row_cnt(H);
  for (uint32_t k = 0; k < n_frames; ++k) {
    const uint32_t frame = first_frame + k * frame_stride;
    for_rows(p, [&](uint32_t y) {
      Counters& cnt = row_cnt[y];
      for (uint32_t x = 0; x < W; ++x) {
        if (!owns_pixel(p, x, y)) continue;
        const uint32_t pixel = y * W + x;
        uint32_t rng = pcg_seed(x, y, W, frame);
        float jx = 0.0f, jy = 0.0f;
        Rand4 j{};
        if (p.rng_mode == 1 && (p.jitter || cam.lens_radius > 0.0f)) j = philox_block(pixel, frame, 0, 2, p.seed);
        if (p.rng_mode == 1 && p.jitter) jx = j.u[0], jy = j.u[1];
        V3 o, d;
        primary_ray(cam, x, y, W, H, jx, jy, j, pixel, frame, p.seed, &o, &d);
        int fi = -1;
        float ft = 0.0f;
        const bool cap = (k == 0) && primary_idx;
        const V3 c = trace_spheres(p, sph, n, o, d, rng, pixel, frame, cnt, cap ? &fi : nullptr, &ft);
        ++cnt.paths;
        if (cap) {
          primary_idx[pixel] = fi;
          if (primary_t) primary_t[pixel] = ft;
        }
        float* a = accum + 4 * (size_t)pixel;  // CPT.cpp:77-80
        a[0] += c.x;
        a[1] += c.y;
        a[2] += c.z;
        a[3] += 1.0f;
      }
    });
  }
  if (counters) {
    twin_counters t{};
    for (const Counters& c : row_cnt) t.paths += c.paths, t.rays += c.rays, t.shade_invocations += c.shades;
    *counters = t;
  }
  return 0;
}

void* twin_mesh_create(const float* vertices, uint32_t nv, const uint32_t* indices, uint32_t nt,
                       const uint32_t* material_id, const twin_material* materials, uint32_t n_materials) {
  (void)nv;
  auto* m = new Mesh();
  m->tris.resize(nt);
  m->mat_id.assign(nt, 0u);
  for (uint32_t i = 0; i < nt; ++i) {
    const float* a = vertices + 3 * (size_t)indices[3 * (size_t)i + 0];
    const float* b = vertices + 3 * (size_t)indices[3 * (size_t)i + 1];
    const float* c = vertices + 3 * (size_t)indices[3 * (size_t)i + 2];
    Tri T;
    T.v0 = v3(a[0], a[1], a[2]);
    T.e1 = v3(b[0] - a[0], b[1] - a[1], b[2] - a[2]);
    T.e2 = v3(c[0] - a[0], c[1] - a[1], c[2] - a[2]);
    m->tris[i] = T;
    if (material_id) m->mat_id[i] = material_id[i];
  }
  if (n_materials) {
    m->mats.assign(materials, materials + n_materials);
  } else {
    twin_material d{};
    d.albedo[0] = d.albedo[1] = d.albedo[2] = 0.7f;
    m->mats.push_back(d);
  }
  for (uint32_t i = 0; i < nt; ++i) {
    if (m->mat_id[i] >= m->mats.size()) m->mat_id[i] = 0;
    const twin_material& mm = m->mats[m->mat_id[i]];
    if (mm.emission[0] > 0.0f || mm.emission[1] > 0.0f || mm.emission[2] > 0.0f) m->lights.push_back(i);
  }
  build_bvh(*m);
  return m;
}
void twin_mesh_destroy(void* mesh) { delete static_cast<Mesh*>(mesh); }

int twin_render_mesh(const twin_params* pp, void* mesh, uint32_t first_frame, uint32_t n_frames, uint32_t frame_stride,
                     float* accum, int32_t* primary_idx, float* primary_t, twin_counters* counters) {
  const twin_params p = *pp;
  const Mesh& m = *static_cast<Mesh*>(mesh);
  const uint32_t W = (uint32_t)p.width, H = (uint32_t)p.height;
  const Camera cam = camera_of(p);
  std::vector<Counters> row_cnt(H);
  for (uint32_t k = 0; k < n_frames; ++k) {
    const uint32_t frame = first_frame + k * frame_stride;
    for_rows(p, [&](uint32_t y) {
      Counters& cnt = row_cnt[y];
      for (uint32_t x = 0; x < W; ++x) {
        if (!owns_pixel(p, x, y)) continue;
        const uint32_t pixel = y * W + x;
        float jx = 0.0f, jy = 0.0f;
        Rand4 j{};
        if (p.jitter || cam.lens_radius > 0.0f) j = philox_block(pixel, frame, 0, 2, p.seed);
        if (p.jitter) jx = j.u[0], jy = j.u[1];
        V3 o, d;
        primary_ray(cam, x, y, W, H, jx, jy, j, pixel, frame, p.seed, &o, &d);
        int32_t fi = -1;
        float ft = 0.0f;
        const bool cap = (k == 0) && primary_idx;
        const V3 c = trace_mesh(p, m, o, d, pixel, frame, cnt, cap ? &fi : nullptr, &ft);
        ++cnt.paths;
        if (cap) {
          primary_idx[pixel] = fi;
          if (primary_t) primary_t[pixel] = ft;
        }
        float* a = accum + 4 * (size_t)pixel;
        a[0] += c.x;
        a[1] += c.y;
        a[2] += c.z;
        a[3] += 1.0f;
      }
    });
  }
  if (counters) {
    twin_counters t{};
    for (const Counters& c : row_cnt)
      t.paths += c.paths, t.rays += c.rays, t.shade_invocations += c.shades, t.nodes_visited += c.nodes,
          t.tris_tested += c.tris;
    *counters = t;
  }
  return 0;
}

int twin_mesh_closest_bruteforce(void* mesh, const float* org, const float* dir, float tnear, int32_t* prim, float* t) {
  const Mesh& m = *static_cast<Mesh*>(mesh);
  MeshHit best{-1, std::numeric_limits<float>::infinity()};
  const V3 o = v3(org[0], org[1], org[2]), d = v3(dir[0], dir[1], dir[2]);
  for (uint32_t i = 0; i < m.tris.size(); ++i) consider(m, i, o, d, tnear, best);
  *prim = best.prim;
  *t = best.t;
  return 0;
}
int twin_mesh_closest(void* mesh, const float* org, const float* dir, float tnear, int32_t* prim, float* t) {
  const Mesh& m = *static_cast<Mesh*>(mesh);
  const MeshHit best = closest_bvh(m, v3(org[0], org[1], org[2]), v3(dir[0], dir[1], dir[2]), tnear, nullptr);
  *prim = best.prim;
  *t = best.t;
  return 0;
}

// CPT.cpp:87-117 + Color.h:7-10
void twin_resolve_rgba8(const float* accum, uint32_t n_pixels, uint32_t frame_count, uint32_t* rgba) {
  for (uint32_t i = 0; i < n_pixels; ++i) {
    float c[4];
    for (int k = 0; k < 4; ++k) {
      float v = accum[4 * (size_t)i + k] / (float)frame_count;
      v = std::clamp(v, 0.0f, 1.0f);
      c[k] = v;
    }
    const uint8_t r = (uint8_t)(c[0] * 255.0f), g = (uint8_t)(c[1] * 255.0f), b = (uint8_t)(c[2] * 255.0f),
                  a = (uint8_t)(c[3] * 255.0f);
    rgba[i] = ((uint32_t)r << 24) | ((uint32_t)g << 16) | ((uint32_t)b << 8) | (uint32_t)a;
  }
}

uint32_t twin_pcg_seed(uint32_t x, uint32_t y, uint32_t width, uint32_t frame) { return pcg_seed(x, y, width, frame); }
float twin_pcg_next(uint32_t* state, uint32_t* raw_word) { return pcg_next(*state, raw_word); }
void twin_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) { philox4x32_10(ctr, key, out); }
void twin_sincos2pi(float u, float* c, float* s) { sincos2pi(u, *c, *s); }
}
