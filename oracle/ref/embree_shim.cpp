// TEST INFRASTRUCTURE (oracle, Tier 1) -- never linked into the product.
//
// Minimal stand-in for the 12 Embree-4 entry points the reference backend uses
// (see embree4/rtcore.h in this directory for the call-site list).  Implements
// closest-hit for RTC_GEOMETRY_TYPE_SPHERE_POINT only, one sphere per geometry
// as the reference creates them (CPUPathTracer.cpp:374-391).
//
// Sphere test = the projection form documented for Embree 4.x sphere points
// (SURVEY.md Appendix B.2), evaluated in scalar FP32 in a fixed operation
// order, compiled with -ffp-contract=off.  That order is the normative one the
// CPU twin (oracle/twin) and the CUDA kernels reproduce bit-for-bit:
//
//   rd2    = 1 / ((dx*dx + dy*dy) + dz*dz)
//   c0     = C - o
//   proj   = ((c0x*dx + c0y*dy) + c0z*dz) * rd2
//   perp   = c0 - proj*d
//   l2     = (px*px + py*py) + pz*pz ;  r2 = r*r ;  miss if l2 > r2
//   td     = sqrtf((r2 - l2) * rd2)
//   t      = proj - td if tnear <= proj-td <= tfar, else proj + td if valid
//   Ng     = (front ? -td : +td) * d - perp          (== P - C, outward)
//
// A candidate replaces the current hit only if its t is strictly smaller, so
// ties keep the geometry attached first.  PARITY UNPINNED (no Embree here).
#include "embree4/rtcore.h"

#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <vector>

struct RTCDeviceTy { int refs = 1; };
struct RTCGeometryTy {
  int refs = 1;
  std::vector<float> vertex;  // float4 per item
  size_t items = 0;
  void* user = nullptr;
};
struct RTCSceneTy {
  int refs = 1;
  std::vector<RTCGeometryTy*> geoms;
  // committed, flattened copy
  std::vector<float> sph;  // x y z r
  std::vector<unsigned> geom_of;
  std::vector<unsigned> prim_of;
};

// ---- instrumentation; read by ref_harness.cpp ----
extern "C" {
uint64_t g_shim_ray_count = 0;
// When non-null, every query whose origin is exactly (0,0,0) is a primary ray
// (the reference casts them in raster order); its result is appended here.
void** g_shim_primary_user = nullptr;  // user-data pointer of the hit geometry (nullptr = miss)
float* g_shim_primary_t = nullptr;
uint64_t g_shim_primary_cap = 0;
uint64_t g_shim_primary_n = 0;
}

extern "C" RTCDevice rtcNewDevice(const char*) { return new RTCDeviceTy(); }
extern "C" void rtcReleaseDevice(RTCDevice d) {
  if (d && --d->refs == 0) delete d;
}
extern "C" RTCScene rtcNewScene(RTCDevice) { return new RTCSceneTy(); }
extern "C" void rtcReleaseScene(RTCScene s) {
  if (!s || --s->refs != 0) return;
  for (auto* g : s->geoms) rtcReleaseGeometry(g);
  delete s;
}
extern "C" RTCGeometry rtcNewGeometry(RTCDevice, enum RTCGeometryType type) {
  if (type != RTC_GEOMETRY_TYPE_SPHERE_POINT) std::abort();
  return new RTCGeometryTy();
}
extern "C" void* rtcSetNewGeometryBuffer(RTCGeometry g, enum RTCBufferType type, unsigned slot, enum RTCFormat fmt,
                                         size_t stride, size_t count) {
  if (type != RTC_BUFFER_TYPE_VERTEX || slot != 0 || fmt != RTC_FORMAT_FLOAT4 || stride != 16) std::abort();
  g->vertex.assign(4 * count, 0.0f);
  g->items = count;
  return g->vertex.data();
}
extern "C" void rtcSetGeometryUserData(RTCGeometry g, void* p) { g->user = p; }
extern "C" void rtcCommitGeometry(RTCGeometry) {}
extern "C" unsigned rtcAttachGeometry(RTCScene s, RTCGeometry g) {
  g->refs++;
  s->geoms.push_back(g);
  return (unsigned)(s->geoms.size() - 1);
}
extern "C" void rtcReleaseGeometry(RTCGeometry g) {
  if (g && --g->refs == 0) delete g;
}
extern "C" void rtcCommitScene(RTCScene s) {
  s->sph.clear();
  s->geom_of.clear();
  s->prim_of.clear();
  for (unsigned gi = 0; gi < s->geoms.size(); ++gi) {
    const RTCGeometryTy* g = s->geoms[gi];
    for (size_t i = 0; i < g->items; ++i) {
      for (int k = 0; k < 4; ++k) s->sph.push_back(g->vertex[4 * i + k]);
      s->geom_of.push_back(gi);
      s->prim_of.push_back((unsigned)i);
    }
  }
}

extern "C" void rtcIntersect1(RTCScene s, struct RTCRayHit* rh, struct RTCIntersectArguments*) {
  ++g_shim_ray_count;
  const float ox = rh->ray.org_x, oy = rh->ray.org_y, oz = rh->ray.org_z;
  const float dx = rh->ray.dir_x, dy = rh->ray.dir_y, dz = rh->ray.dir_z;
  const float tnear = rh->ray.tnear, tfar = rh->ray.tfar;
  const float rd2 = 1.0f / ((dx * dx + dy * dy) + dz * dz);
  bool have = false;
  float best_t = tfar;
  const size_t n = s->geom_of.size();
  for (size_t i = 0; i < n; ++i) {
    const float* c = &s->sph[4 * i];
    const float c0x = c[0] - ox, c0y = c[1] - oy, c0z = c[2] - oz;
    const float proj = ((c0x * dx + c0y * dy) + c0z * dz) * rd2;
    const float px = c0x - proj * dx, py = c0y - proj * dy, pz = c0z - proj * dz;
    const float l2 = (px * px + py * py) + pz * pz;
    const float r2 = c[3] * c[3];
    if (l2 > r2) continue;
    const float td = sqrtf((r2 - l2) * rd2);
    const float tf = proj - td, tb = proj + td;
    const bool vf = (tnear <= tf) && (tf <= tfar);
    const bool vb = (tnear <= tb) && (tb <= tfar);
    if (!vf && !vb) continue;
    const float t = vf ? tf : tb;
    if (have && !(t < best_t)) continue;
    have = true;
    best_t = t;
    const float sgn = vf ? -td : td;
    rh->hit.Ng_x = sgn * dx - px;
    rh->hit.Ng_y = sgn * dy - py;
    rh->hit.Ng_z = sgn * dz - pz;
    rh->hit.u = 0.0f;
    rh->hit.v = 0.0f;
    rh->hit.primID = s->prim_of[i];
    rh->hit.geomID = s->geom_of[i];
    rh->hit.instID[0] = RTC_INVALID_GEOMETRY_ID;
    rh->ray.tfar = t;
  }
  if (g_shim_primary_user && ox == 0.0f && oy == 0.0f && oz == 0.0f && g_shim_primary_n < g_shim_primary_cap) {
    g_shim_primary_user[g_shim_primary_n] = have ? s->geoms[rh->hit.geomID]->user : nullptr;
    g_shim_primary_t[g_shim_primary_n] = have ? best_t : INFINITY;
    ++g_shim_primary_n;
  }
}
