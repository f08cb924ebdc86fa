/* TEST INFRASTRUCTURE (oracle, Tier 1) -- never linked into the product.
 *
 * Stand-in for Intel Embree 4.4.0's <embree4/rtcore.h>, restricted to the 12
 * rtc* entry points the reference backend calls (call sites:
 * libs/render/src/engines/pathtracer/backends/cpu/CPUPathTracer.cpp:167,171,
 * 177,185,188,227,374,378,388,389,390,391).  Embree is an un-vendored binary
 * dependency of the reference (libs/render/CMakeLists.txt:21-29, pinned 4.4.0)
 * and is absent from this image, so the published sphere-point semantics are
 * restated in embree_shim.cpp.  PARITY UNPINNED: no Embree source or golden
 * vector exists under /root/reference to check the stand-in against.
 *
 * Only <stddef.h>, <sys/types.h>, <stdbool.h> are included (like the real
 * header) so that the reference TU's unqualified sqrt/cos/sin/abs calls
 * (CPUPathTracer.cpp:310-320) resolve exactly as they would against Embree.
 */
#pragma once
#include <stddef.h>
#include <sys/types.h>
#include <stdbool.h>

#if defined(__cplusplus)
#define RTC_API_EXTERN_C extern "C"
#define RTC_DEFAULT_NULL = NULL
#else
#define RTC_API_EXTERN_C
#define RTC_DEFAULT_NULL
#endif
#define RTC_ALIGN(n) __attribute__((aligned(n)))
#define RTC_MAX_INSTANCE_LEVEL_COUNT 1
#define RTC_INVALID_GEOMETRY_ID ((unsigned int)-1)

typedef struct RTCDeviceTy* RTCDevice;
typedef struct RTCSceneTy* RTCScene;
typedef struct RTCGeometryTy* RTCGeometry;

enum RTCGeometryType { RTC_GEOMETRY_TYPE_TRIANGLE = 0, RTC_GEOMETRY_TYPE_SPHERE_POINT = 50 };
enum RTCBufferType { RTC_BUFFER_TYPE_INDEX = 0, RTC_BUFFER_TYPE_VERTEX = 1 };
enum RTCFormat { RTC_FORMAT_FLOAT3 = 0x9003, RTC_FORMAT_FLOAT4 = 0x9004 };

struct RTC_ALIGN(16) RTCRay {
  float org_x, org_y, org_z, tnear;
  float dir_x, dir_y, dir_z, time;
  float tfar;
  unsigned int mask, id, flags;
};
struct RTC_ALIGN(16) RTCHit {
  float Ng_x, Ng_y, Ng_z;
  float u, v;
  unsigned int primID, geomID;
  unsigned int instID[RTC_MAX_INSTANCE_LEVEL_COUNT];
};
struct RTC_ALIGN(16) RTCRayHit {
  struct RTCRay ray;
  struct RTCHit hit;
};
struct RTCIntersectArguments;

RTC_API_EXTERN_C RTCDevice rtcNewDevice(const char* config);
RTC_API_EXTERN_C void rtcReleaseDevice(RTCDevice device);
RTC_API_EXTERN_C RTCScene rtcNewScene(RTCDevice device);
RTC_API_EXTERN_C void rtcReleaseScene(RTCScene scene);
RTC_API_EXTERN_C void rtcCommitScene(RTCScene scene);
RTC_API_EXTERN_C RTCGeometry rtcNewGeometry(RTCDevice device, enum RTCGeometryType type);
RTC_API_EXTERN_C void* rtcSetNewGeometryBuffer(RTCGeometry geometry, enum RTCBufferType type, unsigned int slot,
                                               enum RTCFormat format, size_t byteStride, size_t itemCount);
RTC_API_EXTERN_C void rtcSetGeometryUserData(RTCGeometry geometry, void* ptr);
RTC_API_EXTERN_C void rtcCommitGeometry(RTCGeometry geometry);
RTC_API_EXTERN_C unsigned int rtcAttachGeometry(RTCScene scene, RTCGeometry geometry);
RTC_API_EXTERN_C void rtcReleaseGeometry(RTCGeometry geometry);
RTC_API_EXTERN_C void rtcIntersect1(RTCScene scene, struct RTCRayHit* rayhit,
                                    struct RTCIntersectArguments* args RTC_DEFAULT_NULL);
