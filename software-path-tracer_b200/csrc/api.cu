// C ABI of libcudapt.so (include/cudapt.h): context, device buffers, launches.  Host C++ only; the
// kernels live in spheres.cu / wavefront.cu / trace_bvh.cu.  "CPT.cpp" = the reference's
// libs/render/src/engines/pathtracer/backends/cpu/CPUPathTracer.cpp.
#include "../../include/cudapt.h"

#include <chrono>
#include <cstdarg>
#include <cstdio>
#include <algorithm>
#include <cmath>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "env_load.h"
#include "kernels.h"
#include "mesh_host.h"

using namespace cudapt;

struct cudapt_ctx {
  cudapt_config cfg{};
  cudapt_camera cam{};
  cudaStream_t stream = nullptr;
  // CUDA-event pairs around the cudapt_render() calls whose time has not been collected yet: a small ring, so that
  // renders can be queued back to back without the host waiting for the previous one (cudapt_render never blocks
  // unless kTimingRing renders are in flight)
  static constexpr int kTimingRing = 4;
  cudaEvent_t ev0[kTimingRing] = {}, ev1[kTimingRing] = {};
  int timing_head = 0, timing_count = 0;  // oldest uncollected pair, number of uncollected pairs
  std::string err;
  int sm_count = 0;

  uint32_t width = 0, height = 0;
  float4* d_accum = nullptr;
  uint32_t* d_rgba = nullptr;
  uint32_t* d_aov_prim = nullptr;
  float* d_aov_t = nullptr;
  unsigned long long* d_counters = nullptr;
  uint32_t* h_pinned_rgba = nullptr;  // pinned staging for the 8-bit read-back
  size_t pinned_pixels = 0;
  // asynchronous read-back (cudapt_resolve_rgba8_async / cudapt_download_rgba8_async): the copy runs on its own
  // stream, after ev_img_ready (image complete on `stream`), and signals ev_dl_done
  cudaStream_t dl_stream = nullptr;
  cudaEvent_t ev_img_ready = nullptr, ev_dl_done = nullptr;
  bool dl_pending = false;

  // scene
  enum SceneKind { kNone, kSpheres, kMesh } scene = kNone;
  float4* d_spheres = nullptr;
  uint32_t* d_node_id = nullptr;
  uint32_t n_spheres = 0;
  MeshScene mesh;  // host staging + device BVH / wavefront state (mesh_host.h)

  // progressive state (CPT.h:68)
  uint32_t frame_count = 0;
  uint32_t frame_first = 1, frame_stride = 1;
  uint32_t part_rank = 0, part_world = 1, tile_w = 32, tile_h = 32;
  bool counting = false;

  // environment map (octahedral), or nullptr = procedural sky
  float4* d_env = nullptr;
  uint32_t env_size = 0;

  // post-process (resolve stage)
  cudapt_postprocess post{};
  float* d_lum_partials = nullptr;  // auto-exposure scratch: partial sums | [n] = exposure
  uint32_t lum_capacity = 0;
  float last_exposure = 1.0f;

  // stats
  uint64_t launches = 0;
  double render_ms = 0.0, last_render_ms = 0.0, last_trace_ms = 0.0, last_packet_ms = 0.0;
};

namespace {

thread_local std::string g_create_error;

int fail(cudapt_ctx* ctx, int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  if (ctx)
    ctx->err = buf;
  else
    g_create_error = buf;
  return code;
}

#define CU_TRY(ctx, expr)                                                                              \
  do {                                                                                                 \
    cudaError_t e__ = (expr);                                                                          \
    if (e__ != cudaSuccess)                                                                            \
      return fail(ctx, CUDAPT_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, __LINE__); \
  } while (0)

bool valid_config(const cudapt_config& c) {
  return c.max_bounces >= 1 && c.max_bounces <= 64 && c.rr_start >= 0 && c.samples_per_pass >= 1 &&
         c.samples_per_pass <= 1024 && (c.rng_mode == CUDAPT_RNG_REFERENCE_PCG || c.rng_mode == CUDAPT_RNG_PHILOX);
}

CameraParams camera_params(const cudapt_camera& c) {
  CameraParams p;
  p.origin = mk3(c.origin[0], c.origin[1], c.origin[2]);
  p.right = mk3(c.right[0], c.right[1], c.right[2]);
  p.up = mk3(c.up[0], c.up[1], c.up[2]);
  p.forward = mk3(c.forward[0], c.forward[1], c.forward[2]);
  p.tan_half_fov = c.tan_half_fov;
  p.lens_radius = c.lens_radius > 0.0f ? c.lens_radius : 0.0f;
  p.focus_distance = c.focus_distance;
  return p;
}

ImageParams image_params(const cudapt_ctx* ctx) {
  ImageParams im;
  im.width = ctx->width;
  im.height = ctx->height;
  im.inv_width = 1.0f / (float)ctx->width;    // CPT.cpp:54
  im.inv_height = 1.0f / (float)ctx->height;  // CPT.cpp:53
  im.aspect = (float)ctx->width / (float)ctx->height;  // CPT.cpp:65
  im.part_rank = ctx->part_rank;
  im.part_world = ctx->part_world;
  im.tile_w = ctx->tile_w;
  im.tile_h = ctx->tile_h;
  im.tiles_x = (ctx->width + ctx->tile_w - 1) / ctx->tile_w;
  return im;
}

IntegratorParams integrator_params(const cudapt_config& c, const float4* env = nullptr, uint32_t env_size = 0) {
  IntegratorParams ip;
  ip.rng_mode = c.rng_mode;
  ip.max_bounces = c.max_bounces;
  ip.rr_start = c.rr_start;
  ip.jitter = c.jitter;
  ip.nee = c.nee;
  ip.tnear = c.tnear;
  ip.eps = c.eps;
  ip.albedo = c.albedo;
  ip.sky_scale = c.sky_scale;
  ip.seed = c.seed;
  ip.env = env;
  ip.env_size = env ? env_size : 0u;
  return ip;
}

int restart_accumulation(cudapt_ctx* ctx) {  // CPT.cpp:151-154
  if (ctx->d_accum) CU_TRY(ctx, cudaMemsetAsync(ctx->d_accum, 0, sizeof(float4) * (size_t)ctx->width * ctx->height, ctx->stream));
  ctx->frame_count = 0;
  return CUDAPT_OK;
}

// Collects the time of finished renders: waits for the `wait_for` oldest ones, then takes those that have completed.
int collect_timing(cudapt_ctx* ctx, int wait_for) {
  bool any = false;
  while (ctx->timing_count > 0) {
    const int k = ctx->timing_head;
    if (wait_for > 0) {
      CU_TRY(ctx, cudaEventSynchronize(ctx->ev1[k]));
      --wait_for;
    } else if (cudaEventQuery(ctx->ev1[k]) != cudaSuccess) {
      cudaGetLastError();  // cudaErrorNotReady is not an error here
      break;
    }
    float ms = 0.f;
    CU_TRY(ctx, cudaEventElapsedTime(&ms, ctx->ev0[k], ctx->ev1[k]));
    ctx->last_render_ms = ms;
    ctx->render_ms += ms;
    ctx->timing_head = (k + 1) % cudapt_ctx::kTimingRing;
    --ctx->timing_count;
    any = true;
  }
  // per-kernel event pairs of the mesh integrator belong to the most recent render: read them only when nothing is in flight
  if (any && ctx->timing_count == 0) {
    ctx->last_trace_ms = ctx->scene == cudapt_ctx::kMesh ? ctx->mesh.collect_trace_ms() : 0.0;
    ctx->last_packet_ms = ctx->scene == cudapt_ctx::kMesh ? ctx->mesh.collect_packet_ms() : 0.0;
  }
  return CUDAPT_OK;
}
int finish_timing(cudapt_ctx* ctx) { return collect_timing(ctx, ctx->timing_count); }

bool create_timing_events(cudapt_ctx* ctx) {
  for (int k = 0; k < cudapt_ctx::kTimingRing; ++k)
    if (cudaEventCreate(&ctx->ev0[k]) != cudaSuccess || cudaEventCreate(&ctx->ev1[k]) != cudaSuccess) return false;
  return true;
}

void free_spheres(cudapt_ctx* ctx) {
  cudaFree(ctx->d_spheres);
  cudaFree(ctx->d_node_id);
  ctx->d_spheres = nullptr;
  ctx->d_node_id = nullptr;
  ctx->n_spheres = 0;
}

}  // namespace

// for entry points that live in other translation units and have no context (csrc/static_scratch.cu)
namespace cudapt {
void set_thread_error(const char* text) { g_create_error = text ? text : ""; }
}  // namespace cudapt

extern "C" {

int cudapt_abi_version(void) { return CUDAPT_ABI_VERSION; }

void cudapt_default_config(cudapt_config* c) {
  std::memset(c, 0, sizeof(*c));
  c->device = 0;
  c->rng_mode = CUDAPT_RNG_REFERENCE_PCG;
  c->max_bounces = 4;   // CPT.cpp:199
  c->rr_start = 2;      // CPT.cpp:264
  c->tnear = 0.001f;    // CPT.cpp:221
  c->eps = 1e-4f;       // CPT.cpp:277
  c->albedo = 0.7f;     // CPT.cpp:260
  c->sky_scale = 1.0f;
  c->seed = 0;
  c->jitter = 0;
  c->nee = 0;
  c->samples_per_pass = 1;
}

void cudapt_default_camera(cudapt_camera* cam) {  // CPT.cpp:62-73
  std::memset(cam, 0, sizeof(*cam));
  cam->right[0] = 1.0f;
  cam->up[1] = 1.0f;
  cam->forward[2] = 1.0f;
  cam->tan_half_fov = 1.0f;
  cam->focus_distance = 1.0f;
}

// The placeholder render::Camera of the reference (Camera.h:10-38, `#if 0`): setPosition / setTarget / setUp /
// setPerspective(fov).  Basis in double, rounded once: forward = normalize(target - position), right =
// normalize(up x forward), up' = forward x right; with position 0, target +Z, up +Y and 90 degrees this is the
// hard-coded camera of CPT.cpp:62-73 exactly.
int cudapt_camera_look_at(cudapt_camera* cam, const float position[3], const float target[3], const float up[3],
                          float vfov_degrees) {
  if (!cam || !position || !target || !up) return CUDAPT_ERR_INVALID;
  if (!(vfov_degrees > 0.0f && vfov_degrees < 180.0f)) return CUDAPT_ERR_INVALID;
  double f[3], r[3], u[3];
  for (int k = 0; k < 3; ++k) f[k] = (double)target[k] - (double)position[k];
  const double fl = std::sqrt(f[0] * f[0] + f[1] * f[1] + f[2] * f[2]);
  if (!(fl > 0.0)) return CUDAPT_ERR_INVALID;
  for (int k = 0; k < 3; ++k) f[k] /= fl;
  r[0] = (double)up[1] * f[2] - (double)up[2] * f[1];
  r[1] = (double)up[2] * f[0] - (double)up[0] * f[2];
  r[2] = (double)up[0] * f[1] - (double)up[1] * f[0];
  const double rl = std::sqrt(r[0] * r[0] + r[1] * r[1] + r[2] * r[2]);
  if (!(rl > 1e-12)) return CUDAPT_ERR_INVALID;  // up parallel to the view direction
  for (int k = 0; k < 3; ++k) r[k] /= rl;
  u[0] = f[1] * r[2] - f[2] * r[1];
  u[1] = f[2] * r[0] - f[0] * r[2];
  u[2] = f[0] * r[1] - f[1] * r[0];
  for (int k = 0; k < 3; ++k) {
    cam->origin[k] = position[k];
    cam->right[k] = (float)r[k];
    cam->up[k] = (float)u[k];
    cam->forward[k] = (float)f[k];
  }
  cam->lens_radius = 0.0f;
  cam->focus_distance = (float)fl;  // the target is in focus once the caller opens the lens
  cam->tan_half_fov = vfov_degrees == 90.0f ? 1.0f : (float)std::tan((double)vfov_degrees * 0.5 * 3.14159265358979323846 / 180.0);
  return CUDAPT_OK;
}

const char* cudapt_last_error(const cudapt_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

int cudapt_create(const cudapt_config* cfg, cudapt_ctx** out) {
  if (!out) return fail(nullptr, CUDAPT_ERR_INVALID, "cudapt_create: out is NULL");
  *out = nullptr;
  cudapt_config c;
  if (cfg)
    c = *cfg;
  else
    cudapt_default_config(&c);
  if (!valid_config(c)) return fail(nullptr, CUDAPT_ERR_INVALID, "cudapt_create: invalid config");
  int n_dev = 0;
  cudaError_t e = cudaGetDeviceCount(&n_dev);
  if (e != cudaSuccess || n_dev == 0)
    return fail(nullptr, CUDAPT_ERR_CUDA, "cudapt_create: no CUDA device (%s); this backend has no CPU fallback",
                cudaGetErrorString(e));
  if (c.device < 0 || c.device >= n_dev) return fail(nullptr, CUDAPT_ERR_INVALID, "cudapt_create: device %d of %d", c.device, n_dev);
  cudaDeviceProp prop;
  CU_TRY(nullptr, cudaGetDeviceProperties(&prop, c.device));
  if (prop.major != 10)
    return fail(nullptr, CUDAPT_ERR_CUDA, "cudapt_create: device %d is sm_%d%d; libcudapt is built for sm_100a only", c.device,
                prop.major, prop.minor);
  CU_TRY(nullptr, cudaSetDevice(c.device));
  cudapt_ctx* ctx = new (std::nothrow) cudapt_ctx();
  if (!ctx) return fail(nullptr, CUDAPT_ERR_INVALID, "out of host memory");
  ctx->cfg = c;
  ctx->sm_count = prop.multiProcessorCount;
  cudapt_default_camera(&ctx->cam);
  cudapt_default_postprocess(&ctx->post);
  ctx->mesh.sm_count = ctx->sm_count;
  if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess ||
      !create_timing_events(ctx) ||
      cudaMalloc(&ctx->d_counters, sizeof(unsigned long long) * kCntSlots) != cudaSuccess ||
      cudaMemset(ctx->d_counters, 0, sizeof(unsigned long long) * kCntSlots) != cudaSuccess) {
    fail(nullptr, CUDAPT_ERR_CUDA, "cudapt_create: stream/event/counter setup failed: %s", cudaGetErrorString(cudaGetLastError()));
    cudapt_destroy(ctx);
    return CUDAPT_ERR_CUDA;
  }
  *out = ctx;
  return CUDAPT_OK;
}

void cudapt_destroy(cudapt_ctx* ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->cfg.device);
  if (ctx->stream) cudaStreamSynchronize(ctx->stream);
  free_spheres(ctx);
  ctx->mesh.release();
  cudaFree(ctx->d_accum);
  cudaFree(ctx->d_rgba);
  cudaFree(ctx->d_aov_prim);
  cudaFree(ctx->d_aov_t);
  cudaFree(ctx->d_counters);
  cudaFree(ctx->d_lum_partials);
  cudaFree(ctx->d_env);
  if (ctx->h_pinned_rgba) cudaFreeHost(ctx->h_pinned_rgba);
  if (ctx->dl_stream) {
    cudaStreamSynchronize(ctx->dl_stream);
    cudaStreamDestroy(ctx->dl_stream);
  }
  if (ctx->ev_img_ready) cudaEventDestroy(ctx->ev_img_ready);
  if (ctx->ev_dl_done) cudaEventDestroy(ctx->ev_dl_done);
  for (int k = 0; k < cudapt_ctx::kTimingRing; ++k) {
    if (ctx->ev0[k]) cudaEventDestroy(ctx->ev0[k]);
    if (ctx->ev1[k]) cudaEventDestroy(ctx->ev1[k]);
  }
  if (ctx->stream) cudaStreamDestroy(ctx->stream);
  delete ctx;
}

int cudapt_set_resolution(cudapt_ctx* ctx, uint32_t width, uint32_t height) {
  if (!ctx) return CUDAPT_ERR_INVALID;
  if (width == 0 || height == 0 || (uint64_t)width * height > (1ull << 28))
    return fail(ctx, CUDAPT_ERR_INVALID, "cudapt_set_resolution: bad size %ux%u", width, height);
  CU_TRY(ctx, cudaSetDevice(ctx->cfg.device));
  if (width != ctx->width || height != ctx->height || !ctx->d_accum) {  // CPT.cpp:140-149
    CU_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    if (ctx->dl_pending) {  // an asynchronous read-back still reads the old image buffer
      CU_TRY(ctx, cudaEventSynchronize(ctx->ev_dl_done));
      ctx->dl_pending = false;
    }
    cudaFree(ctx->d_accum);
    cudaFree(ctx->d_rgba);
    cudaFree(ctx->d_aov_prim);
    cudaFree(ctx->d_aov_t);
    ctx->d_accum = nullptr, ctx->d_rgba = nullptr, ctx->d_aov_prim = nullptr, ctx->d_aov_t = nullptr;
    const size_t n = (size_t)width * height;
    CU_TRY(ctx, cudaMalloc(&ctx->d_accum, sizeof(float4) * n));
    CU_TRY(ctx, cudaMalloc(&ctx->d_rgba, sizeof(uint32_t) * n));
    CU_TRY(ctx, cudaMalloc(&ctx->d_aov_prim, sizeof(uint32_t) * n));
    CU_TRY(ctx, cudaMalloc(&ctx->d_aov_t, sizeof(float) * n));
    if (ctx->pinned_pixels < n) {
      if (ctx->h_pinned_rgba) cudaFreeHost(ctx->h_pinned_rgba);
      ctx->h_pinned_rgba = nullptr;
      CU_TRY(ctx, cudaMallocHost(&ctx->h_pinned_rgba, sizeof(uint32_t) * n));
      ctx->pinned_pixels = n;
    }
    ctx->width = width;
    ctx->height = height;
    ctx->mesh.invalidate_image_state();
  }
  CU_TRY(ctx, cudaMemsetAsync(ctx->d_aov_prim, 0xFF, sizeof(uint32_t) * (size_t)width * height, ctx->stream));
  CU_TRY(ctx, cudaMemsetAsync(ctx->d_aov_t, 0, sizeof(float) * (size_t)width * height, ctx->stream));
  return restart_accumulation(ctx);
}

int cudapt_set_config(cudapt_ctx* ctx, const cudapt_config* cfg) {
  if (!ctx || !cfg) return CUDAPT_ERR_INVALID;
  if (!valid_config(*cfg)) return fail(ctx, CUDAPT_ERR_INVALID, "cudapt_set_config: invalid config");
  const int dev = ctx->cfg.device;
  const bool spp_changed = cfg->samples_per_pass != ctx->cfg.samples_per_pass;
  ctx->cfg = *cfg;
  ctx->cfg.device = dev;
  if (spp_changed) ctx->mesh.invalidate_image_state();
  CU_TRY(ctx, cudaSetDevice(dev));
  return restart_accumulation(ctx);  // CPT.cpp:132-138
}

int cudapt_set_camera(cudapt_ctx* ctx, const cudapt_camera* cam) {
  if (!ctx || !cam) return CUDAPT_ERR_INVALID;
  ctx->cam = *cam;
  CU_TRY(ctx, cudaSetDevice(ctx->cfg.device));
  return restart_accumulation(ctx);
}

int cudapt_upload_spheres(cudapt_ctx* ctx, const float* center_radius, const uint32_t* node_id, uint32_t n) {
  if (!ctx || (n && !center_radius)) return CUDAPT_ERR_INVALID;
  if ((size_t)n * sizeof(float4) > 200 * 1024)
    return fail(ctx, CUDAPT_ERR_INVALID, "cudapt_upload_spheres: %u spheres exceed the shared-memory scene (12800)", n);
  CU_TRY(ctx, cudaSetDevice(ctx->cfg.device));
  CU_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  free_spheres(ctx);  // rebuild = replace (the reference re-attaches without detaching, CPT.cpp:362-403)
  std::vector<uint32_t> ids(n);
  for (uint32_t i = 0; i < n; ++i) ids[i] = node_id ? node_id[i] : i;
  CU_TRY(ctx, cudaMalloc(&ctx->d_spheres, sizeof(float4) * (n ? n : 1)));
  CU_TRY(ctx, cudaMalloc(&ctx->d_node_id, sizeof(uint32_t) * (n ? n : 1)));
  if (n) {
    CU_TRY(ctx, cudaMemcpy(ctx->d_spheres, center_radius, sizeof(float4) * n, cudaMemcpyHostToDevice));
    CU_TRY(ctx, cudaMemcpy(ctx->d_node_id, ids.data(), sizeof(uint32_t) * n, cudaMemcpyHostToDevice));
  }
  ctx->n_spheres = n;
  ctx->scene = cudapt_ctx::kSpheres;
  return restart_accumulation(ctx);
}

int cudapt_upload_mesh(cudapt_ctx* ctx, const float* vertices, uint32_t nv, const uint32_t* indices, uint32_t nt,
                       const uint32_t* material_id) {
  if (!ctx || !vertices || !indices || nt == 0 || nv == 0) return fail(ctx, CUDAPT_ERR_INVALID, "cudapt_upload_mesh: empty mesh");
  for (size_t i = 0; i < (size_t)nt * 3; ++i)
    if (indices[i] >= nv) return fail(ctx, CUDAPT_ERR_INVALID, "cudapt_upload_mesh: index %u out of range", indices[i]);
  ctx->mesh.set_mesh(vertices, nv, indices, nt, material_id);
  return CUDAPT_OK;
}

int cudapt_upload_materials(cudapt_ctx* ctx, const cudapt_material* materials, uint32_t n) {
  if (!ctx || (n && !materials)) return CUDAPT_ERR_INVALID;
  ctx->mesh.set_materials(materials, n);
  return CUDAPT_OK;
}

int cudapt_build_accel(cudapt_ctx* ctx) {
  if (!ctx) return CUDAPT_ERR_INVALID;
  if (!ctx->mesh.has_host_mesh()) return fail(ctx, CUDAPT_ERR_NO_SCENE, "cudapt_build_accel: no mesh uploaded");
  CU_TRY(ctx, cudaSetDevice(ctx->cfg.device));
  CU_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  std::string err;
  if (!ctx->mesh.build_and_upload(&err)) return fail(ctx, CUDAPT_ERR_CUDA, "cudapt_build_accel: %s", err.c_str());
  ctx->scene = cudapt_ctx::kMesh;
  return restart_accumulation(ctx);
}

int cudapt_build_accel_device(cudapt_ctx* ctx) {
  if (!ctx) return CUDAPT_ERR_INVALID;
  if (!ctx->mesh.has_host_mesh()) return fail(ctx, CUDAPT_ERR_NO_SCENE, "cudapt_build_accel_device: no mesh uploaded");
  CU_TRY(ctx, cudaSetDevice(ctx->cfg.device));
  CU_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  std::string err;
  uint64_t launches = 0;
  const bool ok = ctx->mesh.build_on_device(ctx->stream, &launches, &err);
  ctx->launches += launches;
  if (!ok) return fail(ctx, CUDAPT_ERR_CUDA, "cudapt_build_accel_device: %s", err.c_str());
  ctx->scene = cudapt_ctx::kMesh;
  return restart_accumulation(ctx);
}

// a staged scene (cudapt_stage_scene) takes effect before anything else touches the scene
static int commit_staged_scene(cudapt_ctx* ctx) {
  if (ctx->scene != cudapt_ctx::kMesh || !ctx->mesh.swap_pending()) return CUDAPT_OK;
  std::string err;
  if (!ctx->mesh.commit_staged(ctx->stream, &err)) return fail(ctx, CUDAPT_ERR_CUDA, "staged scene: %s", err.c_str());
  return restart_accumulation(ctx);
}

int cudapt_reupload_scene(cudapt_ctx* ctx) {
  if (!ctx) return CUDAPT_ERR_INVALID;
  if (ctx->scene != cudapt_ctx::kMesh) return fail(ctx, CUDAPT_ERR_NO_SCENE, "cudapt_reupload_scene: no built mesh scene");
  CU_TRY(ctx, cudaSetDevice(ctx->cfg.device));
  if (int rc = commit_staged_scene(ctx)) return rc;
  std::string err;
  if (!ctx->mesh.reupload(ctx->stream, &err)) return fail(ctx, CUDAPT_ERR_CUDA, "cudapt_reupload_scene: %s", err.c_str());
  return restart_accumulation(ctx);
}

int cudapt_stage_scene(cudapt_ctx* ctx) {
  if (!ctx) return CUDAPT_ERR_INVALID;
  if (ctx->scene != cudapt_ctx::kMesh) return fail(ctx, CUDAPT_ERR_NO_SCENE, "cudapt_stage_scene: no built mesh scene");
  CU_TRY(ctx, cudaSetDevice(ctx->cfg.device));
  std::string err;
  if (!ctx->mesh.stage(ctx->stream, &err)) return fail(ctx, CUDAPT_ERR_CUDA, "cudapt_stage_scene: %s", err.c_str());
  return CUDAPT_OK;
}

int cudapt_update_vertices(cudapt_ctx* ctx, const float* vertices, uint32_t nv) {
  if (!ctx || !vertices) return CUDAPT_ERR_INVALID;
  if (ctx->scene != cudapt_ctx::kMesh) return fail(ctx, CUDAPT_ERR_NO_SCENE, "cudapt_update_vertices: no built mesh scene");
  CU_TRY(ctx, cudaSetDevice(ctx->cfg.device));
  if (int rc = commit_staged_scene(ctx)) return rc;
  CU_TRY(ctx, cudaStreamSynchronize(ctx->stream));  // the pinned staging buffer may still be in flight
  std::string err;
  uint64_t launches = 0;
  const bool ok = ctx->mesh.update_vertices(vertices, nv, ctx->stream, &launches, &err);
  ctx->launches += launches;
  if (!ok) return fail(ctx, CUDAPT_ERR_INVALID, "cudapt_update_vertices: %s", err.c_str());
  return restart_accumulation(ctx);
}

int cudapt_upload_environment(cudapt_ctx* ctx, const float* rgb, uint32_t size) {
  if (!ctx) return CUDAPT_ERR_INVALID;
  if (size > 16384u) return fail(ctx, CUDAPT_ERR_INVALID, "cudapt_upload_environment: size %u too large", size);
  CU_TRY(ctx, cudaSetDevice(ctx->cfg.device));
  CU_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  cudaFree(ctx->d_env);
  ctx->d_env = nullptr;
  ctx->env_size = 0;
  if (rgb && size) {
    const size_t n = (size_t)size * size;
    std::vector<float4> texels(n);
    for (size_t i = 0; i < n; ++i) texels[i] = make_float4(rgb[3 * i + 0], rgb[3 * i + 1], rgb[3 * i + 2], 0.0f);
    CU_TRY(ctx, cudaMalloc(&ctx->d_env, sizeof(float4) * n));
    CU_TRY(ctx, cudaMemcpy(ctx->d_env, texels.data(), sizeof(float4) * n, cudaMemcpyHostToDevice));
    ctx->env_size = size;
  }
  return restart_accumulation(ctx);
}

int cudapt_convert_environment(const char* path, uint32_t size, float* octahedral_out) {  // host only, no device
  if (!path || !octahedral_out) return fail(nullptr, CUDAPT_ERR_INVALID, "cudapt_convert_environment: null argument");
  if (size < 2u || size > 16384u) return fail(nullptr, CUDAPT_ERR_INVALID, "cudapt_convert_environment: size %u", size);
  LatLongImage im;
  std::string err;
  if (!load_latlong_image(path, &im, &err)) return fail(nullptr, CUDAPT_ERR_INVALID, "cudapt_convert_environment: %s", err.c_str());
  latlong_to_octahedral(im, size, octahedral_out);
  return CUDAPT_OK;
}

int cudapt_load_environment(cudapt_ctx* ctx, const char* path, uint32_t size, float* octahedral_out) {
  if (!ctx || !path) return CUDAPT_ERR_INVALID;
  if (size < 2u || size > 16384u) return fail(ctx, CUDAPT_ERR_INVALID, "cudapt_load_environment: size %u", size);
  std::vector<float> octa((size_t)size * size * 3);
  if (cudapt_convert_environment(path, size, octa.data()) != CUDAPT_OK)
    return fail(ctx, CUDAPT_ERR_INVALID, "cudapt_load_environment: %s", g_create_error.c_str());
  if (octahedral_out) std::memcpy(octahedral_out, octa.data(), sizeof(float) * octa.size());
  return cudapt_upload_environment(ctx, octa.data(), size);
}

int cudapt_reset(cudapt_ctx* ctx) {
  if (!ctx) return CUDAPT_ERR_INVALID;
  CU_TRY(ctx, cudaSetDevice(ctx->cfg.device));
  int rc = finish_timing(ctx);
  if (rc) return rc;
  CU_TRY(ctx, cudaMemsetAsync(ctx->d_counters, 0, sizeof(unsigned long long) * kCntSlots, ctx->stream));
  ctx->render_ms = 0.0;
  ctx->last_render_ms = 0.0;
  return restart_accumulation(ctx);
}

int cudapt_render(cudapt_ctx* ctx, uint32_t n_frames) {
  if (!ctx) return CUDAPT_ERR_INVALID;
  if (ctx->scene == cudapt_ctx::kNone) return fail(ctx, CUDAPT_ERR_NO_SCENE, "cudapt_render: scene not set before rendering");
  if (!ctx->d_accum) return fail(ctx, CUDAPT_ERR_INVALID, "cudapt_render: resolution not set");
  if (n_frames == 0) return CUDAPT_OK;
  CU_TRY(ctx, cudaSetDevice(ctx->cfg.device));
  // the time of finished renders is collected without waiting; only a full ring makes the host wait for the oldest
  int rc = collect_timing(ctx, ctx->timing_count == cudapt_ctx::kTimingRing ? 1 : 0);
  if (rc) return rc;
  if ((rc = commit_staged_scene(ctx)) != CUDAPT_OK) return rc;
  const int tk = (ctx->timing_head + ctx->timing_count) % cudapt_ctx::kTimingRing;
  const uint32_t first = ctx->frame_first + ctx->frame_count * ctx->frame_stride;  // 1-based (CPT.cpp:61)
  const bool want_aov = ctx->frame_count == 0;
  CU_TRY(ctx, cudaEventRecord(ctx->ev0[tk], ctx->stream));
  if (ctx->scene == cudapt_ctx::kSpheres) {
    SphereLaunch a;
    a.spheres = ctx->d_spheres;
    a.node_id = ctx->d_node_id;
    a.n_spheres = (int)ctx->n_spheres;
    a.accum = ctx->d_accum;
    a.aov_prim = want_aov ? ctx->d_aov_prim : nullptr;
    a.aov_t = want_aov ? ctx->d_aov_t : nullptr;
    a.counters = ctx->d_counters;
    a.first_frame = first;
    a.n_frames = n_frames;
    a.frame_stride = ctx->frame_stride;
    a.cam = camera_params(ctx->cam);
    a.im = image_params(ctx);
    a.ip = integrator_params(ctx->cfg, ctx->d_env, ctx->env_size);
    CU_TRY(ctx, launch_render_spheres(a, ctx->cfg.rng_mode == CUDAPT_RNG_PHILOX, ctx->stream));
    ctx->launches += 1;
  } else {
    if (ctx->cfg.rng_mode != CUDAPT_RNG_PHILOX)
      return fail(ctx, CUDAPT_ERR_INVALID, "cudapt_render: the mesh integrator requires CUDAPT_RNG_PHILOX");
    MeshRenderArgs ra;
    ra.accum = ctx->d_accum;
    ra.aov_prim = want_aov ? ctx->d_aov_prim : nullptr;
    ra.aov_t = want_aov ? ctx->d_aov_t : nullptr;
    ra.counters = ctx->d_counters;
    ra.first_frame = first;
    ra.n_frames = n_frames;
    ra.frame_stride = ctx->frame_stride;
    ra.cam = camera_params(ctx->cam);
    ra.im = image_params(ctx);
    ra.ip = integrator_params(ctx->cfg, ctx->d_env, ctx->env_size);
    ra.samples_per_pass = (uint32_t)ctx->cfg.samples_per_pass;
    ra.counting = ctx->counting;
    std::string err;
    uint64_t launched = 0;
    cudaError_t e = ctx->mesh.render(ra, ctx->stream, &launched, &err);
    if (e != cudaSuccess) return fail(ctx, CUDAPT_ERR_CUDA, "cudapt_render(mesh): %s: %s", err.c_str(), cudaGetErrorString(e));
    ctx->launches += launched;
  }
  CU_TRY(ctx, cudaEventRecord(ctx->ev1[tk], ctx->stream));
  ++ctx->timing_count;
  ctx->frame_count += n_frames;
  return CUDAPT_OK;
}

int cudapt_synchronize(cudapt_ctx* ctx) {
  if (!ctx) return CUDAPT_ERR_INVALID;
  CU_TRY(ctx, cudaSetDevice(ctx->cfg.device));
  CU_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  return finish_timing(ctx);
}

void cudapt_default_postprocess(cudapt_postprocess* pp) {
  std::memset(pp, 0, sizeof(*pp));
  pp->exposure = 1.0f;           // Types.h:87
  pp->tonemap = CUDAPT_TONEMAP_NONE;
  pp->srgb = 0;
  pp->auto_exposure = 0;         // Types.h:88
  pp->target_luminance = 0.18f;  // Types.h:89
}

int cudapt_set_postprocess(cudapt_ctx* ctx, const cudapt_postprocess* pp) {
  if (!ctx || !pp) return CUDAPT_ERR_INVALID;
  if (pp->tonemap < CUDAPT_TONEMAP_NONE || pp->tonemap > CUDAPT_TONEMAP_ACES || !(pp->exposure >= 0.0f) ||
      !(pp->target_luminance > 0.0f))
    return fail(ctx, CUDAPT_ERR_INVALID, "cudapt_set_postprocess: invalid parameters");
  ctx->post = *pp;  // a display setting: the accumulation is NOT restarted
  return CUDAPT_OK;
}

int cudapt_get_exposure(cudapt_ctx* ctx, float* exposure) {
  if (!ctx || !exposure) return CUDAPT_ERR_INVALID;
  *exposure = ctx->last_exposure;
  return CUDAPT_OK;
}

namespace {
// true when the D2H copy can land in the caller's buffer directly: page-locked host memory (cudaMallocHost, or
// registered with cudapt_host_register / cudaHostRegister)
bool host_is_pinned(const void* p) {
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return at.type == cudaMemoryTypeHost;
}

// A pending asynchronous read-back still reads d_rgba: whatever writes the image next waits for it (device side).
int order_after_download(cudapt_ctx* ctx) {
  if (ctx->dl_pending) CU_TRY(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_dl_done, 0));
  return CUDAPT_OK;
}

// resolve (+ auto-exposure) kernels into d_rgba, enqueued on the stream; *d_exposure_out = the device exposure value
// when auto-exposure is on
int enqueue_resolve(cudapt_ctx* ctx, const char* who, const float** d_exposure_out) {
  if (!ctx->d_accum) return fail(ctx, CUDAPT_ERR_INVALID, "%s: resolution not set", who);
  if (ctx->frame_count == 0) return fail(ctx, CUDAPT_ERR_NO_FRAMES, "%s: no frames rendered yet", who);
  CU_TRY(ctx, cudaSetDevice(ctx->cfg.device));
  const int rc = order_after_download(ctx);
  if (rc != CUDAPT_OK) return rc;
  const uint32_t n = ctx->width * ctx->height;
  const cudapt_postprocess& pp = ctx->post;
  const bool plain = pp.exposure == 1.0f && pp.tonemap == CUDAPT_TONEMAP_NONE && !pp.srgb && !pp.auto_exposure;
  const float* d_exposure = nullptr;
  if (plain) {
    CU_TRY(ctx, launch_resolve_rgba8(ctx->d_accum, ctx->d_rgba, n, ctx->frame_count, ctx->stream));  // CPT.cpp:87-117 as is
    ctx->launches += 1;
  } else {
    if (pp.auto_exposure) {
      const uint32_t np = auto_exposure_partials(n);
      if (ctx->lum_capacity < np + 1) {
        cudaFree(ctx->d_lum_partials);
        ctx->d_lum_partials = nullptr;
        CU_TRY(ctx, cudaMalloc(&ctx->d_lum_partials, sizeof(float) * (size_t)(np + 1)));
        ctx->lum_capacity = np + 1;
      }
      CU_TRY(ctx, launch_auto_exposure(ctx->d_accum, n, ctx->frame_count, pp.target_luminance, ctx->d_lum_partials,
                                       ctx->d_lum_partials + np, ctx->stream));
      ctx->launches += 2;
      d_exposure = ctx->d_lum_partials + np;
    }
    CU_TRY(ctx, launch_resolve_post(ctx->d_accum, ctx->d_rgba, n, ctx->frame_count, pp.exposure, d_exposure, pp.tonemap, pp.srgb,
                                    ctx->stream));
    ctx->launches += 1;
  }
  *d_exposure_out = d_exposure;
  return CUDAPT_OK;
}

// resolve (+ auto-exposure) into d_rgba and `dst` (page-locked: the library's staging buffer when null);
// synchronises the stream
int resolve_to_pinned(cudapt_ctx* ctx, const char* who, uint32_t* dst = nullptr) {
  const float* d_exposure = nullptr;
  const int rc = enqueue_resolve(ctx, who, &d_exposure);
  if (rc != CUDAPT_OK) return rc;
  const uint32_t n = ctx->width * ctx->height;
  CU_TRY(ctx, cudaMemcpyAsync(dst ? dst : ctx->h_pinned_rgba, ctx->d_rgba, sizeof(uint32_t) * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
  ctx->last_exposure = ctx->post.exposure;
  if (d_exposure) CU_TRY(ctx, cudaMemcpyAsync(&ctx->last_exposure, d_exposure, sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
  CU_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  return finish_timing(ctx);
}

// D2H of d_rgba into the caller's page-locked buffer on the download stream, after everything queued on `stream`
int enqueue_download(cudapt_ctx* ctx, const char* who, uint32_t* host_out) {
  if (!host_is_pinned(host_out))
    return fail(ctx, CUDAPT_ERR_INVALID, "%s: the destination must be page-locked (cudapt_host_register)", who);
  if (!ctx->dl_stream) {
    CU_TRY(ctx, cudaStreamCreateWithFlags(&ctx->dl_stream, cudaStreamNonBlocking));
    CU_TRY(ctx, cudaEventCreateWithFlags(&ctx->ev_img_ready, cudaEventDisableTiming));
    CU_TRY(ctx, cudaEventCreateWithFlags(&ctx->ev_dl_done, cudaEventDisableTiming));
  }
  CU_TRY(ctx, cudaEventRecord(ctx->ev_img_ready, ctx->stream));
  CU_TRY(ctx, cudaStreamWaitEvent(ctx->dl_stream, ctx->ev_img_ready, 0));
  CU_TRY(ctx, cudaMemcpyAsync(host_out, ctx->d_rgba, sizeof(uint32_t) * (size_t)ctx->width * ctx->height, cudaMemcpyDeviceToHost,
                              ctx->dl_stream));
  CU_TRY(ctx, cudaEventRecord(ctx->ev_dl_done, ctx->dl_stream));
  ctx->dl_pending = true;
  return CUDAPT_OK;
}
}  // namespace

int cudapt_resolve_rgba8(cudapt_ctx* ctx, uint32_t* host_out) {
  if (!ctx || !host_out) return CUDAPT_ERR_INVALID;
  // page-locked destination: the copy engine writes the image where the caller wants it (no staging copy)
  if (host_is_pinned(host_out)) return resolve_to_pinned(ctx, "cudapt_resolve_rgba8", host_out);
  const int rc = resolve_to_pinned(ctx, "cudapt_resolve_rgba8");
  if (rc != CUDAPT_OK) return rc;
  std::memcpy(host_out, ctx->h_pinned_rgba, sizeof(uint32_t) * (size_t)ctx->width * ctx->height);
  return CUDAPT_OK;
}

int cudapt_host_register(cudapt_ctx* ctx, void* host_ptr, size_t bytes) {
  if (!ctx || !host_ptr || bytes == 0) return CUDAPT_ERR_INVALID;
  CU_TRY(ctx, cudaSetDevice(ctx->cfg.device));
  const cudaError_t e = cudaHostRegister(host_ptr, bytes, cudaHostRegisterDefault);
  if (e != cudaSuccess) {
    cudaGetLastError();  // not sticky: the context stays usable
    return fail(ctx, CUDAPT_ERR_CUDA, "cudapt_host_register: %s", cudaGetErrorString(e));
  }
  return CUDAPT_OK;
}
int cudapt_resolve_rgba8_async(cudapt_ctx* ctx, uint32_t* host_out) {
  if (!ctx || !host_out) return CUDAPT_ERR_INVALID;
  if (ctx->post.auto_exposure) return fail(ctx, CUDAPT_ERR_INVALID, "cudapt_resolve_rgba8_async: auto-exposure reads its result back; use cudapt_resolve_rgba8");
  if (!host_is_pinned(host_out))
    return fail(ctx, CUDAPT_ERR_INVALID, "cudapt_resolve_rgba8_async: the destination must be page-locked (cudapt_host_register)");
  const float* d_exposure = nullptr;
  const int rc = enqueue_resolve(ctx, "cudapt_resolve_rgba8_async", &d_exposure);
  if (rc != CUDAPT_OK) return rc;
  ctx->last_exposure = ctx->post.exposure;
  return enqueue_download(ctx, "cudapt_resolve_rgba8_async", host_out);
}

int cudapt_download_rgba8_async(cudapt_ctx* ctx, uint32_t* host_out) {
  if (!ctx || !host_out) return CUDAPT_ERR_INVALID;
  if (!ctx->d_rgba) return fail(ctx, CUDAPT_ERR_INVALID, "cudapt_download_rgba8_async: resolution not set");
  CU_TRY(ctx, cudaSetDevice(ctx->cfg.device));
  return enqueue_download(ctx, "cudapt_download_rgba8_async", host_out);
}

int cudapt_wait_download(cudapt_ctx* ctx) {
  if (!ctx) return CUDAPT_ERR_INVALID;
  if (!ctx->dl_pending) return CUDAPT_OK;
  CU_TRY(ctx, cudaSetDevice(ctx->cfg.device));
  CU_TRY(ctx, cudaEventSynchronize(ctx->ev_dl_done));
  ctx->dl_pending = false;
  return finish_timing(ctx);
}

int cudapt_stream_wait_download(cudapt_ctx* ctx) {
  if (!ctx) return CUDAPT_ERR_INVALID;
  CU_TRY(ctx, cudaSetDevice(ctx->cfg.device));
  return order_after_download(ctx);
}

int cudapt_host_unregister(cudapt_ctx* ctx, void* host_ptr) {
  if (!ctx || !host_ptr) return CUDAPT_ERR_INVALID;
  CU_TRY(ctx, cudaSetDevice(ctx->cfg.device));
  const cudaError_t e = cudaHostUnregister(host_ptr);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return fail(ctx, CUDAPT_ERR_CUDA, "cudapt_host_unregister: %s", cudaGetErrorString(e));
  }
  return CUDAPT_OK;
}

// PNG, 8-bit RGBA, written without a compression library: the zlib stream consists of STORED deflate blocks (RFC 1951
// 3.2.4), which every PNG reader accepts; CRC-32 per chunk (PNG spec 5.5), Adler-32 over the filtered scanlines.
namespace {
uint32_t crc32_update(uint32_t crc, const unsigned char* p, size_t n) {
  static uint32_t table[256];
  static bool ready = false;
  if (!ready) {
    for (uint32_t i = 0; i < 256; ++i) {
      uint32_t c = i;
      for (int k = 0; k < 8; ++k) c = (c & 1u) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
      table[i] = c;
    }
    ready = true;
  }
  for (size_t i = 0; i < n; ++i) crc = table[(crc ^ p[i]) & 0xFFu] ^ (crc >> 8);
  return crc;
}
void put_be32(std::vector<unsigned char>& v, uint32_t x) {
  v.push_back((unsigned char)(x >> 24)), v.push_back((unsigned char)(x >> 16)), v.push_back((unsigned char)(x >> 8)), v.push_back((unsigned char)x);
}
void put_chunk(std::vector<unsigned char>& out, const char* type, const std::vector<unsigned char>& data) {
  put_be32(out, (uint32_t)data.size());
  const size_t start = out.size();
  out.insert(out.end(), type, type + 4);
  out.insert(out.end(), data.begin(), data.end());
  put_be32(out, crc32_update(0xFFFFFFFFu, out.data() + start, out.size() - start) ^ 0xFFFFFFFFu);
}
std::vector<unsigned char> encode_png_rgba8(const uint32_t* packed, uint32_t W, uint32_t H) {
  // filtered scanlines: filter byte 0 + R G B A per pixel (packed word = R<<24 | G<<16 | B<<8 | A, Color.h:9)
  std::vector<unsigned char> raw((size_t)H * (1 + 4 * (size_t)W));
  size_t o = 0;
  for (uint32_t y = 0; y < H; ++y) {
    raw[o++] = 0;
    for (uint32_t x = 0; x < W; ++x) {
      const uint32_t p = packed[(size_t)y * W + x];
      raw[o++] = (unsigned char)(p >> 24), raw[o++] = (unsigned char)(p >> 16), raw[o++] = (unsigned char)(p >> 8), raw[o++] = (unsigned char)p;
    }
  }
  std::vector<unsigned char> z;
  z.push_back(0x78), z.push_back(0x01);  // zlib header: deflate, 32 K window, no preset dictionary
  uint32_t a = 1, b = 0;
  for (size_t pos = 0; pos < raw.size() || pos == 0;) {
    const size_t n = std::min<size_t>(65535, raw.size() - pos);
    z.push_back(pos + n >= raw.size() ? 1 : 0);  // BFINAL, BTYPE = 00 (stored)
    z.push_back((unsigned char)(n & 0xFF)), z.push_back((unsigned char)(n >> 8));
    z.push_back((unsigned char)(~n & 0xFF)), z.push_back((unsigned char)((~n >> 8) & 0xFF));
    z.insert(z.end(), raw.begin() + pos, raw.begin() + pos + n);
    for (size_t i = pos; i < pos + n; ++i) {
      a = (a + raw[i]) % 65521u;
      b = (b + a) % 65521u;
    }
    pos += n;
    if (n == 0) break;
  }
  put_be32(z, (b << 16) | a);
  std::vector<unsigned char> out = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
  std::vector<unsigned char> ihdr;
  put_be32(ihdr, W), put_be32(ihdr, H);
  ihdr.push_back(8), ihdr.push_back(6), ihdr.push_back(0), ihdr.push_back(0), ihdr.push_back(0);  // 8 bit, RGBA, no interlace
  put_chunk(out, "IHDR", ihdr);
  put_chunk(out, "IDAT", z);
  put_chunk(out, "IEND", {});
  return out;
}
}  // namespace

namespace {
// OpenEXR 2, single-part scan-line image, 32-bit FLOAT channels B G R (the file format orders channels by name),
// no compression, one scan line per block, increasing y: linear radiance accum / frames, nothing else.
void put_le(std::vector<unsigned char>& v, const void* p, size_t n) {  // the box is little-endian, as the format
  const unsigned char* b = static_cast<const unsigned char*>(p);
  v.insert(v.end(), b, b + n);
}
void put_i32(std::vector<unsigned char>& v, int32_t x) { put_le(v, &x, 4); }
void put_f32(std::vector<unsigned char>& v, float x) { put_le(v, &x, 4); }
void put_str(std::vector<unsigned char>& v, const char* s) { put_le(v, s, std::strlen(s) + 1); }
void put_attr(std::vector<unsigned char>& v, const char* name, const char* type, const std::vector<unsigned char>& value) {
  put_str(v, name);
  put_str(v, type);
  put_i32(v, (int32_t)value.size());
  v.insert(v.end(), value.begin(), value.end());
}
std::vector<unsigned char> encode_exr_rgb32f(const float* accum_rgba, uint32_t frames, uint32_t W, uint32_t H) {
  std::vector<unsigned char> out;
  const unsigned char magic[4] = {0x76, 0x2f, 0x31, 0x01};
  put_le(out, magic, 4);
  put_i32(out, 2);  // version 2, no flags: single-part scan lines, short names
  std::vector<unsigned char> a;
  for (const char* ch : {"B", "G", "R"}) {
    put_str(a, ch);
    put_i32(a, 2);  // FLOAT
    const unsigned char lin_reserved[4] = {0, 0, 0, 0};
    put_le(a, lin_reserved, 4);
    put_i32(a, 1), put_i32(a, 1);  // x / y sampling
  }
  a.push_back(0);
  put_attr(out, "channels", "chlist", a);
  put_attr(out, "compression", "compression", {0});
  a.clear();
  put_i32(a, 0), put_i32(a, 0), put_i32(a, (int32_t)W - 1), put_i32(a, (int32_t)H - 1);
  put_attr(out, "dataWindow", "box2i", a);
  put_attr(out, "displayWindow", "box2i", a);
  put_attr(out, "lineOrder", "lineOrder", {0});
  a.clear();
  put_f32(a, 1.0f);
  put_attr(out, "pixelAspectRatio", "float", a);
  a.clear();
  put_f32(a, 0.0f), put_f32(a, 0.0f);
  put_attr(out, "screenWindowCenter", "v2f", a);
  a.clear();
  put_f32(a, 1.0f);
  put_attr(out, "screenWindowWidth", "float", a);
  out.push_back(0);  // end of header
  const size_t line_bytes = 3 * sizeof(float) * (size_t)W, block_bytes = 8 + line_bytes;
  const uint64_t first = out.size() + 8 * (uint64_t)H;
  for (uint32_t y = 0; y < H; ++y) {
    const uint64_t off = first + block_bytes * (uint64_t)y;
    put_le(out, &off, 8);
  }
  out.reserve(out.size() + block_bytes * (size_t)H);
  std::vector<float> line(3 * (size_t)W);
  for (uint32_t y = 0; y < H; ++y) {
    put_i32(out, (int32_t)y);
    put_i32(out, (int32_t)line_bytes);
    for (int c = 0; c < 3; ++c)  // B, G, R planes of the line
      for (uint32_t x = 0; x < W; ++x) line[(size_t)c * W + x] = accum_rgba[4 * ((size_t)y * W + x) + (2 - c)] / (float)frames;
    put_le(out, line.data(), line_bytes);
  }
  return out;
}
}  // namespace

int cudapt_write_image(cudapt_ctx* ctx, const char* path) {
  if (!ctx || !path) return CUDAPT_ERR_INVALID;
  const size_t len = std::strlen(path);
  const bool ppm = len > 4 && std::strcmp(path + len - 4, ".ppm") == 0;
  const bool pfm = len > 4 && std::strcmp(path + len - 4, ".pfm") == 0;
  const bool png = len > 4 && std::strcmp(path + len - 4, ".png") == 0;
  const bool exr = len > 4 && std::strcmp(path + len - 4, ".exr") == 0;
  if (!ppm && !pfm && !png && !exr)
    return fail(ctx, CUDAPT_ERR_INVALID, "cudapt_write_image: '%s' is neither .png, .ppm, .pfm nor .exr", path);
  const uint32_t W = ctx->width, H = ctx->height;
  if (png) {
    const int rc = resolve_to_pinned(ctx, "cudapt_write_image");
    if (rc != CUDAPT_OK) return rc;
    const std::vector<unsigned char> bytes = encode_png_rgba8(ctx->h_pinned_rgba, W, H);
    FILE* f = std::fopen(path, "wb");
    if (!f) return fail(ctx, CUDAPT_ERR_INVALID, "cudapt_write_image: cannot open '%s'", path);
    const bool ok = std::fwrite(bytes.data(), 1, bytes.size(), f) == bytes.size();
    std::fclose(f);
    return ok ? CUDAPT_OK : fail(ctx, CUDAPT_ERR_INVALID, "cudapt_write_image: short write to '%s'", path);
  }
  if (ppm) {
    const int rc = resolve_to_pinned(ctx, "cudapt_write_image");
    if (rc != CUDAPT_OK) return rc;
    std::vector<unsigned char> rgb((size_t)W * H * 3);
    for (size_t i = 0; i < (size_t)W * H; ++i) {
      const uint32_t p = ctx->h_pinned_rgba[i];  // R<<24 | G<<16 | B<<8 | A (Color.h:9)
      rgb[3 * i + 0] = (unsigned char)(p >> 24);
      rgb[3 * i + 1] = (unsigned char)(p >> 16);
      rgb[3 * i + 2] = (unsigned char)(p >> 8);
    }
    FILE* f = std::fopen(path, "wb");
    if (!f) return fail(ctx, CUDAPT_ERR_INVALID, "cudapt_write_image: cannot open '%s'", path);
    std::fprintf(f, "P6\n%u %u\n255\n", W, H);
    const bool ok = std::fwrite(rgb.data(), 1, rgb.size(), f) == rgb.size();
    std::fclose(f);
    return ok ? CUDAPT_OK : fail(ctx, CUDAPT_ERR_INVALID, "cudapt_write_image: short write to '%s'", path);
  }
  if (!ctx->d_accum) return fail(ctx, CUDAPT_ERR_INVALID, "cudapt_write_image: resolution not set");
  if (ctx->frame_count == 0) return fail(ctx, CUDAPT_ERR_NO_FRAMES, "cudapt_write_image: no frames rendered yet");
  std::vector<float> acc((size_t)W * H * 4);
  uint32_t fc = 0;
  const int rc = cudapt_read_accum(ctx, acc.data(), &fc);
  if (rc != CUDAPT_OK) return rc;
  if (exr) {
    const std::vector<unsigned char> bytes = encode_exr_rgb32f(acc.data(), fc, W, H);
    FILE* f = std::fopen(path, "wb");
    if (!f) return fail(ctx, CUDAPT_ERR_INVALID, "cudapt_write_image: cannot open '%s'", path);
    const bool ok = std::fwrite(bytes.data(), 1, bytes.size(), f) == bytes.size();
    std::fclose(f);
    return ok ? CUDAPT_OK : fail(ctx, CUDAPT_ERR_INVALID, "cudapt_write_image: short write to '%s'", path);
  }
  std::vector<float> rgb((size_t)W * H * 3);
  for (uint32_t y = 0; y < H; ++y)  // PFM stores the bottom row first
    for (uint32_t x = 0; x < W; ++x)
      for (int c = 0; c < 3; ++c) rgb[3 * ((size_t)(H - 1 - y) * W + x) + c] = acc[4 * ((size_t)y * W + x) + c] / (float)fc;
  FILE* f = std::fopen(path, "wb");
  if (!f) return fail(ctx, CUDAPT_ERR_INVALID, "cudapt_write_image: cannot open '%s'", path);
  std::fprintf(f, "PF\n%u %u\n-1.0\n", W, H);
  const bool ok = std::fwrite(rgb.data(), sizeof(float), rgb.size(), f) == rgb.size();
  std::fclose(f);
  return ok ? CUDAPT_OK : fail(ctx, CUDAPT_ERR_INVALID, "cudapt_write_image: short write to '%s'", path);
}

int cudapt_read_accum(cudapt_ctx* ctx, float* host_out, uint32_t* frame_count) {
  if (!ctx) return CUDAPT_ERR_INVALID;
  if (!ctx->d_accum) return fail(ctx, CUDAPT_ERR_INVALID, "cudapt_read_accum: resolution not set");
  CU_TRY(ctx, cudaSetDevice(ctx->cfg.device));
  CU_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  if (host_out)
    CU_TRY(ctx, cudaMemcpy(host_out, ctx->d_accum, sizeof(float4) * (size_t)ctx->width * ctx->height, cudaMemcpyDeviceToHost));
  if (frame_count) *frame_count = ctx->frame_count;
  return finish_timing(ctx);
}

int cudapt_write_accum(cudapt_ctx* ctx, const float* host_in, uint32_t frame_count) {
  if (!ctx || !host_in) return CUDAPT_ERR_INVALID;
  if (!ctx->d_accum) return fail(ctx, CUDAPT_ERR_INVALID, "cudapt_write_accum: resolution not set");
  CU_TRY(ctx, cudaSetDevice(ctx->cfg.device));
  CU_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  CU_TRY(ctx, cudaMemcpy(ctx->d_accum, host_in, sizeof(float4) * (size_t)ctx->width * ctx->height, cudaMemcpyHostToDevice));
  ctx->frame_count = frame_count;
  return CUDAPT_OK;
}

int cudapt_read_primary_hits(cudapt_ctx* ctx, uint32_t* prim_id, float* t) {
  if (!ctx) return CUDAPT_ERR_INVALID;
  if (!ctx->d_aov_prim) return fail(ctx, CUDAPT_ERR_INVALID, "cudapt_read_primary_hits: resolution not set");
  if (ctx->frame_count == 0) return fail(ctx, CUDAPT_ERR_NO_FRAMES, "cudapt_read_primary_hits: no frames rendered yet");
  CU_TRY(ctx, cudaSetDevice(ctx->cfg.device));
  CU_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  const size_t n = (size_t)ctx->width * ctx->height;
  if (prim_id) CU_TRY(ctx, cudaMemcpy(prim_id, ctx->d_aov_prim, sizeof(uint32_t) * n, cudaMemcpyDeviceToHost));
  if (t) CU_TRY(ctx, cudaMemcpy(t, ctx->d_aov_t, sizeof(float) * n, cudaMemcpyDeviceToHost));
  return finish_timing(ctx);
}

int cudapt_get_stats(cudapt_ctx* ctx, cudapt_stats* out) {
  if (!ctx || !out) return CUDAPT_ERR_INVALID;
  CU_TRY(ctx, cudaSetDevice(ctx->cfg.device));
  CU_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  int rc = finish_timing(ctx);
  if (rc) return rc;
  unsigned long long c[kCntSlots];
  CU_TRY(ctx, cudaMemcpy(c, ctx->d_counters, sizeof(c), cudaMemcpyDeviceToHost));
  std::memset(out, 0, sizeof(*out));
  out->paths = c[kCntPaths];
  out->rays = c[kCntRays];
  out->nodes_visited = c[kCntNodes];
  out->tris_tested = c[kCntTris];
  out->packet_nodes = c[kCntPacketNodes];
  out->packet_tris = c[kCntPacketTris];
  out->packet_bails = c[kCntPacketBail];
  out->packet_rays = c[kCntPacketRays];
  out->shadow_rays = c[kCntShadowRays];
  out->last_packet_ms = ctx->last_packet_ms;
  out->packets = c[kCntPackets];
  out->shade_invocations = c[kCntShades];
  out->kernel_launches = ctx->launches;
  out->render_ms = ctx->render_ms;
  out->last_render_ms = ctx->last_render_ms;
  out->last_trace_ms = ctx->last_trace_ms;
  out->accel_build_ms = ctx->mesh.build_ms;
  out->accel_bytes = ctx->mesh.accel_bytes;
  out->frame_count = ctx->frame_count;
  out->n_triangles = ctx->mesh.n_triangles();
  out->n_nodes = ctx->mesh.n_nodes();
  out->n_spheres = ctx->n_spheres;
  return CUDAPT_OK;
}

int cudapt_set_counting(cudapt_ctx* ctx, int enabled) {
  if (!ctx) return CUDAPT_ERR_INVALID;
  ctx->counting = enabled != 0;
  return CUDAPT_OK;
}

int cudapt_set_tile_partition(cudapt_ctx* ctx, uint32_t rank, uint32_t world, uint32_t tile_w, uint32_t tile_h) {
  if (!ctx) return CUDAPT_ERR_INVALID;
  if (world == 0 || rank >= world || tile_w == 0 || tile_h == 0)
    return fail(ctx, CUDAPT_ERR_INVALID, "cudapt_set_tile_partition: rank %u world %u tile %ux%u", rank, world, tile_w, tile_h);
  ctx->part_rank = rank;
  ctx->part_world = world;
  ctx->tile_w = tile_w;
  ctx->tile_h = tile_h;
  ctx->mesh.invalidate_image_state();
  CU_TRY(ctx, cudaSetDevice(ctx->cfg.device));
  return restart_accumulation(ctx);
}

int cudapt_set_frame_range(cudapt_ctx* ctx, uint32_t first, uint32_t stride) {
  if (!ctx) return CUDAPT_ERR_INVALID;
  if (first == 0 || stride == 0) return fail(ctx, CUDAPT_ERR_INVALID, "cudapt_set_frame_range: first and stride must be >= 1");
  ctx->frame_first = first;
  ctx->frame_stride = stride;
  CU_TRY(ctx, cudaSetDevice(ctx->cfg.device));
  return restart_accumulation(ctx);
}

void* cudapt_accum_device_ptr(cudapt_ctx* ctx) { return ctx ? (void*)ctx->d_accum : nullptr; }
void* cudapt_stream(cudapt_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }

int cudapt_set_frame_count(cudapt_ctx* ctx, uint32_t frame_count) {
  if (!ctx) return CUDAPT_ERR_INVALID;
  ctx->frame_count = frame_count;
  return CUDAPT_OK;
}

}  // extern "C"

// ---- fused multi-GPU read-back tail ------------------------------------------------------------------------
int cudapt_reduce_resolve_peers(cudapt_ctx* ctx, const void* const* accum_by_rank, uint32_t n_ranks, uint32_t rank,
                                uint32_t total_frames, void* root_rgba, void* root_accum) {
  if (!ctx || !accum_by_rank || !root_rgba) return CUDAPT_ERR_INVALID;
  if (!ctx->d_accum) return fail(ctx, CUDAPT_ERR_INVALID, "cudapt_reduce_resolve_peers: resolution not set");
  if (n_ranks < 1 || n_ranks > 16 || rank >= n_ranks) return fail(ctx, CUDAPT_ERR_INVALID, "cudapt_reduce_resolve_peers: rank %u of %u", rank, n_ranks);
  if (total_frames == 0) return fail(ctx, CUDAPT_ERR_NO_FRAMES, "cudapt_reduce_resolve_peers: no frames");
  if (ctx->post.auto_exposure) return fail(ctx, CUDAPT_ERR_INVALID, "cudapt_reduce_resolve_peers: auto-exposure needs the whole image; resolve on the root instead");
  for (uint32_t g = 0; g < n_ranks; ++g)
    if (!accum_by_rank[g]) return fail(ctx, CUDAPT_ERR_INVALID, "cudapt_reduce_resolve_peers: accum_by_rank[%u] is NULL", g);
  CU_TRY(ctx, cudaSetDevice(ctx->cfg.device));
  const uint64_t n = (uint64_t)ctx->width * ctx->height;
  const uint32_t begin = (uint32_t)(n * rank / n_ranks), end = (uint32_t)(n * (rank + 1) / n_ranks);
  CU_TRY(ctx, launch_reduce_resolve_peers(reinterpret_cast<const float4* const*>(accum_by_rank), (int)n_ranks, begin, end, total_frames,
                                          ctx->post.exposure, ctx->post.tonemap, ctx->post.srgb, static_cast<uint32_t*>(root_rgba),
                                          static_cast<float4*>(root_accum), ctx->sm_count, ctx->stream));
  ctx->launches += 1;
  return CUDAPT_OK;
}

void* cudapt_rgba_device_ptr(cudapt_ctx* ctx) { return ctx ? ctx->d_rgba : nullptr; }

int cudapt_download_rgba8(cudapt_ctx* ctx, uint32_t* host_out) {
  if (!ctx || !host_out) return CUDAPT_ERR_INVALID;
  if (!ctx->d_rgba) return fail(ctx, CUDAPT_ERR_INVALID, "cudapt_download_rgba8: resolution not set");
  CU_TRY(ctx, cudaSetDevice(ctx->cfg.device));
  const size_t n = (size_t)ctx->width * ctx->height;
  const bool direct = host_is_pinned(host_out);
  CU_TRY(ctx, cudaMemcpyAsync(direct ? host_out : ctx->h_pinned_rgba, ctx->d_rgba, sizeof(uint32_t) * n, cudaMemcpyDeviceToHost, ctx->stream));
  CU_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  if (!direct) std::memcpy(host_out, ctx->h_pinned_rgba, sizeof(uint32_t) * n);
  return finish_timing(ctx);
}

static_assert(sizeof(cudaIpcMemHandle_t) == CUDAPT_IPC_HANDLE_BYTES, "IPC handle size");
int cudapt_ipc_export(cudapt_ctx* ctx, void* accum_handle, void* rgba_handle) {
  if (!ctx) return CUDAPT_ERR_INVALID;
  if (!ctx->d_accum) return fail(ctx, CUDAPT_ERR_INVALID, "cudapt_ipc_export: resolution not set");
  CU_TRY(ctx, cudaSetDevice(ctx->cfg.device));
  if (accum_handle) CU_TRY(ctx, cudaIpcGetMemHandle(static_cast<cudaIpcMemHandle_t*>(accum_handle), ctx->d_accum));
  if (rgba_handle) CU_TRY(ctx, cudaIpcGetMemHandle(static_cast<cudaIpcMemHandle_t*>(rgba_handle), ctx->d_rgba));
  return CUDAPT_OK;
}
int cudapt_ipc_open(cudapt_ctx* ctx, const void* handle, void** device_ptr) {
  if (!ctx || !handle || !device_ptr) return CUDAPT_ERR_INVALID;
  CU_TRY(ctx, cudaSetDevice(ctx->cfg.device));
  cudaIpcMemHandle_t h;
  std::memcpy(&h, handle, sizeof(h));
  CU_TRY(ctx, cudaIpcOpenMemHandle(device_ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return CUDAPT_OK;
}
int cudapt_ipc_close(cudapt_ctx* ctx, void* device_ptr) {
  if (!ctx || !device_ptr) return CUDAPT_ERR_INVALID;
  CU_TRY(ctx, cudaSetDevice(ctx->cfg.device));
  CU_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  CU_TRY(ctx, cudaIpcCloseMemHandle(device_ptr));
  return CUDAPT_OK;
}

int cudapt_debug_accel_checksum(cudapt_ctx* ctx, uint64_t* nodes_sum, uint64_t* tris_sum) {
  if (!ctx || !nodes_sum || !tris_sum) return CUDAPT_ERR_INVALID;
  if (ctx->scene != cudapt_ctx::kMesh) return fail(ctx, CUDAPT_ERR_NO_SCENE, "cudapt_debug_accel_checksum: no built mesh scene");
  CU_TRY(ctx, cudaSetDevice(ctx->cfg.device));
  std::string err;
  if (!ctx->mesh.accel_checksum(ctx->stream, nodes_sum, tris_sum, &err)) return fail(ctx, CUDAPT_ERR_CUDA, "cudapt_debug_accel_checksum: %s", err.c_str());
  return CUDAPT_OK;
}
