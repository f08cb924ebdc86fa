// TEST INFRASTRUCTURE (oracle, Tier 1) -- never linked into the product.
//
// Headless driver for the UNMODIFIED reference render library.  It makes the
// same five interface calls src/App.cpp makes (create_path_tracer :98,
// set_settings :129, set_scene :130, render() :231, get_render_result() :232)
// and exposes the outcome through a plain C interface for ctypes.
//
// The reference sources are compiled where they lie under /root/reference by
// oracle/Makefile; only this file and the Embree stand-in are ours.  This TU is
// compiled with -fno-access-control so it can read the backend's private float
// accumulation buffer and frame counter (CPUPathTracer.h:68,72) without
// editing reference code.
#include "engines/pathtracer/backends/cpu/CPUPathTracer.h"
#include "render/Log.h"
#include "render/PathTracer.h"
#include "render/Scene.h"
#include "render/Types.h"

#include <chrono>
#include <cstdint>
#include <cstring>
#include <iostream>
#include <memory>
#include <unordered_map>
#include <vector>

extern "C" {
extern uint64_t g_shim_ray_count;
extern void** g_shim_primary_user;
extern float* g_shim_primary_t;
extern uint64_t g_shim_primary_cap;
extern uint64_t g_shim_primary_n;
}

namespace {
struct CoutMute {
  std::streambuf* old;
  CoutMute() : old(std::cout.rdbuf(nullptr)) {}
  ~CoutMute() {
    std::cout.rdbuf(old);
    std::cout.clear();
  }
};

struct RefSession {
  std::unique_ptr<render::PathTracer> tracer;
  std::shared_ptr<render::Scene> scene;
  std::shared_ptr<render::RenderSettings> settings;
  std::unordered_map<const void*, int32_t> index_of;  // SphereObject* -> creation index
  std::vector<int32_t> primary_idx;
  std::vector<float> primary_t;
  uint64_t rays = 0;
  uint32_t w = 0, h = 0;
  bool first = true;
};
}  // namespace

extern "C" {

// spheres: n x {cx, cy, cz, r}, created in this order (App.cpp:101-122 style).
void* ref_create(const float* spheres, int n, int width, int height) {
  CoutMute mute;
  render::Log::set_callback([](render::LogLevel, std::string_view) {});
  auto* s = new RefSession();
  s->tracer = render::PathTracer::create_path_tracer(render::PathTracer::BackendType::CPU_EMBREE);
  s->scene = std::make_shared<render::Scene>();
  for (int i = 0; i < n; ++i) {
    auto* node = s->scene->CreateNode<render::SphereObject>("sphere");
    node->SetRadius(spheres[4 * i + 3]);
    node->SetPosition(glm::vec3(spheres[4 * i + 0], spheres[4 * i + 1], spheres[4 * i + 2]));
    s->index_of[static_cast<const void*>(static_cast<const render::SphereObject*>(node))] = i;
  }
  s->settings = std::make_shared<render::RenderSettings>();
  s->settings->setResolution((uint32_t)width, (uint32_t)height);
  s->tracer->set_settings(s->settings);
  s->tracer->set_scene(s->scene);
  s->w = (uint32_t)width;
  s->h = (uint32_t)height;
  return s;
}

// Runs `frames` render() calls; when resolve_each_frame != 0 also calls
// get_render_result() after each one, as App::run does.  Returns wall seconds.
double ref_render(void* handle, int frames, int resolve_each_frame) {
  auto* s = static_cast<RefSession*>(handle);
  CoutMute mute;
  std::vector<void*> users;
  if (s->first) {
    users.assign((size_t)s->w * s->h, nullptr);
    s->primary_t.assign((size_t)s->w * s->h, 0.0f);
    g_shim_primary_user = users.data();
    g_shim_primary_t = s->primary_t.data();
    g_shim_primary_cap = users.size();
    g_shim_primary_n = 0;
  }
  const uint64_t rays0 = g_shim_ray_count;
  const auto t0 = std::chrono::steady_clock::now();
  for (int f = 0; f < frames; ++f) {
    s->tracer->render();
    if (resolve_each_frame) (void)s->tracer->get_render_result();
    if (s->first) {  // primary rays of frame 1 captured; stop capturing
      g_shim_primary_user = nullptr;
      g_shim_primary_cap = 0;
      s->primary_idx.resize(users.size());
      for (size_t i = 0; i < users.size(); ++i)
        s->primary_idx[i] = users[i] ? s->index_of.at(static_cast<const void*>(users[i])) : -1;
      s->first = false;
    }
  }
  const auto t1 = std::chrono::steady_clock::now();
  s->rays += g_shim_ray_count - rays0;
  return std::chrono::duration<double>(t1 - t0).count();
}

// accum: 4*W*H floats (private m_accumulation_buffer); rgba: W*H (get_render_result()).
void ref_read(void* handle, float* accum, uint32_t* rgba, uint32_t* frame_count) {
  auto* s = static_cast<RefSession*>(handle);
  auto* cpu = static_cast<render::CPUPathTracer*>(s->tracer.get());
  if (accum) std::memcpy(accum, cpu->m_accumulation_buffer.data(), sizeof(float) * cpu->m_accumulation_buffer.size());
  if (frame_count) *frame_count = cpu->m_frameCount;
  if (rgba) {
    const auto& res = s->tracer->get_render_result();
    std::memcpy(rgba, res.image_buffer.data(), sizeof(uint32_t) * res.image_buffer.size());
  }
}

// Primary-hit AOV captured by the stand-in during frame 1: creation index of the sphere (-1 = miss) and t.
void ref_primary(void* handle, int32_t* idx, float* t) {
  auto* s = static_cast<RefSession*>(handle);
  std::memcpy(idx, s->primary_idx.data(), sizeof(int32_t) * s->primary_idx.size());
  std::memcpy(t, s->primary_t.data(), sizeof(float) * s->primary_t.size());
}

uint64_t ref_ray_count(void* handle) { return static_cast<RefSession*>(handle)->rays; }

void ref_destroy(void* handle) {
  CoutMute mute;
  delete static_cast<RefSession*>(handle);
}

const char* ref_backend_name(void* handle) {
  static std::string name;
  name = static_cast<RefSession*>(handle)->tracer->get_backend_name();
  return name.c_str();
}
}
