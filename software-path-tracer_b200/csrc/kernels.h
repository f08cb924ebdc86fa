// Host-visible launch interfaces of the CUDA kernels (internal to libcudapt.so).
#pragma once
#include "common.cuh"

namespace cudapt {

enum CounterSlot : int {
  kCntRays = 0,
  kCntShades = 1,
  kCntPaths = 2,
  kCntNodes = 3,
  kCntTris = 4,
  kCntSlots = 8
};

constexpr int kSphereBlockX = 16;
constexpr int kSphereBlockY = 8;

struct SphereLaunch {
  const float4* spheres;    // {cx, cy, cz, r}
  const uint32_t* node_id;  // caller id per sphere (render::NodeID)
  int n_spheres;
  float4* accum;            // RGBA float accumulation, W*H
  uint32_t* aov_prim;       // primary-hit id of the launch's first frame, or nullptr
  float* aov_t;
  unsigned long long* counters;
  uint32_t first_frame, n_frames, frame_stride;  // 1-based global frame indices (CPT.cpp:61)
  CameraParams cam;
  ImageParams im;
  IntegratorParams ip;
};

cudaError_t launch_render_spheres(const SphereLaunch& a, bool philox, cudaStream_t stream);
cudaError_t launch_resolve_rgba8(const float4* accum, uint32_t* out, uint32_t n_pixels, uint32_t frame_count,
                                 cudaStream_t stream);

}  // namespace cudapt
