// Host side of the mesh (triangle) scene: staging of the uploaded mesh, BVH build + upload, and the
// wavefront render driver.  Internal to libcudapt.so.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "../../include/cudapt.h"
#include "kernels.h"

namespace cudapt {

struct MeshRenderArgs {
  float4* accum;
  uint32_t* aov_prim;
  float* aov_t;
  unsigned long long* counters;
  uint32_t first_frame, n_frames, frame_stride;
  CameraParams cam;
  ImageParams im;
  IntegratorParams ip;
  uint32_t samples_per_pass;
  bool counting;
};

struct MeshDevice;  // device buffers (mesh_host.cu)

class MeshScene {
 public:
  int sm_count = 148;
  double build_ms = 0.0;
  uint64_t accel_bytes = 0;

  void set_mesh(const float* vertices, uint32_t nv, const uint32_t* indices, uint32_t nt, const uint32_t* material_id);
  void set_materials(const cudapt_material* m, uint32_t n);
  bool has_host_mesh() const { return !indices_.empty(); }
  bool build_and_upload(std::string* err);
  bool build_on_device(cudaStream_t stream, uint64_t* launches, std::string* err);  // lbvh_build.cu
  // Same topology, moved vertices: refit the BVH on the host (no rebuild) and upload it again.
  // (on the device, refit.cu; CUDAPT_REFIT_HOST=1 selects the host implementation of the same code).
  bool update_vertices(const float* vertices, uint32_t nv, cudaStream_t stream, uint64_t* launches, std::string* err);
  bool accel_checksum(cudaStream_t stream, uint64_t* nodes_sum, uint64_t* tris_sum, std::string* err);
  double refit_ms = 0.0;
  // H2D copy of the packed scene (nodes, triangles, materials, lights) from pinned host memory.
  bool reupload(cudaStream_t stream, std::string* err);
  // double-buffered upload (cudapt_stage_scene)
  bool stage(cudaStream_t render_stream, std::string* err);
  bool swap_pending() const;
  bool commit_staged(cudaStream_t stream, std::string* err);
  cudaError_t render(const MeshRenderArgs& a, cudaStream_t stream, uint64_t* launches, std::string* err);
  void invalidate_image_state();
  void release();
  uint32_t n_triangles() const { return n_tris_built_; }
  uint32_t n_nodes() const { return n_nodes_built_; }
  // CUDA-event time of the closest-hit traversal launches of the last render() (call after a stream sync).
  double collect_trace_ms();
  double collect_packet_ms();

 private:
  bool refresh_staging(cudaStream_t stream, std::string* err);
  std::vector<float> vertices_;
  std::vector<uint32_t> indices_;
  std::vector<uint32_t> material_id_;
  std::vector<cudapt_material> materials_;
  uint32_t n_tris_built_ = 0, n_nodes_built_ = 0;
  MeshDevice* dev_ = nullptr;
  void next_trace_events(cudaEvent_t* e0, cudaEvent_t* e1);
  std::vector<cudaEvent_t> trace_events_;
  size_t trace_events_used_ = 0;
  std::vector<cudaEvent_t> packet_events_;  // pairs around the camera-ray packet launches
  size_t packet_events_used_ = 0;
};

}  // namespace cudapt
