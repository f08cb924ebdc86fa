// Host driver of the mesh scene: BVH build + upload, wavefront state, per-pass launch sequence.
#include "mesh_host.h"
#include "lbvh_build.h"

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>

#include "bvh8.h"

namespace cudapt {

struct MeshDevice {
  // scene (pinned host staging kept for re-upload)
  Bvh8 bvh;
  std::vector<uint32_t> lights;
  std::vector<cudapt_material> materials;
  uint4* d_nodes = nullptr;
  float4* d_tris = nullptr;
  float4* d_materials = nullptr;
  uint32_t* d_lights = nullptr;
  void* h_pinned = nullptr;  // nodes | tris, pinned, for cudapt_reupload_scene
  size_t pinned_bytes = 0;
  bool staging_stale = false;  // the device copy was refitted in place; refresh h_pinned before re-uploading from it
  // double-buffered upload (cudapt_stage_scene): the NEXT version of nodes | tris is copied on its own stream while
  // the current one is being rendered from; the buffers swap at the next render
  uint4* d_nodes_shadow = nullptr;
  float4* d_tris_shadow = nullptr;
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t ev_copy_done = nullptr;
  cudaEvent_t ev_last_use[2] = {nullptr, nullptr};  // last render that read buffer 0 / 1 (0 = the one built first)
  int current = 0;
  bool swap_pending = false;
  // device refit (refit.cu): caller-order vertices and indices, per-triangle and per-node boxes
  float* d_vertices = nullptr;
  uint32_t* d_indices = nullptr;
  void* d_tri_box = nullptr;
  void* d_node_box = nullptr;
  void* h_pinned_vertices = nullptr;
  uint32_t* d_material_id = nullptr;           // device build only
  LbvhScratch lbvh_scratch;                    // device build only: the builder's temporaries, kept between builds
  uint32_t device_cap_nt = 0, device_cap_nv = 0;  // sizes the device-build buffers were allocated for (0: host build)
  // compacted tile partition (kernels.h PassParams::owned_tiles): the 8x8 slot tiles this rank owns, cached per key
  uint32_t* d_owned_tiles = nullptr;
  uint32_t n_owned_tiles = 0;
  uint32_t owned_key[6] = {0, 0, 0, 0, 0, 0};  // width, height, rank, world, tile_w, tile_h
  // wavefront state
  WavefrontState st{};
  uint32_t state_paths = 0;     // capacity in paths
  uint32_t state_bounces = 0;   // counter slots allocated for this many bounces

  // Copy stream and events of the double-buffered upload.  Created when the scene is BUILT (not at the first stage()),
  // so that every render -- also those issued before the first cudapt_stage_scene -- records its last use of the
  // buffer it read: a stage() that follows waits for that render before it overwrites the other buffer.
  cudaError_t ensure_copy_objects() {
    cudaError_t e;
    if (!copy_stream && (e = cudaStreamCreateWithFlags(&copy_stream, cudaStreamNonBlocking)) != cudaSuccess) return e;
    if (!ev_copy_done && (e = cudaEventCreateWithFlags(&ev_copy_done, cudaEventDisableTiming)) != cudaSuccess) return e;
    for (cudaEvent_t& ev : ev_last_use)
      if (!ev && (e = cudaEventCreateWithFlags(&ev, cudaEventDisableTiming)) != cudaSuccess) return e;
    return cudaSuccess;
  }

  void free_state() {
    cudaFree(st.ray_o), cudaFree(st.ray_d), cudaFree(st.thr), cudaFree(st.L), cudaFree(st.hit);
    cudaFree(st.sh_o), cudaFree(st.sh_d), cudaFree(st.sh_c);
    cudaFree(st.queue[0]), cudaFree(st.queue[1]), cudaFree(st.shadow_queue), cudaFree(st.counters);
    st = WavefrontState{};
    state_paths = 0;
    state_bounces = 0;
  }
  // NOT part of free_state(): render() builds the list and then grows the path state (ensure_state -> free_state), and
  // the list only depends on (resolution, partition), which its key already covers
  void free_partition() {
    cudaFree(d_owned_tiles);
    d_owned_tiles = nullptr;
    n_owned_tiles = 0;
    owned_key[3] = 0;
  }
  void free_scene() {
    cudaFree(d_nodes), cudaFree(d_tris), cudaFree(d_materials), cudaFree(d_lights);
    d_nodes = nullptr, d_tris = nullptr, d_materials = nullptr, d_lights = nullptr;
    cudaFree(d_nodes_shadow), cudaFree(d_tris_shadow);
    d_nodes_shadow = nullptr, d_tris_shadow = nullptr;
    if (copy_stream) cudaStreamDestroy(copy_stream);
    if (ev_copy_done) cudaEventDestroy(ev_copy_done);
    for (cudaEvent_t& e : ev_last_use) {
      if (e) cudaEventDestroy(e);
      e = nullptr;
    }
    copy_stream = nullptr, ev_copy_done = nullptr, current = 0, swap_pending = false;
    cudaFree(d_vertices), cudaFree(d_indices), cudaFree(d_tri_box), cudaFree(d_node_box), cudaFree(d_material_id);
    d_vertices = nullptr, d_indices = nullptr, d_tri_box = nullptr, d_node_box = nullptr, d_material_id = nullptr;
    device_cap_nt = device_cap_nv = 0;
    cudaFree(lbvh_scratch.p);
    lbvh_scratch = LbvhScratch{};
    if (h_pinned_vertices) cudaFreeHost(h_pinned_vertices);
    h_pinned_vertices = nullptr;
    staging_stale = false;
    if (h_pinned) cudaFreeHost(h_pinned);
    h_pinned = nullptr;
    pinned_bytes = 0;
  }
};

void MeshScene::set_mesh(const float* v, uint32_t nv, const uint32_t* idx, uint32_t nt, const uint32_t* mid) {
  vertices_.assign(v, v + 3 * (size_t)nv);
  indices_.assign(idx, idx + 3 * (size_t)nt);
  if (mid)
    material_id_.assign(mid, mid + nt);
  else
    material_id_.assign(nt, 0u);
}

void MeshScene::set_materials(const cudapt_material* m, uint32_t n) { materials_.assign(m, m + n); }

void MeshScene::invalidate_image_state() {
  if (dev_) dev_->free_state();
}

void MeshScene::release() {
  for (cudaEvent_t e : trace_events_) cudaEventDestroy(e);
  trace_events_.clear();
  trace_events_used_ = 0;
  for (cudaEvent_t e : packet_events_) cudaEventDestroy(e);
  packet_events_.clear();
  packet_events_used_ = 0;
  if (dev_) {
    dev_->free_state();
    dev_->free_scene();
    dev_->free_partition();
    delete dev_;
    dev_ = nullptr;
  }
}

#define MH_TRY(expr)                                                    \
  do {                                                                  \
    cudaError_t e__ = (expr);                                           \
    if (e__ != cudaSuccess) {                                           \
      *err = std::string(#expr) + ": " + cudaGetErrorString(e__);       \
      return false;                                                     \
    }                                                                   \
  } while (0)

bool MeshScene::build_and_upload(std::string* err) {
  const auto t0 = std::chrono::steady_clock::now();
  if (!dev_) dev_ = new MeshDevice();
  dev_->free_scene();
  const uint32_t nt = (uint32_t)(indices_.size() / 3), nv = (uint32_t)(vertices_.size() / 3);
  std::vector<cudapt_material> mats = materials_;
  if (mats.empty()) {  // default: the reference's constant 0.7 albedo (CPT.cpp:260) as a Lambert material
    cudapt_material d;
    std::memset(&d, 0, sizeof(d));
    d.albedo[0] = d.albedo[1] = d.albedo[2] = 0.7f;
    mats.push_back(d);
  }
  int threads = (int)std::thread::hardware_concurrency();
  threads = std::max(1, std::min(threads, 32));
  if (!build_bvh8(vertices_.data(), nv, indices_.data(), nt, material_id_.data(), threads, &dev_->bvh, err)) return false;
  const Bvh8& bvh = dev_->bvh;
  // emissive triangles, ascending caller index (the order the NEE light pick is defined on)
  dev_->lights.clear();
  for (uint32_t i = 0; i < nt; ++i) {
    uint32_t m = material_id_[i];
    if (m >= mats.size()) m = 0;
    if (mats[m].emission[0] > 0.f || mats[m].emission[1] > 0.f || mats[m].emission[2] > 0.f)
      dev_->lights.push_back(bvh.tri_of_prim[i]);
  }
  dev_->materials = mats;
  const size_t nb = bvh.nodes.size() * sizeof(Bvh8Node), tb = bvh.tris.size() * sizeof(PackedTri);
  MH_TRY(cudaMallocHost(&dev_->h_pinned, nb + tb));
  dev_->pinned_bytes = nb + tb;
  std::memcpy(dev_->h_pinned, bvh.nodes.data(), nb);
  std::memcpy((char*)dev_->h_pinned + nb, bvh.tris.data(), tb);
  MH_TRY(cudaMalloc(&dev_->d_nodes, nb));
  MH_TRY(cudaMalloc(&dev_->d_tris, tb));
  MH_TRY(cudaMalloc(&dev_->d_materials, mats.size() * sizeof(cudapt_material)));
  MH_TRY(cudaMalloc(&dev_->d_lights, std::max<size_t>(dev_->lights.size(), 1) * sizeof(uint32_t)));
  MH_TRY(dev_->ensure_copy_objects());
  n_tris_built_ = nt;
  n_nodes_built_ = (uint32_t)bvh.nodes.size();
  accel_bytes = nb + tb;
  if (!reupload(nullptr, err)) return false;
  MH_TRY(cudaDeviceSynchronize());
  build_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
  return true;
}

// Device build (lbvh_build.cu): the mesh goes up as it is (vertices, indices, material ids), the tree is built and
// quantised on the GPU, and nodes | tris come BACK into the pinned staging and the host Bvh8 (for re-uploads, the host
// refit path and the light list).  The device arrays the refit needs (vertices, indices, boxes) stay allocated.
bool MeshScene::build_on_device(cudaStream_t stream, uint64_t* launches, std::string* err) {
  const auto t0 = std::chrono::steady_clock::now();
  const bool dbg = std::getenv("CUDAPT_BUILD_DEBUG") != nullptr;
  auto mark = [&](const char* what) {
    if (dbg) std::fprintf(stderr, "[device build] %-28s %8.2f ms\n", what, std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
  };
  if (!dev_) dev_ = new MeshDevice();
  const uint32_t nt = (uint32_t)(indices_.size() / 3), nv = (uint32_t)(vertices_.size() / 3);
  const uint32_t node_cap = nt / 2u + 8u;
  const size_t vb = sizeof(float) * 3 * (size_t)nv, ib = sizeof(uint32_t) * 3 * (size_t)nt;
  const size_t nb_cap = sizeof(Bvh8Node) * (size_t)node_cap, tb = sizeof(PackedTri) * (size_t)nt;
  // A REbuild of a mesh of the same size (the per-frame case) keeps every allocation: pinned memory alone costs ~20 ms
  // to allocate at 1 M triangles, three times the build itself.
  const bool reuse = dev_->device_cap_nt == nt && dev_->device_cap_nv == nv && dev_->d_nodes && dev_->h_pinned;
  if (!reuse) {
    dev_->free_scene();
    MH_TRY(cudaMalloc(&dev_->d_vertices, vb));
    MH_TRY(cudaMalloc(&dev_->d_indices, ib));
    MH_TRY(cudaMalloc(&dev_->d_tri_box, 24 * (size_t)nt));
    MH_TRY(cudaMalloc(&dev_->d_node_box, 24 * (size_t)node_cap));
    MH_TRY(cudaMallocHost(&dev_->h_pinned_vertices, vb));
    MH_TRY(cudaMalloc(&dev_->d_nodes, nb_cap));
    MH_TRY(cudaMalloc(&dev_->d_tris, tb));
    MH_TRY(cudaMalloc(&dev_->d_material_id, sizeof(uint32_t) * (size_t)nt));
    MH_TRY(cudaMallocHost(&dev_->h_pinned, nb_cap + tb));  // sized for the node capacity: the node count changes per build
    dev_->pinned_bytes = nb_cap + tb;
    dev_->device_cap_nt = nt;
    dev_->device_cap_nv = nv;
    MH_TRY(dev_->ensure_copy_objects());
  } else {
    MH_TRY(cudaStreamSynchronize(stream));
    dev_->swap_pending = false;  // a staged copy of the previous tree is dropped (its buffer is capacity-sized, see stage)
    cudaFree(dev_->d_materials), cudaFree(dev_->d_lights);
    dev_->d_materials = nullptr, dev_->d_lights = nullptr;
  }
  std::vector<cudapt_material> mats = materials_;
  if (mats.empty()) {
    cudapt_material d;
    std::memset(&d, 0, sizeof(d));
    d.albedo[0] = d.albedo[1] = d.albedo[2] = 0.7f;
    mats.push_back(d);
  }
  Bvh8& bvh = dev_->bvh;
  for (int k = 0; k < 3; ++k) bvh.world_lo[k] = INFINITY, bvh.world_hi[k] = -INFINITY;
  for (size_t i = 0; i < (size_t)nv; ++i)
    for (int k = 0; k < 3; ++k) {
      bvh.world_lo[k] = std::min(bvh.world_lo[k], vertices_[3 * i + k]);
      bvh.world_hi[k] = std::max(bvh.world_hi[k], vertices_[3 * i + k]);
    }
  std::memcpy(dev_->h_pinned_vertices, vertices_.data(), vb);
  MH_TRY(cudaMemcpyAsync(dev_->d_vertices, dev_->h_pinned_vertices, vb, cudaMemcpyHostToDevice, stream));
  MH_TRY(cudaMemcpyAsync(dev_->d_indices, indices_.data(), ib, cudaMemcpyHostToDevice, stream));
  MH_TRY(cudaMemcpyAsync(dev_->d_material_id, material_id_.data(), sizeof(uint32_t) * (size_t)nt, cudaMemcpyHostToDevice, stream));
  mark("allocations + mesh upload");
  LbvhResult r;
  if (!build_lbvh_device(dev_->d_vertices, dev_->d_indices, dev_->d_material_id, nt, bvh.world_lo, bvh.world_hi, dev_->d_nodes,
                         node_cap, dev_->d_tris, dev_->d_tri_box, dev_->d_node_box, &r, &dev_->lbvh_scratch, launches, stream, err))
    return false;
  mark("tree built on the device");
  bvh.nodes.resize(r.n_nodes);
  bvh.tris.resize(nt);
  bvh.level_start = r.level_start;
  bvh.max_depth = r.max_depth;
  bvh.sah_cost = 0.0;
  // The tree lives on the device; the pinned staging and the host copy follow on demand (re-upload, staging, host
  // refit: refresh_staging) -- except that a scene with emissive materials needs the leaf order of its lights now.
  dev_->staging_stale = true;
  dev_->lights.clear();
  bool any_emissive = false;
  for (const cudapt_material& m : mats) any_emissive = any_emissive || m.emission[0] > 0.f || m.emission[1] > 0.f || m.emission[2] > 0.f;
  if (any_emissive) {
    if (!refresh_staging(stream, err)) return false;
    bvh.tri_of_prim.assign(nt, 0u);
    for (uint32_t k = 0; k < nt; ++k) {
      if (bvh.tris[k].prim >= nt) {
        *err = "internal error: device build produced an invalid triangle record";
        return false;
      }
      bvh.tri_of_prim[bvh.tris[k].prim] = k;
    }
    for (uint32_t i = 0; i < nt; ++i) {
      uint32_t m = material_id_[i];
      if (m >= mats.size()) m = 0;
      if (mats[m].emission[0] > 0.f || mats[m].emission[1] > 0.f || mats[m].emission[2] > 0.f)
        dev_->lights.push_back(bvh.tri_of_prim[i]);
    }
    mark("read-back + light list");
  }
  dev_->materials = mats;
  MH_TRY(cudaMalloc(&dev_->d_materials, mats.size() * sizeof(cudapt_material)));
  MH_TRY(cudaMalloc(&dev_->d_lights, std::max<size_t>(dev_->lights.size(), 1) * sizeof(uint32_t)));
  MH_TRY(cudaMemcpyAsync(dev_->d_materials, mats.data(), mats.size() * sizeof(cudapt_material), cudaMemcpyHostToDevice, stream));
  if (!dev_->lights.empty())
    MH_TRY(cudaMemcpyAsync(dev_->d_lights, dev_->lights.data(), dev_->lights.size() * sizeof(uint32_t), cudaMemcpyHostToDevice,
                           stream));
  MH_TRY(cudaStreamSynchronize(stream));
  n_tris_built_ = nt;
  n_nodes_built_ = r.n_nodes;
  accel_bytes = bvh.nodes.size() * sizeof(Bvh8Node) + tb;
  build_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
  mark("done");
  return true;
}

bool MeshScene::reupload(cudaStream_t stream, std::string* err) {
  if (!dev_ || !dev_->h_pinned) {
    *err = "no built scene";
    return false;
  }
  if (dev_->staging_stale && !refresh_staging(stream, err)) return false;
  const size_t nb = dev_->bvh.nodes.size() * sizeof(Bvh8Node), tb = dev_->bvh.tris.size() * sizeof(PackedTri);
  MH_TRY(cudaMemcpyAsync(dev_->d_nodes, dev_->h_pinned, nb, cudaMemcpyHostToDevice, stream));
  MH_TRY(cudaMemcpyAsync(dev_->d_tris, (char*)dev_->h_pinned + nb, tb, cudaMemcpyHostToDevice, stream));
  MH_TRY(cudaMemcpyAsync(dev_->d_materials, dev_->materials.data(), dev_->materials.size() * sizeof(cudapt_material),
                         cudaMemcpyHostToDevice, stream));
  if (!dev_->lights.empty())
    MH_TRY(cudaMemcpyAsync(dev_->d_lights, dev_->lights.data(), dev_->lights.size() * sizeof(uint32_t), cudaMemcpyHostToDevice,
                           stream));
  return true;
}

// Double-buffered upload: start copying nodes | tris from the pinned staging into the SHADOW device copy on the copy
// stream -- concurrently with whatever the render stream is doing -- and make it the scene at the next render.
bool MeshScene::stage(cudaStream_t render_stream, std::string* err) {
  if (!dev_ || !dev_->h_pinned) {
    *err = "no built scene";
    return false;
  }
  if (dev_->staging_stale && !refresh_staging(render_stream, err)) return false;
  const size_t nb = dev_->bvh.nodes.size() * sizeof(Bvh8Node), tb = dev_->bvh.tris.size() * sizeof(PackedTri);
  if (!dev_->d_nodes_shadow) {
    // after a device build the two node buffers trade places with every swap, and a rebuild writes up to the node
    // CAPACITY into whichever is current
    const size_t nb_alloc = dev_->device_cap_nt ? sizeof(Bvh8Node) * (size_t)(dev_->device_cap_nt / 2u + 8u) : nb;
    MH_TRY(cudaMalloc(&dev_->d_nodes_shadow, nb_alloc));
    MH_TRY(cudaMalloc(&dev_->d_tris_shadow, tb));
  }
  MH_TRY(dev_->ensure_copy_objects());
  // The shadow buffer may still be read by the last render that used it (never by the one in flight now): every
  // render() records ev_last_use[buffer it read] (the events exist from the build on; an event that was never
  // recorded -- the shadow buffer has not been rendered from yet -- is complete, which is then the truth).
  MH_TRY(cudaStreamWaitEvent(dev_->copy_stream, dev_->ev_last_use[dev_->current ^ 1], 0));
  MH_TRY(cudaMemcpyAsync(dev_->d_nodes_shadow, dev_->h_pinned, nb, cudaMemcpyHostToDevice, dev_->copy_stream));
  MH_TRY(cudaMemcpyAsync(dev_->d_tris_shadow, (char*)dev_->h_pinned + nb, tb, cudaMemcpyHostToDevice, dev_->copy_stream));
  MH_TRY(cudaEventRecord(dev_->ev_copy_done, dev_->copy_stream));
  dev_->swap_pending = true;
  return true;
}

bool MeshScene::swap_pending() const { return dev_ && dev_->swap_pending; }

// Makes the staged copy the scene: the render stream waits for the copy, the buffers swap, the small arrays
// (materials, light list) follow on the render stream.
bool MeshScene::commit_staged(cudaStream_t stream, std::string* err) {
  if (!dev_ || !dev_->swap_pending) return true;
  MH_TRY(cudaStreamWaitEvent(stream, dev_->ev_copy_done, 0));
  std::swap(dev_->d_nodes, dev_->d_nodes_shadow);
  std::swap(dev_->d_tris, dev_->d_tris_shadow);
  dev_->current ^= 1;
  dev_->swap_pending = false;
  MH_TRY(cudaMemcpyAsync(dev_->d_materials, dev_->materials.data(), dev_->materials.size() * sizeof(cudapt_material),
                         cudaMemcpyHostToDevice, stream));
  if (!dev_->lights.empty())
    MH_TRY(cudaMemcpyAsync(dev_->d_lights, dev_->lights.data(), dev_->lights.size() * sizeof(uint32_t), cudaMemcpyHostToDevice,
                           stream));
  return true;
}

bool MeshScene::update_vertices(const float* v, uint32_t nv, cudaStream_t stream, uint64_t* launches, std::string* err) {
  if (!dev_ || !dev_->h_pinned) {
    *err = "no built scene";
    return false;
  }
  if ((size_t)nv * 3 != vertices_.size()) {
    *err = "vertex count differs from the uploaded mesh (topology changes need cudapt_upload_mesh + cudapt_build_accel)";
    return false;
  }
  const auto t0 = std::chrono::steady_clock::now();
  vertices_.assign(v, v + 3 * (size_t)nv);
  Bvh8& bvh = dev_->bvh;
  const uint32_t nt = (uint32_t)(indices_.size() / 3);
  const char* force_host = std::getenv("CUDAPT_REFIT_HOST");
  if (force_host && force_host[0] == '1') {
    // host refit (the same code, bvh8_refit.h): kept as the cross-check of the device kernels
    if (dev_->staging_stale && !refresh_staging(stream, err)) return false;
    if (!refit_bvh8(vertices_.data(), nv, indices_.data(), nt, &bvh, err)) return false;
    const size_t nb = bvh.nodes.size() * sizeof(Bvh8Node), tb = bvh.tris.size() * sizeof(PackedTri);
    std::memcpy(dev_->h_pinned, bvh.nodes.data(), nb);
    std::memcpy((char*)dev_->h_pinned + nb, bvh.tris.data(), tb);
    refit_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    build_ms = refit_ms;
    return reupload(stream, err);
  }
  // device refit: only the vertex positions cross PCIe
  const size_t vb = sizeof(float) * 3 * (size_t)nv;
  if (!dev_->d_vertices) {
    MH_TRY(cudaMalloc(&dev_->d_vertices, vb));
    MH_TRY(cudaMalloc(&dev_->d_indices, sizeof(uint32_t) * indices_.size()));
    MH_TRY(cudaMalloc(&dev_->d_tri_box, 24 * (size_t)nt));
    MH_TRY(cudaMalloc(&dev_->d_node_box, 24 * bvh.nodes.size()));
    MH_TRY(cudaMallocHost(&dev_->h_pinned_vertices, vb));
    MH_TRY(cudaMemcpyAsync(dev_->d_indices, indices_.data(), sizeof(uint32_t) * indices_.size(), cudaMemcpyHostToDevice, stream));
  }
  std::memcpy(dev_->h_pinned_vertices, vertices_.data(), vb);
  MH_TRY(cudaMemcpyAsync(dev_->d_vertices, dev_->h_pinned_vertices, vb, cudaMemcpyHostToDevice, stream));
  MH_TRY(launch_refit(dev_->d_vertices, dev_->d_indices, nt, dev_->d_nodes, dev_->d_tris, dev_->d_tri_box, dev_->d_node_box,
                      bvh.level_start.data(), (uint32_t)bvh.level_start.size(), (uint32_t)bvh.nodes.size(), launches, stream));
  MH_TRY(cudaStreamSynchronize(stream));
  dev_->staging_stale = true;
  refit_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
  build_ms = refit_ms;  // reported as accel_build_ms: the cost of bringing the accel up to date
  return true;
}

// After a device refit the pinned staging copy (and the host Bvh8) lag behind the device: fetch them back.
bool MeshScene::refresh_staging(cudaStream_t stream, std::string* err) {
  Bvh8& bvh = dev_->bvh;
  const size_t nb = bvh.nodes.size() * sizeof(Bvh8Node), tb = bvh.tris.size() * sizeof(PackedTri);
  MH_TRY(cudaMemcpyAsync(dev_->h_pinned, dev_->d_nodes, nb, cudaMemcpyDeviceToHost, stream));
  MH_TRY(cudaMemcpyAsync((char*)dev_->h_pinned + nb, dev_->d_tris, tb, cudaMemcpyDeviceToHost, stream));
  MH_TRY(cudaStreamSynchronize(stream));
  std::memcpy(bvh.nodes.data(), dev_->h_pinned, nb);
  std::memcpy(bvh.tris.data(), (char*)dev_->h_pinned + nb, tb);
  dev_->staging_stale = false;
  return true;
}

// Debug/test hook: FNV-1a over the device copies of the node and triangle arrays.
bool MeshScene::accel_checksum(cudaStream_t stream, uint64_t* nodes_sum, uint64_t* tris_sum, std::string* err) {
  if (!dev_ || !dev_->d_nodes) {
    *err = "no built scene";
    return false;
  }
  const size_t nb = dev_->bvh.nodes.size() * sizeof(Bvh8Node), tb = dev_->bvh.tris.size() * sizeof(PackedTri);
  std::vector<unsigned char> buf(std::max(nb, tb));
  auto fnv = [&](size_t n) {
    uint64_t h = 1469598103934665603ull;
    for (size_t i = 0; i < n; ++i) h = (h ^ buf[i]) * 1099511628211ull;
    return h;
  };
  MH_TRY(cudaStreamSynchronize(stream));
  MH_TRY(cudaMemcpy(buf.data(), dev_->d_nodes, nb, cudaMemcpyDeviceToHost));
  *nodes_sum = fnv(nb);
  MH_TRY(cudaMemcpy(buf.data(), dev_->d_tris, tb, cudaMemcpyDeviceToHost));
  *tris_sum = fnv(tb);
  return true;
}

void MeshScene::next_trace_events(cudaEvent_t* e0, cudaEvent_t* e1) {
  *e0 = *e1 = nullptr;
  if (trace_events_used_ + 2 > 8192) return;  // very long renders: the tail is simply not timed per kernel
  while (trace_events_.size() < trace_events_used_ + 2) {
    cudaEvent_t e;
    if (cudaEventCreate(&e) != cudaSuccess) return;
    trace_events_.push_back(e);
  }
  *e0 = trace_events_[trace_events_used_++];
  *e1 = trace_events_[trace_events_used_++];
}

double MeshScene::collect_packet_ms() {
  double total = 0.0;
  for (size_t i = 0; i + 1 < packet_events_used_; i += 2) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, packet_events_[i], packet_events_[i + 1]) == cudaSuccess) total += ms;
  }
  return total;
}

// Sum of the closest-hit traversal launches of the last render(); call after the stream has been synchronised.
double MeshScene::collect_trace_ms() {
  double total = 0.0;
  for (size_t i = 0; i + 1 < trace_events_used_; i += 2) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, trace_events_[i], trace_events_[i + 1]) == cudaSuccess) total += ms;
  }
  return total;
}

static cudaError_t ensure_state(MeshDevice* d, uint32_t paths, uint32_t bounces) {
  if (d->state_paths >= paths && d->state_bounces >= bounces) return cudaSuccess;
  d->free_state();
  WavefrontState& st = d->st;
  cudaError_t e;
#define ST_ALLOC(ptr, bytes) \
  if ((e = cudaMalloc(&ptr, (bytes))) != cudaSuccess) return e
  ST_ALLOC(st.ray_o, sizeof(float4) * (size_t)paths);
  ST_ALLOC(st.ray_d, sizeof(float4) * (size_t)paths);
  ST_ALLOC(st.thr, sizeof(float4) * (size_t)paths);
  ST_ALLOC(st.L, sizeof(float4) * (size_t)paths);
  ST_ALLOC(st.hit, sizeof(uint2) * (size_t)paths);
  ST_ALLOC(st.sh_o, sizeof(float4) * (size_t)paths);
  ST_ALLOC(st.sh_d, sizeof(float4) * (size_t)paths);
  ST_ALLOC(st.sh_c, sizeof(float4) * (size_t)paths);
  ST_ALLOC(st.queue[0], sizeof(uint32_t) * (size_t)paths);
  ST_ALLOC(st.queue[1], sizeof(uint32_t) * (size_t)paths);
  ST_ALLOC(st.shadow_queue, sizeof(uint32_t) * (size_t)paths);
  ST_ALLOC(st.counters, sizeof(uint32_t) * 4 * (size_t)(bounces + 2));
#undef ST_ALLOC
  d->state_paths = paths;
  d->state_bounces = bounces;
  return cudaSuccess;
}

cudaError_t MeshScene::render(const MeshRenderArgs& a, cudaStream_t stream, uint64_t* launches, std::string* err) {
  *launches = 0;
  trace_events_used_ = 0;
  packet_events_used_ = 0;
  if (!dev_ || !dev_->d_nodes) {
    *err = "accel not built";
    return cudaErrorInvalidValue;
  }
  const uint32_t n_pixels = a.im.width * a.im.height;
  // pixel slots of the chunked path-id layout (kernels.h): the image padded to 8x8 tiles
  const uint32_t slot_tiles_x = (a.im.width + 7u) / 8u, slot_tiles_y = (a.im.height + 7u) / 8u;
  const uint32_t n_slots_full = slot_tiles_x * slot_tiles_y * 64u;
  const uint32_t max_b = (uint32_t)a.ip.max_bounces;
  cudaError_t e;
  // Interleaved-tile partition with tiles that are multiples of 8x8 pixels: the passes whose camera rays are traced as
  // packets use slots COMPACTED to the slot tiles this rank owns (list built here once per resolution / partition).
  uint32_t n_slots_dense = 0;
  if (a.im.part_world > 1u && a.im.tile_w % 8u == 0u && a.im.tile_h % 8u == 0u) {
    const uint32_t key[6] = {a.im.width, a.im.height, a.im.part_rank, a.im.part_world, a.im.tile_w, a.im.tile_h};
    if (std::memcmp(key, dev_->owned_key, sizeof(key)) != 0) {
      std::vector<uint32_t> owned;
      for (uint32_t ty = 0; ty < slot_tiles_y; ++ty)
        for (uint32_t tx = 0; tx < slot_tiles_x; ++tx) {
          const uint32_t px = (tx * 8u) / a.im.tile_w, py = (ty * 8u) / a.im.tile_h;  // cudapt_set_tile_partition's rule
          if ((py * a.im.tiles_x + px + py) % a.im.part_world == a.im.part_rank) owned.push_back(ty * slot_tiles_x + tx);
        }
      cudaFree(dev_->d_owned_tiles);
      dev_->d_owned_tiles = nullptr;
      if ((e = cudaMalloc(&dev_->d_owned_tiles, sizeof(uint32_t) * std::max<size_t>(owned.size(), 1))) != cudaSuccess ||
          (!owned.empty() && (e = cudaMemcpyAsync(dev_->d_owned_tiles, owned.data(), sizeof(uint32_t) * owned.size(), cudaMemcpyHostToDevice,
                                                  stream)) != cudaSuccess) ||
          (e = cudaStreamSynchronize(stream)) != cudaSuccess) {  // `owned` goes out of scope
        *err = "owned-tile list";
        return e;
      }
      dev_->n_owned_tiles = (uint32_t)owned.size();
      std::memcpy(dev_->owned_key, key, sizeof(key));
    }
    n_slots_dense = dev_->n_owned_tiles * 64u;  // 0: this rank owns nothing -> the sparse form handles it
  }
  uint32_t S = std::max(1u, std::min(a.samples_per_pass, a.n_frames));
  PassParams probe;  // does a pass of S samples trace its camera rays as packets (and may therefore be compacted)?
  probe.samples = S;
  probe.cam = a.cam;
  probe.im = a.im;
  probe.dense = n_slots_dense != 0u ? 1u : 0u;  // if the pass can be compacted it will be
  const bool first_dense = n_slots_dense != 0u && primary_packets_enabled(probe);
  const uint32_t n_slots = first_dense ? n_slots_dense : n_slots_full;
  // 148 B of state + queues per path: keep a pass within 3/4 of the memory that is free (or already ours), and its
  // path ids within 32 bits.  The driver is asked only when the state would have to GROW: cudaMemGetInfo can take
  // milliseconds, and it sits between the render's first event and its first kernel.
  if ((uint64_t)n_slots * S > dev_->state_paths) {
    size_t free_b = 0, total_b = 0;
    if (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess) {
      const size_t ours = (size_t)dev_->state_paths * 148u;
      const size_t budget = (free_b + ours) / 4u * 3u;
      const size_t fit = budget / ((size_t)n_slots * 148u);
      S = (uint32_t)std::max<size_t>(1, std::min<size_t>(S, fit));
    }
    S = (uint32_t)std::min<uint64_t>(S, 0xFFFFFFFFull / n_slots);
  }
  PassParams p;
  p.st = dev_->st;
  p.scene.nodes = dev_->d_nodes;
  p.scene.tris = dev_->d_tris;
  p.scene.materials = dev_->d_materials;
  p.scene.lights = dev_->d_lights;
  p.scene.n_lights = (uint32_t)dev_->lights.size();
  p.scene.n_materials = (uint32_t)dev_->materials.size();
  p.cam = a.cam;
  p.im = a.im;
  p.ip = a.ip;
  p.n_pixels = n_pixels;
  p.slot_tiles_x = slot_tiles_x;
  p.owned_tiles = dev_->d_owned_tiles;
  p.fd_tiles_x = make_fastdiv(slot_tiles_x);
  p.fd_width = make_fastdiv(a.im.width);
  p.fd_pixels = make_fastdiv(n_pixels);
  p.frame_stride = a.frame_stride;
  p.accum = a.accum;
  p.counters = a.counters;
  const bool nee = a.ip.nee && p.scene.n_lights > 0;
#define R_TRY(expr, what)     \
  if ((e = (expr)) != cudaSuccess) { \
    *err = what;              \
    return e;                 \
  }
  uint32_t done = 0;
  while (done < a.n_frames) {
    const uint32_t s = std::min(S, a.n_frames - done);
    p.samples = s;
    // slots of this pass: compacted to the owned tiles when its camera rays go as packets under a tile partition
    p.dense = n_slots_dense != 0u ? 1u : 0u;  // tentatively: packets of 1..16 samples need every slot owned
    p.dense = primary_packets_enabled(p) ? p.dense : 0u;
    p.n_slots = p.dense ? n_slots_dense : n_slots_full;
    p.fd_slots = make_fastdiv(p.n_slots);
    if ((uint64_t)p.n_slots * s > 0xFFFFFFFFull) {
      *err = "pass too large for 32-bit path ids";
      return cudaErrorInvalidValue;
    }
    if ((e = ensure_state(dev_, p.n_slots * s, max_b)) != cudaSuccess) {  // grows only
      *err = "wavefront state allocation";
      return e;
    }
    p.st = dev_->st;
    p.first_frame = a.first_frame + done * a.frame_stride;
    p.aov_prim = done == 0 ? a.aov_prim : nullptr;
    p.aov_t = done == 0 ? a.aov_t : nullptr;
    R_TRY(cudaMemsetAsync(p.st.counters, 0, sizeof(uint32_t) * 4 * (size_t)(max_b + 2), stream), "counter memset");
    // camera rays: packets (one warp = 32 samples of a pixel; fused raygen) + the per-ray kernel for the rays of
    // packets that gave up, or -- pass size not a multiple of 32 -- raygen + the per-ray kernel
    p.chunked = 0;
    p.explicit_q0 = 0;
    p.s_shift = 0;
    const bool packets = primary_packets_enabled(p);
    // path-id layout: chunked for passes of 32 k samples (one pixel per packet), slot-major for 1..16 (several pixels)
    p.chunked = packets ? (s % 32u == 0u ? 1u : 2u) : 0u;
    if (p.chunked == 1u) p.s_shift = 5;
    while (p.chunked == 2u && (1u << p.s_shift) < s) ++p.s_shift;
    p.explicit_q0 = packets && primary_packet_hitlist(p) ? 1u : 0u;
    uint32_t* fb = p.st.counters + 4 * (size_t)(max_b + 1);  // {fall-back count, fall-back work counter}
    if (packets) {
      cudaEvent_t e0 = nullptr, e1 = nullptr;
      if (packet_events_used_ + 2 <= 1024) {
        while (packet_events_.size() < packet_events_used_ + 2) {
          cudaEvent_t e;
          if (cudaEventCreate(&e) != cudaSuccess) break;
          packet_events_.push_back(e);
        }
        if (packet_events_.size() >= packet_events_used_ + 2) {
          e0 = packet_events_[packet_events_used_++];
          e1 = packet_events_[packet_events_used_++];
        }
      }
      if (e0) cudaEventRecord(e0, stream);
      R_TRY(launch_primary_packets(p, p.st.queue[1], fb, a.counting, sm_count, stream), "primary_packets");
      if (e1) cudaEventRecord(e1, stream);
      if (p.explicit_q0) *launches += 2;  // hit-list scan + expand
    } else {
      R_TRY(launch_raygen(p, sm_count, stream), "raygen");
    }
    ++*launches;
    for (uint32_t b = 0; b < max_b; ++b) {
      const int qin = (int)(b & 1u);
      cudaEvent_t e0 = nullptr, e1 = nullptr;
      next_trace_events(&e0, &e1);  // per-launch timing of the dominant kernel (bench.py roofline)
      if (e0) cudaEventRecord(e0, stream);
      if (b == 0 && packets) {
        R_TRY(launch_trace_queue(p, p.st.queue[1], fb, fb + 1, a.counting, sm_count, stream), "trace_fallback");
      } else {
        R_TRY(launch_trace_closest(p, b, qin, a.counting, sm_count, stream), "trace_closest");
      }
      if (e1) cudaEventRecord(e1, stream);
      R_TRY(launch_shade(p, b, qin, sm_count, stream), "shade");
      *launches += 2;
      if (nee) {
        R_TRY(launch_trace_shadow(p, b, a.counting, sm_count, stream), "trace_shadow");
        ++*launches;
      }
    }
    R_TRY(launch_accumulate(p, sm_count, stream), "accumulate");
    ++*launches;
    done += s;
  }
  R_TRY(dev_->ensure_copy_objects(), "copy stream / events");
  R_TRY(cudaEventRecord(dev_->ev_last_use[dev_->current], stream), "event record");
#undef R_TRY
  return cudaSuccess;
}

}  // namespace cudapt
