#include "mesh_host.h"

namespace cudapt {
struct MeshDevice {};
void MeshScene::set_mesh(const float* v, uint32_t nv, const uint32_t* idx, uint32_t nt, const uint32_t* mid) {
  vertices_.assign(v, v + 3 * (size_t)nv);
  indices_.assign(idx, idx + 3 * (size_t)nt);
  if (mid) material_id_.assign(mid, mid + nt); else material_id_.assign(nt, 0u);
}
void MeshScene::set_materials(const cudapt_material* m, uint32_t n) { materials_.assign(m, m + n); }
bool MeshScene::build_and_upload(std::string* err) { *err = "mesh path not built yet"; return false; }
cudaError_t MeshScene::render(const MeshRenderArgs&, cudaStream_t, uint64_t*, std::string* err) { *err = "mesh path not built yet"; return cudaErrorNotSupported; }
void MeshScene::invalidate_image_state() {}
void MeshScene::release() { delete dev_; dev_ = nullptr; }
}  // namespace cudapt
