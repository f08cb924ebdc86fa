/* TEST INFRASTRUCTURE (oracle, Tier 2: "CPU twin") -- never linked into or called by the product.
 *
 * Plain C interface (for ctypes and the C++ tests) of a CPU restatement of the reference hot path,
 * libs/render/src/engines/pathtracer/backends/cpu/CPUPathTracer.cpp ("CPT.cpp") of
 * imisumi/software-path-tracer, following SURVEY.md Appendix A.
 *
 * Pinning:
 *  - sphere integrator (twin_render_spheres with the default parameters) is pinned against the reference's
 *    own translation units compiled here (oracle/_ref, Tier 1): tests/test_oracle.py requires the float
 *    accumulation buffer to match BIT FOR BIT on the App.cpp:101-122 scene, and against the fixtures in
 *    tests/golden/ generated from Tier 1 by tests/golden/make_golden.py.
 *  - mesh integrator (twin_render_mesh: triangles, materials, NEE, Philox) has NO reference counterpart
 *    (the reference renders spheres only, CPT.cpp:367-398): PARITY UNPINNED by the reference.  It keeps
 *    every reference convention it can (camera mapping, tnear/eps, RR rule, sky, accumulation, 8-bit pack)
 *    and the written spec in DESIGN.md "Mesh integrator numerics" is what the CUDA path is checked against.
 */
#ifndef ORACLE_TWIN_H_
#define ORACLE_TWIN_H_
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct twin_params {
  int32_t width, height;
  int32_t rng_mode;     /* 0 = reference PCG (CPT.cpp:192-195,294-301), 1 = Philox4x32-10 */
  int32_t max_bounces;  /* 4  (CPT.cpp:199) */
  int32_t rr_start;     /* 2  (CPT.cpp:264) */
  float tnear;          /* 1e-3f (CPT.cpp:221) */
  float eps;            /* 1e-4f (CPT.cpp:277) */
  float albedo;         /* 0.7f (CPT.cpp:260) */
  float sky_scale;      /* 1 */
  uint32_t seed;
  int32_t jitter;
  int32_t nee;
  float cam_origin[3], cam_right[3], cam_up[3], cam_forward[3], cam_tan_half_fov;
  int32_t threads;      /* host threads; <= 0 = all */
  /* interleaved-tile partition (mirrors cudapt_set_tile_partition); world <= 1 disables */
  int32_t part_rank, part_world, tile_w, tile_h;
  /* octahedral environment map (mirrors cudapt_upload_environment): env_size x env_size float3 texels, or NULL =
   * the reference's procedural sky (CPT.cpp:286-292) */
  const float* env_rgb;
  int32_t env_size;
  /* thin lens (mirrors cudapt_camera.lens_radius / focus_distance; Philox mode only): 0 = pinhole */
  float cam_lens_radius, cam_focus_distance;
} twin_params;

typedef struct twin_material {
  float albedo[3];
  uint32_t type; /* 0 lambert, 1 mirror, 2 dielectric, 3 ggx-coated */
  float emission[3];
  float param;
} twin_material;

typedef struct twin_counters {
  uint64_t paths, rays, shade_invocations, nodes_visited, tris_tested;
} twin_counters;

void twin_default_params(twin_params* p);

/* Adds frames first_frame .. first_frame + (n_frames-1)*frame_stride (1-based global frame indices) to
 * accum (4*W*H floats, RGBA, caller-initialised).  primary_idx / primary_t (may be NULL) receive the
 * primary hit of the first frame rendered: sphere index (-1 = miss) and t. */
int twin_render_spheres(const twin_params* p, const float* center_radius, int32_t n_spheres, uint32_t first_frame,
                        uint32_t n_frames, uint32_t frame_stride, float* accum, int32_t* primary_idx,
                        float* primary_t, twin_counters* counters);

/* Mesh scene handle: builds the twin's own BVH2 (median split) once. */
void* twin_mesh_create(const float* vertices, uint32_t nv, const uint32_t* indices, uint32_t nt,
                       const uint32_t* material_id, const twin_material* materials, uint32_t n_materials);
void twin_mesh_destroy(void* mesh);
int twin_render_mesh(const twin_params* p, void* mesh, uint32_t first_frame, uint32_t n_frames,
                     uint32_t frame_stride, float* accum, int32_t* primary_idx, float* primary_t,
                     twin_counters* counters);
/* Brute-force closest hit over all triangles (validates the twin's BVH on small scenes). */
int twin_mesh_closest_bruteforce(void* mesh, const float* org, const float* dir, float tnear, int32_t* prim,
                                 float* t);
int twin_mesh_closest(void* mesh, const float* org, const float* dir, float tnear, int32_t* prim, float* t);

/* get_render_result(), CPT.cpp:87-117 + Color.h:7-10. */
void twin_resolve_rgba8(const float* accum, uint32_t n_pixels, uint32_t frame_count, uint32_t* rgba);

/* Raw generators, for known-answer tests. */
uint32_t twin_pcg_seed(uint32_t x, uint32_t y, uint32_t width, uint32_t frame);
float twin_pcg_next(uint32_t* state, uint32_t* raw_word);
void twin_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]);
void twin_sincos2pi(float u, float* c, float* s);

#ifdef __cplusplus
}
#endif
#endif
