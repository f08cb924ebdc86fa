/* cudapt.h -- C ABI of the B200-native path-tracing backend (libcudapt.so).
 *
 * This is the drop-in boundary for the ONE hot path of imisumi/software-path-tracer: the progressive
 * per-pixel path-tracing loop behind render::PathTracer.  Every entry point cites the reference interface
 * it replaces; paths are relative to the reference tree and "CPT.cpp" abbreviates
 * libs/render/src/engines/pathtracer/backends/cpu/CPUPathTracer.cpp.
 *
 * Plain pointers and sizes only.  All functions return 0 on success and a non-zero cudapt_status
 * otherwise; cudapt_last_error() then describes the failure.  There is no CPU fallback: every entry point
 * that computes fails with CUDAPT_ERR_CUDA when no sm_100 device is usable.
 *
 * Threading: one host thread per context (the reference is synchronous and non-reentrant, App.cpp:231).
 * cudapt_render() only enqueues work on the context's stream; every read-back call synchronises.
 */
#ifndef CUDAPT_H_
#define CUDAPT_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CUDAPT_ABI_VERSION 1

typedef struct cudapt_ctx cudapt_ctx;

typedef enum cudapt_status {
  CUDAPT_OK = 0,
  CUDAPT_ERR_INVALID = 1,   /* bad argument / call order (reference: verify() abort, render_assert.h:15-25) */
  CUDAPT_ERR_CUDA = 2,      /* CUDA runtime error; never swallowed */
  CUDAPT_ERR_NO_SCENE = 3,  /* render before any scene upload (reference: verify(m_scene) CPT.cpp:46) */
  CUDAPT_ERR_NO_FRAMES = 4  /* resolve before any frame (reference: assert(m_frameCount>0) CPT.cpp:89) */
} cudapt_status;

/* RNG streams.  REFERENCE_PCG reproduces CPT.cpp:192-195 + :294-301 bit-for-bit (seed x + y*W +
 * frame*982451653u, PCG hash step, float(w)/4294967295.0f).  PHILOX is Philox4x32-10 keyed by
 * (0xB200CAFE, seed) with counter (pixel, sample, bounce, block): any GPU and the CPU oracle replay the
 * same stream for a given (pixel, sample, bounce) irrespective of how work is partitioned. */
typedef enum cudapt_rng_mode { CUDAPT_RNG_REFERENCE_PCG = 0, CUDAPT_RNG_PHILOX = 1 } cudapt_rng_mode;

/* Material models of the mesh integrator (the reference has none: constant 0.7 albedo, CPT.cpp:260). */
typedef enum cudapt_material_type {
  CUDAPT_MAT_LAMBERT = 0,
  CUDAPT_MAT_MIRROR = 1,
  CUDAPT_MAT_DIELECTRIC = 2,
  CUDAPT_MAT_GGX_COATED = 3
} cudapt_material_type;

/* 32-byte material record (two 128-bit loads on the device). */
typedef struct cudapt_material {
  float albedo[3];
  uint32_t type;      /* cudapt_material_type */
  float emission[3];  /* radiance emitted by the front face (geometric-normal side) */
  float param;        /* GGX alpha for GGX_COATED, index of refraction for DIELECTRIC */
} cudapt_material;

/* Pinhole camera.  Defaults reproduce the camera hard-coded in CPT.cpp:62-73 bit-for-bit: origin 0,
 * right +X, up +Y, forward +Z, tan(vfov/2) = 1, pixel-corner sampling, no jitter. */
typedef struct cudapt_camera {
  float origin[3];
  float right[3];
  float up[3];
  float forward[3];
  float tan_half_fov;
  float lens_radius;     /* thin lens (Philox mode only): radius of the aperture disk in the right/up plane; 0 = pinhole */
  float focus_distance;  /* distance along `forward` of the plane that is in focus (default 1; ignored for a pinhole)  */
} cudapt_camera;

typedef struct cudapt_config {
  int32_t device;            /* CUDA device ordinal */
  int32_t rng_mode;          /* cudapt_rng_mode; default REFERENCE_PCG */
  int32_t max_bounces;       /* default 4  (literal at CPT.cpp:199) */
  int32_t rr_start;          /* default 2: Russian roulette once bounce_count > rr_start (CPT.cpp:264) */
  float tnear;               /* default 1e-3f (CPT.cpp:221) */
  float eps;                 /* default 1e-4f normal offset (CPT.cpp:277) */
  float albedo;              /* default 0.7f, sphere integrator only (CPT.cpp:260) */
  float sky_scale;           /* default 1: multiplies the procedural sky (CPT.cpp:286-292) */
  uint32_t seed;             /* Philox key word 1 */
  int32_t jitter;            /* default 0: sub-pixel jitter (Philox mode only); 0 = reference corner sampling */
  int32_t nee;               /* default 0: next-event estimation on emissive triangles (mesh integrator) */
  int32_t samples_per_pass;  /* default 1, <= 1024: samples of every pixel traced together by the wavefront (136 B of path
                                state per sample and pixel: large passes amortise the near-empty tail bounces; a pass
                                that does not fit 3/4 of the free device memory is shrunk; under a tile partition a pass
                                holds only the rank's own pixels, so `world` x more samples fit the same memory) */
} cudapt_config;

typedef struct cudapt_stats {
  uint64_t paths;            /* camera samples traced to termination since the last reset */
  uint64_t rays;             /* closest-hit + shadow queries since the last reset (counted on device) */
  uint64_t nodes_visited;    /* BVH nodes fetched   (only while cudapt_set_counting(ctx,1)) */
  uint64_t tris_tested;      /* triangles tested    (only while cudapt_set_counting(ctx,1)) */
  uint64_t shade_invocations;/* surface interactions shaded */
  uint64_t kernel_launches;  /* kernels launched by this library since cudapt_create (monotonic) */
  double render_ms;          /* CUDA-event time of all cudapt_render() calls since the last reset */
  double last_render_ms;     /* CUDA-event time of the most recent cudapt_render() call */
  double accel_build_ms;     /* host wall time of the last cudapt_build_accel() */
  uint64_t accel_bytes;      /* device bytes of BVH nodes + packed triangles */
  uint32_t frame_count;      /* samples accumulated per pixel (reference m_frameCount, CPT.h:68) */
  uint32_t n_triangles;
  uint32_t n_nodes;
  uint32_t n_spheres;
  double last_trace_ms;      /* CUDA-event time of the closest-hit BVH traversal launches inside the most
                                recent cudapt_render() (mesh integrator; 0 for the sphere megakernel) */
  uint64_t packet_nodes;     /* camera-ray packets (32 samples of a pixel, csrc/primary_packet.cu): node steps,   */
  uint64_t packet_tris;      /* triangle steps (each tests 32 rays) and packets handed to the per-ray kernel --   */
  uint64_t packet_bails;     /* only while counting, like nodes_visited                                            */
  uint64_t packet_rays;      /* camera rays finished by packets (part of .rays; always counted)                    */
  uint64_t shadow_rays;      /* any-hit queries (part of .rays; always counted)                                    */
  double last_packet_ms;     /* CUDA-event time of the packet launches inside the most recent cudapt_render()      */
  uint64_t packets;          /* packets finished (32, 64 or 128 camera rays each; always counted)                  */
} cudapt_stats;

/* ---- lifetime: CPUPathTracer ctor/dtor, CPT.cpp:25-41 ---------------------------------------------- */
void cudapt_default_config(cudapt_config* cfg);
void cudapt_default_camera(cudapt_camera* cam);
/* Fills *cam from the parameters of the reference's placeholder render::Camera (Camera.h:14-20,33-36: setPosition,
 * setTarget, setUp, setPerspective's vertical field of view in degrees).  position 0, target +Z, up +Y, 90 degrees
 * gives cudapt_default_camera() exactly.  CUDAPT_ERR_INVALID for a zero view vector or an up vector parallel to it. */
int cudapt_camera_look_at(cudapt_camera* cam, const float position[3], const float target[3], const float up[3],
                          float vfov_degrees);
int cudapt_create(const cudapt_config* cfg, cudapt_ctx** out);
void cudapt_destroy(cudapt_ctx* ctx);
/* Error text of the last failing call on ctx (ctx may be NULL for a failed cudapt_create). */
const char* cudapt_last_error(const cudapt_ctx* ctx);
int cudapt_abi_version(void);

/* ---- settings: invalidate() size rule, CPT.cpp:140-149 --------------------------------------------- */
/* (Re)allocates accumulation float4[W*H] and result u32[W*H], zeroes them, frame count = 0. */
int cudapt_set_resolution(cudapt_ctx* ctx, uint32_t width, uint32_t height);
/* Replaces the integrator parameters (everything in cfg except .device) and restarts accumulation,
 * like a dirty RenderSettings does (CPT.cpp:132-138). */
int cudapt_set_config(cudapt_ctx* ctx, const cudapt_config* cfg);
int cudapt_set_camera(cudapt_ctx* ctx, const cudapt_camera* cam);

/* ---- scene upload: rebuild_scene(), CPT.cpp:362-403 ------------------------------------------------ */
/* n spheres as {cx,cy,cz,r} (the float4 the reference hands to Embree, CPT.cpp:377-385) plus the
 * caller's stable id per sphere (render::NodeID) reported by cudapt_read_primary_hits.  Replaces any
 * previous scene ("rebuild = replace") and restarts accumulation.  Selects the sphere integrator. */
int cudapt_upload_spheres(cudapt_ctx* ctx, const float* center_radius, const uint32_t* node_id, uint32_t n);
/* Triangle scene (no reference analogue; configs 2-5).  vertices: nv x float3; indices: nt x uint3;
 * material_id: nt (may be NULL = all 0).  Selects the mesh (wavefront) integrator once built. */
int cudapt_upload_mesh(cudapt_ctx* ctx, const float* vertices, uint32_t nv, const uint32_t* indices, uint32_t nt,
                       const uint32_t* material_id);
int cudapt_upload_materials(cudapt_ctx* ctx, const cudapt_material* materials, uint32_t n);
/* Builds the 8-wide quantised BVH on the host (SAH) and uploads it; replaces rtcCommitScene (CPT.cpp:403). */
int cudapt_build_accel(cudapt_ctx* ctx);
/* The same, built ON THE DEVICE (csrc/lbvh_build.cu: Morton codes, radix sort, binary radix tree, collapse to 8-wide
 * nodes, quantisation by the refit kernels): milliseconds instead of the host builder's fraction of a second per
 * million triangles, for scenes whose topology changes (SURVEY.md 8f-3 "refit / rebuild"); the tree is of lower
 * quality (no SAH), images are identical.  Needs at least two triangles. */
int cudapt_build_accel_device(cudapt_ctx* ctx);
/* Copies the already built scene (BVH nodes, packed triangles, materials, light list) from the library's
 * pinned host staging to the device again and restarts accumulation -- the per-frame cost of a scene that
 * changed on the host without changing topology (the reference re-commits its Embree scene whenever
 * Scene::hasChanges(), CPT.cpp:122-131,156-160). */
int cudapt_reupload_scene(cudapt_ctx* ctx);
/* The same copy, DOUBLE-BUFFERED: starts copying the scene into a second device copy on the library's copy stream and
 * returns at once; rendering that is in flight keeps reading the current copy.  The new copy becomes the scene at the
 * next cudapt_render() (or scene call), which waits for the transfer and restarts accumulation.  This is how a host
 * that rewrites the scene every frame overlaps the upload of frame k+1 with the rendering of frame k. */
int cudapt_stage_scene(cudapt_ctx* ctx);
/* Incremental scene update: the triangle mesh keeps its indices and materials, the nv vertices
 * (same count as uploaded) move.  The BVH is REFIT on the host -- tree, child slots and triangle order are kept,
 * every box and quantisation grid is recomputed bottom-up -- and uploaded again; accumulation restarts.  This is
 * what the reference's dirty-bit handshake (Scene::hasChanges()/markChangesProcessed(), Scene.h:207-215,
 * CPT.cpp:122-131,156-160) would trigger if anything ever set the bit again (nothing does), and
 * it fixes the reference's rebuild bug (geometries are never detached, CPT.cpp:362-403): update = replace.
 * Because the closest-hit rule does not depend on the acceleration structure, the image after an update is
 * bit-identical to a fresh build of the moved mesh. */
int cudapt_update_vertices(cudapt_ctx* ctx, const float* vertices, uint32_t nv);
/* Test hook: FNV-1a checksums of the device-resident node and packed-triangle arrays (the refit runs on the device,
 * csrc/refit.cu; with CUDAPT_REFIT_HOST=1 in the environment the host implementation of the same code runs instead
 * and must leave the same bytes). */
int cudapt_debug_accel_checksum(cudapt_ctx* ctx, uint64_t* nodes_sum, uint64_t* tris_sum);

/* Environment map replacing the procedural sky on a miss (reference anchor: the commented-out
 * Scene::setEnvironmentMap, libs/render/src/Scene.cpp:124-169, and the SkyBox of Scene.h:268-278).  rgb is a
 * size x size OCTAHEDRAL map, row-major float3 texels, pole axis +Y: direction d maps to p = (d.x, d.z) /
 * (|d.x|+|d.y|+|d.z|), lower hemisphere folded outwards, texel coordinates (p*0.5+0.5)*size - 0.5, bilinear,
 * clamped.  (Octahedral rather than lat-long so that the lookup needs no atan2/acos and the CPU oracle
 * matches bit for bit.)  Both integrators use it; sky_scale still multiplies it.  rgb = NULL or size = 0 restores
 * the reference's procedural sky (CPT.cpp:286-292).  Restarts accumulation. */
int cudapt_upload_environment(cudapt_ctx* ctx, const float* rgb, uint32_t size);
/* The file form of the above -- what the commented-out Scene::setEnvironmentMap(filepath) did through OpenImageIO
 * (Scene.cpp:124-169).  path is a LATITUDE-LONGITUDE panorama, ".pfm" (portable float map) or ".hdr" / ".pic"
 * (Radiance RGBE, flat or run-length encoded), read without an image library; it is resampled once on the host
 * (double precision, bilinear) into a size x size octahedral map and uploaded.  Panorama convention: column
 * u = 1/2 + atan2(d.x, d.z) / 2pi (the image centre is +Z, the default camera's view direction), row v = acos(d.y) / pi
 * (top row = +Y).  octahedral_out (size*size*3 floats) receives the texels that were uploaded, or NULL.
 * CUDAPT_ERR_INVALID for an unreadable / unsupported file.  Restarts accumulation. */
int cudapt_load_environment(cudapt_ctx* ctx, const char* path, uint32_t size, float* octahedral_out);
/* Only the host half of the above (file -> octahedral texels); needs no device and no context, error text through
 * cudapt_last_error(NULL). */
int cudapt_convert_environment(const char* path, uint32_t size, float* octahedral_out);

/* ---- progressive rendering: render(), CPT.cpp:43-85 ------------------------------------------------ */
/* Zero the accumulation buffer, frame count = 0 (CPT.cpp:151-154). */
int cudapt_reset(cudapt_ctx* ctx);
/* Enqueue n_frames frames of 1 sample per pixel each.  Frame f (1-based, = frame_count+1 as at
 * CPT.cpp:61) seeds the RNG; frame_count advances by n_frames. */
int cudapt_render(cudapt_ctx* ctx, uint32_t n_frames);
int cudapt_synchronize(cudapt_ctx* ctx);

/* ---- read-back: get_render_result(), CPT.cpp:87-117 + Color.h:7-10 --------------------------------- */
/* accum/frame_count -> clamp[0,1] -> (uint8)(c*255) -> R<<24|G<<16|B<<8|A, W*H words, row 0 on top. */
int cudapt_resolve_rgba8(cudapt_ctx* ctx, uint32_t* host_out);
/* Page-lock a caller-owned host buffer (RenderResult::image_buffer, PathTracer.h:23-29) so that
 * cudapt_resolve_rgba8 / cudapt_download_rgba8 / cudapt_read_accum copy straight into it; an ordinary buffer goes
 * through the library's pinned staging + one memcpy (8 MB at 1080p: ~1 ms per read-back).  Thin wrappers over
 * cudaHostRegister / cudaHostUnregister, here so that the C++ host does not have to link the CUDA runtime.
 * Unregister before the buffer is freed or resized. */
int cudapt_host_register(cudapt_ctx* ctx, void* host_ptr, size_t bytes);
int cudapt_host_unregister(cudapt_ctx* ctx, void* host_ptr);
/* Double-buffered read-back for a viewer that shows frame k while frame k+1 renders (App.cpp:231-238 calls render()
 * and get_render_result() back to back every UI frame).  cudapt_resolve_rgba8_async enqueues the resolve on the
 * context's stream and the device-to-host copy on a second stream and returns at once; host_out must be page-locked
 * (CUDAPT_ERR_INVALID otherwise) and holds the image once cudapt_wait_download returns.  Rendering calls made in
 * between overlap the copy; the next call that rewrites the 8-bit image waits for it on the device.  One read-back
 * can be in flight: wait before starting the next one into the same buffer.  Not with auto-exposure (its result is
 * read back synchronously).  cudapt_download_rgba8_async is the same for an image already in the context's 8-bit
 * buffer (cudapt_reduce_resolve_peers on the root); there the ROOT calls cudapt_stream_wait_download before the
 * rendezvous that lets the peers write the next image, which makes its stream wait for the pending copy. */
int cudapt_resolve_rgba8_async(cudapt_ctx* ctx, uint32_t* host_out);
int cudapt_download_rgba8_async(cudapt_ctx* ctx, uint32_t* host_out);
int cudapt_wait_download(cudapt_ctx* ctx);
int cudapt_stream_wait_download(cudapt_ctx* ctx);
/* Post-process stage applied by cudapt_resolve_rgba8 between the division by frame_count and the 8-bit pack
 * (SURVEY.md 8f-2).  Reference anchors: the exposure multiply commented out at CPT.cpp:101-104 with
 * RenderSettings::getExposure()/getAutoExposure()/getTargetLuminance() (Types.h:57-58,67-69), and the
 * acesTonemap / reinhardTonemap / linearToSRGB declarations of libs/render/src/utils/Math.h:25-34 (never
 * defined there).  The defaults (exposure 1, no tone map, linear, no auto-exposure) reproduce CPT.cpp:87-117
 * bit for bit.  Order: c = accum/frames; c *= exposure (auto: target_luminance / exp(mean log luminance));
 * tone map; clamp[0,1]; optional sRGB OETF; (uint8)(c*255).  Alpha is never touched. */
typedef enum cudapt_tonemap { CUDAPT_TONEMAP_NONE = 0, CUDAPT_TONEMAP_REINHARD = 1, CUDAPT_TONEMAP_ACES = 2 } cudapt_tonemap;
typedef struct cudapt_postprocess {
  float exposure;          /* default 1.0f (RenderSettings::m_exposure, Types.h:87) */
  int32_t tonemap;         /* cudapt_tonemap, default NONE */
  int32_t srgb;            /* default 0: linear output as the reference */
  int32_t auto_exposure;   /* default 0 (Types.h:88); 1 = exposure from the log-average luminance, on the device */
  float target_luminance;  /* default 0.18f (Types.h:89) */
} cudapt_postprocess;
void cudapt_default_postprocess(cudapt_postprocess* pp);
int cudapt_set_postprocess(cudapt_ctx* ctx, const cudapt_postprocess* pp);
/* The exposure multiplier the last cudapt_resolve_rgba8 used (the auto-exposure result when enabled). */
int cudapt_get_exposure(cudapt_ctx* ctx, float* exposure);
/* Image writer for headless use (the reference only displays through SDL, App.cpp:236-238):
 * ".png" = 8-bit RGBA and ".ppm" = binary P6 of the post-processed image, ".pfm" = float RGB accum/frame_count
 * (bottom row first, little endian), ".exr" = the same linear floats as an OpenEXR 2 scan-line file (32-bit FLOAT
 * channels B G R, uncompressed), chosen by the file extension.  PNG and EXR are written without a codec library
 * (stored deflate blocks / uncompressed scan lines). */
int cudapt_write_image(cudapt_ctx* ctx, const char* path);

/* Float accumulation (private m_accumulation_buffer, CPT.h:72): 4*W*H floats RGBA, and the frame count. */
int cudapt_read_accum(cudapt_ctx* ctx, float* host_out, uint32_t* frame_count);
/* Restore a previously read state (checkpoint/resume, merging partial renders). */
int cudapt_write_accum(cudapt_ctx* ctx, const float* host_in, uint32_t frame_count);
/* Primary-hit AOV of the first sample: caller id of the primitive hit (sphere node_id, or triangle
 * index; 0xFFFFFFFF = miss) and ray parameter t (+inf on miss).  Either pointer may be NULL. */
int cudapt_read_primary_hits(cudapt_ctx* ctx, uint32_t* prim_id, float* t);
int cudapt_get_stats(cudapt_ctx* ctx, cudapt_stats* out);
/* Enable device-side node/triangle counters (slower instrumented traversal; off by default). */
int cudapt_set_counting(cudapt_ctx* ctx, int enabled);

/* ---- multi-GPU partitioning (no reference analogue) ----------------------------------- */
/* Interleaved image tiles: this context renders only tiles t with (t + tile_row) % world == rank, where
 * tiles are tile_w x tile_h pixels in raster order.  Pixels it does not own stay exactly 0 in the
 * accumulation buffer, so a sum over ranks is bit-identical to a single-GPU render.  world = 1 disables. */
int cudapt_set_tile_partition(cudapt_ctx* ctx, uint32_t rank, uint32_t world, uint32_t tile_w, uint32_t tile_h);
/* Sample ranges: frame k of this context (k = 0,1,..) uses the global 1-based frame index
 * first + k*stride for RNG seeding.  Default first = 1, stride = 1. */
int cudapt_set_frame_range(cudapt_ctx* ctx, uint32_t first, uint32_t stride);
/* Device pointers for the caller's collective (NCCL reduce at read-back boundaries) and timing. */
void* cudapt_accum_device_ptr(cudapt_ctx* ctx);  /* float4[W*H] */
void* cudapt_stream(cudapt_ctx* ctx);            /* cudaStream_t all work is enqueued on */
/* Declare that the accumulation buffer now holds frame_count samples per pixel (after an external reduce). */
int cudapt_set_frame_count(cudapt_ctx* ctx, uint32_t frame_count);

/* ---- fused multi-GPU read-back tail over NVLink peer memory (SURVEY.md 8e "fusion opportunity") ---------- */
/* One kernel per rank does reduce-scatter + resolve + gather: rank r owns pixels [n*r/G, n*(r+1)/G); it loads
 * the float4 accumulation of EVERY rank for those pixels (accum_by_rank[g], g = 0..G-1 in rank order, each a
 * device pointer valid on this context's device: its own buffer, another context's, or a peer mapping obtained
 * with cudapt_ipc_open), adds them in rank order, applies the post-process stage with total_frames and stores
 * the RGBA8 words into root_rgba and -- unless NULL -- the float4 sums into root_accum (both may be peer
 * mappings of the root's buffers; root_accum may alias accum_by_rank[root]).  Auto-exposure is not available
 * here (it needs the whole image).  The caller orders the ranks: every rank's render must have completed before
 * any rank launches this (e.g. a barrier on the streams), and the root reads its buffers after all ranks did. */
int cudapt_reduce_resolve_peers(cudapt_ctx* ctx, const void* const* accum_by_rank, uint32_t n_ranks, uint32_t rank,
                                uint32_t total_frames, void* root_rgba, void* root_accum);
void* cudapt_rgba_device_ptr(cudapt_ctx* ctx);  /* uint32[W*H], the buffer cudapt_resolve_rgba8 fills */
/* Copies the context's RGBA8 device buffer to the host as it is (after peers wrote their slices). */
int cudapt_download_rgba8(cudapt_ctx* ctx, uint32_t* host_out);
/* CUDA IPC plumbing for one-process-per-GPU launches: 64-byte handles of the accumulation and RGBA8 buffers
 * (valid until the next cudapt_set_resolution), and mapping/unmapping of a peer's handle on this device. */
#define CUDAPT_IPC_HANDLE_BYTES 64
int cudapt_ipc_export(cudapt_ctx* ctx, void* accum_handle, void* rgba_handle);
int cudapt_ipc_open(cudapt_ctx* ctx, const void* handle, void** device_ptr);
int cudapt_ipc_close(cudapt_ctx* ctx, void* device_ptr);

/* ---- allocation-free entry point (no counterpart in the reference) ---- */
/* The auto-exposure reduction of cudapt_postprocess.auto_exposure (log-average luminance of accum / frame_count ->
 * exposure = target_luminance / exp(mean log luminance)) for callers that cannot call cudaMalloc: code under CUDA graph
 * capture, tight serving loops, harnesses that forbid it.  It allocates NOTHING (no cudaMalloc / cudaFree, no allocating
 * library routine): its scratch -- one partial sum per 1024 pixels -- lives in a file-scope `static __device__` array of
 * 64 KB that every GPU loading the library holds until exit (csrc/static_scratch.cu), which bounds the image size:
 * n_pixels <= CUDAPT_STATIC_MAX_PIXELS (16 777 216 = two 4K frames); larger sizes are rejected with CUDAPT_ERR_INVALID.
 * The caller passes the input and the output: accum_device = float4[n_pixels] in the memory of `device` (e.g.
 * cudapt_accum_device_ptr), exposure_device_out = one float in device memory; the two kernels are enqueued on `stream`
 * (a cudaStream_t; NULL = the default stream) and the call returns without synchronising.  The result equals, bit for
 * bit, what cudapt_get_exposure reports after an auto-exposure cudapt_resolve_rgba8 of the same buffer.
 * Every call on a device shares the array, which keeps the last call's contents (nothing in it needs clearing: each
 * call overwrites and reads exactly its own [0, n) partials); calls that OVERLAP on the GPU -- different streams without
 * an ordering between them -- race on it.  The first call on a device resolves the array's device address
 * (cudaGetSymbolAddress): make that call outside a graph capture.  Error text through cudapt_last_error(NULL). */
#define CUDAPT_STATIC_MAX_PIXELS 16777216u
int cudapt_auto_exposure_static(int device, const void* accum_device, uint32_t n_pixels, uint32_t frame_count,
                                float target_luminance, float* exposure_device_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* CUDAPT_H_ */
