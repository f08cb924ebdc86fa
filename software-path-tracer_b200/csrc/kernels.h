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
  kCntPacketNodes = 5,  // primary_packet.cu: node / triangle steps per PACKET, packets handed to the per-ray kernel
  kCntPacketTris = 6,
  kCntPacketBail = 7,
  kCntPacketRays = 8,   // camera rays finished by packets (always counted; also included in kCntRays)
  kCntShadowRays = 9,   // any-hit queries (always counted; also included in kCntRays)
  kCntPackets = 10,     // packets finished (32 * R rays each; always counted)
  kCntSlots = 12
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
// post.cu: resolve + exposure / tone map / sRGB, and the auto-exposure reduction
uint32_t auto_exposure_partials(uint32_t n_pixels);
cudaError_t launch_auto_exposure(const float4* accum, uint32_t n_pixels, uint32_t frame_count, float target, float* partials,
                                 float* exposure_out, cudaStream_t stream);
cudaError_t launch_resolve_post(const float4* accum, uint32_t* out, uint32_t n_pixels, uint32_t frame_count, float exposure,
                                const float* exposure_ptr, int tonemap, int srgb, cudaStream_t stream);

// fused multi-GPU tail: sum over ranks (peer loads, rank order) + post-process + RGBA8 store into the root's buffers
cudaError_t launch_reduce_resolve_peers(const float4* const* accum_by_rank, int n_ranks, uint32_t begin, uint32_t end,
                                        uint32_t total_frames, float exposure, int tonemap, int srgb, uint32_t* root_rgba,
                                        float4* root_accum, int sm_count, cudaStream_t stream);

// refit.cu: BVH8 refit on the device (same arithmetic as the host builder, bvh8_refit.h)
cudaError_t launch_refit(const float* d_vertices, const uint32_t* d_indices, uint32_t nt, void* d_nodes, void* d_tris,
                         void* d_tri_box, void* d_node_box, const uint32_t* level_start, uint32_t n_levels, uint32_t n_nodes,
                         uint64_t* launches, cudaStream_t stream);

// ---- mesh wavefront integrator (wavefront.cu, trace_bvh.cu) -------------------------------------------
// Path state lives in HBM as structure-of-arrays indexed by PATH ID = sample_in_pass * n_pixels + pixel;
// queues hold 4-byte path ids, so compaction moves indices only.
struct WavefrontState {
  float4* ray_o;        // {o.xyz, tnear}
  float4* ray_d;        // {d.xyz, tmax}
  float4* thr;          // {throughput.rgb, bits: bounce count | specular flag << 8}
  float4* L;            // {radiance.rgb, -}: per-path sum, added to the accumulation buffer in sample order
  uint2* hit;           // {t bits, leaf-order triangle index (0xFFFFFFFF = miss)}
  float4* sh_o;         // shadow ray {o.xyz, tmax}
  float4* sh_d;         // {d.xyz, -}
  float4* sh_c;         // contribution added to L when unoccluded
  uint32_t* queue[2];   // ping-pong path queues
  uint32_t* shadow_queue;
  uint32_t* counters;   // 4 * (max_bounces + 1) words, zeroed once per pass (see slot helpers below)
};
__host__ __device__ inline uint32_t slot_queue_count(uint32_t bounce) { return 4u * bounce + 0u; }
__host__ __device__ inline uint32_t slot_shadow_count(uint32_t bounce) { return 4u * bounce + 1u; }
__host__ __device__ inline uint32_t slot_trace_work(uint32_t bounce) { return 4u * bounce + 2u; }
__host__ __device__ inline uint32_t slot_shadow_work(uint32_t bounce) { return 4u * bounce + 3u; }

struct MeshSceneDevice {
  const uint4* nodes;        // Bvh8Node as 5 x uint4
  const float4* tris;        // PackedTri as 3 x float4
  const float4* materials;   // cudapt_material as 2 x float4
  const uint32_t* lights;    // leaf-order indices of emissive triangles, ascending caller index
  uint32_t n_lights;
  uint32_t n_materials;
};

// Division of a 32-bit unsigned by a divisor fixed per pass, without the 25-instruction runtime-division sequence
// (Granlund & Montgomery 1994, the round-up form: exact for every 32-bit dividend): q = (t + ((x - t) >> sh1)) >> sh2
// with t = mulhi(m, x), l = ceil(log2 d), m = floor(2^32 (2^l - d) / d) + 1, sh1 = min(l, 1), sh2 = max(l - 1, 0).
// The path-id decode of the shade kernel ran two runtime divisions per path: a quarter of what every round of that
// kernel executes (profiles/r01_notes.md).
struct FastDiv {
  uint32_t d, m, sh1, sh2;
};
inline FastDiv make_fastdiv(uint32_t d) {
  FastDiv f{d ? d : 1u, 1u, 0u, 0u};
  uint32_t l = 0;
  while (l < 32u && (1ull << l) < (unsigned long long)f.d) ++l;
  f.m = (uint32_t)((((1ull << l) - f.d) << 32) / f.d + 1ull);
  f.sh1 = l < 1u ? l : 1u;
  f.sh2 = l > 1u ? l - 1u : 0u;
  return f;
}
__host__ __device__ inline uint32_t fd_div(const FastDiv& f, uint32_t x) {
#ifdef __CUDA_ARCH__
  const uint32_t t = __umulhi(f.m, x);
#else
  const uint32_t t = (uint32_t)(((unsigned long long)f.m * x) >> 32);
#endif
  return (t + ((x - t) >> f.sh1)) >> f.sh2;
}

struct PassParams {
  WavefrontState st;
  MeshSceneDevice scene;
  CameraParams cam;
  ImageParams im;
  IntegratorParams ip;
  uint32_t n_pixels;
  uint32_t samples;          // samples of every pixel in this pass
  uint32_t chunked;          // path id layout, see path_id(): 0 sample-major, 1 chunked (32 k samples), 2 slot-major (1..16 samples)
  uint32_t s_shift;          // ids of one slot and chunk are 2^s_shift apart from the next slot's: 5 (chunked == 1) or log2(samples) (== 2)
  uint32_t explicit_q0;      // packets: queue 0 is the hit list of primary_packet.cu (kList), not the identity
  uint32_t n_slots;          // chunked layout: pixel slots = 8x8 tiles covering the image, >= n_pixels
  uint32_t slot_tiles_x;
  // Interleaved-tile partition whose tiles are multiples of 8x8 pixels, chunked layout: the slots are COMPACTED to the
  // 8x8 slot tiles this rank owns (owned_tiles[k] = index of the k-th owned slot tile in the image's slot-tile grid,
  // ascending), n_slots = 64 * their number.  Every slot is then owned, so a rank's pass holds world x more samples
  // of its own pixels in the same memory, and the camera-ray packets keep their bounce-0 hit list (primary_packet.cu).
  const uint32_t* owned_tiles;  // nullptr: the slots cover the whole image
  uint32_t dense;               // 1: owned_tiles is in use
  FastDiv fd_slots, fd_tiles_x, fd_width, fd_pixels;  // divisions by n_slots, slot_tiles_x, im.width, n_pixels
  uint32_t first_frame;      // global 1-based frame index of sample 0 of this pass
  uint32_t frame_stride;
  float4* accum;
  uint32_t* aov_prim;        // non-null only for the pass that contains the context's first frame
  float* aov_t;
  unsigned long long* counters;  // global stats (CounterSlot)
};

// PATH ID of (pixel, sample-in-pass).  Sample-major (id = s * n_pixels + pixel: a warp of consecutive ids = 32
// neighbouring pixels) or, when the camera rays are traced as packets (primary_packet.cu: one warp = 32 samples of a
// pixel), CHUNKED: id = ((s / 32) * n_slots + slot) * 32 + s % 32, so that a packet's state is contiguous.  `slot`
// numbers the pixels tile by tile (8x8 tiles, row-major inside a tile; the tile grid is padded to the image, slots
// outside it are skipped): consecutive packets, and with them consecutive entries of the bounce-1 queue, come from a
// compact 2-D neighbourhood of the image instead of a scanline, which is what the caches of the per-ray kernel want.
__host__ __device__ inline uint32_t pixel_slot(const PassParams& p, uint32_t x, uint32_t y) {
  return (((y >> 3) * p.slot_tiles_x + (x >> 3)) << 6) | ((y & 7u) << 3) | (x & 7u);
}
__host__ __device__ inline bool slot_pixel(const PassParams& p, uint32_t g, uint32_t& x, uint32_t& y) {
#ifdef __CUDA_ARCH__
  const uint32_t tile = p.dense ? __ldg(p.owned_tiles + (g >> 6)) : (g >> 6);
#else
  const uint32_t tile = g >> 6;
#endif
  const uint32_t ty = fd_div(p.fd_tiles_x, tile), tx = tile - ty * p.slot_tiles_x;
  x = (tx << 3) | (g & 7u);
  y = (ty << 3) | ((g >> 3) & 7u);
  return x < p.im.width && y < p.im.height;
}
// (pixel -> id: sample-major layout, or the chunked layout over the WHOLE image; a compacted pass goes slot -> pixel only)
__host__ __device__ inline uint32_t path_id(const PassParams& p, uint32_t pixel, uint32_t s) {
  if (!p.chunked) return s * p.n_pixels + pixel;
  const uint32_t y = fd_div(p.fd_width, pixel), x = pixel - y * p.im.width;
  return ((((s >> 5) * p.n_slots + pixel_slot(p, x, y)) << p.s_shift) | (s & ((1u << p.s_shift) - 1u)));  // = slot_sample_id()
}
// every slot of the pass belongs to this rank: no partition, or the compacted slots of one
__host__ __device__ inline bool all_slots_owned(const PassParams& p) { return p.im.part_world <= 1u || p.dense != 0u; }
// Path id of (slot g, sample s) in the two slot-based layouts.  chunked == 2, SLOT-MAJOR (passes of 1, 2, 4, 8 or 16
// samples): id = g * S + s, so 32 consecutive ids = 32 / S neighbouring pixels of an 8x8 tile row block x S samples = one
// camera-ray packet of SEVERAL pixels (primary_packet.cu, kMini) -- what a viewer that renders 1 sample per call, or a
// strong-scaled rank with a few samples per pixel, gets.
// ONE formula serves both: with s_shift = 5 it is the chunked id, and in the slot-major layout s < 32, so the chunk term
// vanishes and (g << s_shift) | s is left.
__host__ __device__ inline uint32_t slot_sample_id(const PassParams& p, uint32_t g, uint32_t s) {
  return (((s >> 5) * p.n_slots + g) << p.s_shift) | (s & ((1u << p.s_shift) - 1u));
}
// false: the id is a padding slot of the chunked layout (no pixel)
__host__ __device__ inline bool path_decode(const PassParams& p, uint32_t id, uint32_t& pixel, uint32_t& s) {
  if (p.chunked) {  // both slot-based layouts: in the slot-major one id >> s_shift < n_slots, so chunk = 0 falls out
    const uint32_t gg = id >> p.s_shift, chunk = fd_div(p.fd_slots, gg);
    uint32_t x, y;
    const bool ok = slot_pixel(p, gg - chunk * p.n_slots, x, y);
    pixel = y * p.im.width + x;
    s = (chunk << 5) | (id & ((1u << p.s_shift) - 1u));
    return ok;
  }
  s = fd_div(p.fd_pixels, id);
  pixel = id - s * p.n_pixels;
  return true;
}

constexpr int kTraceBlock = 128;
constexpr int kTraceBlocksPerSM = 8;
constexpr int kShadeBlock = 256;
constexpr int kShadeBlocksPerSM = 4;

cudaError_t launch_raygen(const PassParams& p, int sm_count, cudaStream_t s);
cudaError_t launch_trace_closest(const PassParams& p, uint32_t bounce, int queue_in, bool counting, int sm_count, cudaStream_t s);
// closest hit for an explicit queue of path ids (the packets' fall-back rays)
cudaError_t launch_trace_queue(const PassParams& p, const uint32_t* queue, uint32_t* count, uint32_t* work, bool counting, int sm_count,
                               cudaStream_t s);
// primary_packet.cu: fused raygen + packet traversal of the camera rays of a pass
bool primary_packets_enabled(const PassParams& p);
// the pass size alone allows packets: a multiple of 32 samples (one pixel per packet) or 1, 2, 4, 8, 16 (several pixels)
bool primary_packet_shape(uint32_t samples);
bool primary_packet_hitlist(const PassParams& p);
cudaError_t launch_primary_packets(const PassParams& p, uint32_t* fb_queue, uint32_t* fb_count, bool counting, int sm_count,
                                   cudaStream_t s);
cudaError_t launch_shade(const PassParams& p, uint32_t bounce, int queue_in, int sm_count, cudaStream_t s);
cudaError_t launch_trace_shadow(const PassParams& p, uint32_t bounce, bool counting, int sm_count, cudaStream_t s);
cudaError_t launch_accumulate(const PassParams& p, int sm_count, cudaStream_t s);

}  // namespace cudapt
