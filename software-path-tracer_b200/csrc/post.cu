// K5': resolve with the post-process stage the reference only sketches -- exposure (commented out at
// CPT.cpp:101-104, RenderSettings::getExposure Types.h:67), auto-exposure (Types.h:58,68-69), tone mapping and
// sRGB encoding (declared, never defined: libs/render/src/utils/Math.h:29-34) -- fused into the one pass
// get_render_result() makes over the accumulation buffer (CPT.cpp:87-117).  With the default parameters
// (exposure 1, no tone map, linear output) every pixel equals the reference's resolve bit for bit: c * 1.0f is
// exact and the other stages are skipped.
//
// Arithmetic contract (-fmad=false): each stage is written as separate IEEE multiplies/adds/divides in the
// order of oracle.post_resolve_rgba8 (numpy float32), so exposure, Reinhard and the ACES fit are bit-exact
// against the oracle; powf (sRGB) and logf (auto-exposure) are library functions, hence 1 LSB tolerance there.
#include "kernels.h"

namespace cudapt {
namespace {

__device__ __forceinline__ float tonemap_channel(float c, int mode) {
  if (mode == 1) return c / (1.0f + c);  // Reinhard
  if (mode == 2) {                       // ACES filmic fit (Narkowicz 2015)
    const float num = c * (2.51f * c + 0.03f);
    const float den = c * (2.43f * c + 0.59f) + 0.14f;
    return num / den;
  }
  return c;
}
__device__ __forceinline__ float clamp01(float v) { return (v < 0.0f) ? 0.0f : ((1.0f < v) ? 1.0f : v); }  // std::clamp
__device__ __forceinline__ float srgb_encode(float c) {
  return c <= 0.0031308f ? 12.92f * c : 1.055f * powf(c, 1.0f / 2.4f) - 0.055f;
}

// exposure_ptr != nullptr: the multiplier comes from the auto-exposure pass (device-resident, no host sync)
__global__ void __launch_bounds__(256) resolve_post_kernel(const float4* __restrict__ accum, uint32_t* __restrict__ out,
                                                           uint32_t n_pixels, float frame_count, float exposure,
                                                           const float* __restrict__ exposure_ptr, int tonemap, int srgb) {
  const uint32_t i = blockIdx.x * 256u + threadIdx.x;
  if (i >= n_pixels) return;
  const float e = exposure_ptr ? *exposure_ptr : exposure;
  const float4 a = accum[i];
  float c[3] = {a.x / frame_count, a.y / frame_count, a.z / frame_count};
  uint32_t b[4];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    float v = c[k] * e;
    v = tonemap_channel(v, tonemap);
    v = clamp01(v);
    if (srgb) v = clamp01(srgb_encode(v));
    b[k] = __float2uint_rz(v * 255.0f) & 0xFFu;  // (uint8_t)(v * 255.0f), CPT.cpp:112
  }
  b[3] = __float2uint_rz(clamp01(a.w / frame_count) * 255.0f) & 0xFFu;
  out[i] = (b[0] << 24) | (b[1] << 16) | (b[2] << 8) | b[3];  // Color.h:9
}

// Auto-exposure, pass 1: per-CTA sum of log(max(luminance, 1e-4)) over 1024 pixels, fixed summation order
// (each thread adds its 4 pixels in index order, then a shared-memory tree) -> deterministic partials.
constexpr int kLumBlock = 256, kLumPerThread = 4;
__global__ void __launch_bounds__(kLumBlock) log_luminance_partials_kernel(const float4* __restrict__ accum, uint32_t n_pixels,
                                                                           float frame_count, float* __restrict__ partials) {
  __shared__ float s[kLumBlock];
  const uint32_t base = (blockIdx.x * kLumBlock + threadIdx.x) * kLumPerThread;
  float sum = 0.0f;
#pragma unroll
  for (int k = 0; k < kLumPerThread; ++k) {
    const uint32_t i = base + k;
    if (i < n_pixels) {
      const float4 a = accum[i];
      const float r = a.x / frame_count, g = a.y / frame_count, bl = a.z / frame_count;
      const float lum = (0.2126f * r + 0.7152f * g) + 0.0722f * bl;
      sum += logf(fmaxf(lum, 1e-4f));
    }
  }
  s[threadIdx.x] = sum;
  __syncthreads();
  for (int w = kLumBlock / 2; w > 0; w >>= 1) {
    if ((int)threadIdx.x < w) s[threadIdx.x] += s[threadIdx.x + w];
    __syncthreads();
  }
  if (threadIdx.x == 0) partials[blockIdx.x] = s[0];
}
// pass 2: one CTA adds the partials in double, in index order per thread + tree, and writes
// exposure = target / exp(mean log luminance)
__global__ void __launch_bounds__(256) auto_exposure_kernel(const float* __restrict__ partials, uint32_t n_partials,
                                                            uint32_t n_pixels, float target, float* __restrict__ exposure_out) {
  __shared__ double s[256];
  double sum = 0.0;
  for (uint32_t i = threadIdx.x; i < n_partials; i += 256u) sum += (double)partials[i];
  s[threadIdx.x] = sum;
  __syncthreads();
  for (int w = 128; w > 0; w >>= 1) {
    if ((int)threadIdx.x < w) s[threadIdx.x] += s[threadIdx.x + w];
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    const float mean_log = (float)(s[0] / (double)n_pixels);
    *exposure_out = target / expf(mean_log);
  }
}

// K6': fused multi-GPU read-back tail (SURVEY.md 8e "fusion opportunity", 7.2 K6): reduce-scatter + resolve +
// gather in ONE kernel per rank over NVLink/NVSwitch peer memory.  Rank r owns the pixel slice
// [begin, end); for each of its pixels it LOADS the float4 accumulation of every rank (its own from HBM, the others
// through peer-mapped pointers = NVLink P2P reads), adds them in rank order (deterministic: ((a0 + a1) + a2) + ...,
// the order a sequential host merge would use), runs the post-process stage and STORES the packed RGBA8 word -- and
// optionally the summed float4 -- straight into the root rank's buffers (P2P writes).  Per rank (G-1)/G of one
// buffer crosses NVLink inbound instead of the whole buffer funnelling into the root.
constexpr int kMaxRanks = 16;
struct PeerPtrs {
  const float4* accum[kMaxRanks];
};
__global__ void __launch_bounds__(256) reduce_resolve_peers_kernel(const PeerPtrs peers, int n_ranks, uint32_t begin, uint32_t end,
                                                                   float total_frames, float exposure, int tonemap, int srgb,
                                                                   uint32_t* __restrict__ root_rgba, float4* __restrict__ root_accum) {
  const uint32_t stride = gridDim.x * 256u;
  for (uint32_t i = begin + blockIdx.x * 256u + threadIdx.x; i < end; i += stride) {
    float4 a = peers.accum[0][i];
    for (int g = 1; g < n_ranks; ++g) {
      const float4 b = peers.accum[g][i];
      a.x += b.x, a.y += b.y, a.z += b.z, a.w += b.w;
    }
    if (root_accum) root_accum[i] = a;
    float c[3] = {a.x / total_frames, a.y / total_frames, a.z / total_frames};
    uint32_t b8[4];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      float v = c[k] * exposure;
      v = tonemap_channel(v, tonemap);
      v = clamp01(v);
      if (srgb) v = clamp01(srgb_encode(v));
      b8[k] = __float2uint_rz(v * 255.0f) & 0xFFu;
    }
    b8[3] = __float2uint_rz(clamp01(a.w / total_frames) * 255.0f) & 0xFFu;
    root_rgba[i] = (b8[0] << 24) | (b8[1] << 16) | (b8[2] << 8) | b8[3];
  }
}

}  // namespace

cudaError_t launch_reduce_resolve_peers(const float4* const* accum_by_rank, int n_ranks, uint32_t begin, uint32_t end,
                                        uint32_t total_frames, float exposure, int tonemap, int srgb, uint32_t* root_rgba,
                                        float4* root_accum, int sm_count, cudaStream_t stream) {
  if (n_ranks < 1 || n_ranks > kMaxRanks) return cudaErrorInvalidValue;
  if (end <= begin) return cudaSuccess;
  PeerPtrs p{};
  for (int g = 0; g < n_ranks; ++g) p.accum[g] = accum_by_rank[g];
  const uint32_t want = (end - begin + 255u) / 256u;
  const uint32_t cap = (uint32_t)sm_count * 8u;
  reduce_resolve_peers_kernel<<<want < cap ? want : cap, 256, 0, stream>>>(p, n_ranks, begin, end, (float)total_frames, exposure,
                                                                           tonemap, srgb, root_rgba, root_accum);
  return cudaGetLastError();
}

uint32_t auto_exposure_partials(uint32_t n_pixels) { return (n_pixels + kLumBlock * kLumPerThread - 1) / (kLumBlock * kLumPerThread); }

cudaError_t launch_auto_exposure(const float4* accum, uint32_t n_pixels, uint32_t frame_count, float target, float* partials,
                                 float* exposure_out, cudaStream_t stream) {
  const uint32_t nb = auto_exposure_partials(n_pixels);
  log_luminance_partials_kernel<<<nb, kLumBlock, 0, stream>>>(accum, n_pixels, (float)frame_count, partials);
  auto_exposure_kernel<<<1, 256, 0, stream>>>(partials, nb, n_pixels, target, exposure_out);
  return cudaGetLastError();
}

cudaError_t launch_resolve_post(const float4* accum, uint32_t* out, uint32_t n_pixels, uint32_t frame_count, float exposure,
                                const float* exposure_ptr, int tonemap, int srgb, cudaStream_t stream) {
  resolve_post_kernel<<<(n_pixels + 255u) / 256u, 256, 0, stream>>>(accum, out, n_pixels, (float)frame_count, exposure,
                                                                    exposure_ptr, tonemap, srgb);
  return cudaGetLastError();
}

}  // namespace cudapt
