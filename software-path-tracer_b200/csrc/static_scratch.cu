// cudapt_auto_exposure_static: the auto-exposure reduction (log-average luminance -> exposure multiplier) for callers
// that cannot call cudaMalloc -- code under CUDA graph capture, tight serving loops, harnesses that forbid it.  It has no
// counterpart in the reference.
//
// The existing path (cudapt_resolve_rgba8 with cudapt_postprocess.auto_exposure, csrc/post.cu) keeps its per-CTA partial
// sums in a buffer the context allocates with cudaMalloc.  This variant keeps them in a FILE-SCOPE `static __device__`
// array of a fixed, modest size (64 KB, held by every GPU that loads the library until the process exits) and allocates
// nothing: no cudaMalloc / cudaFree / cudaMallocHost, no allocating library routine.  The caller passes the input (the
// float4 accumulation buffer, e.g. cudapt_accum_device_ptr) and the output (one float in device memory).
//
// The array's name is only a host-side stand-in in host code, not a device address: the address the kernels get comes
// from cudaGetSymbolAddress, called once per device with that device current (below).
//
// Sharing: every call on a device uses the same array, and the array keeps the last call's contents.  Nothing in it has
// to start at zero -- pass 1 overwrites partials [0, n) and pass 2 reads exactly [0, n) -- so stale values beyond n from a
// larger earlier call are never read.  Calls that OVERLAP on the GPU (two streams, or two processes' worth of threads on
// one device) race on the array: serialise them on one stream or with events.
//
// Arithmetic: the two kernels are variants of log_luminance_partials_kernel / auto_exposure_kernel of csrc/post.cu with
// the same operations in the same order (this TU is compiled with the same -fmad=false / IEEE flags), so the exposure
// equals the one cudapt_get_exposure reports after an auto-exposure resolve of the same buffer, bit for bit.
#include <cstdarg>
#include <cstdio>
#include <mutex>

#include "../../include/cudapt.h"
#include "kernels.h"

namespace cudapt {
void set_thread_error(const char* text);  // api.cu: the text cudapt_last_error(NULL) returns on this thread
}

namespace {

constexpr int kLumBlock = 256, kLumPerThread = 4;  // as csrc/post.cu: one partial per 1024 pixels
constexpr uint32_t kMaxPartials = CUDAPT_STATIC_MAX_PIXELS / (kLumBlock * kLumPerThread);
static_assert(CUDAPT_STATIC_MAX_PIXELS % (kLumBlock * kLumPerThread) == 0, "maximum must be a whole number of CTAs");
static_assert(kMaxPartials * sizeof(float) == 64 * 1024, "the documented size of the static scratch is 64 KB");

// THE scratch: one per GPU that loads the library, until exit.
static __device__ __align__(256) float g_static_lum_partials[kMaxPartials];

// pass 1: per-CTA sum of log(max(luminance, 1e-4)) over 1024 pixels, fixed summation order
__global__ void __launch_bounds__(kLumBlock) static_log_luminance_partials_kernel(const float4* __restrict__ accum, uint32_t n_pixels,
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

// pass 2: one CTA adds partials [0, n_partials) in double and writes exposure = target / exp(mean log luminance)
__global__ void __launch_bounds__(256) static_auto_exposure_kernel(const float* __restrict__ partials, uint32_t n_partials,
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

constexpr int kMaxDevices = 64;
std::mutex g_mu;
float* g_scratch_of_device[kMaxDevices] = {};  // device address of g_static_lum_partials, resolved once per device

int fail_static(int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  cudapt::set_thread_error(buf);
  return code;
}

}  // namespace

extern "C" int cudapt_auto_exposure_static(int device, const void* accum_device, uint32_t n_pixels, uint32_t frame_count,
                                           float target_luminance, float* exposure_device_out, void* stream) {
  if (!accum_device || !exposure_device_out)
    return fail_static(CUDAPT_ERR_INVALID, "cudapt_auto_exposure_static: accum_device and exposure_device_out must not be NULL");
  if (device < 0 || device >= kMaxDevices)
    return fail_static(CUDAPT_ERR_INVALID, "cudapt_auto_exposure_static: device ordinal %d outside [0, %d)", device, kMaxDevices);
  if (n_pixels == 0u || frame_count == 0u)
    return fail_static(frame_count == 0u ? CUDAPT_ERR_NO_FRAMES : CUDAPT_ERR_INVALID,
                       "cudapt_auto_exposure_static: n_pixels %u, frame_count %u: nothing to reduce", n_pixels, frame_count);
  if (n_pixels > CUDAPT_STATIC_MAX_PIXELS)
    return fail_static(CUDAPT_ERR_INVALID,
                       "cudapt_auto_exposure_static: n_pixels %u exceeds CUDAPT_STATIC_MAX_PIXELS (%u): the scratch is a fixed 64 KB "
                       "__device__ array; use cudapt_resolve_rgba8 with cudapt_postprocess.auto_exposure for larger images",
                       n_pixels, (unsigned)CUDAPT_STATIC_MAX_PIXELS);
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess) return fail_static(CUDAPT_ERR_CUDA, "cudapt_auto_exposure_static: cudaSetDevice(%d): %s", device, cudaGetErrorString(e));
  float* scratch = nullptr;
  {
    // Set-up, once per device, with that device current: the DEVICE address of the static array.  (In host code the
    // name g_static_lum_partials is a stand-in object; handing &g_static_lum_partials to a kernel would fault or, on a
    // GPU that can reach host memory, silently use the wrong memory.)
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_scratch_of_device[device]) {
      void* p = nullptr;
      e = cudaGetSymbolAddress(&p, g_static_lum_partials);
      if (e != cudaSuccess || !p)
        return fail_static(CUDAPT_ERR_CUDA, "cudapt_auto_exposure_static: cudaGetSymbolAddress on device %d: %s", device,
                           cudaGetErrorString(e));
      g_scratch_of_device[device] = static_cast<float*>(p);
    }
    scratch = g_scratch_of_device[device];
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const uint32_t nb = (n_pixels + kLumBlock * kLumPerThread - 1) / (kLumBlock * kLumPerThread);  // <= kMaxPartials
  static_log_luminance_partials_kernel<<<nb, kLumBlock, 0, s>>>(static_cast<const float4*>(accum_device), n_pixels, (float)frame_count,
                                                                scratch);
  static_auto_exposure_kernel<<<1, 256, 0, s>>>(scratch, nb, n_pixels, target_luminance, exposure_device_out);
  e = cudaGetLastError();
  if (e != cudaSuccess) return fail_static(CUDAPT_ERR_CUDA, "cudapt_auto_exposure_static: launch: %s", cudaGetErrorString(e));
  return CUDAPT_OK;
}
