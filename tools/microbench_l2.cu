// L2 / HBM read bandwidth as the BVH traversal sees it ("L2 bandwidth is not in MEASURED_PEAKS.json
// -- measure it with an L2-resident read microbenchmark and record it").
//   stream : every thread reads consecutive 16-B words (coalesced), the whole working set per pass
//   node   : every THREAD reads one 80-B record (5 x 16 B) at a hashed index -- the per-ray kernel's node fetch
//   tri    : every thread reads one 48-B record (3 x 16 B) at a hashed index -- its triangle fetch
//   warp   : every WARP reads one 80-B record, lanes 0..4 one 16-B word each -- the packet kernel's node fetch
// for working sets from L2-resident (C3: 48 MB triangles + 12 MB nodes) to HBM-bound (C4: 1.1 GB).
// Loads are ld.global.cg (L2 only), so L1 hits do not flatter the numbers.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/microbench_l2 tools/microbench_l2.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

__device__ __forceinline__ unsigned hash32(unsigned x) {
  x ^= x >> 16, x *= 0x7feb352du, x ^= x >> 15, x *= 0x846ca68bu, x ^= x >> 16;
  return x;
}

__global__ void __launch_bounds__(256) stream_kernel(const uint4* __restrict__ p, size_t n16, unsigned* out) {
  unsigned acc = 0;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x) {
    const uint4 v = __ldcg(p + i);
    acc ^= v.x ^ v.y ^ v.z ^ v.w;
  }
  if (acc == 0x12345678u) *out = acc;
}

template <int WORDS>
__global__ void __launch_bounds__(256) gather_kernel(const uint4* __restrict__ p, unsigned n_rec, unsigned per_thread, unsigned seed,
                                                     unsigned* out) {
  unsigned acc = 0;
  const unsigned tid = blockIdx.x * blockDim.x + threadIdx.x;
  unsigned h = hash32(tid * 2654435761u + seed);
  for (unsigned k = 0; k < per_thread; ++k) {
    const unsigned r = (unsigned)(((unsigned long long)h * n_rec) >> 32);
    const uint4* q = p + (size_t)r * WORDS;
#pragma unroll
    for (int w = 0; w < WORDS; ++w) {
      const uint4 v = __ldcg(q + w);
      acc ^= v.x ^ v.y ^ v.z ^ v.w;
    }
    h = hash32(h + acc * 0u + k);  // the next address does not depend on the data: loads of successive records overlap
  }
  if (acc == 0x12345678u) *out = acc;
}

__global__ void __launch_bounds__(256) warp_gather_kernel(const uint4* __restrict__ p, unsigned n_rec, unsigned per_warp, unsigned seed,
                                                          unsigned* out) {
  unsigned acc = 0;
  const unsigned warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31u;
  unsigned h = hash32(warp * 2654435761u + seed);
  for (unsigned k = 0; k < per_warp; ++k) {
    const unsigned r = (unsigned)(((unsigned long long)h * n_rec) >> 32);
    if (lane < 5) {
      const uint4 v = __ldcg(p + (size_t)r * 5 + lane);
      acc ^= v.x ^ v.y ^ v.z ^ v.w;
    }
    h = hash32(h + k);
  }
  if (acc == 0x12345678u) *out = acc;
}

#define CK(x)                                                                   \
  do {                                                                          \
    cudaError_t e = (x);                                                        \
    if (e != cudaSuccess) {                                                     \
      std::fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e));              \
      return 1;                                                                 \
    }                                                                           \
  } while (0)

int main(int argc, char** argv) {
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, 0));
  const int sms = prop.multiProcessorCount;
  std::printf("{\"device\": \"%s\", \"sms\": %d, \"l2_bytes\": %d, \"rows\": [\n", prop.name, sms, prop.l2CacheSize);
  const size_t max_bytes = (size_t)2048 << 20;
  uint4* buf;
  unsigned* out;
  CK(cudaMalloc(&buf, max_bytes));
  CK(cudaMalloc(&out, 4));
  CK(cudaMemset(buf, 1, max_bytes));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  const int sizes_mb[] = {8, 16, 32, 48, 64, 96, 128, 192, 256, 512, 1100, 2048};
  const int n_sizes = (int)(sizeof(sizes_mb) / sizeof(sizes_mb[0]));
  const int grid = sms * 8;
  bool first = true;
  for (int si = 0; si < n_sizes; ++si) {
    const size_t bytes = (size_t)sizes_mb[si] << 20;
    const size_t n16 = bytes / 16;
    for (int mode = 0; mode < 4; ++mode) {
      const unsigned words = mode == 1 ? 5u : (mode == 2 ? 3u : (mode == 3 ? 5u : 1u));
      const unsigned n_rec = (unsigned)(n16 / words);
      // about 4 GB of traffic per timed launch, at least two passes over the working set
      const double target = 4e9;
      int reps = 1;
      unsigned per = 0;
      double moved = 0;
      if (mode == 0) {
        reps = (int)(target / (double)bytes) + 2;
        moved = (double)bytes * reps;
      } else if (mode == 3) {
        per = (unsigned)(target / 80.0 / (grid * 8.0)) + 1u;
        moved = 80.0 * per * grid * 8.0;
      } else {
        per = (unsigned)(target / (16.0 * words) / (grid * 256.0)) + 1u;
        moved = 16.0 * words * per * grid * 256.0;
      }
      float best = 1e30f;
      for (int trial = 0; trial < 4; ++trial) {  // trial 0 warms the L2
        CK(cudaEventRecord(e0));
        if (mode == 0)
          for (int r = 0; r < reps; ++r) stream_kernel<<<grid, 256>>>(buf, n16, out);
        else if (mode == 1)
          gather_kernel<5><<<grid, 256>>>(buf, n_rec, per, 17u + trial, out);
        else if (mode == 2)
          gather_kernel<3><<<grid, 256>>>(buf, n_rec, per, 17u + trial, out);
        else
          warp_gather_kernel<<<grid, 256>>>(buf, n_rec, per, 17u + trial, out);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        if (trial && ms < best) best = ms;
      }
      const char* names[] = {"stream", "node80_per_thread", "tri48_per_thread", "node80_per_warp"};
      std::printf("%s  {\"set_MiB\": %d, \"pattern\": \"%s\", \"GBps\": %.1f, \"ms\": %.3f}", first ? "" : ",\n", sizes_mb[si],
                  names[mode], moved / best * 1e-6, best);
      first = false;
    }
  }
  std::printf("\n]}\n");
  return 0;
}
