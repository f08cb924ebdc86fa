// Byte-pair -> two biased floats: current (2 PRMT) versus proposed (1 PRMT + FFMA.RM + FADD + FFMA), each followed
// by the two slab FFMAs, 4 independent pairs per iteration.  Prints time per pair.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/microbench_pair tools/microbench_pair.cu
#include <cstdio>
#include <cuda_runtime.h>
#define ITER 4096
template <int J> __device__ __forceinline__ float byte_f(unsigned w, unsigned m) {
  unsigned v; asm volatile("prmt.b32 %0, %1, %2, %3;" : "=r"(v) : "r"(w), "r"(m), "n"(0x7540 + J)); return __uint_as_float(v);
}
template <int J> __device__ __forceinline__ float pair_f(unsigned w, unsigned m) {
  unsigned v; asm volatile("prmt.b32 %0, %1, %2, %3;" : "=r"(v) : "r"(w), "r"(m), "n"(J ? 0x7632 : 0x7610)); return __uint_as_float(v);
}
template <int MODE>
__global__ void __launch_bounds__(256) k(float* out, unsigned seed, float adj, float c, unsigned m) {
  unsigned w0 = seed + threadIdx.x, w1 = w0 * 3u, w2 = w0 * 5u, w3 = w0 * 7u;
  float acc0 = 0.f, acc1 = 0.f, acc2 = 0.f, acc3 = 0.f;
#pragma unroll 1
  for (int i = 0; i < ITER; ++i) {
#define PAIR_OLD(w, J, acc) { const float a = byte_f<J>(w, m), b = byte_f<J + 1>(w, m); acc = fmaxf(acc, fmaxf(fmaf(a, adj, c), fmaf(b, adj, c))); }
#define PAIR_NEW(w, J, acc) { const float x = pair_f<J>(w, m); const float B1 = __fmaf_rd(x, 0.00390625f, 8355840.0f); \
    const float b1 = B1 - 8388608.0f; const float B0 = fmaf(b1, -256.0f, x); acc = fmaxf(acc, fmaxf(fmaf(B0, adj, c), fmaf(B1, adj, c))); }
    if (MODE == 0) { PAIR_OLD(w0, 0, acc0) PAIR_OLD(w1, 2, acc1) PAIR_OLD(w2, 0, acc2) PAIR_OLD(w3, 2, acc3) }
    else { PAIR_NEW(w0, 0, acc0) PAIR_NEW(w1, 2, acc1) PAIR_NEW(w2, 0, acc2) PAIR_NEW(w3, 2, acc3) }
    w0 += 0x01010101u; w1 += 0x01010101u; w2 += 0x01010101u; w3 += 0x01010101u;
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc0 + acc1 + acc2 + acc3;
}
template <int MODE> void run(const char* name, float* d) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0), cudaEventCreate(&e1);
  const int blocks = 148 * 8;
  k<MODE><<<blocks, 256>>>(d, 12345u, 0.37f, 1.5f, 0x4B000000u);
  cudaEventRecord(e0);
  k<MODE><<<blocks, 256>>>(d, 12345u, 0.37f, 1.5f, 0x4B000000u);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  const double pairs = (double)blocks * 8 * ITER * 4;
  printf("%-40s %8.3f ms  %6.2f clk/pair/SMSP (at 1.965 GHz)\n", name, ms, ms * 1e-3 * 1.965e9 * 148 * 4 / pairs);
}
int main() {
  float* d; cudaMalloc(&d, 148 * 8 * 256 * 4);
  run<0>("2 PRMT + 2 FFMA (+2 FMNMX)", d);
  run<1>("1 PRMT + RM,FADD,FFMA + 2 FFMA (+2 FMNMX)", d);
  float h[4]; cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost); printf("check %g\n", h[0]);
  return 0;
}
