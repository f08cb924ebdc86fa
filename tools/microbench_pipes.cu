// Issue-rate microbenchmark for the instructions the BVH8 node test is made of (sm_100a): FFMA, FHFMA
// (fma.rn.f32.f16), PRMT, FMNMX, FMNMX3-style chains, and the ALU/FMA mixes we are considering.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/microbench_pipes tools/microbench_pipes.cu
#include <cstdio>
#include <cuda_runtime.h>
#define ITER 4096
template <int MODE>
__global__ void __launch_bounds__(256) k(float* out, unsigned seed, float fa) {
  float a0 = threadIdx.x * 1e-3f, a1 = a0 + 1.f, a2 = a0 + 2.f, a3 = a0 + 3.f, a4 = a0 + 4.f, a5 = a0 + 5.f, a6 = a0 + 6.f, a7 = a0 + 7.f;
  unsigned w = seed + threadIdx.x, m = 0x4B000000u;
  unsigned short ha = (unsigned short)(seed & 0x3fff);
#pragma unroll 1
  for (int i = 0; i < ITER; ++i) {
    if (MODE == 0) {  // 8 FFMA
      a0 = fmaf(a0, fa, 1.f); a1 = fmaf(a1, fa, 1.f); a2 = fmaf(a2, fa, 1.f); a3 = fmaf(a3, fa, 1.f);
      a4 = fmaf(a4, fa, 1.f); a5 = fmaf(a5, fa, 1.f); a6 = fmaf(a6, fa, 1.f); a7 = fmaf(a7, fa, 1.f);
    } else if (MODE == 1) {  // 8 FHFMA
#define FH(x, sel) asm volatile("fma.rn.f32.f16 %0, %1, %2, %0;" : "+f"(x) : "h"((unsigned short)(w >> sel)), "h"(ha));
      FH(a0, 0) FH(a1, 16) FH(a2, 0) FH(a3, 16) FH(a4, 0) FH(a5, 16) FH(a6, 0) FH(a7, 16)
    } else if (MODE == 2) {  // 8 PRMT (dependent on distinct regs) + xor to keep them alive
      unsigned v0, v1, v2, v3, v4, v5, v6, v7;
      asm volatile("prmt.b32 %0, %1, %2, 0x7540;" : "=r"(v0) : "r"(w), "r"(m));
      asm volatile("prmt.b32 %0, %1, %2, 0x7541;" : "=r"(v1) : "r"(w), "r"(m));
      asm volatile("prmt.b32 %0, %1, %2, 0x7542;" : "=r"(v2) : "r"(w), "r"(m));
      asm volatile("prmt.b32 %0, %1, %2, 0x7543;" : "=r"(v3) : "r"(w), "r"(m));
      asm volatile("prmt.b32 %0, %1, %2, 0x7540;" : "=r"(v4) : "r"(v0), "r"(m));
      asm volatile("prmt.b32 %0, %1, %2, 0x7541;" : "=r"(v5) : "r"(v1), "r"(m));
      asm volatile("prmt.b32 %0, %1, %2, 0x7542;" : "=r"(v6) : "r"(v2), "r"(m));
      asm volatile("prmt.b32 %0, %1, %2, 0x7543;" : "=r"(v7) : "r"(v3), "r"(m));
      w = v4 ^ v5 ^ v6 ^ v7;  // 3 LOP3 -> really 2
    } else if (MODE == 3) {  // 8 FMNMX
      a0 = fmaxf(a0, a1); a1 = fminf(a1, a2); a2 = fmaxf(a2, a3); a3 = fminf(a3, a4);
      a4 = fmaxf(a4, a5); a5 = fminf(a5, a6); a6 = fmaxf(a6, a7); a7 = fminf(a7, a0);
    } else if (MODE == 4) {  // 4 PRMT + 4 FFMA interleaved (current child test flavour)
      unsigned v0, v1, v2, v3;
      asm volatile("prmt.b32 %0, %1, %2, 0x7540;" : "=r"(v0) : "r"(w), "r"(m));
      asm volatile("prmt.b32 %0, %1, %2, 0x7541;" : "=r"(v1) : "r"(w), "r"(m));
      asm volatile("prmt.b32 %0, %1, %2, 0x7542;" : "=r"(v2) : "r"(w), "r"(m));
      asm volatile("prmt.b32 %0, %1, %2, 0x7543;" : "=r"(v3) : "r"(w), "r"(m));
      a0 = fmaf(__uint_as_float(v0), fa, a0); a1 = fmaf(__uint_as_float(v1), fa, a1);
      a2 = fmaf(__uint_as_float(v2), fa, a2); a3 = fmaf(__uint_as_float(v3), fa, a3);
    } else if (MODE == 5) {  // 2 PRMT + 4 FHFMA (proposed)
      unsigned p0, p1;
      asm volatile("prmt.b32 %0, %1, %2, 0x4140;" : "=r"(p0) : "r"(w), "r"(0x64646464u));
      asm volatile("prmt.b32 %0, %1, %2, 0x4342;" : "=r"(p1) : "r"(w), "r"(0x64646464u));
#define FH2(x, p, hi) asm volatile("fma.rn.f32.f16 %0, %1, %2, %0;" : "+f"(x) : "h"((unsigned short)(hi ? (p >> 16) : (p & 0xffff))), "h"(ha));
      FH2(a0, p0, 0) FH2(a1, p0, 1) FH2(a2, p1, 0) FH2(a3, p1, 1)
    } else if (MODE == 6) {  // 4 FFMA + 4 FMNMX: cross-pipe dual issue?
      a0 = fmaf(a0, fa, 1.f); a1 = fmaf(a1, fa, 1.f); a2 = fmaf(a2, fa, 1.f); a3 = fmaf(a3, fa, 1.f);
      a4 = fmaxf(a4, a5); a5 = fminf(a5, a6); a6 = fmaxf(a6, a7); a7 = fminf(a7, a4);
    } else if (MODE == 7) {  // 8 IMAD (a*b+c int)
      unsigned* u = &w;
      unsigned b0 = __float_as_uint(a0), b1 = __float_as_uint(a1), b2 = __float_as_uint(a2), b3 = __float_as_uint(a3);
      b0 = b0 * seed + w; b1 = b1 * seed + w; b2 = b2 * seed + w; b3 = b3 * seed + w;
      a0 = __uint_as_float(b0); a1 = __uint_as_float(b1); a2 = __uint_as_float(b2); a3 = __uint_as_float(b3);
      b0 = __float_as_uint(a4); b1 = __float_as_uint(a5); b2 = __float_as_uint(a6); b3 = __float_as_uint(a7);
      b0 = b0 * seed + w; b1 = b1 * seed + w; b2 = b2 * seed + w; b3 = b3 * seed + w;
      a4 = __uint_as_float(b0); a5 = __uint_as_float(b1); a6 = __uint_as_float(b2); a7 = __uint_as_float(b3);
      (void)u;
    } else if (MODE == 8) {  // 4 FFMA2 (fma.rn.f32x2: two FMAs per lane and instruction) on four register pairs
#define F2(lo, hi) { unsigned long long x, y, z, r; \
      asm volatile("mov.b64 %0, {%1, %2};" : "=l"(x) : "f"(lo), "f"(hi)); \
      asm volatile("mov.b64 %0, {%1, %1};" : "=l"(y) : "f"(fa)); \
      asm volatile("mov.b64 %0, {%1, %1};" : "=l"(z) : "f"(1.0f)); \
      asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(x), "l"(y), "l"(z)); \
      asm volatile("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(r)); }
      F2(a0, a1) F2(a2, a3) F2(a4, a5) F2(a6, a7)
    } else if (MODE == 9) {  // 4 PRMT + 2 FFMA2 (the box test's flavour with paired FMAs)
      unsigned v0, v1, v2, v3;
      asm volatile("prmt.b32 %0, %1, %2, 0x7540;" : "=r"(v0) : "r"(w), "r"(m));
      asm volatile("prmt.b32 %0, %1, %2, 0x7541;" : "=r"(v1) : "r"(w), "r"(m));
      asm volatile("prmt.b32 %0, %1, %2, 0x7542;" : "=r"(v2) : "r"(w), "r"(m));
      asm volatile("prmt.b32 %0, %1, %2, 0x7543;" : "=r"(v3) : "r"(w), "r"(m));
#define F2Q(q0, q1, lo, hi) { unsigned long long x, y, z, r; \
      asm volatile("mov.b64 %0, {%1, %2};" : "=l"(x) : "r"(q0), "r"(q1)); \
      asm volatile("mov.b64 %0, {%1, %1};" : "=l"(y) : "f"(fa)); \
      asm volatile("mov.b64 %0, {%1, %2};" : "=l"(z) : "f"(lo), "f"(hi)); \
      asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(x), "l"(y), "l"(z)); \
      asm volatile("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(r)); }
      F2Q(v0, v1, a0, a1) F2Q(v2, v3, a2, a3)
    }
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7 + __uint_as_float(w);
}
template <int MODE>
void run(const char* name, int instr_per_iter, float* d) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0), cudaEventCreate(&e1);
  const int blocks = 148 * 8;
  k<MODE><<<blocks, 256>>>(d, 12345u, 0.999f);
  cudaEventRecord(e0);
  k<MODE><<<blocks, 256>>>(d, 12345u, 0.999f);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  const double warp_instr = (double)blocks * 8 * ITER * instr_per_iter;
  // per SMSP per cycle at 1.965 GHz: 148 SMs x 4 SMSP
  const double per_clk = warp_instr / (ms * 1e-3) / (148.0 * 4 * 1.965e9);
  printf("%-28s %8.3f ms  %6.3f warp-instr/clk/SMSP (counting %d per iteration)\n", name, ms, per_clk, instr_per_iter);
}
int main() {
  float* d;
  cudaMalloc(&d, 148 * 8 * 256 * 4);
  run<0>("8 FFMA", 8, d);
  run<1>("8 FHFMA", 8, d);
  run<2>("8 PRMT (+2 LOP3)", 10, d);
  run<3>("8 FMNMX", 8, d);
  run<4>("4 PRMT + 4 FFMA", 8, d);
  run<5>("2 PRMT + 4 FHFMA", 6, d);
  run<6>("4 FFMA + 4 FMNMX", 8, d);
  run<7>("8 IMAD", 8, d);
  run<8>("4 FFMA2 (= 8 FMA)", 4, d);
  run<9>("4 PRMT + 2 FFMA2", 6, d);
  return 0;
}
