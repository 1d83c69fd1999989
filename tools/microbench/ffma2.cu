// Micro-benchmark: issue cost of packed FP32 (FFMA2 / FADD2 / FMUL2, sm_100a) against scalar FFMA, alone and
// interleaved with ALU (LOP3), MUFU.RSQ and shared-memory loads.  nvcc -arch=sm_100a -O3 -o ffma2 ffma2.cu
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c) { u64 r; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ u64 fadd2(u64 a, u64 b) { u64 r; asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ u64 fmul2(u64 a, u64 b) { u64 r; asm volatile("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ float ffma(float a, float b, float c) { float r; asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r; }
__device__ __forceinline__ unsigned lop(unsigned a, unsigned b) { unsigned r; asm volatile("xor.b32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
__device__ __forceinline__ float rsq(float a) { float r; asm volatile("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a)); return r; }

// MODE: 0 FFMA x8 | 1 FFMA2 x8 | 2 FFMA2 x8 + LOP x8 | 3 FFMA x8 + LOP x8 | 4 FADD2 x8 | 5 FMUL2 x8
//       6 FFMA2 x8 + MUFU x1 | 7 FFMA x8 + MUFU x1 | 8 MUFU x8 | 9 FFMA2 x8 + LDS.128 x2 | 10 FFMA2 x4 + FFMA x4
//       11 FFMA2 x8 + LOP x8 + MUFU x1 + LDS x1 | 12 FFMA2 x8 + LOP x4
template <int MODE>
__global__ void __launch_bounds__(128) k(float* out, int iters, float fa, float fb) {
  __shared__ float4 sm[256];
  sm[threadIdx.x] = make_float4(fa, fb, fa, fb); sm[threadIdx.x + 128] = make_float4(fb, fa, fb, fa);
  __syncthreads();
  float2 a2 = make_float2(fa, fa), b2 = make_float2(fb, fb);
  u64 A = *(u64*)&a2, B = *(u64*)&b2;
  u64 v[8]; float s[8]; unsigned w[8]; float m = fa + threadIdx.x;
  for (int k = 0; k < 8; ++k) { float2 t = make_float2(threadIdx.x + k, threadIdx.x - k); v[k] = *(u64*)&t; s[k] = threadIdx.x * 0.5f + k; w[k] = threadIdx.x * 77u + k; }
  float4 acc = make_float4(0, 0, 0, 0);
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < 4; ++u) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        if (MODE == 0 || MODE == 3 || MODE == 7) s[k] = ffma(s[k], fa, fb);
        if (MODE == 1 || MODE == 2 || MODE == 6 || MODE == 9 || MODE == 11 || MODE == 12) v[k] = ffma2(v[k], A, B);
        if (MODE == 10) { if (k < 4) v[k] = ffma2(v[k], A, B); else s[k] = ffma(s[k], fa, fb); }
        if (MODE == 4) v[k] = fadd2(v[k], B);
        if (MODE == 5) v[k] = fmul2(v[k], A);
        if (MODE == 2 || MODE == 3 || MODE == 11) w[k] = lop(w[k], w[(k + 1) & 7]);
        if (MODE == 12 && k < 4) w[k] = lop(w[k], w[(k + 1) & 7]);
        if (MODE == 8) s[k] = rsq(s[k]);
      }
      if (MODE == 6 || MODE == 7 || MODE == 11) m = rsq(m);
      if (MODE == 9) { float4 t0 = sm[(it + u) & 255], t1 = sm[(it + u + 64) & 255]; acc.x += t0.x; acc.y += t1.y; }
      if (MODE == 11) { float4 t0 = sm[(it + u) & 255]; acc.x += t0.x; }
    }
  }
  float r = m + acc.x + acc.y;
  for (int k = 0; k < 8; ++k) { float2 t = *(float2*)&v[k]; r += t.x + t.y + s[k] + (float)w[k]; }
  if (r == 12345.678f) out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}
template <int MODE>
void run(const char* name, int nvec) {
  float* out; cudaMalloc(&out, 148 * 5 * 128 * 4);
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  int iters = 20000;
  k<MODE><<<148 * 5, 128>>>(out, 100, 1.000001f, 1e-7f);
  cudaEventRecord(a);
  k<MODE><<<148 * 5, 128>>>(out, iters, 1.000001f, 1e-7f);
  cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  // per SMSP: 5 warps, each iters*4 groups; cycles per group per warp-slot
  double groups = (double)iters * 4 * 5;
  double cyc = ms * 1e-3 * clk * 1e3;
  printf("%-44s %8.3f ms  %6.2f SMSP-cycles per group (%d FP instr/group => %.2f cyc per FP instr)\n", name, ms, cyc / groups, nvec, cyc / groups / nvec);
  cudaFree(out);
}
int main() {
  run<0>("FFMA x8", 8);
  run<1>("FFMA2 x8", 8);
  run<4>("FADD2 x8", 8);
  run<5>("FMUL2 x8", 8);
  run<10>("FFMA2 x4 + FFMA x4", 8);
  run<3>("FFMA x8 + LOP x8", 8);
  run<2>("FFMA2 x8 + LOP x8", 8);
  run<12>("FFMA2 x8 + LOP x4", 8);
  run<8>("MUFU.RSQ x8", 8);
  run<7>("FFMA x8 + MUFU x1", 8);
  run<6>("FFMA2 x8 + MUFU x1", 8);
  run<9>("FFMA2 x8 + LDS.128 x2", 8);
  run<11>("FFMA2 x8 + LOP x8 + MUFU x1 + LDS.128 x1", 8);
  return 0;
}
