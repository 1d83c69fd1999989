// Micro-benchmark: throughput of shared-memory integer atomics (32/64-bit) with spread addresses, the
// float->fixed conversions that feed them, and the STS+LDS slot alternative.  nvcc -arch=sm_100a -O3.
#include <cstdio>
#include <cuda_runtime.h>
template <int MODE>
__global__ void k(float* out, int iters, float scale) {
  __shared__ unsigned long long acc64[512];
  __shared__ int acc32[1024];
  __shared__ float slots[4096];
  int tid = threadIdx.x;
  for (int i = tid; i < 512; i += blockDim.x) acc64[i] = 0;
  for (int i = tid; i < 1024; i += blockDim.x) acc32[i] = 0;
  __syncthreads();
  float v = tid * 0.37f + 1.0f;
  unsigned idx = tid * 7u;
  float s = 0.f;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      idx = idx * 1664525u + 1013904223u;
      v = v * 1.0001f + 0.5f;
      if (MODE == 0) atomicAdd(&acc32[(idx >> 8) & 383], __float2int_rn(v * scale));
      if (MODE == 1) atomicAdd(&acc64[(idx >> 8) & 383], (unsigned long long)__float2ll_rn(v * scale));
      if (MODE == 2) { slots[(idx >> 8) & 4095] = v; }
      if (MODE == 3) { s += slots[(idx >> 8) & 4095]; }
      if (MODE == 4) { s += (float)__float2int_rn(v * scale); }
      if (MODE == 5) { s += (float)__float2ll_rn(v * scale); }
      if (MODE == 6) { s += v; }
    }
  }
  __syncthreads();
  if (MODE <= 1 || s == 1234.5f) out[blockIdx.x * blockDim.x + tid] = s + acc32[tid & 1023] + (float)acc64[tid & 511] + slots[tid];
}
template <int MODE>
void run(const char* name) {
  float* out; cudaMalloc(&out, 148 * 4 * 128 * 4);
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  int iters = 20000;
  k<MODE><<<148 * 4, 128>>>(out, 100, 262144.f);
  cudaEventRecord(a);
  k<MODE><<<148 * 4, 128>>>(out, iters, 262144.f);
  cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  double ops = 148.0 * 4 * 4 * iters * 8;  // warp-level ops
  int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  printf("%-28s %8.3f ms  %.2f SM-cycles per warp-op (at %d MHz, 16 warps/SM)\n", name, ms,
         ms * 1e-3 * clk * 1e3 * 148 / ops, clk / 1000);
  cudaFree(out);
}
int main() {
  run<6>("baseline loop (lcg+fma)");
  run<0>("f2i + ATOMS.32 spread");
  run<1>("f2ll + ATOMS.64 spread");
  run<2>("STS.32 scattered");
  run<3>("LDS.32 scattered");
  run<4>("f2i only");
  run<5>("f2ll only");
  return 0;
}
