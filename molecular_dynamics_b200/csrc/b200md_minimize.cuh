// b200md_minimize.cuh -- one iteration of batched FIRE (Bitzek et al., PRL 97, 170201) between two force evaluations,
// as ONE kernel: the surroundings of the step in the reference's tutorial (upstream torchmd.minimizers.minimize_bfgs at
// V_Calculations/Test-1/tutorial.ipynb cell 14 relaxes one replica on the host; this is its batched on-GPU form,
// SURVEY.md 8f N3).  One CTA per replica (grid-stride): a fixed-order block reduction of F.v, |v|^2, |F|^2 and max|F|,
// the per-replica time-step / mixing update, the velocity update with a second reduction for the displacement cap,
// the position update.  HBM-bound: reads F, v, x once (v twice), writes v, x once -- 48 bytes per atom per iteration.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/b200md.h"

namespace b200md {

constexpr int kFireThreads = 256;

struct FireRed { float power, v2, f2, fmax; };

__device__ __forceinline__ float fire_block_sum(float v, float* s, int tid) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if ((tid & 31) == 0) s[tid >> 5] = v;
  __syncthreads();
  float t = 0.0f;
#pragma unroll
  for (int w = 0; w < kFireThreads / 32; ++w) t += s[w];  // fixed order: bit-reproducible
  return t;
}
__device__ __forceinline__ float fire_block_max(float v, float* s, int tid) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  __syncthreads();
  if ((tid & 31) == 0) s[tid >> 5] = v;
  __syncthreads();
  float t = 0.0f;
#pragma unroll
  for (int w = 0; w < kFireThreads / 32; ++w) t = fmaxf(t, s[w]);
  return t;
}

// state[r] = (dt, alpha, steps since the last uphill move, max|F| seen by this call)
__global__ void __launch_bounds__(kFireThreads)
fire_update_kernel(const long long R, const int n3, float* __restrict__ pos, float* __restrict__ vel,
                   const float* __restrict__ forces, float4* __restrict__ state, const b200md_fire_opts o) {
  __shared__ float sred[kFireThreads / 32];
  const int tid = threadIdx.x;
  for (long long r = blockIdx.x; r < R; r += gridDim.x) {
    const float* f = forces + r * n3;
    float* v = vel + r * n3;
    float* x = pos + r * n3;
    float power = 0.0f, v2 = 0.0f, f2 = 0.0f, fm = 0.0f;
    for (int k = tid; k < n3; k += kFireThreads) {
      const float fk = f[k], vk = v[k];
      power = fmaf(fk, vk, power); v2 = fmaf(vk, vk, v2); f2 = fmaf(fk, fk, f2); fm = fmaxf(fm, fabsf(fk));
    }
    power = fire_block_sum(power, sred, tid);
    v2 = fire_block_sum(v2, sred, tid);
    f2 = fire_block_sum(f2, sred, tid);
    fm = fire_block_max(fm, sred, tid);
    float4 st = state[r];
    const float mix = st.y * sqrtf(v2) / fmaxf(sqrtf(f2), 1e-30f);  // alpha |v| / |F|
    const float keep = 1.0f - st.y;
    const bool up = power > 0.0f;
    float npos = up ? st.z + 1.0f : 0.0f;
    float dt = st.x, alpha = st.y;
    if (up && npos > (float)o.n_min) { dt = fminf(dt * o.f_inc, o.dt_max); alpha *= o.f_alpha; }
    if (!up) { dt *= o.f_dec; alpha = o.alpha_start; }
    float smax = 0.0f;
    for (int k = tid; k < n3; k += kFireThreads) {
      const float fk = f[k];
      float vk = up ? fmaf(keep, v[k], mix * fk) : 0.0f;  // v <- (1 - alpha) v + alpha |v| F / |F| ; uphill: stop
      vk = fmaf(dt, fk, vk);
      v[k] = vk;
      smax = fmaxf(smax, fabsf(dt * vk));
    }
    smax = fire_block_max(smax, sred, tid);
    const float scale = dt * (o.max_move / fmaxf(smax, o.max_move));  // no coordinate moves more than max_move
    for (int k = tid; k < n3; k += kFireThreads) x[k] = fmaf(scale, v[k], x[k]);
    if (tid == 0) state[r] = make_float4(dt, alpha, npos, fm);
    __syncthreads();
  }
}

}  // namespace b200md
