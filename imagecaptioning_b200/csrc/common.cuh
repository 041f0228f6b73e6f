// common.cuh -- shared device helpers for libalexcap_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>

namespace alexcap {

#define ACK(expr)                                  \
  do {                                             \
    cudaError_t e__ = (expr);                      \
    if (e__ != cudaSuccess) return e__;            \
  } while (0)

#define ALEXCAP_LAUNCH_CHECK() ACK(cudaGetLastError())

// Number of kernel launches this library has enqueued (host-side counter; bench.py reports it as gpu_launches).
unsigned long long& launch_counter();
#define ALEXCAP_NOTE_LAUNCH() (++alexcap::launch_counter())

// Programmatic dependent launch (PDL).  The per-timestep chain (query GEMM -> attention -> gates GEMM -> LSTM pointwise,
// and its mirror image in the backward pass) is ~140 short kernels; with PDL the next kernel's CTAs are scheduled
// while the previous kernel drains and run their prologue (barrier init, TMEM allocation, tensor-map prefetch)
// before blocking in pdl_wait() until the previous grid has completed and flushed its writes.  Both instructions
// are no-ops in a kernel launched without the attribute.  Rule: no global read of produced data and no global
// write before pdl_wait().
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
bool& pdl_enabled();     // api.cu; ALEXCAP_PDL=0 disables

template <typename... KArgs, typename... Args>
static inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                                     Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

static inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
static inline int64_t ceil_div64(int64_t a, int64_t b) { return (a + b - 1) / b; }
static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// Block-wide reductions; `scratch` holds >= 33 floats of shared memory. All threads get the result.
// blockDim.x must be a multiple of 32 (<= 1024).  Contains two __syncthreads().
__device__ __forceinline__ float block_sum(float v, float* scratch) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  v = warp_sum(v);
  if (lane == 0) scratch[warp] = v;
  __syncthreads();
  if (warp == 0) {
    float r = (lane < nwarp) ? scratch[lane] : 0.f;
    r = warp_sum(r);
    if (lane == 0) scratch[32] = r;
  }
  __syncthreads();
  float out = scratch[32];
  __syncthreads();
  return out;
}
__device__ __forceinline__ float block_max(float v, float* scratch) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  v = warp_max(v);
  if (lane == 0) scratch[warp] = v;
  __syncthreads();
  if (warp == 0) {
    float r = (lane < nwarp) ? scratch[lane] : -INFINITY;
    r = warp_max(r);
    if (lane == 0) scratch[32] = r;
  }
  __syncthreads();
  float out = scratch[32];
  __syncthreads();
  return out;
}

__device__ __forceinline__ float sigmoid_f(float x) { return 1.f / (1.f + expf(-x)); }

// Two-MUFU sigmoid / tanh (ex2.approx + rcp.approx): ~2e-7 absolute error, saturating correctly (e -> inf gives rcp = 0).
__device__ __forceinline__ float sigmoid_fast(float x) {
  float e, r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(-1.4426950408889634f * x));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(e + 1.f));
  return r;
}
__device__ __forceinline__ float tanh_fast2(float x) {
  float e, r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(2.8853900817779268f * x));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(e + 1.f));
  return fmaf(-2.f, r, 1.f);
}

// Exponential-factored tanh (attention scores): tanh(w + q) = 1 - 2 r, r = 1 / (1 + e^{2w} e^{2q}).  EXP2C(x) = 2^(clamp(C x,
// +-60)) with C = 2 log2(e) is what the hoisted Ws = W(features) GEMM stores (activation ACT_EXP2C) and what the per-step
// kernels make of q, so an element costs ONE MUFU op (rcp) instead of two.  The clamp keeps the product of the two factors
// finite and away from 0 * inf; it only changes a result when |w| or |q| exceeds 20.8 AND they nearly cancel.
__device__ __forceinline__ float exp2c(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(fminf(fmaxf(2.8853900817779268f * x, -60.f), 60.f)));
  return y;
}
// epilogue activations of the GEMM engines (enum Act in kernels.h: 1 = tanh, 2 = relu, 3 = exp2c)
__device__ __forceinline__ float act_apply(float x, int act) {
  if (act == 1) return tanhf(x);
  if (act == 2) return fmaxf(x, 0.f);
  if (act == 3) return exp2c(x);
  return x;
}

__device__ __forceinline__ float4 ldg4(const float4* p) { return __ldg(p); }

__device__ __forceinline__ float4 f4_zero() { return make_float4(0.f, 0.f, 0.f, 0.f); }
__device__ __forceinline__ void f4_fma(float4& acc, float s, const float4& v) {
  acc.x = fmaf(s, v.x, acc.x); acc.y = fmaf(s, v.y, acc.y); acc.z = fmaf(s, v.z, acc.z); acc.w = fmaf(s, v.w, acc.w);
}
__device__ __forceinline__ void f4_add(float4& acc, const float4& v) {
  acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
}

}  // namespace alexcap
