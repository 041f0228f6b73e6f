// attention_post.cu -- the two attention kernels that run once per training step, outside the time loop:
// the hoisted dWs pass (backward of W(features), AlexCap/AttentionLanguageModule.py:17) and the doubly-stochastic
// attention regulariser (AlexCap/LSTMwAttentionModel.py:70-71).
#include "common.cuh"
#include "kernels.h"

namespace alexcap {

// ------------------------------------------------------------------------------------------ hoisted dWs
// dWs[b,l,:] = w_v * sum_t de[b,t,l] * (1 - tanh^2(Ws[b,l,:] + q_t[b,:])) : one read of Ws, one write of dWs,
// instead of a read-modify-write of the [B,L,H] accumulator on every reverse step.
constexpr int DWS_ROWS = 4;
__global__ void __launch_bounds__(256) attn_dws_kernel(int L, int H, int S, int B, const float* __restrict__ Ws,
                                                       const float* __restrict__ q, int64_t q_ld,
                                                       const float* __restrict__ wv, const float* __restrict__ de,
                                                       float* __restrict__ dWs) {
  const int b = blockIdx.y, l0 = blockIdx.x * DWS_ROWS, H4 = H >> 2;
  const float* deb = de + (int64_t)b * S * L;
  for (int c = threadIdx.x; c < H4; c += blockDim.x) {
    const float4 v = __ldg(reinterpret_cast<const float4*>(wv) + c);
    float4 w[DWS_ROWS], acc[DWS_ROWS];
#pragma unroll
    for (int r = 0; r < DWS_ROWS; ++r) {
      const int l = l0 + r;
      w[r] = (l < L) ? ldg4(reinterpret_cast<const float4*>(Ws) + ((int64_t)b * L + l) * H4 + c) : f4_zero();
      acc[r] = f4_zero();
    }
    for (int t = 0; t < S; ++t) {
      const float4 qt = __ldg(reinterpret_cast<const float4*>(q + ((int64_t)t * B + b) * q_ld) + c);
#pragma unroll
      for (int r = 0; r < DWS_ROWS; ++r) {
        const int l = l0 + r;
        const float d = (l < L) ? __ldg(deb + (int64_t)t * L + l) : 0.f;
        float a;
        a = tanhf(w[r].x + qt.x); acc[r].x = fmaf(d, 1.f - a * a, acc[r].x);
        a = tanhf(w[r].y + qt.y); acc[r].y = fmaf(d, 1.f - a * a, acc[r].y);
        a = tanhf(w[r].z + qt.z); acc[r].z = fmaf(d, 1.f - a * a, acc[r].z);
        a = tanhf(w[r].w + qt.w); acc[r].w = fmaf(d, 1.f - a * a, acc[r].w);
      }
    }
#pragma unroll
    for (int r = 0; r < DWS_ROWS; ++r) {
      const int l = l0 + r;
      if (l < L)
        reinterpret_cast<float4*>(dWs)[((int64_t)b * L + l) * H4 + c] =
            make_float4(v.x * acc[r].x, v.y * acc[r].y, v.z * acc[r].z, v.w * acc[r].w);
    }
  }
}

cudaError_t attn_dws(cudaStream_t st, int B, int L, int H, int S, const float* Ws, const float* q, int64_t q_ld,
                     const float* wv, const float* de, float* dWs) {
  if ((H & 3) || (q_ld & 3)) return cudaErrorInvalidValue;
  dim3 grid(ceil_div(L, DWS_ROWS), B);
  const int threads = (H / 4 >= 256) ? 256 : ((H / 4 + 31) / 32) * 32;
  ALEXCAP_NOTE_LAUNCH(); attn_dws_kernel<<<grid, threads, 0, st>>>(L, H, S, B, Ws, q, q_ld, wv, de, dWs);
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------ regulariser
// LSTMwAttentionModel.py:70 : ((1 - alphas.sum(1))**2).mean(); its gradient wrt alpha[b,t,l] is the same for all t.
__global__ void __launch_bounds__(256) reg_kernel(int B, int S, int L, float weight, const float* __restrict__ alphas,
                                                  float* __restrict__ r, float* __restrict__ partial) {
  __shared__ float scratch[40];
  const int64_t n = (int64_t)B * L;
  float acc = 0.f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int b = (int)(i / L), l = (int)(i % L);
    float s = 0.f;
    for (int t = 0; t < S; ++t) s += __ldg(alphas + ((int64_t)b * S + t) * L + l);
    const float resid = 1.f - s;
    r[i] = -2.f * weight * resid / (float)n;
    acc = fmaf(resid, resid, acc);
  }
  acc = block_sum(acc, scratch);
  if (threadIdx.x == 0) partial[blockIdx.x] = acc;
}
__global__ void __launch_bounds__(256) reg_final_kernel(int nblk, float scale, const float* __restrict__ partial,
                                                        float* __restrict__ out) {
  __shared__ float scratch[40];
  float acc = 0.f;
  for (int i = threadIdx.x; i < nblk; i += blockDim.x) acc += partial[i];
  acc = block_sum(acc, scratch);
  if (threadIdx.x == 0) out[0] = acc * scale;
}

cudaError_t attn_regulariser(cudaStream_t st, int B, int S, int L, float weight, const float* alphas, float* r,
                             float* loss_reg, float* scratch) {
  const int64_t n = (int64_t)B * L;
  const int nblk = (int)((n + 255) / 256 < 256 ? (n + 255) / 256 : 256);
  ALEXCAP_NOTE_LAUNCH(); reg_kernel<<<nblk, 256, 0, st>>>(B, S, L, weight, alphas, r, scratch);
  ALEXCAP_LAUNCH_CHECK();
  ALEXCAP_NOTE_LAUNCH(); reg_final_kernel<<<1, 256, 0, st>>>(nblk, weight / (float)n, scratch, loss_reg);
  return cudaGetLastError();
}

}  // namespace alexcap
