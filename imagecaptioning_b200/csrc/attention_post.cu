// attention_post.cu -- the two attention kernels that run once per training step, outside the time loop:
// the hoisted dWs pass (backward of W(features), AlexCap/AttentionLanguageModule.py:17) and the doubly-stochastic
// attention regulariser (AlexCap/LSTMwAttentionModel.py:70-71).
#include "common.cuh"
#include "kernels.h"
#include "pipe.cuh"

namespace alexcap {

// ------------------------------------------------------------------------------------------ hoisted dWs
// dWs[b,l,:] = w_v * sum_t de[b,t,l] * (1 - tanh^2(Ws[b,l,:] + q_t[b,:])) : one read of Ws, one write of dWs,
// instead of a read-modify-write of the [B,L,H] accumulator on every reverse step.
// S tanh per element make this kernel MUFU-bound, so the exponential is factored: with r = 1/(1 + e^{2z}),
// 1 - tanh^2 z = 4 r (1 - r) and e^{2(w+q)} = e^{2w} e^{2q}: e^{2w} is computed once per element, e^{2q_t} once per
// (t, column) and shared by the DWS_ROWS rows a thread owns -- 1 + 1/DWS_ROWS MUFU per (element, t) instead of 2.
// (exponents clamped to +-60 so the product of the two factors cannot overflow or hit 0 * inf)
constexpr int DWS_ROWS = 8;
__device__ __forceinline__ float4 exp2c4(const float4 x) {
  float4 y;
  y.x = ex2_ftz(fminf(fmaxf(TANH_C * x.x, -60.f), 60.f)); y.y = ex2_ftz(fminf(fmaxf(TANH_C * x.y, -60.f), 60.f));
  y.z = ex2_ftz(fminf(fmaxf(TANH_C * x.z, -60.f), 60.f)); y.w = ex2_ftz(fminf(fmaxf(TANH_C * x.w, -60.f), 60.f));
  return y;
}
__global__ void __launch_bounds__(256) attn_dws_kernel(int L, int H, int S, int B, const float* __restrict__ Ws,
                                                       const float* __restrict__ q, int64_t q_ld,
                                                       const float* __restrict__ wv, const float* __restrict__ de,
                                                       float* __restrict__ dWs, uint16_t* __restrict__ d_hi,
                                                       uint16_t* __restrict__ d_lo, int64_t d_ld, int ws_exp) {
  const int H4 = H >> 2, gx = (L + DWS_ROWS - 1) / DWS_ROWS, ntiles = gx * B;
  // grid-stride over (image, row group) tiles: the launcher may cap the grid so that the kernel leaves room on every SM for
  // a concurrently running GEMM (see api.cu: dWs is forked next to the lstm dW GEMMs)
  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
  const int b = tile / gx, l0 = (tile % gx) * DWS_ROWS;
  const float* deb = de + (int64_t)b * S * L;
  for (int c = threadIdx.x; c < H4; c += blockDim.x) {
    float4 ew[DWS_ROWS], acc[DWS_ROWS];
#pragma unroll
    for (int r = 0; r < DWS_ROWS; ++r) {
      const int l = l0 + r;
      const float4 wr = (l < L) ? ldg4(reinterpret_cast<const float4*>(Ws) + ((int64_t)b * L + l) * H4 + c) : f4_zero();
      ew[r] = ws_exp ? wr : exp2c4(wr);              // (ws_exp: the hoisted GEMM already stored e^{2w})
      acc[r] = f4_zero();
    }
    for (int t = 0; t < S; ++t) {
      const float4 eq = exp2c4(__ldg(reinterpret_cast<const float4*>(q + ((int64_t)t * B + b) * q_ld) + c));
#pragma unroll
      for (int r = 0; r < DWS_ROWS; ++r) {
        const int l = l0 + r;
        const float d = (l < L) ? __ldg(deb + (int64_t)t * L + l) : 0.f;
        float rr;
        rr = rcp_ftz(fmaf(ew[r].x, eq.x, 1.f)); acc[r].x = fmaf(d, fmaf(-rr, rr, rr), acc[r].x);
        rr = rcp_ftz(fmaf(ew[r].y, eq.y, 1.f)); acc[r].y = fmaf(d, fmaf(-rr, rr, rr), acc[r].y);
        rr = rcp_ftz(fmaf(ew[r].z, eq.z, 1.f)); acc[r].z = fmaf(d, fmaf(-rr, rr, rr), acc[r].z);
        rr = rcp_ftz(fmaf(ew[r].w, eq.w, 1.f)); acc[r].w = fmaf(d, fmaf(-rr, rr, rr), acc[r].w);
      }
    }
    float4 v = __ldg(reinterpret_cast<const float4*>(wv) + c);
    v = make_float4(4.f * v.x, 4.f * v.y, 4.f * v.z, 4.f * v.w);
#pragma unroll
    for (int r = 0; r < DWS_ROWS; ++r) {
      const int l = l0 + r;
      if (l >= L) continue;
      const float4 o = make_float4(v.x * acc[r].x, v.y * acc[r].y, v.z * acc[r].z, v.w * acc[r].w);
      if (dWs) reinterpret_cast<float4*>(dWs)[((int64_t)b * L + l) * H4 + c] = o;
      if (d_hi) {              // bf16 hi (+ lo residual) copies for the pre-split dW_W / dfeat GEMMs: no fp32 round trip
        uint2 h;
        asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(h.x) : "f"(o.y), "f"(o.x));
        asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(h.y) : "f"(o.w), "f"(o.z));
        const int64_t at = ((int64_t)b * L + l) * d_ld + 4 * c;
        *reinterpret_cast<uint2*>(d_hi + at) = h;
        if (d_lo) {
          uint2 lo2;
          asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(lo2.x) : "f"(o.y - __uint_as_float(h.x & 0xffff0000u)), "f"(o.x - __uint_as_float(h.x << 16)));
          asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(lo2.y) : "f"(o.w - __uint_as_float(h.y & 0xffff0000u)), "f"(o.z - __uint_as_float(h.y << 16)));
          *reinterpret_cast<uint2*>(d_lo + at) = lo2;
        }
      }
    }
  }
  }
}

cudaError_t attn_dws(cudaStream_t st, int B, int L, int H, int S, const float* Ws, const float* q, int64_t q_ld,
                     const float* wv, const float* de, float* dWs, uint16_t* d_hi, uint16_t* d_lo, int64_t d_ld, int max_ctas,
                     bool ws_exp) {
  if ((H & 3) || (q_ld & 3) || (d_hi && (d_ld & 3))) return cudaErrorInvalidValue;
  int grid = ceil_div(L, DWS_ROWS) * B;
  if (max_ctas > 0 && grid > max_ctas) grid = max_ctas;
  const int threads = (H / 4 >= 256) ? 256 : ((H / 4 + 31) / 32) * 32;
  ALEXCAP_NOTE_LAUNCH(); attn_dws_kernel<<<grid, threads, 0, st>>>(L, H, S, B, Ws, q, q_ld, wv, de, dWs, d_hi, d_lo, d_ld, ws_exp ? 1 : 0);
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
